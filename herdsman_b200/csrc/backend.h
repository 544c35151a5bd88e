// backend.h -- host runtime behind the C ABI (include/herdb.h): devices, streams, device-resident
// keys, the level-batched circuit executor, the chunk pipeline that moves rows between partition
// files / host buffers and HBM under the kernels, and the task queue that stands in for herdsman's
// out-of-process worker (IWorkerGroup, include/execution/worker/i_worker_group.hpp:12-49).
#pragma once
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/herdb.h"
#include "circuit.h"
#include "kernels.h"

struct ncclComm;

namespace herdb {

struct DeviceKey {
    u32 n = 0;
    u32* bsk = nullptr;          // [n][2][8][2][1024] NTT domain, * 1/N
    u16* ksk = nullptr;          // [N][128][2][ks_stride]
    u32 ks_stride = 0;
};

// One pipelined evaluation's own buffers on one device (a task of either task thread, or a caller of the host-buffer
// entry points): pinned host staging and device row buffers for two row chunks (double buffering), the gate tables of
// its circuit(s), and the events that order uploads, kernels and downloads of consecutive chunks.  Kernels of every
// lane go to the device's one kernel stream: a second kernel stream per lane (consecutive chunks in flight together,
// to fill part-empty grid rounds) was measured and is slower -- the persistent blind-rotation kernel holds every SM, so
// the other stream's key switch cannot start before it ends and the streams serialise, at half the jobs per launch
// (profiles/r02d_two_streams_*.json*).
struct LaneHalf {
    void* h_in = nullptr;  size_t h_in_cap = 0;         // pinned host: framed file bytes / staged caller rows
    void* h_out = nullptr; size_t h_out_cap = 0;
    u16* d_in = nullptr;   size_t d_in_cap = 0;
    u16* d_out = nullptr;  size_t d_out_cap = 0;
    cudaEvent_t ev_in = nullptr, ev_done = nullptr, ev_out = nullptr;
};
struct Lane {
    LaneHalf half[2];
    u16* d_mid[2] = {};    size_t d_mid_cap[2] = {};    // rows between chained map stages (never leave HBM)
    GateDesc* gates = nullptr; size_t gates_cap = 0;
    OutDesc* outs = nullptr;   size_t outs_cap = 0;
    bool busy = false;
};

struct Device {
    int id = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;                // kernels
    cudaStream_t copy = nullptr;                  // uploads that ride under kernels
    cudaStream_t copy_out = nullptr;              // downloads that ride under kernels
    cudaEvent_t ev_piece[3] = {};
    u32 *ftw2 = nullptr, *itw1 = nullptr, *mono = nullptr;
    std::map<std::string, DeviceKey> keys;
    // scratch shared by everything that runs on this device's kernel stream (stream order serialises its users)
    u16* slots = nullptr;  size_t slots_bytes = 0;
    u16* ext = nullptr;    size_t ext_bytes = 0;
    GateDesc* gates = nullptr; size_t gates_cap = 0;   // single-gate batches / device-resident entry points
    OutDesc* outs = nullptr;   size_t outs_cap = 0;
    u16* io_in = nullptr;  size_t io_in_bytes = 0;     // staging of herdb_eval_gate_batch
    u16* io_out = nullptr; size_t io_out_bytes = 0;
    // herdb_eval_gate_batch_dev_noalloc: device addresses of the library's static scratch (looked up once per device)
    u16* noalloc_ext = nullptr; GateDesc* noalloc_gate = nullptr; bool noalloc_ready = false;
    std::vector<Lane*> lanes;
    // timing: CUDA events around the two hot kernels, summed into `stats` once they have completed
    struct Timed { cudaEvent_t a, b; int kind; };
    std::deque<Timed> timed;
    std::vector<cudaEvent_t> free_events;
    herdb_stats stats{};
    std::mutex mu;               // one host thread issues work to a device at a time
};

struct OwnedCircuit {                                // deep copy of a herdb_circuit (the caller's buffers die with the call)
    std::vector<herdb_node> nodes;
    std::vector<herdb_edge> edges;
    std::vector<uint8_t> in_types, out_types;
    std::vector<uint32_t> in_offs;
    void assign(const herdb_circuit* c);
    herdb_circuit view() const;
};

struct InFrame { std::string df; u32 part = 0; uint64_t row_count = 0; };
struct TaskReq {                                     // a MapTask, or a ReduceTask when `reduce` is set
    std::string session, in_df, out_df;
    u32 in_part = 0, out_part = 0;
    uint64_t row_count = 0;
    int schema_type = 1;
    bool reduce = false;
    std::vector<InFrame> inputs;                     // ReduceTask::input_data_frame_ptrs
    OwnedCircuit circuit;
    herdb_done_fn done = nullptr;
    void* user = nullptr;
};

// what the context knows about a session's key beyond the per-device copies
struct KeyInfo {
    u32 n = 0;
    int schema_type = 1;
    bool from_file = false;                          // loaded from <key_dir>/<session>/<schema>.key: revalidated, evictable
    std::string path;
    int64_t size = 0, mtime_ns = 0; uint64_t ino = 0;
    uint64_t last_use = 0;
    int pins = 0;                                    // tasks evaluating with it right now
    size_t dev_bytes = 0;                            // HBM per device
};

}  // namespace herdb

struct herdb_ctx {
    std::vector<herdb::Device*> devs;
    std::string storage_dir, key_dir;
    mutable std::mutex err_mu;
    std::string last_error;
    size_t mem_budget = 0;       // bytes of scratch per device for slots + ext
    size_t io_budget = 0;        // bytes per staging buffer of a lane (rows are chunked to fit)
    size_t key_budget = 0;       // bytes of key material per device before file-backed keys are evicted (LRU)
    uint64_t coalesce_rows = 0;  // queued map tasks of one session with one circuit are evaluated together up to this many rows
    int task_window = 1;         // tasks an IWorkerGroup on this context should keep outstanding (concurrent_workers())
    // keys
    std::mutex key_mu;
    std::map<std::string, herdb::KeyInfo> key_info;
    uint64_t key_clock = 0;
    std::vector<ncclComm*> nccl;                     // one communicator per device, created by the first multi-device key load
    double last_key_load_ms = 0, last_key_bcast_ms = 0;
    // task queue (herdb_map_async / herdb_reduce_async)
    std::mutex q_mu;
    std::condition_variable q_cv, idle_cv;
    std::deque<herdb::TaskReq> queue;
    size_t in_flight = 0;
    int lingering = 0;           // task threads collecting a group of small tasks right now (worker_loop)
    bool stop = false;
    std::vector<std::thread> workers;

    void set_error(const std::string& s) { std::lock_guard<std::mutex> l(err_mu); last_error = s; }
};
