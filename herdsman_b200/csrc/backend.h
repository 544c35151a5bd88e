// backend.h -- host runtime behind the C ABI (include/herdb.h): devices, streams, device-resident
// keys, the level-batched circuit executor and the task queue that stands in for herdsman's
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

namespace herdb {

struct DeviceKey {
    u32 n = 0;
    u32* bsk = nullptr;          // [n][2][8][2][1024] NTT domain, * 1/N
    u16* ksk = nullptr;          // [N][128][2][ks_stride]
    u32 ks_stride = 0;
};

struct Device {
    int id = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy = nullptr;                  // second stream: uploads / downloads that ride under kernels
    cudaEvent_t ev_piece[3] = {};
    u32 *ftw2 = nullptr, *itw1 = nullptr, *mono = nullptr;
    std::map<std::string, DeviceKey> keys;
    // scratch, grown on demand
    u16* slots = nullptr;  size_t slots_bytes = 0;
    u16* ext = nullptr;    size_t ext_bytes = 0;
    GateDesc* gates = nullptr; size_t gates_cap = 0;
    OutDesc* outs = nullptr;   size_t outs_cap = 0;
    u16* io_in = nullptr;  size_t io_in_bytes = 0;     // staging for the host-buffer entry points
    u16* io_out = nullptr; size_t io_out_bytes = 0;
    // timing
    struct Timed { cudaEvent_t a, b; int kind; };
    std::vector<Timed> timed;  size_t timed_used = 0;
    herdb_stats stats{};
    std::mutex mu;               // one host thread drives a device at a time
};

struct InFrame { std::string df; u32 part = 0; uint64_t row_count = 0; };
struct MapTaskReq {                                  // a MapTask, or a ReduceTask when `reduce` is set
    std::string session, in_df, out_df;
    u32 in_part = 0, out_part = 0;
    uint64_t row_count = 0;
    bool reduce = false;
    std::vector<InFrame> inputs;                     // ReduceTask::input_data_frame_ptrs
    CompiledCircuit circuit;
    herdb_done_fn done = nullptr;
    void* user = nullptr;
};

}  // namespace herdb

struct herdb_ctx {
    std::vector<herdb::Device*> devs;
    std::string storage_dir, key_dir;
    mutable std::mutex err_mu;
    std::string last_error;
    size_t mem_budget = 0;       // bytes of scratch per device for slots + ext
    // task queue (herdb_map_async)
    std::mutex q_mu;
    std::condition_variable q_cv, idle_cv;
    std::deque<herdb::MapTaskReq> queue;
    size_t in_flight = 0;
    bool stop = false;
    std::thread worker;

    void set_error(const std::string& s) { std::lock_guard<std::mutex> l(err_mu); last_error = s; }
};
