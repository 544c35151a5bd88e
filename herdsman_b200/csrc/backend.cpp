// backend.cpp -- implementation of the C ABI in include/herdb.h (see backend.h).
#include "backend.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>

#include <nccl.h>
#include <nvtx3/nvToolsExt.h>

#include "host_tables.h"

using namespace herdb;

namespace {

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ctx->set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                    \
            return e_ == cudaErrorMemoryAllocation ? HERDB_ERR_NOMEM : HERDB_ERR_CUDA;             \
        }                                                                                          \
    } while (0)

struct NvtxRange {                                   // shows up in Nsight timelines; costs nothing without a tool attached
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int grow(herdb_ctx* ctx, void** p, size_t* cap, size_t need) {
    if (*cap >= need) return HERDB_OK;
    if (*p) cudaFree(*p);                            // synchronises with everything queued on the device
    *p = nullptr; *cap = 0;
    cudaError_t e = cudaMalloc(p, need);
    if (e != cudaSuccess) { cudaGetLastError(); ctx->set_error(std::string("cudaMalloc: ") + cudaGetErrorString(e)); return HERDB_ERR_NOMEM; }
    *cap = need;
    return HERDB_OK;
}
int grow_pinned(herdb_ctx* ctx, void** p, size_t* cap, size_t need) {
    if (*cap >= need) return HERDB_OK;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; *cap = 0;
    cudaError_t e = cudaHostAlloc(p, need, cudaHostAllocPortable);
    if (e != cudaSuccess) { cudaGetLastError(); ctx->set_error(std::string("cudaHostAlloc: ") + cudaGetErrorString(e)); return HERDB_ERR_NOMEM; }
    *cap = need;
    return HERDB_OK;
}

// ---------------------------------------------------------------------------------------------- kernel timing
// CUDA events around the two hot kernels, on the launching stream; summed into herdb_stats once complete
int timed_begin(herdb_ctx* ctx, Device& d, int kind, cudaStream_t stream) {
    Device::Timed t{}; t.kind = kind;
    for (cudaEvent_t* e : {&t.a, &t.b}) {
        if (!d.free_events.empty()) { *e = d.free_events.back(); d.free_events.pop_back(); }
        else CU(cudaEventCreate(e));
    }
    d.timed.push_back(t);
    CU(cudaEventRecord(t.a, stream));
    return HERDB_OK;
}
int timed_end(herdb_ctx* ctx, Device& d, cudaStream_t stream) {
    CU(cudaEventRecord(d.timed.back().b, stream));
    return HERDB_OK;
}
void account(Device& d, const Device::Timed& t) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) {
        if (t.kind == 0) { d.stats.blind_rotate_ms += ms; d.stats.blind_rotate_launches++; }
        else { d.stats.keyswitch_ms += ms; d.stats.keyswitch_launches++; }
    }
    d.free_events.push_back(t.a); d.free_events.push_back(t.b);
}
void drain_ready(Device& d) {                        // intervals that have completed (no waiting); several streams feed the list
    for (auto it = d.timed.begin(); it != d.timed.end();) {
        if (cudaEventQuery(it->b) == cudaSuccess) { account(d, *it); it = d.timed.erase(it); } else ++it;
    }
    cudaGetLastError();                              // cudaErrorNotReady is not an error
}
int drain(herdb_ctx* ctx, Device& d) {              // the device's own kernel stream (single-stream entry points)
    CU(cudaStreamSynchronize(d.stream));
    drain_ready(d);
    return HERDB_OK;
}

const DeviceKey* find_key(Device& d, const char* session) {
    auto it = d.keys.find(session ? session : "");
    return it == d.keys.end() ? nullptr : &it->second;
}

// ---------------------------------------------------------------------------------------------- one level, one circuit
// one circuit level (or a single-gate pseudo level): blind rotation then key switch, `rows` rows
struct Exec { cudaStream_t stream; u16* slots; u16* ext; };     // where a level runs: kernel stream + scratch
int run_level(herdb_ctx* ctx, Device& d, const Exec& ex, const DeviceKey& key, const GateDesc* d_gates, u32 n_gates,
              const Pool pools[4], u32 rows, u32* d_trace = nullptr, u32 steps = UINT32_MAX, int shape = 0) {
    const u32 jobs = n_gates * rows;
    if (!jobs) return HERDB_OK;
    BrParams bp{};
    bp.bsk = key.bsk; bp.mono = d.mono; bp.ftw2 = d.ftw2; bp.itw1 = d.itw1; bp.gates = d_gates;
    for (int i = 0; i < 4; ++i) bp.pools[i] = pools[i];
    bp.ext = ex.ext; bp.trace = d_trace; bp.n = key.n; bp.steps = std::min(steps, key.n); bp.rows = rows; bp.jobs = jobs;
    bp.job0 = 0; bp.job1 = jobs;
    int launches = 0;
    int rc = timed_begin(ctx, d, 0, ex.stream); if (rc) return rc;
    CU(launch_blind_rotate(bp, d.sm_count, ex.stream, shape, &launches));
    rc = timed_end(ctx, d, ex.stream); if (rc) return rc;
    KsParams kp{};
    kp.ksk = key.ksk; kp.ext = ex.ext; kp.gates = d_gates;
    for (int i = 0; i < 4; ++i) kp.pools[i] = pools[i];
    kp.ks_stride = key.ks_stride; kp.n = key.n; kp.rows = rows; kp.jobs = jobs;
    rc = timed_begin(ctx, d, 1, ex.stream); if (rc) return rc;
    CU(launch_keyswitch(kp, d.sm_count, ex.stream));
    rc = timed_end(ctx, d, ex.stream); if (rc) return rc;
    d.stats.kernel_launches += (uint64_t)launches + 1;
    d.stats.gates_bootstrapped += jobs;
    return HERDB_OK;
}

size_t scratch_per_row(const CompiledCircuit& cc, u32 n) {
    return (size_t)cc.slots * (n + 1) * 2 + (size_t)cc.max_level_width * kExtStride * 2;
}

// all levels of one compiled circuit over `rows` rows that are resident on the device; the gate and output
// tables are already there (d_gates = all levels back to back).  Asynchronous: kernels are queued on d.stream.
int issue_circuit(herdb_ctx* ctx, Device& d, const Exec& ex, const DeviceKey& key, const CompiledCircuit& cc, const GateDesc* d_gates,
                  const OutDesc* d_outs, Pool in, Pool out, bool out_framed, u32 rows) {
    if (!rows) return HERDB_OK;
    const u32 w = key.n + 1;
    Pool pools[4] = {};
    pools[kPoolIn] = in;
    pools[kPoolSlots] = Pool{ex.slots, cc.slots * w, 0};
    pools[kPoolOut] = out;
    size_t off = 0;
    for (auto& L : cc.levels) {
        int rc = run_level(ctx, d, ex, key, d_gates + off, (u32)L.size(), pools, rows); if (rc) return rc;
        off += L.size();
    }
    GatherParams gp{};
    gp.outs = d_outs; for (int i = 0; i < 4; ++i) gp.pools[i] = pools[i];
    gp.n = key.n; gp.rows = rows; gp.out_bits = cc.out_bits;
    gp.framed = out_framed ? 1u : 0u; gp.frame_len = cc.out_bits * w * 2;
    CU(launch_gather_outputs(gp, d.sm_count, ex.stream));
    d.stats.kernel_launches += 1;
    return HERDB_OK;
}

int ensure_scratch(herdb_ctx* ctx, u16** slots, size_t* slots_bytes, u16** ext, size_t* ext_bytes, const CompiledCircuit& cc, u32 n, size_t rows) {
    int rc = grow(ctx, (void**)slots, slots_bytes, std::max<size_t>(16, rows * cc.slots * (n + 1) * 2)); if (rc) return rc;
    return grow(ctx, (void**)ext, ext_bytes, std::max<size_t>(16, rows * cc.max_level_width * kExtStride * 2));
}

int upload_tables(herdb_ctx* ctx, cudaStream_t stream, const std::vector<const CompiledCircuit*>& stages, GateDesc** gates, size_t* gates_cap,
                  OutDesc** outs, size_t* outs_cap, std::vector<size_t>& gate_off, std::vector<size_t>& out_off) {
    std::vector<GateDesc> flat; std::vector<OutDesc> oflat;
    for (const CompiledCircuit* cc : stages) {
        gate_off.push_back(flat.size()); out_off.push_back(oflat.size());
        for (auto& L : cc->levels) flat.insert(flat.end(), L.begin(), L.end());
        oflat.insert(oflat.end(), cc->outs.begin(), cc->outs.end());
    }
    int rc = grow(ctx, (void**)gates, gates_cap, std::max<size_t>(1, flat.size()) * sizeof(GateDesc)); if (rc) return rc;
    rc = grow(ctx, (void**)outs, outs_cap, std::max<size_t>(1, oflat.size()) * sizeof(OutDesc)); if (rc) return rc;
    if (!flat.empty()) CU(cudaMemcpyAsync(*gates, flat.data(), flat.size() * sizeof(GateDesc), cudaMemcpyHostToDevice, stream));
    if (!oflat.empty()) CU(cudaMemcpyAsync(*outs, oflat.data(), oflat.size() * sizeof(OutDesc), cudaMemcpyHostToDevice, stream));
    CU(cudaStreamSynchronize(stream));               // `flat` / `oflat` are pageable host memory
    return HERDB_OK;
}

// ---------------------------------------------------------------------------------------------- keys
int key_check(herdb_ctx* ctx, const herdb_key_header& h, size_t len) {
    if (len < sizeof h || h.magic != HERDB_KEY_MAGIC || h.version != HERDB_KEY_VERSION) { ctx->set_error("key blob: bad magic/version"); return HERDB_ERR_FORMAT; }
    if (h.ring_n != kN || h.Q != kQ || h.q != kLweQ || h.qKS != kQKS || h.baseG_log != HERDB_BASEG_LOG || h.digitsG != kDG ||
        h.baseKS_log != HERDB_BASEKS_LOG || h.digitsKS != kDKS || h.n == 0 || h.n > kLweMax) {
        ctx->set_error("key blob: parameter set not supported by this build (see include/herdb_params.h)"); return HERDB_ERR_FORMAT;
    }
    if (len < herdb_key_blob_size(h.n)) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
    return HERDB_OK;
}

size_t key_dev_bytes(u32 n) {
    const size_t polys = (size_t)n * 2 * kDG2 * 2, ks_rows = (size_t)kN * kBaseKS * kDKS;
    return polys * kN * 4 + ks_rows * (((n + 1 + 7) / 8) * 8) * 2;
}

// raw blob already in device memory -> kernel-ready key.  Queues the conversion kernels on d.stream; the caller
// synchronises (several devices convert at the same time) and then calls install_key_commit.
int install_key_issue(herdb_ctx* ctx, Device& d, const herdb_key_header& h, const uint8_t* d_blob, DeviceKey& k) {
    CU(cudaSetDevice(d.id));
    k = DeviceKey{}; k.n = h.n;
    const u32 polys = h.n * 2 * kDG2 * 2;
    const size_t ks_rows = (size_t)kN * kBaseKS * kDKS;
    k.ks_stride = ((h.n + 1 + 7) / 8) * 8;
    CU(cudaMalloc((void**)&k.bsk, (size_t)polys * kN * 4));
    cudaError_t e = cudaMalloc((void**)&k.ksk, ks_rows * k.ks_stride * 2);
    if (e != cudaSuccess) { cudaFree(k.bsk); k.bsk = nullptr; ctx->set_error("key: out of device memory"); return HERDB_ERR_NOMEM; }
    const u32* raw_bsk = reinterpret_cast<const u32*>(d_blob + sizeof(herdb_key_header));
    const u16* raw_ksk = reinterpret_cast<const u16*>(d_blob + sizeof(herdb_key_header) + (size_t)polys * kN * 4);
    CU(launch_bsk_convert(raw_bsk, k.bsk, d.ftw2, polys, d.sm_count, d.stream));
    CU(launch_ksk_relayout(raw_ksk, k.ksk, h.n, k.ks_stride, ks_rows, d.sm_count, d.stream));
    d.stats.kernel_launches += 3;
    return HERDB_OK;
}
void install_key_commit(Device& d, const char* session, const DeviceKey& k) {
    auto it = d.keys.find(session);
    if (it != d.keys.end()) { cudaFree(it->second.bsk); cudaFree(it->second.ksk); }
    d.keys[session] = k;
}

void drop_device_keys(herdb_ctx* ctx, const std::string& session) {
    for (Device* d : ctx->devs) {
        std::lock_guard<std::mutex> l(d->mu);
        auto it = d->keys.find(session);
        if (it == d->keys.end()) continue;
        cudaSetDevice(d->id);
        cudaStreamSynchronize(d->stream);
        cudaFree(it->second.bsk); cudaFree(it->second.ksk);
        d->keys.erase(it);
    }
}

// key_mu held.  File-backed keys that no task is using are dropped, least recently used first, until the key
// material per device fits the budget again (each STD128 session is 324 MiB on every GPU).
void evict_keys(herdb_ctx* ctx, const std::string& keep) {
    for (;;) {
        size_t total = 0;
        for (auto& kv : ctx->key_info) total += kv.second.dev_bytes;
        if (total <= ctx->key_budget) return;
        auto victim = ctx->key_info.end();
        for (auto it = ctx->key_info.begin(); it != ctx->key_info.end(); ++it)
            if (it->first != keep && it->second.from_file && it->second.pins == 0 &&
                (victim == ctx->key_info.end() || it->second.last_use < victim->second.last_use)) victim = it;
        if (victim == ctx->key_info.end()) return;
        drop_device_keys(ctx, victim->first);
        ctx->key_info.erase(victim);
    }
}

// key_mu held.  The blob goes to the first device from the host and from there to every other device of the
// context in ONE NCCL broadcast over NVLink / NVSwitch; every device then converts its copy
// (coefficient -> NTT domain, gadget-row folding, row padding) at the same time.
int load_key_blob(herdb_ctx* ctx, const char* session, const void* blob, size_t len, KeyInfo meta) {
    NvtxRange range("herdb:key_load");
    herdb_key_header h;
    if (len < sizeof h) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
    memcpy(&h, blob, sizeof h);
    int rc = key_check(ctx, h, len); if (rc) return rc;
    const size_t need = herdb_key_blob_size(h.n);
    const size_t G = ctx->devs.size();
    const double t0 = now_ms();
    std::vector<std::unique_lock<std::mutex>> locks;
    for (Device* d : ctx->devs) locks.emplace_back(d->mu);           // always in device order
    std::vector<uint8_t*> raw(G, nullptr);
    std::vector<DeviceKey> keys(G);
    auto cleanup = [&](bool drop_keys) {
        for (size_t di = 0; di < G; ++di) {
            cudaSetDevice(ctx->devs[di]->id);
            if (raw[di]) cudaFree(raw[di]);
            if (drop_keys) { if (keys[di].bsk) cudaFree(keys[di].bsk); if (keys[di].ksk) cudaFree(keys[di].ksk); }
        }
    };
    auto body = [&]() -> int {
        for (size_t di = 0; di < G; ++di) {
            CU(cudaSetDevice(ctx->devs[di]->id));
            CU(cudaMalloc((void**)&raw[di], need));
        }
        Device& d0 = *ctx->devs[0];
        CU(cudaSetDevice(d0.id));
        CU(cudaMemcpyAsync(raw[0], blob, need, cudaMemcpyHostToDevice, d0.stream));
        d0.stats.h2d_bytes += need;
        double bcast_ms = 0;
        if (G > 1) {
            if (ctx->nccl.empty()) {
                std::vector<int> ids; for (Device* d : ctx->devs) ids.push_back(d->id);
                std::vector<ncclComm_t> comms(G);
                ncclResult_t nr = ncclCommInitAll(comms.data(), (int)G, ids.data());
                if (nr != ncclSuccess) { ctx->set_error(std::string("ncclCommInitAll: ") + ncclGetErrorString(nr)); return HERDB_ERR_CUDA; }
                ctx->nccl.assign(comms.begin(), comms.end());
            }
            CU(cudaStreamSynchronize(d0.stream));
            const double tb = now_ms();
            ncclResult_t nr = ncclGroupStart();
            for (size_t di = 0; di < G && nr == ncclSuccess; ++di)
                nr = ncclBroadcast(raw[di], raw[di], need, ncclUint8, 0, ctx->nccl[di], ctx->devs[di]->stream);
            if (nr == ncclSuccess) nr = ncclGroupEnd(); else ncclGroupEnd();
            if (nr != ncclSuccess) { ctx->set_error(std::string("ncclBroadcast: ") + ncclGetErrorString(nr)); return HERDB_ERR_CUDA; }
            for (size_t di = 0; di < G; ++di) { CU(cudaSetDevice(ctx->devs[di]->id)); CU(cudaStreamSynchronize(ctx->devs[di]->stream)); }
            bcast_ms = now_ms() - tb;
        }
        for (size_t di = 0; di < G; ++di) { int r2 = install_key_issue(ctx, *ctx->devs[di], h, raw[di], keys[di]); if (r2) return r2; }
        for (size_t di = 0; di < G; ++di) { CU(cudaSetDevice(ctx->devs[di]->id)); CU(cudaStreamSynchronize(ctx->devs[di]->stream)); }
        ctx->last_key_bcast_ms = bcast_ms;
        return HERDB_OK;
    };
    rc = body();
    if (rc) { cleanup(true); return rc; }
    for (size_t di = 0; di < G; ++di) install_key_commit(*ctx->devs[di], session, keys[di]);
    cleanup(false);
    ctx->last_key_load_ms = now_ms() - t0;
    for (Device* d : ctx->devs) { d->stats.key_load_ms = ctx->last_key_load_ms; d->stats.key_bcast_ms = ctx->last_key_bcast_ms; }
    locks.clear();
    meta.n = h.n; meta.dev_bytes = key_dev_bytes(h.n); meta.last_use = ++ctx->key_clock;
    auto old = ctx->key_info.find(session);
    if (old != ctx->key_info.end()) meta.pins = old->second.pins;
    ctx->key_info[session] = meta;
    evict_keys(ctx, session);
    return HERDB_OK;
}

std::string key_path(const herdb_ctx* ctx, const std::string& session, int schema_type) {
    // <key_dir>/<session_uuid>/<int(SchemaType)>.key   (key_service.cpp:28-31)
    return ctx->key_dir + "/" + session + "/" + std::to_string(schema_type) + ".key";
}

// key_mu held
int load_key_file(herdb_ctx* ctx, const std::string& session, int schema_type) {
    KeyInfo meta; meta.schema_type = schema_type; meta.from_file = true; meta.path = key_path(ctx, session, schema_type);
    const int fd = open(meta.path.c_str(), O_RDONLY);
    if (fd < 0) { ctx->set_error("cannot open key file " + meta.path); return HERDB_ERR_IO; }
    struct stat st{};
    if (fstat(fd, &st) != 0) { close(fd); ctx->set_error("cannot stat key file " + meta.path); return HERDB_ERR_IO; }
    meta.size = st.st_size; meta.mtime_ns = (int64_t)st.st_mtim.tv_sec * 1000000000ll + st.st_mtim.tv_nsec; meta.ino = st.st_ino;
    std::vector<char> buf((size_t)st.st_size);
    size_t got = 0;
    while (got < buf.size()) {
        const ssize_t r = read(fd, buf.data() + got, buf.size() - got);
        if (r <= 0) break;
        got += (size_t)r;
    }
    close(fd);
    if (got != buf.size()) { ctx->set_error("cannot read key file " + meta.path); return HERDB_ERR_IO; }
    return load_key_blob(ctx, session.c_str(), buf.data(), buf.size(), meta);
}

// A task is about to evaluate with the session's key: make sure the device copies are the key that is on disk NOW.
// KeyService::add_key overwrites the file of a session that uploads a key again and remove_key deletes it
// (key_service.cpp:19-60); a cached copy that no longer matches would give COMPLETED tasks with wrong
// ciphertexts.  The file's size / mtime / inode are compared before every task; a changed file is reloaded, a missing
// one fails the task.  Keys that came through herdb_load_key_mem/_dev have no file and are used as they are.
int pin_session_key(herdb_ctx* ctx, const std::string& session, int schema_type, u32* n_out) {
    std::lock_guard<std::mutex> l(ctx->key_mu);
    auto it = ctx->key_info.find(session);
    bool load = it == ctx->key_info.end();
    if (!load && it->second.from_file) {
        struct stat st{};
        if (it->second.schema_type != schema_type) load = true;
        else if (stat(it->second.path.c_str(), &st) != 0) {
            if (it->second.pins == 0) { drop_device_keys(ctx, session); ctx->key_info.erase(it); }
            ctx->set_error("key file of the session is gone: " + key_path(ctx, session, schema_type));
            return HERDB_ERR_IO;
        } else {
            const int64_t mt = (int64_t)st.st_mtim.tv_sec * 1000000000ll + st.st_mtim.tv_nsec;
            load = st.st_size != it->second.size || mt != it->second.mtime_ns || (uint64_t)st.st_ino != it->second.ino;
        }
        if (load && it->second.pins > 0) {           // another task is evaluating with the old key: let it finish first
            ctx->set_error("key of the session changed while a task of the session is running");
            return HERDB_ERR_IO;
        }
    }
    if (load) {
        int rc = load_key_file(ctx, session, schema_type); if (rc) return rc;
        it = ctx->key_info.find(session);
    }
    it->second.pins++;
    it->second.last_use = ++ctx->key_clock;
    *n_out = it->second.n;
    return HERDB_OK;
}
void unpin_session_key(herdb_ctx* ctx, const std::string& session) {
    std::lock_guard<std::mutex> l(ctx->key_mu);
    auto it = ctx->key_info.find(session);
    if (it != ctx->key_info.end() && it->second.pins > 0) it->second.pins--;
}

// ---------------------------------------------------------------------------------------------- the chunk pipeline
// Rows flow  source -> pinned host buffer -> HBM -> [levels of stage 0 .. stage S-1] -> HBM -> pinned host buffer -> sink
// in chunks, double buffered, one host thread per device:
//   * while the kernels of chunk c run, the thread reads chunk c+1 from the partition file (or the caller's
//     buffer), uploads it on the `copy` stream, and writes chunk c-1 out after its download on `copy_out`;
//   * rows of a partition file are used IN PLACE on the device: the kernels read ciphertexts out of the framed
//     byte image ([u32 length][payload] per row, storage_service.cpp:19-28) and the output kernel writes the
//     framed image of the output partition, so the host touches no row: pread -> cudaMemcpyAsync, and
//     cudaMemcpyAsync -> pwrite.
struct FileSeg { int src_fd; int dst_fd; size_t rows; };    // one task's input partition file, its output file, its rows
struct RowIo {
    const uint8_t* src_mem = nullptr;                        // dense rows in host memory ...
    uint8_t* dst_mem = nullptr;
    std::vector<FileSeg> files;                              // ... or partition files: the rows of several coalesced tasks, concatenated
    bool src_pinned = false, dst_pinned = false;             // caller memory is page-locked: DMA straight from / to it
};
struct EvalPlan {
    std::vector<CompiledCircuit> stages;
    size_t chunk_rows = 1;
    u32 n = 0;
};

bool is_pinned(const void* p) {
    cudaPointerAttributes a{};
    if (!p || cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// compiles the stages for the chunk size one device will run (the level scheduler fills grid rounds for that many rows)
int make_plan(herdb_ctx* ctx, const herdb_circuit* const* circuits, size_t n_stages, size_t n_rows, u32 n, EvalPlan& plan) {
    const size_t G = std::max<size_t>(1, ctx->devs.size());
    const size_t rows_dev = (n_rows + G - 1) / G;
    const size_t w = n + 1;
    plan.n = n;
    plan.stages.assign(n_stages, CompiledCircuit());
    size_t scratch = 1, io_row = 1; u32 widest = 1;
    for (size_t k = 0; k < n_stages; ++k) {
        std::string err;
        if (!circuits[k]) return HERDB_ERR_ARG;
        int rc = compile_circuit(circuits[k], plan.stages[k], err);                       // facts first (as-soon-as-possible levels)
        if (rc) { ctx->set_error("circuit" + (n_stages > 1 ? " of stage " + std::to_string(k) : std::string()) + ": " + err); return rc; }
        if (k && plan.stages[k].in_bits != plan.stages[k - 1].out_bits) {
            ctx->set_error("pipeline: stage " + std::to_string(k) + " does not take the row stage " + std::to_string(k - 1) + " produces");
            return HERDB_ERR_ARG;
        }
        scratch = std::max(scratch, scratch_per_row(plan.stages[k], n) * 5 / 4);           // a moved schedule may need a few more slots
        io_row = std::max(io_row, (size_t)std::max(plan.stages[k].in_bits, plan.stages[k].out_bits) * w * 2 + 4);
        widest = std::max(widest, plan.stages[k].max_level_width);
    }
    size_t chunk = std::max<size_t>(1, std::min(ctx->mem_budget / scratch, ctx->io_budget / io_row));
    chunk = std::min<size_t>(chunk, (size_t)0x7FFFFFFF / widest);
    chunk = std::min(chunk, std::max<size_t>(1, rows_dev));
    const size_t nch = (std::max<size_t>(1, rows_dev) + chunk - 1) / chunk;
    chunk = (std::max<size_t>(1, rows_dev) + nch - 1) / nch;                                // equal chunks: no short last one
    SchedHint hint; hint.rows = chunk; hint.wave = ctx->devs.empty() ? 0 : (u32)ctx->devs[0]->sm_count * kBrCtsPerCta;
    for (size_t k = 0; k < n_stages && n_rows; ++k) {
        std::string err;
        CompiledCircuit cc;
        if (compile_circuit(circuits[k], cc, err, hint) == HERDB_OK && scratch_per_row(cc, n) * chunk <= std::max(ctx->mem_budget, scratch_per_row(cc, n)))
            plan.stages[k] = std::move(cc);
    }
    plan.chunk_rows = chunk;
    return HERDB_OK;
}

Lane* acquire_lane(Device& d) {                      // d.mu held
    for (Lane* L : d.lanes) if (!L->busy) { L->busy = true; return L; }
    Lane* L = new Lane; L->busy = true;
    for (LaneHalf& h : L->half)
        for (cudaEvent_t* e : {&h.ev_in, &h.ev_done, &h.ev_out}) cudaEventCreateWithFlags(e, cudaEventDisableTiming);
    d.lanes.push_back(L);
    return L;
}
void free_lane(Lane* L) {
    for (LaneHalf& h : L->half) {
        if (h.h_in) cudaFreeHost(h.h_in);
        if (h.h_out) cudaFreeHost(h.h_out);
        cudaFree(h.d_in); cudaFree(h.d_out);
        for (cudaEvent_t e : {h.ev_in, h.ev_done, h.ev_out}) if (e) cudaEventDestroy(e);
    }
    cudaFree(L->d_mid[0]); cudaFree(L->d_mid[1]); cudaFree(L->gates); cudaFree(L->outs);
    delete L;
}

int full_pread(int fd, void* buf, size_t len, off_t off) {
    size_t got = 0;
    while (got < len) {
        const ssize_t r = pread(fd, (char*)buf + got, len - got, off + (off_t)got);
        if (r <= 0) return -1;
        got += (size_t)r;
    }
    return 0;
}
int full_pwrite(int fd, const void* buf, size_t len, off_t off) {
    size_t put = 0;
    while (put < len) {
        const ssize_t r = pwrite(fd, (const char*)buf + put, len - put, off + (off_t)put);
        if (r <= 0) return -1;
        put += (size_t)r;
    }
    return 0;
}

// rows [row, row + cnt) of the concatenated file segments <-> buf (`stride` bytes per framed row)
int files_rw(const RowIo& io, size_t row, size_t cnt, size_t stride, void* buf, bool write) {
    size_t seg0 = 0;
    for (const FileSeg& sg : io.files) {
        const size_t lo = std::max(row, seg0), hi = std::min(row + cnt, seg0 + sg.rows);
        if (lo < hi) {
            char* p = (char*)buf + (lo - row) * stride;
            const off_t off = (off_t)((lo - seg0) * stride);
            const int rc = write ? full_pwrite(sg.dst_fd, p, (hi - lo) * stride, off) : full_pread(sg.src_fd, p, (hi - lo) * stride, off);
            if (rc) return rc;
        }
        seg0 += sg.rows;
    }
    return 0;
}

// rows [r0, r0 + rows) of the task on one device
int device_pipeline(herdb_ctx* ctx, Device& d, const char* session, const EvalPlan& plan, const RowIo& io, size_t r0, size_t rows) {
    if (!rows) return HERDB_OK;
    NvtxRange range("herdb:device_slice");
    const size_t S = plan.stages.size(), w = plan.n + 1;
    const CompiledCircuit& first = plan.stages.front(); const CompiledCircuit& last = plan.stages.back();
    const bool in_framed = !io.files.empty(), out_framed = !io.files.empty();
    const size_t in_payload = (size_t)first.in_bits * w * 2, out_payload = (size_t)last.out_bits * w * 2;
    const size_t in_stride = in_payload + (in_framed ? 4 : 0), out_stride = out_payload + (out_framed ? 4 : 0);
    size_t mid_row = 0;
    for (size_t k = 0; k + 1 < S; ++k) mid_row = std::max(mid_row, (size_t)plan.stages[k].out_bits * w * 2);
    // equal chunks (make_plan)
    const size_t chunk = std::max<size_t>(1, std::min(plan.chunk_rows, rows));
    const size_t nch = (rows + chunk - 1) / chunk;
    const bool stage_in = !(io.src_mem && io.src_pinned), stage_out = !(io.dst_mem && io.dst_pinned);

    Lane* L = nullptr;
    const DeviceKey* key = nullptr;
    std::vector<const CompiledCircuit*> stage_ptrs; for (auto& cc : plan.stages) stage_ptrs.push_back(&cc);
    std::vector<size_t> gate_off, out_off;
    {   // buffers, tables, scratch
        std::lock_guard<std::mutex> l(d.mu);
        CU(cudaSetDevice(d.id));
        key = find_key(d, session);
        if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
        L = acquire_lane(d);
        int rc = HERDB_OK;
        for (size_t b = 0; b < std::min<size_t>(2, nch) && !rc; ++b) {
            LaneHalf& h = L->half[b];
            rc = grow(ctx, (void**)&h.d_in, &h.d_in_cap, std::max<size_t>(16, chunk * in_stride));
            if (!rc) rc = grow(ctx, (void**)&h.d_out, &h.d_out_cap, std::max<size_t>(16, chunk * out_stride));
            if (!rc && stage_in) rc = grow_pinned(ctx, &h.h_in, &h.h_in_cap, std::max<size_t>(16, chunk * in_stride));
            if (!rc && stage_out) rc = grow_pinned(ctx, &h.h_out, &h.h_out_cap, std::max<size_t>(16, chunk * out_stride));
        }
        for (int m = 0; m < 2 && !rc && S > 1; ++m) rc = grow(ctx, (void**)&L->d_mid[m], &L->d_mid_cap[m], std::max<size_t>(16, chunk * mid_row));
        for (size_t k = 0; k < S && !rc; ++k) rc = ensure_scratch(ctx, &d.slots, &d.slots_bytes, &d.ext, &d.ext_bytes, plan.stages[k], plan.n, chunk);
        if (!rc) rc = upload_tables(ctx, d.stream, stage_ptrs, &L->gates, &L->gates_cap, &L->outs, &L->outs_cap, gate_off, out_off);
        if (rc) { L->busy = false; return rc; }
    }
    auto release = [&](int rc) {
        std::lock_guard<std::mutex> l(d.mu);
        cudaSetDevice(d.id);
        if (rc) { cudaStreamSynchronize(d.stream); cudaStreamSynchronize(d.copy); cudaStreamSynchronize(d.copy_out); }   // else: every chunk's download has been awaited
        drain_ready(d);
        L->busy = false;
        return rc;
    };

    int rc = HERDB_OK;
    for (size_t c = 0; c <= nch && !rc; ++c) {
        if (c < nch) {
            LaneHalf& h = L->half[c & 1];
            const size_t row = r0 + c * chunk, cnt = std::min(chunk, rows - c * chunk);
            const void* h_src = nullptr;
            if (in_stride) {
                if (stage_in) {
                    if (c >= 2 && cudaEventSynchronize(h.ev_in) != cudaSuccess) { ctx->set_error("upload failed"); rc = HERDB_ERR_CUDA; break; }
                    if (in_framed) {
                        if (files_rw(io, row, cnt, in_stride, h.h_in, false) != 0) { ctx->set_error("short read of the input partition"); rc = HERDB_ERR_IO; break; }
                        for (size_t r = 0; r < cnt; ++r) {                                 // every row announces exactly the circuit's input row
                            uint32_t len; memcpy(&len, (const uint8_t*)h.h_in + r * in_stride, 4);
                            if (len != in_payload) { rc = HERDB_ERR_FORMAT; break; }
                        }
                        if (rc) { ctx->set_error("input partition: row length does not match the circuit's input columns"); break; }
                    } else memcpy(h.h_in, io.src_mem + row * in_stride, cnt * in_stride);
                    h_src = h.h_in;
                } else h_src = io.src_mem + row * in_stride;
            }
            std::lock_guard<std::mutex> l(d.mu);
            NvtxRange chunk_range("herdb:chunk_issue");
            auto issue = [&]() -> int {
                CU(cudaSetDevice(d.id));
                if (c >= 2) CU(cudaStreamWaitEvent(d.copy, h.ev_done, 0));               // kernels of chunk c-2 have read d_in
                if (in_stride) CU(cudaMemcpyAsync(h.d_in, h_src, cnt * in_stride, cudaMemcpyHostToDevice, d.copy));
                CU(cudaEventRecord(h.ev_in, d.copy));
                CU(cudaStreamWaitEvent(d.stream, h.ev_in, 0));
                if (c >= 2) CU(cudaStreamWaitEvent(d.stream, h.ev_out, 0));               // download of chunk c-2 has read d_out
                const Exec ex{d.stream, d.slots, d.ext};
                for (size_t k = 0; k < S; ++k) {
                    const CompiledCircuit& cc = plan.stages[k];
                    Pool in = k == 0 ? Pool{h.d_in + (in_framed ? 2 : 0), (u32)(in_stride / 2), 0} : Pool{L->d_mid[(k - 1) & 1], cc.in_bits * (u32)w, 0};
                    Pool out = k + 1 == S ? Pool{h.d_out + (out_framed ? 2 : 0), (u32)(out_stride / 2), 0} : Pool{L->d_mid[k & 1], cc.out_bits * (u32)w, 0};
                    int r2 = issue_circuit(ctx, d, ex, *key, cc, L->gates + gate_off[k], L->outs + out_off[k], in, out, k + 1 == S && out_framed, (u32)cnt);
                    if (r2) return r2;
                }
                CU(cudaEventRecord(h.ev_done, d.stream));
                CU(cudaStreamWaitEvent(d.copy_out, h.ev_done, 0));
                void* h_dst = stage_out ? h.h_out : (void*)(io.dst_mem + row * out_stride);
                if (out_stride) CU(cudaMemcpyAsync(h_dst, h.d_out, cnt * out_stride, cudaMemcpyDeviceToHost, d.copy_out));
                CU(cudaEventRecord(h.ev_out, d.copy_out));
                d.stats.h2d_bytes += cnt * in_stride; d.stats.d2h_bytes += cnt * out_stride;
                return HERDB_OK;
            };
            rc = issue();
            if (rc) break;
        }
        if (c >= 1) {                                                                      // chunk c-1: wait for its rows, hand them to the sink
            LaneHalf& h = L->half[(c - 1) & 1];
            const size_t row = r0 + (c - 1) * chunk, cnt = std::min(chunk, rows - (c - 1) * chunk);
            if (cudaEventSynchronize(h.ev_out) != cudaSuccess) { ctx->set_error(std::string("evaluation failed: ") + cudaGetErrorString(cudaGetLastError())); rc = HERDB_ERR_CUDA; break; }
            if (out_stride && stage_out) {
                if (out_framed) {
                    if (files_rw(io, row, cnt, out_stride, h.h_out, true) != 0) { ctx->set_error("cannot write the output partition"); rc = HERDB_ERR_IO; break; }
                } else memcpy(io.dst_mem + row * out_stride, h.h_out, cnt * out_stride);
            }
            std::lock_guard<std::mutex> l(d.mu);
            drain_ready(d);
        }
    }
    return release(rc);
}

// n_rows rows through the stages, cut into contiguous per-device slices [k*R/G, (k+1)*R/G): output
// rows land at their input positions, so the result does not depend on the number of devices
int eval_rows_pipelined(herdb_ctx* ctx, const char* session, const EvalPlan& plan, RowIo io, size_t n_rows) {
    if (!n_rows) return HERDB_OK;
    io.src_pinned = is_pinned(io.src_mem); io.dst_pinned = is_pinned(io.dst_mem);
    const size_t G = ctx->devs.size();
    std::vector<int> rcs(G, HERDB_OK);
    auto run_slice = [&](size_t di) {
        const size_t r0 = n_rows * di / G, r1 = n_rows * (di + 1) / G;
        rcs[di] = device_pipeline(ctx, *ctx->devs[di], session, plan, io, r0, r1 - r0);
    };
    if (G == 1) run_slice(0);
    else {
        std::vector<std::thread> th;
        for (size_t di = 0; di < G; ++di) th.emplace_back(run_slice, di);
        for (auto& t : th) t.join();
    }
    for (size_t di = 0; di < G; ++di) {                  // timing intervals of this evaluation are in the stats when the call returns
        Device& d = *ctx->devs[di];
        std::lock_guard<std::mutex> l(d.mu);
        cudaSetDevice(d.id);
        drain_ready(d);
    }
    for (int rc : rcs) if (rc) return rc;
    return HERDB_OK;
}

// the session's key as it is resident (no file check: that is the task seam's business), pinned against eviction and
// unloading for the duration of the evaluation; pair with unpin_session_key
int session_n_pinned(herdb_ctx* ctx, const char* session, u32* n) {
    std::lock_guard<std::mutex> l(ctx->key_mu);
    auto it = ctx->key_info.find(session);
    if (it == ctx->key_info.end()) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    it->second.last_use = ++ctx->key_clock;
    it->second.pins++;
    *n = it->second.n;
    return HERDB_OK;
}

int eval_host_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* const* circuits, size_t n_stages,
                   const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    u32 n = 0;
    EvalPlan plan;
    int rc = session_n_pinned(ctx, session, &n);
    if (rc == HERDB_ERR_NOKEY) {                     // the circuit is still validated (malformed circuits are refused without a key)
        rc = make_plan(ctx, circuits, n_stages, 0, kLweMax, plan);
        return rc ? rc : HERDB_ERR_NOKEY;
    }
    rc = make_plan(ctx, circuits, n_stages, n_rows, n, plan);
    if (!rc && n_rows) {
        if ((!rows_in && plan.stages.front().in_bits) || (!rows_out && plan.stages.back().out_bits)) rc = HERDB_ERR_ARG;
        else {
            RowIo io; io.src_mem = (const uint8_t*)rows_in; io.dst_mem = (uint8_t*)rows_out;
            rc = eval_rows_pipelined(ctx, session, plan, io, n_rows);
        }
    }
    unpin_session_key(ctx, session);
    return rc;
}

}  // namespace

void herdb::OwnedCircuit::assign(const herdb_circuit* c) {
    nodes.assign(c->nodes, c->nodes + c->n_nodes);
    edges.assign(c->edges, c->edges + c->n_edges);
    in_offs.assign(c->in_tuple_offsets, c->in_tuple_offsets + c->n_tuples + 1);
    in_types.assign(c->in_field_types, c->in_field_types + (c->n_tuples ? c->in_tuple_offsets[c->n_tuples] : 0));
    out_types.assign(c->out_field_types, c->out_field_types + c->n_out_fields);
}
herdb_circuit herdb::OwnedCircuit::view() const {
    herdb_circuit c{};
    c.nodes = nodes.data(); c.n_nodes = nodes.size(); c.edges = edges.data(); c.n_edges = edges.size();
    c.in_field_types = in_types.data(); c.in_tuple_offsets = in_offs.data(); c.n_tuples = in_offs.empty() ? 0 : in_offs.size() - 1;
    c.out_field_types = out_types.data(); c.n_out_fields = out_types.size();
    return c;
}

// ================================================================================================ lifecycle
extern "C" const char* herdb_strerror(int code) {
    switch (code) {
    case HERDB_OK: return "ok";
    case HERDB_ERR_ARG: return "bad argument or malformed circuit";
    case HERDB_ERR_CUDA: return "CUDA error (or no CUDA device: this backend has no CPU fallback)";
    case HERDB_ERR_IO: return "key or partition file missing / unreadable";
    case HERDB_ERR_FORMAT: return "key blob or row payload does not match the parameter set";
    case HERDB_ERR_NOKEY: return "no key loaded for the session";
    case HERDB_ERR_NOMEM: return "out of device or host memory";
    case HERDB_ERR_UNSUPPORTED: return "unsupported";
    default: return "unknown error";
    }
}
extern "C" const char* herdb_last_error(const herdb_ctx* ctx) {
    if (!ctx) return "";
    std::lock_guard<std::mutex> l(ctx->err_mu);
    static thread_local std::string copy;
    copy = ctx->last_error;
    return copy.c_str();
}
extern "C" int herdb_device_count(const herdb_ctx* ctx) { return ctx ? (int)ctx->devs.size() : 0; }

static void worker_loop(herdb_ctx* ctx);

static size_t env_mb(const char* name, size_t dflt_mb) {
    const char* v = getenv(name);
    return (size_t)(v && atoll(v) > 0 ? atoll(v) : (long long)dflt_mb) << 20;
}

extern "C" int herdb_create(herdb_ctx** out, const int* device_ids, int n_devices, const char* storage_dir, const char* key_dir) {
    if (!out) return HERDB_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return HERDB_ERR_CUDA;     // no CPU fallback
    herdb_ctx* ctx = new herdb_ctx;
    ctx->storage_dir = storage_dir ? storage_dir : "";
    ctx->key_dir = key_dir ? key_dir : "";
    ctx->mem_budget = env_mb("HERDB_MEM_BUDGET_MB", 16384);
    ctx->io_budget = env_mb("HERDB_IO_CHUNK_MB", 1024);
    ctx->key_budget = env_mb("HERDB_KEY_BUDGET_MB", 32768);
    {
        const char* cr = getenv("HERDB_COALESCE_ROWS");          // 0 switches task coalescing off
        ctx->coalesce_rows = cr ? (uint64_t)std::max<long long>(0, atoll(cr)) : 4096;
        const char* tw = getenv("HERDB_TASK_WINDOW");
        ctx->task_window = std::max(1, std::min(64, tw ? atoi(tw) : 8));
    }
    std::vector<int> ids;
    if (device_ids && n_devices > 0) ids.assign(device_ids, device_ids + n_devices); else ids.push_back(0);
    std::vector<u32> mono, ftw2(kLaneTwWords), itw1(kLaneTwWords);
    build_monomial_table(mono);
    build_lane_twiddles(ftw2.data(), itw1.data());
    auto fail = [&](int rc) { *out = nullptr; herdb_destroy(ctx); return rc; };
    for (size_t k = 0; k < ids.size(); ++k)
        for (size_t j = 0; j < k; ++j) if (ids[j] == ids[k]) return fail(HERDB_ERR_ARG);   // a device may appear once
    for (int id : ids) {
        if (id < 0 || id >= count) { return fail(HERDB_ERR_ARG); }
        Device* d = new Device; d->id = id; ctx->devs.push_back(d);
        if (cudaSetDevice(id) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, id);
        for (cudaStream_t* s : {&d->stream, &d->copy, &d->copy_out})
            if (cudaStreamCreateWithFlags(s, cudaStreamNonBlocking) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        for (auto& e : d->ev_piece)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        if (cudaMalloc((void**)&d->mono, mono.size() * 4) != cudaSuccess || cudaMalloc((void**)&d->ftw2, kLaneTwWords * 4) != cudaSuccess ||
            cudaMalloc((void**)&d->itw1, kLaneTwWords * 4) != cudaSuccess) return fail(HERDB_ERR_NOMEM);
        cudaMemcpy(d->mono, mono.data(), mono.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->ftw2, ftw2.data(), kLaneTwWords * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->itw1, itw1.data(), kLaneTwWords * 4, cudaMemcpyHostToDevice);
    }
    // Two task threads: while one task's kernels run, the next task's key check, partition read and first upload
    // happen; the adapter reports the same number as concurrent_workers() (executor.cpp:96-113 keeps that many in flight).
    const char* tw = getenv("HERDB_TASK_THREADS");
    const int n_workers = std::max(1, std::min(8, tw ? atoi(tw) : 2));
    for (int k = 0; k < n_workers; ++k) ctx->workers.emplace_back(worker_loop, ctx);
    *out = ctx;
    return HERDB_OK;
}

extern "C" int herdb_task_threads(const herdb_ctx* ctx) { return ctx ? (int)ctx->workers.size() : 0; }
extern "C" int herdb_task_window(const herdb_ctx* ctx) { return ctx ? std::max(ctx->task_window, (int)ctx->workers.size()) : 0; }

extern "C" void herdb_destroy(herdb_ctx* ctx) {
    if (!ctx) return;
    {
        std::unique_lock<std::mutex> l(ctx->q_mu);
        ctx->idle_cv.wait(l, [&] { return ctx->queue.empty() && ctx->in_flight == 0; });
        ctx->stop = true;
    }
    ctx->q_cv.notify_all();
    for (auto& t : ctx->workers) if (t.joinable()) t.join();
    for (ncclComm* c : ctx->nccl) ncclCommDestroy(c);
    for (Device* d : ctx->devs) {
        cudaSetDevice(d->id);
        if (d->stream) cudaStreamSynchronize(d->stream);
        if (d->copy) cudaStreamSynchronize(d->copy);
        if (d->copy_out) cudaStreamSynchronize(d->copy_out);
        for (auto& kv : d->keys) { cudaFree(kv.second.bsk); cudaFree(kv.second.ksk); }
        for (auto& t : d->timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
        for (auto& e : d->free_events) cudaEventDestroy(e);
        for (Lane* L : d->lanes) free_lane(L);
        cudaFree(d->mono); cudaFree(d->ftw2); cudaFree(d->itw1); cudaFree(d->slots); cudaFree(d->ext);
        cudaFree(d->gates); cudaFree(d->outs); cudaFree(d->io_in); cudaFree(d->io_out);
        for (cudaStream_t s : {d->stream, d->copy, d->copy_out}) if (s) cudaStreamDestroy(s);
        for (auto& e : d->ev_piece) if (e) cudaEventDestroy(e);
        delete d;
    }
    delete ctx;
}

// ================================================================================================ keys
extern "C" size_t herdb_key_blob_size(uint32_t n) {
    const size_t polys = (size_t)n * 2 * kDG2 * 2, ks_rows = (size_t)kN * kBaseKS * kDKS;
    return sizeof(herdb_key_header) + polys * kN * 4 + ks_rows * (n + 1) * 2;
}

extern "C" int herdb_load_key_dev(herdb_ctx* ctx, const char* session, int device, const void* d_blob, size_t len) {
    if (!ctx || !session || !d_blob || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> kl(ctx->key_mu);
    Device& d = *ctx->devs[device];
    herdb_key_header h;
    {
        std::lock_guard<std::mutex> l(d.mu);
        CU(cudaSetDevice(d.id));
        if (len < sizeof h) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
        CU(cudaMemcpy(&h, d_blob, sizeof h, cudaMemcpyDeviceToHost));
        int rc = key_check(ctx, h, len); if (rc) return rc;
        DeviceKey k;
        rc = install_key_issue(ctx, d, h, static_cast<const uint8_t*>(d_blob), k);
        if (!rc && cudaStreamSynchronize(d.stream) != cudaSuccess) { ctx->set_error("key conversion failed"); rc = HERDB_ERR_CUDA; }
        if (rc) { cudaFree(k.bsk); cudaFree(k.ksk); return rc; }
        install_key_commit(d, session, k);
    }
    KeyInfo meta; meta.n = h.n; meta.dev_bytes = key_dev_bytes(h.n); meta.last_use = ++ctx->key_clock;
    auto old = ctx->key_info.find(session);
    if (old != ctx->key_info.end()) meta.pins = old->second.pins;
    ctx->key_info[session] = meta;
    return HERDB_OK;
}

extern "C" int herdb_load_key_mem(herdb_ctx* ctx, const char* session, const void* blob, size_t len) {
    if (!ctx || !session || !blob) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> kl(ctx->key_mu);
    return load_key_blob(ctx, session, blob, len, KeyInfo());
}

extern "C" int herdb_load_key(herdb_ctx* ctx, const char* session, int schema_type) {
    if (!ctx || !session) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> kl(ctx->key_mu);
    return load_key_file(ctx, session, schema_type);
}

extern "C" int herdb_unload_key(herdb_ctx* ctx, const char* session) {
    if (!ctx || !session) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> kl(ctx->key_mu);
    auto it = ctx->key_info.find(session);
    if (it == ctx->key_info.end()) return HERDB_ERR_NOKEY;
    if (it->second.pins > 0) { ctx->set_error("key is in use by a running task"); return HERDB_ERR_ARG; }
    drop_device_keys(ctx, session);
    ctx->key_info.erase(it);
    return HERDB_OK;
}

extern "C" int herdb_key_lwe_n(herdb_ctx* ctx, const char* session, uint32_t* n_out) {
    if (!ctx || !session || !n_out || ctx->devs.empty()) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> l(ctx->key_mu);
    auto it = ctx->key_info.find(session);
    if (it == ctx->key_info.end()) return HERDB_ERR_NOKEY;
    *n_out = it->second.n;
    return HERDB_OK;
}

extern "C" int herdb_loaded_keys(herdb_ctx* ctx) {
    if (!ctx) return 0;
    std::lock_guard<std::mutex> l(ctx->key_mu);
    return (int)ctx->key_info.size();
}

// ================================================================================================ file-less entry points
// single-gate batches over device buffers: descriptor upload + scratch once, then launches over row ranges
static int gate_batch_prepare(herdb_ctx* ctx, Device& d, int gate, size_t batch, size_t* chunk_out) {
    int rc = grow(ctx, (void**)&d.gates, &d.gates_cap, sizeof(GateDesc)); if (rc) return rc;
    GateDesc g{}; g.gate = (uint8_t)gate; g.pool_a = 0; g.pool_b = 2; g.pool_out = 3;
    CU(cudaMemcpyAsync(d.gates, &g, sizeof g, cudaMemcpyHostToDevice, d.stream));
    CU(cudaStreamSynchronize(d.stream));      // `g` lives on this stack frame
    const size_t per_job = (size_t)kExtStride * 2;
    const size_t chunk = std::min<size_t>(batch, std::max<size_t>(1, std::min<size_t>(ctx->mem_budget / per_job, 0x7FFFFFFF)));
    *chunk_out = chunk;
    return grow(ctx, (void**)&d.ext, &d.ext_bytes, chunk * per_job);
}
static int gate_batch_launch(herdb_ctx* ctx, Device& d, const DeviceKey& key, const u16* d_a, const u16* d_b, u16* d_out,
                             size_t first, size_t count, size_t chunk, int shape = 0) {
    const size_t w = key.n + 1;
    for (size_t r0 = first; r0 < first + count; r0 += chunk) {
        const u32 rows = (u32)std::min(chunk, first + count - r0);
        Pool pools[4] = {};
        pools[0] = Pool{const_cast<u16*>(d_a) + r0 * w, (u32)w, 0};
        pools[2] = Pool{const_cast<u16*>(d_b) + r0 * w, (u32)w, 0};
        pools[3] = Pool{d_out + r0 * w, (u32)w, 0};
        int rc = run_level(ctx, d, Exec{d.stream, d.slots, d.ext}, key, d.gates, 1, pools, rows, nullptr, UINT32_MAX, shape); if (rc) return rc;
    }
    return HERDB_OK;
}

static int gate_batch_dev(herdb_ctx* ctx, const char* session, int device, int gate, const uint16_t* d_a, const uint16_t* d_b,
                          uint16_t* d_out, size_t batch, int shape) {
    if (!ctx || !session || device < 0 || device >= (int)ctx->devs.size() || gate < 0 || gate > HERDB_GATE_XNOR_FAST) return HERDB_ERR_ARG;
    if (!batch) return HERDB_OK;
    if (!d_a || !d_b || !d_out) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const DeviceKey* key = find_key(d, session);
    if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    size_t chunk = 0;
    int rc = gate_batch_prepare(ctx, d, gate, batch, &chunk); if (rc) return rc;
    rc = gate_batch_launch(ctx, d, *key, d_a, d_b, d_out, 0, batch, chunk, shape); if (rc) return rc;
    return drain(ctx, d);
}
extern "C" int herdb_eval_gate_batch_dev(herdb_ctx* ctx, const char* session, int device, int gate,
                                         const uint16_t* d_a, const uint16_t* d_b, uint16_t* d_out, size_t batch) {
    return gate_batch_dev(ctx, session, device, gate, d_a, d_b, d_out, batch, 0);
}
extern "C" int herdb_eval_gate_batch_dev_shape(herdb_ctx* ctx, const char* session, int device, int gate, const uint16_t* d_a,
                                               const uint16_t* d_b, uint16_t* d_out, size_t batch, int cts_per_cta) {
    if (cts_per_cta != 0 && cts_per_cta != 1 && cts_per_cta != 2 && cts_per_cta != 4) return HERDB_ERR_ARG;
    return gate_batch_dev(ctx, session, device, gate, d_a, d_b, d_out, batch, cts_per_cta);
}

// The same single-gate batch with NO allocation anywhere on the call path (no cudaMalloc / cudaFree / event creation):
// the extraction buffer and the gate descriptor are the library's static device arrays (kernels.cu), whose device
// addresses are looked up once per device.  No counterpart in the reference; added
// for callers that cannot allocate (CUDA graph capture, tight serving loops).  Kernel timing events are not recorded on
// this path (creating them allocates); launches and gates are still counted.
extern "C" int herdb_eval_gate_batch_dev_noalloc(herdb_ctx* ctx, const char* session, int device, int gate,
                                                 const uint16_t* d_a, const uint16_t* d_b, uint16_t* d_out, size_t batch) {
    if (!ctx || !session || device < 0 || device >= (int)ctx->devs.size() || gate < 0 || gate > HERDB_GATE_XNOR_FAST) return HERDB_ERR_ARG;
    if (batch > HERDB_NOALLOC_MAX_BATCH) {
        ctx->set_error("herdb_eval_gate_batch_dev_noalloc: batch of " + std::to_string(batch) + " gates exceeds HERDB_NOALLOC_MAX_BATCH = " +
                       std::to_string(HERDB_NOALLOC_MAX_BATCH) + " (the scratch is a fixed array; use herdb_eval_gate_batch_dev, or split the batch)");
        return HERDB_ERR_ARG;
    }
    if (!batch) return HERDB_OK;
    if (!d_a || !d_b || !d_out) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    if (!d.noalloc_ready) {                          // set-up, once per device, with that device current
        CU(noalloc_setup(&d.noalloc_ext, &d.noalloc_gate));
        d.noalloc_ready = true;
    }
    const DeviceKey* key = find_key(d, session);
    if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    const u32 w = key->n + 1, jobs = (u32)batch;
    GateDesc g{}; g.gate = (uint8_t)gate; g.pool_a = 0; g.pool_b = 2; g.pool_out = 3;
    CU(launch_noalloc_set_gate(g, d.stream));        // by value through the kernel parameters: no staging copy
    Pool pools[4] = {};
    pools[0] = Pool{const_cast<u16*>(d_a), w, 0};
    pools[2] = Pool{const_cast<u16*>(d_b), w, 0};
    pools[3] = Pool{d_out, w, 0};
    BrParams bp{};
    bp.bsk = key->bsk; bp.mono = d.mono; bp.ftw2 = d.ftw2; bp.itw1 = d.itw1; bp.gates = d.noalloc_gate;
    for (int i = 0; i < 4; ++i) bp.pools[i] = pools[i];
    bp.ext = d.noalloc_ext; bp.trace = nullptr; bp.n = key->n; bp.steps = key->n; bp.rows = jobs; bp.jobs = jobs; bp.job0 = 0; bp.job1 = jobs;
    int launches = 0;
    CU(launch_blind_rotate(bp, d.sm_count, d.stream, 0, &launches));
    KsParams kp{};
    kp.ksk = key->ksk; kp.ext = d.noalloc_ext; kp.gates = d.noalloc_gate;
    for (int i = 0; i < 4; ++i) kp.pools[i] = pools[i];
    kp.ks_stride = key->ks_stride; kp.n = key->n; kp.rows = jobs; kp.jobs = jobs;
    CU(launch_keyswitch(kp, d.sm_count, d.stream));
    d.stats.kernel_launches += (uint64_t)launches + 2;
    d.stats.gates_bootstrapped += jobs;
    CU(cudaStreamSynchronize(d.stream));
    return HERDB_OK;
}

// Host buffers.  A large batch is cut into a head of one grid round, a middle and a tail of one round (+ the
// remainder): the middle's upload rides the copy stream under the head's kernels and the middle's download
// under the tail's, so only one round's worth of PCIe traffic is exposed.  Pieces are multiples of the round
// size (SM count x ciphertexts per CTA), so the cut adds no part-empty round.
extern "C" int herdb_eval_gate_batch(herdb_ctx* ctx, const char* session, int device, int gate,
                                     const uint16_t* a, const uint16_t* b, uint16_t* out, size_t batch) {
    if (!ctx || !session || device < 0 || device >= (int)ctx->devs.size() || gate < 0 || gate > HERDB_GATE_XNOR_FAST) return HERDB_ERR_ARG;
    if (!batch) return HERDB_OK;
    if (!a || !b || !out) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const DeviceKey* key = find_key(d, session);
    if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    const size_t w = key->n + 1, bytes = batch * w * 2;
    int rc = grow(ctx, (void**)&d.io_in, &d.io_in_bytes, 2 * bytes); if (rc) return rc;
    rc = grow(ctx, (void**)&d.io_out, &d.io_out_bytes, bytes); if (rc) return rc;
    size_t chunk = 0;
    rc = gate_batch_prepare(ctx, d, gate, batch, &chunk); if (rc) return rc;
    u16 *da = d.io_in, *db = d.io_in + batch * w, *dout = d.io_out;
    const size_t wave = (size_t)d.sm_count * kBrCtsPerCta;
    size_t cut[4] = {0, batch, batch, batch};                      // piece p = [cut[p], cut[p+1])
    if (batch >= 4 * wave) { cut[1] = wave; cut[2] = batch - wave - batch % wave; }
    auto h2d = [&](int p, cudaStream_t st) -> cudaError_t {
        const size_t off = cut[p] * w, cnt = (cut[p + 1] - cut[p]) * w * 2;
        if (!cnt) return cudaSuccess;
        cudaError_t e = cudaMemcpyAsync(da + off, a + off, cnt, cudaMemcpyHostToDevice, st);
        return e != cudaSuccess ? e : cudaMemcpyAsync(db + off, b + off, cnt, cudaMemcpyHostToDevice, st);
    };
    auto d2h = [&](int p, cudaStream_t st) -> cudaError_t {
        const size_t off = cut[p] * w, cnt = (cut[p + 1] - cut[p]) * w * 2;
        return cnt ? cudaMemcpyAsync(out + off, dout + off, cnt, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    CU(h2d(0, d.stream));
    rc = gate_batch_launch(ctx, d, *key, da, db, dout, cut[0], cut[1] - cut[0], chunk); if (rc) return rc;
    if (cut[1] < batch) {
        CU(cudaEventRecord(d.ev_piece[0], d.stream));               // head computed
        CU(h2d(1, d.copy)); CU(h2d(2, d.copy));
        CU(cudaEventRecord(d.ev_piece[1], d.copy));                 // middle and tail uploaded
        CU(cudaStreamWaitEvent(d.stream, d.ev_piece[1], 0));
        rc = gate_batch_launch(ctx, d, *key, da, db, dout, cut[1], cut[2] - cut[1], chunk); if (rc) return rc;
        CU(cudaEventRecord(d.ev_piece[2], d.stream));               // middle computed
        rc = gate_batch_launch(ctx, d, *key, da, db, dout, cut[2], cut[3] - cut[2], chunk); if (rc) return rc;
        CU(cudaStreamWaitEvent(d.copy_out, d.ev_piece[0], 0));
        CU(d2h(0, d.copy_out));
        CU(cudaStreamWaitEvent(d.copy_out, d.ev_piece[2], 0));
        CU(d2h(1, d.copy_out));
        CU(d2h(2, d.stream));
        CU(cudaStreamSynchronize(d.copy_out));
    } else {
        CU(d2h(0, d.stream));
    }
    d.stats.h2d_bytes += 2 * bytes; d.stats.d2h_bytes += bytes;
    return drain(ctx, d);
}

extern "C" int herdb_eval_circuit_rows_dev(herdb_ctx* ctx, const char* session, int device, const herdb_circuit* circuit,
                                           const uint16_t* d_rows_in, uint16_t* d_rows_out, size_t n_rows) {
    if (!ctx || !session || !circuit || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    SchedHint hint; hint.rows = n_rows; hint.wave = (u32)ctx->devs[device]->sm_count * kBrCtsPerCta;
    int rc = compile_circuit(circuit, cc, err, hint);
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    if (!n_rows) return HERDB_OK;
    if ((!d_rows_in && cc.in_bits) || (!d_rows_out && cc.out_bits)) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const DeviceKey* key = find_key(d, session);
    if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    const u32 w = key->n + 1;
    std::vector<size_t> goff, ooff;
    rc = upload_tables(ctx, d.stream, {&cc}, &d.gates, &d.gates_cap, &d.outs, &d.outs_cap, goff, ooff); if (rc) return rc;
    // rows are processed in chunks sized so that the slot file + extraction buffer fit the budget
    size_t chunk = std::max<size_t>(1, ctx->mem_budget / std::max<size_t>(1, scratch_per_row(cc, key->n)));
    if (cc.max_level_width) chunk = std::min<size_t>(chunk, (size_t)0x7FFFFFFF / cc.max_level_width);
    chunk = std::min(chunk, n_rows);
    rc = ensure_scratch(ctx, &d.slots, &d.slots_bytes, &d.ext, &d.ext_bytes, cc, key->n, chunk); if (rc) return rc;
    for (size_t r0 = 0; r0 < n_rows; r0 += chunk) {
        const u32 rows = (u32)std::min(chunk, n_rows - r0);
        rc = issue_circuit(ctx, d, Exec{d.stream, d.slots, d.ext}, *key, cc, d.gates, d.outs, Pool{const_cast<u16*>(d_rows_in) + r0 * cc.in_bits * w, cc.in_bits * w, 0},
                           Pool{d_rows_out + r0 * cc.out_bits * w, cc.out_bits * w, 0}, false, rows);
        if (rc) return rc;
    }
    return drain(ctx, d);
}

extern "C" int herdb_eval_pipeline_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* const* circuits, size_t n_stages,
                                        const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    if (!ctx || !session || !circuits || !n_stages) return HERDB_ERR_ARG;
    return eval_host_rows(ctx, session, circuits, n_stages, rows_in, rows_out, n_rows);
}

extern "C" int herdb_eval_circuit_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* circuit,
                                       const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    if (!ctx || !session || !circuit) return HERDB_ERR_ARG;
    return eval_host_rows(ctx, session, &circuit, 1, rows_in, rows_out, n_rows);
}

extern "C" int herdb_circuit_analyze(const herdb_circuit* circuit, herdb_circuit_info* info) {
    if (!circuit || !info) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(circuit, cc, err);
    if (rc) return rc;
    info->in_bits = cc.in_bits; info->out_bits = cc.out_bits; info->boot_gates = cc.boot_gates;
    info->levels = (u32)cc.levels.size(); info->max_level_width = cc.max_level_width; info->slots = cc.slots;
    return HERDB_OK;
}

extern "C" int herdb_circuit_schedule(const herdb_circuit* circuit, uint64_t rows, uint32_t wave_jobs,
                                      herdb_schedule_info* info, uint32_t* level_widths, size_t cap) {
    if (!circuit || !info) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    SchedHint hint; hint.rows = rows; hint.wave = wave_jobs;
    int rc = compile_circuit(circuit, cc, err, hint);
    if (rc) return rc;
    info->levels = (u32)cc.levels.size(); info->max_level_width = cc.max_level_width; info->slots = cc.slots;
    info->policy = cc.policy; info->rounds_asap = cc.rounds_asap; info->rounds = cc.rounds;
    if (level_widths) for (size_t l = 0; l < cc.levels.size() && l < cap; ++l) level_widths[l] = (u32)cc.levels[l].size();
    return HERDB_OK;
}

extern "C" int herdb_circuit_eval_plain(const herdb_circuit* circuit, uint64_t rows, uint32_t wave_jobs,
                                        const uint8_t* in_bits, uint8_t* out_bits) {
    if (!circuit) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    SchedHint hint; hint.rows = rows; hint.wave = wave_jobs;
    int rc = compile_circuit(circuit, cc, err, hint);
    if (rc) return rc;
    if ((!in_bits && cc.in_bits) || (!out_bits && cc.out_bits)) return HERDB_ERR_ARG;
    return eval_compiled_plain(cc, in_bits, out_bits) ? HERDB_ERR_ARG : HERDB_OK;
}

// ================================================================================================ measurement hooks
extern "C" int herdb_get_stats(herdb_ctx* ctx, int device, herdb_stats* out) {
    if (!ctx || !out || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> l(ctx->devs[device]->mu);
    cudaSetDevice(ctx->devs[device]->id);
    drain_ready(*ctx->devs[device]);
    *out = ctx->devs[device]->stats;
    return HERDB_OK;
}
extern "C" int herdb_get_stream(herdb_ctx* ctx, int device, void** stream_out) {
    if (!ctx || !stream_out || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    *stream_out = (void*)ctx->devs[device]->stream;
    return HERDB_OK;
}
extern "C" int herdb_reset_stats(herdb_ctx* ctx) {
    if (!ctx) return HERDB_ERR_ARG;
    for (Device* d : ctx->devs) {
        std::lock_guard<std::mutex> l(d->mu);
        const double kl = d->stats.key_load_ms, kb = d->stats.key_bcast_ms;
        d->stats = herdb_stats{}; d->stats.key_load_ms = kl; d->stats.key_bcast_ms = kb;
    }
    return HERDB_OK;
}
extern "C" int herdb_measure_int_peak(herdb_ctx* ctx, int device, double* imad, double* iadd3, double* imadhi,
                                      double* imadwide, double* mixed) {
    if (!ctx || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const int blocks = d.sm_count * 8, iters = 2048;
    u32* buf = nullptr;
    CU(cudaMalloc((void**)&buf, (size_t)blocks * 256 * 4));
    cudaEvent_t a, b; CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
    double* outs[5] = {imad, iadd3, imadhi, imadwide, mixed};
    for (int mode = 0; mode < 5; ++mode) {
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            CU(cudaEventRecord(a, d.stream));
            CU(launch_int_peak(mode, buf, blocks, iters, d.stream));
            CU(cudaEventRecord(b, d.stream));
            CU(cudaStreamSynchronize(d.stream));
            float ms; CU(cudaEventElapsedTime(&ms, a, b));
            if (rep) best = std::min(best, ms);
        }
        // ops = threads * iters * 64 inner statements; the mixed mode counts its 3 multiplies each
        const double ops = (double)blocks * 256 * iters * 64 * (mode == 4 ? 3.0 : 1.0);
        if (outs[mode]) *outs[mode] = ops / (best * 1e-3) / 1e12;
        d.stats.kernel_launches += 4;
    }
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(buf);
    return HERDB_OK;
}

extern "C" int herdb_measure_l2_bandwidth(herdb_ctx* ctx, int device, double* gbs) {
    if (!ctx || !gbs || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const size_t bytes = (size_t)32 << 20;                           // well inside the 126 MB L2
    const int reps = 64;
    void* buf = nullptr; u32* out = nullptr;
    CU(cudaMalloc(&buf, bytes)); CU(cudaMalloc((void**)&out, 4));
    CU(cudaMemsetAsync(buf, 1, bytes, d.stream));
    cudaEvent_t a, b; CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CU(cudaEventRecord(a, d.stream));
        CU(launch_l2_read(buf, bytes, reps, out, d.sm_count, d.stream));
        CU(cudaEventRecord(b, d.stream));
        CU(cudaStreamSynchronize(d.stream));
        float ms; CU(cudaEventElapsedTime(&ms, a, b));
        if (rep) best = std::min(best, ms);
        d.stats.kernel_launches += 1;
    }
    *gbs = (double)bytes * reps / (best * 1e-3) / 1e9;
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(buf); cudaFree(out);
    return HERDB_OK;
}

// ================================================================================================ the task seam
// Partition files (storage_service.cpp:19-28,99-150): rows concatenated, each row
// [u32 length, host endian][payload].  Our payload = the row's bit-ciphertexts, dense
// (bits * (n+1) u16).  Output is written to "<partition>.tmp" and renamed so that a partial
// file is never observed (the Lambda worker group treats existence as success,
// lambda_http_worker_group.cpp:244-259).
namespace {

std::string partition_path(const herdb_ctx* ctx, const std::string& session, const std::string& df, u32 part) {
    return ctx->storage_dir + "/" + session + "/" + df + "/" + std::to_string(part);
}

struct OutFile {                                     // "<path>.tmp" until commit()
    std::string path, tmp; int fd = -1;
    int open_tmp(herdb_ctx* ctx, const std::string& p) {
        path = p; tmp = p + ".tmp";
        fd = open(tmp.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
        if (fd < 0) { ctx->set_error("cannot create " + tmp); return HERDB_ERR_IO; }
        return HERDB_OK;
    }
    int commit(herdb_ctx* ctx) {
        const bool ok = close(fd) == 0; fd = -1;
        if (!ok || rename(tmp.c_str(), path.c_str()) != 0) { remove(tmp.c_str()); ctx->set_error("cannot write partition " + path); return HERDB_ERR_IO; }
        return HERDB_OK;
    }
    ~OutFile() { if (fd >= 0) { close(fd); remove(tmp.c_str()); } }
};

// MapTask(s): the input partition is streamed through the chunk pipeline, never held in memory as a whole.  `ts` is one
// task, or several queued tasks of one session with the same circuit that the worker coalesced (worker_loop): their
// rows go through the levels as ONE batch -- a 128-row partition alone fills a fraction of a grid round per level --
// while every task still reads its own partition file and writes (and commits) its own.
int run_map_tasks(herdb_ctx* ctx, const std::vector<const TaskReq*>& ts, u32 n) {
    NvtxRange range("herdb:map_task");
    const TaskReq& t0 = *ts[0];
    const herdb_circuit view = t0.circuit.view();
    const herdb_circuit* cp = &view;
    uint64_t total = 0;
    for (const TaskReq* t : ts) total += t->row_count;
    EvalPlan plan;
    int rc = make_plan(ctx, &cp, 1, total, n, plan); if (rc) return rc;
    const size_t w = n + 1, in_stride = (size_t)plan.stages[0].in_bits * w * 2 + 4;
    RowIo io;
    std::deque<OutFile> outs;
    auto close_inputs = [&] { for (const FileSeg& sg : io.files) close(sg.src_fd); };
    for (const TaskReq* t : ts) {
        const std::string in_path = partition_path(ctx, t->session, t->in_df, t->in_part);
        const int fd = open(in_path.c_str(), O_RDONLY);
        if (fd < 0) { close_inputs(); ctx->set_error("cannot open partition " + in_path); return HERDB_ERR_IO; }
        struct stat st{};
        if (fstat(fd, &st) != 0 || (uint64_t)st.st_size != t->row_count * in_stride) {       // row_count rows of the circuit's input row, exactly
            close(fd); close_inputs();
            ctx->set_error("partition " + in_path + ": row count / row length does not match the task (row_count, circuit inputs)");
            return HERDB_ERR_FORMAT;
        }
        outs.emplace_back();
        rc = outs.back().open_tmp(ctx, partition_path(ctx, t->session, t->out_df, t->out_part));
        if (rc) { close(fd); close_inputs(); return rc; }
        io.files.push_back(FileSeg{fd, outs.back().fd, (size_t)t->row_count});
    }
    rc = eval_rows_pipelined(ctx, t0.session.c_str(), plan, io, total);
    close_inputs();
    for (OutFile& o : outs) { if (rc) break; rc = o.commit(ctx); }
    return rc;
}

// a reduce circuit takes two rows of the output layout: inputs[0] == inputs[1] == output column types
int check_reduce_circuit(herdb_ctx* ctx, const herdb_circuit* c) {
    bool ok = c->n_tuples == 2;
    if (ok) {
        const u32 f0 = c->in_tuple_offsets[1] - c->in_tuple_offsets[0], f1 = c->in_tuple_offsets[2] - c->in_tuple_offsets[1];
        ok = f0 == f1 && f0 == c->n_out_fields;
        for (u32 f = 0; ok && f < f0; ++f)
            ok = c->in_field_types[c->in_tuple_offsets[0] + f] == c->out_field_types[f] &&
                 c->in_field_types[c->in_tuple_offsets[1] + f] == c->out_field_types[f];
    }
    if (!ok) ctx->set_error("reduce circuit: needs two input tuples with the column types of the output");
    return ok ? HERDB_OK : HERDB_ERR_ARG;
}

// balanced pairwise fold of `m` rows (out-row layout, contiguous) to one row; every round is one batched
// evaluation: rows 2i and 2i+1 are adjacent in memory, i.e. already the concatenated input row of pair i
int reduce_rows_host(herdb_ctx* ctx, const char* session, const herdb_circuit* circuit, u32 out_bits, std::vector<uint16_t>& cur, size_t m, size_t w) {
    const size_t row = (size_t)out_bits * w;
    std::vector<uint16_t> next;
    while (m > 1) {
        const size_t pairs = m / 2, odd = m & 1;
        next.resize((pairs + odd) * row);
        int rc = eval_host_rows(ctx, session, &circuit, 1, cur.data(), next.data(), pairs);
        if (rc) return rc;
        if (odd) memcpy(next.data() + pairs * row, cur.data() + (m - 1) * row, row * 2);
        cur.swap(next);
        m = pairs + odd;
    }
    cur.resize(row);
    return HERDB_OK;
}

// ReduceTask.  InputDataFramePtr.row_count is a HINT here: the reference's reduce trees label the hidden frames
// of Policy::PARALLEL with row_count = partitions although each of those partitions holds the single row a
// first-level reduce wrote (execution_service.cpp:613-618).  Every input partition is therefore read to its end,
// validating only the per-row length.
int run_reduce_task(herdb_ctx* ctx, const TaskReq& t, u32 n) {
    NvtxRange range("herdb:reduce_task");
    const herdb_circuit view = t.circuit.view();
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(&view, cc, err);
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    const size_t w = n + 1, row_bytes = (size_t)cc.out_bits * w * 2;
    std::vector<uint16_t> rows;
    size_t m = 0;
    for (const InFrame& f : t.inputs) {
        const std::string path = partition_path(ctx, t.session, f.df, f.part);
        uint64_t have = 0;
        rc = herdb_partition_count(path.c_str(), row_bytes, &have);
        if (rc) { ctx->set_error(rc == HERDB_ERR_IO ? "cannot open partition " + path : "partition " + path + ": row length does not match the reduce circuit's columns"); return rc; }
        rows.resize((m + have) * row_bytes / 2);
        rc = herdb_partition_read(path.c_str(), have, row_bytes, rows.data() + m * row_bytes / 2);
        if (rc) { ctx->set_error("cannot read partition " + path); return rc; }
        m += have;
    }
    if (!m) { ctx->set_error("reduce task without input rows"); return HERDB_ERR_ARG; }
    rc = reduce_rows_host(ctx, t.session.c_str(), &view, cc.out_bits, rows, m, w);
    if (rc) return rc;
    const std::string out_path = partition_path(ctx, t.session, t.out_df, t.out_part);
    rc = herdb_partition_write(out_path.c_str(), 1, row_bytes, rows.data());
    if (rc) ctx->set_error("cannot write partition " + out_path);
    return rc;
}

int run_tasks(herdb_ctx* ctx, const std::vector<const TaskReq*>& ts) {      // one task, or a coalesced group of map tasks
    const TaskReq& t = *ts[0];
    u32 n = 0;
    int rc = pin_session_key(ctx, t.session, t.schema_type, &n);
    if (rc) return rc;
    rc = t.reduce ? run_reduce_task(ctx, t, n) : run_map_tasks(ctx, ts, n);
    unpin_session_key(ctx, t.session);
    return rc;
}

bool same_circuit(const OwnedCircuit& a, const OwnedCircuit& b) {
    return a.nodes.size() == b.nodes.size() && a.edges.size() == b.edges.size() && a.in_types == b.in_types && a.out_types == b.out_types &&
           a.in_offs == b.in_offs && (a.nodes.empty() || !memcmp(a.nodes.data(), b.nodes.data(), a.nodes.size() * sizeof(herdb_node))) &&
           (a.edges.empty() || !memcmp(a.edges.data(), b.edges.data(), a.edges.size() * sizeof(herdb_edge)));
}

}  // namespace

extern "C" int herdb_reduce_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* circuit,
                                 const uint16_t* rows_in, size_t n_rows, uint16_t* row_out) {
    if (!ctx || !session || !circuit || !rows_in || !row_out || !n_rows) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(circuit, cc, err);
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    rc = check_reduce_circuit(ctx, circuit); if (rc) return rc;
    uint32_t n = 0;
    rc = herdb_key_lwe_n(ctx, session, &n); if (rc) return rc;
    const size_t w = n + 1, row = (size_t)cc.out_bits * w;
    std::vector<uint16_t> cur(rows_in, rows_in + n_rows * row);
    rc = reduce_rows_host(ctx, session, circuit, cc.out_bits, cur, n_rows, w); if (rc) return rc;
    memcpy(row_out, cur.data(), row * 2);
    return HERDB_OK;
}

extern "C" int herdb_partition_read(const char* path, uint64_t row_count, size_t row_bytes, void* rows_out) {
    if (!path || (!rows_out && row_count && row_bytes)) return HERDB_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return HERDB_ERR_IO;
    int rc = HERDB_OK;
    for (uint64_t r = 0; r < row_count && !rc; ++r) {
        uint32_t len = 0;
        if (fread(&len, sizeof len, 1, f) != 1) rc = HERDB_ERR_FORMAT;                    // fewer rows than row_count
        else if (len != row_bytes) rc = HERDB_ERR_FORMAT;                                 // ragged / foreign payload
        else if (row_bytes && fread(static_cast<uint8_t*>(rows_out) + r * row_bytes, row_bytes, 1, f) != 1) rc = HERDB_ERR_FORMAT;
    }
    if (!rc && fgetc(f) != EOF) rc = HERDB_ERR_FORMAT;                                    // more rows than row_count
    fclose(f);
    return rc;
}

extern "C" int herdb_partition_count(const char* path, size_t row_bytes, uint64_t* rows_out) {
    if (!path || !rows_out) return HERDB_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return HERDB_ERR_IO;
    uint64_t rows = 0; int rc = HERDB_OK;
    for (;;) {
        uint32_t len = 0;
        const size_t got = fread(&len, 1, sizeof len, f);
        if (got == 0) break;                                                              // clean end of file
        if (got != sizeof len || len != row_bytes || (row_bytes && fseek(f, (long)row_bytes, SEEK_CUR) != 0)) { rc = HERDB_ERR_FORMAT; break; }
        ++rows;
    }
    if (!rc) {                                                                            // the last payload must be complete
        struct stat st{};
        if (fstat(fileno(f), &st) != 0 || (uint64_t)st.st_size != rows * (row_bytes + 4)) rc = HERDB_ERR_FORMAT;
    }
    fclose(f);
    *rows_out = rc ? 0 : rows;
    return rc;
}

extern "C" int herdb_partition_write(const char* path, uint64_t row_count, size_t row_bytes, const void* rows) {
    if (!path || (!rows && row_count && row_bytes) || row_bytes > 0xFFFFFFFFull) return HERDB_ERR_ARG;
    const std::string tmp = std::string(path) + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return HERDB_ERR_IO;
    const uint32_t len = (uint32_t)row_bytes;
    bool ok = true;
    for (uint64_t r = 0; r < row_count && ok; ++r) {
        ok = fwrite(&len, sizeof len, 1, f) == 1;
        if (ok && row_bytes) ok = fwrite(static_cast<const uint8_t*>(rows) + r * row_bytes, row_bytes, 1, f) == 1;
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return HERDB_ERR_IO; }
    return HERDB_OK;
}

extern "C" uint64_t herdb_partition_rows(uint64_t rows, uint32_t partitions, uint32_t k) {
    if (!partitions || k >= partitions) return 0;
    return rows / partitions + (k < rows % partitions ? 1 : 0);
}

// Task threads.  A small map task is coalesced with the map tasks queued right behind it that belong to the same session
// and carry the same circuit, up to HERDB_COALESCE_ROWS rows in total: herdsman cuts a frame into partitions
// (storage_service.cpp:121-129) and a partition of a few hundred rows leaves most of a grid round empty at every level.
// The group is evaluated as one batch (run_map_tasks); every task keeps its own files, handle and status.  If the group
// fails, its tasks are run again one by one, so that only the task at fault reports the failure.
static void worker_loop(herdb_ctx* ctx) {
    for (;;) {
        std::vector<TaskReq> group;
        {
            std::unique_lock<std::mutex> l(ctx->q_mu);
            // (while another thread is collecting a group, the tasks that arrive are its to look at first)
            ctx->q_cv.wait(l, [&] { return ctx->stop || (!ctx->queue.empty() && !ctx->lingering); });
            if (ctx->queue.empty()) return;            // stop requested and nothing left
            group.push_back(std::move(ctx->queue.front()));
            ctx->queue.pop_front();
            ctx->in_flight++;
            if (!group[0].reduce && group[0].row_count < ctx->coalesce_rows) {
                uint64_t total = group[0].row_count;
                // wait up to 1 ms for the next task after every arrival, 5 ms in all (a task of a few hundred rows takes ~100 ms)
                const auto t_first = std::chrono::steady_clock::now();
                auto deadline = t_first + std::chrono::milliseconds(1);
                ctx->lingering++;
                for (;;) {
                    if (ctx->queue.empty()) {          // the executor hands its window over task by task: linger for a moment
                        if (ctx->stop || !ctx->q_cv.wait_until(l, deadline, [&] { return ctx->stop || !ctx->queue.empty(); }) || ctx->queue.empty()) break;
                    }
                    const TaskReq& nx = ctx->queue.front();
                    if (nx.reduce || nx.session != group[0].session || nx.schema_type != group[0].schema_type ||
                        total + nx.row_count > ctx->coalesce_rows || !same_circuit(nx.circuit, group[0].circuit)) break;
                    total += nx.row_count;
                    group.push_back(std::move(ctx->queue.front()));
                    ctx->queue.pop_front();
                    ctx->in_flight++;
                    deadline = std::min(std::chrono::steady_clock::now() + std::chrono::milliseconds(1), t_first + std::chrono::milliseconds(5));
                }
                ctx->lingering--;
            }
        }
        ctx->q_cv.notify_all();                        // what this thread left in the queue is for the others
        std::vector<const TaskReq*> ptrs;
        for (const TaskReq& t : group) ptrs.push_back(&t);
        std::vector<int> rcs(group.size(), run_tasks(ctx, ptrs));
        if (group.size() > 1 && rcs[0] != HERDB_OK)
            for (size_t k = 0; k < group.size(); ++k) rcs[k] = run_tasks(ctx, std::vector<const TaskReq*>{&group[k]});
        for (size_t k = 0; k < group.size(); ++k) {
            // CUDA / IO / format errors => ERROR (job fails, executor.cpp:168-178).  Out of device or pinned memory is
            // worth a retry (TIME_OUT => rescheduled up to three times, executor.cpp:136-167): the budgets that size a
            // task's row chunks are halved first, so the retry runs with smaller buffers.
            int status = rcs[k] == HERDB_OK ? HERDB_TASK_COMPLETED : HERDB_TASK_ERROR;
            if (rcs[k] == HERDB_ERR_NOMEM) {
                ctx->mem_budget = std::max<size_t>(ctx->mem_budget / 2, (size_t)64 << 20);
                ctx->io_budget = std::max<size_t>(ctx->io_budget / 2, (size_t)16 << 20);
                status = HERDB_TASK_TIME_OUT;
            }
            if (group[k].done) group[k].done(group[k].user, status);
        }
        {
            std::lock_guard<std::mutex> l(ctx->q_mu);
            ctx->in_flight -= group.size();
        }
        ctx->idle_cv.notify_all();
    }
}

static int enqueue(herdb_ctx* ctx, TaskReq&& t) {
    {
        std::lock_guard<std::mutex> l(ctx->q_mu);
        if (ctx->stop) return HERDB_ERR_ARG;
        ctx->queue.push_back(std::move(t));
    }
    ctx->q_cv.notify_one();
    return HERDB_OK;
}

extern "C" int herdb_map_async(herdb_ctx* ctx, const char* session, const char* in_df, uint32_t in_part, uint64_t row_count,
                               const char* out_df, uint32_t out_part, int schema_type, const herdb_circuit* circuit, herdb_done_fn done, void* user) {
    if (!ctx || !session || !in_df || !out_df || !circuit) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(circuit, cc, err);                     // refused here if malformed; compiled again for the chunk size it runs at
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    TaskReq t;
    t.circuit.assign(circuit);                                      // the circuit is copied during the call
    t.session = session; t.in_df = in_df; t.out_df = out_df; t.in_part = in_part; t.out_part = out_part;
    t.row_count = row_count; t.schema_type = schema_type; t.done = done; t.user = user;
    return enqueue(ctx, std::move(t));
}

extern "C" int herdb_reduce_async(herdb_ctx* ctx, const char* session, const herdb_input_frame* inputs, size_t n_inputs,
                                  const char* out_df, uint32_t out_part, int schema_type, const herdb_circuit* circuit, herdb_done_fn done, void* user) {
    if (!ctx || !session || (!inputs && n_inputs) || !out_df || !circuit) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(circuit, cc, err);
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    rc = check_reduce_circuit(ctx, circuit); if (rc) return rc;
    TaskReq t;
    t.circuit.assign(circuit);
    t.reduce = true; t.session = session; t.out_df = out_df; t.out_part = out_part; t.schema_type = schema_type; t.done = done; t.user = user;
    for (size_t k = 0; k < n_inputs; ++k) {
        if (!inputs[k].df_uuid) return HERDB_ERR_ARG;
        t.inputs.push_back(InFrame{inputs[k].df_uuid, inputs[k].partition, inputs[k].row_count});
    }
    return enqueue(ctx, std::move(t));
}

extern "C" int herdb_wait_idle(herdb_ctx* ctx) {
    if (!ctx) return HERDB_ERR_ARG;
    std::unique_lock<std::mutex> l(ctx->q_mu);
    ctx->idle_cv.wait(l, [&] { return ctx->queue.empty() && ctx->in_flight == 0; });
    return HERDB_OK;
}
