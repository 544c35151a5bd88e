// backend.cpp -- implementation of the C ABI in include/herdb.h (see backend.h).
#include "backend.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>

#include "host_tables.h"

using namespace herdb;

namespace {

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ctx->set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                    \
            return HERDB_ERR_CUDA;                                                                 \
        }                                                                                          \
    } while (0)

int grow(herdb_ctx* ctx, void** p, size_t* cap, size_t need) {
    if (*cap >= need) return HERDB_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *cap = 0;
    cudaError_t e = cudaMalloc(p, need);
    if (e != cudaSuccess) { ctx->set_error(std::string("cudaMalloc: ") + cudaGetErrorString(e)); return HERDB_ERR_NOMEM; }
    *cap = need;
    return HERDB_OK;
}

// CUDA-event timing of the two hot kernels (summed into herdb_stats when the stream is drained)
int timed_begin(herdb_ctx* ctx, Device& d, int kind) {
    if (d.timed_used == d.timed.size()) {
        Device::Timed t{}; t.kind = kind;
        CU(cudaEventCreate(&t.a)); CU(cudaEventCreate(&t.b));
        d.timed.push_back(t);
    }
    d.timed[d.timed_used].kind = kind;
    CU(cudaEventRecord(d.timed[d.timed_used].a, d.stream));
    return HERDB_OK;
}
int timed_end(herdb_ctx* ctx, Device& d) {
    CU(cudaEventRecord(d.timed[d.timed_used].b, d.stream));
    d.timed_used++;
    return HERDB_OK;
}
int drain(herdb_ctx* ctx, Device& d) {
    CU(cudaStreamSynchronize(d.stream));
    for (size_t i = 0; i < d.timed_used; ++i) {
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, d.timed[i].a, d.timed[i].b));
        if (d.timed[i].kind == 0) { d.stats.blind_rotate_ms += ms; d.stats.blind_rotate_launches++; }
        else { d.stats.keyswitch_ms += ms; d.stats.keyswitch_launches++; }
    }
    d.timed_used = 0;
    return HERDB_OK;
}

const DeviceKey* find_key(Device& d, const char* session) {
    auto it = d.keys.find(session ? session : "");
    return it == d.keys.end() ? nullptr : &it->second;
}

// what the level scheduler should plan for: n_rows split evenly over `devices` devices, each launch covering a
// device's share (row chunking by the memory budget only matters far above the row counts where levels have tails)
SchedHint sched_hint(const herdb_ctx* ctx, size_t n_rows, size_t devices) {
    SchedHint h;
    if (ctx->devs.empty() || !devices) return h;
    h.rows = (n_rows + devices - 1) / devices;
    h.wave = (u32)ctx->devs[0]->sm_count * kBrCtsPerCta;
    return h;
}

// one circuit level (or a single-gate pseudo level): blind rotation then key switch, `rows` rows
int run_level(herdb_ctx* ctx, Device& d, const DeviceKey& key, const GateDesc* d_gates, u32 n_gates,
              const Pool pools[4], u32 rows, u32* d_trace = nullptr, u32 steps = UINT32_MAX) {
    const u32 jobs = n_gates * rows;
    if (!jobs) return HERDB_OK;
    BrParams bp{};
    bp.bsk = key.bsk; bp.mono = d.mono; bp.ftw2 = d.ftw2; bp.itw1 = d.itw1; bp.gates = d_gates;
    for (int i = 0; i < 4; ++i) bp.pools[i] = pools[i];
    bp.ext = d.ext; bp.trace = d_trace; bp.n = key.n; bp.steps = std::min(steps, key.n); bp.rows = rows; bp.jobs = jobs;
    int rc = timed_begin(ctx, d, 0); if (rc) return rc;
    CU(launch_blind_rotate(bp, d.sm_count, d.stream));
    rc = timed_end(ctx, d); if (rc) return rc;
    KsParams kp{};
    kp.ksk = key.ksk; kp.ext = d.ext; kp.gates = d_gates;
    for (int i = 0; i < 4; ++i) kp.pools[i] = pools[i];
    kp.ks_stride = key.ks_stride; kp.n = key.n; kp.rows = rows; kp.jobs = jobs;
    rc = timed_begin(ctx, d, 1); if (rc) return rc;
    CU(launch_keyswitch(kp, d.sm_count, d.stream));
    rc = timed_end(ctx, d); if (rc) return rc;
    d.stats.kernel_launches += 2;
    d.stats.gates_bootstrapped += jobs;
    return HERDB_OK;
}

size_t bytes_per_row(const CompiledCircuit& cc, u32 n) {
    return (size_t)cc.slots * (n + 1) * 2 + (size_t)cc.max_level_width * kExtStride * 2;
}

// whole circuit over rows resident on the device (dense [rows][bits][n+1] u16 in and out)
int run_circuit_dev(herdb_ctx* ctx, Device& d, const DeviceKey& key, const CompiledCircuit& cc,
                    const u16* d_rows_in, u16* d_rows_out, size_t n_rows) {
    if (!n_rows) return HERDB_OK;
    const u32 n = key.n, w = n + 1;
    // gate tables of all levels, uploaded once
    size_t total_gates = 0;
    for (auto& L : cc.levels) total_gates += L.size();
    std::vector<GateDesc> flat; flat.reserve(total_gates);
    for (auto& L : cc.levels) flat.insert(flat.end(), L.begin(), L.end());
    int rc = grow(ctx, (void**)&d.gates, &d.gates_cap, std::max<size_t>(1, total_gates) * sizeof(GateDesc)); if (rc) return rc;
    rc = grow(ctx, (void**)&d.outs, &d.outs_cap, std::max<size_t>(1, cc.outs.size()) * sizeof(OutDesc)); if (rc) return rc;
    if (total_gates) CU(cudaMemcpyAsync(d.gates, flat.data(), total_gates * sizeof(GateDesc), cudaMemcpyHostToDevice, d.stream));
    if (!cc.outs.empty()) CU(cudaMemcpyAsync(d.outs, cc.outs.data(), cc.outs.size() * sizeof(OutDesc), cudaMemcpyHostToDevice, d.stream));
    CU(cudaStreamSynchronize(d.stream));      // `flat` is pageable host memory

    // rows are processed in chunks sized so that the slot file + extraction buffer fit the budget
    const size_t bpr = std::max<size_t>(1, bytes_per_row(cc, n));
    size_t chunk = std::max<size_t>(1, ctx->mem_budget / bpr);
    if (cc.max_level_width) chunk = std::min<size_t>(chunk, (size_t)0x7FFFFFFF / cc.max_level_width);
    chunk = std::min(chunk, n_rows);
    rc = grow(ctx, (void**)&d.slots, &d.slots_bytes, std::max<size_t>(16, chunk * cc.slots * w * 2)); if (rc) return rc;
    rc = grow(ctx, (void**)&d.ext, &d.ext_bytes, std::max<size_t>(16, chunk * cc.max_level_width * kExtStride * 2)); if (rc) return rc;

    for (size_t r0 = 0; r0 < n_rows; r0 += chunk) {
        const u32 rows = (u32)std::min(chunk, n_rows - r0);
        Pool pools[4] = {};
        pools[kPoolIn] = Pool{const_cast<u16*>(d_rows_in) + r0 * cc.in_bits * w, cc.in_bits, 0};
        pools[kPoolSlots] = Pool{d.slots, cc.slots, 0};
        pools[kPoolOut] = Pool{d_rows_out + r0 * cc.out_bits * w, cc.out_bits, 0};
        size_t off = 0;
        for (auto& L : cc.levels) {
            rc = run_level(ctx, d, key, d.gates + off, (u32)L.size(), pools, rows); if (rc) return rc;
            off += L.size();
        }
        GatherParams gp{};
        gp.outs = d.outs; for (int i = 0; i < 4; ++i) gp.pools[i] = pools[i];
        gp.rows_out = pools[kPoolOut].base; gp.n = n; gp.rows = rows; gp.out_bits = cc.out_bits;
        CU(launch_gather_outputs(gp, d.sm_count, d.stream));
        d.stats.kernel_launches += 1;
    }
    return HERDB_OK;
}

int key_check(herdb_ctx* ctx, const herdb_key_header& h, size_t len) {
    if (len < sizeof h || h.magic != HERDB_KEY_MAGIC || h.version != HERDB_KEY_VERSION) { ctx->set_error("key blob: bad magic/version"); return HERDB_ERR_FORMAT; }
    if (h.ring_n != kN || h.Q != kQ || h.q != kLweQ || h.qKS != kQKS || h.baseG_log != HERDB_BASEG_LOG || h.digitsG != kDG ||
        h.baseKS_log != HERDB_BASEKS_LOG || h.digitsKS != kDKS || h.n == 0 || h.n > kLweMax) {
        ctx->set_error("key blob: parameter set not supported by this build (see include/herdb_params.h)"); return HERDB_ERR_FORMAT;
    }
    if (len < herdb_client_key_blob_size(h.n)) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
    return HERDB_OK;
}

// raw blob already in device memory -> kernel-ready key
int install_key_from_dev(herdb_ctx* ctx, Device& d, const char* session, const herdb_key_header& h, const uint8_t* d_blob) {
    CU(cudaSetDevice(d.id));
    DeviceKey k; k.n = h.n;
    const u32 polys = h.n * 2 * kDG2 * 2;
    const size_t ks_rows = (size_t)kN * kBaseKS * kDKS;
    k.ks_stride = ((h.n + 1 + 7) / 8) * 8;
    CU(cudaMalloc((void**)&k.bsk, (size_t)polys * kN * 4));
    CU(cudaMalloc((void**)&k.ksk, ks_rows * k.ks_stride * 2));
    const u32* raw_bsk = reinterpret_cast<const u32*>(d_blob + sizeof(herdb_key_header));
    const u16* raw_ksk = reinterpret_cast<const u16*>(d_blob + sizeof(herdb_key_header) + (size_t)polys * kN * 4);
    CU(launch_bsk_convert(raw_bsk, k.bsk, d.ftw2, polys, d.sm_count, d.stream));
    CU(launch_ksk_relayout(raw_ksk, k.ksk, h.n, k.ks_stride, ks_rows, d.sm_count, d.stream));
    CU(cudaStreamSynchronize(d.stream));
    d.stats.kernel_launches += 3;
    auto it = d.keys.find(session);
    if (it != d.keys.end()) { cudaFree(it->second.bsk); cudaFree(it->second.ksk); }
    d.keys[session] = k;
    return HERDB_OK;
}

}  // namespace

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

extern "C" int herdb_create(herdb_ctx** out, const int* device_ids, int n_devices, const char* storage_dir, const char* key_dir) {
    if (!out) return HERDB_ERR_ARG;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return HERDB_ERR_CUDA;     // no CPU fallback
    herdb_ctx* ctx = new herdb_ctx;
    ctx->storage_dir = storage_dir ? storage_dir : "";
    ctx->key_dir = key_dir ? key_dir : "";
    const char* mb = getenv("HERDB_MEM_BUDGET_MB");
    ctx->mem_budget = (size_t)(mb ? atoll(mb) : 16384) << 20;
    std::vector<int> ids;
    if (device_ids && n_devices > 0) ids.assign(device_ids, device_ids + n_devices); else ids.push_back(0);
    std::vector<u32> mono, ftw2(kLaneTwWords), itw1(kLaneTwWords);
    build_monomial_table(mono);
    build_lane_twiddles(ftw2.data(), itw1.data());
    auto fail = [&](int rc) { *out = nullptr; herdb_destroy(ctx); return rc; };
    for (int id : ids) {
        if (id < 0 || id >= count) { return fail(HERDB_ERR_ARG); }
        Device* d = new Device; d->id = id; ctx->devs.push_back(d);
        if (cudaSetDevice(id) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        cudaDeviceGetAttribute(&d->sm_count, cudaDevAttrMultiProcessorCount, id);
        if (cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        if (cudaStreamCreateWithFlags(&d->copy, cudaStreamNonBlocking) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        for (auto& e : d->ev_piece)
            if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return fail(HERDB_ERR_CUDA);
        if (cudaMalloc((void**)&d->mono, mono.size() * 4) != cudaSuccess || cudaMalloc((void**)&d->ftw2, kLaneTwWords * 4) != cudaSuccess ||
            cudaMalloc((void**)&d->itw1, kLaneTwWords * 4) != cudaSuccess) return fail(HERDB_ERR_NOMEM);
        cudaMemcpy(d->mono, mono.data(), mono.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->ftw2, ftw2.data(), kLaneTwWords * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(d->itw1, itw1.data(), kLaneTwWords * 4, cudaMemcpyHostToDevice);
    }
    ctx->worker = std::thread(worker_loop, ctx);
    *out = ctx;
    return HERDB_OK;
}

extern "C" void herdb_destroy(herdb_ctx* ctx) {
    if (!ctx) return;
    {
        std::unique_lock<std::mutex> l(ctx->q_mu);
        ctx->idle_cv.wait(l, [&] { return ctx->queue.empty() && ctx->in_flight == 0; });
        ctx->stop = true;
    }
    ctx->q_cv.notify_all();
    if (ctx->worker.joinable()) ctx->worker.join();
    for (Device* d : ctx->devs) {
        cudaSetDevice(d->id);
        if (d->stream) cudaStreamSynchronize(d->stream);
        for (auto& kv : d->keys) { cudaFree(kv.second.bsk); cudaFree(kv.second.ksk); }
        for (auto& t : d->timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
        cudaFree(d->mono); cudaFree(d->ftw2); cudaFree(d->itw1); cudaFree(d->slots); cudaFree(d->ext);
        cudaFree(d->gates); cudaFree(d->outs); cudaFree(d->io_in); cudaFree(d->io_out);
        if (d->stream) cudaStreamDestroy(d->stream);
        if (d->copy) cudaStreamDestroy(d->copy);
        for (auto& e : d->ev_piece) if (e) cudaEventDestroy(e);
        delete d;
    }
    delete ctx;
}

// ================================================================================================ keys
extern "C" int herdb_load_key_dev(herdb_ctx* ctx, const char* session, int device, const void* d_blob, size_t len) {
    if (!ctx || !session || !d_blob || device < 0 || device >= (int)ctx->devs.size()) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    herdb_key_header h;
    if (len < sizeof h) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
    CU(cudaMemcpy(&h, d_blob, sizeof h, cudaMemcpyDeviceToHost));
    int rc = key_check(ctx, h, len); if (rc) return rc;
    return install_key_from_dev(ctx, d, session, h, static_cast<const uint8_t*>(d_blob));
}

extern "C" int herdb_load_key_mem(herdb_ctx* ctx, const char* session, const void* blob, size_t len) {
    if (!ctx || !session || !blob) return HERDB_ERR_ARG;
    herdb_key_header h;
    if (len < sizeof h) { ctx->set_error("key blob: truncated"); return HERDB_ERR_FORMAT; }
    memcpy(&h, blob, sizeof h);
    int rc = key_check(ctx, h, len); if (rc) return rc;
    const size_t need = herdb_client_key_blob_size(h.n);
    // device 0 receives the blob from the host; further devices receive it GPU-to-GPU
    // (peer copy over NVLink; one process per GPU uses herdb_load_key_dev after an NCCL broadcast)
    uint8_t* d0_raw = nullptr;
    for (size_t di = 0; di < ctx->devs.size(); ++di) {
        Device& d = *ctx->devs[di];
        std::lock_guard<std::mutex> l(d.mu);
        CU(cudaSetDevice(d.id));
        uint8_t* raw = nullptr;
        CU(cudaMalloc((void**)&raw, need));
        cudaError_t e;
        if (di == 0) { e = cudaMemcpy(raw, blob, need, cudaMemcpyHostToDevice); d.stats.h2d_bytes += need; }
        else e = cudaMemcpyPeer(raw, d.id, d0_raw, ctx->devs[0]->id, need);
        if (e != cudaSuccess) { cudaFree(raw); ctx->set_error(std::string("key upload: ") + cudaGetErrorString(e)); return HERDB_ERR_CUDA; }
        rc = install_key_from_dev(ctx, d, session, h, raw);
        if (di == 0 && ctx->devs.size() > 1) d0_raw = raw; else cudaFree(raw);
        if (rc) { if (d0_raw) { cudaSetDevice(ctx->devs[0]->id); cudaFree(d0_raw); } return rc; }
    }
    if (d0_raw) { cudaSetDevice(ctx->devs[0]->id); cudaFree(d0_raw); }
    return HERDB_OK;
}

extern "C" int herdb_load_key(herdb_ctx* ctx, const char* session, int schema_type) {
    if (!ctx || !session) return HERDB_ERR_ARG;
    // <key_dir>/<session_uuid>/<int(SchemaType)>.key   (key_service.cpp:28-31)
    const std::string path = ctx->key_dir + "/" + session + "/" + std::to_string(schema_type) + ".key";
    std::ifstream f(path, std::ios::binary | std::ios::ate);
    if (!f) { ctx->set_error("cannot open key file " + path); return HERDB_ERR_IO; }
    const std::streamsize sz = f.tellg();
    f.seekg(0);
    std::vector<char> buf((size_t)sz);
    if (!f.read(buf.data(), sz)) { ctx->set_error("cannot read key file " + path); return HERDB_ERR_IO; }
    return herdb_load_key_mem(ctx, session, buf.data(), buf.size());
}

extern "C" int herdb_unload_key(herdb_ctx* ctx, const char* session) {
    if (!ctx || !session) return HERDB_ERR_ARG;
    bool found = false;
    for (Device* d : ctx->devs) {
        std::lock_guard<std::mutex> l(d->mu);
        auto it = d->keys.find(session);
        if (it == d->keys.end()) continue;
        cudaSetDevice(d->id);
        cudaFree(it->second.bsk); cudaFree(it->second.ksk);
        d->keys.erase(it); found = true;
    }
    return found ? HERDB_OK : HERDB_ERR_NOKEY;
}

extern "C" int herdb_key_lwe_n(herdb_ctx* ctx, const char* session, uint32_t* n_out) {
    if (!ctx || !session || !n_out || ctx->devs.empty()) return HERDB_ERR_ARG;
    std::lock_guard<std::mutex> l(ctx->devs[0]->mu);
    const DeviceKey* k = find_key(*ctx->devs[0], session);
    if (!k) return HERDB_ERR_NOKEY;
    *n_out = k->n;
    return HERDB_OK;
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
                             size_t first, size_t count, size_t chunk) {
    const size_t w = key.n + 1;
    for (size_t r0 = first; r0 < first + count; r0 += chunk) {
        const u32 rows = (u32)std::min(chunk, first + count - r0);
        Pool pools[4] = {};
        pools[0] = Pool{const_cast<u16*>(d_a) + r0 * w, 1, 0};
        pools[2] = Pool{const_cast<u16*>(d_b) + r0 * w, 1, 0};
        pools[3] = Pool{d_out + r0 * w, 1, 0};
        int rc = run_level(ctx, d, key, d.gates, 1, pools, rows); if (rc) return rc;
    }
    return HERDB_OK;
}

extern "C" int herdb_eval_gate_batch_dev(herdb_ctx* ctx, const char* session, int device, int gate,
                                         const uint16_t* d_a, const uint16_t* d_b, uint16_t* d_out, size_t batch) {
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
    rc = gate_batch_launch(ctx, d, *key, d_a, d_b, d_out, 0, batch, chunk); if (rc) return rc;
    return drain(ctx, d);
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
        CU(cudaStreamWaitEvent(d.copy, d.ev_piece[0], 0));
        CU(d2h(0, d.copy));
        CU(cudaStreamWaitEvent(d.copy, d.ev_piece[2], 0));
        CU(d2h(1, d.copy));
        CU(d2h(2, d.stream));
        CU(cudaStreamSynchronize(d.copy));
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
    int rc = compile_circuit(circuit, cc, err, sched_hint(ctx, n_rows, 1));
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    if (!n_rows) return HERDB_OK;
    if ((!d_rows_in && cc.in_bits) || (!d_rows_out && cc.out_bits)) return HERDB_ERR_ARG;
    Device& d = *ctx->devs[device];
    std::lock_guard<std::mutex> l(d.mu);
    CU(cudaSetDevice(d.id));
    const DeviceKey* key = find_key(d, session);
    if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
    rc = run_circuit_dev(ctx, d, *key, cc, d_rows_in, d_rows_out, n_rows); if (rc) return rc;
    return drain(ctx, d);
}

// host rows -> devices (contiguous row slices, one host thread per device) -> host rows.  `stages` is a chain of
// map circuits: the output row of stage k is the input row of stage k+1 and stays in HBM (two ping-pong buffers).
static int eval_stages_host(herdb_ctx* ctx, const char* session, const std::vector<const CompiledCircuit*>& stages,
                            const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    const size_t G = ctx->devs.size();
    std::vector<int> rcs(G, HERDB_OK);
    auto run_slice = [&](size_t di) {
        Device& d = *ctx->devs[di];
        const size_t r0 = n_rows * di / G, r1 = n_rows * (di + 1) / G, rows = r1 - r0;
        if (!rows) return;
        std::lock_guard<std::mutex> l(d.mu);
        auto body = [&]() -> int {
            CU(cudaSetDevice(d.id));
            const DeviceKey* key = find_key(d, session);
            if (!key) { ctx->set_error("no key for session"); return HERDB_ERR_NOKEY; }
            const size_t w = key->n + 1;
            // bound the staging buffers: process the slice in super-chunks of rows
            const size_t in_row = (size_t)stages.front()->in_bits * w * 2, out_row = (size_t)stages.back()->out_bits * w * 2;
            size_t buf_row = std::max(in_row, out_row), scratch_row = 0;
            for (const CompiledCircuit* cc : stages) {
                buf_row = std::max(buf_row, std::max((size_t)cc->in_bits, (size_t)cc->out_bits) * w * 2);
                scratch_row = std::max(scratch_row, bytes_per_row(*cc, key->n));
            }
            const size_t per_row = 2 * buf_row + scratch_row;
            const size_t step = std::max<size_t>(1, std::min(rows, ctx->mem_budget / std::max<size_t>(1, per_row)));
            int rc = grow(ctx, (void**)&d.io_in, &d.io_in_bytes, std::max<size_t>(16, step * buf_row)); if (rc) return rc;
            rc = grow(ctx, (void**)&d.io_out, &d.io_out_bytes, std::max<size_t>(16, step * buf_row)); if (rc) return rc;
            for (size_t s0 = 0; s0 < rows; s0 += step) {
                const size_t cnt = std::min(step, rows - s0);
                u16 *src = d.io_in, *dst = d.io_out;
                if (in_row) CU(cudaMemcpyAsync(src, (const uint8_t*)rows_in + (r0 + s0) * in_row, cnt * in_row, cudaMemcpyHostToDevice, d.stream));
                d.stats.h2d_bytes += cnt * in_row;
                for (const CompiledCircuit* cc : stages) {
                    rc = run_circuit_dev(ctx, d, *key, *cc, src, dst, cnt); if (rc) return rc;
                    std::swap(src, dst);
                }
                if (out_row) CU(cudaMemcpyAsync((uint8_t*)rows_out + (r0 + s0) * out_row, src, cnt * out_row, cudaMemcpyDeviceToHost, d.stream));
                d.stats.d2h_bytes += cnt * out_row;
                rc = drain(ctx, d); if (rc) return rc;
            }
            return HERDB_OK;
        };
        rcs[di] = body();
    };
    if (G == 1) run_slice(0);
    else {
        std::vector<std::thread> th;
        for (size_t di = 0; di < G; ++di) th.emplace_back(run_slice, di);
        for (auto& t : th) t.join();
    }
    for (int rc : rcs) if (rc) return rc;
    return HERDB_OK;
}

static int eval_rows_host(herdb_ctx* ctx, const char* session, const CompiledCircuit& cc,
                          const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    return eval_stages_host(ctx, session, std::vector<const CompiledCircuit*>{&cc}, rows_in, rows_out, n_rows);
}

extern "C" int herdb_eval_pipeline_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* const* circuits, size_t n_stages,
                                        const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    if (!ctx || !session || !circuits || !n_stages) return HERDB_ERR_ARG;
    std::vector<CompiledCircuit> cc(n_stages);
    std::vector<const CompiledCircuit*> stages;
    for (size_t k = 0; k < n_stages; ++k) {
        std::string err;
        if (!circuits[k]) return HERDB_ERR_ARG;
        int rc = compile_circuit(circuits[k], cc[k], err, sched_hint(ctx, n_rows, ctx->devs.size()));
        if (rc) { ctx->set_error("circuit of stage " + std::to_string(k) + ": " + err); return rc; }
        if (k && cc[k].in_bits != cc[k - 1].out_bits) {
            ctx->set_error("pipeline: stage " + std::to_string(k) + " does not take the row stage " + std::to_string(k - 1) + " produces");
            return HERDB_ERR_ARG;
        }
        stages.push_back(&cc[k]);
    }
    if (!n_rows) return HERDB_OK;
    if ((!rows_in && cc.front().in_bits) || (!rows_out && cc.back().out_bits)) return HERDB_ERR_ARG;
    return eval_stages_host(ctx, session, stages, rows_in, rows_out, n_rows);
}

extern "C" int herdb_eval_circuit_rows(herdb_ctx* ctx, const char* session, const herdb_circuit* circuit,
                                       const uint16_t* rows_in, uint16_t* rows_out, size_t n_rows) {
    if (!ctx || !session || !circuit) return HERDB_ERR_ARG;
    CompiledCircuit cc; std::string err;
    int rc = compile_circuit(circuit, cc, err, sched_hint(ctx, n_rows, ctx->devs.size()));
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    if (!n_rows) return HERDB_OK;
    if ((!rows_in && cc.in_bits) || (!rows_out && cc.out_bits)) return HERDB_ERR_ARG;
    return eval_rows_host(ctx, session, cc, rows_in, rows_out, n_rows);
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
    for (Device* d : ctx->devs) { std::lock_guard<std::mutex> l(d->mu); d->stats = herdb_stats{}; }
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

int read_partition(herdb_ctx* ctx, const std::string& path, uint64_t row_count, size_t row_bytes, std::vector<uint16_t>& rows) {
    rows.resize(row_count * row_bytes / 2);
    const int rc = herdb_partition_read(path.c_str(), row_count, row_bytes, rows.data());
    if (rc == HERDB_ERR_IO) ctx->set_error("cannot open partition " + path);
    else if (rc) ctx->set_error("partition " + path + ": row count / row length does not match the task (row_count, circuit inputs)");
    return rc;
}

int write_partition(herdb_ctx* ctx, const std::string& path, uint64_t row_count, size_t row_bytes, const std::vector<uint16_t>& rows) {
    const int rc = herdb_partition_write(path.c_str(), row_count, row_bytes, rows.data());
    if (rc) ctx->set_error("cannot write partition " + path);
    return rc;
}

int session_lwe_n(herdb_ctx* ctx, const std::string& session, uint32_t* n);

int run_map_task(herdb_ctx* ctx, const MapTaskReq& t) {
    uint32_t n = 0;
    int rc = session_lwe_n(ctx, t.session, &n);
    if (rc) return rc;
    const size_t w = n + 1;
    std::vector<uint16_t> rows_in, rows_out;
    rc = read_partition(ctx, partition_path(ctx, t.session, t.in_df, t.in_part), t.row_count, t.circuit.in_bits * w * 2, rows_in);
    if (rc) return rc;
    rows_out.resize(t.row_count * t.circuit.out_bits * w);
    if (t.row_count) {
        rc = eval_rows_host(ctx, t.session.c_str(), t.circuit, rows_in.data(), rows_out.data(), t.row_count);
        if (rc) return rc;
    }
    return write_partition(ctx, partition_path(ctx, t.session, t.out_df, t.out_part), t.row_count, t.circuit.out_bits * w * 2, rows_out);
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
int reduce_rows_host(herdb_ctx* ctx, const char* session, const CompiledCircuit& cc, std::vector<uint16_t>& cur, size_t m, size_t w) {
    const size_t row = (size_t)cc.out_bits * w;
    std::vector<uint16_t> next;
    while (m > 1) {
        const size_t pairs = m / 2, odd = m & 1;
        next.resize((pairs + odd) * row);
        int rc = eval_rows_host(ctx, session, cc, cur.data(), next.data(), pairs);
        if (rc) return rc;
        if (odd) memcpy(next.data() + pairs * row, cur.data() + (m - 1) * row, row * 2);
        cur.swap(next);
        m = pairs + odd;
    }
    cur.resize(row);
    return HERDB_OK;
}

int session_lwe_n(herdb_ctx* ctx, const std::string& session, uint32_t* n) {
    int rc = herdb_key_lwe_n(ctx, session.c_str(), n);
    if (rc == HERDB_ERR_NOKEY) {                       // first task of a session: pull the key from <key_dir>
        rc = herdb_load_key(ctx, session.c_str(), 1 /* SchemaType::BINFHE, schema_type.hpp:9-13 */);
        if (rc) return rc;
        rc = herdb_key_lwe_n(ctx, session.c_str(), n);
    }
    return rc;
}

int run_reduce_task(herdb_ctx* ctx, const MapTaskReq& t) {
    uint32_t n = 0;
    int rc = session_lwe_n(ctx, t.session, &n);
    if (rc) return rc;
    const size_t w = n + 1, row_bytes = (size_t)t.circuit.out_bits * w * 2;
    std::vector<uint16_t> rows, part;
    size_t m = 0;
    for (const InFrame& f : t.inputs) {
        rc = read_partition(ctx, partition_path(ctx, t.session, f.df, f.part), f.row_count, row_bytes, part);
        if (rc) return rc;
        rows.insert(rows.end(), part.begin(), part.end());
        m += f.row_count;
    }
    if (!m) { ctx->set_error("reduce task without input rows"); return HERDB_ERR_ARG; }
    rc = reduce_rows_host(ctx, t.session.c_str(), t.circuit, rows, m, w);
    if (rc) return rc;
    return write_partition(ctx, partition_path(ctx, t.session, t.out_df, t.out_part), 1, row_bytes, rows);
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
    rc = reduce_rows_host(ctx, session, cc, cur, n_rows, w); if (rc) return rc;
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

static void worker_loop(herdb_ctx* ctx) {
    for (;;) {
        MapTaskReq t;
        {
            std::unique_lock<std::mutex> l(ctx->q_mu);
            ctx->q_cv.wait(l, [&] { return ctx->stop || !ctx->queue.empty(); });
            if (ctx->queue.empty()) return;            // stop requested and nothing left
            t = std::move(ctx->queue.front());
            ctx->queue.pop_front();
            ctx->in_flight++;
        }
        const int rc = t.reduce ? run_reduce_task(ctx, t) : run_map_task(ctx, t);
        // CUDA / IO / format errors => ERROR (job fails, executor.cpp:168-178); TIME_OUT is never produced
        if (t.done) t.done(t.user, rc == HERDB_OK ? HERDB_TASK_COMPLETED : HERDB_TASK_ERROR);
        {
            std::lock_guard<std::mutex> l(ctx->q_mu);
            ctx->in_flight--;
        }
        ctx->idle_cv.notify_all();
    }
}

extern "C" int herdb_map_async(herdb_ctx* ctx, const char* session, const char* in_df, uint32_t in_part, uint64_t row_count,
                               const char* out_df, uint32_t out_part, const herdb_circuit* circuit, herdb_done_fn done, void* user) {
    if (!ctx || !session || !in_df || !out_df || !circuit) return HERDB_ERR_ARG;
    MapTaskReq t;
    std::string err;
    int rc = compile_circuit(circuit, t.circuit, err, sched_hint(ctx, row_count, ctx->devs.size()));   // the circuit is copied (compiled) during the call
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    t.session = session; t.in_df = in_df; t.out_df = out_df; t.in_part = in_part; t.out_part = out_part;
    t.row_count = row_count; t.done = done; t.user = user;
    {
        std::lock_guard<std::mutex> l(ctx->q_mu);
        if (ctx->stop) return HERDB_ERR_ARG;
        ctx->queue.push_back(std::move(t));
    }
    ctx->q_cv.notify_one();
    return HERDB_OK;
}

extern "C" int herdb_reduce_async(herdb_ctx* ctx, const char* session, const herdb_input_frame* inputs, size_t n_inputs,
                                  const char* out_df, uint32_t out_part, const herdb_circuit* circuit, herdb_done_fn done, void* user) {
    if (!ctx || !session || (!inputs && n_inputs) || !out_df || !circuit) return HERDB_ERR_ARG;
    MapTaskReq t;
    std::string err;
    int rc = compile_circuit(circuit, t.circuit, err);
    if (rc) { ctx->set_error("circuit: " + err); return rc; }
    rc = check_reduce_circuit(ctx, circuit); if (rc) return rc;
    t.reduce = true; t.session = session; t.out_df = out_df; t.out_part = out_part; t.done = done; t.user = user;
    for (size_t k = 0; k < n_inputs; ++k) {
        if (!inputs[k].df_uuid) return HERDB_ERR_ARG;
        t.inputs.push_back(InFrame{inputs[k].df_uuid, inputs[k].partition, inputs[k].row_count});
    }
    {
        std::lock_guard<std::mutex> l(ctx->q_mu);
        if (ctx->stop) return HERDB_ERR_ARG;
        ctx->queue.push_back(std::move(t));
    }
    ctx->q_cv.notify_one();
    return HERDB_OK;
}

extern "C" int herdb_wait_idle(herdb_ctx* ctx) {
    if (!ctx) return HERDB_ERR_ARG;
    std::unique_lock<std::mutex> l(ctx->q_mu);
    ctx->idle_cv.wait(l, [&] { return ctx->queue.empty() && ctx->in_flight == 0; });
    return HERDB_OK;
}
