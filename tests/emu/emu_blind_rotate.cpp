// CPU emulation of the blind-rotation CTA body (herdsman_b200/csrc/br_core.cuh): the very
// same source the CUDA kernel compiles, run by 256 host threads with pthread barriers in
// place of __syncthreads/__syncwarp.  TEST ONLY -- lets the indexing, swizzles, lazy
// ranges and table layouts be checked against the oracle without a GPU.
#include <pthread.h>
#include <cstdio>
#include <cstring>
#include <vector>
#include "../../herdsman_b200/csrc/br_core.cuh"
#include "../../herdsman_b200/csrc/host_tables.h"

using namespace herdb;

struct HostEnv {
    u32 tid_;
    pthread_barrier_t* cta;
    pthread_barrier_t* warps;
    u32 tid() const { return tid_; }
    void sync() { pthread_barrier_wait(cta); }
    void syncwarp() { pthread_barrier_wait(&warps[tid_ >> 5]); }
    // cp.async model that exposes both hazards: the destination is poisoned when the copy is issued (a
    // late reader of the old contents sees garbage) and filled only when a wait_group covers its group
    // (an early reader of the new contents sees garbage too)
    struct Pending { u32* dst; const u32* src; u32 group; };
    std::vector<Pending> pend;
    u32 committed = 0;
    void cp16(u32* dst, const u32* src) { for (int k = 0; k < 4; ++k) dst[k] = 0xDEADBEEFu; pend.push_back({dst, src, committed}); }
    void cp_commit() { ++committed; }
    void complete(u32 keep) {
        size_t w = 0;
        for (size_t k = 0; k < pend.size(); ++k) {
            if (pend[k].group + keep < committed) memcpy(pend[k].dst, pend[k].src, 16);
            else pend[w++] = pend[k];
        }
        pend.resize(w);
    }
    void cp_wait1() { complete(1); }
    void cp_wait0() { complete(0); }
    void cp_fence(u32) {}
    // bulk copies: issued by one thread right after a CTA barrier and consumed after the next one, so a plain
    // copy at issue time is ordered correctly by the pthread barriers (the mbarrier wait itself is GPU-only)
    void mbar_init(unsigned long long*, u32) {}
    void bulk_begin(unsigned long long*, u32) {}
    void bulk_copy(void* dst, const void* src, u32 bytes, unsigned long long*) { memcpy(dst, src, bytes); }
    void mbar_wait(unsigned long long*, u32) {}
};

template <int C> struct ThreadArg { HostEnv env; const BrParams* P; BrSmem<C>* S; u32 groups; };

template <int C> static void* thread_main(void* a) {
    ThreadArg<C>* t = (ThreadArg<C>*)a;
    for (u32 g = 0; g < t->groups; ++g) blind_rotate_group<C>(t->env, *t->P, *t->S, g, (g * t->P->steps) & 1);
    return nullptr;
}

template <int C>
static int emu_run(const void* blob, size_t len, int gate, const u16* ct_a, const u16* ct_b,
                   u32 batch, u32 steps, u16* ext_out, u32* trace_out) {
    constexpr int kC = C; constexpr int kThreads = (int)BrShape<C>::kThreads;
    const herdb_key_header* h = (const herdb_key_header*)blob;
    if (len < sizeof *h || h->magic != HERDB_KEY_MAGIC) return -1;
    const u32 n = h->n;
    // bootstrapping key -> kernel NTT convention, scaled by 1/N
    size_t polys = (size_t)n * 2 * kDG2 * 2;
    std::vector<u32> bsk(polys * kN);
    memcpy(bsk.data(), (const uint8_t*)blob + sizeof *h, bsk.size() * 4);
    for (size_t p = 0; p < polys; ++p) {
        ntt_forward_host(&bsk[p * kN]);
        for (u32 x = 0; x < kN; ++x) bsk[p * kN + x] = cmulmod(bsk[p * kN + x], kNInv);
    }
    for (size_t is = 0; is < (size_t)n * 2; ++is)                       // fold the gadget rows (top-digit elimination)
        for (u32 j = 0; j < 2; ++j)
            for (u32 col = 0; col < 2; ++col)
                for (u32 x = 0; x < kN; ++x) {
                    u32 k[4];
                    for (u32 l = 0; l < 4; ++l) k[l] = bsk[((is * kDG2 + j + 2 * l) * 2 + col) * kN + x];
                    fold_top_digit(k);
                    for (u32 l = 0; l < 4; ++l) bsk[((is * kDG2 + j + 2 * l) * 2 + col) * kN + x] = k[l];
                }
    {   // the four points of an aligned group are stored rotated by one word (key_pos, modarith.cuh)
        std::vector<u32> tmp(bsk);
        for (size_t p = 0; p < polys; ++p)
            for (u32 x = 0; x < kN; ++x) bsk[p * kN + key_pos(x)] = tmp[p * kN + x];
    }
    std::vector<u32> mono; build_monomial_table(mono);
    std::vector<u32> ftw2(kLaneTwWords), itw1(kLaneTwWords);
    build_lane_twiddles(ftw2.data(), itw1.data());
    GateDesc gd{}; gd.gate = (uint8_t)gate; gd.pool_a = 0; gd.pool_b = 1; gd.pool_out = 2;
    BrParams P{};
    P.bsk = bsk.data(); P.mono = mono.data(); P.ftw2 = ftw2.data(); P.itw1 = itw1.data(); P.gates = &gd;
    P.pools[0] = Pool{const_cast<u16*>(ct_a), n + 1, 0};
    P.pools[1] = Pool{const_cast<u16*>(ct_b), n + 1, 0};
    std::vector<u16> ext((size_t)batch * kExtStride);
    P.ext = ext.data(); P.trace = trace_out; P.n = n; P.steps = steps; P.rows = batch; P.jobs = batch; P.job0 = 0; P.job1 = batch;
    BrSmem<C>* S = new BrSmem<C>;
    pthread_barrier_t cta, warps[kThreads / 32];
    pthread_barrier_init(&cta, nullptr, kThreads);
    for (auto& w : warps) pthread_barrier_init(&w, nullptr, 32);
    std::vector<ThreadArg<C>> args(kThreads);
    std::vector<pthread_t> th(kThreads);
    u32 groups = (batch + kC - 1) / kC;
    pthread_attr_t attr; pthread_attr_init(&attr); pthread_attr_setstacksize(&attr, 1 << 20);
    for (u32 t = 0; t < (u32)kThreads; ++t) {
        args[t] = ThreadArg<C>{HostEnv{t, &cta, warps}, &P, S, groups};
        pthread_create(&th[t], &attr, thread_main<C>, &args[t]);
    }
    for (auto& t : th) pthread_join(t, nullptr);
    for (u32 b = 0; b < batch; ++b) memcpy(ext_out + (size_t)b * (kN + 1), &ext[(size_t)b * kExtStride], (kN + 1) * 2);
    delete S;
    return 0;
}

extern "C" int emu_blind_rotate(const void* blob, size_t len, int gate, const u16* ct_a, const u16* ct_b,
                                u32 batch, u32 steps, u16* ext_out, u32* trace_out) {
    return emu_run<4>(blob, len, gate, ct_a, ct_b, batch, steps, ext_out, trace_out);
}
// the small-job CTA shapes (C = 2, 1: whole key slice by bulk copy, 8 evaluation points per thread in phase B)
extern "C" int emu_blind_rotate_shape(int shape, const void* blob, size_t len, int gate, const u16* ct_a, const u16* ct_b,
                                      u32 batch, u32 steps, u16* ext_out, u32* trace_out) {
    if (shape == 4) return emu_run<4>(blob, len, gate, ct_a, ct_b, batch, steps, ext_out, trace_out);
    if (shape == 2) return emu_run<2>(blob, len, gate, ct_a, ct_b, batch, steps, ext_out, trace_out);
    if (shape == 1) return emu_run<1>(blob, len, gate, ct_a, ct_b, batch, steps, ext_out, trace_out);
    return -2;
}

// the FP64-pipe fold of phase B (modarith.cuh), host form, for the range / residue property test
extern "C" void emu_fold62_f64(const u64* x, const u32* addend, u32* out, size_t count) {
    for (size_t i = 0; i < count; ++i) out[i] = fold62_f64(x[i], addend[i]);
}
