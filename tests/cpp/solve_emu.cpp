// solve_emu.cpp — CPU check of the on-device solve's LOGIC (csrc/solve_dev.cuh), no GPU involved.
//
// solve_trip / solve_pre are written for one CTA of 128 CUDA threads.  Here the same source is compiled for the host and
// run by 128 std::threads: __syncthreads / __syncwarp / __all_sync become barriers, shared memory an ordinary struct.
// tests/test_solve_emu_cpu.py feeds it the per-trip normal equations of an oracle update and compares state, increments,
// policy and covariance trip by trip with the oracle's solve (which restates esekfom.hpp / laser_mapping.cpp).  What this
// cannot show is anything about the GPU (memory model, scheduling); those are the -m gpu tests' business.
//
//   solve_emu <in.bin> <out.bin>
//   in : int32 mode, max_iter, ntrips, pad; double R, limit[23], x[36], P[23*23]; double sums[ntrips][96]
//   out: per executed trip: double x_after[36], dx_last[23], int32 {search, stopped, iter, t, rematch, trips, converged, n_eff};
//        at the end double P_out[23*23]
#define FLIVO_HOST_EMU
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <atomic>
#include <barrier>
#include <thread>
#include <vector>

#include <cuda_runtime.h>
#ifndef __noinline__
#define __noinline__ __attribute__((noinline))
#endif

// ---- the few CUDA built-ins solve_dev.cuh uses, for 128 host threads ---------------------------------------------
constexpr int EMU_THREADS = 128;
static std::barrier<> g_cta(EMU_THREADS);
static std::barrier<>* g_warp[EMU_THREADS / 32];
static std::atomic<int> g_and{1};
static std::atomic<int> g_warp_and[EMU_THREADS / 32];
static thread_local int t_tid = 0;

static inline void __syncthreads() { g_cta.arrive_and_wait(); }
static inline int __syncthreads_and(int pred) {
    if (!pred) g_and.store(0);
    g_cta.arrive_and_wait();
    const int r = g_and.load();
    g_cta.arrive_and_wait();
    if (t_tid == 0) g_and.store(1);
    g_cta.arrive_and_wait();
    return r;
}
static inline void __syncwarp() { g_warp[t_tid >> 5]->arrive_and_wait(); }
static inline int __all_sync(unsigned, int pred) {
    std::atomic<int>& a = g_warp_and[t_tid >> 5];
    if (!pred) a.store(0);
    __syncwarp();
    const int r = a.load();
    __syncwarp();
    if ((t_tid & 31) == 0) a.store(1);
    __syncwarp();
    return r;
}

static unsigned long long g_shfl[EMU_THREADS / 32][32];
template <class T>
static inline T __shfl_sync(unsigned, T v, int src) {
    static_assert(sizeof(T) <= 8, "shuffle of at most 8 bytes");
    unsigned long long* buf = g_shfl[t_tid >> 5];
    unsigned long long w = 0;
    memcpy(&w, &v, sizeof(T));
    buf[t_tid & 31] = w;
    __syncwarp();
    w = buf[src & 31];
    __syncwarp();
    T out;
    memcpy(&out, &w, sizeof(T));
    return out;
}
static inline int __ffs(int v) { return __builtin_ffs(v); }

#include "../../fast-livo-noted_b200/csrc/solve_dev.cuh"

using namespace flivo;

int main(int argc, char** argv) {
    if (argc < 3) { fprintf(stderr, "usage: solve_emu <in.bin> <out.bin>\n"); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror("open"); return 2; }
    int32_t hdr[4];
    if (fread(hdr, 4, 4, f) != 4) return 2;
    const int mode = hdr[0], max_iter = hdr[1], ntrips = hdr[2];
    static Job job;
    memset(&job, 0, sizeof(job));
    job.mode = mode; job.max_iter = max_iter; job.nranks = 1;
    static double x[36], x_prop[36], P_prop[23 * 23], P_out[23 * 23], sum_out[PSTRIDE];
    bool ok = fread(&job.R, 8, 1, f) == 1 && fread(job.limit, 8, 23, f) == 23 && fread(x, 8, 36, f) == 36 && fread(P_prop, 8, 529, f) == 529;
    std::vector<double> sums((size_t)ntrips * PSTRIDE);
    ok = ok && fread(sums.data(), 8, sums.size(), f) == sums.size();
    fclose(f);
    if (!ok) { fprintf(stderr, "short input\n"); return 2; }
    static Ctl ctl;
    static SolvePre pre;
    static SolveSmem M;
    memset(&ctl, 0, sizeof(ctl));
    SolveParams S;
    memset(&S, 0, sizeof(S));
    S.job = &job; S.phase = SOLVE_PHASE_FULL; S.x = x; S.x_prop = x_prop; S.P_prop = P_prop; S.P_out = P_out; S.sum_out = sum_out;
    S.ctl = &ctl; S.pre = &pre;
    for (auto& b : g_warp) b = new std::barrier<>(32);
    for (auto& a : g_warp_and) a.store(1);
    FILE* o = fopen(argv[2], "wb");
    if (!o) { perror("open out"); return 2; }
    int executed = 0;

    auto cta = [&](int tid) {
        t_tid = tid;
        // ---- what k_ctl_reset does (solve.cu) ----
        for (int e = tid; e < 23 * 23; e += EMU_THREADS) P_out[e] = P_prop[e];
        if (tid < 36) { x_prop[tid] = x[tid]; M.x[tid] = x[tid]; M.xp[tid] = x[tid]; }
        if (tid == 0) { ctl.search = 1; ctl.stopped = 0; ctl.iter = -1; ctl_next_trip(&ctl); }
        __syncthreads();
        solve_pre(S, M, mode, tid, EMU_THREADS, false);
        // ---- the trips: what the solving CTA of k_fused does after it has summed the group records ----
        for (int t = 0; t < ntrips; ++t) {
            __syncthreads();
            if (ctl.stopped) break;
            __syncthreads();
            if (tid < NTERMS) M.sum[tid] = sums[(size_t)t * PSTRIDE + tid];
            __syncthreads();
            solve_trip(S, M, tid, EMU_THREADS);
            __syncthreads();
            if (tid == 0) {
                fwrite(x, 8, 36, o);
                fwrite(ctl.dx_last, 8, 23, o);
                int32_t c[8] = {ctl.search, ctl.stopped, ctl.iter, ctl.t, ctl.rematch, ctl.trips, ctl.converged, ctl.n_eff};
                fwrite(c, 4, 8, o);
                ++executed;
            }
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < EMU_THREADS; ++t) th.emplace_back(cta, t);
    for (auto& t : th) t.join();
    fwrite(P_out, 8, 529, o);
    fclose(o);
    printf("trips %d stopped %d nan_guard %d\n", executed, ctl.stopped, ctl.nan_guard);
    return 0;
}
