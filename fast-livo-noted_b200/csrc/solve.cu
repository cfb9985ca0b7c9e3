// solve.cu — the two stand-alone kernels around solve_dev.cuh.
//
// The solve of a trip normally runs inside k_fused (fused.cu: the CTA that folds the last group record calls
// solve_trip), so a single-GPU or peer-memory update has no solve kernel at all.  What is left here:
//   k_solve_ranks : second half of a trip in the NCCL variant of the sharded update (after ncclAllGather of the ranks'
//                   92-double records): sum in rank order + the same solve_trip;
//   k_ctl_reset   : start of an update: control block, x_prop <- x, P_out <- P_prop, and the state-only half of the first
//                   trip's solve (solve_pre), which the first k_search overlaps (programmatic dependent launch).
#include "solve_dev.cuh"

namespace flivo {

constexpr int SOLVE_THREADS = 128;

__global__ void __launch_bounds__(SOLVE_THREADS) k_solve_ranks(const __grid_constant__ SolveParams S) {
    __shared__ SolveSmem M;
    if (S.ctl->stopped) return;
    solve_trip(S, M, threadIdx.x, SOLVE_THREADS);
}
void launch_solve_ranks(const SolveParams& S, cudaStream_t stream) { k_solve_ranks<<<1, SOLVE_THREADS, 0, stream>>>(S); }

__global__ void __launch_bounds__(SOLVE_THREADS) k_ctl_reset(const __grid_constant__ SolveParams S, double* x_prop, SearchStats* stats) {
    __shared__ SolveSmem M;
    Ctl* ctl = S.ctl;
    const int tid = threadIdx.x;
    const int mode = S.job->mode;
    for (int e = tid; e < 23 * 23; e += SOLVE_THREADS) S.P_out[e] = S.P_prop[e];   // a loop that never stops returns P unchanged
    if (tid < 36) { const double v = S.x[tid]; x_prop[tid] = v; M.x[tid] = v; M.xp[tid] = v; }
    if (tid == 0) {
        ctl->search = 1; ctl->stopped = 0; ctl->iter = -1; ctl->t = 0; ctl->rematch = 0; ctl->trips = 0;
        ctl->searches = 0; ctl->converged = 0; ctl->n_eff = 0; ctl->nan_guard = 0; ctl->res_mean = 0.0;
        ctl_next_trip(ctl);
        for (int i = 0; i < MAX_TRIPS; ++i) ctl->trip_searched[i] = 0;
        for (int i = 0; i < 23; ++i) ctl->dx_last[i] = 0.0;
        if (stats) { stats->far_queries = 0; stats->candidates = 0; stats->sweeps = 0; }
    }
    // the control block is armed: the first trip's k_search (a programmatic dependent of this grid) may start beside the
    // state-only half of the first solve
    solve_publish(ctl, S.job, tid);
    solve_pre(S, M, mode, tid, SOLVE_THREADS, false);
}
void launch_ctl_reset(const SolveParams& S, double* x_prop, SearchStats* stats, cudaStream_t stream) {
    k_ctl_reset<<<1, SOLVE_THREADS, 0, stream>>>(S, x_prop, stats);
}

}  // namespace flivo
