// solve.cuh — parameters of the single-CTA reduce + solve kernel (solve.cu) and host launchers of all kernels.
#pragma once
#include "fused.cuh"

namespace flivo {

enum {
    SOLVE_PHASE_FULL = 0,         // reduce the group partials, then solve (single GPU)
    SOLVE_PHASE_REDUCE_ONLY = 1,  // reduce only: sum_out[92] (flivo_h_share_model; first half of the multi-GPU trip)
    SOLVE_PHASE_FROM_RANKS = 2,   // sum rank_blocks[nranks][96] in rank order, then solve (second half, multi-GPU, NCCL)
    SOLVE_PHASE_PEER = 3          // multi-GPU, fused: reduce, write the record into every peer's exchange buffer over
                                  // NVLink, wait for the peers' records, sum in rank order, solve — one kernel
};

// Exchange buffer of one rank, mapped into every peer (cudaIpc): slots[parity][source rank].
constexpr int PEER_MAX_RANKS = 8;
struct PeerSlot {
    double rec[PSTRIDE];
    unsigned long long flag;      // sequence number of the exchange that filled rec
    unsigned long long pad[7];
};
struct PeerXchg {
    PeerSlot* base[PEER_MAX_RANKS];   // base[r]: rank r's buffer ([2][PEER_MAX_RANKS] slots) as mapped into this process
    unsigned long long* seq;          // this rank's exchange counter (device memory, never reset)
    int rank, nranks;
};

// State-only half of a trip's solve, computed by k_solve_pre on a parallel branch of the update graph while the
// search / fused kernels of the same trip run (it needs x and P_prop, not H^T H).
struct SolvePre {
    double Pt[23 * 23];   // V3: P_prop with the manifold Jacobians applied, T P T^T (esekfom.hpp:1665-1695)
    double dxnew[23];     // V3: J^{-1}-corrected x (-) x_prop (:1646-1686)
    double vec[18];       // V1: state_propagat - state (common_lib.h:576-587)
    double pad;
};

struct SolveParams {
    const double* gpartials;   // [groups * PSTRIDE]
    const Job* job;
    int phase;
    double* x;                 // device flivo_state, in/out
    const double* x_prop;      // device flivo_state
    const double* P_prop;      // n x n
    double* P_out;             // n x n, written when the loop stops
    double* sum_out;           // [96] reduced block (may be null in FULL phase)
    const double* rank_blocks; // [nranks * PSTRIDE]
    Ctl* ctl;
    SolvePre* pre;             // written by k_solve_pre, read by k_solve of the same trip
    PeerXchg peer;             // SOLVE_PHASE_PEER
    unsigned long long* stamps;       // profile mode only: [MAX_TRIPS][16] globaltimer values at the phase boundaries of k_solve
};

void upload_term_table();
void launch_fused(const FusedParams& P, int grid, cudaStream_t stream);
void launch_search(const FusedParams& P, int search_grid, cudaStream_t stream, bool par_probes = false);
int search_ctas_per_sm(bool ivox);
int search_tile_points();
void launch_knn(const MapView& map, const float4* q, int nq, int k, uint32_t max_sq_bits, float* out_xyz, float* out_sqd,
                int* out_id, int* out_count, SearchStats* stats, int grid, cudaStream_t stream);
void launch_transform(const float4* body, int n, const double* state, float4* world, cudaStream_t stream);
void launch_knn_ivox(const MapView& map, const float4* q, int nq, int k, double max_range_sq, float* out_xyz, float* out_sqd,
                     int* out_id, int* out_count, SearchStats* stats, cudaStream_t stream);
void launch_ivox_map_filter(const float4* body, int n, const double* state, const float4* near, const float2* near_meta,
                            const unsigned int* perm, double fs, float4* world_out, unsigned char* flag_add,
                            unsigned char* flag_noneed, cudaStream_t stream);
void launch_solve_pre(const SolveParams& S, cudaStream_t stream);   // before launch_solve of the same trip (FULL / FROM_RANKS)
void launch_solve(const SolveParams& S, cudaStream_t stream);
void launch_ctl_reset(const double* x, double* x_prop, const double* P_prop, double* P_out, Ctl* ctl, SearchStats* stats, cudaStream_t stream);

}  // namespace flivo
