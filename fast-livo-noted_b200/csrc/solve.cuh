// solve.cuh — host launchers of all kernels (the parameter blocks live in fused.cuh).
#pragma once
#include "fused.cuh"

namespace flivo {

void upload_term_table();
void launch_fused(const FusedParams& P, const SolveParams& S, int grid, cudaStream_t stream);
void launch_search(const FusedParams& P, int search_grid, cudaStream_t stream, bool par_probes = false, bool pdl = false);
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
void launch_solve_ranks(const SolveParams& S, cudaStream_t stream);   // SOLVE_PHASE_FROM_RANKS: after the all-gather of the NCCL variant
// start of an update: control block, x_prop <- x, P_out <- P_prop, the first trip's state-only half (S.pre)
void launch_ctl_reset(const SolveParams& S, double* x_prop, SearchStats* stats, cudaStream_t stream);

}  // namespace flivo
