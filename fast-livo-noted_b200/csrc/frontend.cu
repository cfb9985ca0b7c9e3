// frontend.cu — the two steps right before the measurement update, device resident:
//   k_undistort : per-point motion compensation, the back-propagation loop of ImuProcess::UndistortPcl
//                 (src/IMU_Processing.cpp:343-376)
//   voxel grid  : pcl::VoxelGrid::filter as the reference calls it (src/laser_mapping.cpp:129-130, leaf size set at :1111):
//                 centroid of the points of every occupied leaf, output ordered by leaf index.
// PCL is not in the reference tree; the semantics restated here are those of pcl/filters/impl/voxel_grid.hpp
// (applyFilter, downsample_all_data = true, min_points_per_voxel = 0) with float accumulators
// (pcl::CentroidPoint: AccumulatorXYZ, AccumulatorIntensity): "parity unpinned", the oracle defines the bits
// (oracle/flivo_oracle.cpp: orc_voxel_grid), with points of one leaf added in input order.
#include "frontend.cuh"

#include <cub/cub.cuh>

namespace flivo {

// ---- undistortion ------------------------------------------------------------------------------------
// Exp(ang_vel, dt) of include/so3_math.h:30-52 (Rodrigues, identity below 1e-7 rad/s), row-major
__device__ __forceinline__ void d_exp_rate(const double w[3], double dt, double R[9]) {
    const double nrm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    R[0] = 1.0; R[1] = 0.0; R[2] = 0.0; R[3] = 0.0; R[4] = 1.0; R[5] = 0.0; R[6] = 0.0; R[7] = 0.0; R[8] = 1.0;
    if (nrm > 0.0000001) {
        const double ax = w[0] / nrm, ay = w[1] / nrm, az = w[2] / nrm;
        const double K[9] = {0.0, -az, ay, az, 0.0, -ax, -ay, ax, 0.0};
        const double ang = nrm * dt;
        const double s = sin(ang), c1 = 1.0 - cos(ang);
        double cK[9], KK[9];   // Eye3 + sin * K + ((1 - cos) * K) * K, evaluated in the reference's order
#pragma unroll
        for (int e = 0; e < 9; ++e) cK[e] = c1 * K[e];
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) KK[i * 3 + j] = cK[i * 3] * K[j] + cK[i * 3 + 1] * K[3 + j] + cK[i * 3 + 2] * K[6 + j];
#pragma unroll
        for (int e = 0; e < 9; ++e) R[e] = (R[e] + s * K[e]) + KK[e];
    }
}
__device__ __forceinline__ void d_mv3(const double* M, const double v[3], double o[3]) {
    o[0] = M[0] * v[0] + M[1] * v[1] + M[2] * v[2];
    o[1] = M[3] * v[0] + M[4] * v[1] + M[5] * v[2];
    o[2] = M[6] * v[0] + M[7] * v[1] + M[8] * v[2];
}

// One thread per raw point.  The reference walks the scan backwards (points sorted by time offset) through the
// IMU pose segments and compensates a point with the LAST pose whose offset_time is < the point's time
// (strictly: `curvature/1000 > head->offset_time`, :355); the first pose has offset 0, so points at time 0 stay
// as they are, and points later than the last pose are extrapolated from the last segment's head.
__global__ void __launch_bounds__(256) k_undistort(float4* __restrict__ pts, const float* __restrict__ t_ms, size_t n, UndistortParams U) {
    __shared__ double s_off[UNDISTORT_MAX_POSES];
    for (int k = threadIdx.x; k < U.n_poses; k += blockDim.x) s_off[k] = U.poses[k].offset_time;
    __syncthreads();
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double t = (double)t_ms[i] / double(1000);
    // head = largest k <= n_poses-2 with offset_time[k] < t
    int lo = -1;
    {
        int a = 0, b = U.n_poses - 2;   // candidates
        while (a <= b) {
            const int m = (a + b) >> 1;
            if (s_off[m] < t) { lo = m; a = m + 1; } else b = m - 1;
        }
    }
    if (lo < 0) return;   // t <= offset of the first pose: never reached by the reference's loop
    const PoseDev& h = U.poses[lo];
    const double dt = t - h.offset_time;
    double E[9], Ri[9];
    d_exp_rate(h.gyr, dt, E);
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) Ri[r * 3 + c] = h.rot[r * 3] * E[c] + h.rot[r * 3 + 1] * E[3 + c] + h.rot[r * 3 + 2] * E[6 + c];
    double Tei[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) Tei[a] = ((h.pos[a] + h.vel[a] * dt) + 0.5 * h.acc[a] * dt * dt) - U.pos_end[a];
    float4 p = pts[i];
    const double Pi[3] = {(double)p.x, (double)p.y, (double)p.z};
    double a3[3], b3[3], c3[3];
    d_mv3(U.R_LI, Pi, a3);                                   // Lid_rot_to_IMU * P_i + Lid_offset_to_IMU
#pragma unroll
    for (int a = 0; a < 3; ++a) a3[a] += U.t_LI[a];
    d_mv3(Ri, a3, b3);                                       // R_i * (...) + T_ei
#pragma unroll
    for (int a = 0; a < 3; ++a) b3[a] += Tei[a];
    d_mv3(U.extR_Ri, b3, c3);                                // extR_Ri * (...) - exrR_extT
    p.x = (float)(c3[0] - U.exrR_extT[0]);
    p.y = (float)(c3[1] - U.exrR_extT[1]);
    p.z = (float)(c3[2] - U.exrR_extT[2]);
    pts[i] = p;
}
void launch_undistort(float4* pts, const float* t_ms, size_t n, const UndistortParams& U, cudaStream_t stream) {
    if (n == 0) return;
    k_undistort<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(pts, t_ms, n, U);
}

// ---- voxel grid ----------------------------------------------------------------------------------------
// order-preserving float <-> uint for atomicMin/Max
__device__ __forceinline__ unsigned f2ord(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float ord2f(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u); }

__global__ void k_vg_reset(VoxelGridState* st) {
    if (threadIdx.x == 0) {
        for (int a = 0; a < 3; ++a) { st->min_ord[a] = 0xFFFFFFFFu; st->max_ord[a] = 0u; }
        st->n_finite = 0; st->overflow = 0; st->n_out = 0;
    }
}
// getMinMax3D over the finite points
__global__ void __launch_bounds__(256) k_vg_minmax(const float4* __restrict__ pts, size_t n, VoxelGridState* st) {
    unsigned mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
    unsigned cnt = 0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float4 p = pts[i];
        if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
            const unsigned o[3] = {f2ord(p.x), f2ord(p.y), f2ord(p.z)};
#pragma unroll
            for (int a = 0; a < 3; ++a) { mn[a] = min(mn[a], o[a]); mx[a] = max(mx[a], o[a]); }
            ++cnt;
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        mn[a] = __reduce_min_sync(0xFFFFFFFFu, mn[a]);
        mx[a] = __reduce_max_sync(0xFFFFFFFFu, mx[a]);
    }
    cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomicMin(&st->min_ord[a], mn[a]); atomicMax(&st->max_ord[a], mx[a]); }
        if (cnt) atomicAdd(&st->n_finite, cnt);
    }
}
// grid parameters (voxel_grid.hpp applyFilter): min_b = floor(min_p * inv_leaf), div_b = max_b - min_b + 1,
// divb_mul = (1, div_b0, div_b0*div_b1); "leaf size too small" when dx*dy*dz exceeds INT_MAX
__global__ void k_vg_params(VoxelGridState* st, float inv_leaf) {
    if (threadIdx.x != 0) return;
    if (st->n_finite == 0) { st->overflow = 0; for (int a = 0; a < 3; ++a) { st->min_b[a] = 0; st->div_b[a] = 1; } st->mul[0] = 1; st->mul[1] = 1; st->mul[2] = 1; return; }
    long long d[3];
    for (int a = 0; a < 3; ++a) {
        const float mn = ord2f(st->min_ord[a]), mx = ord2f(st->max_ord[a]);
        d[a] = (long long)(__fmul_rn(__fsub_rn(mx, mn), inv_leaf)) + 1;
        st->min_b[a] = (int)floorf(__fmul_rn(mn, inv_leaf));
        const int max_b = (int)floorf(__fmul_rn(mx, inv_leaf));
        st->div_b[a] = max_b - st->min_b[a] + 1;
    }
    st->overflow = (d[0] * d[1] * d[2] > 2147483647ll) ? 1 : 0;
    st->mul[0] = 1; st->mul[1] = st->div_b[0]; st->mul[2] = st->div_b[0] * st->div_b[1];
}
// leaf index of every point; non-finite points get the largest key and drop off the end of the sorted list
__global__ void __launch_bounds__(256) k_vg_keys(const float4* __restrict__ pts, size_t n, const VoxelGridState* __restrict__ st, float inv_leaf,
                                                 unsigned* __restrict__ keys, unsigned* __restrict__ idx) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 p = pts[i];
    unsigned key = 0xFFFFFFFFu;
    if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
        const int i0 = (int)(__fsub_rn(floorf(__fmul_rn(p.x, inv_leaf)), (float)st->min_b[0]));
        const int i1 = (int)(__fsub_rn(floorf(__fmul_rn(p.y, inv_leaf)), (float)st->min_b[1]));
        const int i2 = (int)(__fsub_rn(floorf(__fmul_rn(p.z, inv_leaf)), (float)st->min_b[2]));
        key = (unsigned)(i0 * st->mul[0] + i1 * st->mul[1] + i2 * st->mul[2]);
    }
    keys[i] = key;
    idx[i] = (unsigned)i;
}
// 1 at the first entry of every leaf (sorted order), 0 elsewhere and on the non-finite tail
__global__ void __launch_bounds__(256) k_vg_heads(const unsigned* __restrict__ keys, size_t n, const VoxelGridState* __restrict__ st, unsigned* __restrict__ head) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const size_t nf = st->n_finite;
    head[i] = (i < nf && (i == 0 || keys[i] != keys[i - 1])) ? 1u : 0u;
}
// centroid of one leaf per thread: float sums in input order (the radix sort is stable), then / n
__global__ void __launch_bounds__(128) k_vg_centroids(const float4* __restrict__ pts, const float* __restrict__ t_ms, const unsigned* __restrict__ keys,
                                                      const unsigned* __restrict__ idx, const unsigned* __restrict__ head, const unsigned* __restrict__ pos,
                                                      size_t n, VoxelGridState* st, float4* __restrict__ out, float* __restrict__ out_t) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n || !head[i]) return;
    const size_t nf = st->n_finite;
    const unsigned key = keys[i];
    float sx = 0.f, sy = 0.f, sz = 0.f, si = 0.f, st_ = 0.f;
    unsigned cnt = 0;
    for (size_t j = i; j < nf && keys[j] == key; ++j) {
        const unsigned src = idx[j];
        const float4 p = pts[src];
        sx = __fadd_rn(sx, p.x); sy = __fadd_rn(sy, p.y); sz = __fadd_rn(sz, p.z); si = __fadd_rn(si, p.w);
        if (t_ms) st_ = __fadd_rn(st_, t_ms[src]);
        ++cnt;
    }
    const float fn = (float)cnt;
    const unsigned o = pos[i];
    out[o] = make_float4(__fdiv_rn(sx, fn), __fdiv_rn(sy, fn), __fdiv_rn(sz, fn), __fdiv_rn(si, fn));
    if (out_t) out_t[o] = __fdiv_rn(st_, fn);
    if (i + cnt >= nf) st->n_out = o + 1;   // the last leaf
}
// "Leaf size is too small for the input dataset": PCL warns and returns the input cloud unchanged
__global__ void __launch_bounds__(256) k_vg_passthrough(const float4* __restrict__ pts, const float* __restrict__ t_ms, size_t n, VoxelGridState* st,
                                                        float4* __restrict__ out, float* __restrict__ out_t) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i == 0) st->n_out = (unsigned)n;
    if (i >= n) return;
    out[i] = pts[i];
    if (out_t) out_t[i] = t_ms ? t_ms[i] : 0.f;
}

size_t voxel_grid_temp_bytes(size_t cap) {
    size_t a = 0, b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, a, (unsigned*)nullptr, (unsigned*)nullptr, (unsigned*)nullptr, (unsigned*)nullptr, (int)cap, 0, 32, (cudaStream_t)0);
    cub::DeviceScan::ExclusiveSum(nullptr, b, (unsigned*)nullptr, (unsigned*)nullptr, (int)cap, (cudaStream_t)0);
    return (a > b ? a : b) + 256;
}

// Enqueues the whole filter; the caller reads st->n_out / st->overflow after synchronising.
// Scratch: keys_a/b, idx_a/b, head, pos: [n] u32 each; tmp: voxel_grid_temp_bytes(n).
cudaError_t launch_voxel_grid(const float4* pts, const float* t_ms, size_t n, float leaf, VoxelGridState* st, unsigned* keys_a, unsigned* keys_b,
                              unsigned* idx_a, unsigned* idx_b, unsigned* head, unsigned* pos, void* tmp, size_t tmp_bytes, float4* out, float* out_t,
                              cudaStream_t stream) {
    const float inv_leaf = 1.0f / leaf;   // inverse_leaf_size_ = Array4f::Ones() / leaf_size_
    k_vg_reset<<<1, 32, 0, stream>>>(st);
    if (n == 0) return cudaGetLastError();
    const unsigned nb = (unsigned)((n + 255) / 256);
    k_vg_minmax<<<min(nb, 148u * 8u), 256, 0, stream>>>(pts, n, st);
    k_vg_params<<<1, 32, 0, stream>>>(st, inv_leaf);
    k_vg_keys<<<nb, 256, 0, stream>>>(pts, n, st, inv_leaf, keys_a, idx_a);
    size_t tb = tmp_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tb, keys_a, keys_b, idx_a, idx_b, (int)n, 0, 32, stream);
    if (e != cudaSuccess) return e;
    k_vg_heads<<<nb, 256, 0, stream>>>(keys_b, n, st, head);
    tb = tmp_bytes;
    e = cub::DeviceScan::ExclusiveSum(tmp, tb, head, pos, (int)n, stream);
    if (e != cudaSuccess) return e;
    k_vg_centroids<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(pts, t_ms, keys_b, idx_b, head, pos, n, st, out, out_t);
    return cudaGetLastError();
}
cudaError_t launch_voxel_passthrough(const float4* pts, const float* t_ms, size_t n, VoxelGridState* st, float4* out, float* out_t, cudaStream_t stream) {
    k_vg_passthrough<<<(unsigned)((n + 255) / 256) + 1, 256, 0, stream>>>(pts, t_ms, n, st, out, out_t);
    return cudaGetLastError();
}

}  // namespace flivo
