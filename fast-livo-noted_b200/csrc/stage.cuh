// stage.cuh — scan staging: the per-scan parameter block, the sort key, and the launcher of the staging sort (stage.cu).
//
// A sorted scan is processed in the order of 16 bits of its points' cell Morton code (lidar frame), so that the 128
// points of a search tile query neighbouring map cells.  The order is a STABLE sort by those 16 bits: equal keys keep the
// caller's order, which makes the processing order — and with it every fixed-order sum downstream — a pure function of
// the scan.  (No counterpart in the reference: its loop walks feats_down_body in PCL's voxel-grid order,
// src/laser_mapping.cpp:499-503.)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace flivo {

// Per-scan scalars of the staging sequence, resident in HBM so that the captured graph of the sequence never changes its
// kernel parameters: a call rewrites this block through the graph's own H2D copy node.
struct ScanJob {
    const char* src;      // the scan in the caller's order: n records of `stride` bytes, x y z in the first 12
    unsigned int stride;
    unsigned int n;       // scan points
    unsigned int n_pad;   // n rounded up to SCAN_PAD: the item count the staging graph is built for
    float inv_edge;       // 1 / search cell edge
    unsigned int key_shift;   // the sort key is bits [key_shift, key_shift + 16) of the 63-bit Morton code
    // a scan may arrive in two pieces (flivo_scan_upload_async with few host threads: the head repacked to float4 by the host,
    // the tail as the caller's raw records by DMA): records [0, n_a) at src / stride, records [n_a, n) at src_b / stride_b
    unsigned int n_a;         // = n for a scan in one piece
    unsigned int stride_b;
    const char* src_b;
};
constexpr unsigned int SCAN_PAD = 8192;

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long spread21(unsigned int v) {
    unsigned long long x = v & 0x1FFFFF;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
// Morton code of the point's cell in the lidar frame (a rigid transform preserves locality, so points that are
// neighbours here query neighbouring map cells whatever the state is)
__device__ __forceinline__ unsigned long long scan_morton(float4 p, float inv_edge) {
    const float fx = fminf(fmaxf(floorf(p.x * inv_edge), -1048575.f), 1048575.f);
    const float fy = fminf(fmaxf(floorf(p.y * inv_edge), -1048575.f), 1048575.f);
    const float fz = fminf(fmaxf(floorf(p.z * inv_edge), -1048575.f), 1048575.f);
    const unsigned int ux = (unsigned int)((int)fx + (1 << 20)), uy = (unsigned int)((int)fy + (1 << 20)), uz = (unsigned int)((int)fz + (1 << 20));
    return (spread21(ux) << 2) | (spread21(uy) << 1) | spread21(uz);
}
__device__ __forceinline__ float4 scan_record(const ScanJob* J, size_t i) {
    const char* src = J->src;
    unsigned int stride = J->stride;
    if (i >= J->n_a) { src = J->src_b; stride = J->stride_b; i -= J->n_a; }
    if (stride == 16) return reinterpret_cast<const float4*>(src)[i];
    const float* r = reinterpret_cast<const float*>(src + i * (size_t)stride);
    return make_float4(r[0], r[1], r[2], 0.f);
}
#endif

// Work buffers of the staging sort, owned by the context (sized for its scan capacity `cap`, a multiple of SCAN_PAD).
struct StageWork {
    unsigned short* key16;   // [cap] sort key per point, caller's order
    unsigned int* order1;    // [cap] caller index of the points after the first pass (sorted by the low digit)
    unsigned int* hist;      // [2][tiles][256] tile histograms of the low and of the high digit
};
size_t stage_hist_bytes(size_t n_pad);   // bytes `hist` needs for a scan of n_pad items

// records -> float4 in the caller's order (in4) + the processing order (perm[q] = caller index), the scan gathered into
// that order (out) and the per-point flags of a fresh scan (sel = 1, eff = 0, pflags = 0).  Three kernels, no library call.
// Everything that depends on the scan (source pointer, stride, n) is read from *J on the device; the launch geometry
// depends on n_pad only, so a captured graph of this sequence serves every scan of the same padded size.
void launch_stage_sort(const ScanJob* J, unsigned int n_pad, float4* in4, float4* out, unsigned int* perm,
                       unsigned char* sel, unsigned char* eff, unsigned char* pflags, const StageWork& W, cudaStream_t stream);

}  // namespace flivo
