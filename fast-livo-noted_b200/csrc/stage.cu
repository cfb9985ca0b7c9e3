// stage.cu — the staging sort of a scan (sm_100a): a stable two-pass radix sort (8 + 8 bits) of the scan points by 16 bits of
// their cell's Morton code, fused with the record -> float4 conversion in front and the gather behind.  Three kernels:
//
//   k_stage_keys   records -> in4 (caller's order) + key16; the tile's histogram of the LOW digit -> row `tile` of a
//                  [tiles][256] matrix
//   k_stage_pass<0> every CTA sums the matrix columns itself (points with a smaller digit anywhere + the same digit in the
//                  tiles before it: its bases — no CTA waits for another), ranks its points stably by the low digit ->
//                  position p after pass 1; order1[p] = caller index; the HIGH-digit histogram of the tile p falls into
//                  is counted with global atomics (the counts do not depend on the order the atomics arrive in)
//   k_stage_pass<1> the same ranking by the high digit over order1 -> final position q: perm[q] = caller index,
//                  out[q] = in4[caller index]; the flags of a fresh scan are written on the way
//
// Stable = ties keep the caller's order, so the result is a pure function of the scan (fixed-order sums), and
// it is the same permutation the library sort (cub::DeviceRadixSort::SortPairs on the same 16 bits) produced before.
// Ranking inside a tile: warp w owns R consecutive groups of 32 points; per group the lanes with the same digit find each
// other with match.any, the lowest of them advances the warp's counter of that digit in shared memory, everyone takes
// counter + (equal lanes below); a prefix over the warps' counters per digit then orders the warps.
#include "stage.cuh"
#include <stdlib.h>

namespace flivo {

constexpr int STAGE_THREADS = 512;
constexpr int STAGE_WARPS = STAGE_THREADS / 32;

static inline unsigned int stage_tiles(unsigned int n_pad, int R) {
    const unsigned int tile = STAGE_THREADS * (unsigned int)R;
    return (n_pad + tile - 1) / tile;
}
static inline int stage_rounds(unsigned int n_pad) {   // tile = 1024, 2048 or 8192 points
    static const int forced = getenv("FLIVO_STAGE_R") ? atoi(getenv("FLIVO_STAGE_R")) : 0;   // A/B only
    if (forced == 2 || forced == 4 || forced == 16) return forced;
    return n_pad > (1u << 20) ? 16 : (n_pad > 65536u ? 4 : 2);   // (24 k points: 0.2444 ms per step with 1024-point tiles, 0.2476 with 2048)
}
size_t stage_hist_bytes(size_t n_pad) {
    return 2 * 256 * (size_t)stage_tiles((unsigned int)n_pad, stage_rounds((unsigned int)n_pad)) * sizeof(unsigned int);
}

// Bases of one tile: s_base[d] = (points of the whole scan with a smaller digit) + (points with digit d in the tiles before
// this one), from the [tiles][256] histogram matrix.  Every CTA sums the columns itself — tiles x 256 coalesced loads, 64 per
// thread at 260 k points, all independent — instead of waiting for one CTA to scan the matrix: no serial tail, no tickets.
__device__ __forceinline__ void tile_bases(const unsigned int* __restrict__ hist, unsigned int tiles, unsigned int tile,
                                           unsigned int* s_base /* [256] */, unsigned int (*s_tmp)[2][256] /* [2][2][256] */) {
    const int tid = threadIdx.x, d = tid & 255, half = tid >> 8, lane = tid & 31, w = tid >> 5;
    unsigned int below = 0, total = 0;
#pragma unroll 8
    for (unsigned int t = half; t < tiles; t += 2) {
        const unsigned int v = __ldcg(hist + (size_t)t * 256 + d);
        total += v;
        below += t < tile ? v : 0u;
    }
    s_tmp[half][0][d] = below; s_tmp[half][1][d] = total;
    __syncthreads();
    __shared__ unsigned int s_wsum[8];
    unsigned int inc = 0;
    if (tid < 256) {
        below = s_tmp[0][0][d] + s_tmp[1][0][d];
        total = s_tmp[0][1][d] + s_tmp[1][1][d];
        inc = total;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) s_wsum[w] = inc;
    }
    __syncthreads();
    if (tid < 256) {
        unsigned int pre = 0;
#pragma unroll
        for (int ww = 0; ww < 8; ++ww) pre += ww < w ? s_wsum[ww] : 0u;
        s_base[d] = pre + (inc - total) + below;
    }
    __syncthreads();
}

template <int R>
__global__ void __launch_bounds__(STAGE_THREADS) k_stage_keys(const ScanJob* __restrict__ J, float4* __restrict__ in4, unsigned short* __restrict__ key16,
                                                              unsigned int* hist_lo, unsigned int* hist_hi, unsigned int tiles) {
    __shared__ unsigned int cnt[256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned int n = J->n, shift = J->key_shift;
    const float inv_edge = J->inv_edge;
    if (tid < 256) { cnt[tid] = 0; hist_hi[(size_t)blockIdx.x * 256 + tid] = 0; }   // (the next kernel counts into this row)
    __syncthreads();
    const unsigned int base = blockIdx.x * (STAGE_THREADS * R) + w * (32 * R) + lane;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const unsigned int i = base + r * 32;
        unsigned int dig = 0x100u | lane;
        if (i < n) {
            const float4 p = scan_record(J, i);
            in4[i] = p;
            const unsigned int key = (unsigned int)(scan_morton(p, inv_edge) >> shift) & 0xFFFFu;
            key16[i] = (unsigned short)key;
            dig = key & 255u;
        }
        const unsigned int peers = __match_any_sync(0xFFFFFFFFu, dig);
        if (i < n && lane == __ffs(peers) - 1) atomicAdd(&cnt[dig], (unsigned int)__popc(peers));
    }
    __syncthreads();
    if (tid < 256) hist_lo[(size_t)blockIdx.x * 256 + tid] = cnt[tid];
}

// PASS 0: caller's order -> order after the low digit; PASS 1: that order -> final order, gather, flags
template <int R, int PASS>
__global__ void __launch_bounds__(STAGE_THREADS) k_stage_pass(const ScanJob* __restrict__ J, const unsigned short* __restrict__ key16, const unsigned int* hist,
                                                              unsigned int* order1, unsigned int* hist_hi, unsigned int tiles,
                                                              const float4* __restrict__ in4, float4* __restrict__ out, unsigned int* __restrict__ perm,
                                                              unsigned char* __restrict__ sel, unsigned char* __restrict__ eff, unsigned char* __restrict__ pflags) {
    __shared__ unsigned int cnt[STAGE_WARPS][256];
    __shared__ unsigned int s_base[256];
    __shared__ unsigned int s_tmp[2][2][256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned int n = J->n;
    for (int e = tid; e < STAGE_WARPS * 256; e += STAGE_THREADS) (&cnt[0][0])[e] = 0;
    const unsigned int base = blockIdx.x * (STAGE_THREADS * R) + w * (32 * R) + lane;
    unsigned int item[R], key[R], rank[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const unsigned int p = base + r * 32;
        item[r] = 0; key[r] = 0;
        if (p < n) {
            item[r] = PASS ? __ldcg(order1 + p) : p;
            if (PASS) { sel[p] = 1; eff[p] = 0; pflags[p] = 0; }   // uniform values: written at p, not at the scattered q
        }
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
        if (base + r * 32 < n) key[r] = key16[item[r]];
    tile_bases(hist, tiles, blockIdx.x, s_base, s_tmp);   // (its loads travel together with the key loads above; ends with a barrier)
    const unsigned int lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const bool active = base + r * 32 < n;
        const unsigned int dig = active ? (PASS ? key[r] >> 8 : key[r] & 255u) : (0x100u | lane);
        const unsigned int peers = __match_any_sync(0xFFFFFFFFu, dig);
        const int leader = __ffs(peers) - 1;
        unsigned int c = 0;
        if (active && lane == leader) { c = cnt[w][dig]; cnt[w][dig] = c + __popc(peers); }
        c = __shfl_sync(0xFFFFFFFFu, c, leader);
        rank[r] = c + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    if (tid < 256) {   // per digit: the tile's base in the whole scan, then the warps in order
        unsigned int run = s_base[tid];
#pragma unroll
        for (int ww = 0; ww < STAGE_WARPS; ++ww) { const unsigned int c = cnt[ww][tid]; cnt[ww][tid] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (base + r * 32 >= n) continue;
        const unsigned int dig = PASS ? key[r] >> 8 : key[r] & 255u;
        const unsigned int pos = cnt[w][dig] + rank[r];
        if (PASS == 0) {
            order1[pos] = item[r];
            atomicAdd(hist_hi + (size_t)(pos / (STAGE_THREADS * R)) * 256 + (key[r] >> 8), 1u);
        } else {
            perm[pos] = item[r];
            out[pos] = in4[item[r]];
        }
    }
}

template <int R>
static void stage_sort_r(const ScanJob* J, unsigned int n_pad, float4* in4, float4* out, unsigned int* perm, unsigned char* sel, unsigned char* eff,
                         unsigned char* pflags, const StageWork& W, cudaStream_t s) {
    const unsigned int tiles = stage_tiles(n_pad, R);
    unsigned int* hist_lo = W.hist;
    unsigned int* hist_hi = W.hist + 256 * (size_t)tiles;
    k_stage_keys<R><<<tiles, STAGE_THREADS, 0, s>>>(J, in4, W.key16, hist_lo, hist_hi, tiles);
    k_stage_pass<R, 0><<<tiles, STAGE_THREADS, 0, s>>>(J, W.key16, hist_lo, W.order1, hist_hi, tiles, in4, out, perm, sel, eff, pflags);
    k_stage_pass<R, 1><<<tiles, STAGE_THREADS, 0, s>>>(J, W.key16, hist_hi, W.order1, hist_hi, tiles, in4, out, perm, sel, eff, pflags);
}

void launch_stage_sort(const ScanJob* J, unsigned int n_pad, float4* in4, float4* out, unsigned int* perm,
                       unsigned char* sel, unsigned char* eff, unsigned char* pflags, const StageWork& W, cudaStream_t stream) {
    const int R = stage_rounds(n_pad);
    if (R == 16) stage_sort_r<16>(J, n_pad, in4, out, perm, sel, eff, pflags, W, stream);
    else if (R == 2) stage_sort_r<2>(J, n_pad, in4, out, perm, sel, eff, pflags, W, stream);
    else stage_sort_r<4>(J, n_pad, in4, out, perm, sel, eff, pflags, W, stream);
}

}  // namespace flivo
