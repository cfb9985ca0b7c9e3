// stage.cu — the staging sort of a scan (sm_100a): a stable two-pass radix sort (8 + 8 bits) of the scan points by 16 bits of
// their cell's Morton code, fused with the record -> float4 conversion in front and the gather behind.  Three kernels:
//
//   k_stage_keys   records -> in4 (caller's order) + key16; tile histograms of the LOW digit; the last CTA to finish turns
//                  the digit-major histogram matrix into exclusive bases (one CTA, registers + one block scan per 8 K entries)
//   k_stage_pass<0> stable rank of every point inside its tile by the low digit -> position p after pass 1; order1[p] =
//                  caller index; the HIGH-digit histogram of the tile p falls into is counted with global atomics (the
//                  counts do not depend on the order the atomics arrive in); last CTA: exclusive bases again
//   k_stage_pass<1> the same ranking by the high digit over order1 -> final position q: perm[q] = caller index,
//                  out[q] = in4[caller index]; the flags of a fresh scan are written on the way
//
// Stable = ties keep the caller's order, so the result is a pure function of the scan (fixed-order sums), and
// it is the same permutation the library sort (cub::DeviceRadixSort::SortPairs on the same 16 bits) produced before.
// Ranking inside a tile: warp w owns R consecutive groups of 32 points; per group the lanes with the same digit find each
// other with match.any, the lowest of them advances the warp's counter of that digit in shared memory, everyone takes
// counter + (equal lanes below); a prefix over the warps' counters per digit then orders the warps.
#include "stage.cuh"

namespace flivo {

constexpr int STAGE_THREADS = 512;
constexpr int STAGE_WARPS = STAGE_THREADS / 32;

static inline unsigned int stage_tiles(unsigned int n_pad, int R) {
    const unsigned int tile = STAGE_THREADS * (unsigned int)R;
    return (n_pad + tile - 1) / tile;
}
static inline int stage_rounds(unsigned int n_pad) { return n_pad > (1u << 20) ? 16 : 4; }   // tile = 2048 or 8192 points
size_t stage_hist_bytes(size_t n_pad) {
    return 2 * 256 * (size_t)stage_tiles((unsigned int)n_pad, stage_rounds((unsigned int)n_pad)) * sizeof(unsigned int);
}

// Exclusive prefix sum over h[0..M) in place by ONE CTA of STAGE_THREADS threads, M a multiple of 16.  16 entries per
// thread and round as four 16-byte loads (one L2 round trip), block scan of the per-thread sums, running carry.
__device__ void scan_in_place(unsigned int* h, unsigned int M) {
    __shared__ unsigned int s_warp[STAGE_WARPS];
    __shared__ unsigned int s_carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (unsigned int base = 0; base < M; base += STAGE_THREADS * 16) {
        const unsigned int off = base + (unsigned int)tid * 16;
        uint4 v[4];
        unsigned int sum = 0;
        if (off < M) {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = __ldcg(reinterpret_cast<const uint4*>(h + off) + k);
#pragma unroll
            for (int k = 0; k < 4; ++k) sum += v[k].x + v[k].y + v[k].z + v[k].w;
        }
        unsigned int inc = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned int t = __shfl_up_sync(0xFFFFFFFFu, inc, d); if (lane >= d) inc += t; }
        if (lane == 31) s_warp[w] = inc;
        __syncthreads();
        if (w == 0) {
            const unsigned int t = lane < STAGE_WARPS ? s_warp[lane] : 0;
            unsigned int ti = t;
#pragma unroll
            for (int d = 1; d < STAGE_WARPS; d <<= 1) { const unsigned int u = __shfl_up_sync(0xFFFFFFFFu, ti, d); if (lane >= d) ti += u; }
            if (lane < STAGE_WARPS) s_warp[lane] = ti - t;             // exclusive over warps
        }
        __syncthreads();
        unsigned int run = s_carry + s_warp[w] + (inc - sum);
        if (off < M) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                uint4 o;
                o.x = run; run += v[k].x;
                o.y = run; run += v[k].y;
                o.z = run; run += v[k].z;
                o.w = run; run += v[k].w;
                reinterpret_cast<uint4*>(h + off)[k] = o;
            }
        }
        __syncthreads();
        if (tid == STAGE_THREADS - 1) s_carry = run;   // the last thread's running sum is the total so far (inactive threads add 0)
        __syncthreads();
    }
}

// true in exactly one CTA of the grid: the last one to get here.  Its threads then see everything the others wrote.
__device__ __forceinline__ bool last_cta(unsigned int* ticket) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(ticket, 1u);
        s_last = (t == gridDim.x - 1);
        if (s_last) *ticket = 0;   // ready for the next launch
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0;
}

template <int R>
__global__ void __launch_bounds__(STAGE_THREADS) k_stage_keys(const ScanJob* __restrict__ J, float4* __restrict__ in4, unsigned short* __restrict__ key16,
                                                              unsigned int* hist_lo, unsigned int* hist_hi, unsigned int tiles, unsigned int* ticket) {
    __shared__ unsigned int cnt[256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const unsigned int n = J->n, shift = J->key_shift;
    const float inv_edge = J->inv_edge;
    for (unsigned int e = blockIdx.x * STAGE_THREADS + tid; e < 256 * tiles; e += gridDim.x * STAGE_THREADS) hist_hi[e] = 0;
    if (tid < 256) cnt[tid] = 0;
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
    if (tid < 256) hist_lo[tid * tiles + blockIdx.x] = cnt[tid];
    if (last_cta(ticket)) scan_in_place(hist_lo, 256 * tiles);
}

// PASS 0: caller's order -> order after the low digit; PASS 1: that order -> final order, gather, flags
template <int R, int PASS>
__global__ void __launch_bounds__(STAGE_THREADS) k_stage_pass(const ScanJob* __restrict__ J, const unsigned short* __restrict__ key16, const unsigned int* base_tab,
                                                              unsigned int* order1, unsigned int* hist_hi, unsigned int tiles, unsigned int* ticket,
                                                              const float4* __restrict__ in4, float4* __restrict__ out, unsigned int* __restrict__ perm,
                                                              unsigned char* __restrict__ sel, unsigned char* __restrict__ eff, unsigned char* __restrict__ pflags) {
    __shared__ unsigned int cnt[STAGE_WARPS][256];
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
    __syncthreads();
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
        unsigned int run = __ldcg(base_tab + tid * tiles + blockIdx.x);
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
            atomicAdd(hist_hi + (key[r] >> 8) * tiles + pos / (STAGE_THREADS * R), 1u);
        } else {
            perm[pos] = item[r];
            out[pos] = in4[item[r]];
        }
    }
    if (PASS == 0) {
        if (last_cta(ticket)) scan_in_place(hist_hi, 256 * tiles);
    }
}

template <int R>
static void stage_sort_r(const ScanJob* J, unsigned int n_pad, float4* in4, float4* out, unsigned int* perm, unsigned char* sel, unsigned char* eff,
                         unsigned char* pflags, const StageWork& W, cudaStream_t s) {
    const unsigned int tiles = stage_tiles(n_pad, R);
    unsigned int* hist_lo = W.hist;
    unsigned int* hist_hi = W.hist + 256 * (size_t)tiles;
    k_stage_keys<R><<<tiles, STAGE_THREADS, 0, s>>>(J, in4, W.key16, hist_lo, hist_hi, tiles, W.ticket);
    k_stage_pass<R, 0><<<tiles, STAGE_THREADS, 0, s>>>(J, W.key16, hist_lo, W.order1, hist_hi, tiles, W.ticket + 1, in4, out, perm, sel, eff, pflags);
    k_stage_pass<R, 1><<<tiles, STAGE_THREADS, 0, s>>>(J, W.key16, hist_hi, W.order1, hist_hi, tiles, W.ticket + 1, in4, out, perm, sel, eff, pflags);
}

void launch_stage_sort(const ScanJob* J, unsigned int n_pad, float4* in4, float4* out, unsigned int* perm,
                       unsigned char* sel, unsigned char* eff, unsigned char* pflags, const StageWork& W, cudaStream_t stream) {
    if (stage_rounds(n_pad) == 16) stage_sort_r<16>(J, n_pad, in4, out, perm, sel, eff, pflags, W, stream);
    else stage_sort_r<4>(J, n_pad, in4, out, perm, sel, eff, pflags, W, stream);
}

}  // namespace flivo
