// search.cuh — warp-per-query exact k=5 nearest-neighbour search over the voxel-hash map.
//
// Semantics reproduced: KD_TREE::Nearest_Search / Search (include/ikd-Tree/ikd_Tree.cpp:350-380, :843-986):
// the k nearest live points whose float32 squared distance (calc_dist, :1291-1295) is <= max_dist,
// ascending.  Equal distances at rank k are kept by visiting order in the reference (tree-shape
// dependent); here the canonical total order (distance bits, insertion id) decides, see DESIGN.md.
//
// Exactness argument: cells are unions of voxels whose integer keys come from the same floorf(x/ds)
// the insert path uses, so a point outside the cell block [cq-n, cq+n]^3 is farther than
// n*cell_edge minus a rounding slack from the query; a shell is final when the 5th best distance is
// below that bound (or the bound exceeds max_dist).  Shells: 3^3 cells (fast path, one warp pass),
// 5^3 cells, then rings of 4x4x4-cell blocks through the block directory, then (unbounded searches
// only) a sweep of the whole directory pruned by block distance.
#pragma once
#include "flivo_dev.cuh"

namespace flivo {

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int FAR_MAX_RING = 3;  // block rings tried before the directory sweep

struct Cand {
    uint32_t d;     // squared distance bits (positive floats order as unsigned); INF_BITS = empty
    uint32_t id;    // insertion id (unique)
    uint32_t addr;  // point slot = bucket * 8 + s
};
__device__ __forceinline__ bool cand_less(const Cand& a, const Cand& b) { return a.d < b.d || (a.d == b.d && a.id < b.id); }

// Warp-wide selection state.  c0 carries the current best five (lane r < 5 holds rank r); c1..c3 are the
// staging slots filled by push(); all slot indices are compile-time so everything stays in registers.
struct WarpSelect {
    Cand c0, c1, c2, c3;
    int fill;        // warp-uniform number of staged batches (0..3)
    int count;       // warp-uniform: how many of the best five exist
    uint32_t tau;    // warp-uniform: distance bits of the current 5th best, or the max_dist bound
    uint32_t bound;  // max_dist as float bits (candidates above it are never staged)

    __device__ __forceinline__ void init(uint32_t bound_bits) {
        c0.d = c1.d = c2.d = c3.d = INF_BITS;
        c0.id = c1.id = c2.id = c3.id = 0xFFFFFFFFu;
        c0.addr = c1.addr = c2.addr = c3.addr = 0;
        fill = 0; count = 0; bound = bound_bits; tau = bound_bits;
    }
    // Extract the five smallest of the 4 x 32 staged candidates into c0 of lanes 0..4.
    __device__ __forceinline__ void select(int lane) {
        Cand res; res.d = INF_BITS; res.id = 0xFFFFFFFFu; res.addr = 0;
        int found = 0;
#pragma unroll
        for (int r = 0; r < FLIVO_K_; ++r) {
            Cand lb = c0; int which = 0;
            if (cand_less(c1, lb)) { lb = c1; which = 1; }
            if (cand_less(c2, lb)) { lb = c2; which = 2; }
            if (cand_less(c3, lb)) { lb = c3; which = 3; }
            uint32_t dmin = __reduce_min_sync(FULL, lb.d);
            if (dmin == INF_BITS) break;  // warp-uniform
            uint32_t idmin = __reduce_min_sync(FULL, lb.d == dmin ? lb.id : 0xFFFFFFFFu);
            bool win = (lb.d == dmin) && (lb.id == idmin);
            int wl = __ffs(__ballot_sync(FULL, win)) - 1;
            uint32_t a = __shfl_sync(FULL, lb.addr, wl);
            if (win) {
                if (which == 0) c0.d = INF_BITS; else if (which == 1) c1.d = INF_BITS;
                else if (which == 2) c2.d = INF_BITS; else c3.d = INF_BITS;
            }
            if (lane == r) { res.d = dmin; res.id = idmin; res.addr = a; }
            ++found;
        }
        c0 = res;
        c1.d = c2.d = c3.d = INF_BITS;
        fill = 0;
        count = found;
        uint32_t d5 = __shfl_sync(FULL, c0.d, FLIVO_K_ - 1);
        tau = (found == FLIVO_K_) ? d5 : bound;
    }
    // Stage one candidate per lane (d == INF_BITS for "none").  Warp-collective.
    __device__ __forceinline__ void push(int lane, uint32_t d, uint32_t id, uint32_t addr) {
        if (fill == 0) { c1.d = d; c1.id = id; c1.addr = addr; }
        else if (fill == 1) { c2.d = d; c2.id = id; c2.addr = addr; }
        else { c3.d = d; c3.id = id; c3.addr = addr; }
        if (++fill == 3) select(lane);
    }
    __device__ __forceinline__ void flush(int lane) { if (fill) select(lane); }
    static constexpr int FLIVO_K_ = 5;
};

// squared covered radius (float, conservative) of a cell block reaching n cells beyond the query's cell
__device__ __forceinline__ float covered_sq(float cell_edge, float n_cells, float qabs) {
    double r = (double)n_cells * (double)cell_edge * (1.0 - 1e-6) - ((double)qabs + (double)n_cells * (double)cell_edge) * 4.8e-7;
    if (r <= 0.0) return 0.f;
    return (float)(r * r * (1.0 - 1e-6));
}

// Scan the 8 slots of up to four buckets per step (8 lanes per bucket) and stage the candidates.
// `b` holds one bucket index per lane (NONE32 = nothing); processes the buckets of all lanes whose bit is
// set in `found`, following overflow chains.
__device__ __forceinline__ void scan_found_buckets(const MapView& m, WarpSelect& ws, int lane, unsigned found, uint32_t b,
                                                   float qx, float qy, float qz, unsigned long long& ncand) {
    int nf = __popc(found);
    for (int base = 0; base < nf; base += 4) {
        int my = base + (lane >> 3);
        bool has = my < nf;
        unsigned src = has ? __fns(found, 0, my + 1) : 0u;
        uint32_t bb = __shfl_sync(FULL, b, src & 31);
        if (!has) bb = NONE32;
        while (__any_sync(FULL, bb != NONE32)) {
            uint32_t d = INF_BITS, id = 0xFFFFFFFFu, addr = 0;
            uint32_t nxt = NONE32;
            if (bb != NONE32) {
                addr = bb * BUCKET_SLOTS + (lane & 7);
                float4 p = __ldg(m.buckets + addr);
                int pid = __float_as_int(p.w);
                if (pid >= 0) {
                    float dd = sqdist3(qx, qy, qz, p.x, p.y, p.z);
                    uint32_t db = __float_as_uint(dd);
                    if (db <= ws.tau) { d = db; id = (uint32_t)pid; }
                }
                nxt = __ldg(m.bucket_next + bb);
            }
            ncand += __popc(__ballot_sync(FULL, bb != NONE32));
            ws.push(lane, d, id, addr);
            bb = nxt;
        }
    }
}

// Exact 5-NN of (qx,qy,qz).  On return lanes 0..4 hold rank 0..4 in out (d == INF_BITS when absent);
// returns the number found (warp-uniform).  max_sq_bits: float bits of the largest float <= max_dist
// (INF_BITS for unbounded).
__device__ __forceinline__ int warp_knn5(const MapView& m, float qx, float qy, float qz, uint32_t max_sq_bits, int lane,
                                         Cand& out, SearchStats* stats) {
    WarpSelect ws;
    ws.init(max_sq_bits);
    unsigned long long ncand = 0;
    const int g = m.gshift;
    const int cqx = voxel_key1(qx, m.ds) >> g, cqy = voxel_key1(qy, m.ds) >> g, cqz = voxel_key1(qz, m.ds) >> g;
    const float qabs = fmaxf(fabsf(qx), fmaxf(fabsf(qy), fabsf(qz)));
    const float max_sq = __uint_as_float(max_sq_bits);
    bool far_counted = false;

    // ---- shell 1: 3x3x3 cells; the (at most) 8 distinct blocks are probed by lanes 0..7 -------------
    {
        int bx0 = (cqx - 1) >> 2, by0 = (cqy - 1) >> 2, bz0 = (cqz - 1) >> 2;
        int slot = -1;
        if (lane < 8) {
            int bx = (lane & 4) ? ((cqx + 1) >> 2) : bx0, by = (lane & 2) ? ((cqy + 1) >> 2) : by0, bz = (lane & 1) ? ((cqz + 1) >> 2) : bz0;
            slot = dir_find(m, pack_block(bx, by, bz));
        }
        uint32_t b = NONE32;
        int dx = lane % 3 - 1, dy = (lane / 3) % 3 - 1, dz = lane / 9 - 1;
        int cx = cqx + dx, cy = cqy + dy, cz = cqz + dz;
        int which = (((cx >> 2) != bx0) ? 4 : 0) | (((cy >> 2) != by0) ? 2 : 0) | (((cz >> 2) != bz0) ? 1 : 0);
        int s = __shfl_sync(FULL, slot, which);
        if (lane < 27 && s >= 0) b = __ldg(m.cell_tab + (size_t)s * 64 + cell_off(cx, cy, cz));
        unsigned found = __ballot_sync(FULL, b != NONE32);
        scan_found_buckets(m, ws, lane, found, b, qx, qy, qz, ncand);
        ws.flush(lane);
        float cov = covered_sq(m.cell_edge, 1.f, qabs);
        bool done = (ws.count == WarpSelect::FLIVO_K_ && __uint_as_float(ws.tau) <= cov) || (cov >= max_sq);
        if (done) goto finish;
    }
    far_counted = true;
    // ---- shell 2: 5x5x5 cells minus the inner 27 ---------------------------------------------------
    {
        for (int round = 0; round < 4; ++round) {
            int ci = round * 32 + lane;
            uint32_t b = NONE32;
            if (ci < 125) {
                int dx = ci % 5 - 2, dy = (ci / 5) % 5 - 2, dz = ci / 25 - 2;
                if (max(abs(dx), max(abs(dy), abs(dz))) == 2) b = cell_bucket(m, cqx + dx, cqy + dy, cqz + dz);
            }
            unsigned found = __ballot_sync(FULL, b != NONE32);
            scan_found_buckets(m, ws, lane, found, b, qx, qy, qz, ncand);
        }
        ws.flush(lane);
        float cov = covered_sq(m.cell_edge, 2.f, qabs);
        bool done = (ws.count == WarpSelect::FLIVO_K_ && __uint_as_float(ws.tau) <= cov) || (cov >= max_sq);
        if (done) goto finish;
    }
    // ---- block rings ---------------------------------------------------------------------------------
    {
        const int bqx = cqx >> 2, bqy = cqy >> 2, bqz = cqz >> 2;
        for (int ring = 1; ring <= FAR_MAX_RING; ++ring) {
            int w = 2 * ring + 1, total = w * w * w;
            for (int base = 0; base < total; base += 32) {
                int bi = base + lane;
                int slot = -1;
                int bx = 0, by = 0, bz = 0;
                if (bi < total) {
                    int dx = bi % w - ring, dy = (bi / w) % w - ring, dz = bi / (w * w) - ring;
                    int cheb = max(abs(dx), max(abs(dy), abs(dz)));
                    if (cheb == ring || (ring == 1)) {
                        bx = bqx + dx; by = bqy + dy; bz = bqz + dz;
                        slot = dir_find(m, pack_block(bx, by, bz));
                    }
                }
                unsigned fb = __ballot_sync(FULL, slot >= 0);
                while (fb) {
                    int src = __ffs(fb) - 1;
                    fb &= fb - 1;
                    int s = __shfl_sync(FULL, slot, src);
                    int sbx = __shfl_sync(FULL, bx, src), sby = __shfl_sync(FULL, by, src), sbz = __shfl_sync(FULL, bz, src);
                    // lanes cover the 64 cells of the block, two each; skip cells already visited (Chebyshev <= 2)
                    for (int half = 0; half < 2; ++half) {
                        int off = lane + 32 * half;
                        int cx = sbx * 4 + ((off >> 4) & 3), cy = sby * 4 + ((off >> 2) & 3), cz = sbz * 4 + (off & 3);
                        int cheb = max(abs(cx - cqx), max(abs(cy - cqy), abs(cz - cqz)));
                        uint32_t bb = (cheb > 2) ? __ldg(m.cell_tab + (size_t)s * 64 + off) : NONE32;
                        while (__any_sync(FULL, bb != NONE32)) {
                            for (int sl = 0; sl < BUCKET_SLOTS; ++sl) {
                                uint32_t d = INF_BITS, id = 0xFFFFFFFFu, addr = 0;
                                if (bb != NONE32) {
                                    addr = bb * BUCKET_SLOTS + sl;
                                    float4 p = __ldg(m.buckets + addr);
                                    int pid = __float_as_int(p.w);
                                    if (pid >= 0) {
                                        uint32_t db = __float_as_uint(sqdist3(qx, qy, qz, p.x, p.y, p.z));
                                        if (db <= ws.tau) { d = db; id = (uint32_t)pid; }
                                    }
                                }
                                if (__any_sync(FULL, d != INF_BITS)) ws.push(lane, d, id, addr);
                            }
                            ncand += 8ull * __popc(__ballot_sync(FULL, bb != NONE32));
                            bb = (bb != NONE32) ? __ldg(m.bucket_next + bb) : NONE32;
                        }
                    }
                }
            }
            ws.flush(lane);
            float cov = covered_sq(m.cell_edge, 4.f * ring, qabs);
            bool done = (ws.count == WarpSelect::FLIVO_K_ && __uint_as_float(ws.tau) <= cov) || (cov >= max_sq);
            if (done) goto finish;
        }
        // ---- directory sweep (unbounded searches in sparse regions) --------------------------------
        if (stats && lane == 0) atomicAdd(&stats->sweeps, 1ull);
        const float bedge = m.cell_edge * 4.f;
        for (uint32_t base = 0; base <= m.dir_mask; base += 32) {
            uint32_t s = base + lane;
            unsigned long long key = __ldg(m.dir_keys + s);
            bool want = false;
            int bx = 0, by = 0, bz = 0;
            if (key != DIR_EMPTY) {
                unpack_block(key, bx, by, bz);
                int cheb = max(abs(bx - bqx), max(abs(by - bqy), abs(bz - bqz)));
                if (cheb > FAR_MAX_RING) {
                    // conservative lower bound of the distance from q to the block's box
                    float gap = 0.f;
                    float lo, hi, d1;
                    lo = bx * bedge; hi = lo + bedge; d1 = fmaxf(fmaxf(lo - qx, qx - hi), 0.f); gap += d1 * d1;
                    lo = by * bedge; hi = lo + bedge; d1 = fmaxf(fmaxf(lo - qy, qy - hi), 0.f); gap += d1 * d1;
                    lo = bz * bedge; hi = lo + bedge; d1 = fmaxf(fmaxf(lo - qz, qz - hi), 0.f); gap += d1 * d1;
                    float slack = (qabs + sqrtf(gap) + bedge) * 1e-5f + 1e-6f;
                    float g2 = fmaxf(sqrtf(gap) - slack, 0.f);
                    want = (g2 * g2 * 0.9999f) <= __uint_as_float(ws.tau);
                }
            }
            unsigned fb = __ballot_sync(FULL, want);
            while (fb) {
                int src = __ffs(fb) - 1;
                fb &= fb - 1;
                uint32_t ss = __shfl_sync(FULL, s, src);
                for (int half = 0; half < 2; ++half) {
                    int off = lane + 32 * half;
                    uint32_t bb = __ldg(m.cell_tab + (size_t)ss * 64 + off);
                    while (__any_sync(FULL, bb != NONE32)) {
                        for (int sl = 0; sl < BUCKET_SLOTS; ++sl) {
                            uint32_t d = INF_BITS, id = 0xFFFFFFFFu, addr = 0;
                            if (bb != NONE32) {
                                addr = bb * BUCKET_SLOTS + sl;
                                float4 p = __ldg(m.buckets + addr);
                                int pid = __float_as_int(p.w);
                                if (pid >= 0) {
                                    uint32_t db = __float_as_uint(sqdist3(qx, qy, qz, p.x, p.y, p.z));
                                    if (db <= ws.tau) { d = db; id = (uint32_t)pid; }
                                }
                            }
                            if (__any_sync(FULL, d != INF_BITS)) ws.push(lane, d, id, addr);
                        }
                        ncand += 8ull * __popc(__ballot_sync(FULL, bb != NONE32));
                        bb = (bb != NONE32) ? __ldg(m.bucket_next + bb) : NONE32;
                    }
                }
                ws.flush(lane);  // tighten tau before judging the next block
            }
        }
        ws.flush(lane);
    }
finish:
    if (stats && lane == 0) {
        if (far_counted) atomicAdd(&stats->far_queries, 1ull);
        atomicAdd(&stats->candidates, ncand);
    }
    out = ws.c0;
    return ws.count;
}

// ----------------------------------------------------------------------------------------------------
// Thread-per-query fast path over the 3x3x3 cell shell (the common case on a downsampled map: the 5th
// neighbour lies within one cell edge).  Same semantics and the same "final" test as shell 1 of
// warp_knn5; a query the shell cannot finish is left to the warp-cooperative search above.
// Per thread: <= 8 directory probes, the occupancy bits of each block pick the non-empty cells of the
// shell, every bucket is read as 8 independent float4 loads (one 128-byte line).  Neighbouring threads
// hold neighbouring queries (Morton-ordered scan), so their lines coincide in L1/L2.
// ----------------------------------------------------------------------------------------------------
struct Top5 {
    unsigned long long k[5];   // (distance bits << 32) | id, ascending; ~0 = empty
    uint32_t a[5];             // point slot
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int r = 0; r < 5; ++r) { k[r] = 0xFFFFFFFFFFFFFFFFull; a[r] = 0; }
    }
    // precondition: key < k[4].  Shift-insert: every rank looks at its left neighbour only.
    __device__ __forceinline__ void insert(unsigned long long key, uint32_t addr) {
        const bool c3 = key < k[3], c2 = key < k[2], c1 = key < k[1], c0 = key < k[0];
        k[4] = c3 ? k[3] : key;                  a[4] = c3 ? a[3] : addr;
        k[3] = c3 ? (c2 ? k[2] : key) : k[3];    a[3] = c3 ? (c2 ? a[2] : addr) : a[3];
        k[2] = c2 ? (c1 ? k[1] : key) : k[2];    a[2] = c2 ? (c1 ? a[1] : addr) : a[2];
        k[1] = c1 ? (c0 ? k[0] : key) : k[1];    a[1] = c1 ? (c0 ? a[0] : addr) : a[1];
        k[0] = c0 ? key : k[0];                  a[0] = c0 ? addr : a[0];
    }
};

// 64-bit occupancy mask of the cells (x in xm) x (y in ym) x (z in zm) of one block, xm/ym/zm 4-bit masks
// (cell_off layout: x<<4 | y<<2 | z).  Outer product by two multiplications: ym spread to one bit per nibble times zm gives
// the 16-bit yz plane (the partial products occupy distinct nibbles, so no carries), xm spread to one bit per 16-bit lane
// times that plane gives the block mask.
__device__ __forceinline__ unsigned long long mask3(unsigned xm, unsigned ym, unsigned zm) {
    unsigned ys = (ym | (ym << 6)) & 0x0303u;
    ys = (ys | (ys << 3)) & 0x1111u;
    const unsigned yz = ys * zm;
    const unsigned long long xs = ((unsigned long long)xm * 0x0000200040008001ull) & 0x0001000100010001ull;
    return xs * (unsigned long long)yz;
}

// bits of the cells of block (bx,by,bz) that lie in [cq-1, cq+1]^3 (cell_off layout: x<<4 | y<<2 | z)
__device__ __forceinline__ unsigned axis_bits(int cq, int b) {
    int lo = max(cq - 1, b * 4) - b * 4, hi = min(cq + 1, b * 4 + 3) - b * 4;   // 0 <= lo <= hi <= 3
    return ((2u << hi) - 1u) & ~((1u << lo) - 1u);
}
__device__ __forceinline__ unsigned long long shell_bits(int cqx, int cqy, int cqz, int bx, int by, int bz) {
    return mask3(axis_bits(cqx, bx), axis_bits(cqy, by), axis_bits(cqz, bz));
}

// bits of the cells of block (bx,by,bz) inside the cell box [xl,xh] x [yl,yh] x [zl,zh]
__device__ __forceinline__ unsigned range_bits(int lo_c, int hi_c, int b) {
    const int lo = max(lo_c, b * 4) - b * 4, hi = min(hi_c, b * 4 + 3) - b * 4;
    if (lo > hi) return 0u;
    return ((2u << hi) - 1u) & ~((1u << lo) - 1u);
}
__device__ __forceinline__ unsigned long long box_bits(int xl, int xh, int yl, int yh, int zl, int zh, int bx, int by, int bz) {
    return mask3(range_bits(xl, xh, bx), range_bits(yl, yh, by), range_bits(zl, zh, bz));
}

// Returns true when the shell result is final (then best holds the answer; count = number found).
//
// Two passes.  Pass 1 walks the <= 8 blocks of the shell and writes a per-thread work list of its occupied
// cells into shared memory (s_idx = cell_tab index, s_gap = conservative squared distance from the query to
// the cell's box): the cells of the 2x2x2 octant nearest the query go first (gap 0: always visited), the other
// 19 after.  Pass 2 is ONE flat loop over that list: a cell whose box lies beyond the current fifth distance
// is skipped without touching memory, otherwise its bucket (8 slots = one 128-byte line, plus overflow chain)
// is scanned.  The flat loop keeps the lanes of a warp in step far better than nested block/cell loops.
// s_idx / s_gap: [SHELL_CELLS][blockDim.x], this thread's column = threadIdx.x.
constexpr int SHELL_CELLS = 27;

// PAR_PROBES: the (at most) eight blocks of the shell are probed TOGETHER — eight independent directory loads, then eight
// independent occupancy loads: two global-memory round trips for the whole shell instead of two per block.  Measured: it
// wins when the GPU is under-filled (24 k-point scan: search trip 0.072 -> 0.064 ms) and loses when every SM is full
// (260 k points: 0.149 -> 0.155 ms), so the launcher picks it by scan size.
template <bool PAR_PROBES>
__device__ __forceinline__ bool thread_knn5_shell(const MapView& m, float qx, float qy, float qz, uint32_t max_sq_bits,
                                                  Top5& best, int& count, unsigned& ncand,
                                                  uint32_t* __restrict__ s_idx, float* __restrict__ s_gap, int tstride) {
    best.init();
    const int g = m.gshift;
    const int kx = voxel_key1(qx, m.ds), ky = voxel_key1(qy, m.ds), kz = voxel_key1(qz, m.ds);
    const int cqx = kx >> g, cqy = ky >> g, cqz = kz >> g;
    const float qabs = fmaxf(fabsf(qx), fmaxf(fabsf(qy), fabsf(qz)));
    // per-axis squared gaps to the lower / upper neighbour cell (conservative, see cell_gap_sq) and the nearer side
    const double edge = (double)m.cell_edge, slack = ((double)qabs + 2.0 * edge) * 4.8e-7;
    float glo[3], ghi[3];
    int side[3];
    {
        const float q3[3] = {qx, qy, qz};
        const int c3[3] = {cqx, cqy, cqz};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const double lo = (double)c3[a] * edge;
            double dl = (double)q3[a] - lo, dh = (lo + edge) - (double)q3[a];
            side[a] = (dl < dh) ? -1 : 1;
            dl = fmax(fmax(dl, 0.0) * (1.0 - 1e-6) - slack, 0.0);
            dh = fmax(fmax(dh, 0.0) * (1.0 - 1e-6) - slack, 0.0);
            glo[a] = (float)(dl * dl * (1.0 - 2e-6));
            ghi[a] = (float)(dh * dh * (1.0 - 2e-6));
        }
    }
    // ---- pass 1: work list ------------------------------------------------------------------------------
    // octant = per axis the query's own cell and the neighbour on its nearer side
    const int oxl = min(cqx, cqx + side[0]), oxh = max(cqx, cqx + side[0]);
    const int oyl = min(cqy, cqy + side[1]), oyh = max(cqy, cqy + side[1]);
    const int ozl = min(cqz, cqz + side[2]), ozh = max(cqz, cqz + side[2]);
    int nA = 0, nB = 0;
    const int bx0 = (cqx - 1) >> 2, by0 = (cqy - 1) >> 2, bz0 = (cqz - 1) >> 2;
    const int nbx = ((cqx + 1) >> 2) - bx0, nby = ((cqy + 1) >> 2) - by0, nbz = ((cqz + 1) >> 2) - bz0;   // 0 or 1
    int slots[8];
    unsigned long long occs[8];
    unsigned long long bases[8];
    if (PAR_PROBES) {
        // eight independent record loads (one 32-byte sector each): ONE global-memory round trip for the whole shell
        unsigned long long bkey[8], k0[8];
        uint32_t hs[8];
#pragma unroll
        for (int ib = 0; ib < 8; ++ib) {
            const int ix = ib >> 2, iy = (ib >> 1) & 1, iz = ib & 1;
            const bool valid = !(ix > nbx || iy > nby || iz > nbz);
            bkey[ib] = valid ? pack_block(bx0 + ix, by0 + iy, bz0 + iz) : DIR_EMPTY;
            hs[ib] = hash_block(bkey[ib]) & m.dir_mask;
        }
#pragma unroll
        for (int ib = 0; ib < 8; ++ib) {
            k0[ib] = DIR_EMPTY; occs[ib] = 0ull; bases[ib] = NONE32;   // (base NONE32, chained 0)
            if (bkey[ib] != DIR_EMPTY) ld_rec(m.rec + hs[ib], k0[ib], occs[ib], bases[ib]);
        }
#pragma unroll
        for (int ib = 0; ib < 8; ++ib) {
            int sl = -1;
            if (bkey[ib] != DIR_EMPTY) {
                uint32_t h = hs[ib];
                while (true) {                       // linear probing: the first slot answers almost always
                    if (k0[ib] == bkey[ib]) { sl = (int)h; break; }
                    if (k0[ib] == DIR_EMPTY) break;
                    h = (h + 1) & m.dir_mask;
                    ld_rec(m.rec + h, k0[ib], occs[ib], bases[ib]);
                }
            }
            slots[ib] = sl;
            if (sl < 0) occs[ib] = 0ull;
        }
    }
    // (unrolled only where the probe results sit in register arrays: eight copies of the list construction cost the
    // serial flavour 7 % of the search trip in instruction fetch)
#pragma unroll(PAR_PROBES ? 8 : 1)
    for (int ib = 0; ib < 8; ++ib) {
        const int ix = ib >> 2, iy = (ib >> 1) & 1, iz = ib & 1;
        if (ix > nbx || iy > nby || iz > nbz) continue;
        const int bx = bx0 + ix, by = by0 + iy, bz = bz0 + iz;
        int slot;
        unsigned long long full, occ;   // all occupied cells of the block / those inside the shell
        unsigned long long base;        // {bucket base, chained} of the block's record
        if (PAR_PROBES) {
            slot = slots[ib];
            if (slot < 0) continue;
            full = occs[ib]; base = bases[ib];
        } else {
            slot = rec_find(m, pack_block(bx, by, bz), full, base);   // key, occupancy and bucket base in one load
            if (slot < 0) continue;
        }
        occ = full & shell_bits(cqx, cqy, cqz, bx, by, bz);
        if (!occ) continue;
        const unsigned long long octant = box_bits(oxl, oxh, oyl, oyh, ozl, ozh, bx, by, bz);
        unsigned long long oa = occ & octant, ob = occ & ~octant;
        while (oa) {                                  // nearest octant: always visited, no gap needed
            const int off = __ffsll((long long)oa) - 1;
            oa &= oa - 1;
            s_idx[nA * tstride] = cell_entry(slot, full, base, off);
            ++nA;
        }
        const int ox = bx * 4 - cqx, oy = by * 4 - cqy, oz = bz * 4 - cqz;
        while (ob) {
            const int off = __ffsll((long long)ob) - 1;
            ob &= ob - 1;
            const int dx = ((off >> 4) & 3) + ox, dy = ((off >> 2) & 3) + oy, dz = (off & 3) + oz;
            const float gap = (dx < 0 ? glo[0] : dx > 0 ? ghi[0] : 0.f) + (dy < 0 ? glo[1] : dy > 0 ? ghi[1] : 0.f) +
                              (dz < 0 ? glo[2] : dz > 0 ? ghi[2] : 0.f);
            const int pos = (SHELL_CELLS - 1) - nB;
            ++nB;
            s_idx[pos * tstride] = cell_entry(slot, full, base, off);
            s_gap[pos * tstride] = gap;
        }
    }
    // ---- pass 2: flat loop ------------------------------------------------------------------------------
    // candidates above the bound never enter: the sentinel key of an empty rank is the bound itself
    const unsigned long long bound_key = ((unsigned long long)max_sq_bits << 32) | 0xFFFFFFFFull;
    unsigned long long worst = bound_key;   // key a candidate must not exceed to enter
    int found = 0;
    const int total = nA + nB;
    for (int t = 0; t < total; ++t) {
        const int pos = (t < nA) ? t : SHELL_CELLS - 1 - (t - nA);
        if (t >= nA && s_gap[pos * tstride] > __uint_as_float((uint32_t)(worst >> 32))) continue;
        const uint32_t ent = s_idx[pos * tstride];
        uint32_t b = entry_bucket(m, ent);   // compact block: the entry IS the bucket, no table read
        while (b != NONE32) {
            float4 p[BUCKET_SLOTS];
            ld_bucket(m.buckets + (size_t)b * BUCKET_SLOTS, p);
            const uint32_t nxt = entry_next(m, ent, b);
#pragma unroll
            for (int s = 0; s < BUCKET_SLOTS; ++s) {
                const int pid = __float_as_int(p[s].w);
                if (pid < 0) continue;
                ++ncand;
                const uint32_t db = __float_as_uint(sqdist3(qx, qy, qz, p[s].x, p[s].y, p[s].z));
                const unsigned long long key = ((unsigned long long)db << 32) | (uint32_t)pid;
                if (key <= worst) {
                    best.insert(key, b * BUCKET_SLOTS + s);
                    if (++found >= 5) worst = best.k[4];
                }
            }
            b = nxt;
        }
    }
    count = min(found, 5);
    const float cov = covered_sq(m.cell_edge, 1.f, qabs);
    const float max_sq = __uint_as_float(max_sq_bits);
    return (count == 5 && __uint_as_float((uint32_t)(best.k[4] >> 32)) <= cov) || (cov >= max_sq);
}

// ----------------------------------------------------------------------------------------------------
// Warp-per-query search for the queries the 27-cell shell cannot finish (sparse map regions).
// Cubes of cells of growing radius n = 1, 2, .. around the query's cell; in step n the lanes take the
// 4x4x4-cell blocks that intersect cube n, mask the block's occupancy bits with (cube n minus cube n-1),
// and scan the surviving cells into a PER-LANE sorted top five (no shuffles in the scan).  After each
// step the 32 lists are merged (five rounds of warp-min) to get the global fifth distance, which decides
// whether the cube is final and tightens every lane's bound for the next step.  Cubes beyond
// RING_MAX_CELLS (or unbounded searches in empty space) fall back to warp_knn5's directory sweep.
// Output convention of warp_knn5: lanes 0..4 hold rank 0..4 in `out`; returns the number found.
// ----------------------------------------------------------------------------------------------------
constexpr int RING_MAX_CELLS = 32;   // covers the V1 search range of 24 cells (api.cu: search_bound_bits)
constexpr int RING_SPARSE_MIN_BLOCKS = 64;   // steps with more blocks than this go through the super-block table

__device__ __forceinline__ unsigned axis_bits_n(int cq, int b, int n) {
    int lo = max(cq - n, b * 4) - b * 4, hi = min(cq + n, b * 4 + 3) - b * 4;
    if (lo > hi) return 0u;
    return ((2u << hi) - 1u) & ~((1u << lo) - 1u);
}
__device__ __forceinline__ unsigned long long cube_bits(int cqx, int cqy, int cqz, int bx, int by, int bz, int n) {
    if (n < 0) return 0ull;
    return mask3(axis_bits_n(cqx, bx, n), axis_bits_n(cqy, by, n), axis_bits_n(cqz, bz, n));
}

// conservative lower bound (double) of the squared distance from q to any point stored in cell (cx,cy,cz)
__device__ __forceinline__ double cell_gap_sq(double edge, int cx, int cy, int cz, float qx, float qy, float qz, double slack) {
    double g2 = 0.0, lo, d;
    lo = (double)cx * edge; d = fmax(fmax(lo - (double)qx, (double)qx - (lo + edge)), 0.0); d = fmax(d * (1.0 - 1e-6) - slack, 0.0); g2 += d * d;
    lo = (double)cy * edge; d = fmax(fmax(lo - (double)qy, (double)qy - (lo + edge)), 0.0); d = fmax(d * (1.0 - 1e-6) - slack, 0.0); g2 += d * d;
    lo = (double)cz * edge; d = fmax(fmax(lo - (double)qz, (double)qz - (lo + edge)), 0.0); d = fmax(d * (1.0 - 1e-6) - slack, 0.0); g2 += d * d;
    return g2 * (1.0 - 1e-6);
}

// Per-warp cell queue of warp_knn5_rings (shared memory, provided by the caller): RING_QCAP entries of
// {cell_tab index, packed cell offset from the query's cell}.
constexpr int RING_QCHUNK = 16;                       // cells a lane enqueues per pass
constexpr int RING_QCAP = 32 * RING_QCHUNK + 32;

// scan one queued cell into the lane's sorted top five
__device__ __forceinline__ void ring_scan_cell(const MapView& m, uint2 ent, int bx0, int by0, int bz0, int nby, int nbz,
                                               float qx, float qy, float qz,
                                               double edge, double slack, double gworst, Top5& loc, unsigned long long& lworst,
                                               int& lfound, unsigned& ncand) {
    // ent.x = cell entry (bucket or cell_tab index, see CELL_DIRECT), ent.y = linear index of the block in the step's
    // block box << 6 | cell offset in the block
    const int bi = (int)(ent.y >> 6), off = (int)(ent.y & 63u);
    const int cx = (bx0 + bi / (nby * nbz)) * 4 + ((off >> 4) & 3), cy = (by0 + (bi / nbz) % nby) * 4 + ((off >> 2) & 3),
              cz = (bz0 + bi % nbz) * 4 + (off & 3);
    if (cell_gap_sq(edge, cx, cy, cz, qx, qy, qz, slack) > gworst) return;
    uint32_t b = entry_bucket(m, ent.x);
    while (b != NONE32) {
        float4 pt[BUCKET_SLOTS];
        ld_bucket(m.buckets + (size_t)b * BUCKET_SLOTS, pt);
        const uint32_t nxt = entry_next(m, ent.x, b);
#pragma unroll
        for (int s2 = 0; s2 < BUCKET_SLOTS; ++s2) {
            const int pid = __float_as_int(pt[s2].w);
            if (pid < 0) continue;
            ++ncand;
            const uint32_t db = __float_as_uint(sqdist3(qx, qy, qz, pt[s2].x, pt[s2].y, pt[s2].z));
            const unsigned long long key = ((unsigned long long)db << 32) | (uint32_t)pid;
            if (key <= lworst) {
                loc.insert(key, b * BUCKET_SLOTS + s2);
                if (++lfound >= 5 && loc.k[4] < lworst) lworst = loc.k[4];
            }
        }
        b = nxt;
    }
}

// seed (optional): the sorted best candidates already found in cube 1 by thread_knn5_shell — lane r < seed_cnt passes
// candidate r as (distance bits, id, bucket address); the search then starts at cube 2 (or straight at the cube that
// covers the seeded fifth distance) instead of scanning the 27 inner cells again.
static __device__ __noinline__ int warp_knn5_rings(const MapView m /* by value: a reference would make every caller keep an addressable copy of its whole parameter block */, float qx, float qy, float qz, uint32_t max_sq_bits, int lane,
                                            Cand& out, unsigned& ncand, uint2* __restrict__ queue /* [RING_QCAP], this warp's */,
                                            int seed_cnt = -1, uint32_t seed_d = 0, uint32_t seed_id = 0, uint32_t seed_addr = 0,
                                            int* dbg = nullptr /* steps << 8 | last radius */) {
    Top5 loc;
    loc.init();
    const int g = m.gshift;
    const int cqx = voxel_key1(qx, m.ds) >> g, cqy = voxel_key1(qy, m.ds) >> g, cqz = voxel_key1(qz, m.ds) >> g;
    const float qabs = fmaxf(fabsf(qx), fmaxf(fabsf(qy), fabsf(qz)));
    const float max_sq = __uint_as_float(max_sq_bits);
    const unsigned long long bound_key = ((unsigned long long)max_sq_bits << 32) | 0xFFFFFFFFull;
    unsigned long long lworst = bound_key;   // per lane: key a candidate must not exceed
    double gworst = (double)max_sq;          // warp-uniform: squared distance beyond which a whole cell is skipped
    int lfound = 0;
    const double edge = (double)m.cell_edge;
    // radius schedule: cube 1; then, once five candidates are known, straight to the smallest cube that covers the
    // fifth distance (one more step, guaranteed final); while fewer than five are known: cube 2, then the cube
    // that covers the distance bound.  Each step scans cube(n) minus cube(p).
    int p = -1, n = 1;
    if (seed_cnt >= 0) {   // cube 1 is done
        if (lane < seed_cnt) { loc.insert(((unsigned long long)seed_d << 32) | seed_id, seed_addr); lfound = 1; }
        p = 1;
        if (seed_cnt == 5) {
            const uint32_t d5 = __shfl_sync(FULL, seed_d, 4), i5 = __shfl_sync(FULL, seed_id, 4);
            lworst = ((unsigned long long)d5 << 32) | i5;
            gworst = fmin(gworst, (double)__uint_as_float(d5));
            const float need = __uint_as_float(d5);
            do { ++n; } while (n < RING_MAX_CELLS && covered_sq(m.cell_edge, (float)n, qabs) < need);
        } else {
            n = 2;
        }
    }
    int steps = 0;
    while (true) {
        ++steps;
        if (dbg) *dbg = (steps << 8) | n;
        const int bx0 = (cqx - n) >> 2, by0 = (cqy - n) >> 2, bz0 = (cqz - n) >> 2;
        const int nbx = ((cqx + n) >> 2) - bx0 + 1, nby = ((cqy + n) >> 2) - by0 + 1, nbz = ((cqz + n) >> 2) - bz0 + 1;
        const int nb = nbx * nby * nbz;
        const double slack = ((double)qabs + (double)(n + 1) * edge) * 4.8e-7;
        int qn = 0;   // warp-uniform queue length
        // Block candidates of the step, 32 at a time.  Small cubes: every block of the box, one per lane.  Wide cubes
        // (V1's range of 24 cells: 13^3 = 2197 blocks, mostly absent where the query looks into unmapped space): the
        // SUPER-BLOCKS of the box, one per lane (<= 125 questions), then only the blocks that exist, taken from the
        // lanes' masks bit by bit.
        const bool sparse = nb > RING_SPARSE_MIN_BLOCKS;
        const int bx1 = bx0 + nbx - 1, by1 = by0 + nby - 1, bz1 = bz0 + nbz - 1;
        const int sx0 = bx0 >> 2, sy0 = by0 >> 2, sz0 = bz0 >> 2;
        const int nsy = (by1 >> 2) - sy0 + 1, nsz = (bz1 >> 2) - sz0 + 1;
        const int nsb = ((bx1 >> 2) - sx0 + 1) * nsy * nsz;
        unsigned long long smask = 0ull;   // sparse: blocks of this lane's current super-block still to visit
        int ssx = 0, ssy = 0, ssz = 0;
        for (int base = 0;;) {
            // stage 1: one block per lane -> directory slot and the occupancy bits of its cells in ring n
            int bi = -1;
            int slot = -1, bx = 0, by = 0, bz = 0;
            unsigned long long occ = 0, full = 0, bbase = NONE32;
            if (!sparse) {
                if (base >= nb) break;
                bi = base + lane;
                base += 32;
                if (bi < nb) { bx = bx0 + bi / (nby * nbz); by = by0 + (bi / nbz) % nby; bz = bz0 + bi % nbz; }
                else bi = -1;
            } else {
                if (!__any_sync(FULL, smask != 0ull)) {
                    if (base >= nsb) break;
                    const int si = base + lane;
                    base += 32;
                    if (si < nsb) {
                        ssx = sx0 + si / (nsy * nsz); ssy = sy0 + (si / nsz) % nsy; ssz = sz0 + si % nsz;
                        const unsigned long long inbox = mask3(range_bits(bx0, bx1, ssx), range_bits(by0, by1, ssy), range_bits(bz0, bz1, ssz));
                        smask = inbox ? (sup_find(m, ssx, ssy, ssz) & inbox) : 0ull;
                    }
                    continue;
                }
                if (smask) {
                    const int o = __ffsll((long long)smask) - 1;
                    smask &= smask - 1;
                    bx = ssx * 4 + ((o >> 4) & 3); by = ssy * 4 + ((o >> 2) & 3); bz = ssz * 4 + (o & 3);
                    bi = ((bx - bx0) * nby + (by - by0)) * nbz + (bz - bz0);
                }
            }
            if (bi >= 0) {
                const unsigned long long want = cube_bits(cqx, cqy, cqz, bx, by, bz, n) & ~cube_bits(cqx, cqy, cqz, bx, by, bz, p);
                if (want) {
                    slot = rec_find(m, pack_block(bx, by, bz), full, bbase);
                    if (slot >= 0) occ = full & want;
                }
            }
            // stage 2: the occupied cells of all 32 blocks go through one queue so that every lane scans one cell per
            // round, whatever block it came from
            while (__any_sync(FULL, occ != 0ull)) {
                const int take = min(__popcll(occ), RING_QCHUNK);
                int pre = take;   // inclusive warp scan
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(FULL, pre, o); if (lane >= o) pre += v; }
                const int total = __shfl_sync(FULL, pre, 31);
                int pos = qn + pre - take;
                for (int t = 0; t < take; ++t) {
                    const int off = __ffsll((long long)occ) - 1;
                    occ &= occ - 1;
                    queue[pos++] = make_uint2(cell_entry(slot, full, bbase, off), ((uint32_t)bi << 6) | (uint32_t)off);
                }
                qn += total;
                __syncwarp();
                while (qn >= 32) {
                    qn -= 32;
                    ring_scan_cell(m, queue[qn + lane], bx0, by0, bz0, nby, nbz, qx, qy, qz, edge, slack, gworst, loc, lworst, lfound, ncand);
                }
                __syncwarp();
            }
        }
        if (lane < qn) ring_scan_cell(m, queue[lane], bx0, by0, bz0, nby, nbz, qx, qy, qz, edge, slack, gworst, loc, lworst, lfound, ncand);
        __syncwarp();
        // merge the 32 sorted lists: rank r -> lane r
        int h = 0, count = 0;
        Cand res; res.d = INF_BITS; res.id = 0xFFFFFFFFu; res.addr = 0;
#pragma unroll
        for (int r = 0; r < 5; ++r) {
            unsigned long long head = h == 0 ? loc.k[0] : h == 1 ? loc.k[1] : h == 2 ? loc.k[2] : h == 3 ? loc.k[3] : h == 4 ? loc.k[4] : 0xFFFFFFFFFFFFFFFFull;
            uint32_t ha = h == 0 ? loc.a[0] : h == 1 ? loc.a[1] : h == 2 ? loc.a[2] : h == 3 ? loc.a[3] : h == 4 ? loc.a[4] : 0u;
            const uint32_t hd = (uint32_t)(head >> 32), hi = (uint32_t)head;
            const uint32_t dmin = __reduce_min_sync(FULL, hd);
            if (dmin == 0xFFFFFFFFu) break;  // warp-uniform: nothing left
            const uint32_t idmin = __reduce_min_sync(FULL, hd == dmin ? hi : 0xFFFFFFFFu);
            const bool win = (hd == dmin) && (hi == idmin);
            const int wl = __ffs(__ballot_sync(FULL, win)) - 1;
            const uint32_t wa = __shfl_sync(FULL, ha, wl);
            if (lane == r) { res.d = dmin; res.id = idmin; res.addr = wa; }
            if (win) ++h;
            ++count;
        }
        const uint32_t d5 = __shfl_sync(FULL, res.d, 4);
        const uint32_t i5 = __shfl_sync(FULL, res.id, 4);
        const float cov = covered_sq(m.cell_edge, (float)n, qabs);
        const bool done = (count == 5 && __uint_as_float(d5) <= cov) || (cov >= max_sq);
        if (done) { out = res; return count; }
        if (n >= RING_MAX_CELLS) break;
        p = n;
        if (count == 5) {
            const unsigned long long gk = ((unsigned long long)d5 << 32) | i5;
            if (gk < lworst) lworst = gk;
            gworst = fmin(gworst, (double)__uint_as_float(d5));
            const float need = __uint_as_float(d5);
            do { ++n; } while (n < RING_MAX_CELLS && covered_sq(m.cell_edge, (float)n, qabs) < need);
        } else {
            // fewer than five known: a nearby bound (the V3 gate: <= 8 cells) is covered in one more step; a wide one
            // (V1: 24 cells) by doubling cubes, so that a point a few metres from the map does not pay for the whole range
            int nb_ = n;
            do { ++nb_; } while (nb_ < RING_MAX_CELLS && covered_sq(m.cell_edge, (float)nb_, qabs) < max_sq);
            n = (n < 2) ? 2 : ((nb_ <= 8) ? nb_ : min(2 * n, nb_));
        }
    }
    // not final within RING_MAX_CELLS cells: the generic search (block rings + directory sweep)
    return warp_knn5(m, qx, qy, qz, max_sq_bits, lane, out, nullptr);
}

// ----------------------------------------------------------------------------------------------------
// iVox-compatible search: IVox::GetClosestPoint(pt, closest_pt, max_num = 5, max_range = 5.0)
// (include/ivox3d/ivox3d.h:132-204) over IVoxNode::KNNPointByCondition (include/ivox3d/ivox3d_node.hpp:141-206).
// Candidates = the points of the voxels key + delta, delta running over nearby_grids_ (ivox3d.h:213-243): CENTER,
// NEARBY6, NEARBY18 and NEARBY26 are the offsets with at most 0, 1, 2 and 3 non-zero components (1, 7, 19, 27
// voxels; the visiting order does not matter for the result), whose float squared distance is < max_range^2 compared in double.  The reference truncates every voxel to its K
// nearest (nth_element) and then takes the K nearest of the union: the same SET as the K nearest of all
// candidates.  Its output order is whatever two std::nth_element calls leave (only [0] is the minimum); here the
// list is ascending under the canonical (distance, insertion id) order — stated in DESIGN.md.
// Thread per query; consecutive deltas mostly fall into the same 4x4x4 block, whose directory slot is reused.
// ----------------------------------------------------------------------------------------------------
static __device__ __constant__ signed char c_ivox_delta[27][3] = {   // sorted by the number of non-zero components
    {0, 0, 0}, {-1, 0, 0}, {0, -1, 0}, {0, 0, -1}, {0, 0, 1}, {0, 1, 0}, {1, 0, 0},
    {-1, -1, 0}, {-1, 0, -1}, {-1, 0, 1}, {-1, 1, 0}, {0, -1, -1}, {0, -1, 1}, {0, 1, -1},
    {0, 1, 1}, {1, -1, 0}, {1, 0, -1}, {1, 0, 1}, {1, 1, 0}, {-1, -1, -1}, {-1, -1, 1},
    {-1, 1, -1}, {-1, 1, 1}, {1, -1, -1}, {1, -1, 1}, {1, 1, -1}, {1, 1, 1}};

__device__ __forceinline__ void thread_knn5_ivox(const MapView& m, float qx, float qy, float qz, double max_range_sq,
                                                 Top5& best, int& count, unsigned& ncand) {
    best.init();
    const int kx = voxel_key(m, qx), ky = voxel_key(m, qy), kz = voxel_key(m, qz);
    unsigned long long last_key = DIR_EMPTY, worst = 0xFFFFFFFFFFFFFFFFull;
    int last_slot = -1, found = 0;
    for (int t = 0; t < m.ivox; ++t) {
        const int cx = kx + c_ivox_delta[t][0], cy = ky + c_ivox_delta[t][1], cz = kz + c_ivox_delta[t][2];
        const unsigned long long bk = pack_block(cx >> 2, cy >> 2, cz >> 2);
        if (bk != last_key) { last_key = bk; last_slot = dir_find(m, bk); }
        if (last_slot < 0) continue;
        uint32_t b = __ldg(m.cell_tab + (size_t)last_slot * 64 + cell_off(cx, cy, cz));
        while (b != NONE32) {
            float4 p[BUCKET_SLOTS];
            ld_bucket(m.buckets + (size_t)b * BUCKET_SLOTS, p);
            const uint32_t nxt = __ldg(m.bucket_next + b);
#pragma unroll
            for (int s = 0; s < BUCKET_SLOTS; ++s) {
                const int pid = __float_as_int(p[s].w);
                if (pid < 0) continue;
                ++ncand;
                const float d = sqdist3_eigen(p[s].x, p[s].y, p[s].z, qx, qy, qz);
                if (!((double)d < max_range_sq)) continue;
                const unsigned long long key = ((unsigned long long)__float_as_uint(d) << 32) | (uint32_t)pid;
                if (key <= worst) {
                    best.insert(key, b * BUCKET_SLOTS + s);
                    if (++found >= 5) worst = best.k[4];
                }
            }
            b = nxt;
        }
    }
    count = min(found, 5);
}

}  // namespace flivo
