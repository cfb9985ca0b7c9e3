// flivo_dev.cuh — device-side data layout and strict-float helpers shared by the kernels of
// libflivo_b200.so (sm_100a only).
//
// Every translation unit that includes this file is compiled with -fmad=false: the reference is built
// for baseline x86-64 (no FMA, CMakeLists.txt:7,23), and every float that feeds a comparison (kNN
// distance include/ikd-Tree/ikd_Tree.cpp:1291-1295, plane test include/common_lib.h:694, gates
// src/laser_mapping.cpp:518,533-535) must round exactly as separate mul / add.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace flivo {

// ------------------------------------------------------------------------------------------------
// GPU-resident map ("(a) a GPU-resident map" of north_star): a three-level voxel hash.
//   voxel  : edge ds = KD_TREE::downsample_size; integer key k = floorf(x / ds) per axis — the same float
//            expression the reference uses to form its downsample box (ikd_Tree.cpp:392-397)
//   cell   : 2^gshift voxels per axis (search granularity); at most one live point per voxel after
//            Add_Points(…, true), so a cell of 2x2x2 voxels normally holds <= 8 points = one bucket
//   block  : 4x4x4 cells; the block directory is an open-addressing hash (linear probing) keyed by the
//            packed block coordinates; slot s owns cell_tab[s*64 .. s*64+63] (bucket index per cell)
//            and blk_mask[s] (occupancy bits of its 64 cells)
//   bucket : 8 x float4 {x, y, z, id-as-bits} = one 128-byte line; empty slot <=> id == -1;
//            bucket_next[] chains overflow buckets (dense, not-downsampled Build input)
// Points are never moved once stored; deletes clear the slot in place.
// ------------------------------------------------------------------------------------------------
constexpr unsigned long long DIR_EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr uint32_t NONE32 = 0xFFFFFFFFu;
constexpr int BUCKET_SLOTS = 8;
constexpr int EMPTY_ID = -1;
constexpr int COORD_BIAS = 1 << 20;  // voxel keys are biased into [0, 2^21)
constexpr uint32_t INF_BITS = 0x7F800000u;

// Search record of one directory slot, derived from dir_keys / blk_mask / cell_tab (which stay authoritative for the
// insert / delete pipelines) and kept in step with them by map.cu (k_block_records).  One aligned 32-byte sector answers
// "is this the block, which of its cells are occupied, where are their buckets": a shell probe costs ONE dependent
// round trip per block instead of three (key -> occupancy -> cell table).
//   base != NONE32 : the block is COMPACT — the buckets of its occupied cells are consecutive in cell-offset order, so
//                    bucket(cell off) = base + popc(mask & ((1 << off) - 1)) and the cell table is not read at all.
//                    Every block whose cells were all created by one insert batch is (Build; new ground of a scan).
//   base == NONE32 : cells were added to the block by a later batch; their buckets are looked up in cell_tab.
//   chained        : some cell of the block has an overflow chain (more than 8 live points: dense, not down-sampled
//                    input only).  For the other blocks the search never reads bucket_next.
struct __align__(32) BlockRec {
    unsigned long long key;        // packed block coordinates, DIR_EMPTY = free slot (same as dir_keys[slot])
    unsigned long long mask;       // occupancy bits (same as blk_mask[slot])
    uint32_t base;
    uint32_t chained;
    uint32_t pad[2];
};
// work-list / queue entries of the search name a cell either by its bucket (compact block) or by its cell_tab index;
// CELL_NOCHAIN on a direct entry: the bucket is the cell's only one
constexpr uint32_t CELL_DIRECT = 0x80000000u;
constexpr uint32_t CELL_NOCHAIN = 0x40000000u;
constexpr uint32_t CELL_BUCKET_BITS = 0x3FFFFFFFu;
constexpr size_t DIR_CAP_MAX = 1ull << 25;   // cell_tab indices (slot * 64 + off) must stay below CELL_DIRECT

struct MapView {
    const BlockRec* rec;           // [dir_cap] search records
    unsigned long long* dir_keys;  // [dir_cap]
    uint32_t* cell_tab;            // [dir_cap * 64]
    unsigned long long* blk_mask;  // [dir_cap]
    float4* buckets;               // [bucket_cap * 8]
    uint32_t* bucket_next;         // [bucket_cap]
    // coarse occupancy: a SUPER-BLOCK is 4x4x4 blocks (16 cells a side); sup_keys / sup_mask are an open-addressing table of
    // the same capacity as the directory (packed super-block coordinates -> one bit per block that exists).  A wide search
    // cube asks ~1/64 as many questions of it as of the block directory, and only blocks that exist are then probed.
    unsigned long long* sup_keys;  // [dir_cap]
    unsigned long long* sup_mask;  // [dir_cap]
    uint32_t dir_mask;             // dir_cap - 1 (dir_cap is a power of two)
    uint32_t bucket_cap;
    float ds;                      // downsample_size
    int gshift;                    // log2(cell_ratio)
    float cell_edge;               // ds * 2^gshift (exact)
    // iVox-compatible mode (include/ivox3d/ivox3d.h): ds = Options::resolution_, cell = voxel
    // (gshift 0), voxel key = round(p * inv_resolution_) (Pos2Grid, ivox3d.h:278-281), every point of a voxel kept
    int ivox;                      // 0 = ikd-Tree semantics; else the number of voxels of the nearby stencil (1, 7, 19, 27)
    float inv_res;                 // (float)(1.0 / resolution_) (ivox3d.h:63)
};

// Device-side counters of the map (one struct in global memory).
struct MapCounters {
    unsigned long long live;       // live points
    unsigned int blocks_used;      // directory slots in use
    unsigned int cells_used;
    unsigned int buckets_used;     // bump allocator
    int error;                     // sticky: 1 = coordinate out of range, 2 = directory full, 3 = bucket pool full
    int add_counter;               // tmp_counter of the running Add_Points
    int del_counter;
    unsigned int deferred;         // voxel segments deferred to the serial pass
    unsigned int removed_count;    // entries in the removed-points log
    unsigned int pad[3];
};

struct int3s { int x, y, z; };

// counters of the kNN kernels (search.cuh)
struct SearchStats {
    unsigned long long far_queries;   // queries that left the 27-cell fast path
    unsigned long long candidates;    // candidate points examined
    unsigned long long sweeps;        // queries that needed the directory sweep
};


#ifndef FLIVO_HOST_EMU   // (tests/cpp/solve_emu.cpp compiles solve_dev.cuh for the host: layouts only, no device helpers)
// ---- strict float helpers (explicit rounding; immune to contraction even without -fmad=false) -----
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }

// calc_dist (ikd_Tree.cpp:1291-1295): (dx*dx + dy*dy) + dz*dz in float
__device__ __forceinline__ float sqdist3(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = fsub(ax, bx), dy = fsub(ay, by), dz = fsub(az, bz);
    return fadd(fadd(fmul(dx, dx), fmul(dy, dy)), fmul(dz, dz));
}

// distance2 of the iVox node (ivox3d_node.hpp:13-16): (pt1 - pt2).squaredNorm() on Vector3f; Eigen's unrolled
// reduction of a 3-vector adds x*x + (y*y + z*z) (Eigen is not in the reference tree: order unpinned, the oracle
// restates the same)
__device__ __forceinline__ float sqdist3_eigen(float ax, float ay, float az, float bx, float by, float bz) {
    float dx = fsub(ax, bx), dy = fsub(ay, by), dz = fsub(az, bz);
    return fadd(fmul(dx, dx), fadd(fmul(dy, dy), fmul(dz, dz)));
}

// voxel key: floor(p / downsample_size) as in Box_of_Point.vertex_min (ikd_Tree.cpp:392-396)
__device__ __forceinline__ int voxel_key1(float x, float ds) { return (int)floorf(fdiv(x, ds)); }

// key of the voxel that stores p: Box_of_Point (ikd_Tree.cpp:392-396) or, in iVox mode, IVox::Pos2Grid
// (ivox3d.h:278-281): (pt * inv_resolution_).array().round().cast<int>() — std::round, half away from zero
__device__ __forceinline__ int voxel_key(const MapView& m, float x) {
    return m.ivox ? (int)roundf(fmul(x, m.inv_res)) : (int)floorf(fdiv(x, m.ds));
}

__device__ __forceinline__ bool key_in_range(int k) { return k > -COORD_BIAS + 8 && k < COORD_BIAS - 8; }

// packed block key from block coordinates (each in about [-2^18, 2^18))
__device__ __forceinline__ unsigned long long pack_block(int bx, int by, int bz) {
    return ((unsigned long long)(unsigned)(bx + (1 << 19)) << 40) |
           ((unsigned long long)(unsigned)(by + (1 << 19)) << 20) |
           (unsigned long long)(unsigned)(bz + (1 << 19));
}
__device__ __forceinline__ void unpack_block(unsigned long long k, int& bx, int& by, int& bz) {
    bx = (int)((k >> 40) & 0xFFFFF) - (1 << 19);
    by = (int)((k >> 20) & 0xFFFFF) - (1 << 19);
    bz = (int)(k & 0xFFFFF) - (1 << 19);
}
__device__ __forceinline__ uint32_t hash_block(unsigned long long k) {
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull; k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull; k ^= k >> 33;
    return (uint32_t)k;
}
__device__ __forceinline__ int cell_off(int cx, int cy, int cz) { return ((cx & 3) << 4) | ((cy & 3) << 2) | (cz & 3); }

// directory lookup through the search records: slot index or -1; mask / base of the block on success.  One 32-byte load
// per probe (ld.global.nc.v4.u64 on sm_100: the whole record in one instruction).
// (base comes back with the chained flag in bit 31 .. see rec_base / rec_chained)
__device__ __forceinline__ void ld_rec(const BlockRec* p, unsigned long long& key, unsigned long long& mask, unsigned long long& bc) {
    unsigned long long w3;
    asm("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(key), "=l"(mask), "=l"(bc), "=l"(w3) : "l"(p));
    (void)w3;
}
// One bucket = 8 x float4 = one 128-byte line, fetched by FOUR 256-bit loads.  The search is bound by the L1 wavefront
// queue (every load instruction costs ~2 cycles per distinct line its lanes touch, B300_MICROARCH "LDG"): half the
// load instructions per bucket is half the wavefronts.
__device__ __forceinline__ void ld_bucket(const float4* __restrict__ line, float4 (&p)[BUCKET_SLOTS]) {
#pragma unroll
    for (int h = 0; h < BUCKET_SLOTS / 2; ++h) {
        unsigned long long a, b, c, d;
        asm("ld.global.nc.v4.u64 {%0, %1, %2, %3}, [%4];" : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(line + 2 * h));
        p[2 * h] = make_float4(__uint_as_float((uint32_t)a), __uint_as_float((uint32_t)(a >> 32)), __uint_as_float((uint32_t)b), __uint_as_float((uint32_t)(b >> 32)));
        p[2 * h + 1] = make_float4(__uint_as_float((uint32_t)c), __uint_as_float((uint32_t)(c >> 32)), __uint_as_float((uint32_t)d), __uint_as_float((uint32_t)(d >> 32)));
    }
}
// bc = {base (low word), chained (high word)} as loaded from the record
__device__ __forceinline__ int rec_find(const MapView& m, unsigned long long key, unsigned long long& mask, unsigned long long& bc) {
    uint32_t s = hash_block(key) & m.dir_mask;
    while (true) {
        unsigned long long k;
        ld_rec(m.rec + s, k, mask, bc);
        if (k == key) return (int)s;
        if (k == DIR_EMPTY) { mask = 0ull; return -1; }
        s = (s + 1) & m.dir_mask;
    }
}
// cell `off` of a found block as a work-list entry (see CELL_DIRECT)
__device__ __forceinline__ uint32_t cell_entry(int slot, unsigned long long mask, unsigned long long bc, int off) {
    const uint32_t base = (uint32_t)bc;
    if (base == NONE32) return (uint32_t)slot * 64u + (uint32_t)off;
    return CELL_DIRECT | ((bc >> 32) ? 0u : CELL_NOCHAIN) | (base + (uint32_t)__popcll(mask & ((1ull << off) - 1ull)));
}
__device__ __forceinline__ uint32_t entry_bucket(const MapView& m, uint32_t e) {
    return (e & CELL_DIRECT) ? (e & CELL_BUCKET_BITS) : __ldg(m.cell_tab + e);
}
__device__ __forceinline__ uint32_t entry_next(const MapView& m, uint32_t e, uint32_t b) {
    return ((e & (CELL_DIRECT | CELL_NOCHAIN)) == (CELL_DIRECT | CELL_NOCHAIN)) ? NONE32 : __ldg(m.bucket_next + b);
}

// directory lookup: slot index or -1
__device__ __forceinline__ int dir_find(const MapView& m, unsigned long long key) {
    uint32_t s = hash_block(key) & m.dir_mask;
    while (true) {
        unsigned long long k = __ldg(m.dir_keys + s);
        if (k == key) return (int)s;
        if (k == DIR_EMPTY) return -1;
        s = (s + 1) & m.dir_mask;
    }
}
// find-or-insert; returns slot (never fails while load factor < 1; callers keep it <= 0.5)
// a block came into existence: set its bit in its super-block's mask (the table has the directory's capacity and holds at
// most as many entries as the directory has blocks, so it cannot fill up before the directory does)
__device__ __forceinline__ void sup_mark(const MapView& m, unsigned long long block_key) {
    int bx, by, bz;
    unpack_block(block_key, bx, by, bz);
    const unsigned long long skey = pack_block(bx >> 2, by >> 2, bz >> 2);
    const unsigned long long bit = 1ull << cell_off(bx, by, bz);
    uint32_t s = hash_block(skey ^ 0x9E3779B97F4A7C15ull) & m.dir_mask;
    for (uint32_t probe = 0; probe <= m.dir_mask; ++probe) {
        unsigned long long k = m.sup_keys[s];
        if (k == DIR_EMPTY) k = atomicCAS(m.sup_keys + s, DIR_EMPTY, skey);
        if (k == DIR_EMPTY || k == skey) { atomicOr(m.sup_mask + s, bit); return; }
        s = (s + 1) & m.dir_mask;
    }
}
// bits of the blocks that exist in super-block (sx, sy, sz); 0 when none does
__device__ __forceinline__ unsigned long long sup_find(const MapView& m, int sx, int sy, int sz) {
    const unsigned long long skey = pack_block(sx, sy, sz);
    uint32_t s = hash_block(skey ^ 0x9E3779B97F4A7C15ull) & m.dir_mask;
    while (true) {
        const unsigned long long k = __ldg(m.sup_keys + s);
        if (k == skey) return __ldg(m.sup_mask + s);
        if (k == DIR_EMPTY) return 0ull;
        s = (s + 1) & m.dir_mask;
    }
}
__device__ __forceinline__ int dir_find_or_insert(const MapView& m, unsigned long long key, MapCounters* cnt) {
    uint32_t s = hash_block(key) & m.dir_mask;
    for (uint32_t probe = 0; probe <= m.dir_mask; ++probe) {
        unsigned long long k = m.dir_keys[s];
        if (k == key) return (int)s;
        if (k == DIR_EMPTY) {
            unsigned long long old = atomicCAS(m.dir_keys + s, DIR_EMPTY, key);
            if (old == DIR_EMPTY) { atomicAdd(&cnt->blocks_used, 1u); sup_mark(m, key); return (int)s; }
            if (old == key) return (int)s;
        }
        s = (s + 1) & m.dir_mask;
    }
    atomicExch(&cnt->error, 2);
    return -1;
}
// bucket index of a cell or NONE32
__device__ __forceinline__ uint32_t cell_bucket(const MapView& m, int cx, int cy, int cz) {
    int slot = dir_find(m, pack_block(cx >> 2, cy >> 2, cz >> 2));
    if (slot < 0) return NONE32;
    return __ldg(m.cell_tab + (size_t)slot * 64 + cell_off(cx, cy, cz));
}

// Box of a voxel key exactly as the reference forms it (ikd_Tree.cpp:392-397): min = k*ds, max = min + ds (float)
__device__ __forceinline__ void voxel_box1(int k, float ds, float& lo, float& hi) {
    lo = fmul((float)k, ds);
    hi = fadd(lo, ds);
}
// Search_by_range / Delete_by_range membership (ikd_Tree.cpp:648,1001): min <= p < max
__device__ __forceinline__ bool in_box1(float p, float lo, float hi) { return lo <= p && hi > p; }

// same_point (ikd_Tree.cpp:1287-1289), EPSS = 1e-6 compared in double
__device__ __forceinline__ bool same_point3(float ax, float ay, float az, float bx, float by, float bz) {
    return (double)fabsf(fsub(ax, bx)) < 1e-6 && (double)fabsf(fsub(ay, by)) < 1e-6 && (double)fabsf(fsub(az, bz)) < 1e-6;
}
#endif  // FLIVO_HOST_EMU

}  // namespace flivo
