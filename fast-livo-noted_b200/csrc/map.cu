// map.cu — GPU-resident incremental map with ikd-Tree semantics (sm_100a).
//
// Reference behaviour reproduced here (paths relative to the reference tree):
//   KD_TREE::Build                include/ikd-Tree/ikd_Tree.cpp:337-348   (no downsampling)
//   KD_TREE::Add_Points           :382-457   (per-voxel nearest-to-centre downsample, strict '<', same_point)
//   KD_TREE::Delete_Point_Boxes   :501-521 / Delete_by_range :626-688 (half-open boxes)
//   KD_TREE::Delete_Points        :481-499,  Add_Point_Boxes :459-479, flatten :1238-1263,
//   tree_range :94-118, acquire_removed_points :523-535
//
// Add_Points is sequential in the reference (one range search + delete + insert per point).  Here a
// batch is sorted by voxel key and every voxel segment is decided by one thread, which is exact because
// voxels are independent — EXCEPT where float rounding puts a point inside a neighbouring voxel's box
// ("slop").  Such segments are detected and replayed in original batch order by a
// single warp that executes the reference's per-point logic literally.
#include "map.cuh"

#include <cub/cub.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

namespace flivo {

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            err = std::string(#call) + ": " + cudaGetErrorString(e_);                    \
            return -1;                                                                   \
        }                                                                                \
    } while (0)

// ------------------------------------------------------------------------------------------------
// key helpers
// sort key (63 bits): [bx+2^16:17][by+2^16:17][bz+2^16:17][cell offset in block:6][voxel in cell:6]
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long make_sort_key(int kx, int ky, int kz, int gshift) {
    int gm = (1 << gshift) - 1;
    int cx = kx >> gshift, cy = ky >> gshift, cz = kz >> gshift;
    unsigned long long key = ((unsigned long long)(unsigned)((cx >> 2) + (1 << 16)) << 34) |
                             ((unsigned long long)(unsigned)((cy >> 2) + (1 << 16)) << 17) |
                             (unsigned long long)(unsigned)((cz >> 2) + (1 << 16));
    key = (key << 6) | (unsigned)cell_off(cx, cy, cz);
    key = (key << 6) | (unsigned)(((kx & gm) << 4) | ((ky & gm) << 2) | (kz & gm));
    return key;
}
// Slot-keyed flavour (bit 63 set): [directory slot of the block][cell offset:6][voxel in cell:6].  The block is named by its
// slot in the directory instead of its coordinates, so a batch sorts on 13 + log2(dir_cap) bits (4 radix passes for a scan
// instead of 8); the block must exist before the key is made (k_point_keys creates it), which needs the directory sized
// beforehand — scan-sized batches only.  Both flavours keep block / cell / voxel identity at the same shifts (>> 12, >> 6).
constexpr unsigned long long SLOT_KEY = 1ull << 63;
__device__ __forceinline__ unsigned long long make_slot_key(int slot, int kx, int ky, int kz, int gshift) {
    const int gm = (1 << gshift) - 1;
    return SLOT_KEY | ((unsigned long long)(unsigned)slot << 12) | ((unsigned long long)(unsigned)cell_off(kx >> gshift, ky >> gshift, kz >> gshift) << 6) |
           (unsigned)(((kx & gm) << 4) | ((ky & gm) << 2) | (kz & gm));
}
__device__ __forceinline__ void decode_sort_key(const MapView& m, unsigned long long key, int& kx, int& ky, int& kz) {
    const int gshift = m.gshift;
    int sub = (int)(key & 63), off = (int)((key >> 6) & 63);
    unsigned long long b = key >> 12;
    int bx, by, bz;
    if (key & SLOT_KEY) unpack_block(m.dir_keys[(uint32_t)b & m.dir_mask], bx, by, bz);
    else { bz = (int)(b & 0x1FFFF) - (1 << 16); by = (int)((b >> 17) & 0x1FFFF) - (1 << 16); bx = (int)((b >> 34) & 0x1FFFF) - (1 << 16); }
    int cx = bx * 4 + ((off >> 4) & 3), cy = by * 4 + ((off >> 2) & 3), cz = bz * 4 + (off & 3);
    kx = (cx << gshift) | ((sub >> 4) & 3);
    ky = (cy << gshift) | ((sub >> 2) & 3);
    kz = (cz << gshift) | (sub & 3);
}
__device__ __forceinline__ bool cell_in_range(int c) { return c > -(1 << 18) + 8 && c < (1 << 18) - 8; }
constexpr unsigned long long INVALID_KEY = 0x7FFFFFFFFFFFFFFFull;  // sorts last; skipped by every stage
__device__ __forceinline__ bool point_ok(const MapView& m, float4 p, int& kx, int& ky, int& kz) {
    kx = voxel_key(m, p.x); ky = voxel_key(m, p.y); kz = voxel_key(m, p.z);
    return isfinite(p.x) && isfinite(p.y) && isfinite(p.z) && fabsf(p.x / m.ds) < 1.0e6f && fabsf(p.y / m.ds) < 1.0e6f &&
           fabsf(p.z / m.ds) < 1.0e6f && cell_in_range(kx >> m.gshift) && cell_in_range(ky >> m.gshift) && cell_in_range(kz >> m.gshift);
}

// read-write flavour of the lookups (no __ldg: tables are being mutated by the running pipeline)
__device__ __forceinline__ int dir_find_rw(const MapView& m, unsigned long long key) {
    uint32_t s = hash_block(key) & m.dir_mask;
    while (true) {
        unsigned long long k = ((volatile unsigned long long*)m.dir_keys)[s];
        if (k == key) return (int)s;
        if (k == DIR_EMPTY) return -1;
        s = (s + 1) & m.dir_mask;
    }
}
__device__ __forceinline__ uint32_t cell_bucket_rw(const MapView& m, int cx, int cy, int cz) {
    int slot = dir_find_rw(m, pack_block(cx >> 2, cy >> 2, cz >> 2));
    if (slot < 0) return NONE32;
    return ((volatile uint32_t*)m.cell_tab)[(size_t)slot * 64 + cell_off(cx, cy, cz)];
}
__device__ __forceinline__ float4 ld_point(const MapView& m, size_t slot) {
    return __ldcg(m.buckets + slot);
}
__device__ __forceinline__ float mid1(float lo, float hi) {
    // mid_point = min + (max - min) / 2.0   (double division, narrowed; ikd_Tree.cpp:398-400)
    return (float)((double)lo + (double)fsub(hi, lo) / 2.0);
}
// a coordinate is "slop" when the set of voxel boxes containing it is not exactly {its own key}
__device__ __forceinline__ bool slop_axis(float x, int k, float ds) {
    float lo, hi;
    voxel_box1(k, ds, lo, hi);
    bool own = in_box1(x, lo, hi);
    voxel_box1(k - 1, ds, lo, hi);
    bool left = in_box1(x, lo, hi);
    voxel_box1(k + 1, ds, lo, hi);
    bool right = in_box1(x, lo, hi);
    return !own || left || right;
}

// claim a free slot in the chain starting at bucket b and store p (w = id bits)
__device__ bool claim_slot(const MapView& m, uint32_t b, float4 p, MapCounters* cnt, bool allow_append) {
    for (int guard = 0; guard < (1 << 20); ++guard) {
        for (int s = 0; s < BUCKET_SLOTS; ++s) {
            int* w = reinterpret_cast<int*>(&m.buckets[(size_t)b * BUCKET_SLOTS + s].w);
            if (*((volatile int*)w) == EMPTY_ID) {
                if (atomicCAS(w, EMPTY_ID, -2) == EMPTY_ID) {
                    m.buckets[(size_t)b * BUCKET_SLOTS + s] = p;
                    return true;
                }
            }
        }
        uint32_t nxt = ((volatile uint32_t*)m.bucket_next)[b];
        if (nxt == NONE32) {
            if (!allow_append) { atomicExch(&cnt->error, 3); return false; }
            uint32_t nb = atomicAdd(&cnt->buckets_used, 1u);
            if (nb >= m.bucket_cap) { atomicExch(&cnt->error, 3); return false; }
            uint32_t old = atomicCAS(m.bucket_next + b, NONE32, nb);
            nxt = (old == NONE32) ? nb : old;  // a lost race leaks bucket nb (rare: > 8 live points in one cell)
        }
        b = nxt;
    }
    return false;
}

// ------------------------------------------------------------------------------------------------
// kernels
// ------------------------------------------------------------------------------------------------
__global__ void k_stamp_ids(const float4* __restrict__ in, float4* __restrict__ out, size_t n, int base_id, const BatchJob* job) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (job) {
        base_id = job->base_id;
        if (i >= job->n) { const float qn = __int_as_float(0x7FC00000); out[i] = make_float4(qn, qn, qn, __int_as_float(EMPTY_ID)); return; }   // padding: no stage takes it for a point
    }
    float4 p = in[i];
    p.w = __int_as_float(base_id + (int)i);
    out[i] = p;
}

__global__ void k_point_keys(const float4* __restrict__ pts, size_t n, MapView m, unsigned long long* __restrict__ keys,
                             unsigned int* __restrict__ idx, MapCounters* cnt, const BatchJob* job, int slot_keys) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (job && i >= job->n) { keys[i] = INVALID_KEY; idx[i] = (unsigned int)i; return; }   // padding, not an error
    float4 p = pts[i];
    int kx, ky, kz;
    bool ok = point_ok(m, p, kx, ky, kz);
    if (!ok) atomicExch(&cnt->error, 1);
    unsigned long long key = INVALID_KEY;
    if (ok && slot_keys) {
        const int cx = kx >> m.gshift, cy = ky >> m.gshift, cz = kz >> m.gshift;
        const int slot = dir_find_or_insert(m, pack_block(cx >> 2, cy >> 2, cz >> 2), cnt);   // (a full directory is reported through cnt->error)
        if (slot >= 0) key = make_slot_key(slot, kx, ky, kz, m.gshift);
    } else if (ok) {
        key = make_sort_key(kx, ky, kz, m.gshift);
    }
    keys[i] = key;
    idx[i] = (unsigned int)i;
}

__global__ void k_count_unique(const unsigned long long* __restrict__ keys, size_t n, unsigned int* misc) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    bool valid = keys[i] != INVALID_KEY;
    bool blk_head = valid && ((i == 0) || ((keys[i] >> 12) != (keys[i - 1] >> 12)));
    bool cell_head = valid && ((i == 0) || ((keys[i] >> 6) != (keys[i - 1] >> 6)));
    unsigned act = __activemask();
    unsigned bm = __ballot_sync(act, blk_head), cm = __ballot_sync(act, cell_head);
    int lane = threadIdx.x & 31;
    if (lane == __ffs(act) - 1) {
        if (bm) atomicAdd(&misc[0], __popc(bm));
        if (cm) atomicAdd(&misc[1], __popc(cm));
    }
}

// Cell creation of an insert batch, in three steps so that the NEW cells of the batch receive consecutive buckets in
// sorted-key order (block, then cell offset): a block whose cells all come from one batch is then "compact" — its
// buckets are base + rank of the cell among the block's occupied cells — and the search needs no cell table for it
// (BlockRec, flivo_dev.cuh).
//   k_cells_mark   : one thread per cell head: block slot (created if missing), flag = the cell does not exist yet
//   (exclusive scan of the flags -> rank of every new cell; k_cells_reserve takes the bucket range in one step)
//   k_cells_create : new cell -> bucket = range start + rank; then, for non-downsampled inserts, enough chained
//                    buckets for the incoming points (taken from the bump allocator behind the range)
__global__ void k_cells_mark(const unsigned long long* __restrict__ keys, size_t n, MapView m, MapCounters* cnt,
                             unsigned int* __restrict__ is_new) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    is_new[i] = 0;
    unsigned long long key = keys[i];
    if (key == INVALID_KEY) return;
    if (i != 0 && (key >> 6) == (keys[i - 1] >> 6)) return;
    int kx, ky, kz;
    decode_sort_key(m, key, kx, ky, kz);
    int cx = kx >> m.gshift, cy = ky >> m.gshift, cz = kz >> m.gshift;
    int slot = dir_find_or_insert(m, pack_block(cx >> 2, cy >> 2, cz >> 2), cnt);
    if (slot < 0) return;
    if (m.cell_tab[(size_t)slot * 64 + cell_off(cx, cy, cz)] == NONE32) is_new[i] = 1;
}
// misc[4] = first bucket of the batch's new cells; the bump allocator moves past them
__global__ void k_cells_reserve(const unsigned int* __restrict__ is_new, const unsigned int* __restrict__ rank, size_t n,
                                MapView m, MapCounters* cnt, unsigned int* misc) {
    if (threadIdx.x || blockIdx.x) return;
    const unsigned int total = rank[n - 1] + is_new[n - 1];
    const unsigned int first = cnt->buckets_used;
    misc[4] = first;
    cnt->buckets_used = first + total;
    cnt->cells_used += total;
    if ((size_t)first + total > m.bucket_cap) cnt->error = 3;
}
__global__ void k_cells_create(const unsigned long long* __restrict__ keys, size_t n, MapView m, MapCounters* cnt,
                               const unsigned int* __restrict__ is_new, const unsigned int* __restrict__ rank,
                               const unsigned int* __restrict__ misc, int prealloc_chain) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long key = keys[i];
    if (key == INVALID_KEY) return;
    if (i != 0 && (key >> 6) == (keys[i - 1] >> 6)) return;
    int kx, ky, kz;
    decode_sort_key(m, key, kx, ky, kz);
    int cx = kx >> m.gshift, cy = ky >> m.gshift, cz = kz >> m.gshift;
    int slot = dir_find_rw(m, pack_block(cx >> 2, cy >> 2, cz >> 2));
    if (slot < 0) return;
    int off = cell_off(cx, cy, cz);
    uint32_t* entry = m.cell_tab + (size_t)slot * 64 + off;
    uint32_t b = *entry;
    if (is_new[i]) {
        b = misc[4] + rank[i];
        if (b >= m.bucket_cap) { atomicExch(&cnt->error, 3); return; }
        *entry = b;
        atomicOr(m.blk_mask + slot, 1ull << off);
    }
    if (b == NONE32) return;
    if (!prealloc_chain) return;
    // count incoming points of this cell and free slots of its chain; append what is missing
    size_t j = i + 1;
    while (j < n && (keys[j] >> 6) == (key >> 6)) ++j;
    long need = (long)(j - i);
    uint32_t last = b;
    while (true) {
        for (int s = 0; s < BUCKET_SLOTS; ++s)
            if (__float_as_int(m.buckets[(size_t)last * BUCKET_SLOTS + s].w) < 0) --need;
        uint32_t nxt = m.bucket_next[last];
        if (nxt == NONE32) break;
        last = nxt;
    }
    while (need > 0) {
        uint32_t nb = atomicAdd(&cnt->buckets_used, 1u);
        if (nb >= m.bucket_cap) { atomicExch(&cnt->error, 3); return; }
        m.bucket_next[last] = nb;
        last = nb;
        need -= BUCKET_SLOTS;
    }
}

// search record of one directory slot from the authoritative tables (see BlockRec)
__device__ __forceinline__ void write_record(const MapView& m, BlockRec* rec, uint32_t slot) {
    const unsigned long long key = m.dir_keys[slot], mask = m.blk_mask[slot];
    uint32_t base = NONE32, chained = 0;
    if (key != DIR_EMPTY && mask) {
        const uint32_t* tab = m.cell_tab + (size_t)slot * 64;
        unsigned long long bits = mask;
        base = tab[__ffsll((long long)bits) - 1];
        const uint32_t first = base;
        uint32_t r = 0;
        while (bits) {
            const int off = __ffsll((long long)bits) - 1;
            bits &= bits - 1;
            if (tab[off] != first + r) base = NONE32;
            if (m.bucket_next[tab[off]] != NONE32) chained = 1;
            ++r;
        }
        if (base != NONE32 && (base + 64u) > CELL_BUCKET_BITS) base = NONE32;
    }
    BlockRec v;
    v.key = key; v.mask = mask; v.base = base; v.chained = chained; v.pad[0] = v.pad[1] = 0;
    rec[slot] = v;
}
// after the cells of a batch exist: one thread per block head of the sorted batch refreshes that block's record
__global__ void k_block_records(const unsigned long long* __restrict__ keys, size_t n, MapView m, BlockRec* rec) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long key = keys[i];
    if (key == INVALID_KEY) return;
    if (i != 0 && (key >> 12) == (keys[i - 1] >> 12)) return;
    int kx, ky, kz;
    decode_sort_key(m, key, kx, ky, kz);
    int slot = dir_find_rw(m, pack_block((kx >> m.gshift) >> 2, (ky >> m.gshift) >> 2, (kz >> m.gshift) >> 2));
    if (slot >= 0) write_record(m, rec, (uint32_t)slot);
}
// every slot (after the directory was rehashed into a larger table)
__global__ void k_all_records(MapView m, size_t dir_cap, BlockRec* rec) {
    size_t s = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (s < dir_cap) write_record(m, rec, (uint32_t)s);
}

// non-downsampled insert: every point is stored
__global__ void k_insert_all(const float4* __restrict__ pts, size_t n, MapView m, MapCounters* cnt) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = pts[i];
    int kx, ky, kz;
    if (!point_ok(m, p, kx, ky, kz)) return;
    uint32_t b = cell_bucket_rw(m, kx >> m.gshift, ky >> m.gshift, kz >> m.gshift);
    if (b == NONE32) { atomicExch(&cnt->error, 3); return; }
    if (claim_slot(m, b, p, cnt, true)) atomicAdd(&cnt->live, 1ull);
}

// batch points whose box membership is ambiguous go to the slop list
__global__ void k_mark_slop(const float4* __restrict__ pts, size_t n, MapView m, float4* slop_list, unsigned int* misc) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 p = pts[i];
    int kx, ky, kz;
    if (!point_ok(m, p, kx, ky, kz)) return;
    if (slop_axis(p.x, kx, m.ds) || slop_axis(p.y, ky, m.ds) || slop_axis(p.z, kz, m.ds)) {
        unsigned int s = atomicAdd(&misc[2], 1u);
        if (s < SLOP_LIST_CAP) slop_list[s] = make_float4(p.x, p.y, p.z, __int_as_float((int)i));
    }
}

struct VoxBox {
    float lo[3], hi[3], mid[3];
};
__device__ __forceinline__ VoxBox make_box(int kx, int ky, int kz, float ds) {
    VoxBox b;
    voxel_box1(kx, ds, b.lo[0], b.hi[0]);
    voxel_box1(ky, ds, b.lo[1], b.hi[1]);
    voxel_box1(kz, ds, b.lo[2], b.hi[2]);
    for (int a = 0; a < 3; ++a) b.mid[a] = mid1(b.lo[a], b.hi[a]);
    return b;
}
__device__ __forceinline__ bool in_box3(const VoxBox& b, float x, float y, float z) {
    return in_box1(x, b.lo[0], b.hi[0]) && in_box1(y, b.lo[1], b.hi[1]) && in_box1(z, b.lo[2], b.hi[2]);
}

// Decide every voxel segment (thread per sorted element; only segment heads work).
__global__ void k_ds_decide(const unsigned long long* __restrict__ keys, const unsigned int* __restrict__ sidx,
                            const float4* __restrict__ pts, size_t n, MapView m, SegInfo* seg,
                            const float4* __restrict__ slop_list, const unsigned int* __restrict__ misc,
                            unsigned char* defer_flag, MapCounters* cnt) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long key = keys[i];
    SegInfo si;
    si.keep_id = -1; si.new_idx = -1; si.triggers = 0; si.flags = 0;
    if (key == INVALID_KEY || (i != 0 && key == keys[i - 1])) { seg[i] = si; return; }
    si.flags = 1;
    int kx, ky, kz;
    decode_sort_key(m, key, kx, ky, kz);
    const float ds = m.ds;
    VoxBox box = make_box(kx, ky, kz, ds);
    unsigned int nslop = misc[2];
    bool defer = nslop > SLOP_LIST_CAP;  // list overflow: replay the whole batch serially
    // (a)+(c): ambiguous batch points that belong to, or fall into, this voxel
    if (!defer) {
        for (unsigned int s = 0; s < nslop; ++s) {
            float4 q = slop_list[s];
            bool mine = voxel_key1(q.x, ds) == kx && voxel_key1(q.y, ds) == ky && voxel_key1(q.z, ds) == kz;
            if (mine || in_box3(box, q.x, q.y, q.z)) { defer = true; break; }
        }
    }
    // stored points: those inside the box form Downsample_Storage; (b) ambiguous ones defer the segment
    int c = 0;
    float best_d = INFINITY;
    int best_id = 0x7FFFFFFF;
    float bx = 0.f, by = 0.f, bz = 0.f;
    if (!defer) {
        // The cells come through the blocks' search records (one 32-byte load answers "which cells of this block hold
        // points and where are their buckets"; the cells of one voxel's neighbourhood share one or two blocks).  The records
        // are those of the map BEFORE this batch (k_block_records runs at the end of insert_batch): cells and blocks the
        // batch itself created are missing from them, and they hold no stored point yet.
        int g = m.gshift;
        bool have_rec = false;
        unsigned long long rec_key = 0ull, full = 0ull, bc = 0ull;
        int slot = -1;
        for (int cx = (kx - 1) >> g; cx <= (kx + 1) >> g && !defer; ++cx)
            for (int cy = (ky - 1) >> g; cy <= (ky + 1) >> g && !defer; ++cy)
                for (int cz = (kz - 1) >> g; cz <= (kz + 1) >> g && !defer; ++cz) {
                    const unsigned long long bkey = pack_block(cx >> 2, cy >> 2, cz >> 2);
                    if (!have_rec || bkey != rec_key) { slot = rec_find(m, bkey, full, bc); rec_key = bkey; have_rec = true; }
                    if (slot < 0) continue;
                    const int off = cell_off(cx, cy, cz);
                    if (!((full >> off) & 1ull)) continue;
                    const uint32_t ent = cell_entry(slot, full, bc, off);
                    uint32_t b = entry_bucket(m, ent);
                    while (b != NONE32 && !defer) {
                        float4 line[BUCKET_SLOTS];
                        ld_bucket(m.buckets + (size_t)b * BUCKET_SLOTS, line);
                        const uint32_t nxt = entry_next(m, ent, b);
                        for (int s = 0; s < BUCKET_SLOTS; ++s) {
                            float4 x = line[s];
                            int id = __float_as_int(x.w);
                            if (id < 0) continue;
                            bool inb = in_box3(box, x.x, x.y, x.z);
                            int xkx = voxel_key1(x.x, ds), xky = voxel_key1(x.y, ds), xkz = voxel_key1(x.z, ds);
                            bool mine = xkx == kx && xky == ky && xkz == kz;
                            if (!inb && !mine) continue;
                            if (slop_axis(x.x, xkx, ds) || slop_axis(x.y, xky, ds) || slop_axis(x.z, xkz, ds)) { defer = true; break; }
                            // here: mine && inb
                            ++c;
                            float d = sqdist3(x.x, x.y, x.z, box.mid[0], box.mid[1], box.mid[2]);
                            if (d < best_d || (d == best_d && id < best_id)) { best_d = d; best_id = id; bx = x.x; by = x.y; bz = x.z; }
                        }
                        b = nxt;
                    }
                }
    }
    if (defer) {
        si.flags |= 2;
        for (size_t j = i; j < n && keys[j] == key; ++j) defer_flag[sidx[j]] = 1;
        atomicAdd(&cnt->deferred, 1u);
        seg[i] = si;
        return;
    }
    // replay the segment's points in batch order (ikd_Tree.cpp:403-417)
    bool cur_is_new = false;
    int cur_new_idx = -1;
    bool changed = false;
    for (size_t j = i; j < n && keys[j] == key; ++j) {
        unsigned int pi = sidx[j];
        float4 p = pts[pi];
        float dp = sqdist3(p.x, p.y, p.z, box.mid[0], box.mid[1], box.mid[2]);
        bool stored_wins = (c > 0) && (best_d < dp);
        float rx = stored_wins ? bx : p.x, ry = stored_wins ? by : p.y, rz = stored_wins ? bz : p.z;
        bool trig = (c > 1) || same_point3(p.x, p.y, p.z, rx, ry, rz);
        if (trig) {
            si.triggers++;
            changed = true;
            if (!stored_wins) {
                bx = p.x; by = p.y; bz = p.z; best_d = dp; best_id = __float_as_int(p.w);
                cur_is_new = true; cur_new_idx = (int)pi;
            }
            c = 1;
        }
    }
    if (changed) {
        si.flags |= 4;
        si.keep_id = cur_is_new ? -1 : best_id;
        si.new_idx = cur_is_new ? cur_new_idx : -1;
    }
    seg[i] = si;
}

// clear the stored points a decided segment replaces (all of them live in the segment's own cell)
__global__ void k_ds_clear(const unsigned long long* __restrict__ keys, size_t n, MapView m, const SegInfo* __restrict__ seg,
                           MapCounters* cnt) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    SegInfo si = seg[i];
    if ((si.flags & 7) != 5) return;  // head, not deferred, changed
    int kx, ky, kz;
    decode_sort_key(m, keys[i], kx, ky, kz);
    VoxBox box = make_box(kx, ky, kz, m.ds);
    uint32_t b = cell_bucket_rw(m, kx >> m.gshift, ky >> m.gshift, kz >> m.gshift);
    int cleared = 0;
    while (b != NONE32) {
        for (int s = 0; s < BUCKET_SLOTS; ++s) {
            size_t slot = (size_t)b * BUCKET_SLOTS + s;
            float4 x = ld_point(m, slot);
            int id = __float_as_int(x.w);
            if (id < 0 || id == si.keep_id) continue;
            if (in_box3(box, x.x, x.y, x.z)) {
                m.buckets[slot].w = __int_as_float(EMPTY_ID);
                ++cleared;
            }
        }
        b = m.bucket_next[b];
    }
    if (cleared) atomicAdd(&cnt->live, (unsigned long long)(-(long long)cleared));
    if (si.triggers) atomicAdd(&cnt->add_counter, si.triggers);
}

__global__ void k_ds_store(const unsigned long long* __restrict__ keys, size_t n, const float4* __restrict__ pts, MapView m,
                           const SegInfo* __restrict__ seg, MapCounters* cnt) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    SegInfo si = seg[i];
    if ((si.flags & 7) != 5 || si.new_idx < 0) return;
    int kx, ky, kz;
    decode_sort_key(m, keys[i], kx, ky, kz);
    uint32_t b = cell_bucket_rw(m, kx >> m.gshift, ky >> m.gshift, kz >> m.gshift);
    if (b == NONE32) { atomicExch(&cnt->error, 3); return; }
    if (claim_slot(m, b, pts[si.new_idx], cnt, true)) atomicAdd(&cnt->live, 1ull);
}

// ---- serial replay of deferred points: the reference's per-point logic, literally, by one warp -----
// Every lane scans part of the 27-voxel neighbourhood (cells covering voxel keys K-1..K+1).
__global__ void k_ds_serial(const unsigned int* __restrict__ list, const unsigned int* __restrict__ misc, const float4* __restrict__ pts,
                            MapView m, MapCounters* cnt) {
    const int lane = threadIdx.x;
    const unsigned FULL = 0xFFFFFFFFu;
    unsigned int count = misc[3];
    const float ds = m.ds;
    const int g = m.gshift;
    for (unsigned int t = 0; t < count; ++t) {
        float4 p = pts[list[t]];
        int kx = voxel_key1(p.x, ds), ky = voxel_key1(p.y, ds), kz = voxel_key1(p.z, ds);
        VoxBox box = make_box(kx, ky, kz, ds);
        // up to 8 neighbourhood cells; lane c (<8) owns cell c for the lookup, then 4 lanes share a cell
        int ncx = ((kx + 1) >> g) - ((kx - 1) >> g) + 1, ncy = ((ky + 1) >> g) - ((ky - 1) >> g) + 1,
            ncz = ((kz + 1) >> g) - ((kz - 1) >> g) + 1;
        int ncell = ncx * ncy * ncz;  // 1..27 in general; handled in rounds of 8
        int c = 0;
        float best_d = INFINITY;
        int best_id = 0x7FFFFFFF;
        size_t best_slot = 0;
        for (int pass = 0; pass < 2; ++pass) {
            // pass 0: gather Downsample_Storage statistics; pass 1: clear (if triggered)
            int keep_id = -1;
            bool do_clear = false;
            if (pass == 1) {
                // reduce pass-0 results across the warp
                int ctot = c;
                for (int o = 16; o; o >>= 1) ctot += __shfl_xor_sync(FULL, ctot, o);
                // best = min (dist, id)
                float bd = best_d; int bi = best_id; unsigned long long bs = (unsigned long long)best_slot;
                for (int o = 16; o; o >>= 1) {
                    float od = __shfl_xor_sync(FULL, bd, o);
                    int oi = __shfl_xor_sync(FULL, bi, o);
                    unsigned long long os = __shfl_xor_sync(FULL, bs, o);
                    if (od < bd || (od == bd && oi < bi)) { bd = od; bi = oi; bs = os; }
                }
                float dp = sqdist3(p.x, p.y, p.z, box.mid[0], box.mid[1], box.mid[2]);
                bool stored_wins = (ctot > 0) && (bd < dp);
                float4 w = p;
                if (stored_wins) w = ld_point(m, (size_t)bs);
                bool trig = (ctot > 1) || same_point3(p.x, p.y, p.z, w.x, w.y, w.z);
                if (!trig) break;
                do_clear = true;
                keep_id = stored_wins ? bi : -1;
                c = stored_wins ? 0 : 1;  // reuse c as "must store p" flag after the clear
            }
            int cleared = 0;
            for (int base = 0; base < ncell; base += 8) {
                int ci = base + (lane >> 2);
                uint32_t b = NONE32;
                if (ci < ncell) {
                    int ix = ci / (ncy * ncz), iy = (ci / ncz) % ncy, iz = ci % ncz;
                    b = cell_bucket_rw(m, ((kx - 1) >> g) + ix, ((ky - 1) >> g) + iy, ((kz - 1) >> g) + iz);
                }
                while (__any_sync(FULL, b != NONE32)) {
                    if (b != NONE32) {
                        for (int s = (lane & 3) * 2; s < (lane & 3) * 2 + 2; ++s) {
                            size_t slot = (size_t)b * BUCKET_SLOTS + s;
                            float4 x = ld_point(m, slot);
                            int id = __float_as_int(x.w);
                            if (id < 0 || !in_box3(box, x.x, x.y, x.z)) continue;
                            if (pass == 0) {
                                ++c;
                                float d = sqdist3(x.x, x.y, x.z, box.mid[0], box.mid[1], box.mid[2]);
                                if (d < best_d || (d == best_d && id < best_id)) { best_d = d; best_id = id; best_slot = slot; }
                            } else if (do_clear && id != keep_id) {
                                m.buckets[slot].w = __int_as_float(EMPTY_ID);
                                ++cleared;
                            }
                        }
                        b = ((volatile uint32_t*)m.bucket_next)[b];
                    }
                }
            }
            if (pass == 1) {
                for (int o = 16; o; o >>= 1) cleared += __shfl_xor_sync(FULL, cleared, o);
                __threadfence();
                __syncwarp();
                if (lane == 0) {
                    long long delta = -(long long)cleared;
                    if (c == 1) {  // p itself is the downsample result: store it in the cell of its own key
                        uint32_t b = cell_bucket_rw(m, kx >> g, ky >> g, kz >> g);
                        if (b == NONE32) atomicExch(&cnt->error, 3);
                        else if (claim_slot(m, b, p, cnt, true)) delta += 1;
                    }
                    if (delta) atomicAdd(&cnt->live, (unsigned long long)delta);
                    atomicAdd(&cnt->add_counter, 1);
                }
                __threadfence();
                __syncwarp();
            }
        }
        __syncwarp();
    }
}

// ---- whole-table passes ----------------------------------------------------------------------------
// one warp per directory slot; lane l scans cells l and l+32
template <class F>
__device__ __forceinline__ void for_each_live_in_slot(const MapView& m, size_t slot, int lane, F f) {
    for (int off = lane; off < 64; off += 32) {
        uint32_t b = ((volatile uint32_t*)m.cell_tab)[slot * 64 + off];
        while (b != NONE32) {
            for (int s = 0; s < BUCKET_SLOTS; ++s) {
                size_t ps = (size_t)b * BUCKET_SLOTS + s;
                float4 x = ld_point(m, ps);
                if (__float_as_int(x.w) >= 0) f(x, ps);
            }
            b = ((volatile uint32_t*)m.bucket_next)[b];
        }
    }
}

__global__ void k_delete_boxes(MapView m, size_t dir_cap, const float* __restrict__ boxes, int nb, MapCounters* cnt,
                               float4* removed, unsigned int removed_cap) {
    size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= dir_cap) return;
    unsigned long long key = m.dir_keys[warp];
    if (key == DIR_EMPTY) return;
    // coarse reject: block bounds (generous slack for float rounding) against every box
    int bx, by, bz;
    unpack_block(key, bx, by, bz);
    float edge = m.cell_edge * 4.f;
    float lo[3] = {bx * edge, by * edge, bz * edge};
    bool any = false;
    for (int b = 0; b < nb && !any; ++b) {
        const float* B = boxes + 6 * b;
        bool hit = true;
        for (int a = 0; a < 3; ++a) {
            float slack = 1e-3f * (fabsf(lo[a]) + edge) + 1e-3f;
            if (B[3 + a] < lo[a] - slack || B[a] > lo[a] + edge + slack) hit = false;
        }
        any = hit;
    }
    if (!any) return;
    int deleted = 0;
    for_each_live_in_slot(m, warp, lane, [&](float4 x, size_t ps) {
        for (int b = 0; b < nb; ++b) {
            const float* B = boxes + 6 * b;
            // Delete_by_range point test (ikd_Tree.cpp:648): min <= p && max > p
            if (B[0] <= x.x && B[3] > x.x && B[1] <= x.y && B[4] > x.y && B[2] <= x.z && B[5] > x.z) {
                m.buckets[ps].w = __int_as_float(EMPTY_ID);
                unsigned int r = atomicAdd(&cnt->removed_count, 1u);
                if (r < removed_cap) removed[r] = x;
                ++deleted;
                break;
            }
        }
    });
    for (int o = 16; o; o >>= 1) deleted += __shfl_xor_sync(0xFFFFFFFFu, deleted, o);
    if (lane == 0 && deleted) {
        atomicAdd(&cnt->live, (unsigned long long)(-(long long)deleted));
        atomicAdd(&cnt->del_counter, deleted);
    }
}

__global__ void k_flatten(MapView m, size_t dir_cap, float4* out, unsigned int cap, unsigned int* out_count) {
    size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= dir_cap) return;
    if (m.dir_keys[warp] == DIR_EMPTY) return;
    for_each_live_in_slot(m, warp, lane, [&](float4 x, size_t) {
        unsigned int r = atomicAdd(out_count, 1u);
        if (r < cap) out[r] = x;
    });
}

__device__ __forceinline__ void atomic_min_float(float* addr, float v) {
    // ordered-int trick
    if (v >= 0) atomicMin((int*)addr, __float_as_int(v));
    else atomicMax((unsigned int*)addr, __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_float(float* addr, float v) {
    if (v >= 0) atomicMax((int*)addr, __float_as_int(v));
    else atomicMin((unsigned int*)addr, __float_as_uint(v));
}
__global__ void k_range(MapView m, size_t dir_cap, float* out6) {
    size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= dir_cap) return;
    if (m.dir_keys[warp] == DIR_EMPTY) return;
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for_each_live_in_slot(m, warp, lane, [&](float4 x, size_t) {
        lo[0] = fminf(lo[0], x.x); lo[1] = fminf(lo[1], x.y); lo[2] = fminf(lo[2], x.z);
        hi[0] = fmaxf(hi[0], x.x); hi[1] = fmaxf(hi[1], x.y); hi[2] = fmaxf(hi[2], x.z);
    });
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xFFFFFFFFu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xFFFFFFFFu, hi[a], o));
        }
        if (lane == 0 && lo[a] <= hi[a]) { atomic_min_float(out6 + a, lo[a]); atomic_max_float(out6 + 3 + a, hi[a]); }
    }
}

__global__ void k_rehash(MapView oldm, size_t old_cap, MapView newm, MapCounters* cnt) {
    size_t warp = (blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= old_cap) return;
    unsigned long long key = oldm.dir_keys[warp];
    if (key == DIR_EMPTY) return;
    int slot = 0;
    if (lane == 0) {
        uint32_t s = hash_block(key) & newm.dir_mask;
        while (true) {
            unsigned long long old = atomicCAS(newm.dir_keys + s, DIR_EMPTY, key);
            if (old == DIR_EMPTY) break;
            s = (s + 1) & newm.dir_mask;
        }
        slot = (int)s;
        newm.blk_mask[s] = oldm.blk_mask[warp];
        sup_mark(newm, key);
    }
    slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
    newm.cell_tab[(size_t)slot * 64 + lane] = oldm.cell_tab[warp * 64 + lane];
    newm.cell_tab[(size_t)slot * 64 + 32 + lane] = oldm.cell_tab[warp * 64 + 32 + lane];
}

// Delete_Points: sequential in the reference; one warp replays the list (rarely used API).
__global__ void k_delete_points(const float4* __restrict__ pts, size_t n, MapView m, MapCounters* cnt, float4* removed,
                                unsigned int removed_cap) {
    const int lane = threadIdx.x;
    const unsigned FULL = 0xFFFFFFFFu;
    const int g = m.gshift;
    for (size_t t = 0; t < n; ++t) {
        float4 p = pts[t];
        int kx = voxel_key1(p.x, m.ds), ky = voxel_key1(p.y, m.ds), kz = voxel_key1(p.z, m.ds);
        int ncx = ((kx + 1) >> g) - ((kx - 1) >> g) + 1, ncy = ((ky + 1) >> g) - ((ky - 1) >> g) + 1,
            ncz = ((kz + 1) >> g) - ((kz - 1) >> g) + 1;
        int ncell = ncx * ncy * ncz;
        int best_id = 0x7FFFFFFF;
        unsigned long long best_slot = 0;
        for (int ci = lane; ci < ncell; ci += 32) {
            int ix = ci / (ncy * ncz), iy = (ci / ncz) % ncy, iz = ci % ncz;
            uint32_t b = cell_bucket_rw(m, ((kx - 1) >> g) + ix, ((ky - 1) >> g) + iy, ((kz - 1) >> g) + iz);
            while (b != NONE32) {
                for (int s = 0; s < BUCKET_SLOTS; ++s) {
                    size_t slot = (size_t)b * BUCKET_SLOTS + s;
                    float4 x = ld_point(m, slot);
                    int id = __float_as_int(x.w);
                    if (id >= 0 && id < best_id && same_point3(x.x, x.y, x.z, p.x, p.y, p.z)) { best_id = id; best_slot = slot; }
                }
                b = ((volatile uint32_t*)m.bucket_next)[b];
            }
        }
        for (int o = 16; o; o >>= 1) {
            int oi = __shfl_xor_sync(FULL, best_id, o);
            unsigned long long os = __shfl_xor_sync(FULL, best_slot, o);
            if (oi < best_id) { best_id = oi; best_slot = os; }
        }
        if (lane == 0 && best_id != 0x7FFFFFFF) {
            float4 x = ld_point(m, (size_t)best_slot);
            m.buckets[best_slot].w = __int_as_float(EMPTY_ID);
            unsigned int r = atomicAdd(&cnt->removed_count, 1u);
            if (r < removed_cap) removed[r] = x;
            atomicAdd(&cnt->live, (unsigned long long)(-1ll));
            atomicAdd(&cnt->del_counter, 1);
        }
        __threadfence();
        __syncwarp();
    }
}

// Add_Point_Boxes: pick the logged removed points inside the boxes (compacting the log); they are
// re-inserted by the caller through the non-downsampled insert path with their original ids.
__global__ void k_select_removed(float4* removed, unsigned int count, const float* __restrict__ boxes, int nb,
                                 float4* out, unsigned int* misc /* [0]=picked, [1]=kept */, float4* kept) {
    unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 x = removed[i];
    bool in = false;
    for (int b = 0; b < nb && !in; ++b) {
        const float* B = boxes + 6 * b;
        in = B[0] <= x.x && B[3] > x.x && B[1] <= x.y && B[4] > x.y && B[2] <= x.z && B[5] > x.z;
    }
    if (in) out[atomicAdd(&misc[0], 1u)] = x;
    else kept[atomicAdd(&misc[1], 1u)] = x;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static size_t next_pow2(size_t v) {
    size_t p = 1;
    while (p < v) p <<= 1;
    return p;
}
static inline unsigned nblk(size_t n, int t) { return (unsigned)((n + t - 1) / t); }

MapStore::~MapStore() { release(); }

void MapStore::release() {
    cudaFree(v_.dir_keys); cudaFree(v_.cell_tab); cudaFree(v_.blk_mask); cudaFree(v_.buckets); cudaFree(v_.bucket_next);
    cudaFree(const_cast<BlockRec*>(v_.rec));
    cudaFree(v_.sup_keys); cudaFree(v_.sup_mask);
    v_ = MapView{};
    drop_add_graph();
    if (block_ev_) { cudaEventDestroy(block_ev_); block_ev_ = nullptr; }
    cudaFree(d_bjob_); d_bjob_ = nullptr;
    if (h_bjob_) cudaFreeHost(h_bjob_);
    h_bjob_ = nullptr;
    cudaFree(d_cnt_); d_cnt_ = nullptr;
    if (h_pinned_) cudaFreeHost(h_pinned_);
    h_pinned_ = nullptr;
    cudaFree(d_removed_); d_removed_ = nullptr;
    cudaFree(scratch_pts_); cudaFree(keys_a_); cudaFree(keys_b_); cudaFree(idx_a_); cudaFree(idx_b_);
    cudaFree(defer_list_); cudaFree(defer_flag_); cudaFree(seg_); cudaFree(slop_list_); cudaFree(cub_tmp_); cudaFree(d_misc_);
    scratch_pts_ = nullptr; keys_a_ = keys_b_ = nullptr; idx_a_ = idx_b_ = defer_list_ = nullptr;
    defer_flag_ = nullptr; seg_ = nullptr; slop_list_ = nullptr; cub_tmp_ = nullptr; d_misc_ = nullptr;
    scratch_n_ = 0; cub_tmp_bytes_ = 0; dir_cap_ = bucket_cap_ = 0;
}

int MapStore::alloc_tables(size_t dir_cap, size_t bucket_cap, std::string& err) {
    CK(cudaMalloc(&v_.dir_keys, dir_cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&v_.cell_tab, dir_cap * 64 * sizeof(uint32_t)));
    CK(cudaMalloc(&v_.blk_mask, dir_cap * sizeof(unsigned long long)));
    CK(cudaMalloc((void**)&v_.rec, dir_cap * sizeof(BlockRec)));
    CK(cudaMemsetAsync((void*)v_.rec, 0xFF, dir_cap * sizeof(BlockRec), stream_));   // key = DIR_EMPTY everywhere
    CK(cudaMalloc(&v_.sup_keys, dir_cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&v_.sup_mask, dir_cap * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(v_.sup_keys, 0xFF, dir_cap * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.sup_mask, 0, dir_cap * sizeof(unsigned long long), stream_));
    CK(cudaMalloc(&v_.buckets, bucket_cap * BUCKET_SLOTS * sizeof(float4)));
    CK(cudaMalloc(&v_.bucket_next, bucket_cap * sizeof(uint32_t)));
    CK(cudaMemsetAsync(v_.dir_keys, 0xFF, dir_cap * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.cell_tab, 0xFF, dir_cap * 64 * sizeof(uint32_t), stream_));
    CK(cudaMemsetAsync(v_.blk_mask, 0, dir_cap * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.buckets, 0xFF, bucket_cap * BUCKET_SLOTS * sizeof(float4), stream_));
    CK(cudaMemsetAsync(v_.bucket_next, 0xFF, bucket_cap * sizeof(uint32_t), stream_));
    dir_cap_ = dir_cap;
    bucket_cap_ = bucket_cap;
    v_.dir_mask = (uint32_t)(dir_cap - 1);
    v_.bucket_cap = (uint32_t)bucket_cap;
    return 0;
}

int MapStore::init(cudaStream_t stream, float ds, int cell_ratio, size_t initial_points, int ivox_nearby, std::string& err) {
    stream_ = stream;
    if (!(ds > 0.f)) { err = "downsample_size must be > 0"; return -2; }
    if (ivox_nearby != 0) {
        // iVox mode: ds is IVox::Options::resolution_, one cell per voxel, Pos2Grid keys (flivo_dev.cuh)
        if (ivox_nearby != 1 && ivox_nearby != 7 && ivox_nearby != 19 && ivox_nearby != 27) { err = "ivox_nearby must be 0, 1, 7, 19 or 27"; return -2; }
        cell_ratio = 1;
        v_.ivox = ivox_nearby;
        v_.inv_res = (float)(1.0 / (double)ds);
    }
    if (cell_ratio != 1 && cell_ratio != 2 && cell_ratio != 4) { err = "cell_ratio must be 1, 2 or 4"; return -2; }
    cell_ratio_ = cell_ratio;
    v_.ds = ds;
    v_.gshift = cell_ratio == 1 ? 0 : (cell_ratio == 2 ? 1 : 2);
    v_.cell_edge = ds * (float)cell_ratio;
    if (initial_points == 0) initial_points = 1u << 20;
    // surface-like maps: ~4 points per cell, ~16 cells per block; keep the directory load factor <= ~0.4
    size_t dir_cap = next_pow2(std::max<size_t>(1u << 16, initial_points / 16));
    size_t bucket_cap = std::max<size_t>(1u << 16, initial_points / 2);
    if (alloc_tables(dir_cap, bucket_cap, err)) return -1;
    CK(cudaMalloc(&d_cnt_, sizeof(MapCounters)));
    CK(cudaMemsetAsync(d_cnt_, 0, sizeof(MapCounters), stream_));
    CK(cudaMallocHost(&h_pinned_, sizeof(MapCounters)));
    memset(&h_cnt_, 0, sizeof(h_cnt_));
    removed_cap_ = 1u << 20;
    CK(cudaMalloc(&d_removed_, removed_cap_ * sizeof(float4)));
    CK(cudaMalloc(&d_misc_, 16 * sizeof(unsigned int)));
    CK(cudaMalloc(&slop_list_, SLOP_LIST_CAP * sizeof(float4)));
    next_id_ = 0;
    blocks_ub_ = buckets_ub_ = 0;
    return 0;
}

int MapStore::refresh_counters(std::string& err) {
    CK(cudaMemcpyAsync(h_pinned_, d_cnt_, sizeof(MapCounters), cudaMemcpyDeviceToHost, stream_));
    CK(cudaStreamSynchronize(stream_));
    return finish_counters(err);
}
int MapStore::finish_counters(std::string& err) {
    h_cnt_ = *h_pinned_;
    blocks_ub_ = h_cnt_.blocks_used;
    buckets_ub_ = h_cnt_.buckets_used;
    if (h_cnt_.error) {
        int e = h_cnt_.error;
        err = e == 1 ? "point coordinate not finite or outside the addressable range" :
              e == 2 ? "block directory full" : "bucket pool exhausted";
        cudaMemsetAsync(&d_cnt_->error, 0, sizeof(int), stream_);
        return e == 1 ? -2 : -3;
    }
    return 0;
}

int MapStore::grow_dir(size_t new_cap, std::string& err) {
    MapView oldv = v_;
    size_t old_cap = dir_cap_;
    MapView nv = v_;
    CK(cudaMalloc(&nv.dir_keys, new_cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&nv.cell_tab, new_cap * 64 * sizeof(uint32_t)));
    CK(cudaMalloc(&nv.blk_mask, new_cap * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(nv.dir_keys, 0xFF, new_cap * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(nv.cell_tab, 0xFF, new_cap * 64 * sizeof(uint32_t), stream_));
    CK(cudaMemsetAsync(nv.blk_mask, 0, new_cap * sizeof(unsigned long long), stream_));
    CK(cudaMalloc((void**)&nv.rec, new_cap * sizeof(BlockRec)));
    CK(cudaMalloc(&nv.sup_keys, new_cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&nv.sup_mask, new_cap * sizeof(unsigned long long)));
    CK(cudaMemsetAsync(nv.sup_keys, 0xFF, new_cap * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(nv.sup_mask, 0, new_cap * sizeof(unsigned long long), stream_));
    nv.dir_mask = (uint32_t)(new_cap - 1);
    k_rehash<<<nblk(old_cap * 32, 256), 256, 0, stream_>>>(oldv, old_cap, nv, d_cnt_);
    k_all_records<<<nblk(new_cap, 256), 256, 0, stream_>>>(nv, new_cap, const_cast<BlockRec*>(nv.rec));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(stream_));
    cudaFree(oldv.dir_keys); cudaFree(oldv.cell_tab); cudaFree(oldv.blk_mask); cudaFree(const_cast<BlockRec*>(oldv.rec));
    cudaFree(oldv.sup_keys); cudaFree(oldv.sup_mask);
    v_ = nv;
    dir_cap_ = new_cap;
    return 0;
}

int MapStore::grow_buckets(size_t new_cap, std::string& err) {
    float4* nb = nullptr;
    uint32_t* nn = nullptr;
    CK(cudaMalloc(&nb, new_cap * BUCKET_SLOTS * sizeof(float4)));
    CK(cudaMalloc(&nn, new_cap * sizeof(uint32_t)));
    CK(cudaMemcpyAsync(nb, v_.buckets, bucket_cap_ * BUCKET_SLOTS * sizeof(float4), cudaMemcpyDeviceToDevice, stream_));
    CK(cudaMemcpyAsync(nn, v_.bucket_next, bucket_cap_ * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream_));
    CK(cudaMemsetAsync(nb + bucket_cap_ * BUCKET_SLOTS, 0xFF, (new_cap - bucket_cap_) * BUCKET_SLOTS * sizeof(float4), stream_));
    CK(cudaMemsetAsync(nn + bucket_cap_, 0xFF, (new_cap - bucket_cap_) * sizeof(uint32_t), stream_));
    CK(cudaStreamSynchronize(stream_));
    cudaFree(v_.buckets); cudaFree(v_.bucket_next);
    v_.buckets = nb; v_.bucket_next = nn;
    bucket_cap_ = new_cap;
    v_.bucket_cap = (uint32_t)new_cap;
    return 0;
}

int MapStore::ensure_capacity(size_t new_blocks_ub, size_t new_buckets_ub, std::string& err) {
    size_t need_dir = (blocks_ub_ + new_blocks_ub) * 5 / 2 + 64;  // load factor <= 0.4
    if (need_dir > dir_cap_) {
        size_t nc = next_pow2(need_dir);
        if (nc > DIR_CAP_MAX) { err = "block directory would exceed 2^25 slots"; return -3; }
        if (grow_dir(nc, err)) return -1;
    }
    size_t need_b = buckets_ub_ + new_buckets_ub + 64;
    if (need_b > bucket_cap_) {
        size_t nc = std::max(need_b, bucket_cap_ * 2);
        if (nc >= 0xFFFFFFF0ull) { err = "bucket pool would exceed 2^32 buckets"; return -3; }
        if (grow_buckets(nc, err)) return -1;
    }
    blocks_ub_ += new_blocks_ub;
    buckets_ub_ += new_buckets_ub;
    return 0;
}

int MapStore::ensure_scratch(size_t n, std::string& err) {
    if (n <= scratch_n_) return 0;
    size_t cap = std::max<size_t>(n, 1u << 16);
    cap = cap + cap / 4;
    cudaFree(scratch_pts_); cudaFree(keys_a_); cudaFree(keys_b_); cudaFree(idx_a_); cudaFree(idx_b_);
    cudaFree(defer_list_); cudaFree(defer_flag_); cudaFree(seg_); cudaFree(cub_tmp_);
    CK(cudaMalloc(&scratch_pts_, cap * sizeof(float4)));
    CK(cudaMalloc(&keys_a_, cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&keys_b_, cap * sizeof(unsigned long long)));
    CK(cudaMalloc(&idx_a_, cap * sizeof(unsigned int)));
    CK(cudaMalloc(&idx_b_, cap * sizeof(unsigned int)));
    CK(cudaMalloc(&defer_list_, cap * sizeof(unsigned int)));
    CK(cudaMalloc(&defer_flag_, cap));
    CK(cudaMalloc(&seg_, cap * sizeof(SegInfo)));
    size_t b1 = 0, b2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, b1, keys_a_, keys_b_, idx_a_, idx_b_, (int)std::min<size_t>(cap, 0x7FFFFFFF), 0, 63, stream_);
    cub::DeviceSelect::Flagged(nullptr, b2, thrust::counting_iterator<unsigned int>(0), defer_flag_, defer_list_, d_misc_ + 3,
                               (int)std::min<size_t>(cap, 0x7FFFFFFF), stream_);
    size_t b3 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, b3, idx_a_, idx_b_, (int)std::min<size_t>(cap, 0x7FFFFFFF), stream_);
    cub_tmp_bytes_ = std::max(std::max(b1, b2), b3) + 256;
    CK(cudaMalloc(&cub_tmp_, cub_tmp_bytes_));
    scratch_n_ = cap;
    return 0;
}

int MapStore::stamp_ids(const float4* d_pts, size_t n, int base_id, std::string& err) {
    if (ensure_scratch(n, err)) return -1;
    k_stamp_ids<<<nblk(n, 256), 256, 0, stream_>>>(d_pts, scratch_pts_, n, base_id, nullptr);
    CK(cudaGetLastError());
    return 0;
}

// d_pts_id must be scratch_pts_ or any device buffer that stays valid during the call
int MapStore::insert_batch(const float4* d_pts_id, size_t n, bool downsample_on, int* n_added, std::string& err) {
    if (n_added) *n_added = 0;
    if (n == 0) return 0;
    if (n > 0x7FFFFFF0ull) { err = "batch too large"; return -2; }
    if (ensure_scratch(n, err)) return -1;
    const int T = 256;
    // sizing: exact unique counts for large batches (one extra sync), conservative bound otherwise
    size_t new_blocks = n, new_cells = n;
    if (n > ADD_GRAPH_MAX) {
        k_point_keys<<<nblk(n, T), T, 0, stream_>>>(d_pts_id, n, v_, keys_a_, idx_a_, d_cnt_, nullptr, 0);
        CK(cudaGetLastError());
        size_t tmp = cub_tmp_bytes_;
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_, tmp, keys_a_, keys_b_, idx_a_, idx_b_, (int)n, 0, 63, stream_));
        CK(cudaMemsetAsync(d_misc_, 0, 16 * sizeof(unsigned int), stream_));
        k_count_unique<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, d_misc_);
        CK(cudaGetLastError());
        unsigned int h[2];
        CK(cudaMemcpyAsync(h, d_misc_, sizeof(h), cudaMemcpyDeviceToHost, stream_));
        CK(cudaStreamSynchronize(stream_));
        new_blocks = h[0];
        new_cells = h[1];
    }
    size_t chain_extra = downsample_on ? 64 : n / BUCKET_SLOTS + 64;
    if (ensure_capacity(new_blocks, new_cells + chain_extra, err)) return -1;
    if (enqueue_insert(d_pts_id, n, downsample_on, nullptr, n > ADD_GRAPH_MAX, err)) return -1;
    int rc = refresh_counters(err);
    if (n_added) *n_added = h_cnt_.add_counter;
    CK(cudaMemsetAsync(&d_cnt_->add_counter, 0, sizeof(int), stream_));
    CK(cudaMemsetAsync(&d_cnt_->deferred, 0, sizeof(int), stream_));
    return rc;
}

int MapStore::enqueue_insert(const float4* d_pts_id, size_t n, bool downsample_on, const BatchJob* job, bool keys_sorted, std::string& err) {
    const int T = 256;
    size_t tmp = cub_tmp_bytes_;
    if (!keys_sorted) {   // (a large batch was keyed by coordinates and sorted by the caller already, to size the tables)
        // scan-sized batch: the directory has room for every block it can create (ensure_capacity ran), so the keys name blocks
        // by directory slot and the sort covers 13 + log2(dir_cap) bits (the top one puts the padding's invalid keys last)
        static const bool slot_on = !(getenv("FLIVO_SLOT_KEYS") && atoi(getenv("FLIVO_SLOT_KEYS")) == 0);
        int end_bit = 63;
        if (slot_on) { end_bit = 13; for (size_t c = dir_cap_; c > 1; c >>= 1) ++end_bit; }
        k_point_keys<<<nblk(n, T), T, 0, stream_>>>(d_pts_id, n, v_, keys_a_, idx_a_, d_cnt_, job, slot_on ? 1 : 0);
        CK(cudaGetLastError());
        CK(cub::DeviceRadixSort::SortPairs(cub_tmp_, tmp, keys_a_, keys_b_, idx_a_, idx_b_, (int)n, 0, end_bit, stream_));
    }
    CK(cudaMemsetAsync(d_misc_, 0, 16 * sizeof(unsigned int), stream_));
    // cells of the batch (new ones get consecutive buckets in key order), then the search records of the touched blocks;
    // idx_a_ / keys_a_ are free once the sort has produced keys_b_ / idx_b_
    unsigned int* is_new = idx_a_;
    unsigned int* rank = reinterpret_cast<unsigned int*>(keys_a_);
    k_cells_mark<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, v_, d_cnt_, is_new);
    tmp = cub_tmp_bytes_;
    CK(cub::DeviceScan::ExclusiveSum(cub_tmp_, tmp, is_new, rank, (int)n, stream_));
    k_cells_reserve<<<1, 32, 0, stream_>>>(is_new, rank, n, v_, d_cnt_, d_misc_);
    k_cells_create<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, v_, d_cnt_, is_new, rank, d_misc_, downsample_on ? 0 : 1);
    CK(cudaGetLastError());
    if (!downsample_on) {
        k_insert_all<<<nblk(n, T), T, 0, stream_>>>(d_pts_id, n, v_, d_cnt_);
        CK(cudaGetLastError());
    } else {
        CK(cudaMemsetAsync(defer_flag_, 0, n, stream_));
        k_mark_slop<<<nblk(n, T), T, 0, stream_>>>(d_pts_id, n, v_, slop_list_, d_misc_);
        k_ds_decide<<<nblk(n, 128), 128, 0, stream_>>>(keys_b_, idx_b_, d_pts_id, n, v_, seg_, slop_list_, d_misc_, defer_flag_, d_cnt_);
        k_ds_clear<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, v_, seg_, d_cnt_);
        k_ds_store<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, d_pts_id, v_, seg_, d_cnt_);
        CK(cudaGetLastError());
        tmp = cub_tmp_bytes_;
        CK(cub::DeviceSelect::Flagged(cub_tmp_, tmp, thrust::counting_iterator<unsigned int>(0), defer_flag_, defer_list_,
                                      d_misc_ + 3, (int)n, stream_));
        k_ds_serial<<<1, 32, 0, stream_>>>(defer_list_, d_misc_, d_pts_id, v_, d_cnt_);
        CK(cudaGetLastError());
    }
    // search records of the touched blocks, once the batch is stored (storing may still append overflow buckets)
    k_block_records<<<nblk(n, T), T, 0, stream_>>>(keys_b_, n, v_, const_cast<BlockRec*>(v_.rec));
    CK(cudaGetLastError());
    return 0;
}

int MapStore::wait_stream(std::string& err) {
    if (!block_wait_) { CK(cudaStreamSynchronize(stream_)); return 0; }
    if (!block_ev_) CK(cudaEventCreateWithFlags(&block_ev_, cudaEventBlockingSync | cudaEventDisableTiming));
    CK(cudaEventRecord(block_ev_, stream_));
    CK(cudaEventSynchronize(block_ev_));
    return 0;
}

void MapStore::drop_add_graph() {
    if (add_exec_) { cudaGraphExecDestroy(add_exec_); add_exec_ = nullptr; }
    add_src_ = nullptr; add_npad_ = 0; add_scratch_ = nullptr;
}

// Add_Points(…, true) of a scan-sized batch: the whole sequence (job block H2D, ids, keys, sort, cells, down-sample decision,
// store, records, counters D2H, counter reset) replayed as one graph.  The graph is rebuilt when the padded size, the
// source buffer, the scratch buffers or the map's tables (growth) change.
int MapStore::add_points_graph(const float4* d_pts, size_t n, int* n_added, std::string& err) {
    const size_t n_pad = ((n + ADD_PAD - 1) / ADD_PAD) * ADD_PAD;
    if (ensure_scratch(n_pad, err)) return -1;
    if (ensure_capacity(n_pad, n_pad + 64, err)) return -1;
    if (!d_bjob_) {
        CK(cudaMalloc(&d_bjob_, sizeof(BatchJob)));
        CK(cudaMallocHost(&h_bjob_, sizeof(BatchJob)));
    }
    h_bjob_->n = (unsigned int)n;
    h_bjob_->base_id = next_id_;
    if (!add_exec_ || add_src_ != d_pts || add_npad_ != n_pad || add_scratch_ != scratch_pts_ || memcmp(&add_view_, &v_, sizeof(MapView)) != 0) {
        drop_add_graph();
        cudaGraph_t g = nullptr;
        CK(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
        cudaError_t ce = cudaMemcpyAsync(d_bjob_, h_bjob_, sizeof(BatchJob), cudaMemcpyHostToDevice, stream_);
        k_stamp_ids<<<nblk(n_pad, 256), 256, 0, stream_>>>(d_pts, scratch_pts_, n_pad, 0, d_bjob_);
        std::string cerr;
        int rc = enqueue_insert(scratch_pts_, n_pad, true, d_bjob_, false, cerr);
        if (ce == cudaSuccess) ce = cudaMemcpyAsync(h_pinned_, d_cnt_, sizeof(MapCounters), cudaMemcpyDeviceToHost, stream_);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(&d_cnt_->add_counter, 0, sizeof(int), stream_);
        if (ce == cudaSuccess) ce = cudaMemsetAsync(&d_cnt_->deferred, 0, sizeof(int), stream_);
        cudaError_t ce2 = cudaStreamEndCapture(stream_, &g);
        if (rc || ce != cudaSuccess || ce2 != cudaSuccess) {
            if (g) cudaGraphDestroy(g);
            cudaGetLastError();
            err = "capture of the Add_Points graph: " + (rc ? cerr : std::string(cudaGetErrorString(ce != cudaSuccess ? ce : ce2)));
            return -4;   // the caller falls back to plain launches
        }
        ce = cudaGraphInstantiate(&add_exec_, g, 0);
        cudaGraphDestroy(g);
        if (ce != cudaSuccess) { add_exec_ = nullptr; cudaGetLastError(); err = std::string("cudaGraphInstantiate (Add_Points): ") + cudaGetErrorString(ce); return -4; }
        add_src_ = d_pts; add_npad_ = n_pad; add_scratch_ = scratch_pts_; add_view_ = v_;
    }
    CK(cudaGraphLaunch(add_exec_, stream_));
    if (wait_stream(err)) return -1;
    next_id_ += (int)n;
    int rc = finish_counters(err);
    if (n_added) *n_added = h_cnt_.add_counter;
    return rc;
}

int MapStore::build(const float4* d_pts, size_t n, std::string& err) {
    // KD_TREE::Build deletes the old tree first (ikd_Tree.cpp:338-340)
    CK(cudaMemsetAsync(v_.dir_keys, 0xFF, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.cell_tab, 0xFF, dir_cap_ * 64 * sizeof(uint32_t), stream_));
    CK(cudaMemsetAsync(v_.blk_mask, 0, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync((void*)v_.rec, 0xFF, dir_cap_ * sizeof(BlockRec), stream_));
    CK(cudaMemsetAsync(v_.sup_keys, 0xFF, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.sup_mask, 0, dir_cap_ * sizeof(unsigned long long), stream_));
    size_t used = std::min<size_t>(bucket_cap_, (size_t)buckets_ub_ + 64);
    CK(cudaMemsetAsync(v_.buckets, 0xFF, used * BUCKET_SLOTS * sizeof(float4), stream_));
    CK(cudaMemsetAsync(v_.bucket_next, 0xFF, used * sizeof(uint32_t), stream_));
    CK(cudaMemsetAsync(d_cnt_, 0, sizeof(MapCounters), stream_));
    memset(&h_cnt_, 0, sizeof(h_cnt_));
    blocks_ub_ = buckets_ub_ = 0;
    next_id_ = 0;
    if (n == 0) return 0;
    if (stamp_ids(d_pts, n, 0, err)) return -1;
    next_id_ = (int)n;
    return insert_batch(scratch_pts_, n, false, nullptr, err);
}

int MapStore::add_points(const float4* d_pts, size_t n, bool downsample_on, int* n_added, std::string& err) {
    if (n_added) *n_added = 0;
    if (n == 0) return 0;
    if ((size_t)next_id_ + n > 0x7FFFFFF0ull) { err = "insertion ids exhausted (2^31)"; return -3; }
    if (v_.ivox && downsample_on) { err = "iVox mode: AddPoints keeps every point (no voxel down-sampling)"; return -2; }
    static const bool graph_on = !(getenv("FLIVO_ADD_GRAPH") && atoi(getenv("FLIVO_ADD_GRAPH")) == 0);
    if (downsample_on && graph_on && n <= ADD_GRAPH_MAX && add_graph_fail_ < 3) {
        const int rc = add_points_graph(d_pts, n, n_added, err);
        if (rc != -4) { add_graph_fail_ = 0; return rc; }
        ++add_graph_fail_;   // could not be captured: plain launches (and no further attempts after three failures in a row)
        err.clear();
    }
    if (stamp_ids(d_pts, n, next_id_, err)) return -1;
    next_id_ += (int)n;
    return insert_batch(scratch_pts_, n, downsample_on, n_added, err);
}

int MapStore::delete_boxes(const float* h_boxes6, size_t nb, int* n_deleted, std::string& err) {
    if (n_deleted) *n_deleted = 0;
    if (nb == 0) return 0;
    float* d_boxes = nullptr;
    CK(cudaMalloc(&d_boxes, nb * 6 * sizeof(float)));
    CK(cudaMemcpyAsync(d_boxes, h_boxes6, nb * 6 * sizeof(float), cudaMemcpyHostToDevice, stream_));
    k_delete_boxes<<<nblk(dir_cap_ * 32, 256), 256, 0, stream_>>>(v_, dir_cap_, d_boxes, (int)nb, d_cnt_, d_removed_, (unsigned)removed_cap_);
    CK(cudaGetLastError());
    int rc = refresh_counters(err);
    cudaFree(d_boxes);
    if (n_deleted) *n_deleted = h_cnt_.del_counter;
    int zero = 0;
    CK(cudaMemcpyAsync(&d_cnt_->del_counter, &zero, sizeof(int), cudaMemcpyHostToDevice, stream_));
    return rc;
}

int MapStore::delete_points(const float4* d_pts, size_t n, int* n_deleted, std::string& err) {
    if (n_deleted) *n_deleted = 0;
    if (n == 0) return 0;
    if (v_.ivox) { err = "iVox mode: Delete_Points is an ikd-Tree operation"; return -2; }
    k_delete_points<<<1, 32, 0, stream_>>>(d_pts, n, v_, d_cnt_, d_removed_, (unsigned)removed_cap_);
    CK(cudaGetLastError());
    int rc = refresh_counters(err);
    if (n_deleted) *n_deleted = h_cnt_.del_counter;
    int zero = 0;
    CK(cudaMemcpyAsync(&d_cnt_->del_counter, &zero, sizeof(int), cudaMemcpyHostToDevice, stream_));
    return rc;
}

int MapStore::add_boxes(const float* h_boxes6, size_t nb, int* n_revived, std::string& err) {
    if (n_revived) *n_revived = 0;
    if (refresh_counters(err)) return -1;
    unsigned int count = std::min<unsigned int>(h_cnt_.removed_count, (unsigned)removed_cap_);
    if (nb == 0 || count == 0) return 0;
    float* d_boxes = nullptr;
    float4 *picked = nullptr, *kept = nullptr;
    CK(cudaMalloc(&d_boxes, nb * 6 * sizeof(float)));
    CK(cudaMalloc(&picked, (size_t)count * sizeof(float4)));
    CK(cudaMalloc(&kept, (size_t)count * sizeof(float4)));
    CK(cudaMemcpyAsync(d_boxes, h_boxes6, nb * 6 * sizeof(float), cudaMemcpyHostToDevice, stream_));
    CK(cudaMemsetAsync(d_misc_, 0, 16 * sizeof(unsigned int), stream_));
    k_select_removed<<<nblk(count, 256), 256, 0, stream_>>>(d_removed_, count, d_boxes, (int)nb, picked, d_misc_, kept);
    CK(cudaGetLastError());
    unsigned int h[2];
    CK(cudaMemcpyAsync(h, d_misc_, sizeof(h), cudaMemcpyDeviceToHost, stream_));
    CK(cudaStreamSynchronize(stream_));
    CK(cudaMemcpyAsync(d_removed_, kept, (size_t)h[1] * sizeof(float4), cudaMemcpyDeviceToDevice, stream_));
    CK(cudaMemcpyAsync(&d_cnt_->removed_count, &h[1], sizeof(unsigned int), cudaMemcpyHostToDevice, stream_));
    int rc = 0;
    if (h[0]) rc = insert_batch(picked, h[0], false, nullptr, err);  // ids travel in .w
    CK(cudaStreamSynchronize(stream_));
    cudaFree(d_boxes); cudaFree(picked); cudaFree(kept);
    if (n_revived) *n_revived = (int)h[0];
    return rc;
}

int MapStore::flatten(float* h_xyz, int* h_id, size_t cap, size_t* n, std::string& err) {
    if (refresh_counters(err)) return -1;
    size_t live = (size_t)h_cnt_.live;
    if (n) *n = live;
    if (cap == 0 || live == 0 || (!h_xyz && !h_id)) return 0;
    float4* d_out = nullptr;
    CK(cudaMalloc(&d_out, live * sizeof(float4)));
    CK(cudaMemsetAsync(d_misc_, 0, 16 * sizeof(unsigned int), stream_));
    k_flatten<<<nblk(dir_cap_ * 32, 256), 256, 0, stream_>>>(v_, dir_cap_, d_out, (unsigned)live, d_misc_);
    CK(cudaGetLastError());
    std::vector<float4> h(live);
    CK(cudaMemcpyAsync(h.data(), d_out, live * sizeof(float4), cudaMemcpyDeviceToHost, stream_));
    CK(cudaStreamSynchronize(stream_));
    cudaFree(d_out);
    size_t m = std::min(cap, live);
    for (size_t i = 0; i < m; ++i) {
        if (h_xyz) { h_xyz[3 * i] = h[i].x; h_xyz[3 * i + 1] = h[i].y; h_xyz[3 * i + 2] = h[i].z; }
        if (h_id) memcpy(&h_id[i], &h[i].w, 4);
    }
    return 0;
}

int MapStore::range(float out6[6], std::string& err) {
    float init[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
    float* d = nullptr;
    CK(cudaMalloc(&d, sizeof(init)));
    CK(cudaMemcpyAsync(d, init, sizeof(init), cudaMemcpyHostToDevice, stream_));
    k_range<<<nblk(dir_cap_ * 32, 256), 256, 0, stream_>>>(v_, dir_cap_, d);
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out6, d, sizeof(init), cudaMemcpyDeviceToHost, stream_));
    CK(cudaStreamSynchronize(stream_));
    cudaFree(d);
    if (!(out6[0] <= out6[3])) for (int i = 0; i < 6; ++i) out6[i] = 0.f;  // empty tree: memset 0 (ikd_Tree.cpp:104-106)
    return 0;
}

int MapStore::acquire_removed(float* h_xyz, size_t cap, size_t* n, std::string& err) {
    if (refresh_counters(err)) return -1;
    size_t count = std::min<size_t>(h_cnt_.removed_count, removed_cap_);
    if (n) *n = count;
    if (cap == 0 || !h_xyz) return 0;   // size query only; the log is kept
    std::vector<float4> h(count);
    if (count) {
        CK(cudaMemcpyAsync(h.data(), d_removed_, count * sizeof(float4), cudaMemcpyDeviceToHost, stream_));
        CK(cudaStreamSynchronize(stream_));
    }
    size_t m = std::min(cap, count);
    for (size_t i = 0; i < m; ++i) { h_xyz[3 * i] = h[i].x; h_xyz[3 * i + 1] = h[i].y; h_xyz[3 * i + 2] = h[i].z; }
    unsigned int zero = 0;
    CK(cudaMemcpyAsync(&d_cnt_->removed_count, &zero, sizeof(unsigned int), cudaMemcpyHostToDevice, stream_));
    CK(cudaStreamSynchronize(stream_));
    return 0;
}

int MapStore::set_downsample(float ds, std::string& err) {
    if (!(ds > 0.f)) { err = "downsample_size must be > 0"; return -2; }
    if (v_.ivox) { err = "iVox mode: the voxel resolution is fixed at context creation"; return -2; }
    if (ds == v_.ds) return 0;
    if (refresh_counters(err)) return -1;
    size_t live = (size_t)h_cnt_.live;
    float4* d_all = nullptr;
    if (live) {
        CK(cudaMalloc(&d_all, live * sizeof(float4)));
        CK(cudaMemsetAsync(d_misc_, 0, 16 * sizeof(unsigned int), stream_));
        k_flatten<<<nblk(dir_cap_ * 32, 256), 256, 0, stream_>>>(v_, dir_cap_, d_all, (unsigned)live, d_misc_);
        CK(cudaGetLastError());
    }
    int keep_next = next_id_;
    v_.ds = ds;
    v_.cell_edge = ds * (float)cell_ratio_;
    std::string e2;
    // re-index: clear tables, re-insert the live points with their ids
    CK(cudaMemsetAsync(v_.dir_keys, 0xFF, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.cell_tab, 0xFF, dir_cap_ * 64 * sizeof(uint32_t), stream_));
    CK(cudaMemsetAsync(v_.blk_mask, 0, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync((void*)v_.rec, 0xFF, dir_cap_ * sizeof(BlockRec), stream_));
    CK(cudaMemsetAsync(v_.sup_keys, 0xFF, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.sup_mask, 0, dir_cap_ * sizeof(unsigned long long), stream_));
    CK(cudaMemsetAsync(v_.buckets, 0xFF, bucket_cap_ * BUCKET_SLOTS * sizeof(float4), stream_));
    CK(cudaMemsetAsync(v_.bucket_next, 0xFF, bucket_cap_ * sizeof(uint32_t), stream_));
    unsigned int rc_keep = h_cnt_.removed_count;
    CK(cudaMemsetAsync(d_cnt_, 0, sizeof(MapCounters), stream_));
    CK(cudaMemcpyAsync(&d_cnt_->removed_count, &rc_keep, sizeof(unsigned int), cudaMemcpyHostToDevice, stream_));
    blocks_ub_ = buckets_ub_ = 0;
    int rc = 0;
    if (live) rc = insert_batch(d_all, live, false, nullptr, err);
    CK(cudaStreamSynchronize(stream_));
    cudaFree(d_all);
    next_id_ = keep_next;
    return rc;
}

}  // namespace flivo
