// api.cu — the C ABI of libflivo_b200.so (include/flivo.h): context, host<->device staging, launch sequences.
//
// The reference has no FFI layer; each entry point stands in for the C++ member cited beside its
// declaration in include/flivo.h.  Everything here is plumbing: the arithmetic lives in map.cu (GPU map),
// search.cuh + fused.cu (per-trip fused kernel) and solve.cu (on-device reduce + solve + policy).
// There is no CPU fallback: without a CUDA device every call fails with FLIVO_E_CUDA.
#include "../../include/flivo.h"
#include "map.cuh"
#include "solve.cuh"
#include "stage.cuh"
#include "frontend.cuh"

#include <cub/cub.cuh>
#include <dlfcn.h>
#include <omp.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <string>
#include <vector>

using namespace flivo;

// ---- NCCL through dlopen (no link-time dependency; the process may already carry torch's libnccl) -------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId_t;
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(ncclUniqueId_t*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId_t, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load(std::string& err) {
        if (h) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* nm : names) { h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) { err = std::string("dlopen libnccl failed: ") + dlerror(); return false; }
        GetUniqueId = (decltype(GetUniqueId))dlsym(h, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(h, "ncclCommInitRank");
        CommDestroy = (decltype(CommDestroy))dlsym(h, "ncclCommDestroy");
        AllGather = (decltype(AllGather))dlsym(h, "ncclAllGather");
        GetErrorString = (decltype(GetErrorString))dlsym(h, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !CommDestroy || !AllGather) { err = "libnccl lacks a required symbol"; return false; }
        return true;
    }
};
static NcclApi g_nccl;
constexpr int NCCL_DOUBLE = 8;  // ncclFloat64 (nccl.h ncclDataType_t)

// Device block of one update.  [job .. x] is the H2D prefix of a call, [x .. stats] the D2H range.
struct DevBlock {
    Job job;
    double P_prop[23 * 23];
    double x[36];
    double P_out[23 * 23];
    Ctl ctl;
    SearchStats stats;
    double x_prop[36];
    double sum[PSTRIDE];
};

constexpr int SSTAMP_CTAS = 4096;

// One of the two landing buffers of flivo_scan_upload_async: while the update of the current scan runs on the context's
// stream, the next scan travels host -> device on the copy stream into the other slot.
struct ScanSlot {
    char* d_in = nullptr;           // device copy of the caller's records (their own stride when the source is pinned
    size_t d_bytes = 0;             //   memory, else the float4 repack)
    float4* h_pin = nullptr;        // pinned repack buffer (pageable sources only)
    size_t h_n = 0;
    cudaEvent_t copied = nullptr;   // the H2D copy into d_in has finished (copy stream)
    cudaEvent_t consumed = nullptr; // the staging sequence that reads d_in has run (context stream)
    bool copy_pending = false, filled = false;
    size_t n = 0, stride = 0;
    size_t n_a = 0, stride_b = 0;   // split upload: records [0, n_a) are float4 at d_in, [n_a, n) raw records of stride_b behind them
};

struct flivo_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    MapStore map;
    std::string err;
    flivo_options opt{};
    int sm_count = 148;
    // host staging (pinned)
    float4* h_stage = nullptr;
    size_t h_stage_n = 0;
    cudaEvent_t stage_ev = nullptr;   // last H2D copy out of h_stage
    bool stage_pending = false;
    double* h_small = nullptr;      // pinned: 36 state + 36 + 529 P + ... (2048 doubles)
    // generic device staging for map / knn inputs
    float4* d_stage = nullptr;
    size_t d_stage_n = 0;
    // scan
    size_t scan_n = 0, scan_cap = 0;
    bool scan_external = false;
    bool scan_sorted = false;
    float4* d_body = nullptr;       // processing order: d_body_own or the caller's device pointer
    float4* d_body_own = nullptr;
    float4* d_body_in = nullptr;    // caller's order, float4 (written by k_scan_prepare)
    ScanSlot slot[2];               // flivo_scan_upload_async landing buffers; `back` is the one the next upload fills
    int back = 0;
    cudaStream_t copy = nullptr;    // H2D copies of scans, beside the update running on `stream`
    // a committed / set scan is staged (keys, sort, gather, flag reset) lazily, right before its first consumer, so that
    // the staging graph and the update graph are launched back to back from inside one call
    bool stage_todo = false;
    const char* todo_src = nullptr, *todo_src_b = nullptr;
    size_t todo_stride = 0, todo_stride_b = 0, todo_n_a = 0;
    int todo_slot = -1;
    cudaEvent_t ev_stage0 = nullptr, ev_upd0 = nullptr, ev_upd1 = nullptr;   // in-graph time stamps (external event nodes)
    // an update launched by flivo_iekf_update_begin and not yet collected by flivo_iekf_update_end
    struct { bool active = false; int np = 0, trips = 0, launches = 0; bool timed = false, staged = false; } upd;
    int graph_skip = 0;             // calls left on plain launches after the graph cache kept missing
    unsigned int useq = 0;          // updates started on this context (Job::useq)
    bool block_wait = false;        // the update's final wait sleeps on block_ev instead of spinning (set by flivo_frame)
    cudaEvent_t block_ev = nullptr;
    double frame_s[5] = {0, 0, 0, 0, 0};   // FLIVO_FRAME_TRACE=1: host seconds spent in flivo_frame's commit / begin / upload / end / map_add
    long frame_calls = 0;
    unsigned int* d_perm = nullptr; // processing index -> caller's index
    unsigned long long *d_skey_a = nullptr, *d_skey_b = nullptr;
    unsigned int* d_sidx_a = nullptr;
    int stage_launches = 0;                   // own kernels the last staging sequence launched
    void* d_cub = nullptr;
    size_t cub_bytes = 0;
    ScanJob* d_sjob = nullptr;          // staging scalars (device) and their pinned mirror
    ScanJob* h_sjob = nullptr;
    cudaGraphExec_t sort_exec = nullptr;   // captured staging sequence, keyed by (n_pad, sort bits)
    unsigned int sort_npad = 0;
    cudaEvent_t sjob_ev = nullptr;      // the last staging graph has consumed h_sjob
    bool sjob_pending = false;
    float4* d_near = nullptr;
    float2* d_near_meta = nullptr;
    unsigned char *d_sel = nullptr, *d_eff = nullptr;
    float4 *d_normvec = nullptr, *d_world = nullptr;
    double *d_partials = nullptr, *d_gpartials = nullptr;
    unsigned int* d_group_cnt = nullptr;
    size_t group_slots = 0;
    unsigned long long* d_far_list = nullptr;
    float4* d_plane = nullptr;
    unsigned char* d_pflags = nullptr;
    // front end (raw scan -> undistort -> voxel grid), see frontend.cu
    float4* d_raw = nullptr;        // x, y, z, intensity
    float* d_raw_t = nullptr;       // per-point time offset in ms (PointXYZINormal::curvature)
    size_t raw_n = 0, raw_cap = 0;
    bool raw_has_t = false;
    float4* d_ds = nullptr;         // voxel-grid output (x, y, z, intensity), leaf order
    float* d_ds_t = nullptr;
    unsigned *d_vg_keys_a = nullptr, *d_vg_keys_b = nullptr, *d_vg_idx_a = nullptr, *d_vg_idx_b = nullptr, *d_vg_head = nullptr, *d_vg_pos = nullptr;
    void* d_vg_tmp = nullptr;
    size_t vg_tmp_bytes = 0;
    VoxelGridState* d_vg_state = nullptr;
    PoseDev* d_poses = nullptr;
    void* d_out = nullptr;          // un-permute scratch for optional outputs
    size_t d_out_bytes = 0;
    // loop state
    DevBlock* d_blk = nullptr;      // job + state + P + ctl + stats in one allocation (see DevBlock)
    DevBlock* h_blk = nullptr;      // pinned mirror
    Ctl* d_ctl = nullptr;           // the pointers below point into d_blk
    SearchStats* d_stats = nullptr;
    unsigned long long* d_stamps = nullptr;   // [MAX_TRIPS][16] k_solve phase stamps (profile mode)
    unsigned long long* d_sstamps = nullptr;  // [SSTAMP_CTAS][4] k_search per-CTA stamps (profile mode)
    SolvePre* d_pre = nullptr;                // state-only half of a trip's solve (solve_pre -> next trip's solve_trip)
    double *d_x = nullptr, *d_xprop = nullptr, *d_Pprop = nullptr, *d_Pout = nullptr, *d_sum = nullptr, *d_rank_blocks = nullptr;
    // captured launch sequence of one update (CUDA graph), re-captured only when a kernel parameter changes
    cudaGraphExec_t graph_exec = nullptr;
    FusedParams graph_P{};
    SolveParams graph_S{};
    int search_resident = 0;        // CTAs of k_search resident on the whole GPU
    int graph_search_grid = 0, graph_fused_grid = 0, graph_trips = 0, graph_churn = 0;
    cudaEvent_t ev[2 * MAX_TRIPS + 2] = {};
    flivo_stats stats{};
    // GNSS observation of the next updates (flivo_set_gnss)
    bool gnss_on = false;
    double gnss[9] = {};
    // iVox mode (flivo_options::ivox_nearby != 0)
    double ivox_filter = 0.0;                 // filter_size_map_min of map_incremental's add rule
    unsigned char* d_ivox_flags = nullptr;    // [2][cap] add / no-need-downsample flags in the caller's order
    float4* d_ivox_sel = nullptr;             // [2][cap] the two selected point lists
    size_t ivox_cap = 0;
    // multi-GPU
    ncclComm_t comm = nullptr;
    int rank = 0, nranks = 1;
    // peer-memory exchange (NVLink): own buffer + the peers' buffers mapped through cudaIpc
    PeerSlot* d_xchg = nullptr;
    unsigned long long* d_xseq = nullptr;
    PeerSlot* peer_base[PEER_MAX_RANKS] = {};
    bool peer_on = false;
    bool peer_local = false;                  // peers attached by device pointer (same process), nothing to close
    bool peer_broken = false;                 // an exchange timed out: updates are refused until flivo_peer_reset
    unsigned long long peer_timeout_ns = 2000000000ull;
};

#define CKC(call)                                                                        \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);               \
            return FLIVO_E_CUDA;                                                         \
        }                                                                                \
    } while (0)
#define NEED(cond, code, msg) do { if (!(cond)) { if (ctx) ctx->err = (msg); return (code); } } while (0)

static int map_rc(flivo_ctx* ctx, int rc) {
    if (rc == 0) return FLIVO_OK;
    if (rc == -2) return FLIVO_E_ARG;
    if (rc == -3) return FLIVO_E_CAPACITY;
    return FLIVO_E_CUDA;
}

// strided host xyz -> pinned float4 -> device buffer (w = 0).  The repack of the caller's AoS (48-byte
// pcl::PointXYZINormal in the reference) runs on a few host threads, chunk by chunk, each chunk's H2D copy
// overlapping the repack of the next; the call does not wait for the copies (stream order protects the
// consumers, stage_ev protects the pinned buffer against the next call).
static int host_threads() {
    static int t = 0;
    if (t == 0) {
        const char* e = getenv("FLIVO_HOST_THREADS");
        if (e) {
            t = atoi(e);
        } else {
            // up to 16 repack threads (the repack is host-memory bound: a 260 k-point 48-byte AoS scan stages in ~125 us at
            // 16 threads, ~220 us at 8; a single parallel region with a copy-issuing thread was measured slower than
            // this chunk loop: profiles/r1_ab_launch_structure.txt), but never more than this process's share of the host when several ranks
            // (one per GPU, LOCAL_WORLD_SIZE of torchrun) stage scans at the same time: oversubscribed OpenMP
            // teams spin against each other.  (The reference's own loop runs MP_PROC_NUM = 4, CMakeLists.txt:33-36.)
            const char* lw = getenv("LOCAL_WORLD_SIZE");
            const int ranks = lw ? std::max(1, atoi(lw)) : 1;
            t = std::min(16, std::max(1, omp_get_num_procs() / ranks));
        }
        if (t < 1) t = 1;
        if (t > omp_get_num_procs()) t = omp_get_num_procs();
    }
    return t;
}
static int stage_points(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes, float4* dst) {
    if (n == 0) return FLIVO_OK;
    if (ctx->stage_pending) { CKC(cudaEventSynchronize(ctx->stage_ev)); ctx->stage_pending = false; }
    if (n > ctx->h_stage_n) {
        if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
        ctx->h_stage = nullptr; ctx->h_stage_n = 0;
        size_t cap = std::max<size_t>(n + n / 4, 1u << 16);
        CKC(cudaMallocHost(&ctx->h_stage, cap * sizeof(float4)));
        ctx->h_stage_n = cap;
    }
    const char* src = reinterpret_cast<const char*>(xyz);
    float4* h = ctx->h_stage;
    const size_t CH = 1u << 15;
    const int nt = host_threads();
    for (size_t c0 = 0; c0 < n; c0 += CH) {
        const size_t m = std::min(CH, n - c0);
        if (stride_bytes == 16) {
            memcpy(h + c0, src + c0 * 16, m * 16);
        } else {
#pragma omp parallel for num_threads(nt) schedule(static) if (m >= 8192 && nt > 1)
            for (long i = 0; i < (long)m; ++i) {
                const float* p = reinterpret_cast<const float*>(src + (c0 + (size_t)i) * stride_bytes);
                h[c0 + i] = make_float4(p[0], p[1], p[2], 0.f);
            }
        }
        CKC(cudaMemcpyAsync(dst + c0, h + c0, m * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    }
    CKC(cudaEventRecord(ctx->stage_ev, ctx->stream));
    ctx->stage_pending = true;
    return FLIVO_OK;
}
static int ensure_dstage(flivo_ctx* ctx, size_t n) {
    if (n <= ctx->d_stage_n) return FLIVO_OK;
    cudaFree(ctx->d_stage); ctx->d_stage = nullptr; ctx->d_stage_n = 0;
    size_t cap = std::max<size_t>(n + n / 4, 1u << 16);
    CKC(cudaMalloc(&ctx->d_stage, cap * sizeof(float4)));
    ctx->d_stage_n = cap;
    return FLIVO_OK;
}
static bool stride_ok(size_t s) { return s >= 12 && (s % 4) == 0; }

// ---- small kernels of the plumbing layer ------------------------------------------------------------------
__global__ void k_ctl_set(Ctl* ctl, int search) {
    if (threadIdx.x == 0) { ctl->search = search; ctl->stopped = 0; ctl_next_trip(ctl); }
}
// caller's records (any stride: the 48-byte PointXYZINormal AoS arrives as it is) -> float4 in the caller's order, plus,
// when the scan is sorted (keys != nullptr), the Morton key of the point's cell and the identity index
__global__ void k_scan_prepare(const ScanJob* __restrict__ J, float4* __restrict__ in4, unsigned long long* keys, unsigned int* idx) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const unsigned n = J->n;
    if (i >= J->n_pad) return;
    if (keys) idx[i] = (unsigned int)i;
    if (i >= n) { if (keys) keys[i] = 0xFFFFFFFFFFFFFFFFull; return; }   // padding: largest key, and (stable sort) behind every real point
    const float4 p = scan_record(J, i);
    in4[i] = p;
    if (keys) keys[i] = scan_morton(p, J->inv_edge);
}
// gather the scan into processing order and (when the arrays are given) reset the per-point flags of a fresh scan in
// the same pass: sel = 1, eff = 0, pflags = 0 (see finish_scan)
__global__ void k_gather4(const ScanJob* __restrict__ J, const float4* __restrict__ in4, const unsigned int* __restrict__ perm, float4* __restrict__ out,
                          unsigned char* __restrict__ sel, unsigned char* __restrict__ eff, unsigned char* __restrict__ pflags) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= J->n) return;
    out[i] = in4[perm[i]];
    sel[i] = 1; eff[i] = 0; pflags[i] = 0;
}
// out[perm[i]*elems + e] = in[i*elems + e]   (processing order -> caller's order); perm == nullptr: identity
template <class T>
__global__ void k_scatter(const T* __restrict__ in, const unsigned int* __restrict__ perm, size_t n, int elems, T* __restrict__ out) {
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * (size_t)elems) return;
    size_t i = t / elems;
    int e = (int)(t % elems);
    size_t o = perm ? perm[i] : i;
    out[o * elems + e] = in[t];
}
// neighbour ids / squared distances from the cached Nearest_Points (x,y,z,id) and the query
__global__ void k_near_unpack(const float4* __restrict__ near, const float4* __restrict__ world, const unsigned int* __restrict__ perm,
                              size_t n, int* out_id, float* out_sqd, int ivox) {
    size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (t >= n * 5) return;
    size_t i = t / 5;
    int r = (int)(t % 5);
    float4 p = near[t];
    int id = __float_as_int(p.w);
    size_t o = perm ? perm[i] : i;
    if (out_id) out_id[o * 5 + r] = id;
    if (out_sqd) {
        float4 q = world[i];
        out_sqd[o * 5 + r] = id >= 0 ? (ivox ? sqdist3_eigen(p.x, p.y, p.z, q.x, q.y, q.z) : sqdist3(q.x, q.y, q.z, p.x, p.y, p.z)) : INFINITY;
    }
}
static inline unsigned nblk(size_t n, int t) { return (unsigned)((n + t - 1) / t); }

// ---- context ------------------------------------------------------------------------------------------------
extern "C" const char* flivo_version(void) { return "flivo_b200 0.1 (sm_100a)"; }
extern "C" const char* flivo_last_error(const flivo_ctx* ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }
extern "C" void* flivo_ctx_stream(flivo_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

extern "C" int flivo_sync(flivo_ctx* ctx) {
    NEED(ctx, FLIVO_E_ARG, "null ctx");
    CKC(cudaSetDevice(ctx->device));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

static int ctx_init(flivo_ctx* ctx, int device, const flivo_options* opt) {
    int ndev = 0;
    CKC(cudaGetDeviceCount(&ndev));
    NEED(ndev > 0 && device >= 0 && device < ndev, FLIVO_E_CUDA, "no such CUDA device (libflivo_b200 has no CPU fallback)");
    ctx->device = device;
    CKC(cudaSetDevice(device));
    cudaDeviceProp prop;
    CKC(cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    CKC(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    flivo_options o{};
    if (opt) o = *opt;
    if (!(o.downsample_size > 0.f)) o.downsample_size = 0.2f;
    if (o.cell_ratio == 0) o.cell_ratio = 2;
    ctx->opt = o;
    if (o.ivox_nearby) { o.cell_ratio = 1; ctx->ivox_filter = (double)o.downsample_size; ctx->opt = o; }
    int rc = ctx->map.init(ctx->stream, o.downsample_size, o.cell_ratio, (size_t)o.initial_points, o.ivox_nearby, ctx->err);
    if (rc) return map_rc(ctx, rc);
    upload_term_table();
    CKC(cudaGetLastError());
    CKC(cudaMalloc(&ctx->d_blk, sizeof(DevBlock)));
    CKC(cudaMalloc(&ctx->d_sstamps, ((size_t)SSTAMP_CTAS * 4 + 32) * sizeof(unsigned long long)));
    CKC(cudaMalloc(&ctx->d_pre, sizeof(SolvePre)));
    CKC(cudaMalloc(&ctx->d_stamps, MAX_TRIPS * 16 * sizeof(unsigned long long)));
    CKC(cudaMemset(ctx->d_stamps, 0, MAX_TRIPS * 16 * sizeof(unsigned long long)));
    CKC(cudaMemsetAsync(ctx->d_blk, 0, sizeof(DevBlock), ctx->stream));
    CKC(cudaMallocHost(&ctx->h_blk, sizeof(DevBlock)));
    memset(ctx->h_blk, 0, sizeof(DevBlock));
    ctx->d_ctl = &ctx->d_blk->ctl; ctx->d_stats = &ctx->d_blk->stats;
    ctx->d_x = ctx->d_blk->x; ctx->d_xprop = ctx->d_blk->x_prop; ctx->d_Pprop = ctx->d_blk->P_prop; ctx->d_Pout = ctx->d_blk->P_out;
    ctx->d_sum = ctx->d_blk->sum;
    CKC(cudaMallocHost(&ctx->h_small, 2048 * sizeof(double)));
    CKC(cudaMalloc(&ctx->d_sjob, sizeof(ScanJob)));
    CKC(cudaMallocHost(&ctx->h_sjob, sizeof(ScanJob)));
    CKC(cudaEventCreateWithFlags(&ctx->sjob_ev, cudaEventDisableTiming));
    for (auto& e : ctx->ev) CKC(cudaEventCreate(&e));
    CKC(cudaEventCreateWithFlags(&ctx->stage_ev, cudaEventDisableTiming));
    CKC(cudaStreamCreateWithFlags(&ctx->copy, cudaStreamNonBlocking));
    for (auto& sl : ctx->slot) {
        CKC(cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming));
        CKC(cudaEventCreateWithFlags(&sl.consumed, cudaEventDisableTiming));
    }
    CKC(cudaEventCreate(&ctx->ev_stage0)); CKC(cudaEventCreate(&ctx->ev_upd0)); CKC(cudaEventCreate(&ctx->ev_upd1));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

extern "C" int flivo_ctx_create(int device, const flivo_options* opt, flivo_ctx** out) {
    if (!out) return FLIVO_E_ARG;
    *out = nullptr;
    flivo_ctx* ctx = new (std::nothrow) flivo_ctx();
    if (!ctx) return FLIVO_E_CAPACITY;
    int rc = ctx_init(ctx, device, opt);
    if (rc != FLIVO_OK) {
        // keep the message reachable for the caller: a failed create returns no ctx, so print it
        fprintf(stderr, "flivo_ctx_create: %s\n", ctx->err.c_str());
        flivo_ctx_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return FLIVO_OK;
}

static void free_scan(flivo_ctx* ctx) {
    if (ctx->sort_exec) { cudaGraphExecDestroy(ctx->sort_exec); ctx->sort_exec = nullptr; ctx->sort_npad = 0; }   // holds the old pointers
    cudaFree(ctx->d_body_own);
    cudaFree(ctx->d_body_in); cudaFree(ctx->d_perm); cudaFree(ctx->d_skey_a); cudaFree(ctx->d_skey_b); cudaFree(ctx->d_sidx_a);
    cudaFree(ctx->d_cub); cudaFree(ctx->d_near); cudaFree(ctx->d_near_meta); cudaFree(ctx->d_sel); cudaFree(ctx->d_eff);
    cudaFree(ctx->d_normvec); cudaFree(ctx->d_world); cudaFree(ctx->d_partials); cudaFree(ctx->d_gpartials); cudaFree(ctx->d_group_cnt); cudaFree(ctx->d_far_list); cudaFree(ctx->d_plane); cudaFree(ctx->d_pflags);
    ctx->d_body = ctx->d_body_own = ctx->d_body_in = nullptr; ctx->d_perm = nullptr; ctx->d_skey_a = ctx->d_skey_b = nullptr; ctx->d_sidx_a = nullptr;
    ctx->d_cub = nullptr; ctx->d_near = nullptr; ctx->d_near_meta = nullptr; ctx->d_sel = ctx->d_eff = nullptr;
    ctx->d_normvec = ctx->d_world = nullptr; ctx->d_partials = ctx->d_gpartials = nullptr; ctx->d_group_cnt = nullptr; ctx->d_far_list = nullptr; ctx->d_plane = nullptr; ctx->d_pflags = nullptr;
    ctx->scan_cap = 0; ctx->scan_n = 0; ctx->cub_bytes = 0;
}

static void free_raw(flivo_ctx* ctx);
extern "C" void flivo_ctx_destroy(flivo_ctx* ctx) {
    if (ctx && ctx->frame_calls > 0)
        fprintf(stderr, "[flivo_frame trace] %ld frames, host us per frame: commit %.1f  update_begin %.1f  upload_next %.1f  update_end(wait) %.1f  map_add_scan %.1f\n",
                ctx->frame_calls, 1e6 * ctx->frame_s[0] / ctx->frame_calls, 1e6 * ctx->frame_s[1] / ctx->frame_calls, 1e6 * ctx->frame_s[2] / ctx->frame_calls,
                1e6 * ctx->frame_s[3] / ctx->frame_calls, 1e6 * ctx->frame_s[4] / ctx->frame_calls);
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(ctx->comm);
    ctx->map.release();
    free_scan(ctx);
    free_raw(ctx);
    cudaFree(ctx->d_vg_state); cudaFree(ctx->d_poses);
    if (ctx->peer_on && !ctx->peer_local) for (int r = 0; r < ctx->nranks; ++r) if (r != ctx->rank && ctx->peer_base[r]) cudaIpcCloseMemHandle(ctx->peer_base[r]);
    cudaFree(ctx->d_xchg); cudaFree(ctx->d_xseq);
    cudaFree(ctx->d_ivox_flags); cudaFree(ctx->d_ivox_sel);
    if (ctx->graph_exec) cudaGraphExecDestroy(ctx->graph_exec);
    cudaFree(ctx->d_stage); cudaFree(ctx->d_out); cudaFree(ctx->d_blk); cudaFree(ctx->d_stamps); cudaFree(ctx->d_sstamps); cudaFree(ctx->d_pre);
    cudaFree(ctx->d_rank_blocks);
    if (ctx->h_blk) cudaFreeHost(ctx->h_blk);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->h_small) cudaFreeHost(ctx->h_small);
    cudaFree(ctx->d_sjob);
    if (ctx->h_sjob) cudaFreeHost(ctx->h_sjob);
    if (ctx->sjob_ev) cudaEventDestroy(ctx->sjob_ev);
    if (ctx->block_ev) cudaEventDestroy(ctx->block_ev);
    for (auto& e : ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->stage_ev) cudaEventDestroy(ctx->stage_ev);
    if (ctx->copy) { cudaStreamSynchronize(ctx->copy); cudaStreamDestroy(ctx->copy); }
    for (auto& sl : ctx->slot) {
        cudaFree(sl.d_in);
        if (sl.h_pin) cudaFreeHost(sl.h_pin);
        if (sl.copied) cudaEventDestroy(sl.copied);
        if (sl.consumed) cudaEventDestroy(sl.consumed);
    }
    for (cudaEvent_t e : {ctx->ev_stage0, ctx->ev_upd0, ctx->ev_upd1}) if (e) cudaEventDestroy(e);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// ---- map: ikd-Tree API surface ------------------------------------------------------------------------------
#define ENTER()                                                      \
    NEED(ctx, FLIVO_E_ARG, "null ctx");                              \
    CKC(cudaSetDevice(ctx->device));                                 \
    ctx->err.clear()

extern "C" int flivo_map_set_downsample(flivo_ctx* ctx, float downsample_size) {
    ENTER();
    if (ctx->opt.ivox_nearby) {
        NEED(downsample_size > 0.f, FLIVO_E_ARG, "filter_size_map_min must be > 0");
        ctx->ivox_filter = (double)downsample_size;
        return FLIVO_OK;
    }
    return map_rc(ctx, ctx->map.set_downsample(downsample_size, ctx->err));
}

extern "C" int flivo_map_build(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes) {
    ENTER();
    NEED(n == 0 || (xyz && stride_ok(stride_bytes)), FLIVO_E_ARG, "bad xyz / stride");
    NEED(n < 0x7FFFFFF0ull, FLIVO_E_ARG, "too many points");
    int rc = ensure_dstage(ctx, n);
    if (rc) return rc;
    if ((rc = stage_points(ctx, xyz, n, stride_bytes, ctx->d_stage))) return rc;
    return map_rc(ctx, ctx->map.build(ctx->d_stage, n, ctx->err));
}

extern "C" int flivo_map_add_points(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes, int downsample_on, int32_t* n_added) {
    ENTER();
    if (n_added) *n_added = 0;
    NEED(n == 0 || (xyz && stride_ok(stride_bytes)), FLIVO_E_ARG, "bad xyz / stride");
    int rc = ensure_dstage(ctx, n);
    if (rc) return rc;
    if ((rc = stage_points(ctx, xyz, n, stride_bytes, ctx->d_stage))) return rc;
    int added = 0;
    rc = ctx->map.add_points(ctx->d_stage, n, downsample_on != 0, &added, ctx->err);
    if (n_added) *n_added = added;
    return map_rc(ctx, rc);
}

extern "C" int flivo_map_delete_boxes(flivo_ctx* ctx, const float* boxes6, size_t nboxes, int32_t* n_deleted) {
    ENTER();
    if (n_deleted) *n_deleted = 0;
    NEED(nboxes == 0 || boxes6, FLIVO_E_ARG, "null boxes");
    int del = 0;
    int rc = ctx->map.delete_boxes(boxes6, nboxes, &del, ctx->err);
    if (n_deleted) *n_deleted = del;
    return map_rc(ctx, rc);
}

extern "C" int flivo_map_add_boxes(flivo_ctx* ctx, const float* boxes6, size_t nboxes, int32_t* n_revived) {
    ENTER();
    if (n_revived) *n_revived = 0;
    NEED(nboxes == 0 || boxes6, FLIVO_E_ARG, "null boxes");
    int rev = 0;
    int rc = ctx->map.add_boxes(boxes6, nboxes, &rev, ctx->err);
    if (n_revived) *n_revived = rev;
    return map_rc(ctx, rc);
}

extern "C" int flivo_map_delete_points(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes, int32_t* n_deleted) {
    ENTER();
    if (n_deleted) *n_deleted = 0;
    NEED(n == 0 || (xyz && stride_ok(stride_bytes)), FLIVO_E_ARG, "bad xyz / stride");
    int rc = ensure_dstage(ctx, n);
    if (rc) return rc;
    if ((rc = stage_points(ctx, xyz, n, stride_bytes, ctx->d_stage))) return rc;
    int del = 0;
    rc = ctx->map.delete_points(ctx->d_stage, n, &del, ctx->err);
    if (n_deleted) *n_deleted = del;
    return map_rc(ctx, rc);
}

extern "C" int flivo_map_size(flivo_ctx* ctx, int64_t* n_live) {
    ENTER();
    NEED(n_live, FLIVO_E_ARG, "null out");
    int rc = ctx->map.refresh_counters(ctx->err);
    *n_live = ctx->map.live();
    return map_rc(ctx, rc);
}

extern "C" int flivo_map_range(flivo_ctx* ctx, float out6[6]) {
    ENTER();
    NEED(out6, FLIVO_E_ARG, "null out");
    return map_rc(ctx, ctx->map.range(out6, ctx->err));
}

extern "C" int flivo_map_flatten(flivo_ctx* ctx, float* out_xyz, int32_t* out_id, size_t cap, size_t* n) {
    ENTER();
    return map_rc(ctx, ctx->map.flatten(out_xyz, out_id, cap, n, ctx->err));
}

extern "C" int flivo_map_acquire_removed(flivo_ctx* ctx, float* out_xyz, size_t cap, size_t* n) {
    ENTER();
    return map_rc(ctx, ctx->map.acquire_removed(out_xyz, cap, n, ctx->err));
}

// largest float <= max_dist as bits (the reference compares the float distance against a double bound)
static bool bound_bits(double max_dist, uint32_t* bits) {
    if (!(max_dist >= 0.0)) return false;  // negative or NaN: nothing can match
    float f = (float)max_dist;
    if ((double)f > max_dist) f = nextafterf(f, -INFINITY);
    memcpy(bits, &f, 4);
    return true;
}

static int fetch_stats(flivo_ctx* ctx) {
    CKC(cudaMemcpyAsync(&ctx->h_blk->stats, ctx->d_stats, sizeof(SearchStats), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    ctx->stats.last_far_queries = (int64_t)ctx->h_blk->stats.far_queries;
    ctx->stats.last_candidates = (int64_t)ctx->h_blk->stats.candidates;
    return FLIVO_OK;
}

extern "C" int flivo_knn(flivo_ctx* ctx, const float* q_xyz, size_t nq, size_t stride_bytes, int k, double max_dist,
                         float* out_xyz, float* out_sqd, int32_t* out_id, int32_t* out_count) {
    ENTER();
    NEED(k >= 1 && k <= FLIVO_K, FLIVO_E_ARG, "k must be in 1..5");
    NEED(nq == 0 || (q_xyz && stride_ok(stride_bytes)), FLIVO_E_ARG, "bad queries / stride");
    NEED(nq < 0x7FFFFFF0ull / 8, FLIVO_E_ARG, "too many queries");
    if (nq == 0) return FLIVO_OK;
    uint32_t bits = 0;
    if (!bound_bits(max_dist, &bits)) {
        for (size_t i = 0; i < nq * (size_t)k; ++i) {
            if (out_id) out_id[i] = -1;
            if (out_sqd) out_sqd[i] = INFINITY;
            if (out_xyz) out_xyz[3 * i] = out_xyz[3 * i + 1] = out_xyz[3 * i + 2] = 0.f;
        }
        if (out_count) memset(out_count, 0, nq * sizeof(int32_t));
        return FLIVO_OK;
    }
    int rc = ensure_dstage(ctx, nq);
    if (rc) return rc;
    if ((rc = stage_points(ctx, q_xyz, nq, stride_bytes, ctx->d_stage))) return rc;
    size_t per = (size_t)k * (3 + 1 + 1) * 4 + 4;
    size_t bytes = nq * per + 1024;
    if (bytes > ctx->d_out_bytes) {
        cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->d_out_bytes = 0;
        CKC(cudaMalloc(&ctx->d_out, bytes + bytes / 4));
        ctx->d_out_bytes = bytes + bytes / 4;
    }
    float* d_xyz = (float*)ctx->d_out;
    float* d_sqd = d_xyz + nq * k * 3;
    int* d_id = (int*)(d_sqd + nq * k);
    int* d_cnt = d_id + nq * k;
    CKC(cudaMemsetAsync(ctx->d_stats, 0, sizeof(SearchStats), ctx->stream));
    int grid = (int)std::min<size_t>((nq + 3) / 4, (size_t)ctx->sm_count * 16);
    if (ctx->opt.ivox_nearby) launch_knn_ivox(ctx->map.view(), ctx->d_stage, (int)nq, k, max_dist, d_xyz, d_sqd, d_id, d_cnt, ctx->d_stats, ctx->stream);
    else launch_knn(ctx->map.view(), ctx->d_stage, (int)nq, k, bits, d_xyz, d_sqd, d_id, d_cnt, ctx->d_stats, grid, ctx->stream);
    CKC(cudaGetLastError());
    if (out_xyz) CKC(cudaMemcpyAsync(out_xyz, d_xyz, nq * k * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (out_sqd) CKC(cudaMemcpyAsync(out_sqd, d_sqd, nq * k * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (out_id) CKC(cudaMemcpyAsync(out_id, d_id, nq * k * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (out_count) CKC(cudaMemcpyAsync(out_count, d_cnt, nq * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    return fetch_stats(ctx);
}

// ---- batched Nearest_Search without any allocation -------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels as flivo_knn (k_knn /
// k_knn_ivox), but the staged queries and the result rows live in ONE file-scope device array instead of the context's
// growable cudaMalloc buffers, and the host copies go straight from / to the caller's memory (no pinned staging buffer).
//   * limit: FLIVO_KNN_NOALLOC_MAX_QUERIES queries per call (beyond it: FLIVO_E_CAPACITY, nothing launched);
//   * the array (about 3.8 MB) is held by every GPU that loads the library until the process exits;
//   * every call on a device shares it and it keeps the last call's contents.  Nothing in it has to start at zero: the
//     kernel writes every result row of the call (xyz, distance, id, count for all nq queries) before it is read back.
//     Calls that overlap on one GPU (two contexts, two threads) race on it: serialise them.
// In host code the array's name is only a stand-in: the device address comes from cudaGetSymbolAddress, looked up once
// per device with that device current.
constexpr size_t KNN_NA_MAXQ = FLIVO_KNN_NOALLOC_MAX_QUERIES;
constexpr size_t KNN_NA_ROW = (size_t)FLIVO_K * (3 + 1 + 1) * 4 + 4;            // result bytes per query at k = FLIVO_K
constexpr size_t KNN_NA_BYTES = KNN_NA_MAXQ * (sizeof(float4) + KNN_NA_ROW) + 1024;
static __device__ __align__(256) unsigned char g_knn_noalloc_scratch[KNN_NA_BYTES];
constexpr int KNN_NA_MAXDEV = 64;
static void* g_knn_noalloc_ptr[KNN_NA_MAXDEV] = {};

extern "C" int flivo_knn_noalloc(flivo_ctx* ctx, const float* q_xyz, size_t nq, size_t stride_bytes, int k, double max_dist,
                                 float* out_xyz, float* out_sqd, int32_t* out_id, int32_t* out_count) {
    ENTER();   // (makes the context's device current)
    NEED(k >= 1 && k <= FLIVO_K, FLIVO_E_ARG, "k must be in 1..5");
    NEED(nq == 0 || (q_xyz && stride_ok(stride_bytes)), FLIVO_E_ARG, "bad queries / stride");
    NEED(nq <= KNN_NA_MAXQ, FLIVO_E_CAPACITY, "flivo_knn_noalloc: more than FLIVO_KNN_NOALLOC_MAX_QUERIES (32768) queries; split the batch or use flivo_knn");
    NEED(ctx->device >= 0 && ctx->device < KNN_NA_MAXDEV, FLIVO_E_ARG, "device index out of range");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is in flight (flivo_iekf_update_end first)");
    if (nq == 0) return FLIVO_OK;
    uint32_t bits = 0;
    if (!bound_bits(max_dist, &bits)) {   // negative or NaN bound: nothing can match
        for (size_t i = 0; i < nq * (size_t)k; ++i) {
            if (out_id) out_id[i] = -1;
            if (out_sqd) out_sqd[i] = INFINITY;
            if (out_xyz) out_xyz[3 * i] = out_xyz[3 * i + 1] = out_xyz[3 * i + 2] = 0.f;
        }
        if (out_count) memset(out_count, 0, nq * sizeof(int32_t));
        return FLIVO_OK;
    }
    if (!g_knn_noalloc_ptr[ctx->device]) {   // set-up, once per device, this device current
        void* p = nullptr;
        CKC(cudaGetSymbolAddress(&p, g_knn_noalloc_scratch));
        NEED(p != nullptr, FLIVO_E_CUDA, "cudaGetSymbolAddress returned no address");
        g_knn_noalloc_ptr[ctx->device] = p;
    }
    unsigned char* base = static_cast<unsigned char*>(g_knn_noalloc_ptr[ctx->device]);
    float4* d_q = reinterpret_cast<float4*>(base);
    float* d_xyz = reinterpret_cast<float*>(base + KNN_NA_MAXQ * sizeof(float4));
    float* d_sqd = d_xyz + nq * k * 3;
    int* d_id = reinterpret_cast<int*>(d_sqd + nq * k);
    int* d_cnt = d_id + nq * k;
    cudaStream_t s = ctx->stream;
    // queries: x y z of every record -> the first 12 bytes of a float4 slot (w is never read by the kernels)
    if (stride_bytes == 16) CKC(cudaMemcpyAsync(d_q, q_xyz, nq * 16, cudaMemcpyHostToDevice, s));
    else CKC(cudaMemcpy2DAsync(d_q, 16, q_xyz, stride_bytes, 12, nq, cudaMemcpyHostToDevice, s));
    const int grid = (int)std::min<size_t>((nq + 3) / 4, (size_t)ctx->sm_count * 16);
    if (ctx->opt.ivox_nearby) launch_knn_ivox(ctx->map.view(), d_q, (int)nq, k, max_dist, d_xyz, d_sqd, d_id, d_cnt, nullptr, s);
    else launch_knn(ctx->map.view(), d_q, (int)nq, k, bits, d_xyz, d_sqd, d_id, d_cnt, nullptr, grid, s);
    CKC(cudaGetLastError());
    if (out_xyz) CKC(cudaMemcpyAsync(out_xyz, d_xyz, nq * k * 3 * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (out_sqd) CKC(cudaMemcpyAsync(out_sqd, d_sqd, nq * k * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (out_id) CKC(cudaMemcpyAsync(out_id, d_id, nq * k * sizeof(int), cudaMemcpyDeviceToHost, s));
    if (out_count) CKC(cudaMemcpyAsync(out_count, d_cnt, nq * sizeof(int), cudaMemcpyDeviceToHost, s));
    CKC(cudaStreamSynchronize(s));
    return FLIVO_OK;
}

// ---- scan ---------------------------------------------------------------------------------------------------
static int ensure_scan(flivo_ctx* ctx, size_t n) {
    if (n > ctx->scan_cap) {
        free_scan(ctx);
        size_t cap = std::max<size_t>(n + n / 8, 1u << 15);
        cap = ((cap + SCAN_PAD - 1) / SCAN_PAD) * SCAN_PAD;   // the staging sort runs on n rounded up to SCAN_PAD
        size_t units = (cap + UNIT_POINTS - 1) / UNIT_POINTS, groups = (units + GROUP_UNITS - 1) / GROUP_UNITS;
        CKC(cudaMalloc(&ctx->d_body_in, cap * sizeof(float4)));
        CKC(cudaMalloc(&ctx->d_body_own, cap * sizeof(float4)));
        CKC(cudaMalloc(&ctx->d_perm, cap * sizeof(unsigned int)));
        CKC(cudaMalloc(&ctx->d_skey_a, cap * sizeof(unsigned long long)));
        CKC(cudaMalloc(&ctx->d_skey_b, cap * sizeof(unsigned long long)));
        CKC(cudaMalloc(&ctx->d_sidx_a, cap * sizeof(unsigned int)));
        size_t b = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, b, ctx->d_skey_a, ctx->d_skey_b, ctx->d_sidx_a, ctx->d_perm, (int)cap, 0, 63, ctx->stream);
        ctx->cub_bytes = b + 256;
        CKC(cudaMalloc(&ctx->d_cub, ctx->cub_bytes));
        CKC(cudaMalloc(&ctx->d_near, cap * 5 * sizeof(float4)));
        CKC(cudaMalloc(&ctx->d_near_meta, cap * sizeof(float2)));
        CKC(cudaMalloc(&ctx->d_sel, cap));
        CKC(cudaMalloc(&ctx->d_eff, cap));
        CKC(cudaMalloc(&ctx->d_normvec, cap * sizeof(float4)));
        CKC(cudaMalloc(&ctx->d_world, cap * sizeof(float4)));
        const size_t max_grid = (size_t)ctx->sm_count * 8 + 256;   // >= any fused_grid()
        groups = max_grid / FOLD_CTAS + 1;
        CKC(cudaMalloc(&ctx->d_partials, max_grid * PSTRIDE * sizeof(double)));   // one record per CTA of k_fused
        CKC(cudaMalloc(&ctx->d_gpartials, groups * PSTRIDE * sizeof(double)));   // one record per FOLD_CTAS CTAs
        CKC(cudaMalloc(&ctx->d_group_cnt, (groups + 1) * sizeof(unsigned int)));   // + the ticket of the group folders
        CKC(cudaMalloc(&ctx->d_far_list, cap * sizeof(unsigned long long)));
        CKC(cudaMemsetAsync(ctx->d_far_list, 0, cap * sizeof(unsigned long long), ctx->stream));   // tag 0 = never valid
        CKC(cudaMalloc(&ctx->d_plane, cap * sizeof(float4)));
        CKC(cudaMalloc(&ctx->d_pflags, cap));
        CKC(cudaMemsetAsync(ctx->d_group_cnt, 0, (groups + 1) * sizeof(unsigned int), ctx->stream));
        ctx->group_slots = groups;
        ctx->scan_cap = cap;
    }
    return FLIVO_OK;
}

// A new current scan: n records of `stride` bytes at `src` (device memory; slot >= 0: the landing buffer they sit in).
// Nothing is launched here: the staging sequence runs right before the scan's first consumer (ensure_staged), so that the
// staging graph and the update graph are launched back to back from inside one call.
static int finish_scan(flivo_ctx* ctx, const char* src, size_t stride, size_t n, bool external, int slot) {
    ctx->scan_external = external;
    ctx->scan_sorted = ctx->opt.sort_scan != 0;
    ctx->scan_n = n;
    ctx->stage_todo = true;
    ctx->todo_src = src; ctx->todo_stride = stride; ctx->todo_slot = slot;
    ctx->todo_n_a = n; ctx->todo_src_b = src; ctx->todo_stride_b = stride;   // one piece (flivo_scan_commit overrides for a split upload)
    ctx->d_body = ctx->d_body_own;
    if (!ctx->scan_sorted && external && stride == 16) ctx->d_body = (float4*)src;   // used in place, no copy
    return FLIVO_OK;
}

// The staging sequence of the current scan, if it has not run yet:
//   sorted scans : the staging sort of stage.cu (records -> float4 in the caller's order + 16-bit keys, two stable 8-bit
//                  passes, gather into processing order + flag reset: three own kernels) as ONE captured graph (H2D of the ScanJob block included): a single
//                  launch, no launch gaps between the short kernels.  The graph depends only on the padded item count and
//                  the device pointers of this context; source pointer, stride and n travel in ScanJob.
//   unsorted     : k_scan_prepare alone (nothing at all for a float4 device pointer, which is used in place).
// *staged tells the caller whether anything was enqueued (its time stamp ev_stage0 is then valid).
static int ensure_staged(flivo_ctx* ctx, bool* staged = nullptr) {
    if (staged) *staged = false;
    if (!ctx->stage_todo) return FLIVO_OK;
    ctx->stage_todo = false;
    const size_t n = ctx->scan_n;
    const bool sort = ctx->scan_sorted;
    cudaStream_t s = ctx->stream;
    if (ctx->todo_slot >= 0) CKC(cudaStreamWaitEvent(s, ctx->slot[ctx->todo_slot].copied, 0));   // the records have landed
    const bool in_place = !sort && ctx->d_body != ctx->d_body_own;
    if (!in_place) {
        // The sort key: 16 of the Morton bits — cell bits 2..6 of x and y, 2..7 of z.  The two lowest bits per axis carry no
        // locality a 128-point tile can use; the bits above repeat every 51 m (102 m in z), so a tile may mix two or three
        // far-apart regions, which costs less than a third radix pass (0.5257 vs 0.5279 ms per 260 k-point update, 0.325 vs
        // 0.334 at 24 k points; dropping fine bits instead — 9..25 — loses 8 us of search time).  FLIVO_SORT_BITS=lo,hi
        // overrides.  Two implementations of the same stable sort (identical permutation): the own three-kernel sort of
        // stage.cu for scans up to STAGE_OWN_MAX points — no library kernel on the step, and faster (step time own / library:
        // 24 k points 0.244 / 0.262 ms, 100 k 0.337 / 0.351, 260 k 0.470 / 0.479) — and the library's onesweep radix sort
        // above that (4 M points: 3.22 / 3.10 ms) — profiles/r2_ab_stage_sort.txt.  FLIVO_SORT=own|cub forces one of them;
        // a key width other than 16 bits always takes the library sort.
        static int sb0 = -1, sb1 = -1, sort_mode = 0;   // 0: by size, 1: own, 2: library
        if (sb0 < 0) {
            const char* e = getenv("FLIVO_SORT_BITS");
            if (!e || sscanf(e, "%d,%d", &sb0, &sb1) != 2) { sb0 = 6; sb1 = 22; }
            const char* m = getenv("FLIVO_SORT");
            sort_mode = (sb1 - sb0 != 16) ? 2 : (m && !strcmp(m, "cub")) ? 2 : (m && !strcmp(m, "own")) ? 1 : 0;
        }
        const unsigned n_pad = sort ? (unsigned)(((n + SCAN_PAD - 1) / SCAN_PAD) * SCAN_PAD) : (unsigned)n;
        constexpr unsigned STAGE_OWN_MAX = 1u << 20;
        const bool own_sort = sort_mode == 1 || (sort_mode == 0 && n_pad <= STAGE_OWN_MAX);
        if (ctx->sjob_pending) { CKC(cudaEventSynchronize(ctx->sjob_ev)); ctx->sjob_pending = false; }
        ScanJob* J = ctx->h_sjob;
        J->src = ctx->todo_src; J->stride = (unsigned)ctx->todo_stride; J->n = (unsigned)n; J->n_pad = n_pad;
        J->n_a = (unsigned)ctx->todo_n_a; J->src_b = ctx->todo_src_b; J->stride_b = (unsigned)ctx->todo_stride_b;
        J->inv_edge = 1.0f / ctx->map.view().cell_edge;
        J->key_shift = (unsigned)sb0;
        if (sort) {
            if (!ctx->sort_exec || ctx->sort_npad != n_pad) {
                if (ctx->sort_exec) { cudaGraphExecDestroy(ctx->sort_exec); ctx->sort_exec = nullptr; }
                cudaGraph_t g = nullptr;
                CKC(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
                cudaError_t ce = cudaEventRecordWithFlags(ctx->ev_stage0, s, cudaEventRecordExternal);
                if (ce == cudaSuccess) ce = cudaMemcpyAsync(ctx->d_sjob, ctx->h_sjob, sizeof(ScanJob), cudaMemcpyHostToDevice, s);
                if (own_sort) {
                    // (the work buffers alias the library sort's key / index arrays: 8 bytes per point cover the key16 array,
                    // and the two histogram matrices take n_pad bytes of the alternate key array)
                    StageWork W;
                    W.key16 = reinterpret_cast<unsigned short*>(ctx->d_skey_a);
                    W.order1 = ctx->d_sidx_a;
                    W.hist = reinterpret_cast<unsigned int*>(ctx->d_skey_b);
                    if (stage_hist_bytes(n_pad) > (size_t)n_pad * sizeof(unsigned long long)) ce = cudaErrorInvalidValue;
                    if (ce == cudaSuccess) launch_stage_sort(ctx->d_sjob, n_pad, ctx->d_body_in, ctx->d_body_own, ctx->d_perm, ctx->d_sel, ctx->d_eff, ctx->d_pflags, W, s);
                } else {
                    k_scan_prepare<<<nblk(n_pad, 256), 256, 0, s>>>(ctx->d_sjob, ctx->d_body_in, ctx->d_skey_a, ctx->d_sidx_a);
                    size_t tmp = ctx->cub_bytes;
                    if (ce == cudaSuccess) ce = cub::DeviceRadixSort::SortPairs(ctx->d_cub, tmp, ctx->d_skey_a, ctx->d_skey_b, ctx->d_sidx_a, ctx->d_perm, (int)n_pad, sb0, sb1, s);
                    k_gather4<<<nblk(n_pad, 256), 256, 0, s>>>(ctx->d_sjob, ctx->d_body_in, ctx->d_perm, ctx->d_body_own, ctx->d_sel, ctx->d_eff, ctx->d_pflags);
                }
                cudaError_t ce2 = cudaStreamEndCapture(s, &g);
                if (ce != cudaSuccess || ce2 != cudaSuccess) {
                    if (g) cudaGraphDestroy(g);
                    ctx->err = std::string("capture of the scan staging graph: ") + cudaGetErrorString(ce != cudaSuccess ? ce : ce2);
                    return FLIVO_E_CUDA;
                }
                ce = cudaGraphInstantiate(&ctx->sort_exec, g, 0);
                cudaGraphDestroy(g);
                if (ce != cudaSuccess) { ctx->sort_exec = nullptr; ctx->err = std::string("cudaGraphInstantiate (scan staging): ") + cudaGetErrorString(ce); return FLIVO_E_CUDA; }
                ctx->sort_npad = n_pad;
            }
            CKC(cudaGraphLaunch(ctx->sort_exec, s));
        } else {
            CKC(cudaEventRecord(ctx->ev_stage0, s));
            CKC(cudaMemcpyAsync(ctx->d_sjob, ctx->h_sjob, sizeof(ScanJob), cudaMemcpyHostToDevice, s));
            k_scan_prepare<<<nblk(n, 256), 256, 0, s>>>(ctx->d_sjob, ctx->d_body_own, nullptr, nullptr);
            CKC(cudaGetLastError());
        }
        CKC(cudaEventRecord(ctx->sjob_ev, s));
        ctx->sjob_pending = true;
        ctx->stage_launches = sort ? (own_sort ? 3 : 2) : 1;
        if (staged) *staged = true;
    }
    if (ctx->todo_slot >= 0) CKC(cudaEventRecord(ctx->slot[ctx->todo_slot].consumed, s));   // the slot may be refilled
    // point_selected_surf starts true (V3 global array semantics are re-established by the first search trip); no cached
    // neighbours yet: a re-use trip before any search selects nothing.  (Sorted scans: done inside k_gather4.)
    if (!sort) {
        CKC(cudaMemsetAsync(ctx->d_sel, 1, n, s));
        CKC(cudaMemsetAsync(ctx->d_eff, 0, n, s));
        CKC(cudaMemsetAsync(ctx->d_pflags, 0, n, s));
    }
    // iVox mode reads the cached count outside the search (empty-result rule of GetClosestPoint, map_incremental's add rule)
    if (ctx->opt.ivox_nearby) {
        CKC(cudaMemsetAsync(ctx->d_near_meta, 0, n * sizeof(float2), s));
        CKC(cudaMemsetAsync(ctx->d_near, 0xFF, n * 5 * sizeof(float4), s));   // id -1 everywhere: a point that never finds a candidate reports an empty list
    }
    return FLIVO_OK;
}

// ---- host -> device scan transfer, double buffered ------------------------------------------------------------------
static bool host_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// feats_down_body of the NEXT frame (src/laser_mapping.cpp:129-130,165-167 hands one over per frame): the records travel on
// the copy stream into the back slot while the update of the current scan runs.  Pinned / registered sources of packed
// xyz or float4 records (cudaHostRegister, flivo_host_register) are copied as they are — one DMA, no host work; the xyz
// gather happens on the device (k_scan_prepare).  Everything else (pageable memory, or wide records such as the 48-byte
// PointXYZINormal) is repacked to float4 into a pinned buffer by a few host threads first (chunked, each chunk's copy
// overlapping the repack of the next); with the update of the current scan in flight that work is hidden.
extern "C" int flivo_scan_upload_async(flivo_ctx* ctx, const float* body_xyz, size_t n, size_t stride_bytes) {
    ENTER();
    NEED(n > 0 && body_xyz && stride_ok(stride_bytes), FLIVO_E_ARG, "bad scan / stride");
    NEED(n < 0x7FFFFFF0ull / 8, FLIVO_E_ARG, "scan too large");
    ScanSlot& sl = ctx->slot[ctx->back];
    // whole records by DMA only when that does not inflate the PCIe traffic: a pinned float4 / packed-xyz buffer goes as it
    // is; the 48-byte PointXYZINormal records would triple the bytes (measured on B200 / PCIe 5: 12.5 MB per 260 k-point scan
    // took longer than the whole update and made the pipelined step copy-bound, 0.74 ms against 0.49), so they are repacked
    // to float4 by host threads — which now happens while the previous update runs.  The exception is a process with few host
    // threads to itself (8 ranks on one 32-core host: 4 each): there the repack, not the link, is the bottleneck — measured
    // at N = 8: 1.13 ms per pipelined step against 0.48 ms of update — and pinned wide records travel as they are
    // (FLIVO_RAW_DMA=1 / 0 forces / forbids it).
    static const int raw_mode = getenv("FLIVO_RAW_DMA") ? atoi(getenv("FLIVO_RAW_DMA")) : -1;
    // (Letting the copy engine gather the xyz itself — a 2-D copy of 12-byte rows at the record pitch — was measured too: 1.2 G
    // points/s pipelined at N = 1 against 2.07 G for the repack and 1.52 G for the raw records; profiles/r2_ab_upload.txt.)
    // Best with few threads is BOTH at once (FLIVO_RAW_SPLIT = per cent of the scan repacked, default 50): the copy engine
    // moves the tail of the scan raw while the host threads repack its head, each doing about half of what bounds it.
    // (by team size when not forced: 4 threads repack half, 8 threads four fifths, more than 8 everything)
    static const int split_pct = getenv("FLIVO_RAW_SPLIT") ? std::min(100, std::max(0, atoi(getenv("FLIVO_RAW_SPLIT")))) : (host_threads() <= 4 ? 50 : 80);
    const bool few = raw_mode < 0 && host_threads() <= 8 && n >= 65536;
    const bool wide_pinned = stride_bytes > 16 && (raw_mode == 1 || few || (raw_mode == 2)) && host_pinned(body_xyz);
    const bool pinned = (stride_bytes <= 16 && host_pinned(body_xyz)) || (wide_pinned && (raw_mode == 1 || split_pct == 0));
    const bool split = wide_pinned && !pinned;
    const size_t n_a = split ? std::min(n, ((n * (size_t)split_pct / 100 + 63) / 64) * 64) : n;   // points repacked by the host
    const size_t stride = pinned ? stride_bytes : 16;
    const size_t bytes = split ? n_a * 16 + (n - n_a) * stride_bytes : n * stride;
    if (sl.copy_pending) { CKC(cudaEventSynchronize(sl.copied)); sl.copy_pending = false; }
    if (bytes > sl.d_bytes) {
        CKC(cudaEventSynchronize(sl.consumed));                 // (no-op for a never recorded event)
        cudaFree(sl.d_in); sl.d_in = nullptr; sl.d_bytes = 0;
        const size_t cap = std::max<size_t>(bytes + bytes / 8, 1u << 20);
        CKC(cudaMalloc(&sl.d_in, cap));
        sl.d_bytes = cap;
    }
    CKC(cudaStreamWaitEvent(ctx->copy, sl.consumed, 0));        // the staging that read this slot's previous content has run
    if (split && n_a < n)   // the raw tail first: it travels while the host repacks the head
        CKC(cudaMemcpyAsync(sl.d_in + n_a * 16, reinterpret_cast<const char*>(body_xyz) + n_a * stride_bytes, (n - n_a) * stride_bytes,
                            cudaMemcpyHostToDevice, ctx->copy));
    if (pinned) {
        CKC(cudaMemcpyAsync(sl.d_in, body_xyz, bytes, cudaMemcpyHostToDevice, ctx->copy));
    } else {
        const size_t n = n_a;   // (the part the host repacks: all of the scan unless the upload is split)
        if (n > sl.h_n) {
            if (sl.h_pin) cudaFreeHost(sl.h_pin);
            sl.h_pin = nullptr; sl.h_n = 0;
            const size_t cap = std::max<size_t>(n + n / 4, 1u << 16);
            CKC(cudaMallocHost(&sl.h_pin, cap * sizeof(float4)));
            sl.h_n = cap;
        }
        const char* src = reinterpret_cast<const char*>(body_xyz);
        float4* h = sl.h_pin;
        const size_t CH = 1u << 15;
        const int nt = host_threads();
        for (size_t c0 = 0; c0 < n; c0 += CH) {
            const size_t m = std::min(CH, n - c0);
            if (stride_bytes == 16) {
                memcpy(h + c0, src + c0 * 16, m * 16);
            } else {
                // a thread team only for large scans: a 24 k-point repack is ~50 us on one thread, and replay drives several
                // contexts from several host threads at once (8 teams of 16 threads were measured to halve the throughput)
#pragma omp parallel for num_threads(nt) schedule(static) if (n >= 65536 && m >= 8192 && nt > 1)
                for (long i = 0; i < (long)m; ++i) {
                    const float* p = reinterpret_cast<const float*>(src + (c0 + (size_t)i) * stride_bytes);
                    h[c0 + i] = make_float4(p[0], p[1], p[2], 0.f);
                }
            }
            CKC(cudaMemcpyAsync(sl.d_in + c0 * 16, h + c0, m * sizeof(float4), cudaMemcpyHostToDevice, ctx->copy));
        }
    }
    CKC(cudaEventRecord(sl.copied, ctx->copy));
    sl.copy_pending = true; sl.filled = true; sl.n = n; sl.stride = stride;
    sl.n_a = split ? n_a : n; sl.stride_b = split ? stride_bytes : stride;
    return FLIVO_OK;
}

// The scan uploaded last becomes the current scan (the update after this call works on it).
extern "C" int flivo_scan_commit(flivo_ctx* ctx) {
    ENTER();
    ScanSlot& sl = ctx->slot[ctx->back];
    NEED(sl.filled, FLIVO_E_STATE, "no scan uploaded since the last commit");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is in flight (flivo_iekf_update_end first)");
    int rc = ensure_scan(ctx, sl.n);
    if (rc) return rc;
    sl.filled = false;
    const int slot = ctx->back;
    ctx->back ^= 1;
    rc = finish_scan(ctx, sl.d_in, sl.stride, sl.n, false, slot);
    if (sl.n_a < sl.n) { ctx->todo_n_a = sl.n_a; ctx->todo_src_b = sl.d_in + sl.n_a * 16; ctx->todo_stride_b = sl.stride_b; }
    return rc;
}

extern "C" int flivo_scan_upload(flivo_ctx* ctx, const float* body_xyz, size_t n, size_t stride_bytes) {
    int rc = flivo_scan_upload_async(ctx, body_xyz, n, stride_bytes);
    if (rc != FLIVO_OK) return rc;
    return flivo_scan_commit(ctx);
}

// cudaHostRegister / cudaHostUnregister for callers without the CUDA headers: a registered scan buffer is copied by DMA
// as it is (flivo_scan_upload_async), with no host-side repack.
extern "C" int flivo_host_register(void* ptr, size_t bytes) {
    if (!ptr || !bytes) return FLIVO_E_ARG;
    return cudaHostRegister(ptr, bytes, cudaHostRegisterPortable) == cudaSuccess ? FLIVO_OK : (cudaGetLastError(), FLIVO_E_CUDA);
}
extern "C" int flivo_host_unregister(void* ptr) {
    if (!ptr) return FLIVO_E_ARG;
    return cudaHostUnregister(ptr) == cudaSuccess ? FLIVO_OK : (cudaGetLastError(), FLIVO_E_CUDA);
}

extern "C" int flivo_scan_set_device(flivo_ctx* ctx, const void* dev_xyzw, size_t n) {
    ENTER();
    NEED(n > 0 && dev_xyzw, FLIVO_E_ARG, "bad scan");
    NEED(n < 0x7FFFFFF0ull / 8, FLIVO_E_ARG, "scan too large");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is in flight (flivo_iekf_update_end first)");
    int rc = ensure_scan(ctx, n);
    if (rc) return rc;
    return finish_scan(ctx, (const char*)dev_xyzw, 16, n, true, -1);
}

static FusedParams make_fused(flivo_ctx* ctx) {
    FusedParams P;
    memset(&P, 0, sizeof(P));   // padding too: the captured-graph cache compares parameter blocks bytewise
    P.map = ctx->map.view();
    P.body = ctx->d_body;
    P.job = &ctx->d_blk->job;
    P.state = ctx->d_x;
    P.near = ctx->d_near; P.near_meta = ctx->d_near_meta; P.sel = ctx->d_sel; P.eff = ctx->d_eff;
    P.plane = ctx->d_plane; P.pflags = ctx->d_pflags;
    P.normvec = ctx->d_normvec; P.world = ctx->d_world;
    P.partials = ctx->d_partials; P.gpartials = ctx->d_gpartials; P.group_cnt = ctx->d_group_cnt; P.far_list = ctx->d_far_list;
    P.total_cnt = ctx->d_group_cnt + ctx->group_slots;
    P.ctl = ctx->d_ctl; P.stats = ctx->d_stats;
    return P;
}
static int fused_grid(flivo_ctx* ctx);
// per-call scalars -> the pinned mirror of the device block (copied with the state by the caller)
static void fill_job(flivo_ctx* ctx, int mode, int max_iter, int force_search, double R, const double* limit23, uint32_t max_sq_bits, int aux_out) {
    Job& j = ctx->h_blk->job;
    j.n = (int)ctx->scan_n;
    j.units = (int)((ctx->scan_n + UNIT_POINTS - 1) / UNIT_POINTS);
    j.groups = (fused_grid(ctx) + FOLD_CTAS - 1) / FOLD_CTAS;   // group records of k_fused summed by k_solve
    j.mode = mode; j.max_iter = max_iter; j.force_search = force_search; j.nranks = ctx->nranks; j.max_sq_bits = max_sq_bits; j.aux_out = aux_out;
    j.useq = (++ctx->useq) & 0xFFFFFFu;
    j.R = R;
    for (int i = 0; i < 23; ++i) j.limit[i] = limit23 ? limit23[i] : 0.001;
    j.gnss_on = (ctx->gnss_on && mode == FLIVO_MODE_V1) ? 1 : 0; j.gnss_pad = 0;
    for (int i = 0; i < 3; ++i) { j.gnss_pos[i] = ctx->gnss[i]; j.gnss_lever[i] = ctx->gnss[3 + i]; j.gnss_std[i] = ctx->gnss[6 + i]; }
}
// launch grids are rounded up to capacity buckets so that scans of similar size share one captured graph;
// surplus CTAs see i >= job->n (k_search) or an exhausted unit counter (k_fused) and leave at once
static int search_grid(flivo_ctx* ctx) {
    const int tp = search_tile_points();
    const int tiles = (int)((ctx->scan_n + tp - 1) / tp);
    if (ctx->search_resident <= 0) ctx->search_resident = ctx->sm_count * search_ctas_per_sm(ctx->opt.ivox_nearby != 0);
    return std::max(1, std::min(((tiles + 255) / 256) * 256, ctx->search_resident));   // tiles are dealt dynamically
}
// scans whose tiles fill at most half of the resident CTAs leave the GPU under-filled: the search is then bound by each
// thread's chain of dependent loads and the probe-all-blocks-at-once flavour wins (search.cuh, PAR_PROBES).  A function
// of the search grid alone, so the captured graph's cache key covers it.
static bool par_probes(flivo_ctx* ctx, int sgrid) {
    static int force = -2;   // FLIVO_PAR_PROBES=0/1 overrides the choice (development A/B)
    if (force == -2) { const char* e = getenv("FLIVO_PAR_PROBES"); force = e ? atoi(e) : -1; }
    if (force >= 0) return force != 0;
    return sgrid * 2 <= ctx->search_resident;
}
static int fused_grid(flivo_ctx* ctx) {
    const int units = (int)((ctx->scan_n + UNIT_POINTS - 1) / UNIT_POINTS);
    return std::max(1, std::min(((units + 255) / 256) * 256, ctx->sm_count * 8));
}
// the fused path searches with the reference's own far-neighbour gate (pointSearchSqDis[4] > 5, src/laser_mapping.cpp:518)
// as the distance bound: a point whose 5th neighbour lies beyond it is masked either way
static uint32_t gate_bits() { float f = 5.0f; uint32_t b; memcpy(&b, &f, 4); return b; }
// ... in V3 only.  There point_selected_surf is a persistent array, so a point gated out on a search trip stays out until
// the next search (origin:963).  In V1 (the compiled path) the flags are a LOCAL vector<bool>(n, true) rebuilt on every call
// (src/laser_mapping.cpp:490) and the gate is applied on search trips only (:518): on re-use trips a point whose fifth
// neighbour lies beyond the gate still enters with its cached Nearest_Points if esti_plane accepts them.  V1 therefore
// keeps neighbours beyond the gate and records the gate in PF_GATE alone — up to a range of FLIVO_V1_RANGE_CELLS search
// cells (24: 9.6 m at filter_size_map 0.2, 24 m at 0.5).  Stated deviation: the reference's Nearest_Search is unbounded
// (max_dist = INFINITY, ikd_Tree.h:174); a point whose fifth neighbour lies beyond that range counts as having fewer than
// five here (masked), where the reference would fit a plane through points that far away.  The bound is what keeps a
// scan that looks into unmapped space fast: an unbounded search ends in a sweep of the whole block directory per query
// (measured on a map grown from one scan: 5.6 ms per 24 k-point update; profiles/r2_ab_search.txt).  iVox mode bounds the
// range itself.
static uint32_t search_bound_bits(const flivo_ctx* ctx, int mode) {
    if (mode != FLIVO_MODE_V1 || ctx->opt.ivox_nearby) return gate_bits();
    const float R = (float)FLIVO_V1_RANGE_CELLS * ctx->map.view().cell_edge;
    float b = R * R;
    if (!(b > 5.0f)) b = 5.0f;
    uint32_t bits;
    memcpy(&bits, &b, 4);
    return bits;
}

// copy a per-point device array to the host in the caller's order
template <class T>
static int fetch_rows(flivo_ctx* ctx, const T* d_src, int elems, T* h_dst) {
    size_t n = ctx->scan_n, bytes = n * elems * sizeof(T);
    const T* src = d_src;
    if (ctx->scan_sorted) {
        if (bytes > ctx->d_out_bytes) {
            cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->d_out_bytes = 0;
            CKC(cudaMalloc(&ctx->d_out, bytes + bytes / 4));
            ctx->d_out_bytes = bytes + bytes / 4;
        }
        k_scatter<T><<<nblk(n * elems, 256), 256, 0, ctx->stream>>>(d_src, ctx->d_perm, n, elems, (T*)ctx->d_out);
        CKC(cudaGetLastError());
        src = (const T*)ctx->d_out;
    }
    CKC(cudaMemcpyAsync(h_dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

// ---- measurement update -----------------------------------------------------------------------------------------
static void expand_sum(const double* sum, double* HTH, double* HTz) {
    int t = 0;
    for (int i = 0; i < 12; ++i)
        for (int j = i; j < 12; ++j) { HTH[i * 12 + j] = sum[t]; HTH[j * 12 + i] = sum[t]; ++t; }
    for (int i = 0; i < 12; ++i) HTz[i] = sum[78 + i];
}

extern "C" int flivo_h_share_model(flivo_ctx* ctx, const flivo_state* st, int mode, int do_search,
                                   double HTH[144], double HTz[12], int32_t* n_eff, double* res_sum,
                                   uint8_t* selected, uint8_t* eff, int32_t* nbr_id, float* nbr_sqd,
                                   float* plane, float* world) {
    ENTER();
    NEED(st && HTH && HTz, FLIVO_E_ARG, "null state / outputs");
    NEED(mode == FLIVO_MODE_V3 || mode == FLIVO_MODE_V1, FLIVO_E_ARG, "bad mode");
    NEED(ctx->scan_n > 0, FLIVO_E_STATE, "no scan uploaded");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is in flight (flivo_iekf_update_end first)");
    { int rcs = ensure_staged(ctx); if (rcs) return rcs; }
    fill_job(ctx, mode, 1, 0, 0.001, nullptr, search_bound_bits(ctx, mode), 1);
    memcpy(ctx->h_blk->x, st, sizeof(flivo_state));
    CKC(cudaMemcpyAsync(&ctx->d_blk->job, &ctx->h_blk->job, sizeof(Job), cudaMemcpyHostToDevice, ctx->stream));
    CKC(cudaMemcpyAsync(ctx->d_x, ctx->h_blk->x, sizeof(flivo_state), cudaMemcpyHostToDevice, ctx->stream));
    k_ctl_set<<<1, 32, 0, ctx->stream>>>(ctx->d_ctl, do_search ? 1 : 0);
    if (do_search) CKC(cudaMemsetAsync(ctx->d_stats, 0, sizeof(SearchStats), ctx->stream));
    FusedParams P = make_fused(ctx);
    if (do_search) { const int sg = search_grid(ctx); launch_search(P, sg, ctx->stream, par_probes(ctx, sg)); }
    SolveParams S;
    memset(&S, 0, sizeof(S));
    S.job = &ctx->d_blk->job; S.phase = SOLVE_PHASE_REDUCE_ONLY;
    S.sum_out = ctx->d_sum; S.ctl = ctx->d_ctl;
    launch_fused(P, S, fused_grid(ctx), ctx->stream);   // the CTA that folds the last record writes the sums (no solve)
    CKC(cudaGetLastError());
    double* hs = ctx->h_small + 64;
    CKC(cudaMemcpyAsync(hs, ctx->d_sum, NTERMS * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    expand_sum(hs, HTH, HTz);
    if (res_sum) *res_sum = hs[90];
    if (n_eff) *n_eff = (int32_t)hs[91];
    int rc;
    if (selected && (rc = fetch_rows<unsigned char>(ctx, ctx->d_sel, 1, selected))) return rc;
    if (eff && (rc = fetch_rows<unsigned char>(ctx, ctx->d_eff, 1, eff))) return rc;
    if (plane && (rc = fetch_rows<float>(ctx, (const float*)ctx->d_normvec, 4, plane))) return rc;
    if (world) {
        std::vector<float> tmp(ctx->scan_n * 4);
        if ((rc = fetch_rows<float>(ctx, (const float*)ctx->d_world, 4, tmp.data()))) return rc;
        for (size_t i = 0; i < ctx->scan_n; ++i) { world[3 * i] = tmp[4 * i]; world[3 * i + 1] = tmp[4 * i + 1]; world[3 * i + 2] = tmp[4 * i + 2]; }
    }
    if (nbr_id || nbr_sqd) {
        size_t n = ctx->scan_n, bytes = n * 5 * 8;
        if (bytes > ctx->d_out_bytes) {
            cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->d_out_bytes = 0;
            CKC(cudaMalloc(&ctx->d_out, bytes + bytes / 4));
            ctx->d_out_bytes = bytes + bytes / 4;
        }
        int* d_id = (int*)ctx->d_out;
        float* d_sq = (float*)(d_id + n * 5);
        k_near_unpack<<<nblk(n * 5, 256), 256, 0, ctx->stream>>>(ctx->d_near, ctx->d_world, ctx->scan_sorted ? ctx->d_perm : nullptr, n, d_id, d_sq, ctx->opt.ivox_nearby);
        CKC(cudaGetLastError());
        if (nbr_id) CKC(cudaMemcpyAsync(nbr_id, d_id, n * 5 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        if (nbr_sqd) CKC(cudaMemcpyAsync(nbr_sqd, d_sq, n * 5 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        CKC(cudaStreamSynchronize(ctx->stream));
    }
    if (do_search) return fetch_stats(ctx);
    return FLIVO_OK;
}

extern "C" int flivo_repack_v1(const double HTH[144], const double HTz[12], double laser_point_cov, double HPH[81], double HPL[9]) {
    if (!HTH || !HTz || !HPH || !HPL || !(laser_point_cov > 0.0)) return FLIVO_E_ARG;
    // Hsub = [A | n] = V3 columns [3:6 | 0:3]; HPH.block<6,6> += Hsub^T (1/R) Hsub (src/laser_mapping.cpp:584-592)
    static const int map6[6] = {3, 4, 5, 0, 1, 2};
    const double w = 1.0 / laser_point_cov;
    for (int i = 0; i < 81; ++i) HPH[i] = 0.0;
    for (int i = 0; i < 9; ++i) HPL[i] = 0.0;
    for (int a = 0; a < 6; ++a) {
        for (int b = 0; b < 6; ++b) HPH[a * 9 + b] = HTH[map6[a] * 12 + map6[b]] * w;
        HPL[a] = HTz[map6[a]] * w;
    }
    return FLIVO_OK;
}

// One update as a CUDA graph: the fixed sequence of max_iter+1 trips captured from the stream; kernels of skipped
// phases (re-use trips, trips after the stop) exit at once.  (A device-side WHILE/IF conditional graph was measured
// slower, profiles/r1_ab_launch_structure.txt.)
constexpr size_t PDL_FUSED_MAX_POINTS = 65536;
static int build_update_graph(flivo_ctx* ctx, const FusedParams& P, const SolveParams& S0, int sgrid, int fgrid, int trips) {
    cudaStream_t s = ctx->stream;
    cudaGraph_t g = nullptr;
    CKC(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    const SolveParams& S = S0;   // S0.phase: FULL, or PEER when the peer-memory exchange is attached
    // the call's inputs (job + P_prop + x, from the pinned mirror), the control-block reset and, at the end, the results
    // (x, P_out, ctl, stats -> pinned mirror) are nodes of the same graph: one launch per update.  A trip is two kernels:
    // the solve runs inside k_fused (the CTA that folds the last record).
    // The first and last nodes are event records (external events): the device time of an update is read between two
    // stamps taken by the GPU itself, so it does not move when the host thread that launched the graph is descheduled.
    cudaError_t cio = cudaEventRecordWithFlags(ctx->ev_upd0, s, cudaEventRecordExternal);
    if (cio == cudaSuccess) cio = cudaMemcpyAsync(ctx->d_blk, ctx->h_blk, offsetof(DevBlock, P_out), cudaMemcpyHostToDevice, s);
    launch_ctl_reset(S, ctx->d_blk->x_prop, ctx->d_stats, s);
    // k_search is a programmatic dependent of the kernel before it (k_ctl_reset, then the previous trip's k_fused): it starts
    // as soon as that kernel's solving CTA has published state + control block, i.e. beside the covariance update / the next
    // trip's state-only half, and its launch latency is off the critical path (FLIVO_PDL=0 turns it off)
    static int pdl_on = -1;
    if (pdl_on < 0) { const char* e = getenv("FLIVO_PDL"); pdl_on = e ? atoi(e) : 1; }          // A/B: 3 = k_search only, 2 = k_search from the second trip on only
    for (int t = 0; t < trips && t <= MAX_TRIPS; ++t) {
        FusedParams Pt = P;
        Pt.trip = t; Pt.pdl = (pdl_on == 1 || pdl_on == 3 || pdl_on == 4 || (pdl_on == 2 && t > 0)) ? 1 : 0;     // (trip 0 is the dependent of k_ctl_reset)
        // k_fused behind its own trip's k_search (early on re-use trips): pays for small scans, where the per-trip serial chain
        // dominates (24 k points: 0.268 -> 0.255 ms per step), and costs a little once the early CTAs crowd the solving CTA's
        // SM (100 k: +2.5 us, 260 k: +4 us; profiles/r2_ab_pdl.txt) — so by scan size.  FLIVO_PDL=4 forces it, 3 turns it off.
        Pt.pdl_fused = ((pdl_on == 1 && ctx->scan_n <= PDL_FUSED_MAX_POINTS) || pdl_on == 4) ? 1 : 0;
        launch_search(Pt, sgrid, s, par_probes(ctx, sgrid), Pt.pdl != 0);
        launch_fused(Pt, S, fgrid, s);
    }
    if (cio == cudaSuccess)
        cio = cudaMemcpyAsync(ctx->h_blk->x, ctx->d_blk->x, offsetof(DevBlock, x_prop) - offsetof(DevBlock, x), cudaMemcpyDeviceToHost, s);
    if (cio == cudaSuccess) cio = cudaEventRecordWithFlags(ctx->ev_upd1, s, cudaEventRecordExternal);
    cudaError_t ce = cudaStreamEndCapture(s, &g);
    if (ce == cudaSuccess && cio != cudaSuccess) { ce = cio; if (g) cudaGraphDestroy(g); g = nullptr; }
    if (ce != cudaSuccess) { ctx->err = std::string("cudaStreamEndCapture: ") + cudaGetErrorString(ce); return FLIVO_E_CUDA; }
    ce = cudaGraphInstantiate(&ctx->graph_exec, g, 0);
    cudaGraphDestroy(g);
    if (ce != cudaSuccess) { ctx->graph_exec = nullptr; ctx->err = std::string("cudaGraphInstantiate: ") + cudaGetErrorString(ce); return FLIVO_E_CUDA; }
    return FLIVO_OK;
}

// The update in two halves, so that the caller can do host work (stage the next scan: flivo_scan_upload_async) while
// the GPU runs this one: _begin launches everything and returns, _end waits and hands the results back.
extern "C" int flivo_iekf_update_begin(flivo_ctx* ctx, const flivo_state* inout, const double* Pm, const flivo_cfg* cfg) {
    ENTER();
    NEED(inout && Pm && cfg, FLIVO_E_ARG, "null state / P / cfg");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is already in flight (flivo_iekf_update_end first)");
    NEED(cfg->mode == FLIVO_MODE_V3 || cfg->mode == FLIVO_MODE_V1, FLIVO_E_ARG, "bad mode");
    NEED(cfg->max_iter >= 1 && cfg->max_iter + 1 <= MAX_TRIPS, FLIVO_E_ARG, "max_iter must be in 1..15");
    NEED(cfg->laser_point_cov > 0.0, FLIVO_E_ARG, "laser_point_cov must be > 0");
    NEED(ctx->scan_n > 0, FLIVO_E_STATE, "no scan uploaded");
    NEED(!(ctx->peer_on && ctx->peer_broken), FLIVO_E_STATE, "an earlier peer exchange timed out: call flivo_peer_reset on every rank (after a barrier) first");
    const int np = cfg->mode == FLIVO_MODE_V1 ? 18 : 23;
    const int trips = cfg->max_iter + 1;  // loop index -1 .. max_iter-1
    const bool timed = (cfg->flags & FLIVO_CFG_PROFILE) != 0;   // per-trip events, plain launches (no graph)
    cudaStream_t s = ctx->stream;
    DevBlock* h = ctx->h_blk;
    bool staged = false;
    { int rcs = ensure_staged(ctx, &staged); if (rcs) return rcs; }   // staging graph and update graph go out back to back
    fill_job(ctx, cfg->mode, cfg->max_iter, cfg->force_search, cfg->laser_point_cov, cfg->limit, search_bound_bits(ctx, cfg->mode), 0);
    memcpy(h->x, inout, sizeof(flivo_state));
    memset(h->P_prop, 0, sizeof(h->P_prop));
    memcpy(h->P_prop, Pm, sizeof(double) * np * np);
    bool use_graph = !timed && (!ctx->comm || ctx->peer_on);   // an NCCL call between the kernels is not captured
    FusedParams P = make_fused(ctx);
    if (timed && search_grid(ctx) <= SSTAMP_CTAS) {
        CKC(cudaMemsetAsync(ctx->d_sstamps, 0, ((size_t)SSTAMP_CTAS * 4 + 32) * sizeof(unsigned long long), s));
        P.search_stamps = ctx->d_sstamps;
    }
    SolveParams S;
    memset(&S, 0, sizeof(S));
    S.job = &ctx->d_blk->job;
    S.x = ctx->d_x; S.x_prop = ctx->d_xprop; S.P_prop = ctx->d_Pprop; S.P_out = ctx->d_Pout;
    S.sum_out = ctx->d_sum; S.rank_blocks = ctx->d_rank_blocks; S.ctl = ctx->d_ctl;
    S.stamps = timed ? ctx->d_stamps : nullptr;
    S.pre = ctx->d_pre;
    if (ctx->peer_on) {
        for (int r = 0; r < ctx->nranks; ++r) S.peer.base[r] = ctx->peer_base[r];
        S.peer.seq = ctx->d_xseq; S.peer.rank = ctx->rank; S.peer.nranks = ctx->nranks; S.peer.timeout_ns = ctx->peer_timeout_ns;
    }
    const int full_phase = ctx->peer_on ? SOLVE_PHASE_PEER : SOLVE_PHASE_FULL;
    const int sgrid = search_grid(ctx), fgrid = fused_grid(ctx);
    int launches = 1;
    auto launch_trip = [&](int t, bool with_events) -> int {
        if (with_events) CKC(cudaEventRecord(ctx->ev[1 + 2 * t], s));
        launch_search(P, sgrid, s, par_probes(ctx, sgrid));   // exits at once on re-use trips
        if (with_events) CKC(cudaEventRecord(ctx->ev[2 + 2 * t], s));
        if (ctx->comm && !ctx->peer_on) {
            S.phase = SOLVE_PHASE_REDUCE_ONLY;          // k_fused leaves this rank's sums in d_sum
            launch_fused(P, S, fgrid, s);
            int nrc = g_nccl.AllGather(ctx->d_sum, ctx->d_rank_blocks, PSTRIDE, NCCL_DOUBLE, ctx->comm, s);
            if (nrc != 0) { ctx->err = std::string("ncclAllGather: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "error"); return FLIVO_E_NCCL; }
            S.phase = SOLVE_PHASE_FROM_RANKS;
            launch_solve_ranks(S, s);
            launches += 4;
        } else {
            S.phase = full_phase;
            launch_fused(P, S, fgrid, s);              // Gram + (in its last CTA) the solve
            launches += 2;
        }
        return FLIVO_OK;
    };
    if (use_graph) {
        // the sequence depends only on (P, S, grids, trips): replay the captured graph unless one of them changed
        S.phase = full_phase;
        const bool same = ctx->graph_exec && memcmp(&P, &ctx->graph_P, sizeof(P)) == 0 && memcmp(&S, &ctx->graph_S, sizeof(S)) == 0 &&
                          sgrid == ctx->graph_search_grid && fgrid == ctx->graph_fused_grid && trips == ctx->graph_trips;
        if (!same) {
            // parameters that keep changing (e.g. a new external unsorted scan pointer every call) would re-capture on every
            // call: after three misses in a row fall back to plain launches for 16 calls, then try the graph again
            if (ctx->graph_skip > 0) {
                --ctx->graph_skip;
                use_graph = false;
            } else if (++ctx->graph_churn > 3) {
                use_graph = false;
                ctx->graph_churn = 0; ctx->graph_skip = 16;
            } else {
                if (ctx->graph_exec) { cudaGraphExecDestroy(ctx->graph_exec); ctx->graph_exec = nullptr; }
                int rc = build_update_graph(ctx, P, S, sgrid, fgrid, trips);
                if (rc != FLIVO_OK) return rc;
                ctx->graph_P = P; ctx->graph_S = S; ctx->graph_search_grid = sgrid; ctx->graph_fused_grid = fgrid; ctx->graph_trips = trips;
                launches = 1;
            }
        } else {
            ctx->graph_churn = 0;
        }
    }
    if (use_graph) {
        CKC(cudaGraphLaunch(ctx->graph_exec, s));   // H2D of the inputs, control reset, the trips, D2H of the results
        launches = 1 + trips * 2;
    } else {
        CKC(cudaEventRecord(ctx->ev_upd0, s));
        CKC(cudaMemcpyAsync(ctx->d_blk, h, offsetof(DevBlock, P_out), cudaMemcpyHostToDevice, s));   // job + P_prop + x
        S.phase = full_phase;
        launch_ctl_reset(S, ctx->d_blk->x_prop, ctx->d_stats, s);
        launches = 1;
        for (int t = 0; t < trips; ++t) { int rc = launch_trip(t, timed); if (rc != FLIVO_OK) return rc; }
    }
    CKC(cudaGetLastError());
    if (!use_graph) {
        CKC(cudaEventRecord(ctx->ev[1 + 2 * trips], s));
        CKC(cudaMemcpyAsync(h->x, ctx->d_blk->x, offsetof(DevBlock, x_prop) - offsetof(DevBlock, x), cudaMemcpyDeviceToHost, s));   // x, P_out, ctl, stats
        CKC(cudaEventRecord(ctx->ev_upd1, s));
    }
    ctx->upd.active = true; ctx->upd.np = np; ctx->upd.trips = trips; ctx->upd.launches = launches + (staged ? ctx->stage_launches : 0);
    ctx->upd.timed = timed; ctx->upd.staged = staged;
    return FLIVO_OK;
}

extern "C" int flivo_iekf_update_end(flivo_ctx* ctx, flivo_state* inout, double* Pm, flivo_report* rep) {
    ENTER();
    NEED(inout && Pm, FLIVO_E_ARG, "null state / P");
    NEED(ctx->upd.active, FLIVO_E_STATE, "no update in flight (flivo_iekf_update_begin first)");
    ctx->upd.active = false;
    const int np = ctx->upd.np, trips = ctx->upd.trips, launches = ctx->upd.launches;
    const bool timed = ctx->upd.timed;
    DevBlock* h = ctx->h_blk;
    if (ctx->block_wait) {   // (flivo_frame with more driving threads than cores: sleep instead of spinning)
        if (!ctx->block_ev) CKC(cudaEventCreateWithFlags(&ctx->block_ev, cudaEventBlockingSync | cudaEventDisableTiming));
        CKC(cudaEventRecord(ctx->block_ev, ctx->stream));
        CKC(cudaEventSynchronize(ctx->block_ev));
    } else {
        CKC(cudaStreamSynchronize(ctx->stream));
    }
    const Ctl& c = h->ctl;
    if (c.nan_guard != 2) {   // (a timed-out peer exchange may have stopped the loop after some trips: nothing is handed back)
        memcpy(inout, h->x, sizeof(flivo_state));
        memcpy(Pm, h->P_out, sizeof(double) * np * np);
    }
    ctx->stats.last_far_queries = (int64_t)h->stats.far_queries;
    ctx->stats.last_candidates = (int64_t)h->stats.candidates;
    for (int t = 0; t < 16; ++t) { ctx->stats.last_fused_ms[t] = 0.f; ctx->stats.last_solve_ms[t] = 0.f; ctx->stats.last_trip_searched[t] = 0; }
    for (int t = 0; t < trips && t < 16; ++t) {
        if (timed) {
            // (profile mode) last_fused_ms = the whole trip: k_search + k_fused incl. its solve; last_solve_ms = the k_fused part
            cudaEventElapsedTime(&ctx->stats.last_fused_ms[t], ctx->ev[1 + 2 * t], ctx->ev[3 + 2 * t]);
            cudaEventElapsedTime(&ctx->stats.last_solve_ms[t], ctx->ev[2 + 2 * t], ctx->ev[3 + 2 * t]);
        }
        ctx->stats.last_trip_searched[t] = t < c.trips ? c.trip_searched[t] : 0;
    }
    if (rep) {
        memset(rep, 0, sizeof(*rep));
        rep->trips = c.trips; rep->searches = c.searches; rep->n_eff = c.n_eff; rep->converged = c.converged;
        rep->res_mean = c.res_mean;
        for (int i = 0; i < 23; ++i) rep->dx_last[i] = c.dx_last[i];
        // device time from the first staging node (when this update staged its scan) to the last node of the update
        cudaEventElapsedTime(&rep->gpu_ms, ctx->upd.staged ? ctx->ev_stage0 : ctx->ev_upd0, ctx->ev_upd1);
        rep->kernel_launches = launches;   // own kernels: k_ctl_reset + 2 per trip (+ the staging kernels when the scan was staged here)
    }
    if (c.nan_guard == 2) {
        ctx->peer_broken = true;
        ctx->err = "peer exchange timed out: update skipped; the ranks are out of step until every rank calls flivo_peer_reset (after a barrier)";
        return FLIVO_E_STATE;
    }
    if (c.nan_guard) { ctx->err = "non-finite normal equations: update skipped"; return FLIVO_E_STATE; }
    return FLIVO_OK;
}

extern "C" int flivo_iekf_update(flivo_ctx* ctx, flivo_state* inout, double* Pm, const flivo_cfg* cfg, flivo_report* rep) {
    int rc = flivo_iekf_update_begin(ctx, inout, Pm, cfg);
    if (rc != FLIVO_OK) return rc;
    return flivo_iekf_update_end(ctx, inout, Pm, rep);
}

// iVox branch of map_incremental (src/laser_mapping.cpp:340-387); the state is already in d_x
static int ivox_add_scan(flivo_ctx* ctx, int32_t* n_added) {
    const size_t n = ctx->scan_n;
    if (n > ctx->ivox_cap) {
        cudaFree(ctx->d_ivox_flags); cudaFree(ctx->d_ivox_sel);
        ctx->d_ivox_flags = nullptr; ctx->d_ivox_sel = nullptr; ctx->ivox_cap = 0;
        const size_t cap = n + n / 8;
        CKC(cudaMalloc(&ctx->d_ivox_flags, 2 * cap));
        CKC(cudaMalloc(&ctx->d_ivox_sel, (2 * cap + 4) * sizeof(float4)));
        ctx->ivox_cap = cap;
    }
    const size_t cap = ctx->ivox_cap;
    unsigned char *f_add = ctx->d_ivox_flags, *f_non = ctx->d_ivox_flags + cap;
    float4 *sel_add = ctx->d_ivox_sel, *sel_non = ctx->d_ivox_sel + cap;
    int* d_counts = reinterpret_cast<int*>(ctx->d_ivox_sel + 2 * cap);
    launch_ivox_map_filter(ctx->d_body, (int)n, ctx->d_x, ctx->d_near, ctx->d_near_meta, ctx->scan_sorted ? ctx->d_perm : nullptr,
                           ctx->ivox_filter, ctx->d_world, f_add, f_non, ctx->stream);
    CKC(cudaGetLastError());
    size_t tmp = ctx->cub_bytes;
    size_t need = 0;
    CKC(cub::DeviceSelect::Flagged(nullptr, need, ctx->d_world, f_add, sel_add, d_counts, (int)n, ctx->stream));
    if (need > tmp) {
        if (ctx->sort_exec) { cudaGraphExecDestroy(ctx->sort_exec); ctx->sort_exec = nullptr; ctx->sort_npad = 0; }   // captured with the old buffer
        cudaFree(ctx->d_cub); ctx->d_cub = nullptr; ctx->cub_bytes = 0;
        CKC(cudaMalloc(&ctx->d_cub, need + 256));
        ctx->cub_bytes = need + 256;
    }
    tmp = ctx->cub_bytes;
    CKC(cub::DeviceSelect::Flagged(ctx->d_cub, tmp, ctx->d_world, f_add, sel_add, d_counts, (int)n, ctx->stream));
    tmp = ctx->cub_bytes;
    CKC(cub::DeviceSelect::Flagged(ctx->d_cub, tmp, ctx->d_world, f_non, sel_non, d_counts + 1, (int)n, ctx->stream));
    int h[2] = {0, 0};
    CKC(cudaMemcpyAsync(h, d_counts, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    // ivox_->AddPoints(points_to_add); ivox_->AddPoints(point_no_need_downsample); (:386-387) — one concatenated batch
    // stores the same points with the same insertion ids
    if (h[1] > 0) CKC(cudaMemcpyAsync(sel_add + h[0], sel_non, (size_t)h[1] * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
    const size_t total = (size_t)h[0] + (size_t)h[1];
    int added = 0;
    int rc = total ? ctx->map.add_points(sel_add, total, false, &added, ctx->err) : 0;
    if (n_added) *n_added = (int32_t)total;
    return map_rc(ctx, rc);
}

extern "C" int flivo_set_gnss(flivo_ctx* ctx, const double pos_global[3], const double antlever[3], const double std3[3]) {
    ENTER();
    if (!pos_global) { ctx->gnss_on = false; return FLIVO_OK; }
    NEED(antlever && std3, FLIVO_E_ARG, "null antlever / std");
    NEED(std3[0] > 0.0 && std3[1] > 0.0 && std3[2] > 0.0, FLIVO_E_ARG, "gnss std must be > 0");
    for (int i = 0; i < 3; ++i) { ctx->gnss[i] = pos_global[i]; ctx->gnss[3 + i] = antlever[i]; ctx->gnss[6 + i] = std3[i]; }
    ctx->gnss_on = true;
    return FLIVO_OK;
}

extern "C" int flivo_map_add_scan(flivo_ctx* ctx, const flivo_state* st, int32_t* n_added) {
    ENTER();
    if (n_added) *n_added = 0;
    NEED(st, FLIVO_E_ARG, "null state");
    NEED(ctx->scan_n > 0, FLIVO_E_STATE, "no scan uploaded");
    NEED(!ctx->upd.active, FLIVO_E_STATE, "an update is in flight (flivo_iekf_update_end first)");
    { int rcs = ensure_staged(ctx); if (rcs) return rcs; }
    memcpy(ctx->h_blk->x, st, sizeof(flivo_state));
    CKC(cudaMemcpyAsync(ctx->d_x, ctx->h_blk->x, sizeof(flivo_state), cudaMemcpyHostToDevice, ctx->stream));
    if (ctx->opt.ivox_nearby) return ivox_add_scan(ctx, n_added);
    // Add_Points is order dependent only through ties; keep the caller's order (batch index = insertion id order)
    const float4* src = ctx->scan_sorted ? ctx->d_body_in : ctx->d_body;   // (k_scan_prepare keeps the caller-order float4 copy)
    launch_transform(src, (int)ctx->scan_n, ctx->d_x, ctx->d_world, ctx->stream);
    CKC(cudaGetLastError());
    int added = 0;
    int rc = ctx->map.add_points(ctx->d_world, ctx->scan_n, true, &added, ctx->err);
    if (n_added) *n_added = added;
    return map_rc(ctx, rc);
}

// One frame of LaserMapping::Run's step 5 (src/laser_mapping.cpp:165-251) in one call, for replay loops and bindings whose
// per-call overhead matters: the scan staged by the previous frame (or by flivo_scan_upload_async) becomes current, the
// update runs, the NEXT frame's scan is staged meanwhile, then map_incremental with the posterior.
extern "C" int flivo_frame(flivo_ctx* ctx, flivo_state* x, double* P, const flivo_cfg* cfg, const float* next_body_xyz, size_t n_next,
                           size_t stride_bytes, flivo_report* rep, int32_t* n_added) {
    static const bool trace = getenv("FLIVO_FRAME_TRACE") != nullptr;   // where a frame's host time goes (printed by flivo_ctx_destroy)
    // A replay drives several contexts from as many host threads; where a process has few cores to itself (eight ranks on
    // one host) spinning waits starve each other, so the frame's two waits sleep on an event instead (FLIVO_BLOCKING_SYNC=0/1
    // overrides; the price is the wake-up latency, which a loop with several frames in flight does not see).
    static const int block_mode = getenv("FLIVO_BLOCKING_SYNC") ? atoi(getenv("FLIVO_BLOCKING_SYNC")) : (host_threads() <= 4 ? 1 : 0);
    if (ctx) { ctx->block_wait = block_mode != 0; ctx->map.set_blocking_wait(block_mode != 0); }
    struct Unblock { flivo_ctx* c; ~Unblock() { if (c) { c->block_wait = false; c->map.set_blocking_wait(false); } } } unblock{ctx};
    using clk = std::chrono::steady_clock;
    clk::time_point t0;
    int phase = 0;
    auto lap = [&]() { if (trace && ctx) { const clk::time_point t1 = clk::now(); ctx->frame_s[phase] += std::chrono::duration<double>(t1 - t0).count(); t0 = t1; } ++phase; };
    if (trace) t0 = clk::now();
    int rc = flivo_scan_commit(ctx);
    if (rc != FLIVO_OK) return rc;
    lap();
    if ((rc = flivo_iekf_update_begin(ctx, x, P, cfg)) != FLIVO_OK) return rc;
    lap();
    int rc_up = FLIVO_OK;
    if (next_body_xyz && n_next) rc_up = flivo_scan_upload_async(ctx, next_body_xyz, n_next, stride_bytes);
    lap();
    if ((rc = flivo_iekf_update_end(ctx, x, P, rep)) != FLIVO_OK) return rc;
    lap();
    if (rc_up != FLIVO_OK) return rc_up;
    rc = flivo_map_add_scan(ctx, x, n_added);
    lap();
    if (trace) ++ctx->frame_calls;
    return rc;
}

// ---- multi-GPU ----------------------------------------------------------------------------------------------
extern "C" int flivo_comm_unique_id(void* out128) {
    if (!out128) return FLIVO_E_ARG;
    std::string err;
    if (!g_nccl.load(err)) { fprintf(stderr, "flivo_comm_unique_id: %s\n", err.c_str()); return FLIVO_E_NCCL; }
    ncclUniqueId_t id;
    if (g_nccl.GetUniqueId(&id) != 0) return FLIVO_E_NCCL;
    memcpy(out128, &id, 128);
    return FLIVO_OK;
}

extern "C" int flivo_comm_init(flivo_ctx* ctx, const void* unique_id128, int rank, int nranks) {
    ENTER();
    NEED(unique_id128 && nranks >= 1 && rank >= 0 && rank < nranks, FLIVO_E_ARG, "bad rank / nranks / id");
    NEED(!ctx->comm, FLIVO_E_STATE, "communicator already attached");
    if (!g_nccl.load(ctx->err)) return FLIVO_E_NCCL;
    ncclUniqueId_t id;
    memcpy(&id, unique_id128, 128);
    int rc = g_nccl.CommInitRank(&ctx->comm, nranks, id, rank);
    if (rc != 0) { ctx->comm = nullptr; ctx->err = std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"); return FLIVO_E_NCCL; }
    ctx->rank = rank; ctx->nranks = nranks;
    cudaFree(ctx->d_rank_blocks); ctx->d_rank_blocks = nullptr;
    CKC(cudaMalloc(&ctx->d_rank_blocks, (size_t)nranks * PSTRIDE * sizeof(double)));
    CKC(cudaMemsetAsync(ctx->d_rank_blocks, 0, (size_t)nranks * PSTRIDE * sizeof(double), ctx->stream));
    return FLIVO_OK;
}

extern "C" int flivo_comm_destroy(flivo_ctx* ctx) {
    ENTER();
    if (ctx->comm) {
        CKC(cudaStreamSynchronize(ctx->stream));
        g_nccl.CommDestroy(ctx->comm);
        ctx->comm = nullptr;
    }
    ctx->rank = 0; ctx->nranks = 1;
    return FLIVO_OK;
}

// ---- peer-memory exchange over NVLink / NVSwitch (replaces the NCCL all-gather of the sharded update) ---------------
// Every rank allocates one exchange buffer ([2 parities][PEER_MAX_RANKS] records + flags), exports it as a cudaIpc
// handle, and maps the buffers of all ranks; k_solve (SOLVE_PHASE_PEER) then writes its 96-double record straight into
// every peer's buffer, publishes a sequence number, waits for the peers' sequence numbers in its own buffer and sums
// the records in rank order — compute, exchange and solve in one kernel, and the whole update stays one CUDA graph.
extern "C" int flivo_peer_export(flivo_ctx* ctx, void* handle_out64) {
    ENTER();
    NEED(handle_out64, FLIVO_E_ARG, "null handle");
    static_assert(sizeof(cudaIpcMemHandle_t) == FLIVO_PEER_HANDLE_BYTES, "ipc handle size");
    if (!ctx->d_xchg) {
        const size_t bytes = sizeof(PeerSlot) * 2 * PEER_MAX_RANKS;
        CKC(cudaMalloc(&ctx->d_xchg, bytes));
        CKC(cudaMemset(ctx->d_xchg, 0, bytes));
        CKC(cudaMalloc(&ctx->d_xseq, sizeof(unsigned long long)));
        CKC(cudaMemset(ctx->d_xseq, 0, sizeof(unsigned long long)));
    }
    cudaIpcMemHandle_t h;
    CKC(cudaIpcGetMemHandle(&h, ctx->d_xchg));
    memcpy(handle_out64, &h, sizeof(h));
    return FLIVO_OK;
}

extern "C" int flivo_peer_attach(flivo_ctx* ctx, const void* handles, int rank, int nranks) {
    ENTER();
    NEED(handles && nranks >= 1 && nranks <= PEER_MAX_RANKS && rank >= 0 && rank < nranks, FLIVO_E_ARG, "bad rank / nranks (<= 8) / handles");
    NEED(ctx->d_xchg, FLIVO_E_STATE, "call flivo_peer_export first");
    NEED(!ctx->peer_on, FLIVO_E_STATE, "peers already attached");
    const char* hp = static_cast<const char*>(handles);
    for (int r = 0; r < nranks; ++r) {
        if (r == rank) { ctx->peer_base[r] = ctx->d_xchg; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, hp + (size_t)r * sizeof(h), sizeof(h));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            ctx->err = std::string("cudaIpcOpenMemHandle (rank ") + std::to_string(r) + "): " + cudaGetErrorString(e);
            for (int q = 0; q < r; ++q) if (q != rank && ctx->peer_base[q]) { cudaIpcCloseMemHandle(ctx->peer_base[q]); ctx->peer_base[q] = nullptr; }
            return FLIVO_E_CUDA;
        }
        ctx->peer_base[r] = static_cast<PeerSlot*>(p);
    }
    ctx->rank = rank; ctx->nranks = nranks;
    ctx->peer_on = true; ctx->peer_local = false; ctx->peer_broken = false;
    return FLIVO_OK;
}

// Same exchange between contexts of ONE process (several GPUs driven by one process, or several ranks on one GPU as the
// tests and the sanitizer runs do): cudaIpc handles cannot be opened by the process that exported them, so the ranks
// pass each other their buffers as device pointers.  With distinct devices, peer access must be enabled by the caller.
extern "C" int flivo_peer_local_base(flivo_ctx* ctx, void** base_out) {
    ENTER();
    NEED(base_out, FLIVO_E_ARG, "null out");
    char h[FLIVO_PEER_HANDLE_BYTES];
    int rc = flivo_peer_export(ctx, h);   // allocates the buffer on first use
    if (rc != FLIVO_OK) return rc;
    *base_out = ctx->d_xchg;
    return FLIVO_OK;
}
extern "C" int flivo_peer_attach_local(flivo_ctx* ctx, void* const* bases, int rank, int nranks) {
    ENTER();
    NEED(bases && nranks >= 1 && nranks <= PEER_MAX_RANKS && rank >= 0 && rank < nranks, FLIVO_E_ARG, "bad rank / nranks (<= 8) / bases");
    NEED(ctx->d_xchg, FLIVO_E_STATE, "call flivo_peer_local_base first");
    NEED(!ctx->peer_on, FLIVO_E_STATE, "peers already attached");
    NEED(bases[rank] == (void*)ctx->d_xchg, FLIVO_E_ARG, "bases[rank] is not this context's buffer");
    for (int r = 0; r < nranks; ++r) { NEED(bases[r], FLIVO_E_ARG, "null base"); ctx->peer_base[r] = static_cast<PeerSlot*>(bases[r]); }
    ctx->rank = rank; ctx->nranks = nranks;
    ctx->peer_on = true; ctx->peer_local = true; ctx->peer_broken = false;
    return FLIVO_OK;
}

// Re-arm the exchange after a time-out (or before reusing a set of ranks): zeroes this rank's sequence counter and its
// exchange buffer.  Every rank must call it, with a barrier of the caller's before (no update in flight anywhere) and
// after (no rank starts an update while another still clears its buffer).
extern "C" int flivo_peer_reset(flivo_ctx* ctx) {
    ENTER();
    NEED(ctx->d_xchg, FLIVO_E_STATE, "no exchange buffer (flivo_peer_export first)");
    CKC(cudaStreamSynchronize(ctx->stream));
    CKC(cudaMemsetAsync(ctx->d_xchg, 0, sizeof(PeerSlot) * 2 * PEER_MAX_RANKS, ctx->stream));
    CKC(cudaMemsetAsync(ctx->d_xseq, 0, sizeof(unsigned long long), ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    ctx->peer_broken = false;
    return FLIVO_OK;
}

// How long a rank waits for a peer's record before it gives the update up (default 2 s).
extern "C" int flivo_peer_set_timeout(flivo_ctx* ctx, double seconds) {
    ENTER();
    NEED(seconds > 0.0 && seconds < 3600.0, FLIVO_E_ARG, "timeout must be in (0, 3600) s");
    ctx->peer_timeout_ns = (unsigned long long)(seconds * 1e9);
    return FLIVO_OK;
}

extern "C" int flivo_peer_detach(flivo_ctx* ctx) {
    ENTER();
    if (ctx->peer_on) {
        CKC(cudaStreamSynchronize(ctx->stream));
        for (int r = 0; r < ctx->nranks; ++r) {
            if (!ctx->peer_local && r != ctx->rank && ctx->peer_base[r]) cudaIpcCloseMemHandle(ctx->peer_base[r]);
            ctx->peer_base[r] = nullptr;
        }
        ctx->peer_on = false; ctx->peer_local = false;
        if (!ctx->comm) { ctx->rank = 0; ctx->nranks = 1; }
    }
    return FLIVO_OK;
}

extern "C" int flivo_get_stats(flivo_ctx* ctx, flivo_stats* out) {
    ENTER();
    NEED(out, FLIVO_E_ARG, "null out");
    int rc = ctx->map.refresh_counters(ctx->err);
    const MapCounters& c = ctx->map.counters();
    ctx->stats.map_live = (int64_t)c.live;
    ctx->stats.map_cells = c.cells_used;
    ctx->stats.map_blocks = c.blocks_used;
    ctx->stats.dir_capacity = (int64_t)ctx->map.dir_cap();
    ctx->stats.bucket_capacity = (int64_t)ctx->map.bucket_cap();
    *out = ctx->stats;
    return map_rc(ctx, rc);
}

// Diagnostic: globaltimer stamps (ns) at the phase boundaries of k_solve for each trip of the last update that ran
// with FLIVO_CFG_PROFILE.  out: [trips][16]; entries 0..9 are used (see k_solve), 0 where a phase did not run.
extern "C" int flivo_debug_solve_stamps(flivo_ctx* ctx, uint64_t* out, int trips_cap) {
    if (!ctx || !out || trips_cap <= 0) return FLIVO_E_ARG;
    if (trips_cap > MAX_TRIPS) trips_cap = MAX_TRIPS;
    CKC(cudaMemcpyAsync(out, ctx->d_stamps, (size_t)trips_cap * 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

// =====================================================================================================================
// Front end of the path: raw scan -> motion compensation -> voxel-grid down-sampling, on the device
// =====================================================================================================================
static void free_raw(flivo_ctx* ctx) {
    cudaFree(ctx->d_raw); cudaFree(ctx->d_raw_t); cudaFree(ctx->d_ds); cudaFree(ctx->d_ds_t);
    cudaFree(ctx->d_vg_keys_a); cudaFree(ctx->d_vg_keys_b); cudaFree(ctx->d_vg_idx_a); cudaFree(ctx->d_vg_idx_b);
    cudaFree(ctx->d_vg_head); cudaFree(ctx->d_vg_pos); cudaFree(ctx->d_vg_tmp);
    ctx->d_raw = ctx->d_ds = nullptr; ctx->d_raw_t = ctx->d_ds_t = nullptr;
    ctx->d_vg_keys_a = ctx->d_vg_keys_b = ctx->d_vg_idx_a = ctx->d_vg_idx_b = ctx->d_vg_head = ctx->d_vg_pos = nullptr;
    ctx->d_vg_tmp = nullptr; ctx->vg_tmp_bytes = 0; ctx->raw_cap = 0; ctx->raw_n = 0;
}
static int ensure_raw(flivo_ctx* ctx, size_t n) {
    if (!ctx->d_vg_state) {
        CKC(cudaMalloc(&ctx->d_vg_state, sizeof(VoxelGridState)));
        CKC(cudaMemset(ctx->d_vg_state, 0, sizeof(VoxelGridState)));
        CKC(cudaMalloc(&ctx->d_poses, UNDISTORT_MAX_POSES * sizeof(PoseDev)));
    }
    if (n <= ctx->raw_cap) return FLIVO_OK;
    CKC(cudaStreamSynchronize(ctx->stream));
    free_raw(ctx);
    const size_t cap = std::max<size_t>(n + n / 8, 1u << 16);
    CKC(cudaMalloc(&ctx->d_raw, cap * sizeof(float4)));
    CKC(cudaMalloc(&ctx->d_raw_t, cap * sizeof(float)));
    CKC(cudaMalloc(&ctx->d_ds, cap * sizeof(float4)));
    CKC(cudaMalloc(&ctx->d_ds_t, cap * sizeof(float)));
    unsigned** u[6] = {&ctx->d_vg_keys_a, &ctx->d_vg_keys_b, &ctx->d_vg_idx_a, &ctx->d_vg_idx_b, &ctx->d_vg_head, &ctx->d_vg_pos};
    for (auto pp : u) CKC(cudaMalloc(pp, cap * sizeof(unsigned)));
    ctx->vg_tmp_bytes = voxel_grid_temp_bytes(cap);
    CKC(cudaMalloc(&ctx->d_vg_tmp, ctx->vg_tmp_bytes));
    ctx->raw_cap = cap;
    return FLIVO_OK;
}

// lidar_meas.lidar->points as the pre-processing step leaves them (PointXYZINormal: xyz at byte 0, intensity at 32,
// curvature = time offset in ms at 36): copied to the device once; the following two calls work in HBM.
extern "C" int flivo_raw_upload(flivo_ctx* ctx, const void* pts, size_t n, size_t stride_bytes, int time_offset_bytes, int intensity_offset_bytes) {
    ENTER();
    NEED(n > 0 && pts && stride_ok(stride_bytes), FLIVO_E_ARG, "bad raw scan / stride");
    NEED(n < 0x7FFFFFF0ull / 8, FLIVO_E_ARG, "raw scan too large");
    NEED(time_offset_bytes < 0 || ((size_t)time_offset_bytes + 4 <= stride_bytes && time_offset_bytes % 4 == 0), FLIVO_E_ARG, "bad time offset");
    NEED(intensity_offset_bytes < 0 || ((size_t)intensity_offset_bytes + 4 <= stride_bytes && intensity_offset_bytes % 4 == 0), FLIVO_E_ARG, "bad intensity offset");
    int rc = ensure_raw(ctx, n);
    if (rc) return rc;
    if (ctx->stage_pending) { CKC(cudaEventSynchronize(ctx->stage_ev)); ctx->stage_pending = false; }
    const size_t need = n + (n + 3) / 4;   // float4 points followed by the float times in the same pinned buffer
    if (need > ctx->h_stage_n) {
        if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
        ctx->h_stage = nullptr; ctx->h_stage_n = 0;
        const size_t cap = std::max<size_t>(need + need / 4, 1u << 16);
        CKC(cudaMallocHost(&ctx->h_stage, cap * sizeof(float4)));
        ctx->h_stage_n = cap;
    }
    const char* src = reinterpret_cast<const char*>(pts);
    float4* h = ctx->h_stage;
    float* ht = reinterpret_cast<float*>(ctx->h_stage + n);
    const int nt = host_threads();
    const size_t CH = 1u << 16;
    for (size_t c0 = 0; c0 < n; c0 += CH) {
        const size_t m = std::min(CH, n - c0);
#pragma omp parallel for num_threads(nt) schedule(static) if (m >= 8192 && nt > 1)
        for (long i = 0; i < (long)m; ++i) {
            const char* rec = src + (c0 + (size_t)i) * stride_bytes;
            const float* p = reinterpret_cast<const float*>(rec);
            const float inten = intensity_offset_bytes >= 0 ? *reinterpret_cast<const float*>(rec + intensity_offset_bytes) : 0.f;
            h[c0 + i] = make_float4(p[0], p[1], p[2], inten);
            ht[c0 + i] = time_offset_bytes >= 0 ? *reinterpret_cast<const float*>(rec + time_offset_bytes) : 0.f;
        }
        CKC(cudaMemcpyAsync(ctx->d_raw + c0, h + c0, m * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
        CKC(cudaMemcpyAsync(ctx->d_raw_t + c0, ht + c0, m * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    }
    CKC(cudaEventRecord(ctx->stage_ev, ctx->stream));
    ctx->stage_pending = true;
    ctx->raw_n = n;
    ctx->raw_has_t = time_offset_bytes >= 0;
    return FLIVO_OK;
}

// Same, from device pointers (n float4 x,y,z,intensity and n float times in ms, may be NULL): copied device to device.
extern "C" int flivo_raw_set_device(flivo_ctx* ctx, const void* dev_xyzi, const void* dev_time_ms, size_t n) {
    ENTER();
    NEED(n > 0 && dev_xyzi, FLIVO_E_ARG, "bad raw scan");
    NEED(n < 0x7FFFFFF0ull / 8, FLIVO_E_ARG, "raw scan too large");
    int rc = ensure_raw(ctx, n);
    if (rc) return rc;
    CKC(cudaMemcpyAsync(ctx->d_raw, dev_xyzi, n * sizeof(float4), cudaMemcpyDeviceToDevice, ctx->stream));
    if (dev_time_ms) CKC(cudaMemcpyAsync(ctx->d_raw_t, dev_time_ms, n * sizeof(float), cudaMemcpyDeviceToDevice, ctx->stream));
    ctx->raw_n = n;
    ctx->raw_has_t = dev_time_ms != nullptr;
    return FLIVO_OK;
}

// The per-point loop of ImuProcess::UndistortPcl (src/IMU_Processing.cpp:336-376), in place on the raw scan.
extern "C" int flivo_raw_undistort(flivo_ctx* ctx, const flivo_pose6d* poses, int n_poses, const double rot_end[9], const double pos_end[3],
                                   const double R_LI[9], const double t_LI[3]) {
    ENTER();
    NEED(ctx->raw_n > 0, FLIVO_E_STATE, "no raw scan uploaded");
    NEED(ctx->raw_has_t, FLIVO_E_STATE, "raw scan has no per-point times");
    NEED(poses && n_poses >= 2 && n_poses <= UNDISTORT_MAX_POSES, FLIVO_E_ARG, "need 2..256 IMU poses");
    NEED(rot_end && pos_end && R_LI && t_LI, FLIVO_E_ARG, "null argument");
    static_assert(sizeof(flivo_pose6d) == sizeof(PoseDev), "pose layout");
    UndistortParams U;
    memset(&U, 0, sizeof(U));
    U.poses = ctx->d_poses; U.n_poses = n_poses;
    // (a copy from pageable host memory is staged by the driver before the call returns: the caller's array is free at once)
    CKC(cudaMemcpyAsync(ctx->d_poses, poses, (size_t)n_poses * sizeof(PoseDev), cudaMemcpyHostToDevice, ctx->stream));
    for (int i = 0; i < 3; ++i) { U.pos_end[i] = pos_end[i]; U.t_LI[i] = t_LI[i]; }
    for (int i = 0; i < 9; ++i) U.R_LI[i] = R_LI[i];
    // extR_Ri = R_LI^T * rot_end^T ; exrR_extT = R_LI^T * t_LI   (:336-337)
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            double sacc = 0.0;
            for (int k = 0; k < 3; ++k) sacc += R_LI[k * 3 + r] * rot_end[c * 3 + k];
            U.extR_Ri[r * 3 + c] = sacc;
        }
        double t = 0.0;
        for (int k = 0; k < 3; ++k) t += R_LI[k * 3 + r] * t_LI[k];
        U.exrR_extT[r] = t;
    }
    launch_undistort(ctx->d_raw, ctx->d_raw_t, ctx->raw_n, U, ctx->stream);
    CKC(cudaGetLastError());
    return FLIVO_OK;
}

// downSizeFilterSurf.setInputCloud(feats_undistort); downSizeFilterSurf.filter(*feats_down_body) (src/laser_mapping.cpp:129-130)
// on the raw scan in HBM; the result becomes the current scan exactly as after flivo_scan_upload (feats_down_body).
extern "C" int flivo_raw_downsample(flivo_ctx* ctx, float leaf_size, size_t* n_out) {
    ENTER();
    NEED(ctx->raw_n > 0, FLIVO_E_STATE, "no raw scan uploaded");
    NEED(leaf_size > 0.f, FLIVO_E_ARG, "leaf size must be positive");
    const size_t n = ctx->raw_n;
    const float* tms = ctx->raw_has_t ? ctx->d_raw_t : nullptr;
    CKC(launch_voxel_grid(ctx->d_raw, tms, n, leaf_size, ctx->d_vg_state, ctx->d_vg_keys_a, ctx->d_vg_keys_b, ctx->d_vg_idx_a, ctx->d_vg_idx_b,
                          ctx->d_vg_head, ctx->d_vg_pos, ctx->d_vg_tmp, ctx->vg_tmp_bytes, ctx->d_ds, ctx->d_ds_t, ctx->stream));
    VoxelGridState* hs = reinterpret_cast<VoxelGridState*>(ctx->h_small);
    CKC(cudaMemcpyAsync(hs, ctx->d_vg_state, sizeof(VoxelGridState), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    if (hs->overflow) {   // PCL: "Leaf size is too small for the input dataset. Integer indices would overflow." -> output = input
        CKC(launch_voxel_passthrough(ctx->d_raw, tms, n, ctx->d_vg_state, ctx->d_ds, ctx->d_ds_t, ctx->stream));
        hs->n_out = (unsigned)n;
    }
    const size_t m = hs->n_out;
    if (n_out) *n_out = m;
    if (m == 0) { ctx->scan_n = 0; return FLIVO_OK; }
    int rc = ensure_scan(ctx, m);
    if (rc) return rc;
    return finish_scan(ctx, (const char*)ctx->d_ds, 16, m, true, -1);
}

// feats_down_body back to the host in the caller's (= leaf) order: n x (x, y, z, intensity)
extern "C" int flivo_scan_download(flivo_ctx* ctx, float* out_xyzi, size_t cap, size_t* n_out) {
    ENTER();
    NEED(out_xyzi || cap == 0, FLIVO_E_ARG, "null output");
    const size_t n = ctx->scan_n;
    if (n_out) *n_out = n;
    if (n == 0 || cap == 0) return FLIVO_OK;
    NEED(cap >= n, FLIVO_E_CAPACITY, "output too small");
    { int rcs = ensure_staged(ctx); if (rcs) return rcs; }
    const float4* src = ctx->d_body;
    if (ctx->scan_sorted) {   // processing order -> caller's order
        const size_t bytes = n * sizeof(float4);
        if (bytes > ctx->d_out_bytes) {
            cudaFree(ctx->d_out); ctx->d_out = nullptr; ctx->d_out_bytes = 0;
            CKC(cudaMalloc(&ctx->d_out, bytes + bytes / 4));
            ctx->d_out_bytes = bytes + bytes / 4;
        }
        k_scatter<float4><<<nblk(n, 256), 256, 0, ctx->stream>>>(ctx->d_body, ctx->d_perm, n, 1, (float4*)ctx->d_out);
        CKC(cudaGetLastError());
        src = (const float4*)ctx->d_out;
    }
    CKC(cudaMemcpyAsync(out_xyzi, src, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

// the raw scan (after flivo_raw_undistort: feats_undistort) back to the host: n x (x, y, z, intensity)
extern "C" int flivo_raw_download(flivo_ctx* ctx, float* out_xyzi, size_t cap, size_t* n_out) {
    ENTER();
    NEED(out_xyzi || cap == 0, FLIVO_E_ARG, "null output");
    const size_t n = ctx->raw_n;
    if (n_out) *n_out = n;
    if (n == 0 || cap == 0) return FLIVO_OK;
    NEED(cap >= n, FLIVO_E_CAPACITY, "output too small");
    CKC(cudaMemcpyAsync(out_xyzi, ctx->d_raw, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}

extern "C" int flivo_debug_scan_order(flivo_ctx* ctx, uint32_t* out, size_t cap) {
    ENTER();
    NEED(out, FLIVO_E_ARG, "null output");
    NEED(ctx->scan_n > 0, FLIVO_E_STATE, "no scan uploaded");
    NEED(cap >= ctx->scan_n, FLIVO_E_ARG, "output smaller than the scan");
    { int rcs = ensure_staged(ctx); if (rcs) return rcs; }
    if (ctx->scan_sorted) {
        CKC(cudaMemcpyAsync(out, ctx->d_perm, ctx->scan_n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        CKC(cudaStreamSynchronize(ctx->stream));
    } else {
        for (size_t i = 0; i < ctx->scan_n; ++i) out[i] = (uint32_t)i;
    }
    return FLIVO_OK;
}

// Diagnostic: per-CTA globaltimer stamps of the FIRST search trip of the last profile-mode update:
// out[cta][4] = {start, end of phase 1 (tiles), end, far queries finished by the CTA}; *n_ctas = launch grid.
extern "C" int flivo_debug_search_stamps(flivo_ctx* ctx, uint64_t* out, int ctas_cap, int* n_ctas) {
    if (!ctx || !out || ctas_cap <= 0) return FLIVO_E_ARG;
    const int g = search_grid(ctx);
    if (n_ctas) *n_ctas = g;
    const int m = std::min(std::min(g, ctas_cap), SSTAMP_CTAS);
    CKC(cudaMemcpyAsync(out, ctx->d_sstamps, (size_t)m * 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (ctas_cap >= SSTAMP_CTAS + 8)   // far-search totals ride behind the per-CTA rows: out[SSTAMP_CTAS*4 .. +32]
        CKC(cudaMemcpyAsync(out + (size_t)SSTAMP_CTAS * 4, ctx->d_sstamps + (size_t)SSTAMP_CTAS * 4, 32 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CKC(cudaStreamSynchronize(ctx->stream));
    return FLIVO_OK;
}
