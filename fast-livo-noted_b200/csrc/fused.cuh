// fused.cuh — parameters shared by the fused per-trip kernel, the batched kNN kernel and the solve kernels.
#pragma once
#include "flivo_dev.cuh"

namespace flivo {

constexpr int UNIT_WARPS = 4;                    // warps per CTA of the fused kernel
constexpr int UNIT_POINTS = UNIT_WARPS * 32;     // scan points per work unit
constexpr int NTERMS = 92;                       // 78 (H^T H upper) + 12 (H^T z) + res_sum + n_eff
constexpr int PSTRIDE = 96;                      // doubles per partial record
constexpr int GROUP_UNITS = 32;                  // (capacity sizing of group_cnt)
constexpr int FOLD_CTAS = 32;                    // CTA records of k_fused folded into one group record by the last CTA of the group
constexpr int MAX_TRIPS = 16;

// Device-resident control block of one iterated update (written by the solve kernel, read by the fused kernel).
struct Ctl {
    int search;        // 1: this trip runs the kNN (nearest_search_en / dyn_share.converge)
    int stopped;       // 1: loop finished; later kernels of the captured sequence exit immediately
    int iter;          // reference loop index (starts at -1)
    int t;             // V3 converge counter t (esekfom.hpp:1631)
    int rematch;       // V1 rematch_num (src/laser_mapping.cpp:166)
    int trips;         // h_share_model calls executed
    int searches;
    int converged;
    int n_eff;
    int nan_guard;     // 1: a non-finite normal-equation block was seen (update skipped); 2: a peer exchange timed out
    int trip_searched[MAX_TRIPS];
    double res_mean;
    double dx_last[23];
    // work counters of the search kernel of the current trip (see k_search); ctl_next_trip() rearms them
    unsigned int tile_counter;   // next tile of 128 scan points to deal
    unsigned int tiles_done;     // tiles whose far points have all been published
    unsigned int far_count;      // far_list entries published
    unsigned int far_taken;      // far_list tickets handed to consumer warps
    unsigned int epoch;          // bumped whenever the counters are rearmed; far_list entries carry epoch+1 as their tag
    unsigned int ready;          // (update sequence number << 8) | trips whose solve has published state + control block
                                 // (release store; k_search's PDL entry acquires it)
};
#if defined(__CUDACC__) || defined(FLIVO_HOST_EMU)
__device__ __forceinline__ void ctl_next_trip(Ctl* ctl) {
    ctl->tile_counter = 0; ctl->tiles_done = 0; ctl->far_count = 0; ctl->far_taken = 0;
    ctl->epoch += 1;
}
#endif

// Per-call scalars, resident in HBM so that the captured launch sequence (CUDA graph) of an update never changes
// its kernel parameters: a call only rewrites this block (one H2D copy together with the state and P).
struct Job {
    int n;                     // scan points
    int units, groups;         // work units of 128 points / group records k_solve sums (= ceil(fused grid / FOLD_CTAS))
    int mode;                  // FLIVO_MODE_*
    int max_iter;
    int force_search;
    int nranks;
    int aux_out;               // 1: k_fused also writes eff / normvec / world (flivo_h_share_model outputs)
    unsigned int useq;         // sequence number of this update (24 bits used): tells a fresh ctl->ready from the previous update's
    uint32_t max_sq_bits;      // kNN distance bound as float bits (INF_BITS = unbounded)
    double R;                  // laser_point_cov
    double limit[23];
    // GNSS position observation folded into the V1 normal equations (src/laser_mapping.cpp:612-617,
    // GNSSProcessing::computeH src/GNSS_Processing.cpp:204-237); see flivo_set_gnss
    int gnss_on, gnss_pad;
    double gnss_pos[3];        // antenna position in the global frame (rot_enu2global_^T * blh + delta_pos_)
    double gnss_lever[3];      // antlever_
    double gnss_std[3];        // gnss_.std
};

// What the CTA of k_fused that folds the last group record does with the trip's sums (solve_dev.cuh: solve_trip)
enum {
    SOLVE_PHASE_FULL = 0,         // solve (single GPU)
    SOLVE_PHASE_REDUCE_ONLY = 1,  // sum_out[92] only (flivo_h_share_model; first half of the multi-GPU trip, NCCL variant)
    SOLVE_PHASE_FROM_RANKS = 2,   // sum rank_blocks[nranks][96] in rank order, then solve (k_solve_ranks: second half of
                                  // the multi-GPU trip, NCCL variant)
    SOLVE_PHASE_PEER = 3          // multi-GPU, fused: write the record into every peer's exchange buffer over NVLink,
                                  // wait for the peers' records, sum in rank order, solve — inside k_fused
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
    unsigned long long* seq;          // this rank's exchange counter (device memory; reset only by flivo_peer_reset)
    unsigned long long timeout_ns;    // a peer's record is awaited this long before the update is given up (nan_guard = 2)
    int rank, nranks;
};

// State-only half of a trip's solve (it needs x and P_prop, not H^T H): prepared at the end of the previous trip's
// solve (for the first trip: by k_ctl_reset), solve_dev.cuh: solve_pre.
struct SolvePre {
    double Pt[23 * 23];   // V3: P_prop with the manifold Jacobians applied, T P T^T (esekfom.hpp:1665-1695)
    double dxnew[23];     // V3: J^{-1}-corrected x (-) x_prop (:1646-1686)
    double vec[18];       // V1: state_propagat - state (common_lib.h:576-587)
    double pad;
};

struct SolveParams {
    const Job* job;
    int phase;
    double* x;                 // device flivo_state, in/out
    const double* x_prop;      // device flivo_state
    const double* P_prop;      // n x n
    double* P_out;             // n x n, written when the loop stops
    double* sum_out;           // [96] reduced block (may be null in FULL phase)
    const double* rank_blocks; // [nranks * PSTRIDE]
    Ctl* ctl;
    SolvePre* pre;             // written by solve_pre, read by the next trip's solve
    PeerXchg peer;             // SOLVE_PHASE_PEER
    unsigned long long* stamps;       // profile mode only: [MAX_TRIPS][16] globaltimer values at the phase boundaries of the solve
};

enum { PF_HAVE5 = 1,    // five neighbours exist (points_near.size() >= NUM_MATCH_POINTS)
       PF_GATE = 2,     // !(pointSearchSqDis[4] > 5)            (src/laser_mapping.cpp:518)
       PF_PLANE = 4 };  // esti_plane accepted the five points  (include/common_lib.h:694-699)

struct FusedParams {
    MapView map;
    const float4* body;        // scan, lidar frame, processing order
    const Job* job;
    const double* state;       // device flivo_state (36 doubles): rot, pos, offR, offT, ...
    float4* near;              // [n*5] cached Nearest_Points (x,y,z,id)
    float2* near_meta;         // [n] {count as float bits, 5th squared distance}
    float4* plane;             // [n] esti_plane of the cached neighbours (a, b, c, d): a function of the neighbours only,
                               //     so it is fitted once per search and reused by every trip until the next search
    unsigned char* pflags;     // [n] PF_* of the cached neighbours
    unsigned char* sel;        // [n] point_selected_surf (persistent across trips in V3 mode)
    unsigned char* eff;        // [n] rows that entered H
    float4* normvec;           // [n] (nx,ny,nz,pd2)
    float4* world;             // [n] feats_down_world
    unsigned long long* far_list;   // [n] (tag << 32 | point) of points the 27-cell shell could not finish: queue between the
                               //     two phases of k_search, entries valid when tag == ctl->epoch + 1
    double* partials;          // [fused grid * PSTRIDE] one record per CTA of k_fused
    double* gpartials;         // [groups * PSTRIDE] records folded FOLD_CTAS at a time (fixed order)
    unsigned int* group_cnt;   // [groups] arrival tickets; the folding CTA resets its entry
    unsigned int* total_cnt;   // arrival ticket of the group folders: the last one sums the group records and solves
    Ctl* ctl;
    SearchStats* stats;
    int trip;                  // position of this launch in the update's trip sequence (k_search under PDL waits for ctl->ready >= trip)
    int pdl;                   // 1: k_search was launched as a programmatic dependent of the previous trip's k_fused
    int pdl_fused;            // 1: k_fused is a programmatic dependent of this trip's k_search, which lets it go early on re-use trips
    unsigned long long* search_stamps;   // profile mode only: [grid][4] globaltimer of CTA start / end of phase 1 / end, far queries done
};

}  // namespace flivo
