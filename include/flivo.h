/*
 * flivo.h — C ABI of libflivo_b200.so: the B200-native (sm_100a) LIO measurement-update hot path of
 * FAST-LIVO and the GPU-resident map that replaces ikd-Tree underneath it.
 *
 * The reference has no FFI layer: the boundary is two C++ classes linked into one
 * executable.  Every entry point below therefore cites the reference C++ member it stands in for
 * (paths relative to the reference tree); the ROS-free drop-in classes KD_TREE / LaserMapping in
 * fast-livo-noted_b200/host/ forward to these calls, and INTEGRATION.md shows the binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; caller owns every host buffer, the library owns device memory;
 *   - every call returns an int status: 0 = FLIVO_OK, negative = error (flivo_last_error(ctx) has text);
 *   - no exceptions cross the ABI; no CPU fallback exists: without a CUDA device every call that
 *     needs one fails with FLIVO_E_CUDA;
 *   - point inputs are strided float triples: stride_bytes = 12 for packed xyz, 16 for float4,
 *     48 for the reference's pcl::PointXYZINormal AoS (include/ikd-Tree/ikd_Tree.h:21-22);
 *   - one ctx = one device + one stream; calls on one ctx are not re-entrant, distinct ctxs are independent.
 */
#ifndef FLIVO_H_
#define FLIVO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLIVO_OK          0
#define FLIVO_E_CUDA     -1   /* CUDA runtime / no device */
#define FLIVO_E_ARG      -2   /* bad argument */
#define FLIVO_E_CAPACITY -3   /* a device table could not grow */
#define FLIVO_E_NCCL     -4   /* communicator setup / collective failed */
#define FLIVO_E_STATE    -5   /* call order (e.g. update before scan upload / map build) */

#define FLIVO_K           5   /* NUM_MATCH_POINTS, include/common_lib.h:37 */

/* Jacobian/solve flavour.  V3 = src/origin_laserMapping.cpp:916-1048 + IKFoM
 * esekf::update_iterated_dyn_share_modified (1x12 row, 23-DOF manifold state; what north_star names).
 * V1 = the compiled LaserMapping::h_share_model / Run (src/laser_mapping.cpp:485-645,178-238;
 * 1x6 row = columns [3:6|0:3] of the V3 row, 18-DOF StatesGroup). */
#define FLIVO_MODE_V3     0
#define FLIVO_MODE_V1     1
/* V1 keeps a point's five neighbours beyond the far-neighbour gate (they re-enter on re-use trips, src/laser_mapping.cpp:490,518)
 * up to this many search cells (cell = cell_ratio * downsample_size); farther than that a point counts as having fewer than
 * five neighbours.  The reference's search is unbounded: stated deviation, see DESIGN.md §4. */
#define FLIVO_V1_RANGE_CELLS 24

typedef struct flivo_ctx flivo_ctx;

/* Flat state, row-major rotations.  Superset of StatesGroup (include/common_lib.h:518-603) and
 * state_ikfom (include/use-ikfom.hpp:12-21). 36 doubles. */
typedef struct flivo_state {
    double rot[9];   /* state.rot_end            / s.rot            */
    double pos[3];   /* state.pos_end            / s.pos            */
    double offR[9];  /* Lidar_rot_to_IMU         / s.offset_R_L_I   */
    double offT[3];  /* Lidar_offset_to_IMU      / s.offset_T_L_I   */
    double vel[3];
    double bg[3];
    double ba[3];
    double grav[3];
} flivo_state;

typedef struct flivo_options {
    float    downsample_size;   /* KD_TREE box_length / set_downsample_param (ikd_Tree.h:167,171); default 0.2 */
    int32_t  cell_ratio;        /* search-grid cell edge = cell_ratio * downsample_size; 0 -> 2 */
    uint64_t initial_points;    /* sizing hint for the device tables; 0 -> 1<<20 */
    int32_t  sort_scan;         /* 1: process scan points in Morton order of their cells (results are
                                   reported in the caller's order either way); 0: caller's order */
    int32_t  ivox_nearby;       /* 0: ikd-Tree semantics (exact kNN, Add_Points voxel rule).  FLIVO_IVOX_*: the map
                                   behaves as faster_lio::IVox<3, DEFAULT> (include/ivox3d/ivox3d.h, the structure
                                   the reference builds when USE_ikdtree is off, CMakeLists.txt:15): downsample_size
                                   is Options::resolution_ (ivox_grid_resolution, src/laser_mapping.cpp:1021) and
                                   the value is the size of the nearby stencil (ivox_nearby_type, :1022-1035) */
} flivo_options;

/* IVox::NearbyType (include/ivox3d/ivox3d.h:46-51) as the number of voxels searched */
enum { FLIVO_IVOX_OFF = 0, FLIVO_IVOX_CENTER = 1, FLIVO_IVOX_NEARBY6 = 7, FLIVO_IVOX_NEARBY18 = 19, FLIVO_IVOX_NEARBY26 = 27 };

typedef struct flivo_cfg {
    int32_t mode;               /* FLIVO_MODE_V3 / FLIVO_MODE_V1 */
    int32_t max_iter;           /* NUM_MAX_ITERATIONS (src/laser_mapping.cpp:968); the loop runs i = -1 .. max_iter-1 */
    double  laser_point_cov;    /* LASER_POINT_COV (src/laser_mapping.cpp:975), the R of esekfom.hpp:1619 */
    double  limit[23];          /* V3 per-component convergence limit (epsi, src/origin_laserMapping.cpp:1231-1233) */
    int32_t force_search;       /* 1: kNN on every trip (SURVEY §3.2 "worst case"); 0: reference re-match policy */
    int32_t flags;              /* FLIVO_CFG_* */
} flivo_cfg;
/* per-trip CUDA events (flivo_stats.last_fused_ms / last_solve_ms) and plain launches instead of the captured graph */
#define FLIVO_CFG_PROFILE 1

typedef struct flivo_report {
    int32_t trips;              /* h_share_model calls executed */
    int32_t searches;           /* trips that ran the kNN */
    int32_t n_eff;              /* effct_feat_num of the last trip (src/laser_mapping.cpp:548-559) */
    int32_t converged;          /* V1 flg_EKF_converged / V3 t>0 at exit */
    double  res_mean;           /* res_mean_last of the last trip (:561) */
    double  dx_last[23];        /* last state increment (V1: first 18 entries, order rot,pos,vel,bg,ba,grav) */
    float   gpu_ms;             /* device time of the whole loop (CUDA events on the ctx stream) */
    int32_t kernel_launches;    /* kernels launched by this call */
} flivo_report;

/* ---- context ------------------------------------------------------------------------------- */
/* KD_TREE::KD_TREE / InitializeKDTree (include/ikd-Tree/ikd_Tree.cpp:9-16,40-44) */
int  flivo_ctx_create(int device, const flivo_options* opt, flivo_ctx** out);
void flivo_ctx_destroy(flivo_ctx* ctx);           /* KD_TREE::~KD_TREE (:18-25) */
const char* flivo_last_error(const flivo_ctx* ctx);
const char* flivo_version(void);
/* cudaStream_t of the ctx as an opaque pointer (for callers that time with their own CUDA events) */
void* flivo_ctx_stream(flivo_ctx* ctx);
int  flivo_sync(flivo_ctx* ctx);

/* ---- map: ikd-Tree API surface (include/ikd-Tree/ikd_Tree.h:166-188) ------------------------- */
/* KD_TREE::set_downsample_param (ikd_Tree.cpp:36-38). Rebuilds the voxel index when the map is non-empty.
 * iVox mode: sets filter_size_map_min of map_incremental's add rule (src/laser_mapping.cpp:353-375; default = the
 * voxel resolution); the voxel resolution itself is fixed at context creation. */
int flivo_map_set_downsample(flivo_ctx* ctx, float downsample_size);
/* KD_TREE::Build (ikd_Tree.cpp:337-348): replaces the map; no downsampling; ids = 0..n-1. */
int flivo_map_build(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes);
/* KD_TREE::Add_Points (ikd_Tree.cpp:382-457). Every offered point consumes one insertion id.
 * *n_added = the reference's return value (tmp_counter). */
int flivo_map_add_points(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes,
                         int downsample_on, int32_t* n_added);
/* KD_TREE::Delete_Point_Boxes (ikd_Tree.cpp:501-521, half-open boxes :648). boxes6 = nboxes x {min xyz, max xyz}. */
int flivo_map_delete_boxes(flivo_ctx* ctx, const float* boxes6, size_t nboxes, int32_t* n_deleted);
/* KD_TREE::Add_Point_Boxes (ikd_Tree.cpp:459-479): revives points removed by delete_boxes/delete_points
 * that have not yet been handed out by flivo_map_acquire_removed. */
int flivo_map_add_boxes(flivo_ctx* ctx, const float* boxes6, size_t nboxes, int32_t* n_revived);
/* KD_TREE::Delete_Points (ikd_Tree.cpp:481-499; same_point EPSS 1e-6 :1287-1289). */
int flivo_map_delete_points(flivo_ctx* ctx, const float* xyz, size_t n, size_t stride_bytes, int32_t* n_deleted);
/* KD_TREE::size / validnum (ikd_Tree.cpp:74-136): both report live points (the reference's size() also
 * counts lazily deleted nodes that a rebuild has not purged yet — tree-state dependent). */
int flivo_map_size(flivo_ctx* ctx, int64_t* n_live);
/* KD_TREE::tree_range (ikd_Tree.cpp:94-118): out6 = {min xyz, max xyz} over live points (zeros if empty). */
int flivo_map_range(flivo_ctx* ctx, float out6[6]);
/* KD_TREE::flatten(Root_Node, Storage, NOT_RECORD) (ikd_Tree.cpp:1238-1263): live points as x,y,z and id.
 * Call with cap = 0 to obtain *n only. Order is unspecified. */
int flivo_map_flatten(flivo_ctx* ctx, float* out_xyz, int32_t* out_id, size_t cap, size_t* n);
/* KD_TREE::acquire_removed_points (ikd_Tree.cpp:523-535): points removed by box/point deletes since the last call. */
int flivo_map_acquire_removed(flivo_ctx* ctx, float* out_xyz, size_t cap, size_t* n);

/* Batched KD_TREE::Nearest_Search (ikd_Tree.cpp:350-380; Search :843-986): exact k nearest live points
 * with squared distance <= max_dist, ascending by (float32 squared distance, insertion id).
 * k <= FLIVO_K.  Rows are padded with id -1 / +inf when fewer than k exist.  out_xyz/out_sqd/out_id may be NULL.
 * iVox mode: batched IVox::GetClosestPoint(pt, closest_pt, max_num = k, max_range) (include/ivox3d/ivox3d.h:132-204):
 * the k nearest among the points of the nearby voxels with squared distance < max_dist (= max_range * max_range,
 * strict), ascending by (distance, insertion id); the reference leaves the order after element 0 to std::nth_element. */
int flivo_knn(flivo_ctx* ctx, const float* q_xyz, size_t nq, size_t stride_bytes, int k, double max_dist,
              float* out_xyz /* nq*k*3 */, float* out_sqd /* nq*k */, int32_t* out_id /* nq*k */,
              int32_t* out_count /* nq */);

/* flivo_knn without any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same search,
 * same arguments and results as flivo_knn (the search statistics of flivo_get_stats are not updated), but the staged
 * queries and result rows live in one static device array of the library (about 3.8 MB, held by every GPU that loads
 * the library until the process exits) and the call performs no cudaMalloc / cudaMallocHost / cudaFree.
 * At most FLIVO_KNN_NOALLOC_MAX_QUERIES queries per call: more -> FLIVO_E_CAPACITY, nothing launched.  All calls on a
 * device share the array (it keeps the last call's contents; every row a call reads back is written by that call first):
 * calls that overlap on one GPU race on it, so serialise them. */
#define FLIVO_KNN_NOALLOC_MAX_QUERIES 32768
int flivo_knn_noalloc(flivo_ctx* ctx, const float* q_xyz, size_t nq, size_t stride_bytes, int k, double max_dist,
                      float* out_xyz /* nq*k*3 */, float* out_sqd /* nq*k */, int32_t* out_id /* nq*k */,
                      int32_t* out_count /* nq */);

/* ---- scan + measurement update ---------------------------------------------------------------- */
/* feats_down_body (src/laser_mapping.cpp:129-130): lidar-frame scan; copied to the device. */
int flivo_scan_upload(flivo_ctx* ctx, const float* body_xyz, size_t n, size_t stride_bytes);
/* Double-buffered form (the reference hands one feats_down_body over per frame, src/laser_mapping.cpp:129-130,165-167):
 * flivo_scan_upload_async copies the NEXT scan into the context's back buffer on a copy stream and returns without
 * waiting — the current scan and an update in flight are not touched; the caller's buffer must stay valid until the
 * following flivo_scan_commit (or any later synchronous call) returns.  A source in pinned / registered host memory
 * travels as it is, whole records, by one DMA (the xyz gather happens on the device); a pageable source is first
 * repacked by a few host threads.  flivo_scan_commit makes the uploaded scan the current one.
 * flivo_scan_upload == flivo_scan_upload_async + flivo_scan_commit. */
int flivo_scan_upload_async(flivo_ctx* ctx, const float* body_xyz, size_t n, size_t stride_bytes);
int flivo_scan_commit(flivo_ctx* ctx);
/* cudaHostRegister / cudaHostUnregister for callers without the CUDA headers (e.g. the PointVector the node reuses). */
int flivo_host_register(void* ptr, size_t bytes);
int flivo_host_unregister(void* ptr);
/* Same, from a device pointer to n float4 (x,y,z,ignored) already resident in HBM (no copy through the host). */
int flivo_scan_set_device(flivo_ctx* ctx, const void* dev_xyzw, size_t n);

/* ---- front end of the path: the two steps right before it, device resident --------- */
/* msg/Pose6D.msg as filled by set_pose6d (include/common_lib.h:618-633): pose of the IMU at each IMU sample of the scan */
typedef struct flivo_pose6d {
    double offset_time;   /* seconds since the first lidar point of the scan */
    double acc[3];        /* total acceleration, global frame */
    double gyr[3];        /* unbiased angular velocity, body frame */
    double vel[3];
    double pos[3];
    double rot[9];        /* row-major */
} flivo_pose6d;
/* lidar_meas.lidar->points (the raw, time-stamped scan): xyz at byte 0 of every record; time_offset_bytes = byte offset
 * of the per-point time in milliseconds (pcl::PointXYZINormal::curvature = 36), < 0: none; intensity_offset_bytes
 * (PointXYZINormal::intensity = 32) or < 0.  Copied to the device once. */
int flivo_raw_upload(flivo_ctx* ctx, const void* pts, size_t n, size_t stride_bytes, int time_offset_bytes, int intensity_offset_bytes);
/* Same from device memory: n float4 (x,y,z,intensity) and n float times in ms (may be NULL). */
int flivo_raw_set_device(flivo_ctx* ctx, const void* dev_xyzi, const void* dev_time_ms, size_t n);
/* The per-point back-propagation loop of ImuProcess::UndistortPcl (src/IMU_Processing.cpp:336-376): every point is moved
 * to the scan-end frame with the last IMU pose before its time stamp; in place on the raw scan (-> feats_undistort).
 * poses: IMUpose (offset_time ascending, first = 0); rot_end / pos_end: the propagated state at the scan end;
 * R_LI / t_LI: Lid_rot_to_IMU / Lid_offset_to_IMU.  Points must be ordered by time like the reference's. */
int flivo_raw_undistort(flivo_ctx* ctx, const flivo_pose6d* poses, int n_poses, const double rot_end[9], const double pos_end[3],
                        const double R_LI[9], const double t_LI[3]);
/* downSizeFilterSurf.filter(*feats_down_body) (pcl::VoxelGrid, src/laser_mapping.cpp:129-130, leaf size :1111): centroid
 * (x, y, z, intensity, time) of the points of every occupied leaf, ordered by leaf index.  The result becomes the
 * current scan as if passed to flivo_scan_upload; *n_out = its size. */
int flivo_raw_downsample(flivo_ctx* ctx, float leaf_size, size_t* n_out);
/* Read back n x (x, y, z, intensity): the raw scan (feats_undistort) / the current scan (feats_down_body), caller's order. */
int flivo_raw_download(flivo_ctx* ctx, float* out_xyzi, size_t cap, size_t* n_out);
int flivo_scan_download(flivo_ctx* ctx, float* out_xyzi, size_t cap, size_t* n_out);

/* One LaserMapping::h_share_model call (src/laser_mapping.cpp:485-645 / src/origin_laserMapping.cpp:916-1048)
 * at state `st`: world transform, kNN when do_search (else the cached Nearest_Points), esti_plane, gates,
 * 1x12 rows, H^T H (12x12 row-major) and H^T z.  Optional outputs may be NULL:
 *   selected n  : point_selected_surf after this call;  eff n : rows that entered H (:552)
 *   nbr_id n*5  : neighbour ids (ascending distance; -1 padded); nbr_sqd n*5
 *   plane n*4   : normvec (nx,ny,nz,pd2), zeros for unselected points;  world n*3 : feats_down_world */
int flivo_h_share_model(flivo_ctx* ctx, const flivo_state* st, int mode, int do_search,
                        double HTH[144], double HTz[12], int32_t* n_eff, double* res_sum,
                        uint8_t* selected, uint8_t* eff, int32_t* nbr_id, float* nbr_sqd,
                        float* plane, float* world);
/* V1 repack of the same sums: HPH 9x9 row-major, HPL 9 (src/laser_mapping.cpp:564-593). */
int flivo_repack_v1(const double HTH[144], const double HTz[12], double laser_point_cov,
                    double HPH[81], double HPL[9]);

/* The whole iterated update on the device, no host round trip between trips:
 * V3: esekf::update_iterated_dyn_share_modified (include/IKFoM_toolkit/esekfom/esekfom.hpp:1619-1928), P 23x23;
 * V1: the loop in LaserMapping::Run (src/laser_mapping.cpp:165-242), P = state.cov 18x18.  Row-major, in/out. */
int flivo_iekf_update(flivo_ctx* ctx, flivo_state* inout, double* P, const flivo_cfg* cfg, flivo_report* rep);
/* The same update in two halves: _begin launches it and returns at once, _end waits and hands back state, P and the
 * report.  Between the two the caller may stage the NEXT frame's scan (flivo_scan_upload_async), which then travels to
 * the device while this update runs; no other call on the ctx is allowed in between. */
int flivo_iekf_update_begin(flivo_ctx* ctx, const flivo_state* x, const double* P, const flivo_cfg* cfg);
int flivo_iekf_update_end(flivo_ctx* ctx, flivo_state* x_out, double* P_out, flivo_report* rep);

/* One frame of step 5 of LaserMapping::Run (src/laser_mapping.cpp:165-251) in ONE call: flivo_scan_commit (the scan staged
 * earlier becomes current), flivo_iekf_update_begin, flivo_scan_upload_async of the NEXT frame's scan (next_body_xyz may be
 * NULL), flivo_iekf_update_end, flivo_map_add_scan with the posterior.  x and P are in/out as in flivo_iekf_update. */
int flivo_frame(flivo_ctx* ctx, flivo_state* x, double* P, const flivo_cfg* cfg, const float* next_body_xyz, size_t n_next,
                size_t stride_bytes, flivo_report* rep, int32_t* n_added);

/* The GNSS position observation of the compiled (V1) path: `if (bgnss_en && p_gnss->new_gnss_) { computeH(...);
 * HPH.block<6,6>(0,0) += gnss_HPH; HPL.block<6,1>(0,0) += gnss_HPL; }` (src/laser_mapping.cpp:612-617) with
 * GNSSProcessing::computeH (src/GNSS_Processing.cpp:204-237) evaluated on the device at every trip's iterate.
 * pos_global = rot_enu2global_^T * gnss_.blh + delta_pos_ (:206-207, host side), antlever = antlever_, std3 = gnss_.std.
 * Applies to every following FLIVO_MODE_V1 update until cleared with pos_global = NULL (new_gnss_ = false). */
int flivo_set_gnss(flivo_ctx* ctx, const double pos_global[3], const double antlever[3], const double std3[3]);

/* LaserMapping::map_incremental (src/laser_mapping.cpp:329-389, USE_ikdtree branch :384): world-transform the
 * uploaded scan with `st` on the device and Add_Points(..., true) without a host round trip.
 * iVox mode: the other branch of the same function (:340-387): per point, from the neighbours cached by the last
 * search of this scan, add / add without down-sampling / drop; then AddPoints(points_to_add) and
 * AddPoints(point_no_need_downsample).  A scan that was never searched is added whole (Nearest_Points empty). */
int flivo_map_add_scan(flivo_ctx* ctx, const flivo_state* st, int32_t* n_added);

/* ---- multi-GPU (BASELINE configs 4-5; no counterpart in the reference) -------------------------- */
/* Shard a mega-scan over ranks: every rank holds the full map and a slice of the scan; each trip the
 * 92-double normal-equation block is combined across ranks in fixed rank order (bit-identical on every
 * rank) before the replicated on-device solve.  unique_id: 128 bytes from flivo_comm_unique_id on rank 0. */
int flivo_comm_unique_id(void* out128);
int flivo_comm_init(flivo_ctx* ctx, const void* unique_id128, int rank, int nranks);
int flivo_comm_destroy(flivo_ctx* ctx);
/* Peer-memory exchange over NVLink / NVSwitch for the sharded update (one process per GPU, one node, <= 8 ranks):
 * replaces the NCCL all-gather by stores into the peers' memory from inside the solve kernel, so that reduce, exchange and
 * solve are one kernel and the update one CUDA graph.  Every rank: flivo_peer_export -> the caller gathers the 64-byte
 * handles of all ranks in rank order (any transport, e.g. torch.distributed.all_gather_object) -> flivo_peer_attach. */
#define FLIVO_PEER_HANDLE_BYTES 64
int flivo_peer_export(flivo_ctx* ctx, void* handle_out64);
int flivo_peer_attach(flivo_ctx* ctx, const void* handles /* nranks * 64 bytes */, int rank, int nranks);
int flivo_peer_detach(flivo_ctx* ctx);
/* The same exchange between contexts of ONE process (one process driving several GPUs, or several ranks on one GPU as the
 * tests do): cudaIpc handles cannot be opened by the exporting process, so the ranks pass their buffers as device
 * pointers: flivo_peer_local_base on every context -> flivo_peer_attach_local(bases in rank order).  Across devices the
 * caller enables peer access. */
int flivo_peer_local_base(flivo_ctx* ctx, void** base_out);
int flivo_peer_attach_local(flivo_ctx* ctx, void* const* bases /* nranks */, int rank, int nranks);
/* A rank waits `seconds` (default 2) for a peer's record, then gives the update up: FLIVO_E_STATE, state and P untouched,
 * and every later sharded update is refused until flivo_peer_reset.  flivo_peer_reset zeroes this rank's exchange counter
 * and buffer; every rank calls it, between two barriers of the caller's (no update in flight, none started early). */
int flivo_peer_set_timeout(flivo_ctx* ctx, double seconds);
int flivo_peer_reset(flivo_ctx* ctx);

/* ---- introspection used by bench.py ------------------------------------------------------------ */
typedef struct flivo_stats {
    int64_t map_live;           /* live points */
    int64_t map_cells;          /* occupied search cells */
    int64_t map_blocks;         /* occupied 4x4x4-cell blocks */
    int64_t dir_capacity;       /* block-directory slots */
    int64_t bucket_capacity;    /* point buckets allocated */
    int64_t last_far_queries;   /* queries of the last search trip that left the 27-cell fast path */
    int64_t last_candidates;    /* candidate points examined by the last search trip */
    float   last_fused_ms[16];  /* per-trip device time of the fused kernel in the last flivo_iekf_update */
    float   last_solve_ms[16];
    int32_t last_trip_searched[16];
} flivo_stats;
int flivo_get_stats(flivo_ctx* ctx, flivo_stats* out);
/* Diagnostic only (no reference counterpart): device globaltimer stamps (ns) at the phase boundaries of the solve
 * kernel, [trips_cap][16], of the last flivo_iekf_update that ran with FLIVO_CFG_PROFILE. */
int flivo_debug_solve_stamps(flivo_ctx* ctx, uint64_t* out, int trips_cap);
/* Same for the search kernel: per CTA {start, end of the tile phase, end, far queries finished} of the first search trip. */
int flivo_debug_search_stamps(flivo_ctx* ctx, uint64_t* out, int ctas_cap, int* n_ctas);
/* Diagnostic only: the processing order of the current scan (stages it if that has not happened yet):
 * out[q] = index in the caller's order of the point processed q-th; identity for an unsorted scan.  The order is the
 * stable sort of the scan by bits [6, 22) of the Morton code of each point's search cell in the lidar frame
 * (csrc/stage.cuh); the reference walks feats_down_body in PCL's voxel-grid order (src/laser_mapping.cpp:499-503). */
int flivo_debug_scan_order(flivo_ctx* ctx, uint32_t* out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif /* FLIVO_H_ */
