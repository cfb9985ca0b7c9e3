// ============================================================================
// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// CPU restatement ("oracle") of the FAST-LIVO LIO measurement-update hot path.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load this; the product (fast-livo-noted_b200/) never includes, links or calls it.
//
// Parity status: the kNN / Add_Points / Delete_Point_Boxes half is PINNED against the
// reference's own ikd_Tree.cpp compiled unmodified into oracle/_ref (see oracle/Makefile,
// tests/test_oracle_vs_ikdtree.py).  The Eigen-dependent half (esti_plane's
// colPivHouseholderQr, the 23x23 / 18x18 inverses) is "parity unpinned": Eigen is not in
// the reference tree nor in this image, the reference has no tests or golden vectors
// (SURVEY.md §4), so the operation order below is a restatement of Eigen's published
// algorithm and DEFINES the expected bits.
//
// All citations are relative to /root/reference/.
// ============================================================================
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

// Flat state (row-major 3x3 rotations).  Superset of StatesGroup (include/common_lib.h:518-603)
// and state_ikfom (include/use-ikfom.hpp:12-21).
typedef struct orc_state {
    double rot[9];    // R   (state.rot_end / s.rot)
    double pos[3];    // t   (state.pos_end / s.pos)
    double offR[9];   // R_LI (Lidar_rot_to_IMU / s.offset_R_L_I)
    double offT[3];   // t_LI (Lidar_offset_to_IMU / s.offset_T_L_I)
    double vel[3];
    double bg[3];
    double ba[3];
    double grav[3];
} orc_state;

enum { ORC_MODE_V3 = 0, ORC_MODE_V1 = 1 };

// a1: pointBodyToWorld (src/laser_mapping.cpp:662-671; V3 src/origin_laserMapping.cpp:935-940)
void orc_body_to_world(const float* body_xyz, size_t n, size_t stride_floats,
                       const orc_state* st, float* world_xyz /* n*3 */);

// a2/a3: exact kNN with ikd-Tree Nearest_Search semantics (include/ikd-Tree/ikd_Tree.cpp:350-380,
// :843-986, calc_dist :1291-1295) by brute force under the canonical total order
// (dist_f32 ascending, then insertion id ascending).  alive==NULL means all alive.
void orc_knn_bruteforce(const float* map_xyz, const int32_t* map_id, const uint8_t* alive, size_t m,
                        const float* q_xyz, size_t nq, int k, double max_dist,
                        int32_t* out_id /* nq*k, -1 padded */, float* out_sqd /* nq*k */,
                        float* out_xyz /* nq*k*3 */, int32_t* out_count /* nq */,
                        int64_t* tie_straddles /* may be NULL: #queries whose rank-k and rank-k+1 distances tie */);

// a4: esti_plane<float> (include/common_lib.h:670-702).  near5: 5*3 floats.  Returns 1 if plane accepted.
int orc_esti_plane(const float* near5, float threshold, float pabcd[4]);

// a1..a8 for one h_share_model trip, neighbours given (already searched or cached).
//   near_xyz: n*5*3, near_cnt: n, near_d5: n (5th squared distance, valid when near_cnt==5)
//   searched: 1 if this trip performed the search (the far-neighbour gate :518 is then applied)
//   mask_io:  V3: in = point_selected_surf from the previous trip (ignored when searched), out = this trip's
//             V1: input ignored (vector re-initialised to true each call, src/laser_mapping.cpp:490)
//   eff_out:  point_selected_surf[i] && res_last[i] <= 2.0 (the rows that enter H, :552)
//   outputs: HTH 12x12 row-major, HTz 12, n_eff, res_sum; normvec n*4 (nx,ny,nz,pd2) (zeros when masked)
//            world_xyz n*3; borderline: count of points within 1e-5 of a gate threshold (may be NULL)
void orc_h_share_model(const float* body_xyz, size_t n, size_t stride_floats, const orc_state* st,
                       const float* near_xyz, const int32_t* near_cnt, const float* near_d5,
                       int searched, int mode, uint8_t* mask_io, uint8_t* eff_out, float* normvec, float* world_xyz,
                       double* HTH, double* HTz, int64_t* n_eff, double* res_sum, int64_t* borderline);

// V1 repack (src/laser_mapping.cpp:564-593): HPH 9x9 row-major, HPL 9 from the 12x12 block.
void orc_repack_v1(const double* HTH, const double* HTz, double laser_point_cov, double* HPH81, double* HPL9);

// a9 (V3): one trip of esekf::update_iterated_dyn_share_modified
// (include/IKFoM_toolkit/esekfom/esekfom.hpp:1619-1928).  P matrices 23x23 row-major.
//   x (in/out), x_prop, P_prop in; P_out written only when *stopped becomes 1.
//   iter is the reference loop index i (starts at -1); t_io is the converge counter t.
//   converge_out = dyn_share.converge for the next trip (1 => next trip searches).
void orc_solve_v3(orc_state* x, const orc_state* x_prop, const double* P_prop, const double* HTH,
                  const double* HTz, double R, int iter, int max_iter, const double* limit23,
                  int* t_io, int* converge_out, int* stopped, double* dx_out23, double* P_out);

// a9 (V1): one trip of the loop body in LaserMapping::Run (src/laser_mapping.cpp:178-238).
//   cov 18x18 row-major (in/out; updated when *stopped becomes 1).  HPH 9x9, HPL 9.
//   rematch_io: rematch_num; search_next: nearest_search_en for the next trip.
void orc_solve_v1(orc_state* x, const orc_state* x_prop, double* cov18, const double* HPH81,
                  const double* HPL9, int iterCount, int max_iter, int* rematch_io, int* search_next,
                  int* stopped, double* solution18);

// a12: KD_TREE::Add_Points voxel-downsample rule (include/ikd-Tree/ikd_Tree.cpp:382-457) restated
// over a flat point list.  Handle-based tiny map used by tests; ids are insertion ids.
typedef struct orc_map orc_map;
orc_map* orc_map_create(float downsample_size);
void orc_map_destroy(orc_map*);
void orc_map_build(orc_map*, const float* xyz, size_t n);                       // :337-348 (no downsample)
int  orc_map_add_points(orc_map*, const float* xyz, size_t n, int downsample_on); // returns tmp_counter
int  orc_map_delete_boxes(orc_map*, const float* boxes6, size_t nboxes);        // :501-521, :626-688 half-open
size_t orc_map_live(orc_map*, float* xyz /* cap*3 or NULL */, int32_t* ids /* or NULL */, size_t cap);

// ---- front end ----
// msg/Pose6D.msg (set_pose6d, include/common_lib.h:618-633)
typedef struct orc_pose6d {
    double offset_time, acc[3], gyr[3], vel[3], pos[3], rot[9];
} orc_pose6d;
// ImuProcess::UndistortPcl, backward loop (src/IMU_Processing.cpp:336-376), in place
void orc_undistort(float* pts, size_t n, size_t stride_floats, const float* t_ms, const orc_pose6d* IMUpose, int n_poses,
                   const double rot_end[9], const double pos_end[3], const double R_LI[9], const double t_LI[3]);
// pcl::VoxelGrid::filter (src/laser_mapping.cpp:129-130); returns the number of output points
long orc_voxel_grid(const float* xyzi, const float* t_ms, size_t n, float leaf, float* out_xyzi, float* out_t, size_t cap);

// ---- iVox branch (oracle/ivox_oracle.cpp) ----
// faster_lio::IVox<3, DEFAULT> (include/ivox3d/ivox3d.h): nearby = 1, 7, 19, 27 voxels (CENTER, NEARBY6/18/26);
// capacity 0 -> Options::capacity_ default (1 000 000 voxels)
typedef struct orc_ivox orc_ivox;
orc_ivox* orc_ivox_create(float resolution, int nearby, size_t capacity);
void orc_ivox_destroy(orc_ivox*);
void orc_ivox_add_points(orc_ivox*, const float* xyz, size_t n);                       // AddPoints (:247-276)
size_t orc_ivox_num_grids(orc_ivox*);                                                  // NumValidGrids (:206-209)
size_t orc_ivox_evicted(orc_ivox*);
size_t orc_ivox_points(orc_ivox*, float* xyz, int32_t* ids, size_t cap);
// batched GetClosestPoint(pt, closest_pt, k, max_range) (:132-204); out_count = -1: no candidate (list untouched)
void orc_ivox_knn(orc_ivox*, const float* q_xyz, size_t nq, int k, double max_range,
                  int32_t* out_id, float* out_sqd, float* out_xyz, int32_t* out_count, int64_t* tie_straddles);
// map_incremental's add rule (src/laser_mapping.cpp:340-381): 1 = points_to_add, 2 = point_no_need_downsample, 0 = dropped
void orc_ivox_map_filter(const float* world_xyz, size_t n, const float* near_xyz, const int32_t* near_cnt,
                         double filter_size_map_min, uint8_t* flag_out);

// GNSSProcessing::computeH (src/GNSS_Processing.cpp:204-237): the 6x6 / 6 block added to HPH / HPL (src/laser_mapping.cpp:612-617)
void orc_gnss_block(const double* rot, const double* pos, const double* gnss_pos, const double* antlever,
                    const double* std3, double* HTH36, double* HTL6);

#ifdef __cplusplus
}
#endif
