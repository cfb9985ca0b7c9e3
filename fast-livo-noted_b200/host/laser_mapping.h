// laser_mapping.h — ROS-free drop-in for the LIO half of the reference's LaserMapping
// (include/laser_mapping.h:62-241, src/laser_mapping.cpp).  Class, member and method names follow the
// reference so that its call sites read the same; what is kept is exactly the hot path:
//   Run()             step 5 of LaserMapping::Run (src/laser_mapping.cpp:129-251): first scan -> ikdtree.Build (from
//                     feats_down_body as it is, > 5 points, as the reference does),
//                     otherwise the iterated update (on the device, one call) and map_incremental()
//   h_share_model()   src/laser_mapping.cpp:485-645: one trip; returns the 9x9 HPH / 9 HPL of the compiled V1 path
//   map_incremental() src/laser_mapping.cpp:329-389, ikd-Tree branch (:384)
//   downsample_on_device()   downSizeFilterSurf.filter (:129-130) on the raw scan already in HBM (after
//                            ImuProcess::UndistortPcl of host/imu_processing.h, or flivo_raw_upload)
//   lasermap_fov_segment()   the sliding local map of the upstream-style main (src/origin_laserMapping.cpp:367-429):
//                            which boxes leave the cube around the lidar, ikdtree.Delete_Point_Boxes, removed-point cache
// Everything ROS (InitROS, callbacks, publishers), IMU propagation, VIO, GNSS and loop closure stay with the
// caller: it fills `state` (the propagated prior) and `feats_down_body` (the down-sampled, undistorted scan in the
// lidar frame) before Run(), exactly the two inputs the reference's step 5 reads.
// Header-only; link with -lflivo_b200.
#pragma once
#include <string.h>
#include <stdexcept>
#include <string>
#include <vector>

#include "ikd_Tree.h"

struct LaserMappingOptions {          // the rosparams the path reads (src/laser_mapping.cpp:945-1116)
    int    max_iteration = 4;         // NUM_MAX_ITERATIONS (:968)
    double laser_point_cov = 0.001;   // LASER_POINT_COV (:975)
    double filter_size_map = 0.2;     // filter_size_map_min (:983) -> KD_TREE downsample size
    double filter_size_surf = 0.5;    // filter_size_surf_min (:982) -> leaf of downSizeFilterSurf (:1111)
    double cube_side_length = 200;    // cube_len (:984)
    float  det_range = 300.0f;        // DET_RANGE (src/origin_laserMapping.cpp:89)
    double extrinsic_T[3] = {0, 0, 0};                   // Lidar_offset_to_IMU
    double extrinsic_R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; // Lidar_rot_to_IMU
    int    device = 0;
};

class LaserMapping {
public:
    explicit LaserMapping(const LaserMappingOptions& opt = LaserMappingOptions())
        : NUM_MAX_ITERATIONS(opt.max_iteration), LASER_POINT_COV(opt.laser_point_cov), filter_size_map_min(opt.filter_size_map),
          filter_size_surf_min(opt.filter_size_surf), cube_len(opt.cube_side_length), DET_RANGE(opt.det_range),
          ikdtree(0.5f, 0.6f, (float)opt.filter_size_map, opt.device) {
        memcpy(Lidar_offset_to_IMU, opt.extrinsic_T, sizeof(Lidar_offset_to_IMU));
        memcpy(Lidar_rot_to_IMU, opt.extrinsic_R, sizeof(Lidar_rot_to_IMU));
    }

    // One LiDAR frame through step 5 (src/laser_mapping.cpp:129-251).  Returns false when the frame only built the map.
    bool Run() {
        if (!scan_on_device_) {
            feats_down_size = (int)feats_down_body.size();
            if (feats_down_size == 0) return false;
            upload_scan();
        }
        if (feats_down_size == 0) return false;
        scan_on_device_ = false;   // consumed by this frame
        state_propagat = state;                                                      // :59
        if (ikdtree.Root_Node == nullptr) {
            // :134-142: the first usable scan becomes the map exactly as the reference builds it — from feats_down_body,
            // i.e. in the LIDAR frame (the node starts at the identity pose and ignores the extrinsic here), more than 5
            // points required, set_downsample_param first — and the frame ends without an update.
            if (feats_down_size > 5) {
                if ((int)feats_down_body.size() != feats_down_size) fetch_scan_to_host();   // scan came from downsample_on_device
                ikdtree.set_downsample_param((float)filter_size_map_min);
                ikdtree.Build(feats_down_body);
            }
            return false;
        }
        // :165-242, whole loop on the device (V1: 18-DOF StatesGroup, 1x6 rows, rematch policy)
        flivo_state st = to_flivo(state);
        flivo_cfg cfg{};
        cfg.mode = FLIVO_MODE_V1;
        cfg.max_iter = NUM_MAX_ITERATIONS;
        cfg.laser_point_cov = LASER_POINT_COV;
        for (double& l : cfg.limit) l = 0.001;
        flivo_report rep{};
        ck(flivo_iekf_update(ikdtree.ctx(), &st, state.cov, &cfg, &rep), "Run/update");
        from_flivo(st, state);
        effct_feat_num = rep.n_eff;
        res_mean_last = rep.res_mean;
        flg_EKF_converged = rep.converged != 0;
        last_report = rep;
        map_incremental();                                                           // :251
        return true;
    }

    // src/laser_mapping.cpp:485-645: one trip at the current `state`; nearest_search_en selects search vs cached
    // Nearest_Points.  HPH is 9x9 row-major, HPL has 9 entries (only the leading 6x6 / 6 are populated, :564-593).
    void h_share_model(std::vector<double>& HPH, std::vector<double>& HPL) {
        if (!scan_uploaded_) upload_scan();
        flivo_state st = to_flivo(state);
        double HTH[144], HTz[12], res_sum = 0;
        int32_t ne = 0;
        ck(flivo_h_share_model(ikdtree.ctx(), &st, FLIVO_MODE_V1, nearest_search_en ? 1 : 0, HTH, HTz, &ne, &res_sum, nullptr, nullptr,
                               nullptr, nullptr, nullptr, nullptr), "h_share_model");
        HPH.assign(81, 0.0);
        HPL.assign(9, 0.0);
        flivo_repack_v1(HTH, HTz, LASER_POINT_COV, HPH.data(), HPL.data());
        effct_feat_num = ne;
        res_mean_last = ne > 0 ? res_sum / ne : 0.0;
    }

    // src/laser_mapping.cpp:329-389 (ikd-Tree branch): world-transform with the FINAL state, Add_Points(..., true)
    void map_incremental() {
        if (!scan_uploaded_) upload_scan();
        flivo_state st = to_flivo(state);
        int32_t added = 0;
        ck(flivo_map_add_scan(ikdtree.ctx(), &st, &added), "map_incremental");
        kdtree_incremental_added = added;
        if (added > 0 && ikdtree.Root_Node == nullptr) { PointVector none; ikdtree.Add_Points(none, true); }
    }

    // downSizeFilterSurf.setInputCloud(feats_undistort); downSizeFilterSurf.filter(*feats_down_body) (:129-130) on the raw
    // scan resident in HBM; the result is the scan of the next Run().  want_host_copy fills feats_down_body as well.
    int downsample_on_device(bool want_host_copy = false) {
        size_t n = 0;
        ck(flivo_raw_downsample(ikdtree.ctx(), (float)filter_size_surf_min, &n), "downsample");
        feats_down_size = (int)n;
        scan_uploaded_ = true;
        scan_on_device_ = true;
        feats_down_body.clear();
        if (want_host_copy && n) fetch_scan_to_host();
        return (int)n;
    }
    // feats_down_body <- the current device-resident scan (x, y, z, intensity), caller's order
    void fetch_scan_to_host() {
        size_t n = (size_t)feats_down_size;
        std::vector<float> buf(n * 4);
        ck(flivo_scan_download(ikdtree.ctx(), buf.data(), n, &n), "scan download");
        feats_down_body.assign(n, PointType());
        for (size_t i = 0; i < n; ++i) {
            feats_down_body[i].x = buf[4 * i]; feats_down_body[i].y = buf[4 * i + 1]; feats_down_body[i].z = buf[4 * i + 2];
            feats_down_body[i].intensity = buf[4 * i + 3];
        }
    }

    // The sliding local map of the upstream-style main (behaviour of src/origin_laserMapping.cpp:367-429).  pos_LiD: position
    // of the lidar in the world (state.pos_end in the non-IKFOM branch, :381).  The map is kept inside an axis-aligned cube
    // of side cube_len; when the lidar comes within MOV_THRESHOLD * DET_RANGE of a face the cube slides away from that face
    // by slide_step() and the slab of the old cube that is no longer covered is deleted from the map, one slab per axis
    // that moved (each slab spans the OLD cube on the other two axes, as the reference's boxes do).
    static float MOV_THRESHOLD() { return 1.5f; }                                    // :96
    // -1: slide toward the lower face, +1: toward the upper face, 0: stay (the lower face wins when both are near)
    int slide_direction(int axis, const double pos_LiD[3]) const {
        const float reach = MOV_THRESHOLD() * DET_RANGE;
        const float to_lower = fabs(pos_LiD[axis] - LocalMap_Points.vertex_min[axis]);
        const float to_upper = fabs(pos_LiD[axis] - LocalMap_Points.vertex_max[axis]);
        return to_lower <= reach ? -1 : (to_upper <= reach ? +1 : 0);
    }
    // 90 % of the room the cube has beyond the detection reach, but at least DET_RANGE * (MOV_THRESHOLD - 1)   (:403)
    float slide_step() const {
        const double room = (cube_len - 2.0 * MOV_THRESHOLD() * DET_RANGE) * 0.5 * 0.9;
        const double least = double(DET_RANGE * (MOV_THRESHOLD() - 1));
        return (float)(room > least ? room : least);
    }
    void lasermap_fov_segment(const double pos_LiD[3]) {
        cub_needrm.clear();
        kdtree_delete_counter = 0;
        if (!Localmap_Initialized) {                       // first call: centre the cube on the lidar, nothing to delete
            for (int axis = 0; axis < 3; ++axis) {
                LocalMap_Points.vertex_min[axis] = pos_LiD[axis] - cube_len / 2.0;
                LocalMap_Points.vertex_max[axis] = pos_LiD[axis] + cube_len / 2.0;
            }
            Localmap_Initialized = true;
            return;
        }
        int dir[3];
        for (int axis = 0; axis < 3; ++axis) dir[axis] = slide_direction(axis, pos_LiD);
        if (!dir[0] && !dir[1] && !dir[2]) return;
        const float step = slide_step();
        const BoxPointType before = LocalMap_Points;
        for (int axis = 0; axis < 3; ++axis) {
            if (!dir[axis]) continue;
            BoxPointType slab = before;                    // what the cube leaves behind along this axis
            if (dir[axis] < 0) slab.vertex_min[axis] = before.vertex_max[axis] - step;
            else slab.vertex_max[axis] = before.vertex_min[axis] + step;
            cub_needrm.push_back(slab);
            if (dir[axis] < 0) { LocalMap_Points.vertex_min[axis] -= step; LocalMap_Points.vertex_max[axis] -= step; }
            else { LocalMap_Points.vertex_min[axis] += step; LocalMap_Points.vertex_max[axis] += step; }
        }
        points_cache_collect();                            // (:422) hand out what earlier deletes removed
        kdtree_delete_counter = ikdtree.Delete_Point_Boxes(cub_needrm);              // (:424)
    }
    // src/origin_laserMapping.cpp:332-337
    void points_cache_collect() {
        PointVector points_history;
        ikdtree.acquire_removed_points(points_history);
        points_cache_size = (int)points_history.size();
    }

    void Finish() {}

    // ---- data members the reference's step 5 reads / writes (same names) ------------------------------------
    StatesGroup state, state_propagat;
    PointVector feats_down_body;           // input of Run(): down-sampled scan, lidar frame
    int feats_down_size = 0;
    double Lidar_offset_to_IMU[3];
    double Lidar_rot_to_IMU[9];
    int NUM_MAX_ITERATIONS;
    double LASER_POINT_COV;
    double filter_size_map_min;
    double filter_size_surf_min;
    double cube_len;
    float DET_RANGE;
    BoxPointType LocalMap_Points{};
    bool Localmap_Initialized = false;
    std::vector<BoxPointType> cub_needrm;
    int kdtree_delete_counter = 0;
    int points_cache_size = 0;
    bool nearest_search_en = true;
    bool flg_EKF_inited = true, flg_EKF_converged = false;
    int effct_feat_num = 0;
    double res_mean_last = 0.0;
    int kdtree_incremental_added = 0;
    flivo_report last_report{};
    KD_TREE ikdtree;

    // call after changing feats_down_body outside Run()
    void upload_scan() {
        if (feats_down_body.empty()) throw std::runtime_error("LaserMapping: empty feats_down_body");
        ck(flivo_scan_upload(ikdtree.ctx(), &feats_down_body[0].x, feats_down_body.size(), sizeof(PointType)), "scan upload");
        scan_uploaded_ = true;
    }

private:
    flivo_state to_flivo(const StatesGroup& s) const {
        flivo_state f{};
        memcpy(f.rot, s.rot_end, sizeof(f.rot)); memcpy(f.pos, s.pos_end, sizeof(f.pos));
        memcpy(f.offR, Lidar_rot_to_IMU, sizeof(f.offR)); memcpy(f.offT, Lidar_offset_to_IMU, sizeof(f.offT));
        memcpy(f.vel, s.vel_end, sizeof(f.vel)); memcpy(f.bg, s.bias_g, sizeof(f.bg)); memcpy(f.ba, s.bias_a, sizeof(f.ba));
        memcpy(f.grav, s.gravity, sizeof(f.grav));
        return f;
    }
    static void from_flivo(const flivo_state& f, StatesGroup& s) {
        memcpy(s.rot_end, f.rot, sizeof(f.rot)); memcpy(s.pos_end, f.pos, sizeof(f.pos));
        memcpy(s.vel_end, f.vel, sizeof(f.vel)); memcpy(s.bias_g, f.bg, sizeof(f.bg)); memcpy(s.bias_a, f.ba, sizeof(f.ba));
        memcpy(s.gravity, f.grav, sizeof(f.grav));
    }
    void ck(int rc, const char* what) {
        if (rc != FLIVO_OK) throw std::runtime_error(std::string("LaserMapping::") + what + ": " + flivo_last_error(ikdtree.ctx()));
    }
    bool scan_uploaded_ = false;
    bool scan_on_device_ = false;   // the current scan came from downsample_on_device(): feats_down_body may be empty
};
