// laser_mapping.h — ROS-free drop-in for the LIO half of the reference's LaserMapping
// (include/laser_mapping.h:62-241, src/laser_mapping.cpp).  Class, member and method names follow the
// reference so that its call sites read the same; what is kept is exactly the hot path:
//   Run()             step 5 of LaserMapping::Run (src/laser_mapping.cpp:129-251): first scan -> ikdtree.Build,
//                     otherwise the iterated update (on the device, one call) and map_incremental()
//   h_share_model()   src/laser_mapping.cpp:485-645: one trip; returns the 9x9 HPH / 9 HPL of the compiled V1 path
//   map_incremental() src/laser_mapping.cpp:329-389, ikd-Tree branch (:384)
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
    double extrinsic_T[3] = {0, 0, 0};                   // Lidar_offset_to_IMU
    double extrinsic_R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; // Lidar_rot_to_IMU
    int    device = 0;
};

class LaserMapping {
public:
    explicit LaserMapping(const LaserMappingOptions& opt = LaserMappingOptions())
        : NUM_MAX_ITERATIONS(opt.max_iteration), LASER_POINT_COV(opt.laser_point_cov), filter_size_map_min(opt.filter_size_map),
          ikdtree(0.5f, 0.6f, (float)opt.filter_size_map, opt.device) {
        memcpy(Lidar_offset_to_IMU, opt.extrinsic_T, sizeof(Lidar_offset_to_IMU));
        memcpy(Lidar_rot_to_IMU, opt.extrinsic_R, sizeof(Lidar_rot_to_IMU));
    }

    // One LiDAR frame through step 5 (src/laser_mapping.cpp:129-251).  Returns false when the frame only built the map.
    bool Run() {
        feats_down_size = (int)feats_down_body.size();
        if (feats_down_size == 0) return false;
        state_propagat = state;                                                      // :59
        upload_scan();
        if (ikdtree.Root_Node == nullptr) {                                          // :134-142 first scan: world points, Build
            flivo_state st = to_flivo(state);
            std::vector<float> w((size_t)feats_down_size * 3);
            double HTH[144], HTz[12];
            // an empty map yields no neighbours; the trip is used only for its world-transformed points
            ck(flivo_h_share_model(ikdtree.ctx(), &st, FLIVO_MODE_V1, 1, HTH, HTz, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr,
                                   nullptr, w.data()), "Run/first scan");
            PointVector world(feats_down_size);
            for (int i = 0; i < feats_down_size; ++i) { world[i] = feats_down_body[i]; world[i].x = w[3 * i]; world[i].y = w[3 * i + 1]; world[i].z = w[3 * i + 2]; }
            ikdtree.Build(world);
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
};
