// C++ caller of the drop-in classes, written the way the reference's own call sites use them
// (src/laser_mapping.cpp:134-143,165-251): KD_TREE Build / Nearest_Search / Add_Points / Delete_Point_Boxes and
// LaserMapping::Run over a few frames of a synthetic scene.  Reads the scene from a flat binary file written by
// tests/test_dropin_cpp.py and prints results as "key value..." lines for the Python side to compare with the oracle.
//   test_dropin <scene.bin>         run on cuda:0
//   test_dropin --no-gpu            only check that construction fails loudly without a device
//   test_dropin --frontend <in.bin> <out.bin>   ImuProcess::UndistortPcl -> LaserMapping::downsample_on_device -> Run,
//                                   then a walk through lasermap_fov_segment (tests/test_frontend_cpp.py)
#include <stdio.h>
#include <string.h>
#include <stdexcept>
#include <vector>

#include "imu_processing.h"
#include "ivox3d.h"
#include "laser_mapping.h"

static std::vector<float> read_f32(FILE* f, size_t n) {
    std::vector<float> v(n);
    if (fread(v.data(), 4, n, f) != n) throw std::runtime_error("short read");
    return v;
}
static std::vector<double> read_f64(FILE* f, size_t n) {
    std::vector<double> v(n);
    if (fread(v.data(), 8, n, f) != n) throw std::runtime_error("short read");
    return v;
}

// front end through the drop-in classes, the way the upstream-style main chains them (src/origin_laserMapping.cpp:1290-1376)
static int run_frontend(const char* in_path, const char* out_path) {
    FILE* f = fopen(in_path, "rb");
    if (!f) { perror("open"); return 2; }
    int32_t hdr[4];   // n_raw, n_imu, n_map, n_walk
    if (fread(hdr, 4, 4, f) != 4) return 2;
    std::vector<double> par = read_f64(f, 5);   // ds_map, leaf, cube_len, det_range, lidar_beg_time
    std::vector<double> x = read_f64(f, 36);
    std::vector<double> imu = read_f64(f, (size_t)(hdr[1] + 1) * 7);   // last_imu_ first, then the frame's samples
    std::vector<float> raw = read_f32(f, (size_t)hdr[0] * 5);
    std::vector<float> map = read_f32(f, (size_t)hdr[2] * 3);
    std::vector<double> walk = read_f64(f, (size_t)hdr[3] * 3);
    fclose(f);
    FILE* o = fopen(out_path, "wb");
    if (!o) { perror("open out"); return 2; }
    try {
        LaserMappingOptions opt;
        opt.filter_size_map = par[0]; opt.filter_size_surf = par[1]; opt.cube_side_length = par[2]; opt.det_range = (float)par[3];
        memcpy(opt.extrinsic_R, &x[12], 72); memcpy(opt.extrinsic_T, &x[21], 24);
        LaserMapping lm(opt);
        ImuProcess p_imu;
        p_imu.set_extrinsic(opt.extrinsic_T, opt.extrinsic_R);
        auto to_msg = [&](int k) { ImuMsg m; m.stamp = imu[7 * k]; memcpy(m.angular_velocity, &imu[7 * k + 1], 24); memcpy(m.linear_acceleration, &imu[7 * k + 4], 24); return m; };
        p_imu.last_imu_ = to_msg(0);
        p_imu.last_lidar_end_time_ = par[4];
        memcpy(p_imu.mean_acc, &x[33], 24);   // |gravity|-sized mean acceleration: scale factor G_m_s2/|mean_acc| ~ 1
        LidarMeasureGroup meas;
        meas.lidar_beg_time = par[4];
        for (int k = 1; k <= hdr[1]; ++k) meas.imu.push_back(to_msg(k));
        meas.lidar.resize(hdr[0]);
        for (int i = 0; i < hdr[0]; ++i) {
            meas.lidar[i].x = raw[5 * i]; meas.lidar[i].y = raw[5 * i + 1]; meas.lidar[i].z = raw[5 * i + 2];
            meas.lidar[i].intensity = raw[5 * i + 3]; meas.lidar[i].curvature = raw[5 * i + 4];
        }
        memcpy(lm.state.rot_end, &x[0], 72); memcpy(lm.state.pos_end, &x[9], 24);
        memcpy(lm.state.vel_end, &x[24], 24); memcpy(lm.state.bias_g, &x[27], 24); memcpy(lm.state.bias_a, &x[30], 24);
        memcpy(lm.state.gravity, &x[33], 24);
        p_imu.UndistortPcl(lm.ikdtree.ctx(), meas, lm.state);
        // out: n_poses, poses, state (rot9 pos3 vel3), cov 324, undistorted n*4, n_down, down n*4
        int32_t np = (int32_t)p_imu.IMUpose.size();
        fwrite(&np, 4, 1, o);
        fwrite(p_imu.IMUpose.data(), sizeof(flivo_pose6d), np, o);
        fwrite(lm.state.rot_end, 8, 9, o); fwrite(lm.state.pos_end, 8, 3, o); fwrite(lm.state.vel_end, 8, 3, o);
        fwrite(lm.state.cov, 8, 324, o);
        PointVector und;
        p_imu.download_undistorted(lm.ikdtree.ctx(), und, meas);
        for (auto& q : und) { float v[4] = {q.x, q.y, q.z, q.intensity}; fwrite(v, 4, 4, o); }
        int32_t nd = lm.downsample_on_device(true);
        fwrite(&nd, 4, 1, o);
        for (auto& q : lm.feats_down_body) { float v[4] = {q.x, q.y, q.z, q.intensity}; fwrite(v, 4, 4, o); }
        printf("frontend raw %d poses %d down %d\n", hdr[0], np, nd);
        // first frame builds the map from the down-sampled scan (device-resident scan, no host copy needed)
        bool updated = lm.Run();
        printf("run updated %d size %d root %d\n", (int)updated, lm.ikdtree.size(), lm.ikdtree.Root_Node != nullptr);
        // sliding local map over a separate, larger cloud
        PointVector cloud(hdr[2]);
        for (int i = 0; i < hdr[2]; ++i) { cloud[i].x = map[3 * i]; cloud[i].y = map[3 * i + 1]; cloud[i].z = map[3 * i + 2]; }
        lm.ikdtree.Build(cloud);
        for (int k = 0; k < hdr[3]; ++k) {
            lm.lasermap_fov_segment(&walk[3 * k]);
            printf("fov %d boxes %zu deleted %d valid %d cache %d box %.9g %.9g %.9g %.9g %.9g %.9g\n", k, lm.cub_needrm.size(), lm.kdtree_delete_counter,
                   lm.ikdtree.validnum(), lm.points_cache_size, lm.LocalMap_Points.vertex_min[0], lm.LocalMap_Points.vertex_min[1],
                   lm.LocalMap_Points.vertex_min[2], lm.LocalMap_Points.vertex_max[0], lm.LocalMap_Points.vertex_max[1], lm.LocalMap_Points.vertex_max[2]);
        }
    } catch (const std::exception& e) {
        printf("error: %s\n", e.what());
        fclose(o);
        return 1;
    }
    fclose(o);
    return 0;
}

// faster_lio::IVox drop-in (host/ivox3d.h), used the way src/laser_mapping.cpp:145-148,520 uses ivox_: AddPoints in two
// batches, then GetClosestPoint for every query (single-query form for the first few, batched form for all)
static int run_ivox(const char* in_path, const char* out_path) {
    FILE* f = fopen(in_path, "rb");
    if (!f) { perror("open"); return 2; }
    int32_t hdr[4];   // n_map, n_queries, nearby (0, 6, 18, 26), unused
    if (fread(hdr, 4, 4, f) != 4) return 2;
    const double res = read_f64(f, 1)[0];
    std::vector<float> map = read_f32(f, (size_t)hdr[0] * 3), qs = read_f32(f, (size_t)hdr[1] * 3);
    fclose(f);
    using IVoxType = faster_lio::IVox<3, faster_lio::IVoxNodeType::DEFAULT, PointType>;   // include/laser_mapping.h:65
    try {
        IVoxType::Options o;
        o.resolution_ = (float)res;
        o.nearby_type_ = hdr[2] == 0 ? IVoxType::NearbyType::CENTER : hdr[2] == 6 ? IVoxType::NearbyType::NEARBY6
                       : hdr[2] == 18 ? IVoxType::NearbyType::NEARBY18 : IVoxType::NearbyType::NEARBY26;
        IVoxType ivox(o);
        PointVector a(hdr[0] / 2), b(hdr[0] - hdr[0] / 2), q(hdr[1]);
        for (int i = 0; i < hdr[0]; ++i) {
            PointType& p = i < hdr[0] / 2 ? a[i] : b[i - hdr[0] / 2];
            p.x = map[3 * i]; p.y = map[3 * i + 1]; p.z = map[3 * i + 2];
        }
        for (int i = 0; i < hdr[1]; ++i) { q[i].x = qs[3 * i]; q[i].y = qs[3 * i + 1]; q[i].z = qs[3 * i + 2]; }
        ivox.AddPoints(a);
        ivox.AddPoints(b);
        std::vector<float> xyz, sqd;
        std::vector<int32_t> ids, cnt;
        ivox.GetClosestPointBatch(q, 5, 5.0, xyz, sqd, ids, cnt);
        // the single-query form must agree with the batch rows and leave the list untouched when nothing is near
        int single_ok = 1;
        for (int i = 0; i < hdr[1] && i < 64; ++i) {
            PointVector near(1);
            near[0].x = 123.f;
            const bool got = ivox.GetClosestPoint(q[i], near, 5);
            if (got != (cnt[i] > 0)) single_ok = 0;
            if (!got && !(near.size() == 1 && near[0].x == 123.f)) single_ok = 0;
            if (got) {
                if ((int)near.size() != cnt[i]) single_ok = 0;
                for (int r = 0; r < cnt[i] && r < (int)near.size(); ++r)
                    if (near[r].x != xyz[(i * 5 + r) * 3] || near[r].y != xyz[(i * 5 + r) * 3 + 1] || near[r].z != xyz[(i * 5 + r) * 3 + 2]) single_ok = 0;
            }
        }
        FILE* out = fopen(out_path, "wb");
        if (!out) { perror("open out"); return 2; }
        int32_t oh[4] = {(int32_t)ivox.NumPoints(), (int32_t)ivox.NumValidGrids(), single_ok, 0};
        fwrite(oh, 4, 4, out);
        fwrite(ids.data(), 4, ids.size(), out);
        fwrite(cnt.data(), 4, cnt.size(), out);
        fwrite(sqd.data(), 4, sqd.size(), out);
        fclose(out);
    } catch (const std::exception& e) {
        printf("error: %s\n", e.what());
        return 3;
    }
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 4 && strcmp(argv[1], "--frontend") == 0) return run_frontend(argv[2], argv[3]);
    if (argc >= 4 && strcmp(argv[1], "--ivox") == 0) return run_ivox(argv[2], argv[3]);
    if (argc >= 2 && strcmp(argv[1], "--no-gpu") == 0) {
        try {
            KD_TREE t(0.5f, 0.6f, 0.5f);
            printf("constructed\n");      // a GPU is present
            return 0;
        } catch (const std::exception& e) {
            printf("refused: %s\n", e.what());
            return 3;
        }
    }
    if (argc < 2) { fprintf(stderr, "usage: test_dropin <scene.bin> | --no-gpu\n"); return 2; }
    FILE* f = fopen(argv[1], "rb");
    if (!f) { perror("open"); return 2; }
    int32_t hdr[4];   // n_map, n_scan, n_frames, max_iter
    if (fread(hdr, 4, 4, f) != 4) return 2;
    double ds = read_f64(f, 1)[0];
    std::vector<float> map = read_f32(f, (size_t)hdr[0] * 3);
    LaserMappingOptions opt;
    opt.max_iteration = hdr[3];
    opt.filter_size_map = ds;
    std::vector<double> ext = read_f64(f, 3);
    memcpy(opt.extrinsic_T, ext.data(), 24);
    try {
        LaserMapping lm(opt);
        // --- KD_TREE surface, as src/laser_mapping.cpp:134-143 uses it ---
        PointVector cloud(hdr[0]);
        for (int i = 0; i < hdr[0]; ++i) { cloud[i].x = map[3 * i]; cloud[i].y = map[3 * i + 1]; cloud[i].z = map[3 * i + 2]; }
        lm.ikdtree.Build(cloud);
        printf("size %d root %d\n", lm.ikdtree.size(), lm.ikdtree.Root_Node != nullptr);
        PointVector near;
        std::vector<float> sq;
        lm.ikdtree.Nearest_Search(cloud[7], 5, near, sq);
        printf("nn5");
        for (size_t i = 0; i < near.size(); ++i) printf(" %.9g %.9g %.9g %.9g", near[i].x, near[i].y, near[i].z, sq[i]);
        printf("\n");
        // --- frames through LaserMapping::Run ---
        for (int fr = 0; fr < hdr[2]; ++fr) {
            std::vector<double> x = read_f64(f, 36);
            std::vector<float> body = read_f32(f, (size_t)hdr[1] * 3);
            memcpy(lm.state.rot_end, &x[0], 72); memcpy(lm.state.pos_end, &x[9], 24);
            memcpy(lm.state.vel_end, &x[24], 24); memcpy(lm.state.bias_g, &x[27], 24); memcpy(lm.state.bias_a, &x[30], 24);
            memcpy(lm.state.gravity, &x[33], 24);
            StatesGroup fresh;
            memcpy(lm.state.cov, fresh.cov, sizeof(fresh.cov));
            for (int i = 0; i < 18; ++i) lm.state.cov[i * 18 + i] = 1e-4;
            lm.feats_down_body.resize(hdr[1]);
            for (int i = 0; i < hdr[1]; ++i) { lm.feats_down_body[i].x = body[3 * i]; lm.feats_down_body[i].y = body[3 * i + 1]; lm.feats_down_body[i].z = body[3 * i + 2]; }
            bool updated = lm.Run();
            printf("frame %d updated %d trips %d searches %d n_eff %d added %d size %d\n", fr, (int)updated, lm.last_report.trips,
                   lm.last_report.searches, lm.effct_feat_num, lm.kdtree_incremental_added, lm.ikdtree.size());
            printf("state");
            for (int i = 0; i < 9; ++i) printf(" %.17g", lm.state.rot_end[i]);
            for (int i = 0; i < 3; ++i) printf(" %.17g", lm.state.pos_end[i]);
            printf("\n");
        }
        // one explicit trip through h_share_model (V1 9x9 / 9 outputs)
        std::vector<double> HPH, HPL;
        lm.nearest_search_en = true;
        lm.h_share_model(HPH, HPL);
        printf("hph00 %.17g hpl0 %.17g n_eff %d\n", HPH[0], HPL[0], lm.effct_feat_num);
        std::vector<BoxPointType> boxes(1);
        for (int i = 0; i < 3; ++i) { boxes[0].vertex_min[i] = -3.f; boxes[0].vertex_max[i] = 3.f; }
        int del = lm.ikdtree.Delete_Point_Boxes(boxes);
        PointVector removed;
        lm.ikdtree.acquire_removed_points(removed);
        printf("deleted %d removed %zu size %d\n", del, removed.size(), lm.ikdtree.size());
    } catch (const std::exception& e) {
        printf("error: %s\n", e.what());
        fclose(f);
        return 1;
    }
    fclose(f);
    return 0;
}
