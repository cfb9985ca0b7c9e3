// C++ caller of the drop-in classes, written the way the reference's own call sites use them
// (src/laser_mapping.cpp:134-143,165-251): KD_TREE Build / Nearest_Search / Add_Points / Delete_Point_Boxes and
// LaserMapping::Run over a few frames of a synthetic scene.  Reads the scene from a flat binary file written by
// tests/test_dropin_cpp.py and prints results as "key value..." lines for the Python side to compare with the oracle.
//   test_dropin <scene.bin>         run on cuda:0
//   test_dropin --no-gpu            only check that construction fails loudly without a device
#include <stdio.h>
#include <string.h>
#include <stdexcept>
#include <vector>

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

int main(int argc, char** argv) {
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
