// ============================================================================
// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
// C-ABI wrapper around the reference's OWN ikd-Tree, compiled UNMODIFIED from
// /root/reference/include/ikd-Tree/ikd_Tree.cpp (never copied into this repo) against
// oracle/shim/*.  Built by oracle/Makefile into oracle/_ref/libflivo_ref.so.
// Used (a) to pin the restated kNN / Add_Points / Delete_Point_Boxes semantics and
// (b) as the timed CPU baseline ("kind": "reference" for the kNN half): the reference's
// per-point OpenMP loop (src/laser_mapping.cpp:499-545) around KD_TREE::Nearest_Search plus the
// restated plane fit / Jacobian / solve of flivo_oracle.cpp.
// ============================================================================
#include "ikd_Tree.h"          // from /root/reference/include/ikd-Tree (via -I)
#include "flivo_oracle.h"

#include <omp.h>
#include <stdlib.h>
#include <new>
#include <string.h>
#include <vector>

namespace {
inline float id_as_float(int32_t id) { float f; memcpy(&f, &id, 4); return f; }
inline int32_t float_as_id(float f) { int32_t i; memcpy(&i, &f, 4); return i; }

struct RefTree {
    KD_TREE tree;
    int32_t next_id = 0;
    // per-scan caches for ref_iekf_update (Nearest_Points, src/laser_mapping.cpp:165)
    std::vector<PointVector> nearest;
    RefTree(float ds) : tree(0.5f, 0.6f, ds) {}
};

inline PointType make_point(const float* p, int32_t id) {
    PointType q;
    q.x = p[0]; q.y = p[1]; q.z = p[2];
    q.normal_x = id_as_float(id);  // payload survives because the tree copies whole structs
    return q;
}
}  // namespace

extern "C" {

// KD_TREE leaves Rebuild_Ptr (ikd_Tree.h:130) without an initialiser and relies on static (zeroed) storage
// in the reference (a global / class member of a zero-initialised object); emulate that with calloc +
// placement new so the unmodified code sees the same initial state.
void* ref_tree_create(float downsample_size) {
    void* mem = calloc(1, sizeof(RefTree));
    return new (mem) RefTree(downsample_size);
}
// The reference's rebuild-thread entry point (ikd_Tree.cpp:183-186, `void* multi_thread_ptr`) has no
// return statement; under gcc >= 8 with optimisation that is compiled as unreachable code and the
// thread crashes when ~KD_TREE() asks it to terminate.  The wrapper therefore never destructs a tree:
// it releases the nodes through the public API (Build of an empty cloud, :337-341) and parks the
// object (and its idle thread) until process exit.
void ref_tree_destroy(void* h) {
    RefTree* t = static_cast<RefTree*>(h);
    t->tree.Build(PointVector());
}

// KD_TREE::Build (ikd_Tree.cpp:337-348); ids 0..n-1
void ref_tree_build(void* h, const float* xyz, size_t n) {
    RefTree* t = static_cast<RefTree*>(h);
    PointVector pv(n);
    for (size_t i = 0; i < n; ++i) pv[i] = make_point(xyz + 3 * i, (int32_t)i);
    t->next_id = (int32_t)n;
    t->tree.Build(pv);
}
// KD_TREE::Add_Points (ikd_Tree.cpp:382-457); every offered point consumes one id
int ref_tree_add_points(void* h, const float* xyz, size_t n, int downsample_on) {
    RefTree* t = static_cast<RefTree*>(h);
    PointVector pv(n);
    for (size_t i = 0; i < n; ++i) pv[i] = make_point(xyz + 3 * i, t->next_id++);
    return t->tree.Add_Points(pv, downsample_on != 0);
}
// KD_TREE::Delete_Point_Boxes (ikd_Tree.cpp:501-521)
int ref_tree_delete_boxes(void* h, const float* boxes6, size_t nb) {
    RefTree* t = static_cast<RefTree*>(h);
    std::vector<BoxPointType> bv(nb);
    for (size_t i = 0; i < nb; ++i)
        for (int c = 0; c < 3; ++c) { bv[i].vertex_min[c] = boxes6[6 * i + c]; bv[i].vertex_max[c] = boxes6[6 * i + 3 + c]; }
    return t->tree.Delete_Point_Boxes(bv);
}
int ref_tree_size(void* h) { return static_cast<RefTree*>(h)->tree.size(); }
int ref_tree_validnum(void* h) { return static_cast<RefTree*>(h)->tree.validnum(); }
// KD_TREE::flatten(Root_Node, storage, NOT_RECORD) (ikd_Tree.cpp:1238-1263)
size_t ref_tree_flatten(void* h, float* xyz, int32_t* ids, size_t cap) {
    RefTree* t = static_cast<RefTree*>(h);
    PointVector pv;
    t->tree.flatten(t->tree.Root_Node, pv, NOT_RECORD);
    size_t n = pv.size() < cap ? pv.size() : cap;
    for (size_t i = 0; i < n; ++i) {
        if (xyz) { xyz[3 * i] = pv[i].x; xyz[3 * i + 1] = pv[i].y; xyz[3 * i + 2] = pv[i].z; }
        if (ids) ids[i] = float_as_id(pv[i].normal_x);
    }
    return pv.size();
}
// Batched KD_TREE::Nearest_Search (ikd_Tree.cpp:350-380) under the reference's OpenMP pattern.
void ref_tree_knn(void* h, const float* q_xyz, size_t nq, int k, double max_dist, int nthreads,
                  int32_t* out_id, float* out_sqd, float* out_xyz, int32_t* out_count) {
    RefTree* t = static_cast<RefTree*>(h);
    if (nthreads > 0) omp_set_num_threads(nthreads);
#pragma omp parallel for
    for (long i = 0; i < (long)nq; ++i) {
        PointType q = make_point(q_xyz + 3 * i, -1);
        PointVector near;
        std::vector<float> sq;
        t->tree.Nearest_Search(q, k, near, sq, max_dist);
        int found = (int)near.size();
        out_count[i] = found;
        for (int s = 0; s < k; ++s) {
            bool ok = s < found;
            out_id[i * k + s] = ok ? float_as_id(near[s].normal_x) : -1;
            out_sqd[i * k + s] = ok ? sq[s] : INFINITY;
            if (out_xyz) {
                out_xyz[(i * k + s) * 3 + 0] = ok ? near[s].x : 0.f;
                out_xyz[(i * k + s) * 3 + 1] = ok ? near[s].y : 0.f;
                out_xyz[(i * k + s) * 3 + 2] = ok ? near[s].z : 0.f;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Whole LIO measurement update on the CPU: the reference's IEKF loop structure
// (V1: src/laser_mapping.cpp:165-242; V3: esekfom.hpp:1619-1928) with the reference's own ikd-Tree
// for Nearest_Search, `#pragma omp parallel for` with `nthreads` threads (MP_PROC_NUM,
// CMakeLists.txt:27-46), and the restated plane fit / Jacobian / solve.
//   P: 23x23 (V3) or 18x18 (V1), in/out.   timings: [0]=kNN loop, [1]=plane+Jacobian+Gram, [2]=solve (seconds)
//   force_search: 1 => kNN every trip ("worst case" label in SURVEY §3.2)
// Returns number of trips executed.
// ---------------------------------------------------------------------------------------------
int ref_iekf_update(void* h, const float* body_xyz, size_t n, orc_state* x, double* P, int mode,
                    int max_iter, double laser_point_cov, const double* limit23, int nthreads,
                    int force_search, double* timings, int* searches_out, int64_t* n_eff_out,
                    uint8_t* eff_mask_out /* n, last trip */, int32_t* near_id_out /* n*5, last search */) {
    RefTree* t = static_cast<RefTree*>(h);
    if (nthreads > 0) omp_set_num_threads(nthreads);
    orc_state x_prop = *x;
    const int np = (mode == ORC_MODE_V1) ? 18 : 23;
    std::vector<double> P_prop(P, P + np * np);
    std::vector<float> world(n * 3), near_xyz(n * 15), near_d5(n), normvec(n * 4);
    std::vector<int32_t> near_cnt(n), near_id(n * 5, -1);
    std::vector<uint8_t> mask(n, 1), eff(n, 0);
    double HTH[144], HTz[12], HPH[81], HPL[9], res_sum;
    int64_t n_eff = 0;
    int trips = 0, searches = 0;
    int search = 1, tcount = 0, rematch = 0, stopped = 0;
    if (timings) timings[0] = timings[1] = timings[2] = 0.0;
    for (int it = -1; it < max_iter && !stopped; ++it) {
        double t0 = omp_get_wtime();
        if (search || force_search) {
            orc_body_to_world(body_xyz, n, 3, x, world.data());
#pragma omp parallel for
            for (long i = 0; i < (long)n; ++i) {
                PointType q = make_point(&world[3 * i], -1);
                PointVector near;
                std::vector<float> sq(5);
                t->tree.Nearest_Search(q, 5, near, sq);
                int found = (int)near.size();
                near_cnt[i] = found;
                near_d5[i] = found >= 5 ? sq[4] : INFINITY;
                for (int s = 0; s < 5; ++s) {
                    bool ok = s < found;
                    near_xyz[i * 15 + 3 * s + 0] = ok ? near[s].x : 0.f;
                    near_xyz[i * 15 + 3 * s + 1] = ok ? near[s].y : 0.f;
                    near_xyz[i * 15 + 3 * s + 2] = ok ? near[s].z : 0.f;
                    near_id[i * 5 + s] = ok ? float_as_id(near[s].normal_x) : -1;
                }
            }
            searches++;
        }
        double t1 = omp_get_wtime();
        orc_h_share_model(body_xyz, n, 3, x, near_xyz.data(), near_cnt.data(), near_d5.data(),
                          (search || force_search) ? 1 : 0, mode, mask.data(), eff.data(), normvec.data(),
                          world.data(), HTH, HTz, &n_eff, &res_sum, nullptr);
        double t2 = omp_get_wtime();
        if (mode == ORC_MODE_V1) {
            orc_repack_v1(HTH, HTz, laser_point_cov, HPH, HPL);
            orc_solve_v1(x, &x_prop, P, HPH, HPL, it, max_iter, &rematch, &search, &stopped, nullptr);
        } else {
            std::vector<double> P_out(np * np);
            orc_solve_v3(x, &x_prop, P_prop.data(), HTH, HTz, laser_point_cov, it, max_iter, limit23,
                         &tcount, &search, &stopped, nullptr, P_out.data());
            if (stopped) memcpy(P, P_out.data(), sizeof(double) * np * np);
        }
        double t3 = omp_get_wtime();
        if (timings) { timings[0] += t1 - t0; timings[1] += t2 - t1; timings[2] += t3 - t2; }
        trips++;
    }
    if (searches_out) *searches_out = searches;
    if (n_eff_out) *n_eff_out = n_eff;
    if (eff_mask_out) memcpy(eff_mask_out, eff.data(), n);
    if (near_id_out) memcpy(near_id_out, near_id.data(), sizeof(int32_t) * n * 5);
    return trips;
}

}  // extern "C"
