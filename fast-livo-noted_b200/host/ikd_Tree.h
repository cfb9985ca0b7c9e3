// ikd_Tree.h — ROS/PCL-free drop-in for the reference's KD_TREE (include/ikd-Tree/ikd_Tree.h:166-188).
// Same class name, same public methods and argument meaning; the k-d tree, its rebuild thread and its
// operation logger are gone: every call forwards to the GPU-resident map behind the C ABI (include/flivo.h).
// Header-only; link with -lflivo_b200.
#pragma once
#include <math.h>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/flivo.h"
#include "flivo_types.h"

class KD_TREE {
public:
    // ikd_Tree.cpp:9-16.  delete_param / balance_param steer re-balancing of a tree that no longer exists.
    KD_TREE(float delete_param = 0.5f, float balance_param = 0.6f, float box_length = 0.2f, int device = 0) {
        (void)delete_param; (void)balance_param;
        flivo_options opt{};
        opt.downsample_size = box_length;
        opt.sort_scan = 1;
        int rc = flivo_ctx_create(device, &opt, &ctx_);
        if (rc != FLIVO_OK) throw std::runtime_error("KD_TREE: flivo_ctx_create failed (libflivo_b200 needs a CUDA device; there is no CPU fallback)");
    }
    ~KD_TREE() { flivo_ctx_destroy(ctx_); }
    KD_TREE(const KD_TREE&) = delete;
    KD_TREE& operator=(const KD_TREE&) = delete;

    void Set_delete_criterion_param(float) {}
    void Set_balance_criterion_param(float) {}
    void set_downsample_param(float box_length) { ck(flivo_map_set_downsample(ctx_, box_length), "set_downsample_param"); }
    void InitializeKDTree(float delete_param = 0.5f, float balance_param = 0.7f, float box_length = 0.2f) {
        (void)delete_param; (void)balance_param;
        set_downsample_param(box_length);
    }
    int size() { int64_t n = 0; ck(flivo_map_size(ctx_, &n), "size"); return (int)n; }
    int validnum() { return size(); }
    void root_alpha(float& alpha_bal, float& alpha_del) { alpha_bal = 0.5f; alpha_del = 0.f; }   // a hash map is always "balanced"

    // ikd_Tree.cpp:337-348 (argument by value, as in the reference)
    void Build(PointVector point_cloud) {
        ck(flivo_map_build(ctx_, point_cloud.empty() ? nullptr : &point_cloud[0].x, point_cloud.size(), sizeof(PointType)), "Build");
        Root_Node = point_cloud.empty() ? nullptr : sentinel();
    }
    // ikd_Tree.cpp:350-380.  One query = one kernel launch: kept for API parity (and made thread-safe for the
    // reference's OpenMP callers); hot paths use Nearest_Search_Batch or flivo_iekf_update.
    void Nearest_Search(PointType point, int k_nearest, PointVector& Nearest_Points, std::vector<float>& Point_Distance,
                        double max_dist = INFINITY) {
        std::lock_guard<std::mutex> lock(mu_);
        if (k_nearest > FLIVO_K) throw std::invalid_argument("KD_TREE::Nearest_Search: k_nearest > 5");
        float xyz[FLIVO_K * 3], sqd[FLIVO_K];
        int32_t ids[FLIVO_K], cnt = 0;
        ck(flivo_knn(ctx_, &point.x, 1, sizeof(PointType), k_nearest, max_dist, xyz, sqd, ids, &cnt), "Nearest_Search");
        Nearest_Points.clear();
        Point_Distance.clear();
        for (int i = 0; i < cnt; ++i) {
            PointType p;
            p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2];
            Nearest_Points.push_back(p);
            Point_Distance.push_back(sqd[i]);
        }
    }
    // batched form: row i of the outputs belongs to queries[i]; rows are padded with id -1 / +inf
    void Nearest_Search_Batch(const PointVector& queries, int k_nearest, std::vector<float>& xyz, std::vector<float>& sqd,
                              std::vector<int32_t>& ids, std::vector<int32_t>& counts, double max_dist = INFINITY) {
        std::lock_guard<std::mutex> lock(mu_);
        size_t n = queries.size();
        xyz.assign(n * k_nearest * 3, 0.f); sqd.assign(n * k_nearest, INFINITY); ids.assign(n * k_nearest, -1); counts.assign(n, 0);
        if (n == 0) return;
        ck(flivo_knn(ctx_, &queries[0].x, n, sizeof(PointType), k_nearest, max_dist, xyz.data(), sqd.data(), ids.data(), counts.data()),
           "Nearest_Search_Batch");
    }
    // ikd_Tree.cpp:382-457
    int Add_Points(PointVector& PointToAdd, bool downsample_on) {
        int32_t added = 0;
        ck(flivo_map_add_points(ctx_, PointToAdd.empty() ? nullptr : &PointToAdd[0].x, PointToAdd.size(), sizeof(PointType),
                                downsample_on ? 1 : 0, &added), "Add_Points");
        if (added > 0 || size() > 0) Root_Node = sentinel();
        return added;
    }
    void Add_Point_Boxes(std::vector<BoxPointType>& BoxPoints) {
        ck(flivo_map_add_boxes(ctx_, BoxPoints.empty() ? nullptr : BoxPoints[0].vertex_min, BoxPoints.size(), nullptr), "Add_Point_Boxes");
    }
    void Delete_Points(PointVector& PointToDel) {
        ck(flivo_map_delete_points(ctx_, PointToDel.empty() ? nullptr : &PointToDel[0].x, PointToDel.size(), sizeof(PointType), nullptr), "Delete_Points");
    }
    int Delete_Point_Boxes(std::vector<BoxPointType>& BoxPoints) {
        int32_t n = 0;
        ck(flivo_map_delete_boxes(ctx_, BoxPoints.empty() ? nullptr : BoxPoints[0].vertex_min, BoxPoints.size(), &n), "Delete_Point_Boxes");
        return n;
    }
    // ikd_Tree.cpp:1238-1263: callers pass Root_Node; only NOT_RECORD semantics exist here (order unspecified)
    void flatten(KD_TREE_NODE* root, PointVector& Storage, delete_point_storage_set storage_type) {
        (void)storage_type;
        Storage.clear();
        if (root == nullptr) return;
        size_t n = 0;
        ck(flivo_map_flatten(ctx_, nullptr, nullptr, 0, &n), "flatten");
        std::vector<float> xyz(n * 3);
        if (n) ck(flivo_map_flatten(ctx_, xyz.data(), nullptr, n, &n), "flatten");
        Storage.resize(n);
        for (size_t i = 0; i < n; ++i) { Storage[i].x = xyz[3 * i]; Storage[i].y = xyz[3 * i + 1]; Storage[i].z = xyz[3 * i + 2]; }
    }
    void acquire_removed_points(PointVector& removed_points) {
        size_t n = 0;
        ck(flivo_map_acquire_removed(ctx_, nullptr, 0, &n), "acquire_removed_points");
        std::vector<float> xyz(n * 3);
        if (n) ck(flivo_map_acquire_removed(ctx_, xyz.data(), n, &n), "acquire_removed_points");
        removed_points.resize(n);
        for (size_t i = 0; i < n; ++i) { removed_points[i].x = xyz[3 * i]; removed_points[i].y = xyz[3 * i + 1]; removed_points[i].z = xyz[3 * i + 2]; }
    }
    BoxPointType tree_range() {
        BoxPointType b;
        float r[6];
        ck(flivo_map_range(ctx_, r), "tree_range");
        for (int i = 0; i < 3; ++i) { b.vertex_min[i] = r[i]; b.vertex_max[i] = r[3 + i]; }
        return b;
    }
    PointVector PCL_Storage;
    KD_TREE_NODE* Root_Node = nullptr;   // nullptr <=> map empty (src/laser_mapping.cpp:134 tests exactly this)
    int max_queue_size = 0;

    flivo_ctx* ctx() { return ctx_; }    // for LaserMapping (scan upload + on-device update share the map's context)

private:
    static KD_TREE_NODE* sentinel() { static char s; return reinterpret_cast<KD_TREE_NODE*>(&s); }
    void ck(int rc, const char* what) {
        if (rc != FLIVO_OK) throw std::runtime_error(std::string("KD_TREE::") + what + ": " + flivo_last_error(ctx_));
    }
    flivo_ctx* ctx_ = nullptr;
    std::mutex mu_;
};
