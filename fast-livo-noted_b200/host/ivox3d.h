// ivox3d.h — ROS/PCL/Eigen-free drop-in for the reference's faster_lio::IVox<3, IVoxNodeType::DEFAULT, PointType>
// (include/ivox3d/ivox3d.h:38-106; the local map the reference builds when USE_ikdtree is off, CMakeLists.txt:15,
// include/laser_mapping.h:65).  Same namespace, class name, Options members and public methods; the
// unordered_map / LRU list of voxels is gone: every call forwards to the GPU-resident voxel hash behind the C ABI
// (include/flivo.h) created in iVox mode (flivo_options::ivox_nearby).  Header-only; link with -lflivo_b200.
//
// Stated differences:
//   * GetClosestPoint returns its list ascending by (distance, insertion id); the reference leaves everything after
//     element 0 in the order std::nth_element produced.
//   * no LRU eviction at Options::capacity_ voxels (ivox3d.h:262-265): the device tables grow instead.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/flivo.h"
#include "flivo_types.h"

namespace faster_lio {

enum class IVoxNodeType { DEFAULT, PHC };   // PHC (ivox3d_node.hpp:64-102) is not built: the reference instantiates DEFAULT

template <int dim = 3, IVoxNodeType node_type = IVoxNodeType::DEFAULT, typename PointT = PointType>
class IVox {
    static_assert(dim == 3 && node_type == IVoxNodeType::DEFAULT, "only IVox<3, DEFAULT> exists on the device");
public:
    using PointVector = ::PointVector;
    enum class NearbyType { CENTER, NEARBY6, NEARBY18, NEARBY26 };
    struct Options {
        float resolution_ = 0.2f;
        float inv_resolution_ = 10.0f;
        NearbyType nearby_type_ = NearbyType::NEARBY6;
        std::size_t capacity_ = 1000000;
        int device_ = 0;            // (new) CUDA device of the map
    };

    explicit IVox(Options options) : options_(options) {
        options_.inv_resolution_ = (float)(1.0 / options_.resolution_);
        flivo_options opt{};
        opt.downsample_size = options_.resolution_;
        opt.sort_scan = 1;
        opt.ivox_nearby = options_.nearby_type_ == NearbyType::CENTER ? FLIVO_IVOX_CENTER
                        : options_.nearby_type_ == NearbyType::NEARBY6 ? FLIVO_IVOX_NEARBY6
                        : options_.nearby_type_ == NearbyType::NEARBY18 ? FLIVO_IVOX_NEARBY18 : FLIVO_IVOX_NEARBY26;
        if (flivo_ctx_create(options_.device_, &opt, &ctx_) != FLIVO_OK)
            throw std::runtime_error("IVox: flivo_ctx_create failed (libflivo_b200 needs a CUDA device; there is no CPU fallback)");
    }
    ~IVox() { flivo_ctx_destroy(ctx_); }
    IVox(const IVox&) = delete;
    IVox& operator=(const IVox&) = delete;

    // ivox3d.h:247-276
    void AddPoints(const PointVector& points_to_add) {
        int32_t added = 0;
        ck(flivo_map_add_points(ctx_, points_to_add.empty() ? nullptr : &points_to_add[0].x, points_to_add.size(), sizeof(PointT), 0, &added), "AddPoints");
    }
    // ivox3d.h:108-130: the single nearest point
    bool GetClosestPoint(const PointT& pt, PointT& closest_pt) {
        float xyz[3], sqd;
        int32_t id, cnt = 0;
        ck(flivo_knn(ctx_, &pt.x, 1, sizeof(PointT), 1, INFINITY, xyz, &sqd, &id, &cnt), "GetClosestPoint");
        if (cnt < 1) return false;
        closest_pt = PointT();
        closest_pt.x = xyz[0]; closest_pt.y = xyz[1]; closest_pt.z = xyz[2];
        return true;
    }
    // ivox3d.h:132-204.  As in the reference, closest_pt is left untouched when no candidate exists.
    bool GetClosestPoint(const PointT& pt, PointVector& closest_pt, int max_num = 5, double max_range = 5.0) {
        if (max_num > FLIVO_K) throw std::invalid_argument("IVox::GetClosestPoint: max_num > 5");
        float xyz[FLIVO_K * 3], sqd[FLIVO_K];
        int32_t ids[FLIVO_K], cnt = 0;
        ck(flivo_knn(ctx_, &pt.x, 1, sizeof(PointT), max_num, max_range * max_range, xyz, sqd, ids, &cnt), "GetClosestPoint");
        if (cnt == 0) return false;
        closest_pt.clear();
        for (int i = 0; i < cnt; ++i) {
            PointT p;
            p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2];
            closest_pt.push_back(p);
        }
        return true;
    }
    // ivox3d.h:245-262: nearest point of every cloud point (default-constructed where none exists)
    bool GetClosestPoint(const PointVector& cloud, PointVector& closest_cloud) {
        const size_t n = cloud.size();
        closest_cloud.assign(n, PointT());
        if (n == 0) return true;
        std::vector<float> xyz(n * 3), sqd(n);
        std::vector<int32_t> ids(n), cnt(n);
        ck(flivo_knn(ctx_, &cloud[0].x, n, sizeof(PointT), 1, INFINITY, xyz.data(), sqd.data(), ids.data(), cnt.data()), "GetClosestPoint");
        for (size_t i = 0; i < n; ++i)
            if (cnt[i] > 0) { closest_cloud[i].x = xyz[3 * i]; closest_cloud[i].y = xyz[3 * i + 1]; closest_cloud[i].z = xyz[3 * i + 2]; }
        return true;
    }
    // batched form of the K-nearest query: row i of the outputs belongs to queries[i] (rows padded with id -1 / +inf)
    void GetClosestPointBatch(const PointVector& queries, int max_num, double max_range, std::vector<float>& xyz,
                              std::vector<float>& sqd, std::vector<int32_t>& ids, std::vector<int32_t>& counts) {
        const size_t n = queries.size();
        xyz.assign(n * max_num * 3, 0.f); sqd.assign(n * max_num, INFINITY); ids.assign(n * max_num, -1); counts.assign(n, 0);
        if (n) ck(flivo_knn(ctx_, &queries[0].x, n, sizeof(PointT), max_num, max_range * max_range, xyz.data(), sqd.data(), ids.data(), counts.data()), "GetClosestPointBatch");
    }
    size_t NumPoints() const { int64_t n = 0; flivo_map_size(ctx_, &n); return (size_t)n; }
    // ivox3d.h:206-209
    size_t NumValidGrids() const { flivo_stats s{}; flivo_get_stats(ctx_, &s); return (size_t)s.map_cells; }

    flivo_ctx* ctx() { return ctx_; }   // for LaserMapping: scan upload / update / map_incremental run on the same context

private:
    void ck(int rc, const char* what) const {
        if (rc != FLIVO_OK) throw std::runtime_error(std::string("IVox::") + what + ": " + flivo_last_error(ctx_));
    }
    Options options_;
    flivo_ctx* ctx_ = nullptr;
};

}  // namespace faster_lio
