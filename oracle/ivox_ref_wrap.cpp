// TEST INFRASTRUCTURE (oracle/).  C wrapper around the reference's OWN faster_lio::IVox<3, DEFAULT, pcl::PointXYZINormal>
// (include/ivox3d/ivox3d.h, ivox3d_node.hpp, eigen_types.h), compiled unmodified from where they lie under /root/reference
// against the stand-ins of oracle/shim (Eigen column vectors, pcl point types, glog): oracle/_ref/libivox_ref.so.
// It exists to pin oracle/ivox_oracle.cpp (the restatement) on the real container logic — voxel keys, neighbour stencils,
// per-voxel truncation, nth_element selection, LRU eviction at capacity_.  The three Eigen calls with arithmetic in them
// (squaredNorm, array().round(), cast<int>()) are the shim's reading of Eigen (oracle/shim/Eigen/eigen_min.h).
// The single-point GetClosestPoint(pt, closest_pt) (ivox3d.h:108-130) calls NodeType::NNPoint, which the DEFAULT node
// does not have: it cannot be instantiated for the configuration the reference builds, so it is not wrapped.
#include <cstdint>
#include <vector>

#include <pcl/point_types.h>
#include <unordered_map>   // (ivox3d.h uses std::unordered_map / std::string / std::cout without including their headers)
#include <string>
#include <iostream>
#include <numeric>
#include <chrono>
#include "ivox3d/ivox3d.h"

namespace {
using Pt = pcl::PointXYZINormal;
using IVoxT = faster_lio::IVox<3, faster_lio::IVoxNodeType::DEFAULT, Pt>;
}

extern "C" {

void* ref_ivox_create(float resolution, int nearby /* 0 CENTER, 1 NEARBY6, 2 NEARBY18, 3 NEARBY26 */, long long capacity) {
    IVoxT::Options o;
    o.resolution_ = resolution;
    o.nearby_type_ = nearby == 0 ? IVoxT::NearbyType::CENTER : nearby == 1 ? IVoxT::NearbyType::NEARBY6
                   : nearby == 2 ? IVoxT::NearbyType::NEARBY18 : IVoxT::NearbyType::NEARBY26;
    o.capacity_ = (std::size_t)capacity;
    return new IVoxT(o);
}
void ref_ivox_destroy(void* h) { delete static_cast<IVoxT*>(h); }

void ref_ivox_add_points(void* h, const float* xyz, long long n) {
    IVoxT::PointVector pts((std::size_t)n);
    for (long long i = 0; i < n; ++i) { pts[i].x = xyz[3 * i]; pts[i].y = xyz[3 * i + 1]; pts[i].z = xyz[3 * i + 2]; }
    static_cast<IVoxT*>(h)->AddPoints(pts);
}
long long ref_ivox_num_grids(void* h) { return (long long)static_cast<IVoxT*>(h)->NumValidGrids(); }

// GetClosestPoint(pt, closest_pt, max_num, max_range) per query (ivox3d.h:132-204).  out_xyz: [nq][max_num][3] in the order
// the reference returned them; out_count[i] = -1 when it returned false (closest_pt untouched), else closest_pt.size().
void ref_ivox_knn(void* h, const float* q, long long nq, int max_num, double max_range, float* out_xyz, int32_t* out_count) {
    IVoxT* iv = static_cast<IVoxT*>(h);
    IVoxT::PointVector res;
    for (long long i = 0; i < nq; ++i) {
        Pt p;
        p.x = q[3 * i]; p.y = q[3 * i + 1]; p.z = q[3 * i + 2];
        res.clear();
        const bool ok = iv->GetClosestPoint(p, res, max_num, max_range);
        out_count[i] = ok ? (int32_t)res.size() : -1;
        for (int r = 0; r < max_num; ++r) {
            const bool have = ok && r < (int)res.size();
            out_xyz[((std::size_t)i * max_num + r) * 3 + 0] = have ? res[r].x : 0.f;
            out_xyz[((std::size_t)i * max_num + r) * 3 + 1] = have ? res[r].y : 0.f;
            out_xyz[((std::size_t)i * max_num + r) * 3 + 2] = have ? res[r].z : 0.f;
        }
    }
}

}  // extern "C"
