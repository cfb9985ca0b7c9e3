// TEST INFRASTRUCTURE (oracle/): stand-in for <pcl/common/centroid.h>.  ivox3d_node.hpp includes it for
// pcl::CentroidPoint, used only by the PHC node (ivox3d_node.hpp:223-236), which the reference never instantiates.
#pragma once
#include <cassert>
#include <cstddef>
#include "../point_types.h"
namespace pcl {
template <class PointT>
struct CentroidPoint {
    double sx = 0, sy = 0, sz = 0;
    std::size_t n = 0;
    void add(const PointT& p) { sx += p.x; sy += p.y; sz += p.z; ++n; }
    template <class OutT> void get(OutT& o) const { if (n) { o.x = (float)(sx / n); o.y = (float)(sy / n); o.z = (float)(sz / n); } }
};
}  // namespace pcl
