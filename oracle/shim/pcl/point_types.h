// TEST INFRASTRUCTURE (oracle/): minimal stand-in for <pcl/point_types.h> so that the
// reference's include/ikd-Tree/ikd_Tree.{h,cpp} can be compiled UNMODIFIED from
// /root/reference into oracle/_ref/ (see oracle/Makefile).  PCL itself is not installed.
// Only the layout facts ikd-Tree relies on are reproduced: pcl::PointXYZINormal is a
// 48-byte, 16-byte-aligned POD whose x,y,z lead the struct (reference
// include/ikd-Tree/ikd_Tree.h:21, include/common_lib.h:48).
#pragma once
#include <math.h>
#include <cmath>
#include <string.h>
#include <unistd.h>
#include <algorithm>
#include <vector>

#include "../Eigen/eigen_min.h"

namespace pcl {
// getVector3fMap(): the x, y, z view the reference's include/ivox3d headers use (ivox3d_node.hpp:14, ivox3d.h:264)
struct alignas(16) PointXYZINormal {
    float x = 0.f, y = 0.f, z = 0.f, _pad0 = 1.f;
    float normal_x = 0.f, normal_y = 0.f, normal_z = 0.f, _pad1 = 0.f;
    float intensity = 0.f, curvature = 0.f, _pad2 = 0.f, _pad3 = 0.f;
    Eigen::Map3f getVector3fMap() { return Eigen::Map3f{&x}; }
    Eigen::Map3f getVector3fMap() const { return Eigen::Map3f{const_cast<float*>(&x)}; }
};
static_assert(sizeof(PointXYZINormal) == 48, "PointXYZINormal must be 48 bytes");
struct alignas(16) PointXYZ {
    float x = 0.f, y = 0.f, z = 0.f, _pad0 = 1.f;
    Eigen::Map3f getVector3fMap() { return Eigen::Map3f{&x}; }
    Eigen::Map3f getVector3fMap() const { return Eigen::Map3f{const_cast<float*>(&x)}; }
};
struct alignas(16) PointXYZI {
    float x = 0.f, y = 0.f, z = 0.f, _pad0 = 1.f;
    float intensity = 0.f, _pad1 = 0.f, _pad2 = 0.f, _pad3 = 0.f;
    Eigen::Map3f getVector3fMap() { return Eigen::Map3f{&x}; }
    Eigen::Map3f getVector3fMap() const { return Eigen::Map3f{const_cast<float*>(&x)}; }
};
}  // namespace pcl
