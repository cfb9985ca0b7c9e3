// flivo_types.h — plain types the ROS-free drop-in classes share with their callers.
//
// The reference builds on PCL and Eigen (include/ikd-Tree/ikd_Tree.h:2-4,21-22; include/common_lib.h:518-603),
// neither of which exists in this image.  The layouts below are byte-compatible with the reference's:
//   PointType   == pcl::PointXYZINormal (48 bytes, 16-byte aligned: x y z _ | nx ny nz _ | intensity curvature _ _)
//   BoxPointType as include/ikd-Tree/ikd_Tree.h:63-66
// so a maintainer who has PCL/Eigen defines FLIVO_USE_PCL and the same headers compile against the real types
// (see INTEGRATION.md).  Rotations are row-major 3x3, matching flivo_state.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <cmath>
#include <new>
#include <vector>

#ifdef FLIVO_USE_PCL
#include <pcl/point_types.h>
#include <Eigen/StdVector>
typedef pcl::PointXYZINormal PointType;
typedef std::vector<PointType, Eigen::aligned_allocator<PointType>> PointVector;
#else
struct alignas(16) PointType {
    float x = 0.f, y = 0.f, z = 0.f, data_pad = 1.f;
    float normal_x = 0.f, normal_y = 0.f, normal_z = 0.f, normal_pad = 0.f;
    float intensity = 0.f, curvature = 0.f, tail_pad[2] = {0.f, 0.f};
};
template <class T>
struct FlivoAlignedAllocator {
    typedef T value_type;
    FlivoAlignedAllocator() = default;
    template <class U> FlivoAlignedAllocator(const FlivoAlignedAllocator<U>&) {}
    T* allocate(size_t n) {
        void* p = nullptr;
        if (posix_memalign(&p, 16, n * sizeof(T) ? n * sizeof(T) : 16) != 0) throw std::bad_alloc();
        return static_cast<T*>(p);
    }
    void deallocate(T* p, size_t) { free(p); }
    template <class U> bool operator==(const FlivoAlignedAllocator<U>&) const { return true; }
    template <class U> bool operator!=(const FlivoAlignedAllocator<U>&) const { return false; }
};
typedef std::vector<PointType, FlivoAlignedAllocator<PointType>> PointVector;
#endif
static_assert(sizeof(PointType) == 48, "PointType must keep the 48-byte pcl::PointXYZINormal layout");

struct BoxPointType {
    float vertex_min[3];
    float vertex_max[3];
};

enum delete_point_storage_set { NOT_RECORD, DELETE_POINTS_REC, MULTI_THREAD_REC };

struct KD_TREE_NODE;   // opaque: the GPU map has no nodes; Root_Node is only a "map is built" sentinel

// 18-DOF state of the compiled estimator (StatesGroup, include/common_lib.h:518-603), Eigen-free.
struct StatesGroup {
    double rot_end[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    double pos_end[3] = {0, 0, 0};
    double vel_end[3] = {0, 0, 0};
    double bias_g[3] = {0, 0, 0};
    double bias_a[3] = {0, 0, 0};
    double gravity[3] = {0, 0, -9.805};
    double cov[18 * 18];                 // row-major, order rot, pos, vel, bg, ba, grav
    StatesGroup() {
        for (int i = 0; i < 18 * 18; ++i) cov[i] = 0.0;
        for (int i = 0; i < 18; ++i) cov[i * 18 + i] = 0.001;   // INIT_COV, include/common_lib.h:36
    }
};
