// TEST INFRASTRUCTURE (oracle/): the few pieces of Eigen that the reference's include/ivox3d/{ivox3d.h,
// ivox3d_node.hpp,eigen_types.h} use, so that those headers compile UNMODIFIED from /root/reference into
// oracle/_ref/libivox_ref.so (oracle/Makefile).  Eigen itself is not installed.  Only fixed-size column vectors
// exist; everything else is a declaration for the aliases of eigen_types.h.  Arithmetic follows Eigen 3.x:
//   squaredNorm of a 3-vector = x*x + (y*y + z*z)   (redux_novec_unroller splits [0,3) into [0,1) and [1,3))
//   array().round()           = std::round per coefficient (half away from zero)
//   cast<int>()               = static_cast per coefficient
// These three are MY reading of Eigen, not Eigen: what the compiled header pins is the container logic of
// IVox (voxel keys, neighbour stencils, per-voxel truncation, nth_element selection, LRU eviction).
#pragma once
#include <cmath>
#include <cstddef>

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

namespace Eigen {
template <class T, int R, int C = 1>
struct Matrix;   // only column vectors are defined below

template <class T, int N>
struct Matrix<T, N, 1> {
    T v[N];
    Matrix() { for (int i = 0; i < N; ++i) v[i] = T(0); }
    Matrix(T a, T b) { static_assert(N == 2, "size"); v[0] = a; v[1] = b; }
    Matrix(T a, T b, T c) { static_assert(N == 3, "size"); v[0] = a; v[1] = b; v[2] = c; }
    static Matrix Zero() { return Matrix(); }
    T& operator[](int i) { return v[i]; }
    const T& operator[](int i) const { return v[i]; }
    Matrix operator+(const Matrix& o) const { Matrix r; for (int i = 0; i < N; ++i) r.v[i] = v[i] + o.v[i]; return r; }
    Matrix operator-(const Matrix& o) const { Matrix r; for (int i = 0; i < N; ++i) r.v[i] = v[i] - o.v[i]; return r; }
    Matrix operator*(T s) const { Matrix r; for (int i = 0; i < N; ++i) r.v[i] = v[i] * s; return r; }
    bool operator==(const Matrix& o) const { for (int i = 0; i < N; ++i) if (!(v[i] == o.v[i])) return false; return true; }
    bool operator!=(const Matrix& o) const { return !(*this == o); }
    T squaredNorm() const {
        if (N == 3) return v[0] * v[0] + (v[1] * v[1] + v[2] * v[2]);
        T s = v[0] * v[0];
        for (int i = 1; i < N; ++i) s = s + v[i] * v[i];
        return s;
    }
    const Matrix& array() const { return *this; }
    Matrix round() const { Matrix r; for (int i = 0; i < N; ++i) r.v[i] = std::round(v[i]); return r; }
    template <class U>
    Matrix<U, N, 1> cast() const { Matrix<U, N, 1> r; for (int i = 0; i < N; ++i) r.v[i] = static_cast<U>(v[i]); return r; }
};

// what pcl's getVector3fMap() returns: a view of the point's x, y, z
struct Map3f {
    float* p;
    operator Matrix<float, 3, 1>() const { return Matrix<float, 3, 1>(p[0], p[1], p[2]); }
    Map3f& operator=(const Matrix<float, 3, 1>& m) { p[0] = m[0]; p[1] = m[1]; p[2] = m[2]; return *this; }
    Matrix<float, 3, 1> operator-(const Map3f& o) const { return Matrix<float, 3, 1>(p[0] - o.p[0], p[1] - o.p[1], p[2] - o.p[2]); }
    Matrix<float, 3, 1> operator-(const Matrix<float, 3, 1>& o) const { return Matrix<float, 3, 1>(p[0] - o[0], p[1] - o[1], p[2] - o[2]); }
};

template <class T> struct Quaternion;
typedef Matrix<int, 2, 1> Vector2i;
typedef Matrix<int, 3, 1> Vector3i;
typedef Matrix<double, 2, 1> Vector2d;
typedef Matrix<float, 2, 1> Vector2f;
typedef Matrix<double, 3, 1> Vector3d;
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<double, 3, 3> Matrix3d;
typedef Matrix<float, 3, 3> Matrix3f;
typedef Matrix<double, 4, 4> Matrix4d;
typedef Matrix<float, 4, 4> Matrix4f;
typedef Quaternion<double> Quaterniond;
typedef Quaternion<float> Quaternionf;
}  // namespace Eigen
