// ============================================================================
// TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See flivo_oracle.h for scope and parity status.
// Build: g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC  (no -march=native, no -ffast-math:
// the reference is built -O3 without FMA, CMakeLists.txt:7,23; every float that feeds a
// comparison is evaluated here in the reference's source order with separate mul/add).
// All citations are relative to /root/reference/.
// ============================================================================
#include "flivo_oracle.h"

#include <math.h>
#include <string.h>
#include <float.h>
#include <algorithm>
#include <array>
#include <unordered_map>
#include <vector>

namespace {

// ---- tiny fixed-size helpers (double) --------------------------------------------------------
// Eigen evaluates a fixed-size 3-term reduction as x0 + (x1 + x2) (Eigen/src/Core/Redux.h,
// redux_novec_unroller splits [0,3) into [0,1) and [1,3)).
inline double dot3(const double* a, const double* b) { return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]); }
inline void matvec3(const double* M, const double* v, double* o) {  // row-major M
    o[0] = dot3(M + 0, v); o[1] = dot3(M + 3, v); o[2] = dot3(M + 6, v);
}
inline void matTvec3(const double* M, const double* v, double* o) {
    for (int c = 0; c < 3; ++c) o[c] = M[c] * v[0] + (M[3 + c] * v[1] + M[6 + c] * v[2]);
}
inline void cross3(const double* a, const double* b, double* o) {  // [a]x b
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}

// a1: V3D p_global(state.rot_end * (Lidar_rot_to_IMU*p_body + Lidar_offset_to_IMU) + state.pos_end)
// src/laser_mapping.cpp:664-669; result narrowed to float on assignment to PointType.
inline void body_to_world_1(const float* pb, const orc_state* st, float* pw) {
    double p[3] = {(double)pb[0], (double)pb[1], (double)pb[2]}, a[3], b[3];
    matvec3(st->offR, p, a);
    a[0] += st->offT[0]; a[1] += st->offT[1]; a[2] += st->offT[2];
    matvec3(st->rot, a, b);
    pw[0] = (float)(b[0] + st->pos[0]);
    pw[1] = (float)(b[1] + st->pos[1]);
    pw[2] = (float)(b[2] + st->pos[2]);
}

// calc_dist, include/ikd-Tree/ikd_Tree.cpp:1291-1295 (float, left-to-right, no FMA)
inline float sqdist(const float* a, const float* b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    return dx * dx + dy * dy + dz * dz;
}

}  // namespace

extern "C" void orc_body_to_world(const float* body_xyz, size_t n, size_t stride_floats,
                                  const orc_state* st, float* world_xyz) {
#pragma omp parallel for schedule(static)
    for (long i = 0; i < (long)n; ++i) body_to_world_1(body_xyz + i * stride_floats, st, world_xyz + 3 * i);
}

// ---------------------------------------------------------------------------------------------
// a2/a3 brute-force kNN.  Membership rule of KD_TREE::Search (ikd_Tree.cpp:858-864): a point enters
// iff d <= max_dist and (fewer than k held or d < current k-th) — i.e. the exact k nearest; equal
// distances at rank k are resolved in the reference by visiting order (tree-shape dependent), here
// by the canonical order (d, id).  Output ascending (ikd_Tree.cpp:371-378).
// ---------------------------------------------------------------------------------------------
extern "C" void orc_knn_bruteforce(const float* map_xyz, const int32_t* map_id, const uint8_t* alive, size_t m,
                                   const float* q_xyz, size_t nq, int k, double max_dist,
                                   int32_t* out_id, float* out_sqd, float* out_xyz, int32_t* out_count,
                                   int64_t* tie_straddles) {
    int64_t ties = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : ties)
    for (long qi = 0; qi < (long)nq; ++qi) {
        const float* q = q_xyz + 3 * qi;
        std::vector<float> bd(k + 1, INFINITY);
        std::vector<int32_t> bi(k + 1, INT32_MAX);
        std::vector<long> bj(k + 1, -1);
        int cnt = 0;  // number held (<= k+1; slot k is the runner-up, for tie statistics)
        for (size_t j = 0; j < m; ++j) {
            if (alive && !alive[j]) continue;
            float d = sqdist(q, map_xyz + 3 * j);
            if (!((double)d <= max_dist)) continue;
            int32_t id = map_id ? map_id[j] : (int32_t)j;
            // insertion into a (k+1)-long sorted list under (d, id)
            int pos = cnt < k + 1 ? cnt : k + 1;
            while (pos > 0 && (d < bd[pos - 1] || (d == bd[pos - 1] && id < bi[pos - 1]))) --pos;
            if (pos > k) continue;
            int last = cnt < k + 1 ? cnt : k;
            for (int s = last; s > pos; --s) { bd[s] = bd[s - 1]; bi[s] = bi[s - 1]; bj[s] = bj[s - 1]; }
            bd[pos] = d; bi[pos] = id; bj[pos] = (long)j;
            if (cnt < k + 1) ++cnt;
        }
        int found = cnt < k ? cnt : k;
        out_count[qi] = found;
        for (int s = 0; s < k; ++s) {
            bool ok = s < found;
            out_id[qi * k + s] = ok ? bi[s] : -1;
            out_sqd[qi * k + s] = ok ? bd[s] : INFINITY;
            for (int c = 0; c < 3; ++c) out_xyz[(qi * k + s) * 3 + c] = ok ? map_xyz[3 * bj[s] + c] : 0.f;
        }
        if (cnt == k + 1 && bd[k] == bd[k - 1]) ties += 1;
    }
    if (tie_straddles) *tie_straddles = ties;
}

// ---------------------------------------------------------------------------------------------
// a4 esti_plane<float> (include/common_lib.h:670-702):
//   A (5x3) rows = neighbours, b = -1;  normvec = A.colPivHouseholderQr().solve(b);
//   n = |normvec|; pabcd = (normvec/n, 1/n); reject if any |pabcd . (p,1)| > threshold.
// Eigen is absent, so the QR is a restatement of Eigen's ColPivHouseholderQR::computeInPlace /
// _solve_impl (column norms kept as norms, LAPACK-style norm downdate, nonzero-pivot threshold,
// Householder "essential part" convention), all in float, every 5-element reduction summed
// left-to-right, every 3-element one as x0 + (x1 + x2).  PARITY UNPINNED vs Eigen at the ulp level.
// ---------------------------------------------------------------------------------------------
namespace {
inline float norm_tail(const float* col, int from) {  // sqrt(sum_{i>=from} col[i]^2), 5 rows
    float s = 0.f;
    bool first = true;
    for (int i = from; i < 5; ++i) {
        float t = col[i] * col[i];
        if (first) { s = t; first = false; } else s = s + t;
    }
    return sqrtf(s);
}
}  // namespace

extern "C" int orc_esti_plane(const float* near5, float threshold, float pabcd[4]) {
    float A[3][5];  // column-major: A[c][r]
    for (int r = 0; r < 5; ++r)
        for (int c = 0; c < 3; ++c) A[c][r] = near5[3 * r + c];
    float cvec[5] = {-1.f, -1.f, -1.f, -1.f, -1.f};
    float hcoef[3] = {0.f, 0.f, 0.f};
    int perm[3] = {0, 1, 2};  // perm[i] = original column now at position i
    float nrmU[3], nrmD[3];
    for (int c = 0; c < 3; ++c) { nrmD[c] = norm_tail(A[c], 0); nrmU[c] = nrmD[c]; }
    float maxn = nrmU[0];
    if (nrmU[1] > maxn) maxn = nrmU[1];
    if (nrmU[2] > maxn) maxn = nrmU[2];
    const float eps = FLT_EPSILON;
    float th = maxn * eps;
    const float threshold_helper = (th * th) / 5.0f;
    const float downdate_thr = sqrtf(eps);
    int nonzero_pivots = 3;
    for (int k = 0; k < 3; ++k) {
        int big = k;
        for (int j = k + 1; j < 3; ++j)
            if (nrmU[j] > nrmU[big]) big = j;  // first maximum, as maxCoeff(&idx)
        float big_sq = nrmU[big] * nrmU[big];
        if (nonzero_pivots == 3 && big_sq < threshold_helper * (float)(5 - k)) nonzero_pivots = k;
        if (big != k) {
            for (int r = 0; r < 5; ++r) std::swap(A[k][r], A[big][r]);
            std::swap(nrmU[k], nrmU[big]);
            std::swap(nrmD[k], nrmD[big]);
            std::swap(perm[k], perm[big]);
        }
        // makeHouseholderInPlace on A[k][k..4]
        float tailSq = 0.f;
        {
            bool first = true;
            for (int i = k + 1; i < 5; ++i) {
                float t = A[k][i] * A[k][i];
                if (first) { tailSq = t; first = false; } else tailSq = tailSq + t;
            }
        }
        float c0 = A[k][k], beta, tau;
        if (tailSq <= FLT_MIN) {
            tau = 0.f; beta = c0;
            for (int i = k + 1; i < 5; ++i) A[k][i] = 0.f;
        } else {
            beta = sqrtf(c0 * c0 + tailSq);
            if (c0 >= 0.f) beta = -beta;
            float den = c0 - beta;
            for (int i = k + 1; i < 5; ++i) A[k][i] = A[k][i] / den;
            tau = (beta - c0) / beta;
        }
        A[k][k] = beta;
        hcoef[k] = tau;
        // applyHouseholderOnTheLeft to the trailing columns
        if (tau != 0.f) {
            for (int j = k + 1; j < 3; ++j) {
                float tmp = 0.f;
                bool first = true;
                for (int i = k + 1; i < 5; ++i) {
                    float t = A[k][i] * A[j][i];
                    if (first) { tmp = t; first = false; } else tmp = tmp + t;
                }
                tmp = tmp + A[j][k];
                A[j][k] = A[j][k] - tau * tmp;
                for (int i = k + 1; i < 5; ++i) A[j][i] = A[j][i] - (tau * A[k][i]) * tmp;
            }
        }
        // LAPACK-style column-norm downdate
        for (int j = k + 1; j < 3; ++j) {
            if (nrmU[j] != 0.f) {
                float temp = fabsf(A[j][k]) / nrmU[j];
                temp = (1.f + temp) * (1.f - temp);
                temp = temp < 0.f ? 0.f : temp;
                float ratio = nrmU[j] / nrmD[j];
                float temp2 = temp * (ratio * ratio);
                if (temp2 <= downdate_thr) {
                    nrmD[j] = norm_tail(A[j], k + 1);
                    nrmU[j] = nrmD[j];
                } else {
                    nrmU[j] = nrmU[j] * sqrtf(temp);
                }
            }
        }
    }
    // solve: c = Q^T b (first nonzero_pivots reflectors), back-substitute, un-permute
    float x[3] = {0.f, 0.f, 0.f};
    if (nonzero_pivots > 0) {
        for (int k = 0; k < nonzero_pivots; ++k) {
            float tau = hcoef[k];
            if (tau != 0.f) {
                float tmp = 0.f;
                bool first = true;
                for (int i = k + 1; i < 5; ++i) {
                    float t = A[k][i] * cvec[i];
                    if (first) { tmp = t; first = false; } else tmp = tmp + t;
                }
                tmp = tmp + cvec[k];
                cvec[k] = cvec[k] - tau * tmp;
                for (int i = k + 1; i < 5; ++i) cvec[i] = cvec[i] - (tau * A[k][i]) * tmp;
            }
        }
        for (int i = nonzero_pivots - 1; i >= 0; --i) {
            cvec[i] = cvec[i] / A[i][i];
            for (int j = 0; j < i; ++j) cvec[j] = cvec[j] - cvec[i] * A[i][j];
        }
        for (int i = 0; i < nonzero_pivots; ++i) x[perm[i]] = cvec[i];
    }
    float n = sqrtf(x[0] * x[0] + (x[1] * x[1] + x[2] * x[2]));
    pabcd[0] = x[0] / n;
    pabcd[1] = x[1] / n;
    pabcd[2] = x[2] / n;
    pabcd[3] = (float)(1.0 / (double)n);  // "1.0 / n" is evaluated in double, common_lib.h:691
    for (int j = 0; j < 5; ++j) {
        const float* p = near5 + 3 * j;
        float v = pabcd[0] * p[0] + pabcd[1] * p[1] + pabcd[2] * p[2] + pabcd[3];
        if (fabsf(v) > threshold) return 0;
    }
    return 1;
}

// ---------------------------------------------------------------------------------------------
// One h_share_model trip (front half: src/laser_mapping.cpp:503-561 == src/origin_laserMapping.cpp:928-1007;
// row: src/origin_laserMapping.cpp:1018-1044; Gram: esekfom.hpp:1781,1801).
// ---------------------------------------------------------------------------------------------
extern "C" void orc_h_share_model(const float* body_xyz, size_t n, size_t stride_floats, const orc_state* st,
                                  const float* near_xyz, const int32_t* near_cnt, const float* near_d5,
                                  int searched, int mode, uint8_t* mask_io, uint8_t* eff_out, float* normvec, float* world_xyz,
                                  double* HTH, double* HTz, int64_t* n_eff, double* res_sum, int64_t* borderline) {
    std::vector<float> res_last(n, 1000.f);
    int64_t border = 0;
#pragma omp parallel for schedule(static) reduction(+ : border)
    for (long i = 0; i < (long)n; ++i) {
        const float* pb = body_xyz + i * stride_floats;
        float pw[3];
        body_to_world_1(pb, st, pw);
        world_xyz[3 * i + 0] = pw[0]; world_xyz[3 * i + 1] = pw[1]; world_xyz[3 * i + 2] = pw[2];
        normvec[4 * i + 0] = normvec[4 * i + 1] = normvec[4 * i + 2] = normvec[4 * i + 3] = 0.f;
        bool sel;
        if (mode == ORC_MODE_V1) sel = true;           // vector<bool>(n, true) each call, :490
        else sel = mask_io[i] != 0;                    // global point_selected_surf[], origin:963
        bool have5 = near_cnt[i] >= 5;
        if (searched) {
            // point_selected_surf[i] = pointSearchSqDis[4] > 5 ? false : true  (:518 / origin:963);
            // fewer than 5 neighbours is UB in the reference (§3.3) and is defined here as "not selected".
            sel = have5 && !(near_d5[i] > 5.f);
            if (have5 && fabsf(near_d5[i] - 5.f) < 1e-5f) border += 1;
        }
        if (!sel || !have5) { mask_io[i] = 0; continue; }
        sel = false;
        float pabcd[4];
        if (orc_esti_plane(near_xyz + 15 * i, 0.1f, pabcd)) {
            float pd2 = pabcd[0] * pw[0] + pabcd[1] * pw[1] + pabcd[2] * pw[2] + pabcd[3];
            // float s = 1 - 0.9 * fabs(pd2) / sqrt(p_body.norm());   (double arithmetic, narrowed)
            double pbd[3] = {(double)pb[0], (double)pb[1], (double)pb[2]};
            double nrm = sqrt(pbd[0] * pbd[0] + (pbd[1] * pbd[1] + pbd[2] * pbd[2]));
            float s = (float)(1.0 - 0.9 * (double)fabsf(pd2) / sqrt(nrm));
            if (fabs((double)s - 0.9) < 1e-5) border += 1;
            if ((double)s > 0.9) {
                sel = true;
                normvec[4 * i + 0] = pabcd[0]; normvec[4 * i + 1] = pabcd[1];
                normvec[4 * i + 2] = pabcd[2]; normvec[4 * i + 3] = pd2;
                res_last[i] = fabsf(pd2);
            }
        }
        mask_io[i] = sel ? 1 : 0;
    }
    // compaction predicate (:552): point_selected_surf[i] && res_last[i] <= 2.0
    // Jacobian row + normal equations, serial, in index order (:569-593 / origin:1018-1044, esekfom.hpp:1781)
    for (int a = 0; a < 144; ++a) HTH[a] = 0.0;
    for (int a = 0; a < 12; ++a) HTz[a] = 0.0;
    int64_t ne = 0;
    double rs = 0.0;
    double RLIt_unused[1]; (void)RLIt_unused;
    for (size_t i = 0; i < n; ++i) {
        bool eff = mask_io[i] && (res_last[i] <= 2.0f);
        eff_out[i] = eff ? 1 : 0;
        if (!eff) continue;
        const float* pb = body_xyz + i * stride_floats;
        double p[3] = {(double)pb[0], (double)pb[1], (double)pb[2]};
        double pt[3];
        matvec3(st->offR, p, pt);
        pt[0] += st->offT[0]; pt[1] += st->offT[1]; pt[2] += st->offT[2];
        double nv[3] = {(double)normvec[4 * i], (double)normvec[4 * i + 1], (double)normvec[4 * i + 2]};
        double C[3], A[3], B[3], t[3];
        matTvec3(st->rot, nv, C);     // s.rot.conjugate() * norm_vec
        cross3(pt, C, A);             // point_crossmat * C
        matTvec3(st->offR, C, t);     // s.offset_R_L_I.conjugate() * C
        cross3(p, t, B);              // point_be_crossmat * (...)
        double h[12] = {nv[0], nv[1], nv[2], A[0], A[1], A[2], B[0], B[1], B[2], C[0], C[1], C[2]};
        double z = -(double)normvec[4 * i + 3];
        for (int a = 0; a < 12; ++a) {
            for (int b = 0; b < 12; ++b) HTH[a * 12 + b] += h[a] * h[b];
            HTz[a] += h[a] * z;
        }
        rs += (double)res_last[i];
        ne += 1;
    }
    *n_eff = ne;
    *res_sum = rs;
    if (borderline) *borderline = border;
}

// V1: Hsub = [A | n] (1x6), HPH.block<6,6> += Hsub^T (1/R) Hsub, HPL.block<6,1> += Hsub^T (1/R) err
// (src/laser_mapping.cpp:584-592).  V1 columns [0:3|3:6] = V3 columns [3:6|0:3].
extern "C" void orc_repack_v1(const double* HTH, const double* HTz, double laser_point_cov, double* HPH81, double* HPL9) {
    static const int map6[6] = {3, 4, 5, 0, 1, 2};
    for (int i = 0; i < 81; ++i) HPH81[i] = 0.0;
    for (int i = 0; i < 9; ++i) HPL9[i] = 0.0;
    double w = 1.0 / laser_point_cov;
    for (int a = 0; a < 6; ++a) {
        for (int b = 0; b < 6; ++b) HPH81[a * 9 + b] = HTH[map6[a] * 12 + map6[b]] * w;
        HPL9[a] = HTz[map6[a]] * w;
    }
}

// ---------------------------------------------------------------------------------------------
// Dense helpers for the solves: inverse by LU with partial pivoting (what Eigen's .inverse() uses for
// sizes > 4), restated.  PARITY UNPINNED vs Eigen at the ulp level; tolerance-checked in tests.
// ---------------------------------------------------------------------------------------------
namespace {

bool mat_inverse(const double* Ain, int n, double* Ainv) {
    std::vector<double> a(Ain, Ain + (size_t)n * n);
    std::vector<int> piv(n);
    for (int k = 0; k < n; ++k) {
        int p = k;
        double best = fabs(a[k * n + k]);
        for (int i = k + 1; i < n; ++i) {
            double v = fabs(a[i * n + k]);
            if (v > best) { best = v; p = i; }
        }
        piv[k] = p;
        if (best == 0.0) return false;
        if (p != k)
            for (int j = 0; j < n; ++j) std::swap(a[k * n + j], a[p * n + j]);
        double inv = 1.0 / a[k * n + k];
        for (int i = k + 1; i < n; ++i) a[i * n + k] *= inv;
        for (int i = k + 1; i < n; ++i) {
            double f = a[i * n + k];
            for (int j = k + 1; j < n; ++j) a[i * n + j] -= f * a[k * n + j];
        }
    }
    // solve A X = I column by column
    for (int c = 0; c < n; ++c) {
        std::vector<double> b(n, 0.0);
        b[c] = 1.0;
        for (int k = 0; k < n; ++k)
            if (piv[k] != k) std::swap(b[k], b[piv[k]]);
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < i; ++j) b[i] -= a[i * n + j] * b[j];
        for (int i = n - 1; i >= 0; --i) {
            for (int j = i + 1; j < n; ++j) b[i] -= a[i * n + j] * b[j];
            b[i] /= a[i * n + i];
        }
        for (int i = 0; i < n; ++i) Ainv[i * n + c] = b[i];
    }
    return true;
}

// ---- quaternion / manifold helpers (IKFoM) ---------------------------------------------------
struct Quat { double w, x, y, z; };

Quat quat_from_R(const double* R) {  // Eigen::Quaternion(matrix) algorithm (Shepperd), row-major R
    Quat q;
    double t = R[0] + R[4] + R[8];
    if (t > 0.0) {
        t = sqrt(t + 1.0);
        q.w = 0.5 * t;
        t = 0.5 / t;
        q.x = (R[7] - R[5]) * t; q.y = (R[2] - R[6]) * t; q.z = (R[3] - R[1]) * t;
    } else {
        int i = 0;
        if (R[4] > R[0]) i = 1;
        if (R[8] > R[i * 3 + i]) i = 2;
        int j = (i + 1) % 3, k = (j + 1) % 3;
        t = sqrt(R[i * 3 + i] - R[j * 3 + j] - R[k * 3 + k] + 1.0);
        double v[3];
        v[i] = 0.5 * t;
        t = 0.5 / t;
        q.w = (R[k * 3 + j] - R[j * 3 + k]) * t;
        v[j] = (R[j * 3 + i] + R[i * 3 + j]) * t;
        v[k] = (R[k * 3 + i] + R[i * 3 + k]) * t;
        q.x = v[0]; q.y = v[1]; q.z = v[2];
    }
    return q;
}
void quat_to_R(const Quat& q, double* R) {  // Eigen toRotationMatrix
    double tx = 2.0 * q.x, ty = 2.0 * q.y, tz = 2.0 * q.z;
    double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
    double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
    double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
    R[0] = 1.0 - (tyy + tzz); R[1] = txy - twz;         R[2] = txz + twy;
    R[3] = txy + twz;         R[4] = 1.0 - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy;         R[7] = tyz + twx;         R[8] = 1.0 - (txx + tyy);
}
Quat quat_mul(const Quat& a, const Quat& b) {
    Quat r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    return r;
}
Quat quat_conj(const Quat& a) { return Quat{a.w, -a.x, -a.y, -a.z}; }

// MTK::cos_sinc_sqrt, mtk/src/mtkmath.hpp:139-171
void cos_sinc_sqrt(double x2, double* c, double* s) {
    const double taylor_0_bound = DBL_EPSILON;
    const double taylor_2_bound = sqrt(taylor_0_bound);
    const double taylor_n_bound = sqrt(taylor_2_bound);
    if (x2 >= taylor_n_bound) {
        double x = sqrt(x2);
        *c = cos(x); *s = sin(x) / x;
        return;
    }
    static const double inv[] = {1 / 3., 1 / 4., 1 / 5., 1 / 6., 1 / 7., 1 / 8., 1 / 9.};
    double cosi = 1., sinc = 1;
    double term = -1 / 2. * x2;
    for (int i = 0; i < 3; ++i) {
        cosi += term;
        term *= inv[2 * i];
        sinc += term;
        term *= -inv[2 * i + 1] * x2;
    }
    *c = cosi; *s = sinc;
}
// MTK::exp, mtkmath.hpp:249-256, as used by SO3::exp (SOn.hpp:293-297) with scale/2
Quat so3_exp(const double* v, double scale_half) {
    double n2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    double c, s;
    cos_sinc_sqrt(scale_half * scale_half * n2, &c, &s);
    double mult = s * scale_half;
    return Quat{c, mult * v[0], mult * v[1], mult * v[2]};
}
// SO3::log -> MTK::log(res, w, vec, scale=2, plus_minus_periodicity=true), mtkmath.hpp:268-290
void so3_log(const Quat& q, double* res) {
    double nv = sqrt(q.x * q.x + q.y * q.y + q.z * q.z);
    if (nv < 1e-11) nv = 1e-11;
    double s = 2.0 / nv * atan(nv / q.w);
    res[0] = s * q.x; res[1] = s * q.y; res[2] = s * q.z;
}
// MTK::A_matrix, mtkmath.hpp:236-247 (row-major out)
void A_matrix(const double* v, double* A) {
    double sq = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    double nrm = sqrt(sq);
    for (int i = 0; i < 9; ++i) A[i] = 0.0;
    A[0] = A[4] = A[8] = 1.0;
    if (nrm < 1e-11) return;
    double H[9] = {0, -v[2], v[1], v[2], 0, -v[0], -v[1], v[0], 0};
    double H2[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) H2[i * 3 + j] = H[i * 3] * H[j] + H[i * 3 + 1] * H[3 + j] + H[i * 3 + 2] * H[6 + j];
    double a = (1 - cos(nrm)) / sq, b = (1 - sin(nrm) / nrm) / sq;
    for (int i = 0; i < 9; ++i) A[i] += a * H[i] + b * H2[i];
}
const double S2_LEN = 98090.0 / 10000.0;  // MTK::S2<double, 98090, 10000, 1>, use-ikfom.hpp:8
// S2_Bx for S2_typ == 1 (mtk/types/S2.hpp:213-231); Bx is 3x2 row-major
void s2_Bx(const double* vec, double* Bx) {
    const double len = S2_LEN;
    if (vec[0] + len > 1e-11) {
        double d = len + vec[0];
        Bx[0] = -vec[1];                   Bx[1] = -vec[2];
        Bx[2] = len - vec[1] * vec[1] / d; Bx[3] = -vec[2] * vec[1] / d;
        Bx[4] = -vec[2] * vec[1] / d;      Bx[5] = len - vec[2] * vec[2] / d;
        for (int i = 0; i < 6; ++i) Bx[i] /= len;
    } else {
        for (int i = 0; i < 6; ++i) Bx[i] = 0.0;
        Bx[1 * 2 + 1] = -1; Bx[2 * 2 + 0] = 1;
    }
}
void hat3(const double* v, double* H) {
    H[0] = 0; H[1] = -v[2]; H[2] = v[1]; H[3] = v[2]; H[4] = 0; H[5] = -v[0]; H[6] = -v[1]; H[7] = v[0]; H[8] = 0;
}
// S2::boxminus, S2.hpp:136-160
void s2_boxminus(const double* vec, const double* other, double* res2) {
    double cr[3];
    cross3(vec, other, cr);
    double v_sin = sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
    double v_cos = vec[0] * other[0] + vec[1] * other[1] + vec[2] * other[2];
    double theta = atan2(v_sin, v_cos);
    if (v_sin < 1e-11) {
        if (fabs(theta) > 1e-11) { res2[0] = 3.1415926; res2[1] = 0; }
        else { res2[0] = 0; res2[1] = 0; }
    } else {
        double Bx[6];
        s2_Bx(other, Bx);
        double hv[3];
        cross3(other, vec, hv);  // hat(other.vec) * vec
        double f = theta / v_sin;
        for (int c = 0; c < 2; ++c) res2[c] = f * (Bx[0 * 2 + c] * hv[0] + Bx[1 * 2 + c] * hv[1] + Bx[2 * 2 + c] * hv[2]);
    }
}
// S2::boxplus, S2.hpp:128-134
void s2_boxplus(double* vec, const double* delta2) {
    double Bx[6];
    s2_Bx(vec, Bx);
    double Bu[3];
    for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
    Quat q = so3_exp(Bu, 0.5);
    double R[9], o[3];
    quat_to_R(q, R);
    matvec3(R, vec, o);
    vec[0] = o[0]; vec[1] = o[1]; vec[2] = o[2];
}
// S2_Nx_yy, S2.hpp:259-264: 1/len/len * Bx^T * hat(vec)   (2x3 row-major)
void s2_Nx_yy(const double* vec, double* Nx) {
    double Bx[6], H[9];
    s2_Bx(vec, Bx);
    hat3(vec, H);
    for (int r = 0; r < 2; ++r)
        for (int c = 0; c < 3; ++c)
            Nx[r * 3 + c] = 1 / S2_LEN / S2_LEN * (Bx[0 * 2 + r] * H[0 * 3 + c] + Bx[1 * 2 + r] * H[1 * 3 + c] + Bx[2 * 2 + r] * H[2 * 3 + c]);
}
// S2_Mx, S2.hpp:266-280.  NB "scalar(1/2)" there is integer division == 0, so exp_delta is the
// identity rotation; reproduced.
void s2_Mx(const double* vec, const double* delta2, double* Mx /*3x2*/) {
    double Bx[6], H[9];
    s2_Bx(vec, Bx);
    hat3(vec, H);
    double dn = sqrt(delta2[0] * delta2[0] + delta2[1] * delta2[1]);
    if (dn < 1e-11) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 2; ++c)
                Mx[r * 2 + c] = -(H[r * 3] * Bx[c] + H[r * 3 + 1] * Bx[2 + c] + H[r * 3 + 2] * Bx[4 + c]);
    } else {
        double Bu[3];
        for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
        Quat q = so3_exp(Bu, 0.0);
        double Rm[9], Am[9];
        quat_to_R(q, Rm);
        A_matrix(Bu, Am);
        // -R * hat(vec) * A^T * Bx
        double T1[9], T2[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) T1[i * 3 + j] = Rm[i * 3] * H[j] + Rm[i * 3 + 1] * H[3 + j] + Rm[i * 3 + 2] * H[6 + j];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) T2[i * 3 + j] = T1[i * 3] * Am[j * 3] + T1[i * 3 + 1] * Am[j * 3 + 1] + T1[i * 3 + 2] * Am[j * 3 + 2];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 2; ++c)
                Mx[r * 2 + c] = -(T2[r * 3] * Bx[c] + T2[r * 3 + 1] * Bx[2 + c] + T2[r * 3 + 2] * Bx[4 + c]);
    }
}

const int NV3 = 23;
// rows idx..idx+d-1 of M (n x n) <- J (d x d) * rows
void rot_rows(double* M, int n, int idx, int d, const double* J) {
    for (int c = 0; c < n; ++c) {
        double t[3];
        for (int r = 0; r < d; ++r) {
            double s = 0;
            for (int k = 0; k < d; ++k) s += J[r * d + k] * M[(idx + k) * n + c];
            t[r] = s;
        }
        for (int r = 0; r < d; ++r) M[(idx + r) * n + c] = t[r];
    }
}
// cols idx.. of M <- cols * J^T
void rot_cols(double* M, int n, int idx, int d, const double* J) {
    for (int r = 0; r < n; ++r) {
        double t[3];
        for (int c = 0; c < d; ++c) {
            double s = 0;
            for (int k = 0; k < d; ++k) s += M[r * n + idx + k] * J[c * d + k];
            t[c] = s;
        }
        for (int c = 0; c < d; ++c) M[r * n + idx + c] = t[c];
    }
}
// x (-) y for state_ikfom: boxminus per sub-manifold in declaration order (build_manifold.hpp:198-200)
void state_boxminus(const orc_state* x, const orc_state* y, double* d /*23*/) {
    for (int i = 0; i < 3; ++i) d[i] = x->pos[i] - y->pos[i];
    { Quat a = quat_from_R(x->rot), b = quat_from_R(y->rot); so3_log(quat_mul(quat_conj(b), a), d + 3); }
    { Quat a = quat_from_R(x->offR), b = quat_from_R(y->offR); so3_log(quat_mul(quat_conj(b), a), d + 6); }
    for (int i = 0; i < 3; ++i) d[9 + i] = x->offT[i] - y->offT[i];
    for (int i = 0; i < 3; ++i) d[12 + i] = x->vel[i] - y->vel[i];
    for (int i = 0; i < 3; ++i) d[15 + i] = x->bg[i] - y->bg[i];
    for (int i = 0; i < 3; ++i) d[18 + i] = x->ba[i] - y->ba[i];
    s2_boxminus(x->grav, y->grav, d + 21);
}
void state_boxplus(orc_state* x, const double* d) {
    for (int i = 0; i < 3; ++i) x->pos[i] += d[i];
    { Quat q = quat_mul(quat_from_R(x->rot), so3_exp(d + 3, 0.5)); quat_to_R(q, x->rot); }
    { Quat q = quat_mul(quat_from_R(x->offR), so3_exp(d + 6, 0.5)); quat_to_R(q, x->offR); }
    for (int i = 0; i < 3; ++i) x->offT[i] += d[9 + i];
    for (int i = 0; i < 3; ++i) x->vel[i] += d[12 + i];
    for (int i = 0; i < 3; ++i) x->bg[i] += d[15 + i];
    for (int i = 0; i < 3; ++i) x->ba[i] += d[18 + i];
    s2_boxplus(x->grav, d + 21);
}

}  // namespace

// esekfom.hpp:1633-1927, the dof_Measurement >= n branch (:1779-1806).  With fewer than 23 effective
// points the reference takes the algebraically identical gain form (:1716-1739, matrix inversion
// lemma); this restatement always uses the information form.
// NOTE the reference keeps rot / offset_R_L_I as unit quaternions; the flat state carries rotation
// matrices, converted on entry/exit (identical up to double rounding).
extern "C" void orc_solve_v3(orc_state* x, const orc_state* x_prop, const double* P_prop, const double* HTH,
                             const double* HTz, double R, int iter, int max_iter, const double* limit23,
                             int* t_io, int* converge_out, int* stopped, double* dx_out23, double* P_out) {
    const int n = NV3;
    double dx[23], dx_new[23];
    state_boxminus(x, x_prop, dx);
    memcpy(dx_new, dx, sizeof(dx));
    std::vector<double> P(P_prop, P_prop + n * n);
    const int so3_idx[2] = {3, 6};
    for (int s = 0; s < 2; ++s) {
        int idx = so3_idx[s];
        double A[9], J[9];
        A_matrix(dx + idx, A);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) J[i * 3 + j] = A[j * 3 + i];
        double t[3];
        for (int r = 0; r < 3; ++r) t[r] = J[r * 3] * dx_new[idx] + J[r * 3 + 1] * dx_new[idx + 1] + J[r * 3 + 2] * dx_new[idx + 2];
        for (int r = 0; r < 3; ++r) dx_new[idx + r] = t[r];
        rot_rows(P.data(), n, idx, 3, J);
        rot_cols(P.data(), n, idx, 3, J);
    }
    {
        int idx = 21;
        double Nx[6], Mx[6], J2[4];
        s2_Nx_yy(x->grav, Nx);
        s2_Mx(x_prop->grav, dx + idx, Mx);
        for (int r = 0; r < 2; ++r)
            for (int c = 0; c < 2; ++c) J2[r * 2 + c] = Nx[r * 3] * Mx[c] + Nx[r * 3 + 1] * Mx[2 + c] + Nx[r * 3 + 2] * Mx[4 + c];
        double t0 = J2[0] * dx_new[idx] + J2[1] * dx_new[idx + 1];
        double t1 = J2[2] * dx_new[idx] + J2[3] * dx_new[idx + 1];
        dx_new[idx] = t0; dx_new[idx + 1] = t1;
        rot_rows(P.data(), n, idx, 2, J2);
        rot_cols(P.data(), n, idx, 2, J2);
    }
    std::vector<double> PR(n * n), P_temp(n * n), P_inv(n * n);
    for (int i = 0; i < n * n; ++i) PR[i] = P[i] / R;
    mat_inverse(PR.data(), n, P_temp.data());
    for (int a = 0; a < 12; ++a)
        for (int b = 0; b < 12; ++b) P_temp[a * n + b] += HTH[a * 12 + b];
    mat_inverse(P_temp.data(), n, P_inv.data());
    double K_h[23];
    std::vector<double> K_x(n * n, 0.0);
    for (int i = 0; i < n; ++i) {
        double s = 0;
        for (int a = 0; a < 12; ++a) s += P_inv[i * n + a] * HTz[a];
        K_h[i] = s;
        for (int b = 0; b < 12; ++b) {
            double v = 0;
            for (int a = 0; a < 12; ++a) v += P_inv[i * n + a] * HTH[a * 12 + b];
            K_x[i * n + b] = v;
        }
    }
    double dx_[23];
    for (int i = 0; i < n; ++i) {
        double s = K_h[i];
        for (int j = 0; j < n; ++j) s += (K_x[i * n + j] - (i == j ? 1.0 : 0.0)) * dx_new[j];
        dx_[i] = s;
    }
    state_boxplus(x, dx_);
    int converge = 1;
    for (int i = 0; i < n; ++i)
        if (fabs(dx_[i]) > limit23[i]) { converge = 0; break; }
    int t = *t_io;
    if (converge) t++;
    if (!t && iter == max_iter - 2) converge = 1;
    *t_io = t;
    *converge_out = converge;
    if (dx_out23) memcpy(dx_out23, dx_, sizeof(dx_));
    *stopped = 0;
    if (t > 1 || iter == max_iter - 1) {
        std::vector<double> L(P);
        for (int s = 0; s < 2; ++s) {
            int idx = so3_idx[s];
            double A[9], J[9];
            A_matrix(dx_ + idx, A);
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) J[i * 3 + j] = A[j * 3 + i];
            // L rows(idx) = J * P rows(idx)  (P as it is now, :1839-1841)
            for (int c = 0; c < n; ++c)
                for (int r = 0; r < 3; ++r)
                    L[(idx + r) * n + c] = J[r * 3] * P[idx * n + c] + J[r * 3 + 1] * P[(idx + 1) * n + c] + J[r * 3 + 2] * P[(idx + 2) * n + c];
            for (int c = 0; c < 12; ++c) {
                double t3[3];
                for (int r = 0; r < 3; ++r)
                    t3[r] = J[r * 3] * K_x[idx * n + c] + J[r * 3 + 1] * K_x[(idx + 1) * n + c] + J[r * 3 + 2] * K_x[(idx + 2) * n + c];
                for (int r = 0; r < 3; ++r) K_x[(idx + r) * n + c] = t3[r];
            }
            rot_cols(L.data(), n, idx, 3, J);
            rot_cols(P.data(), n, idx, 3, J);
        }
        {
            int idx = 21;
            double Nx[6], Mx[6], J2[4];
            s2_Nx_yy(x->grav, Nx);
            s2_Mx(x_prop->grav, dx_ + idx, Mx);
            for (int r = 0; r < 2; ++r)
                for (int c = 0; c < 2; ++c) J2[r * 2 + c] = Nx[r * 3] * Mx[c] + Nx[r * 3 + 1] * Mx[2 + c] + Nx[r * 3 + 2] * Mx[4 + c];
            for (int c = 0; c < n; ++c)
                for (int r = 0; r < 2; ++r) L[(idx + r) * n + c] = J2[r * 2] * P[idx * n + c] + J2[r * 2 + 1] * P[(idx + 1) * n + c];
            for (int c = 0; c < 12; ++c) {
                double a0 = J2[0] * K_x[idx * n + c] + J2[1] * K_x[(idx + 1) * n + c];
                double a1 = J2[2] * K_x[idx * n + c] + J2[3] * K_x[(idx + 1) * n + c];
                K_x[idx * n + c] = a0; K_x[(idx + 1) * n + c] = a1;
            }
            rot_cols(L.data(), n, idx, 2, J2);
            rot_cols(P.data(), n, idx, 2, J2);
        }
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                double s = 0;
                for (int a = 0; a < 12; ++a) s += K_x[i * n + a] * P[a * n + j];
                P_out[i * n + j] = L[i * n + j] - s;
            }
        *stopped = 1;
    }
}

// ---------------------------------------------------------------------------------------------
// V1 loop body, src/laser_mapping.cpp:178-238; StatesGroup algebra include/common_lib.h:552-587,
// Exp / Log include/so3_math.h:62-91.  18-DOF order: rot, pos, vel, bg, ba, grav.
// ---------------------------------------------------------------------------------------------
namespace {
void so3_Exp_v1(double v1, double v2, double v3, double* R) {  // so3_math.h:62-82
    double nrm = sqrt(v1 * v1 + v2 * v2 + v3 * v3);
    for (int i = 0; i < 9; ++i) R[i] = 0;
    R[0] = R[4] = R[8] = 1;
    if (nrm > 0.00001) {
        double r[3] = {v1 / nrm, v2 / nrm, v3 / nrm};
        double K[9] = {0, -r[2], r[1], r[2], 0, -r[0], -r[1], r[0], 0}, K2[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) K2[i * 3 + j] = K[i * 3] * K[j] + K[i * 3 + 1] * K[3 + j] + K[i * 3 + 2] * K[6 + j];
        double s = sin(nrm), c = 1.0 - cos(nrm);
        for (int i = 0; i < 9; ++i) R[i] += s * K[i] + c * K2[i];
    }
}
void so3_Log_v1(const double* R, double* o) {  // so3_math.h:85-91
    double tr = R[0] + R[4] + R[8];
    double theta = (tr > 3.0 - 1e-6) ? 0.0 : acos(0.5 * (tr - 1));
    double K[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    double f = (fabs(theta) < 0.001) ? 0.5 : (0.5 * theta / sin(theta));
    for (int i = 0; i < 3; ++i) o[i] = f * K[i];
}
}  // namespace

extern "C" void orc_solve_v1(orc_state* x, const orc_state* x_prop, double* cov, const double* HPH81,
                             const double* HPL9, int iterCount, int max_iter, int* rematch_io, int* search_next,
                             int* stopped, double* solution18) {
    const int n = 18;
    std::vector<double> HTH(n * n, 0.0), cinv(n * n), S(n * n), K1(n * n), G(n * n, 0.0);
    for (int a = 0; a < 9; ++a)
        for (int b = 0; b < 9; ++b) HTH[a * n + b] = HPH81[a * 9 + b];
    mat_inverse(cov, n, cinv.data());
    for (int i = 0; i < n * n; ++i) S[i] = HTH[i] + cinv[i];
    mat_inverse(S.data(), n, K1.data());
    for (int i = 0; i < n; ++i)
        for (int b = 0; b < 9; ++b) {
            double s = 0;
            for (int a = 0; a < 9; ++a) s += K1[i * n + a] * HTH[a * n + b];
            G[i * n + b] = s;
        }
    // vec = state_propagat - state   (common_lib.h:576-587)
    double vec[18];
    {
        double rotd[9];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                rotd[i * 3 + j] = x->rot[0 * 3 + i] * x_prop->rot[0 * 3 + j] + x->rot[1 * 3 + i] * x_prop->rot[1 * 3 + j] + x->rot[2 * 3 + i] * x_prop->rot[2 * 3 + j];
        so3_Log_v1(rotd, vec);
        for (int i = 0; i < 3; ++i) {
            vec[3 + i] = x_prop->pos[i] - x->pos[i];
            vec[6 + i] = x_prop->vel[i] - x->vel[i];
            vec[9 + i] = x_prop->bg[i] - x->bg[i];
            vec[12 + i] = x_prop->ba[i] - x->ba[i];
            vec[15 + i] = x_prop->grav[i] - x->grav[i];
        }
    }
    double sol[18];
    for (int i = 0; i < n; ++i) {
        double s1 = 0, s2 = 0;
        for (int a = 0; a < 9; ++a) { s1 += K1[i * n + a] * HPL9[a]; s2 += G[i * n + a] * vec[a]; }
        sol[i] = s1 + vec[i] - s2;
    }
    // state += solution (common_lib.h:565-574)
    {
        double E[9], Rn[9];
        so3_Exp_v1(sol[0], sol[1], sol[2], E);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Rn[i * 3 + j] = x->rot[i * 3] * E[j] + x->rot[i * 3 + 1] * E[3 + j] + x->rot[i * 3 + 2] * E[6 + j];
        memcpy(x->rot, Rn, sizeof(Rn));
        for (int i = 0; i < 3; ++i) {
            x->pos[i] += sol[3 + i]; x->vel[i] += sol[6 + i]; x->bg[i] += sol[9 + i];
            x->ba[i] += sol[12 + i]; x->grav[i] += sol[15 + i];
        }
    }
    if (solution18) memcpy(solution18, sol, sizeof(sol));
    const double R2D = 180.0 / 3.14159265358979323846;  // so3/common_lib PI_M
    double rn = sqrt(sol[0] * sol[0] + sol[1] * sol[1] + sol[2] * sol[2]);
    double tn = sqrt(sol[3] * sol[3] + sol[4] * sol[4] + sol[5] * sol[5]);
    bool converged = (rn * R2D < 0.01) && (tn * 100 < 0.015);
    int rematch = *rematch_io;
    int nse = 0;
    if (converged || ((rematch == 0) && (iterCount == (max_iter - 2)))) { nse = 1; rematch++; }
    *rematch_io = rematch;
    *search_next = nse;
    *stopped = 0;
    if (rematch >= 2 || (iterCount == max_iter - 1)) {
        std::vector<double> nc(n * n);
        for (int i = 0; i < n; ++i)
            for (int j = 0; j < n; ++j) {
                double s = 0;
                for (int k = 0; k < n; ++k) s += ((i == k ? 1.0 : 0.0) - G[i * n + k]) * cov[k * n + j];
                nc[i * n + j] = s;
            }
        memcpy(cov, nc.data(), sizeof(double) * n * n);
        *stopped = 1;
    }
}

// ---------------------------------------------------------------------------------------------
// a12/a13: Add_Points / Delete_Point_Boxes over a flat list with a voxel index.
// Box of a point: vertex_min = floor(p/ds)*ds, vertex_max = vertex_min + ds (float, ikd_Tree.cpp:392-397);
// mid = min + (max-min)/2.0 (double arithmetic narrowed to float, :398-400);
// Search_by_range membership: min <= q < max per axis (:1001).
// ---------------------------------------------------------------------------------------------
struct orc_map {
    float ds;
    std::vector<float> xyz;
    std::vector<int32_t> id;
    std::vector<uint8_t> alive;
    int32_t next_id = 0;
    // voxel index over integer keys floor(q/ds); box membership is re-tested with float compares, and the
    // 27 surrounding keys are visited because float rounding lets a point sit in a neighbouring key's box.
    struct KeyHash {
        size_t operator()(const std::array<int32_t, 3>& k) const {
            uint64_t h = (uint64_t)(uint32_t)k[0] * 0x9E3779B97F4A7C15ull;
            h ^= ((uint64_t)(uint32_t)k[1] + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full + (h << 6) + (h >> 2);
            h ^= ((uint64_t)(uint32_t)k[2] + 0x165667B1ull) * 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
            return (size_t)h;
        }
    };
    std::unordered_map<std::array<int32_t, 3>, std::vector<int32_t>, KeyHash> vox;  // key -> slots (alive or dead)
    std::array<int32_t, 3> key_of(const float* p) const {
        return {(int32_t)floorf(p[0] / ds), (int32_t)floorf(p[1] / ds), (int32_t)floorf(p[2] / ds)};
    }
    int32_t push(const float* p, int32_t pid) {
        int32_t slot = (int32_t)id.size();
        xyz.push_back(p[0]); xyz.push_back(p[1]); xyz.push_back(p[2]);
        id.push_back(pid);
        alive.push_back(1);
        vox[key_of(p)].push_back(slot);
        return slot;
    }
};

extern "C" orc_map* orc_map_create(float downsample_size) {
    orc_map* m = new orc_map();
    m->ds = downsample_size;
    return m;
}
extern "C" void orc_map_destroy(orc_map* m) { delete m; }
extern "C" void orc_map_build(orc_map* m, const float* xyz, size_t n) {
    m->xyz.clear(); m->id.clear(); m->alive.clear(); m->vox.clear();
    m->next_id = 0;
    for (size_t i = 0; i < n; ++i) m->push(xyz + 3 * i, m->next_id++);
}
extern "C" int orc_map_add_points(orc_map* m, const float* xyz, size_t n, int downsample_on) {
    int counter = 0;
    const float ds = m->ds;
    for (size_t i = 0; i < n; ++i) {
        const float* p = xyz + 3 * i;
        int32_t pid = m->next_id++;
        if (!downsample_on) { m->push(p, pid); continue; }
        float bmin[3], bmax[3], mid[3];
        for (int c = 0; c < 3; ++c) {
            bmin[c] = floorf(p[c] / ds) * ds;
            bmax[c] = bmin[c] + ds;
            mid[c] = (float)((double)bmin[c] + (double)(bmax[c] - bmin[c]) / 2.0);
        }
        std::array<int32_t, 3> k = m->key_of(p);
        std::vector<int32_t> storage;  // live slots inside the box
        for (int dx = -1; dx <= 1; ++dx)
            for (int dy = -1; dy <= 1; ++dy)
                for (int dz = -1; dz <= 1; ++dz) {
                    auto it = m->vox.find({k[0] + dx, k[1] + dy, k[2] + dz});
                    if (it == m->vox.end()) continue;
                    for (int32_t s : it->second) {
                        if (!m->alive[s]) continue;
                        const float* q = &m->xyz[3 * s];
                        if (bmin[0] <= q[0] && bmax[0] > q[0] && bmin[1] <= q[1] && bmax[1] > q[1] && bmin[2] <= q[2] && bmax[2] > q[2])
                            storage.push_back(s);
                    }
                }
        // ties inside Downsample_Storage resolve by tree traversal order in the reference; canonical here: lowest id first
        std::sort(storage.begin(), storage.end(), [&](int32_t a, int32_t b) { return m->id[a] < m->id[b]; });
        float min_dist = sqdist(p, mid);
        int32_t best = -1;  // -1 == the new point
        for (int32_t s : storage) {
            float d = sqdist(&m->xyz[3 * s], mid);
            if (d < min_dist) { min_dist = d; best = s; }
        }
        const float* w = best < 0 ? p : &m->xyz[3 * best];
        bool same = fabsf(p[0] - w[0]) < 1e-6 && fabsf(p[1] - w[1]) < 1e-6 && fabsf(p[2] - w[2]) < 1e-6;  // EPSS, :1287-1289
        if (storage.size() > 1 || same) {
            float wv[3] = {w[0], w[1], w[2]};
            int32_t wid = best < 0 ? pid : m->id[best];
            for (int32_t s : storage) m->alive[s] = 0;
            m->push(wv, wid);
            counter++;
        }
    }
    return counter;
}
extern "C" int orc_map_delete_boxes(orc_map* m, const float* boxes6, size_t nboxes) {
    int counter = 0;
    for (size_t b = 0; b < nboxes; ++b) {
        const float* mn = boxes6 + 6 * b;
        const float* mx = mn + 3;
        for (size_t s = 0; s < m->id.size(); ++s) {
            if (!m->alive[s]) continue;
            const float* q = &m->xyz[3 * s];
            if (mn[0] <= q[0] && mx[0] > q[0] && mn[1] <= q[1] && mx[1] > q[1] && mn[2] <= q[2] && mx[2] > q[2]) {
                m->alive[s] = 0;
                counter++;
            }
        }
    }
    return counter;
}
extern "C" size_t orc_map_live(orc_map* m, float* xyz, int32_t* ids, size_t cap) {
    size_t c = 0;
    for (size_t s = 0; s < m->id.size(); ++s) {
        if (!m->alive[s]) continue;
        if (c < cap) {
            if (xyz) { xyz[3 * c] = m->xyz[3 * s]; xyz[3 * c + 1] = m->xyz[3 * s + 1]; xyz[3 * c + 2] = m->xyz[3 * s + 2]; }
            if (ids) ids[c] = m->id[s];
        }
        ++c;
    }
    return c;
}

// =====================================================================================================================
// Front end: motion compensation, voxel-grid down-sampling, local-map sliding
// =====================================================================================================================
namespace {
// Exp(ang_vel, dt), include/so3_math.h:30-52; row-major 3x3
void exp_rate(const double w[3], double dt, double R[9]) {
    const double nrm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    for (int e = 0; e < 9; ++e) R[e] = (e % 4 == 0) ? 1.0 : 0.0;
    if (nrm > 0.0000001) {
        const double a[3] = {w[0] / nrm, w[1] / nrm, w[2] / nrm};
        const double K[9] = {0.0, -a[2], a[1], a[2], 0.0, -a[0], -a[1], a[0], 0.0};
        const double ang = nrm * dt;
        const double s = sin(ang), c1 = 1.0 - cos(ang);
        double cK[9], KK[9];
        for (int e = 0; e < 9; ++e) cK[e] = c1 * K[e];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) KK[i * 3 + j] = cK[i * 3] * K[j] + cK[i * 3 + 1] * K[3 + j] + cK[i * 3 + 2] * K[6 + j];
        for (int e = 0; e < 9; ++e) R[e] = (R[e] + s * K[e]) + KK[e];
    }
}
void mv3(const double* M, const double v[3], double o[3]) {
    for (int r = 0; r < 3; ++r) o[r] = M[r * 3] * v[0] + M[r * 3 + 1] * v[1] + M[r * 3 + 2] * v[2];
}
}  // namespace

// The backward loop of ImuProcess::UndistortPcl, src/IMU_Processing.cpp:336-376, restated literally (two nested
// loops walking the scan from its last point, `break` at the first point).  pts: n records of stride_floats floats,
// xyz first, modified in place; t_ms[i] = curvature of point i.
extern "C" void orc_undistort(float* pts, size_t n, size_t stride_floats, const float* t_ms, const orc_pose6d* IMUpose, int n_poses,
                              const double rot_end[9], const double pos_end[3], const double R_LI[9], const double t_LI[3]) {
    double extR_Ri[9], exrR_extT[3];
    for (int r = 0; r < 3; ++r) {   // Lid_rot_to_IMU^T * rot_end^T ; Lid_rot_to_IMU^T * Lid_offset_to_IMU  (:336-337)
        for (int c = 0; c < 3; ++c) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += R_LI[k * 3 + r] * rot_end[c * 3 + k];
            extR_Ri[r * 3 + c] = s;
        }
        double t = 0.0;
        for (int k = 0; k < 3; ++k) t += R_LI[k * 3 + r] * t_LI[k];
        exrR_extT[r] = t;
    }
    if (n < 1) return;                                            // :339
    long it_pcl = (long)n - 1;                                    // :342
    for (int it_kp = n_poses - 1; it_kp != 0; it_kp--) {          // :343
        const orc_pose6d* head = &IMUpose[it_kp - 1];
        for (; (double)t_ms[it_pcl] / double(1000) > head->offset_time; it_pcl--) {   // :354
            const double dt = (double)t_ms[it_pcl] / double(1000) - head->offset_time;
            double E[9], R_i[9], T_ei[3];
            exp_rate(head->gyr, dt, E);
            for (int r = 0; r < 3; ++r)
                for (int c = 0; c < 3; ++c) R_i[r * 3 + c] = head->rot[r * 3] * E[c] + head->rot[r * 3 + 1] * E[3 + c] + head->rot[r * 3 + 2] * E[6 + c];
            for (int a = 0; a < 3; ++a) T_ei[a] = ((head->pos[a] + head->vel[a] * dt) + 0.5 * head->acc[a] * dt * dt) - pos_end[a];
            float* p = pts + (size_t)it_pcl * stride_floats;
            const double P_i[3] = {(double)p[0], (double)p[1], (double)p[2]};
            double a3[3], b3[3], c3[3];
            mv3(R_LI, P_i, a3);
            for (int a = 0; a < 3; ++a) a3[a] += t_LI[a];
            mv3(R_i, a3, b3);
            for (int a = 0; a < 3; ++a) b3[a] += T_ei[a];
            mv3(extR_Ri, b3, c3);
            p[0] = (float)(c3[0] - exrR_extT[0]);
            p[1] = (float)(c3[1] - exrR_extT[1]);
            p[2] = (float)(c3[2] - exrR_extT[2]);
            if (it_pcl == 0) break;                               // :373
        }
    }
}

// pcl::VoxelGrid<PointXYZINormal>::applyFilter as called at src/laser_mapping.cpp:129-130 (PCL is not in the reference tree:
// restated from pcl/filters/impl/voxel_grid.hpp; downsample_all_data = true, min_points_per_voxel = 0, no field filter).
// in: n x (x, y, z, intensity) and n times (may be NULL).  out buffers hold up to `cap` leaves.  Returns the number of
// output points; when the leaf is too small for the extent PCL returns the input unchanged — same here.
// Points of one leaf are accumulated in input order (PCL sorts (leaf, index) pairs with std::sort, which leaves the order
// inside a leaf unspecified: parity unpinned at the ulp level of the float sums).
extern "C" long orc_voxel_grid(const float* xyzi, const float* t_ms, size_t n, float leaf, float* out_xyzi, float* out_t, size_t cap) {
    const float inv = 1.0f / leaf;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    size_t nfin = 0;
    for (size_t i = 0; i < n; ++i) {
        const float* p = xyzi + 4 * i;
        if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) continue;
        for (int a = 0; a < 3; ++a) { mn[a] = std::min(mn[a], p[a]); mx[a] = std::max(mx[a], p[a]); }
        ++nfin;
    }
    auto passthrough = [&]() -> long {
        for (size_t i = 0; i < n && i < cap; ++i) {
            memcpy(out_xyzi + 4 * i, xyzi + 4 * i, 16);
            if (out_t) out_t[i] = t_ms ? t_ms[i] : 0.f;
        }
        return (long)n;
    };
    if (nfin == 0) return 0;
    const int64_t dx = (int64_t)((mx[0] - mn[0]) * inv) + 1, dy = (int64_t)((mx[1] - mn[1]) * inv) + 1, dz = (int64_t)((mx[2] - mn[2]) * inv) + 1;
    if (dx * dy * dz > (int64_t)INT32_MAX) return passthrough();
    int min_b[3], div_b[3];
    for (int a = 0; a < 3; ++a) {
        min_b[a] = (int)floorf(mn[a] * inv);
        const int max_b = (int)floorf(mx[a] * inv);
        div_b[a] = max_b - min_b[a] + 1;
    }
    const int mul[3] = {1, div_b[0], div_b[0] * div_b[1]};
    std::vector<std::pair<unsigned, unsigned>> iv;   // (leaf index, point index)
    iv.reserve(nfin);
    for (size_t i = 0; i < n; ++i) {
        const float* p = xyzi + 4 * i;
        if (!std::isfinite(p[0]) || !std::isfinite(p[1]) || !std::isfinite(p[2])) continue;
        const int i0 = (int)(floorf(p[0] * inv) - (float)min_b[0]);
        const int i1 = (int)(floorf(p[1] * inv) - (float)min_b[1]);
        const int i2 = (int)(floorf(p[2] * inv) - (float)min_b[2]);
        iv.emplace_back((unsigned)(i0 * mul[0] + i1 * mul[1] + i2 * mul[2]), (unsigned)i);
    }
    std::stable_sort(iv.begin(), iv.end(), [](const std::pair<unsigned, unsigned>& a, const std::pair<unsigned, unsigned>& b) { return a.first < b.first; });
    size_t out = 0;
    for (size_t s = 0; s < iv.size();) {
        size_t e = s;
        float sx = 0.f, sy = 0.f, sz = 0.f, si = 0.f, st = 0.f;
        while (e < iv.size() && iv[e].first == iv[s].first) {
            const float* p = xyzi + 4 * (size_t)iv[e].second;
            sx += p[0]; sy += p[1]; sz += p[2]; si += p[3];
            if (t_ms) st += t_ms[iv[e].second];
            ++e;
        }
        const float fn = (float)(e - s);
        if (out < cap) {
            out_xyzi[4 * out] = sx / fn; out_xyzi[4 * out + 1] = sy / fn; out_xyzi[4 * out + 2] = sz / fn; out_xyzi[4 * out + 3] = si / fn;
            if (out_t) out_t[out] = st / fn;
        }
        ++out;
        s = e;
    }
    return (long)out;
}
