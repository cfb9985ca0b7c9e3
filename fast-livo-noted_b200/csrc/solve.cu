// solve.cu — "(c) the 12x12 solve and state/covariance update kept on-device in double precision".
//
// One single-CTA kernel per trip closes the loop on the device:
//   * fixed-order final reduction of the fused kernel's group partials into H^T H (12x12) and H^T z;
//   * V3: one trip of esekf::update_iterated_dyn_share_modified
//         (reference include/IKFoM_toolkit/esekfom/esekfom.hpp:1619-1928) on the 23-DOF manifold state
//         (include/use-ikfom.hpp:12-21; SO3 / S2 algebra mtk/types/SOn.hpp:233-238, S2.hpp:128-280,
//         A_matrix mtk/src/mtkmath.hpp:236-247);
//   * V1: one trip of the loop body of LaserMapping::Run (src/laser_mapping.cpp:178-238) on the 18-DOF
//         StatesGroup (include/common_lib.h:552-587, Exp/Log include/so3_math.h:62-91);
//   * the re-match / stop policy, written to the device control block the next fused
//     launch reads — no host round trip between trips.
//
// Algebra note: the reference forms P_inv = ((P/R)^-1 + [HTH 0;0 0])^-1 with two dense
// n x n inverses and then only uses its first 12 columns.  By the matrix-inversion lemma those columns are
// W = Q[:, :12] (I + HTH Q[:12,:12])^-1 with Q = P/R: one 12x12 (V1: 6x6) solve, better conditioned, and
// identical in exact arithmetic.  Parity with the restated reference is tolerance-checked (dx <= 1e-6).
#include "fused.cuh"
#include "search.cuh"
#include "solve.cuh"

namespace flivo {

// ---- small dense helpers ------------------------------------------------------------------------------
__device__ void d_hat(const double* v, double* H) {
    H[0] = 0; H[1] = -v[2]; H[2] = v[1]; H[3] = v[2]; H[4] = 0; H[5] = -v[0]; H[6] = -v[1]; H[7] = v[0]; H[8] = 0;
}
__device__ void d_mat3mul(const double* A, const double* B, double* C) {
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}
// MTK::A_matrix (mtkmath.hpp:236-247)
__device__ __noinline__ void d_A_matrix(const double* v, double* A) {
    double sq = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    double nrm = sqrt(sq);
    for (int i = 0; i < 9; ++i) A[i] = 0.0;
    A[0] = A[4] = A[8] = 1.0;
    if (nrm < 1e-11) return;
    double H[9], H2[9];
    d_hat(v, H);
    d_mat3mul(H, H, H2);
    double a = (1 - cos(nrm)) / sq, b = (1 - sin(nrm) / nrm) / sq;
    for (int i = 0; i < 9; ++i) A[i] += a * H[i] + b * H2[i];
}
// rotation-matrix form of SO3::exp (SOn.hpp:293-297 via MTK::exp / cos_sinc_sqrt, mtkmath.hpp:139-171,249-256):
// Rodrigues with series near zero.  The reference composes unit quaternions; identical up to rounding.
__device__ __noinline__ void d_so3_exp(const double* v, double* R) {
    double t2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    double a, b;  // sin(t)/t, (1-cos t)/t^2
    if (t2 < 1e-8) {
        a = 1.0 - t2 / 6.0 + t2 * t2 / 120.0;
        b = 0.5 - t2 / 24.0 + t2 * t2 / 720.0;
    } else {
        double t = sqrt(t2);
        a = sin(t) / t;
        b = (1.0 - cos(t)) / t2;
    }
    double K[9], K2[9];
    d_hat(v, K);
    d_mat3mul(K, K, K2);
    for (int i = 0; i < 9; ++i) R[i] = a * K[i] + b * K2[i];
    R[0] += 1.0; R[4] += 1.0; R[8] += 1.0;
}
// SO3 log of a rotation matrix, accurate for small angles (SOn.hpp:300-304 / MTK::log computes the same
// vector from the unit quaternion: 2*atan(|v|/w)/|v| * v).
__device__ __noinline__ void d_so3_log(const double* R, double* o) {
    double vx = 0.5 * (R[7] - R[5]), vy = 0.5 * (R[2] - R[6]), vz = 0.5 * (R[3] - R[1]);  // sin(t) * axis
    double s = sqrt(vx * vx + vy * vy + vz * vz);
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
    double f;
    if (s < 1e-11) f = 1.0;
    else f = atan2(s, c) / s;
    o[0] = f * vx; o[1] = f * vy; o[2] = f * vz;
}
// V1 Exp / Log (include/so3_math.h:62-91)
__device__ __noinline__ void d_Exp_v1(double v1, double v2, double v3, double* R) {
    double nrm = sqrt(v1 * v1 + v2 * v2 + v3 * v3);
    for (int i = 0; i < 9; ++i) R[i] = 0;
    R[0] = R[4] = R[8] = 1;
    if (nrm > 0.00001) {
        double r[3] = {v1 / nrm, v2 / nrm, v3 / nrm};
        double K[9], K2[9];
        d_hat(r, K);
        d_mat3mul(K, K, K2);
        double s = sin(nrm), c = 1.0 - cos(nrm);
        for (int i = 0; i < 9; ++i) R[i] += s * K[i] + c * K2[i];
    }
}
__device__ __noinline__ void d_Log_v1(const double* R, double* o) {
    double tr = R[0] + R[4] + R[8];
    double theta = (tr > 3.0 - 1e-6) ? 0.0 : acos(0.5 * (tr - 1));
    double K[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    double f = (fabs(theta) < 0.001) ? 0.5 : (0.5 * theta / sin(theta));
    for (int i = 0; i < 3; ++i) o[i] = f * K[i];
}

// ---- S2 (gravity) algebra, MTK::S2<double, 98090, 10000, 1> (use-ikfom.hpp:8; S2.hpp) ----------------------
#define S2_LEN (98090.0 / 10000.0)
__device__ __noinline__ void d_s2_Bx(const double* vec, double* Bx) {  // 3x2 row-major, S2_typ == 1 (S2.hpp:213-231)
    const double len = S2_LEN;
    if (vec[0] + len > 1e-11) {
        double d = len + vec[0];
        Bx[0] = -vec[1];                   Bx[1] = -vec[2];
        Bx[2] = len - vec[1] * vec[1] / d; Bx[3] = -vec[2] * vec[1] / d;
        Bx[4] = -vec[2] * vec[1] / d;      Bx[5] = len - vec[2] * vec[2] / d;
        for (int i = 0; i < 6; ++i) Bx[i] /= len;
    } else {
        for (int i = 0; i < 6; ++i) Bx[i] = 0.0;
        Bx[3] = -1; Bx[4] = 1;
    }
}
__device__ void d_cross(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __noinline__ void d_s2_boxminus(const double* vec, const double* other, double* res2) {  // S2.hpp:136-160
    double cr[3];
    d_cross(vec, other, cr);
    double v_sin = sqrt(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
    double v_cos = vec[0] * other[0] + vec[1] * other[1] + vec[2] * other[2];
    double theta = atan2(v_sin, v_cos);
    if (v_sin < 1e-11) {
        if (fabs(theta) > 1e-11) { res2[0] = 3.1415926; res2[1] = 0; }
        else { res2[0] = 0; res2[1] = 0; }
    } else {
        double Bx[6], hv[3];
        d_s2_Bx(other, Bx);
        d_cross(other, vec, hv);
        double f = theta / v_sin;
        for (int c = 0; c < 2; ++c) res2[c] = f * (Bx[c] * hv[0] + Bx[2 + c] * hv[1] + Bx[4 + c] * hv[2]);
    }
}
__device__ __noinline__ void d_s2_boxplus(double* vec, const double* delta2) {  // S2.hpp:128-134
    double Bx[6], Bu[3], R[9], o[3];
    d_s2_Bx(vec, Bx);
    for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
    d_so3_exp(Bu, R);
    for (int r = 0; r < 3; ++r) o[r] = R[r * 3] * vec[0] + R[r * 3 + 1] * vec[1] + R[r * 3 + 2] * vec[2];
    vec[0] = o[0]; vec[1] = o[1]; vec[2] = o[2];
}
__device__ __noinline__ void d_s2_Nx_yy(const double* vec, double* Nx) {  // 2x3, S2.hpp:259-264
    double Bx[6], H[9];
    d_s2_Bx(vec, Bx);
    d_hat(vec, H);
    for (int r = 0; r < 2; ++r)
        for (int c = 0; c < 3; ++c)
            Nx[r * 3 + c] = 1 / S2_LEN / S2_LEN * (Bx[r] * H[c] + Bx[2 + r] * H[3 + c] + Bx[4 + r] * H[6 + c]);
}
// S2_Mx (S2.hpp:266-280).  "scalar(1/2)" there is integer division == 0, so exp_delta is the identity.
__device__ __noinline__ void d_s2_Mx(const double* vec, const double* delta2, double* Mx /*3x2*/) {
    double Bx[6], H[9];
    d_s2_Bx(vec, Bx);
    d_hat(vec, H);
    double dn = sqrt(delta2[0] * delta2[0] + delta2[1] * delta2[1]);
    if (dn < 1e-11) {
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 2; ++c) Mx[r * 2 + c] = -(H[r * 3] * Bx[c] + H[r * 3 + 1] * Bx[2 + c] + H[r * 3 + 2] * Bx[4 + c]);
    } else {
        double Bu[3], Am[9], T2[9];
        for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
        d_A_matrix(Bu, Am);
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) T2[i * 3 + j] = H[i * 3] * Am[j * 3] + H[i * 3 + 1] * Am[j * 3 + 1] + H[i * 3 + 2] * Am[j * 3 + 2];
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 2; ++c) Mx[r * 2 + c] = -(T2[r * 3] * Bx[c] + T2[r * 3 + 1] * Bx[2 + c] + T2[r * 3 + 2] * Bx[4 + c]);
    }
}

// state field offsets inside flivo_state (doubles)
enum { S_ROT = 0, S_POS = 9, S_OFFR = 12, S_OFFT = 21, S_VEL = 24, S_BG = 27, S_BA = 30, S_GRAV = 33 };

// x (-) y, state_ikfom order pos, rot, offR, offT, vel, bg, ba, grav(S2).  The three parts touch disjoint
// fields (0: pos, rot; 1: offR, offT, vel, bg, ba; 2: grav) so three threads of different warps run them at once.
__device__ __noinline__ void d_boxminus_v3_part(const double* x, const double* y, double* d, int part) {
    double Rd[9];
    if (part == 0) {
        for (int i = 0; i < 3; ++i) d[i] = x[S_POS + i] - y[S_POS + i];
        // other.conjugate() * this  ->  y.rot^T * x.rot
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Rd[i * 3 + j] = y[S_ROT + i] * x[S_ROT + j] + y[S_ROT + 3 + i] * x[S_ROT + 3 + j] + y[S_ROT + 6 + i] * x[S_ROT + 6 + j];
        d_so3_log(Rd, d + 3);
    } else if (part == 1) {
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) Rd[i * 3 + j] = y[S_OFFR + i] * x[S_OFFR + j] + y[S_OFFR + 3 + i] * x[S_OFFR + 3 + j] + y[S_OFFR + 6 + i] * x[S_OFFR + 6 + j];
        d_so3_log(Rd, d + 6);
        for (int i = 0; i < 3; ++i) d[9 + i] = x[S_OFFT + i] - y[S_OFFT + i];
        for (int i = 0; i < 3; ++i) d[12 + i] = x[S_VEL + i] - y[S_VEL + i];
        for (int i = 0; i < 3; ++i) d[15 + i] = x[S_BG + i] - y[S_BG + i];
        for (int i = 0; i < 3; ++i) d[18 + i] = x[S_BA + i] - y[S_BA + i];
    } else {
        d_s2_boxminus(x + S_GRAV, y + S_GRAV, d + 21);
    }
}
__device__ __noinline__ void d_boxplus_v3_part(double* x, const double* d, int part) {
    double E[9], Rn[9];
    if (part == 0) {
        for (int i = 0; i < 3; ++i) x[S_POS + i] += d[i];
        d_so3_exp(d + 3, E);
        d_mat3mul(x + S_ROT, E, Rn);
        for (int i = 0; i < 9; ++i) x[S_ROT + i] = Rn[i];
    } else if (part == 1) {
        d_so3_exp(d + 6, E);
        d_mat3mul(x + S_OFFR, E, Rn);
        for (int i = 0; i < 9; ++i) x[S_OFFR + i] = Rn[i];
        for (int i = 0; i < 3; ++i) x[S_OFFT + i] += d[9 + i];
        for (int i = 0; i < 3; ++i) x[S_VEL + i] += d[12 + i];
        for (int i = 0; i < 3; ++i) x[S_BG + i] += d[15 + i];
        for (int i = 0; i < 3; ++i) x[S_BA + i] += d[18 + i];
    } else {
        d_s2_boxplus(x + S_GRAV, d + 21);
    }
}
// J = A_matrix(v)^T and seg <- J * seg  (esekfom.hpp:1663-1672)
__device__ __noinline__ void d_so3_jac(const double* v, double* J, double* seg) {
    double A[9];
    d_A_matrix(v, A);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) J[i * 3 + j] = A[j * 3 + i];
    if (seg) {
        double t0 = J[0] * seg[0] + J[1] * seg[1] + J[2] * seg[2];
        double t1 = J[3] * seg[0] + J[4] * seg[1] + J[5] * seg[2];
        double t2 = J[6] * seg[0] + J[7] * seg[1] + J[8] * seg[2];
        seg[0] = t0; seg[1] = t1; seg[2] = t2;
    }
}
// J2 = Nx(x.grav) * Mx(x_prop.grav, d2) and seg <- J2 * seg  (esekfom.hpp:1684-1695)
__device__ __noinline__ void d_s2_jac(const double* xg, const double* xpg, const double* d2, double* J2, double* seg) {
    double Nx[6], Mx[6];
    d_s2_Nx_yy(xg, Nx);
    d_s2_Mx(xpg, d2, Mx);
    for (int r = 0; r < 2; ++r) for (int c = 0; c < 2; ++c) J2[r * 2 + c] = Nx[r * 3] * Mx[c] + Nx[r * 3 + 1] * Mx[2 + c] + Nx[r * 3 + 2] * Mx[4 + c];
    if (seg) {
        double t0 = J2[0] * seg[0] + J2[1] * seg[1];
        double t1 = J2[2] * seg[0] + J2[3] * seg[1];
        seg[0] = t0; seg[1] = t1;
    }
}

// T = blockdiag(I3, J3, J6, I.., J2) stored sparsely: apply to rows / columns of an n x n matrix in shared memory
struct BlockT {
    double J3[9], J6[9], J2[4];
};
// M <- T * M  (rows 3..5, 6..8, 21..22); all threads participate, caller syncs
__device__ void apply_T_rows(double* M, int n, int ncols, const BlockT& T, int tid, int nthreads) {
    for (int c = tid; c < ncols; c += nthreads) {
        double a0 = M[3 * n + c], a1 = M[4 * n + c], a2 = M[5 * n + c];
        M[3 * n + c] = T.J3[0] * a0 + T.J3[1] * a1 + T.J3[2] * a2;
        M[4 * n + c] = T.J3[3] * a0 + T.J3[4] * a1 + T.J3[5] * a2;
        M[5 * n + c] = T.J3[6] * a0 + T.J3[7] * a1 + T.J3[8] * a2;
        a0 = M[6 * n + c]; a1 = M[7 * n + c]; a2 = M[8 * n + c];
        M[6 * n + c] = T.J6[0] * a0 + T.J6[1] * a1 + T.J6[2] * a2;
        M[7 * n + c] = T.J6[3] * a0 + T.J6[4] * a1 + T.J6[5] * a2;
        M[8 * n + c] = T.J6[6] * a0 + T.J6[7] * a1 + T.J6[8] * a2;
        a0 = M[21 * n + c]; a1 = M[22 * n + c];
        M[21 * n + c] = T.J2[0] * a0 + T.J2[1] * a1;
        M[22 * n + c] = T.J2[2] * a0 + T.J2[3] * a1;
    }
}
// M <- M * T^T  (columns 3..5, 6..8, 21..22)
__device__ void apply_T_cols(double* M, int n, const BlockT& T, int tid, int nthreads) {
    for (int r = tid; r < n; r += nthreads) {
        double* row = M + r * n;
        double a0 = row[3], a1 = row[4], a2 = row[5];
        row[3] = a0 * T.J3[0] + a1 * T.J3[1] + a2 * T.J3[2];
        row[4] = a0 * T.J3[3] + a1 * T.J3[4] + a2 * T.J3[5];
        row[5] = a0 * T.J3[6] + a1 * T.J3[7] + a2 * T.J3[8];
        a0 = row[6]; a1 = row[7]; a2 = row[8];
        row[6] = a0 * T.J6[0] + a1 * T.J6[1] + a2 * T.J6[2];
        row[7] = a0 * T.J6[3] + a1 * T.J6[4] + a2 * T.J6[5];
        row[8] = a0 * T.J6[6] + a1 * T.J6[7] + a2 * T.J6[8];
        a0 = row[21]; a1 = row[22];
        row[21] = a0 * T.J2[0] + a1 * T.J2[1];
        row[22] = a0 * T.J2[2] + a1 * T.J2[3];
    }
}

// Solve S^T X = B by Gauss-Jordan with partial pivoting, the whole CTA on the m x w array [S^T | B] (w = m + nrhs)
// in shared memory.  k_solve is one cold CTA: its time is dependent-instruction latency, so the step of one pivot is
// arranged as ONE barrier interval.  Every thread owns one element, computes its new value from the OLD matrix (the
// row swap p <-> k is applied by index) and writes it to the other of two buffers; the thread -> column mapping
// rotates with the pivot so that the next pivot column always lands in lanes 0..m-1 of warp 0, which pick the next
// pivot row straight from registers (three hardware warp reductions: non-negative doubles order like their bit
// patterns) before the barrier.  Same operations per entry as the textbook order: pivot row scaled by 1/piv (a
// division), other rows  a_ic - a_ik * (scaled pivot row)_c;  first maximum wins pivot ties.
// Returns the buffer that holds [ . | X ]; *ok = false on a zero pivot.
__device__ __forceinline__ void gj_pick(bool cand, double v, int lane, int* s_piv, double* s_mag) {
    const unsigned FULLM = 0xFFFFFFFFu;
    const unsigned long long bits = cand ? (unsigned long long)__double_as_longlong(fabs(v)) : 0ull;
    const unsigned hi = (unsigned)(bits >> 32), lo = (unsigned)bits;
    const unsigned mhi = __reduce_max_sync(FULLM, cand ? hi : 0u);
    const unsigned mlo = __reduce_max_sync(FULLM, (cand && hi == mhi) ? lo : 0u);
    const unsigned idx = __reduce_min_sync(FULLM, (cand && hi == mhi && lo == mlo) ? (unsigned)lane : 0xFFFFFFFFu);
    if (lane == 0) { *s_piv = (int)idx; *s_mag = __longlong_as_double((long long)(((unsigned long long)mhi << 32) | mlo)); }
}
__device__ const double* block_gauss_jordan(double* A0, double* A1, int m, int w, double* s_mag /*[2]*/, int* s_piv /*[2]*/, int tid, bool* ok_out) {
    bool ok = true;
    if (tid < 32) gj_pick(tid < m, tid < m ? A0[tid * w] : 0.0, tid, &s_piv[0], &s_mag[0]);
    __syncthreads();
    const bool active = tid < m * w;
    const int slot = tid / m, i = tid - slot * m;
    double* cur = A0;
    double* nxt = A1;
    for (int k = 0; k < m; ++k) {
        const int p = s_piv[k & 1];
        if (s_mag[k & 1] == 0.0) ok = false;
        double nv = 0.0;
        if (active) {
            int c = slot + k + 1;                                  // slot 0 <-> column k+1, the next pivot column
            if (c >= w) c -= w;
            const double prow = cur[p * w + c] / cur[p * w + k];   // scaled pivot row
            const int src = (i == k) ? p : ((i == p) ? k : i);     // row that sits at position i after the swap
            nv = (i == k) ? prow : (cur[src * w + c] - cur[src * w + k] * prow);
            nxt[i * w + c] = nv;
        }
        if (tid < 32 && k + 1 < m) gj_pick(tid < m && tid > k, nv, tid, &s_piv[(k + 1) & 1], &s_mag[(k + 1) & 1]);
        __syncthreads();
        double* t = cur; cur = nxt; nxt = t;
    }
    *ok_out = ok;
    return cur;
}

// ---- fixed-order final reduction of the group partials ------------------------------------------------------
// out92[t] = sum over groups; deterministic: slices of groups are summed sequentially by fixed threads, then the
// slice sums are added in slice order.
__device__ void reduce_groups(const double* __restrict__ gpartials, int groups, double* s_slices /*[8][96]*/, double* out92,
                              int tid, int nthreads) {
    const int NS = 8;
    int slice = tid / PSTRIDE, t = tid % PSTRIDE;
    if (slice < NS && t < NTERMS) {
        int g0 = (int)(((long long)groups * slice) / NS), g1 = (int)(((long long)groups * (slice + 1)) / NS);
        double s = 0.0;
        for (int g = g0; g < g1; ++g) s += __ldcg(gpartials + (size_t)g * PSTRIDE + t);
        s_slices[slice * PSTRIDE + t] = s;
    }
    __syncthreads();
    if (tid < NTERMS) {
        double s = 0.0;
        for (int sl = 0; sl < NS; ++sl) s += s_slices[sl * PSTRIDE + tid];
        out92[tid] = s;
    }
    __syncthreads();
}

constexpr int SOLVE_THREADS = 768;  // 8 slices x 96

__device__ __forceinline__ void stamp(const SolveParams& S, const Ctl* ctl, int k, int back = 0) {
    if (S.stamps && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        int trip = ctl->trips - back;   // back = 1 once this trip has been counted
        trip = trip < 0 ? 0 : (trip < MAX_TRIPS ? trip : MAX_TRIPS - 1);
        S.stamps[trip * 16 + k] = t;
    }
}

__global__ void __launch_bounds__(SOLVE_THREADS) k_solve(SolveParams S) {
    __shared__ double s_slices[8 * PSTRIDE];
    __shared__ double s_sum[PSTRIDE];
    __shared__ double s_H[144];
    __shared__ double s_z[12];
    __shared__ double s_P[23 * 23];
    __shared__ double s_L[23 * 23];
    __shared__ double s_Aug[12 * 35], s_Aug2[12 * 35];   // Gauss-Jordan ping-pong
    __shared__ double s_W[23 * 12];
    __shared__ double s_Kx[23 * 12];
    __shared__ double s_x[36], s_xp[36];
    __shared__ double s_dxnew[23], s_dxo[23], s_Kh[23];
    __shared__ BlockT s_T;
    __shared__ int s_flag[4];  // [0]=stop, [1]=ok, [2]=pivot row
    __shared__ double s_gjf[16];

    Ctl* ctl = S.ctl;
    const int tid = threadIdx.x, NT = blockDim.x;
    stamp(S, ctl, 0);
    // the state, P and the state-only half prepared by k_solve_pre do not depend on the sums, nor on the exit test: all of
    // their global loads are issued here, together with the test's own words, and land in registers while the group
    // records are being summed (P: both flavours' sources are read, the mode picks one afterwards)
    const int stopped = ctl->stopped;
    const int mode = S.job->mode;
    double r_P1 = 0.0, r_P3 = 0.0, r_x = 0.0, r_xp = 0.0, r_dx = 0.0;
    if (S.phase != SOLVE_PHASE_REDUCE_ONLY) {
        if (tid < 18 * 18) r_P1 = S.P_prop[tid];
        if (tid < 23 * 23) r_P3 = S.pre->Pt[tid];
        if (tid < 36) { r_x = S.x[tid]; r_xp = S.x_prop[tid]; }
        if (tid < 23) r_dx = S.pre->dxnew[tid];
    }
    // the V3 loop decision is taken by warp 4: lane i owns component i's convergence limit, lane 0 the loop counters
    double r_lim = 0.0;
    int c_t = 0, c_iter = 0, c_trips = 0, c_search = 0, c_searches = 0, j_max_iter = 0, j_force = 0;
    int c_rematch = 0;
    if (S.phase != SOLVE_PHASE_REDUCE_ONLY) {
        if ((tid >> 5) == 4 && tid - 128 < 23) r_lim = S.job->limit[tid - 128];
        if (tid == 128 || tid == 0) {   // (thread 0 takes the V1 decision)
            c_t = ctl->t; c_iter = ctl->iter; c_trips = ctl->trips; c_search = ctl->search; c_searches = ctl->searches;
            c_rematch = ctl->rematch;
            j_max_iter = S.job->max_iter; j_force = S.job->force_search;
        }
    }
    if (S.phase != SOLVE_PHASE_REDUCE_ONLY && stopped) return;
    const int n = (mode == 1) ? 18 : 23;
    const double r_P = (mode == 1) ? r_P1 : r_P3;

    // ---- 0. normal equations -------------------------------------------------------------------------
    if (S.phase == SOLVE_PHASE_PEER) {
        // ---- fused reduce + exchange over peer memory ------------------------------------------------------
        // Every rank holds the same state, so all ranks run the same trips and their exchange counters agree.
        __shared__ unsigned long long s_seq;
        __shared__ int s_timeout;
        reduce_groups(S.gpartials, S.job->groups, s_slices, s_sum, tid, NT);
        if (tid == 0) { s_seq = *S.peer.seq + 1ull; s_timeout = 0; }
        __syncthreads();
        const unsigned long long seq = s_seq;
        const int par = (int)(seq & 1ull), me = S.peer.rank, nr = S.peer.nranks;
        for (int e = tid; e < nr * PSTRIDE; e += NT) {                 // my record -> every rank's buffer (NVLink stores)
            const int r = e / PSTRIDE, t = e - r * PSTRIDE;
            S.peer.base[r][par * PEER_MAX_RANKS + me].rec[t] = (t < NTERMS) ? s_sum[t] : 0.0;
        }
        __threadfence_system();
        __syncthreads();
        if (tid < nr) {
            unsigned long long* f = &S.peer.base[tid][par * PEER_MAX_RANKS + me].flag;
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"(seq) : "memory");
            // wait for rank `tid`'s record in my own buffer (bounded: a lost peer must not hang the GPU)
            const unsigned long long* g = &S.peer.base[me][par * PEER_MAX_RANKS + tid].flag;
            unsigned long long t0, t1, v;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            for (;;) {
                asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(g) : "memory");
                if (v == seq) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ull) { s_timeout = 1; break; }      // 2 s
            }
        }
        __syncthreads();
        if (tid < NTERMS) {                                             // fixed rank order: bit-identical on all ranks
            double s = 0.0;
            for (int r = 0; r < nr; ++r) {
                double v;
                asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(&S.peer.base[me][par * PEER_MAX_RANKS + r].rec[tid]) : "memory");
                s += v;
            }
            s_sum[tid] = s_timeout ? __longlong_as_double(0x7FF8000000000000ll) : s;   // NaN -> the guard below stops the update
        }
        if (tid == 0) *S.peer.seq = seq;
        __syncthreads();
    } else if (S.phase == SOLVE_PHASE_FROM_RANKS) {
        // multi-GPU: every rank's 92-term block was gathered; sum in rank order (bit-identical on all ranks)
        if (tid < NTERMS) {
            double s = 0.0;
            for (int r = 0; r < S.job->nranks; ++r) s += S.rank_blocks[(size_t)r * PSTRIDE + tid];
            s_sum[tid] = s;
        }
        __syncthreads();
    } else {
        reduce_groups(S.gpartials, S.job->groups, s_slices, s_sum, tid, NT);
    }
    stamp(S, ctl, 1);
    if (S.phase == SOLVE_PHASE_REDUCE_ONLY) {
        if (tid < NTERMS) S.sum_out[tid] = s_sum[tid];
        if (tid == 0) { ctl_next_trip(ctl); }
        return;
    }
    if (tid < 144) {
        int a = tid / 12, b = tid % 12;
        int i = a < b ? a : b, j = a < b ? b : a;
        int t = i * 12 - (i * (i - 1)) / 2 + (j - i);  // index of (i, j), i <= j, in the 78 upper terms
        s_H[tid] = s_sum[t];
    }
    if (tid < 12) s_z[tid] = s_sum[78 + tid];
    if (tid < 36) { s_x[tid] = r_x; s_xp[tid] = r_xp; }
    {   // the state-only half of this trip was prepared by k_solve_pre while the search / fused kernels ran
        const double* Psrc = (mode == 1) ? S.P_prop : S.pre->Pt;
        if (tid < n * n) s_P[tid] = r_P;
        for (int e = tid + NT; e < n * n; e += NT) s_P[e] = Psrc[e];   // (only when launched with fewer than n*n threads)
        if (tid < 23) s_dxnew[tid] = r_dx;
    }
    __syncthreads();
    if (tid == 0) {
        int ne = (int)s_sum[91];
        ctl->n_eff = ne;
        ctl->res_mean = ne > 0 ? s_sum[90] / (double)ne : 0.0;
    }
    if (S.sum_out && tid < NTERMS) S.sum_out[tid] = s_sum[tid];
    const int all_finite = __syncthreads_and(tid >= NTERMS || isfinite(s_sum[tid]));
    if (!all_finite) {  // NaN guard: leave x and P untouched and stop the loop
        if (tid == 0) { ctl->nan_guard = 1; ctl->stopped = 1; ctl->trips += 1; ctl_next_trip(ctl); }
        return;
    }

    stamp(S, ctl, 2);
    stamp(S, ctl, 3);   // (3, 4: the manifold / T P T^T phases moved to k_solve_pre)
    stamp(S, ctl, 4);
    if (mode == 0) {
        // ================= V3: esekfom.hpp:1633-1927 ==================================================
        // Aug = [S^T | Q1^T],  S = I + HTH * Q11,  Q = P / R
        const double invR = 1.0 / S.job->R;
        for (int e = tid; e < 12 * 35; e += NT) {
            int r = e / 35, c = e % 35;
            double v;
            if (c < 12) {  // S^T[r][c] = S[c][r] = delta + sum_k H[c][k] Q[k][r]
                double s = 0.0;
                for (int k = 0; k < 12; ++k) s += s_H[c * 12 + k] * (s_P[k * 23 + r] * invR);
                v = s + (r == c ? 1.0 : 0.0);
            } else {       // Q1^T[r][i] = Q[i][r]
                int i = c - 12;
                v = s_P[i * 23 + r] * invR;
            }
            s_Aug[e] = v;
        }
        __syncthreads();
        stamp(S, ctl, 5);
        bool gj_ok;
        const double* X = block_gauss_jordan(s_Aug, s_Aug2, 12, 35, s_gjf, &s_flag[2], tid, &gj_ok);
        if (tid == 0) s_flag[1] = gj_ok ? 1 : 0;
        stamp(S, ctl, 6);
        // W[i][a] = X[a][i];  K_h = W * HTz (:1801);  K_x[:, :12] = W * HTH (:1806)
        for (int e = tid; e < 23 * 12; e += NT) { int i = e / 12, a = e % 12; s_W[e] = X[a * 35 + 12 + i]; }
        __syncthreads();
        for (int e = tid; e < 23 * 12; e += NT) {
            int i = e / 12, b = e % 12;
            double s = 0.0;
            for (int a = 0; a < 12; ++a) s += s_W[i * 12 + a] * s_H[a * 12 + b];
            s_Kx[e] = s;
        }
        if (tid < 23) {
            double s = 0.0;
            for (int a = 0; a < 12; ++a) s += s_W[tid * 12 + a] * s_z[a];
            s_Kh[tid] = s;
        }
        __syncthreads();
        if (tid < 23) {                                                       // :1812
            double s = s_Kh[tid];
            for (int j = 0; j < 12; ++j) s += s_Kx[tid * 12 + j] * s_dxnew[j];
            s_dxo[tid] = s - s_dxnew[tid];
        }
        __syncthreads();
        stamp(S, ctl, 7);
        if ((tid & 31) == 0 && tid < 96) d_boxplus_v3_part(s_x, s_dxo, tid >> 5);   // :1814
        if ((tid >> 5) == 4) {
            const int lane = tid & 31;
            // esekfom.hpp:1816-1823: every component of dx below its limit (one lane each; the limits were loaded up front)
            int converge = __all_sync(0xffffffffu, lane >= 23 || !(fabs(s_dxo[lane < 23 ? lane : 0]) > r_lim));
            if (lane < 23) ctl->dx_last[lane] = s_dxo[lane];
            if (lane == 0) {
            int t = c_t, it = c_iter;
            if (converge) t++;
            if (!t && it == j_max_iter - 2) converge = 1;                         // :1825-1828
            int stop = (t > 1 || it == j_max_iter - 1) ? 1 : 0;                    // :1831
            ctl->t = t;
            int trip = c_trips;
            if (trip < MAX_TRIPS) ctl->trip_searched[trip] = (c_search || j_force) ? 1 : 0;
            ctl->searches = c_searches + ((c_search || j_force) ? 1 : 0);
            ctl->trips = trip + 1;
            ctl->search = converge;
            ctl->iter = it + 1;
            ctl->converged = t > 0 ? 1 : 0;
            ctl->stopped = stop;
            ctl_next_trip(ctl);
            s_flag[0] = stop;
            }
        }
        __syncthreads();
        stamp(S, ctl, 8, 1);
        if (s_flag[0] && (tid & 31) == 0 && tid < 96) {
            const int part = tid >> 5;
            if (part == 0) d_so3_jac(s_dxo + 3, s_T.J3, nullptr);             // :1842
            else if (part == 1) d_so3_jac(s_dxo + 6, s_T.J6, nullptr);
            else d_s2_jac(s_x + S_GRAV, s_xp + S_GRAV, s_dxo + 21, s_T.J2, nullptr);   // :1877-1879 (x_ already updated)
        }
        __syncthreads();
        if (tid < 36) S.x[tid] = s_x[tid];
        if (s_flag[0]) {
            // L = T' P T'^T, P <- P T'^T, K_x <- T' K_x, P_out = L - K_x[:, :12] P[:12, :]   (:1833-1921)
            apply_T_cols(s_P, 23, s_T, tid, NT);
            __syncthreads();
            for (int e = tid; e < 23 * 23; e += NT) s_L[e] = s_P[e];
            __syncthreads();
            apply_T_rows(s_L, 23, 23, s_T, tid, NT);
            apply_T_rows(s_Kx, 12, 12, s_T, tid, NT);
            __syncthreads();
            for (int e = tid; e < 23 * 23; e += NT) {
                int i = e / 23, j = e % 23;
                double s = 0.0;
                for (int a = 0; a < 12; ++a) s += s_Kx[i * 12 + a] * s_P[a * 23 + j];
                S.P_out[e] = s_L[e] - s;
            }
        }
        stamp(S, ctl, 9, 1);
    } else {
        // ================= V1: src/laser_mapping.cpp:178-238 ==========================================
        // HPH (6x6) = Hsub^T (1/R) Hsub with Hsub = [A | n] = V3 columns [3:6 | 0:3]  (:584-586)
        __shared__ double s_H6[36], s_z6[6], s_vec[18], s_sol[18], s_G[18 * 6];
        const double w = 1.0 / S.job->R;
        if (tid < 36) {
            const int map6[6] = {3, 4, 5, 0, 1, 2};
            int a = tid / 6, b = tid % 6;
            s_H6[tid] = s_H[map6[a] * 12 + map6[b]] * w;
            if (tid < 6) s_z6[tid] = s_z[map6[tid]] * w;
        }
        if (tid < 18) s_vec[tid] = S.pre->vec[tid];   // state_propagat - state, from k_solve_pre
        __syncthreads();
        if (S.job->gnss_on) {
            // HPH.block<6,6>(0,0) += gnss_HPH; HPL.block<6,1>(0,0) += gnss_HPL (src/laser_mapping.cpp:612-617) with
            // computeH (src/GNSS_Processing.cpp:204-237) evaluated at the current iterate: Hsub = [R [l]x | -I] (3x6),
            // error = p_gnss - pos + R l, dropped when |error| > 2 m, weights diag(1/std0, 1/std1, 100/std2)
            if (tid == 0) {
                const double* R = s_x + S_ROT;
                const double* l = S.job->gnss_lever;
                const double K[9] = {0.0, -l[2], l[1], l[2], 0.0, -l[0], -l[1], l[0], 0.0};   // SKEW_SYM_MATRX (so3_math.h:9)
                double Hs[18], RK[9], err[3];
                d_mat3mul(R, K, RK);
                for (int r = 0; r < 3; ++r) {
                    for (int c = 0; c < 3; ++c) { Hs[r * 6 + c] = RK[r * 3 + c]; Hs[r * 6 + 3 + c] = (r == c) ? -1.0 : 0.0; }
                    err[r] = S.job->gnss_pos[r] - s_x[S_POS + r] + (R[r * 3] * l[0] + R[r * 3 + 1] * l[1] + R[r * 3 + 2] * l[2]);
                }
                if (!(sqrt(err[0] * err[0] + err[1] * err[1] + err[2] * err[2]) > 2.0)) {
                    const double wg[3] = {1.0 / S.job->gnss_std[0], 1.0 / S.job->gnss_std[1], 100.0 / S.job->gnss_std[2]};
                    for (int a = 0; a < 6; ++a) {
                        for (int b = 0; b < 6; ++b) {
                            double acc = 0.0;
                            for (int k = 0; k < 3; ++k) acc += (Hs[k * 6 + a] * wg[k]) * Hs[k * 6 + b];
                            s_H6[a * 6 + b] += acc;
                        }
                        double acc = 0.0;
                        for (int k = 0; k < 3; ++k) acc += (Hs[k * 6 + a] * wg[k]) * err[k];
                        s_z6[a] += acc;
                    }
                }
            }
            __syncthreads();
        }
        // Aug (6 x 24) = [S^T | C1^T],  S = I + H6 * C11,  C = state.cov
        for (int e = tid; e < 6 * 24; e += NT) {
            int r = e / 24, c = e % 24;
            double v;
            if (c < 6) {
                double s = 0.0;
                for (int k = 0; k < 6; ++k) s += s_H6[c * 6 + k] * s_P[k * 18 + r];
                v = s + (r == c ? 1.0 : 0.0);
            } else {
                v = s_P[(c - 6) * 18 + r];
            }
            s_Aug[e] = v;
        }
        __syncthreads();
        bool gj_ok;
        const double* X = block_gauss_jordan(s_Aug, s_Aug2, 6, 24, s_gjf, &s_flag[2], tid, &gj_ok);
        if (tid == 0) s_flag[1] = gj_ok ? 1 : 0;
        for (int e = tid; e < 18 * 6; e += NT) { int i = e / 6, a = e % 6; s_W[e] = X[a * 24 + 6 + i]; }  // K_1[:, :6]
        __syncthreads();
        for (int e = tid; e < 18 * 6; e += NT) {                             // G[:, :6] = K_1[:, :6] * HPH (:191)
            int i = e / 6, b = e % 6;
            double s = 0.0;
            for (int a = 0; a < 6; ++a) s += s_W[i * 6 + a] * s_H6[a * 6 + b];
            s_G[e] = s;
        }
        __syncthreads();
        if (tid < 18) {                                                      // :193
            double s1 = 0.0, s2 = 0.0;
            for (int a = 0; a < 6; ++a) { s1 += s_W[tid * 6 + a] * s_z6[a]; s2 += s_G[tid * 6 + a] * s_vec[a]; }
            s_sol[tid] = s1 + s_vec[tid] - s2;
        }
        __syncthreads();
        if (tid == 0) {
            double E[9], Rn[9];
            d_Exp_v1(s_sol[0], s_sol[1], s_sol[2], E);                        // state += solution (:204)
            d_mat3mul(s_x + S_ROT, E, Rn);
            for (int i = 0; i < 9; ++i) s_x[S_ROT + i] = Rn[i];
            for (int i = 0; i < 3; ++i) {
                s_x[S_POS + i] += s_sol[3 + i]; s_x[S_VEL + i] += s_sol[6 + i]; s_x[S_BG + i] += s_sol[9 + i];
                s_x[S_BA + i] += s_sol[12 + i]; s_x[S_GRAV + i] += s_sol[15 + i];
            }
            const double R2D = 180.0 / 3.14159265358979323846;
            double rn = sqrt(s_sol[0] * s_sol[0] + s_sol[1] * s_sol[1] + s_sol[2] * s_sol[2]);
            double tn = sqrt(s_sol[3] * s_sol[3] + s_sol[4] * s_sol[4] + s_sol[5] * s_sol[5]);
            int converged = ((rn * R2D < 0.01) && (tn * 100 < 0.015)) ? 1 : 0;   // :209-212
            int it = c_iter, rematch = c_rematch, nse = 0;
            if (converged || ((rematch == 0) && (it == (j_max_iter - 2)))) { nse = 1; rematch++; }   // :217-222
            int stop = (rematch >= 2 || (it == j_max_iter - 1)) ? 1 : 0;          // :224
            int trip = c_trips;
            if (trip < MAX_TRIPS) ctl->trip_searched[trip] = (c_search || j_force) ? 1 : 0;
            ctl->searches = c_searches + ((c_search || j_force) ? 1 : 0);
            ctl->trips = trip + 1;
            ctl->rematch = rematch;
            ctl->search = nse;
            ctl->iter = it + 1;
            ctl->converged = converged;
            ctl->stopped = stop;
            ctl_next_trip(ctl);
            for (int i = 0; i < 18; ++i) ctl->dx_last[i] = s_sol[i];
            s_flag[0] = stop;
        }
        __syncthreads();
        if (tid < 36) S.x[tid] = s_x[tid];
        if (s_flag[0]) {                                                     // state.cov = (I - G) * state.cov (:229)
            for (int e = tid; e < 18 * 18; e += NT) {
                int i = e / 18, j = e % 18;
                double s = 0.0;
                for (int a = 0; a < 6; ++a) s += s_G[i * 6 + a] * s_P[a * 18 + j];
                S.P_out[e] = s_P[e] - s;
            }
        }
    }
}


// ---- state-only half of the solve (see SolvePre) ---------------------------------------------------------------
__global__ void __launch_bounds__(256) k_solve_pre(SolveParams S) {
    __shared__ double s_P[23 * 23];
    __shared__ double s_x[36], s_xp[36];
    __shared__ double s_dx[23], s_dxnew[23];
    __shared__ BlockT s_T;
    const Ctl* ctl = S.ctl;
    if (ctl->stopped) return;
    const int tid = threadIdx.x, NT = blockDim.x;
    const int mode = S.job->mode;
    if (tid < 36) { s_x[tid] = S.x[tid]; s_xp[tid] = S.x_prop[tid]; }
    if (mode == 0) for (int e = tid; e < 23 * 23; e += NT) s_P[e] = S.P_prop[e];
    __syncthreads();
    if (mode == 0) {
        if ((tid & 31) == 0 && tid < 96) {
            const int part = tid >> 5;                                        // three warps, one lane each
            d_boxminus_v3_part(s_x, s_xp, s_dx, part);                        // esekfom.hpp:1646
            if (part == 0) {
                for (int i = 0; i < 6; ++i) s_dxnew[i] = s_dx[i];
                d_so3_jac(s_dx + 3, s_T.J3, s_dxnew + 3);                     // :1663
            } else if (part == 1) {
                for (int i = 6; i < 21; ++i) s_dxnew[i] = s_dx[i];
                d_so3_jac(s_dx + 6, s_T.J6, s_dxnew + 6);
            } else {
                s_dxnew[21] = s_dx[21]; s_dxnew[22] = s_dx[22];
                d_s2_jac(s_x + S_GRAV, s_xp + S_GRAV, s_dx + 21, s_T.J2, s_dxnew + 21);   // :1684-1686
            }
        }
        __syncthreads();
        apply_T_rows(s_P, 23, 23, s_T, tid, NT);                              // :1665-1672, :1688-1695
        __syncthreads();
        apply_T_cols(s_P, 23, s_T, tid, NT);
        __syncthreads();
        for (int e = tid; e < 23 * 23; e += NT) S.pre->Pt[e] = s_P[e];
        if (tid < 23) S.pre->dxnew[tid] = s_dxnew[tid];
    } else if (tid == 0) {
        // vec = state_propagat - state (common_lib.h:576-587)
        double rotd[9], vec[18];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j)
                rotd[i * 3 + j] = s_x[S_ROT + i] * s_xp[S_ROT + j] + s_x[S_ROT + 3 + i] * s_xp[S_ROT + 3 + j] + s_x[S_ROT + 6 + i] * s_xp[S_ROT + 6 + j];
        d_Log_v1(rotd, vec);
        for (int i = 0; i < 3; ++i) {
            vec[3 + i] = s_xp[S_POS + i] - s_x[S_POS + i];
            vec[6 + i] = s_xp[S_VEL + i] - s_x[S_VEL + i];
            vec[9 + i] = s_xp[S_BG + i] - s_x[S_BG + i];
            vec[12 + i] = s_xp[S_BA + i] - s_x[S_BA + i];
            vec[15 + i] = s_xp[S_GRAV + i] - s_x[S_GRAV + i];
        }
        for (int i = 0; i < 18; ++i) S.pre->vec[i] = vec[i];
    }
}
void launch_solve_pre(const SolveParams& S, cudaStream_t stream) { k_solve_pre<<<1, 256, 0, stream>>>(S); }

void launch_solve(const SolveParams& S, cudaStream_t stream) { k_solve<<<1, SOLVE_THREADS, 0, stream>>>(S); }

// start of an update: reset the control block, x_prop <- x, P_out <- P_prop (so a loop that never stops returns P unchanged)
__global__ void k_ctl_reset(const double* x, double* x_prop, const double* P_prop, double* P_out, Ctl* ctl, SearchStats* stats) {
    for (int e = threadIdx.x; e < 23 * 23; e += blockDim.x) P_out[e] = P_prop[e];
    if (threadIdx.x < 36) x_prop[threadIdx.x] = x[threadIdx.x];
    if (threadIdx.x == 0) {
        ctl->search = 1; ctl->stopped = 0; ctl->iter = -1; ctl->t = 0; ctl->rematch = 0; ctl->trips = 0;
        ctl->searches = 0; ctl->converged = 0; ctl->n_eff = 0; ctl->nan_guard = 0; ctl->res_mean = 0.0;
        ctl_next_trip(ctl);
        for (int i = 0; i < MAX_TRIPS; ++i) ctl->trip_searched[i] = 0;
        for (int i = 0; i < 23; ++i) ctl->dx_last[i] = 0.0;
        if (stats) { stats->far_queries = 0; stats->candidates = 0; stats->sweeps = 0; }
    }
}
void launch_ctl_reset(const double* x, double* x_prop, const double* P_prop, double* P_out, Ctl* ctl, SearchStats* stats, cudaStream_t stream) {
    k_ctl_reset<<<1, 256, 0, stream>>>(x, x_prop, P_prop, P_out, ctl, stats);
}

}  // namespace flivo
