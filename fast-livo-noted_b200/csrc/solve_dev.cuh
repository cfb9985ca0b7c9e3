// solve_dev.cuh — "(c) the 12x12 solve and state/covariance update kept on-device in double precision", as device
// functions.  The solve of a trip has no kernel of its own: it runs in the CTA of k_fused that folds the last group
// record (fused.cu), so a trip is k_search -> k_fused and the loop never leaves the device.  The same code runs from
// the small stand-alone kernels of solve.cu (NCCL variant of the sharded update, control-block reset).
//
//   * V3: one trip of esekf::update_iterated_dyn_share_modified
//         (reference include/IKFoM_toolkit/esekfom/esekfom.hpp:1619-1928) on the 23-DOF manifold state
//         (include/use-ikfom.hpp:12-21; SO3 / S2 algebra mtk/types/SOn.hpp:233-238, S2.hpp:128-280,
//         A_matrix mtk/src/mtkmath.hpp:236-247);
//   * V1: one trip of the loop body of LaserMapping::Run (src/laser_mapping.cpp:178-238) on the 18-DOF
//         StatesGroup (include/common_lib.h:552-587, Exp/Log include/so3_math.h:62-91);
//   * the re-match / stop policy, written to the device control block the next kernels read.
//
// Algebra note: the reference forms P_inv = ((P/R)^-1 + [HTH 0;0 0])^-1 with two dense
// n x n inverses and then only uses its first 12 columns.  By the matrix-inversion lemma those columns are
// W = Q[:, :12] (I + HTH Q[:12,:12])^-1 with Q = P/R: one 12x12 (V1: 6x6) solve, better conditioned, and
// identical in exact arithmetic.  Parity with the restated reference is tolerance-checked (dx <= 1e-6).
//
// Latency notes: the solve is one CTA of dependent FP64 work, so it is written for a short chain, not for
// throughput: Gauss-Jordan by ONE warp on the augmented matrix in shared memory (lane per column, no block
// barrier inside the 12 pivots), every 3x3 / S2 helper inlined (no stack frames), the state-only half of the NEXT
// trip (x (-) x_prop, manifold Jacobians, T P T^T) prepared at the end of this one.
#pragma once
#include "fused.cuh"
#include "solve.cuh"

namespace flivo {

#define SDEV static __device__ __forceinline__
// The solve is executed ONCE per launch by one CTA: its instruction lines are cold, and a helper shared by several call
// sites / several single-lane warps (SO(3) exp / log, A_matrix, the S2 charts with their sin / cos / atan2 bodies) is
// fetched once when it is a real function but once per copy when inlined — in theory.  Measured (profiles/r2_ab_solve.txt)
// the inlined build is the faster one (update phase 2.8 vs 3.4 us, whole solve 23-28 vs 27-34 us): default.
// -DFLIVO_SOLVE_INLINE=0 builds the out-of-line flavour.
#if !defined(FLIVO_SOLVE_INLINE) || FLIVO_SOLVE_INLINE
#define SHELP static __device__ __forceinline__
#else
#define SHELP static __device__ __noinline__
#endif

// ---- small dense helpers ------------------------------------------------------------------------------
SDEV void d_hat(const double* v, double* H) {
    H[0] = 0; H[1] = -v[2]; H[2] = v[1]; H[3] = v[2]; H[4] = 0; H[5] = -v[0]; H[6] = -v[1]; H[7] = v[0]; H[8] = 0;
}
SDEV void d_mat3mul(const double* A, const double* B, double* C) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}
// MTK::A_matrix (mtkmath.hpp:236-247)
SHELP void d_A_matrix(const double* v, double* A) {
    double sq = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    double nrm = sqrt(sq);
#pragma unroll
    for (int i = 0; i < 9; ++i) A[i] = 0.0;
    A[0] = A[4] = A[8] = 1.0;
    if (nrm < 1e-11) return;
    double H[9], H2[9];
    d_hat(v, H);
    d_mat3mul(H, H, H2);
    double a = (1 - cos(nrm)) / sq, b = (1 - sin(nrm) / nrm) / sq;
#pragma unroll
    for (int i = 0; i < 9; ++i) A[i] += a * H[i] + b * H2[i];
}
// rotation-matrix form of SO3::exp (SOn.hpp:293-297 via MTK::exp / cos_sinc_sqrt, mtkmath.hpp:139-171,249-256):
// Rodrigues with series near zero.  The reference composes unit quaternions; identical up to rounding.
SHELP void d_so3_exp(const double* v, double* R) {
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
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = a * K[i] + b * K2[i];
    R[0] += 1.0; R[4] += 1.0; R[8] += 1.0;
}
// SO3 log of a rotation matrix, accurate for small angles (SOn.hpp:300-304 / MTK::log computes the same
// vector from the unit quaternion: 2*atan(|v|/w)/|v| * v).
SHELP void d_so3_log(const double* R, double* o) {
    double vx = 0.5 * (R[7] - R[5]), vy = 0.5 * (R[2] - R[6]), vz = 0.5 * (R[3] - R[1]);  // sin(t) * axis
    double s = sqrt(vx * vx + vy * vy + vz * vz);
    double c = 0.5 * (R[0] + R[4] + R[8] - 1.0);
    double f;
    if (s < 1e-11) f = 1.0;
    else f = atan2(s, c) / s;
    o[0] = f * vx; o[1] = f * vy; o[2] = f * vz;
}
// V1 Exp / Log (include/so3_math.h:62-91)
SHELP void d_Exp_v1(double v1, double v2, double v3, double* R) {
    double nrm = sqrt(v1 * v1 + v2 * v2 + v3 * v3);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = 0;
    R[0] = R[4] = R[8] = 1;
    if (nrm > 0.00001) {
        double r[3] = {v1 / nrm, v2 / nrm, v3 / nrm};
        double K[9], K2[9];
        d_hat(r, K);
        d_mat3mul(K, K, K2);
        double s = sin(nrm), c = 1.0 - cos(nrm);
#pragma unroll
        for (int i = 0; i < 9; ++i) R[i] += s * K[i] + c * K2[i];
    }
}
SHELP void d_Log_v1(const double* R, double* o) {
    double tr = R[0] + R[4] + R[8];
    double theta = (tr > 3.0 - 1e-6) ? 0.0 : acos(0.5 * (tr - 1));
    double K[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]};
    double f = (fabs(theta) < 0.001) ? 0.5 : (0.5 * theta / sin(theta));
#pragma unroll
    for (int i = 0; i < 3; ++i) o[i] = f * K[i];
}

// ---- S2 (gravity) algebra, MTK::S2<double, 98090, 10000, 1> (use-ikfom.hpp:8; S2.hpp) ----------------------
#define S2_LEN (98090.0 / 10000.0)
SDEV void d_s2_Bx(const double* vec, double* Bx) {  // 3x2 row-major, S2_typ == 1 (S2.hpp:213-231)
    const double len = S2_LEN;
    if (vec[0] + len > 1e-11) {
        double d = len + vec[0];
        Bx[0] = -vec[1];                   Bx[1] = -vec[2];
        Bx[2] = len - vec[1] * vec[1] / d; Bx[3] = -vec[2] * vec[1] / d;
        Bx[4] = -vec[2] * vec[1] / d;      Bx[5] = len - vec[2] * vec[2] / d;
#pragma unroll
        for (int i = 0; i < 6; ++i) Bx[i] /= len;
    } else {
#pragma unroll
        for (int i = 0; i < 6; ++i) Bx[i] = 0.0;
        Bx[3] = -1; Bx[4] = 1;
    }
}
SDEV void d_cross(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
SHELP void d_s2_boxminus(const double* vec, const double* other, double* res2) {  // S2.hpp:136-160
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
#pragma unroll
        for (int c = 0; c < 2; ++c) res2[c] = f * (Bx[c] * hv[0] + Bx[2 + c] * hv[1] + Bx[4 + c] * hv[2]);
    }
}
SHELP void d_s2_boxplus(double* vec, const double* delta2) {  // S2.hpp:128-134
    double Bx[6], Bu[3], R[9], o[3];
    d_s2_Bx(vec, Bx);
#pragma unroll
    for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
    d_so3_exp(Bu, R);
#pragma unroll
    for (int r = 0; r < 3; ++r) o[r] = R[r * 3] * vec[0] + R[r * 3 + 1] * vec[1] + R[r * 3 + 2] * vec[2];
    vec[0] = o[0]; vec[1] = o[1]; vec[2] = o[2];
}
SHELP void d_s2_Nx_yy(const double* vec, double* Nx) {  // 2x3, S2.hpp:259-264
    double Bx[6], H[9];
    d_s2_Bx(vec, Bx);
    d_hat(vec, H);
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c)
            Nx[r * 3 + c] = 1 / S2_LEN / S2_LEN * (Bx[r] * H[c] + Bx[2 + r] * H[3 + c] + Bx[4 + r] * H[6 + c]);
}
// S2_Mx (S2.hpp:266-280).  "scalar(1/2)" there is integer division == 0, so exp_delta is the identity.
SHELP void d_s2_Mx(const double* vec, const double* delta2, double* Mx /*3x2*/) {
    double Bx[6], H[9];
    d_s2_Bx(vec, Bx);
    d_hat(vec, H);
    double dn = sqrt(delta2[0] * delta2[0] + delta2[1] * delta2[1]);
    if (dn < 1e-11) {
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 2; ++c) Mx[r * 2 + c] = -(H[r * 3] * Bx[c] + H[r * 3 + 1] * Bx[2 + c] + H[r * 3 + 2] * Bx[4 + c]);
    } else {
        double Bu[3], Am[9], T2[9];
#pragma unroll
        for (int r = 0; r < 3; ++r) Bu[r] = Bx[r * 2] * delta2[0] + Bx[r * 2 + 1] * delta2[1];
        d_A_matrix(Bu, Am);
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = 0; j < 3; ++j) T2[i * 3 + j] = H[i * 3] * Am[j * 3] + H[i * 3 + 1] * Am[j * 3 + 1] + H[i * 3 + 2] * Am[j * 3 + 2];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 2; ++c) Mx[r * 2 + c] = -(T2[r * 3] * Bx[c] + T2[r * 3 + 1] * Bx[2 + c] + T2[r * 3 + 2] * Bx[4 + c]);
    }
}

// state field offsets inside flivo_state (doubles)
enum { S_ROT = 0, S_POS = 9, S_OFFR = 12, S_OFFT = 21, S_VEL = 24, S_BG = 27, S_BA = 30, S_GRAV = 33 };

// R = A^T B for row-major 3x3 stored at a and b
SDEV void d_mat3_AtB(const double* a, const double* b, double* Rd) {
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) Rd[i * 3 + j] = a[i] * b[j] + a[3 + i] * b[3 + j] + a[6 + i] * b[6 + j];
}

// x (-) y, state_ikfom order pos, rot, offR, offT, vel, bg, ba, grav(S2).  The three parts touch disjoint
// fields (0: pos, rot; 1: offR, offT, vel, bg, ba; 2: grav) so three lanes of different warps run them at once.
SHELP void d_boxminus_v3_part(const double* x, const double* y, double* d, int part) {
    double Rd[9];
    if (part == 0) {
#pragma unroll
        for (int i = 0; i < 3; ++i) d[i] = x[S_POS + i] - y[S_POS + i];
        d_mat3_AtB(y + S_ROT, x + S_ROT, Rd);   // other.conjugate() * this  ->  y.rot^T * x.rot
        d_so3_log(Rd, d + 3);
    } else if (part == 1) {
        d_mat3_AtB(y + S_OFFR, x + S_OFFR, Rd);
        d_so3_log(Rd, d + 6);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            d[9 + i] = x[S_OFFT + i] - y[S_OFFT + i];
            d[12 + i] = x[S_VEL + i] - y[S_VEL + i];
            d[15 + i] = x[S_BG + i] - y[S_BG + i];
            d[18 + i] = x[S_BA + i] - y[S_BA + i];
        }
    } else {
        d_s2_boxminus(x + S_GRAV, y + S_GRAV, d + 21);
    }
}
SHELP void d_boxplus_v3_part(double* x, const double* d, int part) {
    double E[9], Rn[9];
    if (part == 0) {
#pragma unroll
        for (int i = 0; i < 3; ++i) x[S_POS + i] += d[i];
        d_so3_exp(d + 3, E);
        d_mat3mul(x + S_ROT, E, Rn);
#pragma unroll
        for (int i = 0; i < 9; ++i) x[S_ROT + i] = Rn[i];
    } else if (part == 1) {
        d_so3_exp(d + 6, E);
        d_mat3mul(x + S_OFFR, E, Rn);
#pragma unroll
        for (int i = 0; i < 9; ++i) x[S_OFFR + i] = Rn[i];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            x[S_OFFT + i] += d[9 + i];
            x[S_VEL + i] += d[12 + i];
            x[S_BG + i] += d[15 + i];
            x[S_BA + i] += d[18 + i];
        }
    } else {
        d_s2_boxplus(x + S_GRAV, d + 21);
    }
}
// J = A_matrix(v)^T and seg <- J * seg  (esekfom.hpp:1663-1672)
SHELP void d_so3_jac(const double* v, double* J, double* seg) {
    double A[9];
    d_A_matrix(v, A);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) J[i * 3 + j] = A[j * 3 + i];
    if (seg) {
        double t0 = J[0] * seg[0] + J[1] * seg[1] + J[2] * seg[2];
        double t1 = J[3] * seg[0] + J[4] * seg[1] + J[5] * seg[2];
        double t2 = J[6] * seg[0] + J[7] * seg[1] + J[8] * seg[2];
        seg[0] = t0; seg[1] = t1; seg[2] = t2;
    }
}
// J2 = Nx(x.grav) * Mx(x_prop.grav, d2) and seg <- J2 * seg  (esekfom.hpp:1684-1695)
SHELP void d_s2_jac(const double* xg, const double* xpg, const double* d2, double* J2, double* seg) {
    double Nx[6], Mx[6];
    d_s2_Nx_yy(xg, Nx);
    d_s2_Mx(xpg, d2, Mx);
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int c = 0; c < 2; ++c) J2[r * 2 + c] = Nx[r * 3] * Mx[c] + Nx[r * 3 + 1] * Mx[2 + c] + Nx[r * 3 + 2] * Mx[4 + c];
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
SDEV void apply_T_rows(double* M, int n, int ncols, const BlockT& T, int tid, int nthreads) {
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
SDEV void apply_T_cols(double* M, int n, const BlockT& T, int tid, int nthreads) {
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

// Solve S Y = B by Gauss-Jordan with partial pivoting on the M x W array A = [S | B] (W = M + nrhs <= 32) in shared
// memory, by ONE warp with the matrix in REGISTERS: lane c holds column c (M doubles).  On return columns M..W-1 of A
// hold Y.  The routine is a chain of dependent instructions executed by a single warp, so what counts is the number of
// instructions per pivot, not parallel width (a first version that kept the matrix in shared memory, 13 elements per
// lane and pivot, took 15.8 us for 12 pivots on B200; the block-wide version of round 1, one barrier per pivot, 6.7 us):
//   * pivot k: lane k finds the pivot row of its own column among the rows not yet used — INTEGER compares of the
//     magnitudes' bit patterns (non-negative doubles order like their bits), two interleaved half-chains; first
//     maximum wins ties — and the warp learns the row by one shuffle;
//   * rows are never swapped: the pivot row stays where it is and is marked used; pos[i] remembers which unknown row i
//     ended up solving, and the final store puts it there;
//   * every lane fetches column k (the multipliers) by M shuffles, scales its pivot-row entry by ONE reciprocal
//     (correctly rounded) and updates its M - 1 other entries.
// Same operations per entry as the textbook order.  Returns false on a zero pivot (the caller then leaves state and
// covariance untouched).
SDEV unsigned long long d_mag_bits(double v) {
#ifdef __CUDA_ARCH__
    return (unsigned long long)__double_as_longlong(v) & 0x7FFFFFFFFFFFFFFFull;
#else
    unsigned long long b;
    memcpy(&b, &v, 8);
    return b & 0x7FFFFFFFFFFFFFFFull;
#endif
}
SDEV double d_rcp(double v) {
#ifdef __CUDA_ARCH__
    return __drcp_rn(v);
#else
    return 1.0 / v;
#endif
}
template <int M, int W>
SDEV bool warp_gauss_jordan(double* A, int lane) {
    static_assert(W <= 32 && M >= 2 && M <= 16, "one column per lane");
    constexpr unsigned FULLW = 0xFFFFFFFFu;
    constexpr int HALF = M / 2;
    double a[M];
    int pos[M];
#pragma unroll
    for (int i = 0; i < M; ++i) { a[i] = lane < W ? A[i * W + lane] : 0.0; pos[i] = 0; }
    unsigned used = 0u;
    bool ok = true;
#pragma unroll 1
    for (int k = 0; k < M; ++k) {
        // pivot row of column k (computed by every lane on its own column; lane k's answer counts).  The magnitudes are
        // compared by the HIGH words of their bit patterns (sign stripped): a pivot within 2^-20 of the largest entry is as
        // good as the largest; the value travels with the key so that the reciprocal can start as soon as the row is known
        unsigned ba = 0u, bb = 0u;
        int pa = __ffs(~used) - 1, pb = pa;   // (a row that is still free, should every candidate be zero)
        double va = 0.0, vb = 0.0;
#pragma unroll
        for (int i = 0; i < HALF; ++i) {
            const unsigned key = ((used >> i) & 1u) ? 0u : (unsigned)(d_mag_bits(a[i]) >> 32);
            if (key > ba) { ba = key; pa = i; va = a[i]; }
        }
#pragma unroll
        for (int i = HALF; i < M; ++i) {
            const unsigned key = ((used >> i) & 1u) ? 0u : (unsigned)(d_mag_bits(a[i]) >> 32);
            if (key > bb) { bb = key; pb = i; vb = a[i]; }
        }
        const bool second = bb > ba;
        const int pz = __shfl_sync(FULLW, (second ? pb : pa) | (((second ? bb : ba) == 0u) ? 256 : 0), k);
        const double r = d_rcp(__shfl_sync(FULLW, second ? vb : va, k));
        const int p = pz & 255;
        if (pz & 256) ok = false;
        used |= 1u << p;
        double ap = a[0];                      // this lane's entry of the pivot row (selected while the reciprocal runs)
#pragma unroll
        for (int i = 1; i < M; ++i) if (i == p) ap = a[i];
        const double prow = ap * r;
#pragma unroll
        for (int i = 0; i < M; ++i) {
            const double m = __shfl_sync(FULLW, a[i], k);   // column k, row i
            if (i == p) { a[i] = prow; pos[i] = k; }
            else a[i] = a[i] - m * prow;
        }
    }
    if (lane >= M && lane < W) {
#pragma unroll
        for (int i = 0; i < M; ++i) A[pos[i] * W + lane] = a[i];
    }
    __syncwarp();
    return ok;
}

// shared-memory workspace of one solve (aliased onto k_fused's Gram staging buffers in the CTA that runs it)
struct SolveSmem {
    double sum[PSTRIDE];          // reduced normal equations: 78 upper H^T H, 12 H^T z, res_sum, n_eff
    double H[144], z[12];
    double P[23 * 23], L[23 * 23];
    double Aug[12 * 25];          // [S | H | z] of the trip's 12x12 system (V1: 6 x 13)
    double Kx[23 * 12];
    double x[36], xp[36];
    double dxnew[24], dxo[24], Kh[24], dx[24];
    double H6[36], z6[6], vec[18], sol[18], G[18 * 6];
    BlockT T;
    int flag[4];                  // [0] = stop, [1] = gj ok
    unsigned long long seq;
    int timeout;
};

SDEV void solve_stamp(const SolveParams& S, int trip, int k) {
#ifndef FLIVO_HOST_EMU
    if (S.stamps && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        trip = trip < 0 ? 0 : (trip < MAX_TRIPS ? trip : MAX_TRIPS - 1);
        S.stamps[trip * 16 + k] = t;
    }
#endif
}

// The state and the control block of the trip are final: publish them (release) and let the next trip's k_search start
// while this CTA goes on with the covariance update or the next trip's state-only half.  k_search, launched as a
// programmatic dependent of this grid, waits for ctl->ready (acquire) before it reads either (fused.cu).
SDEV void solve_publish(Ctl* ctl, const Job* job, int tid) {
#ifndef FLIVO_HOST_EMU
    __syncthreads();                      // x (36 threads) and the control block (decision lane) are written
    if (tid == 0) {
        __threadfence();
        const unsigned v = (job->useq << 8) | ((unsigned)ctl->trips & 255u);
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(&ctl->ready), "r"(v) : "memory");
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

// ---- state-only half of the NEXT trip's solve: needs x and P_prop, not H^T H -------------------------------------
// V3: dx = x (-) x_prop, the SO(3) / S2 Jacobians and T P T^T (esekfom.hpp:1646-1695) -> S.pre->{Pt, dxnew};
// V1: vec = state_propagat - state (common_lib.h:576-587) -> S.pre->vec.  M.x / M.xp hold the two states.
// All threads of the CTA call it (NT >= 96); ends with a barrier.
// p_loaded: M.L already holds P_prop (solve_trip fetches it at its very start, long before it gets here).
static __device__ __noinline__ void solve_pre(const SolveParams& S, SolveSmem& M, int mode, int tid, int NT, bool p_loaded) {
    if (mode == 0) {
        if (!p_loaded) for (int e = tid; e < 23 * 23; e += NT) M.L[e] = S.P_prop[e];
        if ((tid & 31) == 0 && tid < 96) {
            const int part = tid >> 5;                                        // three warps, one lane each
            d_boxminus_v3_part(M.x, M.xp, M.dx, part);                        // esekfom.hpp:1646
            if (part == 0) {
                for (int i = 0; i < 6; ++i) M.dxnew[i] = M.dx[i];
                d_so3_jac(M.dx + 3, M.T.J3, M.dxnew + 3);                     // :1663
            } else if (part == 1) {
                for (int i = 6; i < 21; ++i) M.dxnew[i] = M.dx[i];
                d_so3_jac(M.dx + 6, M.T.J6, M.dxnew + 6);
            } else {
                M.dxnew[21] = M.dx[21]; M.dxnew[22] = M.dx[22];
                d_s2_jac(M.x + S_GRAV, M.xp + S_GRAV, M.dx + 21, M.T.J2, M.dxnew + 21);   // :1684-1686
            }
        }
        __syncthreads();
        apply_T_rows(M.L, 23, 23, M.T, tid, NT);                              // :1665-1672, :1688-1695
        __syncthreads();
        apply_T_cols(M.L, 23, M.T, tid, NT);
        __syncthreads();
        for (int e = tid; e < 23 * 23; e += NT) S.pre->Pt[e] = M.L[e];
        if (tid < 23) S.pre->dxnew[tid] = M.dxnew[tid];
    } else if (tid == 0) {
        double rotd[9], vec[18];
        d_mat3_AtB(M.x + S_ROT, M.xp + S_ROT, rotd);
        d_Log_v1(rotd, vec);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            vec[3 + i] = M.xp[S_POS + i] - M.x[S_POS + i];
            vec[6 + i] = M.xp[S_VEL + i] - M.x[S_VEL + i];
            vec[9 + i] = M.xp[S_BG + i] - M.x[S_BG + i];
            vec[12 + i] = M.xp[S_BA + i] - M.x[S_BA + i];
            vec[15 + i] = M.xp[S_GRAV + i] - M.x[S_GRAV + i];
        }
        for (int i = 0; i < 18; ++i) S.pre->vec[i] = vec[i];
    }
    __syncthreads();
}

// ---- one trip's solve ------------------------------------------------------------------------------------------------
// On entry M.sum[0..NTERMS) holds THIS rank's reduced sums (phase FROM_RANKS: ignored, the gathered rank blocks are
// summed instead) and the CTA is past a barrier.  NT >= 128 threads (the decision warp is warp 3).
static __device__ __noinline__ void solve_trip(const SolveParams& S, SolveSmem& M, int tid, int NT) {
    Ctl* ctl = S.ctl;
    const int lane = tid & 31, warp = tid >> 5;
    // the state, P and the state-only half prepared at the end of the previous trip do not depend on the sums: their
    // global loads are issued first and land while the exchange / unpacking runs
    const int mode = S.job->mode;
    const int n = (mode == 1) ? 18 : 23;
    int c_t = 0, c_iter = 0, c_trips = 0, c_search = 0, c_searches = 0, c_rematch = 0, j_max_iter = 0, j_force = 0;
    double r_lim = 0.0;
    if (S.phase != SOLVE_PHASE_REDUCE_ONLY) {
        if (tid < 36) { M.x[tid] = S.x[tid]; M.xp[tid] = S.x_prop[tid]; }
        const double* Psrc = (mode == 1) ? S.P_prop : S.pre->Pt;
        for (int e = tid; e < n * n; e += NT) M.P[e] = Psrc[e];
        if (mode == 0) {
            if (tid < 23) M.dxnew[tid] = S.pre->dxnew[tid];
            for (int e = tid; e < 23 * 23; e += NT) M.L[e] = S.P_prop[e];   // for solve_pre at the end of this trip
        }
        else if (tid < 18) M.vec[tid] = S.pre->vec[tid];
        if (warp == 3 && lane < 23) r_lim = S.job->limit[lane];
        if (tid == 96 || tid == 0) {   // lane 0 of the decision warp (V3) / thread 0 (V1)
            c_t = ctl->t; c_iter = ctl->iter; c_trips = ctl->trips; c_search = ctl->search; c_searches = ctl->searches;
            c_rematch = ctl->rematch;
            j_max_iter = S.job->max_iter; j_force = S.job->force_search;
        }
    }
    const int trip_no = S.stamps ? ctl->trips : 0;
    solve_stamp(S, trip_no, 0);

    // ---- 0. normal equations of all ranks ----------------------------------------------------------------------
#ifndef FLIVO_HOST_EMU
    if (S.phase == SOLVE_PHASE_PEER) {
        // fused reduce + exchange over peer memory: every rank holds the same state, so all ranks run the same trips
        // and their exchange counters agree.  My record goes straight into every rank's buffer (NVLink stores), then a
        // sequence number (release); the peers' records are awaited in my own buffer (acquire) and summed in rank order.
        if (tid == 0) { M.seq = *S.peer.seq + 1ull; M.timeout = 0; }
        __syncthreads();
        const unsigned long long seq = M.seq;
        const int par = (int)(seq & 1ull), me = S.peer.rank, nr = S.peer.nranks;
        for (int e = tid; e < nr * PSTRIDE; e += NT) {
            const int r = e / PSTRIDE, t = e - r * PSTRIDE;
            S.peer.base[r][par * PEER_MAX_RANKS + me].rec[t] = (t < NTERMS) ? M.sum[t] : 0.0;
        }
        __threadfence_system();
        __syncthreads();
        solve_stamp(S, trip_no, 10);
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
                if (t1 - t0 > S.peer.timeout_ns) { M.timeout = 1; break; }
            }
        }
        __syncthreads();
        solve_stamp(S, trip_no, 11);
        if (tid < NTERMS) {                                             // fixed rank order: bit-identical on all ranks
            double s = 0.0;
            for (int r = 0; r < nr; ++r) {
                double v;
                asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(&S.peer.base[me][par * PEER_MAX_RANKS + r].rec[tid]) : "memory");
                s += v;
            }
            M.sum[tid] = s;
        }
        // a timed-out exchange is not committed: the ranks are out of step and stay stopped until the caller
        // re-synchronises them (flivo_peer_reset after a barrier); nan_guard = 2 tells the host
        if (tid == 0 && !M.timeout) *S.peer.seq = seq;
        __syncthreads();
        if (M.timeout) {
            if (tid == 0) { ctl->nan_guard = 2; ctl->stopped = 1; ctl->trips += 1; ctl_next_trip(ctl); }
            return;
        }
    } else
#endif
    if (S.phase == SOLVE_PHASE_FROM_RANKS) {
        // multi-GPU, NCCL variant: every rank's 92-term block was gathered; sum in rank order
        if (tid < NTERMS) {
            double s = 0.0;
            for (int r = 0; r < S.job->nranks; ++r) s += S.rank_blocks[(size_t)r * PSTRIDE + tid];
            M.sum[tid] = s;
        }
        __syncthreads();
    }
    solve_stamp(S, trip_no, 1);
    if (S.sum_out && tid < NTERMS) S.sum_out[tid] = M.sum[tid];
    if (S.phase == SOLVE_PHASE_REDUCE_ONLY) {
        if (tid == 0) ctl_next_trip(ctl);
        return;
    }
    for (int e = tid; e < 144; e += NT) {   // (NT = 128 < 144: a plain `tid < 144` would leave the last 16 entries stale)
        const int a = e / 12, b = e % 12;
        const int i = a < b ? a : b, j = a < b ? b : a;
        M.H[e] = M.sum[i * 12 - (i * (i - 1)) / 2 + (j - i)];   // index of (i, j), i <= j, in the 78 upper terms
    }
    if (tid < 12) M.z[tid] = M.sum[78 + tid];
    if (tid == 0) {
        const int ne = (int)M.sum[91];
        ctl->n_eff = ne;
        ctl->res_mean = ne > 0 ? M.sum[90] / (double)ne : 0.0;
    }
    const int all_finite = __syncthreads_and(tid >= NTERMS || isfinite(M.sum[tid]));
    if (!all_finite) {  // NaN guard: leave x and P untouched and stop the loop
        if (tid == 0) { ctl->nan_guard = 1; ctl->stopped = 1; ctl->trips += 1; ctl_next_trip(ctl); }
        return;
    }
    solve_stamp(S, trip_no, 2);

    if (mode == 0) {
        // ================= V3: esekfom.hpp:1633-1927 ==================================================
        // The gain needs W = Q1 S^-1 (23 x 12), S = I + HTH * Q11, Q = P / R, only through K_x = W HTH and K_h = W HTz:
        // solve S Y = [HTH | HTz] (13 right-hand sides: the 12 x 25 system fits one warp, a column per lane) and form
        // K_x | K_h = Q1 Y afterwards with all threads.  Aug = [S | HTH | HTz].
        const double invR = 1.0 / S.job->R;
        for (int e = tid; e < 12 * 25; e += NT) {
            const int r = e / 25, c = e % 25;
            double v;
            if (c < 12) {  // S[r][c] = delta + sum_k H[r][k] Q[k][c]
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 12; ++k) s += M.H[r * 12 + k] * (M.P[k * 23 + c] * invR);
                v = s + (r == c ? 1.0 : 0.0);
            } else {
                v = c < 24 ? M.H[r * 12 + (c - 12)] : M.z[r];
            }
            M.Aug[e] = v;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 3);
        if (warp == 0) {
            const bool ok = warp_gauss_jordan<12, 25>(M.Aug, lane);
            if (lane == 0) M.flag[1] = ok ? 1 : 0;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 4);
        if (!M.flag[1]) {   // singular system (cannot happen for SPD H and P; NaN guard of SURVEY §5): x and P untouched
            if (tid == 0) { ctl->nan_guard = 1; ctl->stopped = 1; ctl->trips += 1; ctl_next_trip(ctl); }
            return;
        }
        // Y = Aug[:, 12:25];  K_x[:, :12] = Q1 Y[:, :12] (:1806);  K_h = Q1 Y[:, 12] (:1801)
        for (int e = tid; e < 23 * 13; e += NT) {
            const int i = e / 13, b = e % 13;
            double s = 0.0;
#pragma unroll
            for (int a = 0; a < 12; ++a) s += (M.P[i * 23 + a] * invR) * M.Aug[a * 25 + 12 + b];
            if (b < 12) M.Kx[i * 12 + b] = s;
            else M.Kh[i] = s;
        }
        __syncthreads();
        if (tid < 23) {                                                       // :1812
            double s = M.Kh[tid];
#pragma unroll
            for (int j = 0; j < 12; ++j) s += M.Kx[tid * 12 + j] * M.dxnew[j];
            M.dxo[tid] = s - M.dxnew[tid];
        }
        __syncthreads();
        solve_stamp(S, trip_no, 5);
        if (lane == 0 && warp < 3) d_boxplus_v3_part(M.x, M.dxo, warp);       // :1814
        if (warp == 3) {
            // esekfom.hpp:1816-1823: every component of dx below its limit (one lane each; the limits were loaded up front)
            int converge = __all_sync(0xffffffffu, lane >= 23 || !(fabs(M.dxo[lane < 23 ? lane : 0]) > r_lim));
            if (lane < 23) ctl->dx_last[lane] = M.dxo[lane];
            if (lane == 0) {
                int t = c_t;
                const int it = c_iter;
                if (converge) t++;
                if (!t && it == j_max_iter - 2) converge = 1;                         // :1825-1828
                const int stop = (t > 1 || it == j_max_iter - 1) ? 1 : 0;              // :1831
                ctl->t = t;
                const int trip = c_trips;
                if (trip < MAX_TRIPS) ctl->trip_searched[trip] = (c_search || j_force) ? 1 : 0;
                ctl->searches = c_searches + ((c_search || j_force) ? 1 : 0);
                ctl->trips = trip + 1;
                ctl->search = converge;
                ctl->iter = it + 1;
                ctl->converged = t > 0 ? 1 : 0;
                ctl->stopped = stop;
                ctl_next_trip(ctl);
                M.flag[0] = stop;
            }
        }
        __syncthreads();
        solve_stamp(S, trip_no, 6);
        if (tid < 36) S.x[tid] = M.x[tid];
        solve_publish(ctl, S.job, tid);
        if (M.flag[0]) {
            if (lane == 0 && warp < 3) {
                if (warp == 0) d_so3_jac(M.dxo + 3, M.T.J3, nullptr);             // :1842
                else if (warp == 1) d_so3_jac(M.dxo + 6, M.T.J6, nullptr);
                else d_s2_jac(M.x + S_GRAV, M.xp + S_GRAV, M.dxo + 21, M.T.J2, nullptr);   // :1877-1879 (x_ already updated)
            }
            __syncthreads();
            // L = T' P T'^T, P <- P T'^T, K_x <- T' K_x, P_out = L - K_x[:, :12] P[:12, :]   (:1833-1921)
            apply_T_cols(M.P, 23, M.T, tid, NT);
            __syncthreads();
            for (int e = tid; e < 23 * 23; e += NT) M.L[e] = M.P[e];
            __syncthreads();
            apply_T_rows(M.L, 23, 23, M.T, tid, NT);
            apply_T_rows(M.Kx, 12, 12, M.T, tid, NT);
            __syncthreads();
            for (int e = tid; e < 23 * 23; e += NT) {
                const int i = e / 23, j = e % 23;
                double s = 0.0;
#pragma unroll
                for (int a = 0; a < 12; ++a) s += M.Kx[i * 12 + a] * M.P[a * 23 + j];
                S.P_out[e] = M.L[e] - s;
            }
            solve_stamp(S, trip_no, 7);
        } else {
            solve_pre(S, M, 0, tid, NT, true);     // the next trip's state-only half, while nothing else can run anyway
            solve_stamp(S, trip_no, 7);
        }
    } else {
        // ================= V1: src/laser_mapping.cpp:178-238 ==========================================
        // HPH (6x6) = Hsub^T (1/R) Hsub with Hsub = [A | n] = V3 columns [3:6 | 0:3]  (:584-586)
        const double w = 1.0 / S.job->R;
        if (tid < 36) {
            const int a = tid / 6, b = tid % 6;
            const int ma = a < 3 ? a + 3 : a - 3, mb = b < 3 ? b + 3 : b - 3;
            M.H6[tid] = M.H[ma * 12 + mb] * w;
            if (tid < 6) M.z6[tid] = M.z[tid < 3 ? tid + 3 : tid - 3] * w;
        }
        __syncthreads();
        if (S.job->gnss_on) {
            // HPH.block<6,6>(0,0) += gnss_HPH; HPL.block<6,1>(0,0) += gnss_HPL (src/laser_mapping.cpp:612-617) with
            // computeH (src/GNSS_Processing.cpp:204-237) evaluated at the current iterate: Hsub = [R [l]x | -I] (3x6),
            // error = p_gnss - pos + R l, dropped when |error| > 2 m, weights diag(1/std0, 1/std1, 100/std2)
            if (tid == 0) {
                const double* R = M.x + S_ROT;
                const double* l = S.job->gnss_lever;
                const double K[9] = {0.0, -l[2], l[1], l[2], 0.0, -l[0], -l[1], l[0], 0.0};   // SKEW_SYM_MATRX (so3_math.h:9)
                double Hs[18], RK[9], err[3];
                d_mat3mul(R, K, RK);
                for (int r = 0; r < 3; ++r) {
                    for (int c = 0; c < 3; ++c) { Hs[r * 6 + c] = RK[r * 3 + c]; Hs[r * 6 + 3 + c] = (r == c) ? -1.0 : 0.0; }
                    err[r] = S.job->gnss_pos[r] - M.x[S_POS + r] + (R[r * 3] * l[0] + R[r * 3 + 1] * l[1] + R[r * 3 + 2] * l[2]);
                }
                if (!(sqrt(err[0] * err[0] + err[1] * err[1] + err[2] * err[2]) > 2.0)) {
                    const double wg[3] = {1.0 / S.job->gnss_std[0], 1.0 / S.job->gnss_std[1], 100.0 / S.job->gnss_std[2]};
                    for (int a = 0; a < 6; ++a) {
                        for (int b = 0; b < 6; ++b) {
                            double acc = 0.0;
                            for (int k = 0; k < 3; ++k) acc += (Hs[k * 6 + a] * wg[k]) * Hs[k * 6 + b];
                            M.H6[a * 6 + b] += acc;
                        }
                        double acc = 0.0;
                        for (int k = 0; k < 3; ++k) acc += (Hs[k * 6 + a] * wg[k]) * err[k];
                        M.z6[a] += acc;
                    }
                }
            }
            __syncthreads();
        }
        // K_1[:, :6] = C1 S^-1 with S = I + H6 * C11, C = state.cov, enters only through G = K_1 HPH and K_1 HPL:
        // solve S Y = [H6 | z6] (6 x 13 system), then G | K_1 HPL = C1 Y.  Aug = [S | H6 | z6].
        for (int e = tid; e < 6 * 13; e += NT) {
            const int r = e / 13, c = e % 13;
            double v;
            if (c < 6) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < 6; ++k) s += M.H6[r * 6 + k] * M.P[k * 18 + c];
                v = s + (r == c ? 1.0 : 0.0);
            } else {
                v = c < 12 ? M.H6[r * 6 + (c - 6)] : M.z6[r];
            }
            M.Aug[e] = v;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 3);
        if (warp == 0) {
            const bool ok = warp_gauss_jordan<6, 13>(M.Aug, lane);
            if (lane == 0) M.flag[1] = ok ? 1 : 0;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 4);
        if (!M.flag[1]) {
            if (tid == 0) { ctl->nan_guard = 1; ctl->stopped = 1; ctl->trips += 1; ctl_next_trip(ctl); }
            return;
        }
        for (int e = tid; e < 18 * 7; e += NT) {                             // G[:, :6] = K_1[:, :6] * HPH (:191);  Kh = K_1[:, :6] * HPL
            const int i = e / 7, b = e % 7;
            double s = 0.0;
#pragma unroll
            for (int a = 0; a < 6; ++a) s += M.P[i * 18 + a] * M.Aug[a * 13 + 6 + b];
            if (b < 6) M.G[i * 6 + b] = s;
            else M.Kh[i] = s;
        }
        __syncthreads();
        if (tid < 18) {                                                      // :193
            double s2 = 0.0;
#pragma unroll
            for (int a = 0; a < 6; ++a) s2 += M.G[tid * 6 + a] * M.vec[a];
            M.sol[tid] = M.Kh[tid] + M.vec[tid] - s2;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 5);
        if (tid == 0) {
            double E[9], Rn[9];
            d_Exp_v1(M.sol[0], M.sol[1], M.sol[2], E);                        // state += solution (:204)
            d_mat3mul(M.x + S_ROT, E, Rn);
            for (int i = 0; i < 9; ++i) M.x[S_ROT + i] = Rn[i];
            for (int i = 0; i < 3; ++i) {
                M.x[S_POS + i] += M.sol[3 + i]; M.x[S_VEL + i] += M.sol[6 + i]; M.x[S_BG + i] += M.sol[9 + i];
                M.x[S_BA + i] += M.sol[12 + i]; M.x[S_GRAV + i] += M.sol[15 + i];
            }
            const double R2D = 180.0 / 3.14159265358979323846;
            const double rn = sqrt(M.sol[0] * M.sol[0] + M.sol[1] * M.sol[1] + M.sol[2] * M.sol[2]);
            const double tn = sqrt(M.sol[3] * M.sol[3] + M.sol[4] * M.sol[4] + M.sol[5] * M.sol[5]);
            const int converged = ((rn * R2D < 0.01) && (tn * 100 < 0.015)) ? 1 : 0;   // :209-212
            const int it = c_iter;
            int rematch = c_rematch, nse = 0;
            if (converged || ((rematch == 0) && (it == (j_max_iter - 2)))) { nse = 1; rematch++; }   // :217-222
            const int stop = (rematch >= 2 || (it == j_max_iter - 1)) ? 1 : 0;          // :224
            const int trip = c_trips;
            if (trip < MAX_TRIPS) ctl->trip_searched[trip] = (c_search || j_force) ? 1 : 0;
            ctl->searches = c_searches + ((c_search || j_force) ? 1 : 0);
            ctl->trips = trip + 1;
            ctl->rematch = rematch;
            ctl->search = nse;
            ctl->iter = it + 1;
            ctl->converged = converged;
            ctl->stopped = stop;
            ctl_next_trip(ctl);
            for (int i = 0; i < 18; ++i) ctl->dx_last[i] = M.sol[i];
            M.flag[0] = stop;
        }
        __syncthreads();
        solve_stamp(S, trip_no, 6);
        if (tid < 36) S.x[tid] = M.x[tid];
        solve_publish(ctl, S.job, tid);
        if (M.flag[0]) {                                                     // state.cov = (I - G) * state.cov (:229)
            for (int e = tid; e < 18 * 18; e += NT) {
                const int i = e / 18, j = e % 18;
                double s = 0.0;
#pragma unroll
                for (int a = 0; a < 6; ++a) s += M.G[i * 6 + a] * M.P[a * 18 + j];
                S.P_out[e] = M.P[e] - s;
            }
        } else {
            solve_pre(S, M, 1, tid, NT, false);
        }
        solve_stamp(S, trip_no, 7);
    }
}

#undef SDEV
#undef SHELP
}  // namespace flivo
