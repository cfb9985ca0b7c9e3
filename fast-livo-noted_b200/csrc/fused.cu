// fused.cu — the per-trip fused kernel of the LIO measurement update (sm_100a) and the batched kNN kernel.
//
// One launch = one LaserMapping::h_share_model call (reference src/laser_mapping.cpp:485-645 and
// src/origin_laserMapping.cpp:916-1048):
//   a1 world transform (:662-671)            -> per lane, double, narrowed to float
//   a2 exact 5-NN (ikd_Tree.cpp:350-380)     -> thread per point over the 27-cell shell of the voxel-hash map,
//                                               warp-cooperative search for the rare far queries (search.cuh)
//   a3 far-neighbour gate (:518)             -> per lane
//   a4 esti_plane (common_lib.h:670-702)     -> per lane, float, no FMA, fixed operation order
//   a5 residual + range gate (:532-543)      -> per lane
//   a6 compaction (:547-561)                 -> replaced by masked accumulation; mask and n_eff are outputs
//   a7 1x12 Jacobian row (origin:1018-1044)  -> per lane, double
//   a8 H^T H / H^T z (esekfom.hpp:1781,1801) -> FP64 tensor-core MMA (DMMA.8x8x4) Gram in registers per warp, fixed-order
//                                               CTA and final reductions (no floating-point atomics)
// Compiled with -fmad=false: see flivo_dev.cuh.
#include "fused.cuh"
#include "search.cuh"
#include "solve_dev.cuh"

namespace flivo {

// (a, b) of the 78 upper-triangle terms, then (a, 12) for H^T z
__constant__ unsigned char c_term_a[90];
__constant__ unsigned char c_term_b[90];

void upload_term_table() {
    unsigned char a[90], b[90];
    int t = 0;
    for (int i = 0; i < 12; ++i)
        for (int j = i; j < 12; ++j) { a[t] = (unsigned char)i; b[t] = (unsigned char)j; ++t; }
    for (int i = 0; i < 12; ++i) { a[t] = (unsigned char)i; b[t] = 12; ++t; }
    cudaMemcpyToSymbol(c_term_a, a, 90);
    cudaMemcpyToSymbol(c_term_b, b, 90);
}

// ---- esti_plane<float> -----------------------------------------------------------------------------
// Same operation order as the CPU oracle's restatement of Eigen's ColPivHouseholderQR:
// 5-element reductions left to right, 3-element ones as x0 + (x1 + x2).  A is column-major A[c][r].
__device__ __forceinline__ float norm_tail5(const float* col, int from) {
    float s = col[from] * col[from];
    for (int i = from + 1; i < 5; ++i) s = s + col[i] * col[i];
    return sqrtf(s);
}

__device__ bool esti_plane_dev(const float4 nb[5], float threshold, float pabcd[4]) {
    float A[3][5];
#pragma unroll
    for (int r = 0; r < 5; ++r) { A[0][r] = nb[r].x; A[1][r] = nb[r].y; A[2][r] = nb[r].z; }
    float cvec[5] = {-1.f, -1.f, -1.f, -1.f, -1.f};
    float hcoef[3] = {0.f, 0.f, 0.f};
    int perm[3] = {0, 1, 2};
    float nrmU[3], nrmD[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) { nrmD[c] = norm_tail5(A[c], 0); nrmU[c] = nrmD[c]; }
    float maxn = nrmU[0];
    if (nrmU[1] > maxn) maxn = nrmU[1];
    if (nrmU[2] > maxn) maxn = nrmU[2];
    const float eps = 1.1920928955078125e-07f;  // FLT_EPSILON
    float th = maxn * eps;
    const float threshold_helper = (th * th) / 5.0f;
    const float downdate_thr = sqrtf(eps);
    int nonzero_pivots = 3;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        int big = k;
#pragma unroll
        for (int j = k + 1; j < 3; ++j)
            if (nrmU[j] > nrmU[big]) big = j;
        float big_sq = nrmU[big] * nrmU[big];
        if (nonzero_pivots == 3 && big_sq < threshold_helper * (float)(5 - k)) nonzero_pivots = k;
        if (big != k) {
            // big in {k+1, .., 2}: swap with compile-time indices to keep everything in registers
#pragma unroll
            for (int j = k + 1; j < 3; ++j) {
                if (j == big) {
#pragma unroll
                    for (int r = 0; r < 5; ++r) { float t = A[k][r]; A[k][r] = A[j][r]; A[j][r] = t; }
                    float t = nrmU[k]; nrmU[k] = nrmU[j]; nrmU[j] = t;
                    t = nrmD[k]; nrmD[k] = nrmD[j]; nrmD[j] = t;
                    int ti = perm[k]; perm[k] = perm[j]; perm[j] = ti;
                }
            }
        }
        float tailSq = 0.f;
        if (k + 1 < 5) {
            tailSq = A[k][k + 1] * A[k][k + 1];
#pragma unroll
            for (int i = k + 2; i < 5; ++i) tailSq = tailSq + A[k][i] * A[k][i];
        }
        float c0 = A[k][k], beta, tau;
        if (tailSq <= 1.17549435e-38f) {  // FLT_MIN
            tau = 0.f; beta = c0;
#pragma unroll
            for (int i = k + 1; i < 5; ++i) A[k][i] = 0.f;
        } else {
            beta = sqrtf(c0 * c0 + tailSq);
            if (c0 >= 0.f) beta = -beta;
            float den = c0 - beta;
#pragma unroll
            for (int i = k + 1; i < 5; ++i) A[k][i] = A[k][i] / den;
            tau = (beta - c0) / beta;
        }
        A[k][k] = beta;
        hcoef[k] = tau;
        if (tau != 0.f) {
#pragma unroll
            for (int j = k + 1; j < 3; ++j) {
                float tmp = A[k][k + 1] * A[j][k + 1];
#pragma unroll
                for (int i = k + 2; i < 5; ++i) tmp = tmp + A[k][i] * A[j][i];
                tmp = tmp + A[j][k];
                A[j][k] = A[j][k] - tau * tmp;
#pragma unroll
                for (int i = k + 1; i < 5; ++i) A[j][i] = A[j][i] - (tau * A[k][i]) * tmp;
            }
        }
#pragma unroll
        for (int j = k + 1; j < 3; ++j) {
            if (nrmU[j] != 0.f) {
                float temp = fabsf(A[j][k]) / nrmU[j];
                temp = (1.f + temp) * (1.f - temp);
                temp = temp < 0.f ? 0.f : temp;
                float ratio = nrmU[j] / nrmD[j];
                float temp2 = temp * (ratio * ratio);
                if (temp2 <= downdate_thr) {
                    nrmD[j] = norm_tail5(A[j], k + 1);
                    nrmU[j] = nrmD[j];
                } else {
                    nrmU[j] = nrmU[j] * sqrtf(temp);
                }
            }
        }
    }
    float x[3] = {0.f, 0.f, 0.f};
    if (nonzero_pivots > 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (k < nonzero_pivots) {
                float tau = hcoef[k];
                if (tau != 0.f) {
                    float tmp = A[k][k + 1] * cvec[k + 1];
#pragma unroll
                    for (int i = k + 2; i < 5; ++i) tmp = tmp + A[k][i] * cvec[i];
                    tmp = tmp + cvec[k];
                    cvec[k] = cvec[k] - tau * tmp;
#pragma unroll
                    for (int i = k + 1; i < 5; ++i) cvec[i] = cvec[i] - (tau * A[k][i]) * tmp;
                }
            }
        }
#pragma unroll
        for (int i = 2; i >= 0; --i) {
            if (i < nonzero_pivots) {
                cvec[i] = cvec[i] / A[i][i];
#pragma unroll
                for (int j = 0; j < i; ++j) cvec[j] = cvec[j] - cvec[i] * A[i][j];
            }
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (i < nonzero_pivots) {
                // x[perm[i]] = cvec[i] with compile-time destinations
                if (perm[i] == 0) x[0] = cvec[i]; else if (perm[i] == 1) x[1] = cvec[i]; else x[2] = cvec[i];
            }
        }
    }
    float n = sqrtf(x[0] * x[0] + (x[1] * x[1] + x[2] * x[2]));
    pabcd[0] = x[0] / n;
    pabcd[1] = x[1] / n;
    pabcd[2] = x[2] / n;
    pabcd[3] = (float)(1.0 / (double)n);  // "1.0 / n" in double, common_lib.h:691
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        float v = pabcd[0] * nb[j].x + pabcd[1] * nb[j].y + pabcd[2] * nb[j].z + pabcd[3];
        if (fabsf(v) > threshold) ok = false;
    }
    return ok;
}

__device__ __forceinline__ double dot3d(const double* a, double b0, double b1, double b2) { return a[0] * b0 + (a[1] * b1 + a[2] * b2); }

// ----------------------------------------------------------------------------------------------------
// fused kernel: a1 (re-transform), a3/a5 gates, a7 row, a8 Gram — everything that depends on the state.
//
// Work item = a tile of 32 consecutive scan points per warp, dealt statically (tile = global warp id + k * warps),
// so the summation order is fixed by (n, grid) alone and results are bit-reproducible.  Each lane builds the row
// of its point: H[p] = (h[0..11], z, eff ? 1 : 0, |pd2|, 0) — 16 columns; the Gram G = H^T H over the tile is
// accumulated with FP64 tensor-core MMAs (mma.sync m8n8k4 = SASS DMMA.8x8x4; the only dense contraction on the
// path) into three 8x8 register blocks per warp that live across ALL tiles of the warp:
//     G00 = rows 0..7 x cols 0..7,  G08 = rows 0..7 x cols 8..15,  G88 = rows 8..15 x cols 8..15
// so H^T H (78 upper terms), H^T z = G[:,12], n_eff = G[13][13] and res_sum = G[13][14] come out of the same
// accumulation, with no per-tile reduction, no barrier and no atomics.  One CTA-level fold at the end writes one
// 96-double record per CTA; the records are folded 16 at a time by the last CTA of each group to arrive, and the CTA
// that folds the LAST group record sums the group records (fixed order) and runs the trip's solve on the spot
// (solve_dev.cuh: reduce, exchange with the peers in sharded mode, 12x12 solve, state update, stop policy, and the
// state-only half of the next trip) — a trip has no solve kernel and no kernel boundary between Gram and solve.
// Staging: the tile's rows go through shared memory transposed (Ht[col][point], row pitch 36 doubles) so that the
// MMA fragment loads are bank-conflict-free at the 2-wavefront minimum of a 256-byte request.
// ----------------------------------------------------------------------------------------------------
constexpr int HT_PITCH = 36;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

constexpr size_t FUSED_SMEM_DOUBLES = (sizeof(SolveSmem) + 7) / 8 > (size_t)UNIT_WARPS * 16 * HT_PITCH ? (sizeof(SolveSmem) + 7) / 8 : (size_t)UNIT_WARPS * 16 * HT_PITCH;

#ifndef FLIVO_FUSED_CTAS_PER_SM
#define FLIVO_FUSED_CTAS_PER_SM 8   // 64 registers; the solve inside it spills under this cap (A/B: profiles/r2_ab_solve.txt)
#endif
__global__ void __launch_bounds__(UNIT_POINTS, FLIVO_FUSED_CTAS_PER_SM) k_fused(const __grid_constant__ FusedParams P, const __grid_constant__ SolveParams SP) {
    // per warp: Ht[16][36]; reused for the final fold and, in the one CTA that solves, as the solve's workspace
    __shared__ __align__(16) double s_raw[FUSED_SMEM_DOUBLES];
    double (*s_ht)[16 * HT_PITCH] = reinterpret_cast<double (*)[16 * HT_PITCH]>(s_raw);
    __shared__ double s_state[36];

    Ctl* ctl = P.ctl;
    // Programmatic dependent launch (P.pdl_fused): on a re-use trip this trip's k_search lets its dependents go as soon as it
    // has seen the previous solve's release of state + control block, so this grid may run beside the tail of the previous
    // trip's k_fused (state-only half of the next solve / covariance update).  The per-point part below needs only what
    // that release published — acquired here by one thread per CTA, read with L1-bypassing loads — and the scan's cached
    // planes; what the tail still writes (SolvePre, P_out) is read by the solving CTA alone, behind griddepcontrol.wait.
    // On a search trip k_search never releases its dependents early, and the wait right below makes its results visible.
    if (P.pdl_fused) {
        if (threadIdx.x == 0) {
            const unsigned useq = __ldcg(&P.job->useq) & 0xFFFFFFu;
            for (;;) {
                unsigned r, st;
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(r) : "l"(&ctl->ready) : "memory");
                if ((r >> 8) != useq) continue;
                if ((r & 255u) >= (unsigned)P.trip) break;
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(st) : "l"(&ctl->stopped) : "memory");
                if (st) break;
            }
        }
        __syncthreads();
    }
    // every word the prologue needs is fetched before the exit test: independent loads, one global-memory latency
    const int2 ss = __ldcg(reinterpret_cast<const int2*>(&ctl->search));   // {search, stopped}
    const int n_pts = P.job->n, mode = P.job->mode, aux_out = P.job->aux_out, fsearch = P.job->force_search;
    double st0 = 0.0;
    if (threadIdx.x < 36) st0 = __ldcg(P.state + threadIdx.x);
    if (ss.y) return;
    const bool search = (ss.x != 0) || (fsearch != 0);
    if (P.pdl_fused && search) asm volatile("griddepcontrol.wait;" ::: "memory");   // this trip's neighbours, planes and flags
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x < 36) s_state[threadIdx.x] = st0;
    __syncthreads();
    const double* Rm = s_state;          // rot
    const double* tw = s_state + 9;      // pos
    const double* Rl = s_state + 12;     // offR
    const double* tl = s_state + 21;     // offT
    double* ht = s_ht[warp];

    double g00a = 0.0, g00b = 0.0, g08a = 0.0, g08b = 0.0, g88a = 0.0, g88b = 0.0;
    const int n_tiles = (n_pts + 31) >> 5;
    const int total_warps = gridDim.x * UNIT_WARPS;
    for (int tile = blockIdx.x * UNIT_WARPS + warp; tile < n_tiles; tile += total_warps) {
        const int i = tile * 32 + lane;
        const bool have = i < n_pts;

        // ---- a1: world transform --------------------------------------------------------------------
        float4 pb = have ? __ldg(P.body + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        float4 pl4 = make_float4(0.f, 0.f, 0.f, 0.f);
        unsigned pf = 0, selprev = 0;
        if (have) { pf = P.pflags[i]; pl4 = __ldg(P.plane + i); selprev = P.sel[i]; }
        double pd[3] = {(double)pb.x, (double)pb.y, (double)pb.z};
        double pl[3];  // R_LI p + t_LI
        pl[0] = dot3d(Rl + 0, pd[0], pd[1], pd[2]) + tl[0];
        pl[1] = dot3d(Rl + 3, pd[0], pd[1], pd[2]) + tl[1];
        pl[2] = dot3d(Rl + 6, pd[0], pd[1], pd[2]) + tl[2];
        float wx = (float)(dot3d(Rm + 0, pl[0], pl[1], pl[2]) + tw[0]);
        float wy = (float)(dot3d(Rm + 3, pl[0], pl[1], pl[2]) + tw[1]);
        float wz = (float)(dot3d(Rm + 6, pl[0], pl[1], pl[2]) + tw[2]);
        if (have && aux_out) P.world[i] = make_float4(wx, wy, wz, pb.w);

        // ---- a2..a5: the neighbours enter only through esti_plane, fitted once per search by k_search / k_far
        // (P.plane, P.pflags); what depends on the state is re-evaluated here on every trip.
        bool sel;
        if (mode == 1) sel = true;                        // V1: vector<bool>(n, true) each call (:490)
        else sel = have && (selprev != 0);                  // V3: persistent flag (origin:963)
        const bool have5 = (pf & PF_HAVE5) != 0;
        if (search) sel = have5 && (pf & PF_GATE);          // pointSearchSqDis[4] > 5 ? false : true (:518)
        bool eff = false;
        float4 nv = make_float4(0.f, 0.f, 0.f, 0.f);
        float res = 0.f;
        if (sel && have5) {
            sel = false;
            if (pf & PF_PLANE) {                            // esti_plane(pabcd, points_near, 0.1f) (:530)
                float pd2 = pl4.x * wx + pl4.y * wy + pl4.z * wz + pl4.w;
                double nrm = sqrt(pd[0] * pd[0] + (pd[1] * pd[1] + pd[2] * pd[2]));
                float s = (float)(1.0 - 0.9 * (double)fabsf(pd2) / sqrt(nrm));
                if ((double)s > 0.9) {
                    sel = true;
                    nv = make_float4(pl4.x, pl4.y, pl4.z, pd2);
                    res = fabsf(pd2);
                    eff = res <= 2.0f;                       // res_last[i] <= 2.0 (:552)
                }
            }
        } else {
            sel = false;
        }
        if (have) {
            P.sel[i] = sel ? 1 : 0;
            if (aux_out) { P.eff[i] = eff ? 1 : 0; P.normvec[i] = nv; }
        }

        // ---- a7: Jacobian row (origin:1018-1044), transposed into shared memory -------------------------
        __syncwarp();   // the previous tile's fragment loads are done
        if (eff) {
            double n0 = (double)nv.x, n1 = (double)nv.y, n2 = (double)nv.z;
            double C0 = Rm[0] * n0 + (Rm[3] * n1 + Rm[6] * n2);   // R^T n
            double C1 = Rm[1] * n0 + (Rm[4] * n1 + Rm[7] * n2);
            double C2 = Rm[2] * n0 + (Rm[5] * n1 + Rm[8] * n2);
            double T0 = Rl[0] * C0 + (Rl[3] * C1 + Rl[6] * C2);   // R_LI^T C
            double T1 = Rl[1] * C0 + (Rl[4] * C1 + Rl[7] * C2);
            double T2 = Rl[2] * C0 + (Rl[5] * C1 + Rl[8] * C2);
            ht[0 * HT_PITCH + lane] = n0; ht[1 * HT_PITCH + lane] = n1; ht[2 * HT_PITCH + lane] = n2;
            ht[3 * HT_PITCH + lane] = pl[1] * C2 - pl[2] * C1;     // A = [R_LI p + t_LI]x C
            ht[4 * HT_PITCH + lane] = pl[2] * C0 - pl[0] * C2;
            ht[5 * HT_PITCH + lane] = pl[0] * C1 - pl[1] * C0;
            ht[6 * HT_PITCH + lane] = pd[1] * T2 - pd[2] * T1;     // B = [p]x R_LI^T C
            ht[7 * HT_PITCH + lane] = pd[2] * T0 - pd[0] * T2;
            ht[8 * HT_PITCH + lane] = pd[0] * T1 - pd[1] * T0;
            ht[9 * HT_PITCH + lane] = C0; ht[10 * HT_PITCH + lane] = C1; ht[11 * HT_PITCH + lane] = C2;
            ht[12 * HT_PITCH + lane] = -(double)nv.w;              // z = -pd2
            ht[13 * HT_PITCH + lane] = 1.0;                        // -> n_eff
            ht[14 * HT_PITCH + lane] = (double)res;                // -> res_sum
            ht[15 * HT_PITCH + lane] = 0.0;
        } else {
#pragma unroll
            for (int c = 0; c < 16; ++c) ht[c * HT_PITCH + lane] = 0.0;
        }
        __syncwarp();

        // ---- a8: G += H^T H over the tile's 32 points, 8 k-steps of 4 points ------------------------------
        if (__any_sync(FULL, eff)) {
            const int fr = lane >> 2, fk = lane & 3;
#pragma unroll
            for (int ks = 0; ks < 8; ++ks) {
                const double f0 = ht[fr * HT_PITCH + ks * 4 + fk];          // H[4 ks + fk][fr]
                const double f8 = ht[(8 + fr) * HT_PITCH + ks * 4 + fk];    // H[4 ks + fk][8 + fr]
                dmma884(g00a, g00b, f0, f0);
                dmma884(g08a, g08b, f0, f8);
                dmma884(g88a, g88b, f8, f8);
            }
        }
    }

    // ---- fold the CTA's warps (fixed order) into one 96-double record ----------------------------------------
    // accumulator fragment: lane holds G[lane/4][2*(lane%4) + {0,1}] of each 8x8 block
    __syncthreads();
    {
        double* w = s_ht[warp];   // [3][64]
        const int e = (lane >> 2) * 8 + 2 * (lane & 3);
        w[e] = g00a; w[e + 1] = g00b;
        w[64 + e] = g08a; w[64 + e + 1] = g08b;
        w[128 + e] = g88a; w[128 + e + 1] = g88b;
    }
    __syncthreads();
    if (threadIdx.x < NTERMS) {
        const int t = threadIdx.x;
        int r, c;   // position in the 16x16 Gram
        if (t < 78) { r = c_term_a[t]; c = c_term_b[t]; }
        else if (t < 90) { r = t - 78; c = 12; }
        else if (t == 90) { r = 13; c = 14; }                // res_sum
        else { r = 13; c = 13; }                             // n_eff
        const int blk = (r < 8) ? ((c < 8) ? 0 : 1) : 2;
        const int e = blk * 64 + (r & 7) * 8 + (c & 7);
        double sum = s_ht[0][e];
#pragma unroll
        for (int wv = 1; wv < UNIT_WARPS; ++wv) sum += s_ht[wv][e];
        P.partials[(size_t)blockIdx.x * PSTRIDE + t] = sum;
    }
    // ---- second level: the last CTA of each group of FOLD_CTAS to arrive adds the group's records in CTA order
    // (the result does not depend on which CTA that is), so the single-CTA solve reads grid/16 records, not grid.
    __shared__ unsigned s_last;
    __threadfence();
    __syncthreads();
    const unsigned grp = blockIdx.x / FOLD_CTAS;
    const unsigned members = min((unsigned)FOLD_CTAS, gridDim.x - grp * FOLD_CTAS);
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(P.group_cnt + grp, 1u);
        s_last = (ticket == members - 1) ? 1u : 0u;
        if (s_last) P.group_cnt[grp] = 0;   // ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < NTERMS) {
        const double* src = P.partials + (size_t)grp * FOLD_CTAS * PSTRIDE + threadIdx.x;
        double v[FOLD_CTAS];
#pragma unroll
        for (int m = 0; m < FOLD_CTAS; ++m) v[m] = (m < (int)members) ? __ldcg(src + (size_t)m * PSTRIDE) : 0.0;
        double sum = v[0];
#pragma unroll
        for (int m = 1; m < FOLD_CTAS; ++m) if (m < (int)members) sum += v[m];
        P.gpartials[(size_t)grp * PSTRIDE + threadIdx.x] = sum;
    }
    // ---- third level: the group folder that arrives last sums the group records in group order and solves ----------
    const unsigned ngroups = (gridDim.x + FOLD_CTAS - 1) / FOLD_CTAS;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(P.total_cnt, 1u);
        s_last = (ticket == ngroups - 1) ? 1u : 0u;
        if (s_last) *P.total_cnt = 0;       // ready for the next launch
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    SolveSmem& M = *reinterpret_cast<SolveSmem*>(s_raw);
    if (threadIdx.x < NTERMS) {
        // group order, sixteen independent loads in flight at a time (absent records add +0.0, which changes nothing)
        const double* src = P.gpartials + threadIdx.x;
        double sum = 0.0;
        for (unsigned g0 = 0; g0 < ngroups; g0 += 16) {
            double v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = (g0 + j < ngroups) ? __ldcg(src + (size_t)(g0 + j) * PSTRIDE) : 0.0;
#pragma unroll
            for (int j = 0; j < 16; ++j) sum += v[j];
        }
        M.sum[threadIdx.x] = sum;
    }
    if (P.pdl_fused) asm volatile("griddepcontrol.wait;" ::: "memory");   // the previous trip's k_fused is complete: SolvePre / P_out are final
    __syncthreads();
    solve_trip(SP, M, threadIdx.x, UNIT_POINTS);
}

// ----------------------------------------------------------------------------------------------------
// search kernels of a search trip (a1 + a2): refresh the Nearest_Points cache (P.near / P.near_meta)
//   k_search : thread per point over the 27-cell shell; points it cannot finish go to P.far_list
//   k_far    : warp per far point (growing cubes of cells), spread over the whole grid so sparse map
//              regions do not serialise inside one CTA
// Both exit at once on re-use trips (ctl->search == 0 and no force_search) and after the loop stopped.
// ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ void world_of(const double* __restrict__ st, float4 pb, float& wx, float& wy, float& wz) {
    double pd0 = (double)pb.x, pd1 = (double)pb.y, pd2 = (double)pb.z;
    double l0 = dot3d(st + 12, pd0, pd1, pd2) + st[21], l1 = dot3d(st + 15, pd0, pd1, pd2) + st[22], l2 = dot3d(st + 18, pd0, pd1, pd2) + st[23];
    wx = (float)(dot3d(st + 0, l0, l1, l2) + st[9]);
    wy = (float)(dot3d(st + 3, l0, l1, l2) + st[10]);
    wz = (float)(dot3d(st + 6, l0, l1, l2) + st[11]);
}

// a3 + a4 on the freshly found neighbours: flags and plane of point i (consumed by k_fused on every trip
// until the next search)
__device__ __forceinline__ void store_plane(const FusedParams& P, int i, const float4 nbs[5], int cnt, float d5) {
    unsigned pf = 0;
    float pabcd[4] = {0.f, 0.f, 0.f, 0.f};
    if (cnt >= 5) {
        pf |= PF_HAVE5;
        // the iVox branch has no far-neighbour gate: GetClosestPoint bounds the range itself (src/laser_mapping.cpp:516-521)
        if (P.map.ivox || !(d5 > 5.f)) pf |= PF_GATE;
        if (esti_plane_dev(nbs, 0.1f, pabcd)) pf |= PF_PLANE;
    }
    P.plane[i] = make_float4(pabcd[0], pabcd[1], pabcd[2], pabcd[3]);
    P.pflags[i] = (unsigned char)pf;
}

constexpr double IVOX_MAX_RANGE_SQ = 25.0;   // max_range = 5.0 default of GetClosestPoint (ivox3d.h:79), squared in double (ivox3d_node.hpp:162)
#ifndef FLIVO_SEARCH_THREADS
#define FLIVO_SEARCH_THREADS 128   // points per tile = threads per CTA (64 measured: profiles/r2_ab_search.txt)
#endif
constexpr int SEARCH_THREADS = FLIVO_SEARCH_THREADS;
constexpr int SEARCH_WARPS = SEARCH_THREADS / 32;
// register budget of k_search: 5 CTAs/SM = 96 registers (68 B of spills); 6 (80 registers, 140 B) and 4 (116, none) measured
// 1-2 % slower, 7 (72 registers, 112 B) 4 % slower (profiles/r1_ab_launch_structure.txt)
#ifndef FLIVO_SEARCH_CTAS
#define FLIVO_SEARCH_CTAS 5
#endif
constexpr int SEARCH_CTAS_PER_SM = FLIVO_SEARCH_CTAS;
constexpr int FAR_BATCH = 32;   // far points a warp finishes before its lanes fit their planes (one lane per point)

// shared memory of k_search: the two phases of a CTA never overlap, so their buffers share the space
union SearchSmem {
    struct {
        uint32_t idx[SHELL_CELLS * SEARCH_THREADS];   // per-thread work lists of thread_knn5_shell
        float gap[SHELL_CELLS * SEARCH_THREADS];
    } p1;
    struct {
        uint2 queue[SEARCH_WARPS][RING_QCAP];         // per-warp cell queue of warp_knn5_rings
        float4 near[SEARCH_WARPS][FAR_BATCH][5];      // neighbours of the warp's current batch of far points
    } p2;
};

__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One kernel per search trip (a1 + a2 + a3/a4 of the plane cache).  Two phases, work dealt dynamically:
//   phase 1  a CTA takes tiles of 128 consecutive (spatially sorted) scan points from ctl->tile_counter; thread per
//            point over the 27-cell shell; points the shell cannot finish are published to P.far_list;
//   phase 2  once the tiles are dealt out, every warp turns consumer: it takes tickets from ctl->far_taken and runs
//            the warp-per-query growing-cube search on the published far points, while other CTAs may still be in
//            phase 1.  A ticket whose entry is not there yet is awaited until every tile has been published
//            (ctl->tiles_done); only CTAs that hold a tile are ever waited for and those are running, so the wait
//            cannot deadlock whatever the grid size.  Entries carry the epoch of the trip, so the list never needs
//            clearing.  After FAR_BATCH finished queries the warp's lanes fit the planes of the batch, one lane each.
// Exits at once on re-use trips (ctl->search == 0 and no force_search) and after the loop stopped.
// IVOX: the iVox-compatible flavour (thread_knn5_ivox, no far phase) is a separate instantiation so that the exact-search
// kernel does not carry its code: k_search is instruction-fetch sensitive (ncu: no_instruction stalls).
// PAR: thread_knn5_shell<PAR_PROBES>, chosen by the launcher for scans that leave the GPU under-filled.
template <bool IVOX, bool PAR>
__global__ void __launch_bounds__(SEARCH_THREADS, IVOX ? 6 : SEARCH_CTAS_PER_SM) k_search(FusedParams P) {
    __shared__ double s_state[24];
    __shared__ SearchSmem S;
    __shared__ unsigned s_tile;
    __shared__ unsigned s_bid[SEARCH_WARPS][FAR_BATCH];
    __shared__ float2 s_bmeta[SEARCH_WARPS][FAR_BATCH];
    Ctl* ctl = P.ctl;
    // Programmatic dependent launch: this grid may have started while the previous trip's k_fused still runs its last CTA
    // (covariance update / next trip's state-only half).  What it needs from that trip — state and control block — is
    // published by solve_publish with a release store of ctl->ready; one thread per CTA acquires it, the barrier hands the
    // ordering to the others, and the loads below bypass L1 (__ldcg).  A loop that has stopped never advances `ready`.  The
    // word carries the update's sequence number, so the first trip (dependent of k_ctl_reset) cannot take the previous
    // update's final word — or its `stopped` — for its own.
    if (P.pdl) {
        if (threadIdx.x == 0) {
            const unsigned useq = __ldcg(&P.job->useq) & 0xFFFFFFu;
            for (;;) {
                unsigned r, st;
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(r) : "l"(&ctl->ready) : "memory");
                if ((r >> 8) != useq) continue;              // still the previous update's word: k_ctl_reset has not armed this one
                if ((r & 255u) >= (unsigned)P.trip) break;
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(st) : "l"(&ctl->stopped) : "memory");
                if (st) break;
            }
        }
        __syncthreads();
    }
    // exit test of re-use trips and stopped trips; every word of the prologue is fetched ahead of it by independent
    // loads (one global-memory latency instead of three)
    const int2 ss = __ldcg(reinterpret_cast<const int2*>(&ctl->search));   // {search, stopped}: the first two words of Ctl
    const int fs = P.job->force_search;
    const int n = P.job->n;
    const uint32_t max_sq_bits = P.job->max_sq_bits;
    const unsigned tag = __ldcg(&ctl->epoch) + 1u;
    const int searches_before = __ldcg(&ctl->searches);
    double st0 = 0.0;
    if (threadIdx.x < 24) st0 = __ldcg(P.state + threadIdx.x);
    if (ss.y || !(ss.x != 0 || fs != 0)) {
        // Nothing to search.  This trip's k_fused may start now (it is a programmatic dependent of this grid when P.pdl_fused
        // is set) — but this grid must not COMPLETE before the previous trip's k_fused has: the solving CTA of the next
        // k_fused takes the completion of this grid as the completion of that one.
        if (threadIdx.x == 0) {
            asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
            asm volatile("griddepcontrol.wait;" ::: "memory");
        }
        return;
    }
    unsigned long long t_start = 0, t_p1 = 0;
    const bool stamp_on = P.search_stamps != nullptr && searches_before == 0;   // first search trip of the update
    if (stamp_on && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
    if (threadIdx.x < 24) s_state[threadIdx.x] = st0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned n_tiles = (unsigned)((n + SEARCH_THREADS - 1) / SEARCH_THREADS);
    unsigned nc = 0, nfar = 0;

    // ---- phase 1 ------------------------------------------------------------------------------------
    bool had_tile = false;
    for (;;) {
        __syncthreads();   // the tile's far points are published (and s_tile has been read by everyone)
        if (threadIdx.x == 0) {
            if (had_tile) atomicAdd(&ctl->tiles_done, 1u);
            s_tile = atomicAdd(&ctl->tile_counter, 1u);
        }
        __syncthreads();
        const unsigned tile = s_tile;
        if (tile >= n_tiles) break;
        had_tile = true;
        const int i = (int)(tile * SEARCH_THREADS) + threadIdx.x;
        const bool have = i < n;
        bool fin = true;
        if (have) {
            float wx, wy, wz;
            world_of(s_state, __ldg(P.body + i), wx, wy, wz);
            Top5 best;
            int cnt = 0;
            bool keep_old = false;
            if (IVOX) {
                thread_knn5_ivox(P.map, wx, wy, wz, IVOX_MAX_RANGE_SQ, best, cnt, nc);
                // no candidate at all: GetClosestPoint returns before touching closest_pt (ivox3d.h:168-170), so the
                // point keeps the neighbours of the previous search of this scan
                keep_old = cnt == 0;
            } else {
                fin = thread_knn5_shell<PAR>(P.map, wx, wy, wz, max_sq_bits, best, cnt, nc, S.p1.idx + threadIdx.x, S.p1.gap + threadIdx.x, SEARCH_THREADS);
            }
            if (keep_old) {
            } else if (fin) {
                float4 nbs[5];
#pragma unroll
                for (int r = 0; r < 5; ++r) {
                    nbs[r] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
                    if (r < cnt) nbs[r] = __ldg(P.map.buckets + best.a[r]);
                    P.near[(size_t)i * 5 + r] = nbs[r];
                }
                const float d5 = (cnt >= 5) ? __uint_as_float((uint32_t)(best.k[4] >> 32)) : INFINITY;
                P.near_meta[i] = make_float2(__int_as_float(cnt), d5);
                store_plane(P, i, nbs, cnt, d5);
            } else {
                // hand the cube-1 result to the warp that will finish this point: (distance bits, id, address) per rank,
                // parked in the point's own slots of P.near until the final neighbours overwrite them
                uint4* seed = reinterpret_cast<uint4*>(P.near + (size_t)i * 5);
#pragma unroll
                for (int r = 0; r < 5; ++r)
                    seed[r] = (r < cnt) ? make_uint4((uint32_t)(best.k[r] >> 32), (uint32_t)best.k[r], best.a[r], 0u) : make_uint4(INF_BITS, 0xFFFFFFFFu, 0u, 0u);
                P.near_meta[i] = make_float2(__int_as_float(cnt), INFINITY);
                __threadfence();   // the seed is visible before the queue entry that announces the point
            }
        }
        // publish the far points: one atomic per warp
        const unsigned pend = __ballot_sync(FULL, have && !fin);
        if (pend) {
            unsigned basep = 0;
            if (lane == 0) basep = atomicAdd(&ctl->far_count, (unsigned)__popc(pend));
            basep = __shfl_sync(FULL, basep, 0);
            if (have && !fin) {
                P.far_list[basep + __popc(pend & ((1u << lane) - 1u))] = ((unsigned long long)tag << 32) | (unsigned)i;
                __threadfence();   // before this CTA counts the tile as done
            }
            nfar += (unsigned)__popc(pend);
        }
    }

    if (stamp_on && threadIdx.x == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_p1));
    unsigned nfar_done = 0;
    if (!IVOX) {   // (the iVox search is always final: nothing is ever published to far_list)
    // ---- phase 2 (the barrier that ended phase 1 makes the shared buffers free) ---------------------------
    uint2* queue = S.p2.queue[warp];
    int nb = 0;
    for (;;) {
        unsigned long long v = 0ull;
        if (lane == 0) {
            const unsigned e = atomicAdd(&ctl->far_taken, 1u);
            if (e < (unsigned)n) {
                for (;;) {
                    v = ld_volatile_u64(P.far_list + e);
                    if ((unsigned)(v >> 32) == tag) break;
                    if (ld_volatile_u32(&ctl->tiles_done) >= n_tiles) {   // nothing more will be published
                        __threadfence();
                        v = ld_volatile_u64(P.far_list + e);
                        break;
                    }
                    __nanosleep(256);
                }
            }
        }
        v = __shfl_sync(FULL, v, 0);
        if ((unsigned)(v >> 32) != tag) break;
        const unsigned i = (unsigned)v;
        __threadfence();   // the producer's seed stores precede the entry this warp just saw
        float wx, wy, wz;
        world_of(s_state, __ldg(P.body + i), wx, wy, wz);
        uint4 sd = make_uint4(INF_BITS, 0xFFFFFFFFu, 0u, 0u);
        if (lane < 5) sd = __ldcg(reinterpret_cast<const uint4*>(P.near + (size_t)i * 5) + lane);
        const int seed_cnt = __float_as_int(__ldcg(&P.near_meta[i].x));
        Cand out;
        int dbg = 0;
        unsigned long long tq0 = 0;
        if (stamp_on) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tq0));
        const int cnt = warp_knn5_rings(P.map, wx, wy, wz, max_sq_bits, lane, out, nc, queue, seed_cnt, sd.x, sd.y, sd.z, stamp_on ? &dbg : nullptr);
        if (stamp_on && lane == 0) {
            unsigned long long tq1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tq1));
            unsigned long long* G = P.search_stamps + (size_t)4096 * 4;
            atomicAdd(G + 0, tq1 - tq0); atomicAdd(G + 1, 1ull); atomicAdd(G + 2, (unsigned long long)(dbg >> 8));
            atomicAdd(G + 3 + min(dbg & 255, 12), 1ull);
            atomicAdd(G + 16 + min(max(seed_cnt, 0), 5), 1ull);
            atomicMax(G + 22, tq1 - tq0);
            atomicAdd(G + 23 + min(dbg & 255, 8), tq1 - tq0);   // duration by final radius
        }
        if (lane < 5) {
            float4 p = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            if (out.d != INF_BITS) p = __ldg(P.map.buckets + out.addr);
            P.near[(size_t)i * 5 + lane] = p;
            S.p2.near[warp][nb][lane] = p;
        }
        const uint32_t dlast = __shfl_sync(FULL, out.d, 4);
        if (lane == 0) {
            const float2 meta = make_float2(__int_as_float(cnt), (cnt >= 5) ? __uint_as_float(dlast) : INFINITY);
            P.near_meta[i] = meta;
            s_bid[warp][nb] = i; s_bmeta[warp][nb] = meta;
        }
        ++nb; ++nfar_done;
        if (nb == FAR_BATCH) {
            __syncwarp();
            float4 nbs[5];
#pragma unroll
            for (int r = 0; r < 5; ++r) nbs[r] = S.p2.near[warp][lane][r];
            store_plane(P, (int)s_bid[warp][lane], nbs, __float_as_int(s_bmeta[warp][lane].x), s_bmeta[warp][lane].y);
            nb = 0;
            __syncwarp();
        }
    }
    __syncwarp();
    if (lane < nb) {
        float4 nbs[5];
#pragma unroll
        for (int r = 0; r < 5; ++r) nbs[r] = S.p2.near[warp][lane][r];
        store_plane(P, (int)s_bid[warp][lane], nbs, __float_as_int(s_bmeta[warp][lane].x), s_bmeta[warp][lane].y);
    }
    }   // !IVOX
    if (stamp_on) {
        __syncthreads();
        if (lane == 0 && nfar_done) atomicAdd(&P.search_stamps[blockIdx.x * 4 + 3], (unsigned long long)nfar_done);
        if (threadIdx.x == 0) {
            unsigned long long t_end;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
            P.search_stamps[blockIdx.x * 4 + 0] = t_start; P.search_stamps[blockIdx.x * 4 + 1] = t_p1; P.search_stamps[blockIdx.x * 4 + 2] = t_end;
        }
    }
    if (P.stats) {
        const unsigned tot = __reduce_add_sync(FULL, nc);
        if (lane == 0) {
            if (nfar) atomicAdd(&P.stats->far_queries, (unsigned long long)nfar);
            if (tot) atomicAdd(&P.stats->candidates, (unsigned long long)tot);
        }
    }
    if (threadIdx.x == 0) asm volatile("griddepcontrol.wait;" ::: "memory");   // see the early exit above
}

// ----------------------------------------------------------------------------------------------------
// batched Nearest_Search for the C ABI (flivo_knn): one warp per query
// ----------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_knn(MapView map, const float4* __restrict__ q, int nq, int k, uint32_t max_sq_bits,
                                             float* out_xyz, float* out_sqd, int* out_id, int* out_count, SearchStats* stats) {
    __shared__ uint2 s_queue[4][RING_QCAP];
    const int lane = threadIdx.x & 31;
    const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int i = wid; i < nq; i += nwarps) {
        float4 p = __ldg(q + i);
        Cand out;
        unsigned nc = 0;
        int cnt = warp_knn5_rings(map, p.x, p.y, p.z, max_sq_bits, lane, out, nc, s_queue[threadIdx.x >> 5]);
        const unsigned tot = __reduce_add_sync(FULL, nc);
        if (stats && lane == 0) atomicAdd(&stats->candidates, (unsigned long long)tot);
        int found = min(cnt, k);
        if (lane < k) {
            bool ok = lane < found;
            float4 np = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            if (ok) np = __ldg(map.buckets + out.addr);
            if (out_xyz) { out_xyz[((size_t)i * k + lane) * 3 + 0] = np.x; out_xyz[((size_t)i * k + lane) * 3 + 1] = np.y; out_xyz[((size_t)i * k + lane) * 3 + 2] = np.z; }
            if (out_sqd) out_sqd[(size_t)i * k + lane] = ok ? __uint_as_float(out.d) : INFINITY;
            if (out_id) out_id[(size_t)i * k + lane] = ok ? (int)out.id : -1;
        }
        if (lane == 0 && out_count) out_count[i] = found;
    }
}

// iVox mode: thread per query, IVox::GetClosestPoint semantics; max_range_sq = max_range * max_range (strict <)
__global__ void __launch_bounds__(128) k_knn_ivox(MapView map, const float4* __restrict__ q, int nq, int k, double max_range_sq,
                                                  float* out_xyz, float* out_sqd, int* out_id, int* out_count, SearchStats* stats) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned nc = 0;
    if (i < nq) {
        const float4 p = __ldg(q + i);
        Top5 best;
        int cnt = 0;
        thread_knn5_ivox(map, p.x, p.y, p.z, max_range_sq, best, cnt, nc);
        const int found = min(cnt, k);
        for (int r = 0; r < k; ++r) {
            const bool ok = r < found;
            float4 np = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
            if (ok) np = __ldg(map.buckets + best.a[r]);
            const size_t o = (size_t)i * k + r;
            if (out_xyz) { out_xyz[o * 3 + 0] = np.x; out_xyz[o * 3 + 1] = np.y; out_xyz[o * 3 + 2] = np.z; }
            if (out_sqd) out_sqd[o] = ok ? __uint_as_float((uint32_t)(best.k[r] >> 32)) : INFINITY;
            if (out_id) out_id[o] = ok ? (int)(uint32_t)best.k[r] : -1;
        }
        if (out_count) out_count[i] = found;
    }
    const unsigned tot = __reduce_add_sync(FULL, nc);
    if (stats && (threadIdx.x & 31) == 0 && tot) atomicAdd(&stats->candidates, (unsigned long long)tot);
}

// world transform only (map_incremental, src/laser_mapping.cpp:343-346): body -> world float4 (w = 0)
__global__ void k_transform(const float4* __restrict__ body, int n, const double* __restrict__ state, float4* __restrict__ world) {
    __shared__ double s[36];
    if (threadIdx.x < 36) s[threadIdx.x] = state[threadIdx.x];
    __syncthreads();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 pb = __ldg(body + i);
    double pd0 = (double)pb.x, pd1 = (double)pb.y, pd2 = (double)pb.z;
    double l0 = dot3d(s + 12, pd0, pd1, pd2) + s[21], l1 = dot3d(s + 15, pd0, pd1, pd2) + s[22], l2 = dot3d(s + 18, pd0, pd1, pd2) + s[23];
    world[i] = make_float4((float)(dot3d(s + 0, l0, l1, l2) + s[9]), (float)(dot3d(s + 3, l0, l1, l2) + s[10]),
                           (float)(dot3d(s + 6, l0, l1, l2) + s[11]), 0.f);
}

// map_incremental of the iVox branch (src/laser_mapping.cpp:329-389): world transform with the final state, then per
// point the decision "add / add without down-sampling / drop" from its cached Nearest_Points and the centre of its
// filter_size_map_min voxel.  Thread per point in processing order; results land at the caller's index (perm).
//   flag 1 = points_to_add, 2 = point_no_need_downsample, 0 = dropped
// Eigen float vector arithmetic restated: (pw / fs).floor() + 0.5) * fs in float; calc_dist(Vector3f, Vector3f) =
// (p1 - p2).norm() (include/common_lib.h:85).
__global__ void k_ivox_map_filter(const float4* __restrict__ body, int n, const double* __restrict__ state,
                                  const float4* __restrict__ near, const float2* __restrict__ near_meta,
                                  const unsigned int* __restrict__ perm, double fs, float4* __restrict__ world_out,
                                  unsigned char* __restrict__ flag_add, unsigned char* __restrict__ flag_noneed) {
    __shared__ double s[24];
    if (threadIdx.x < 24) s[threadIdx.x] = state[threadIdx.x];
    __syncthreads();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    float wx, wy, wz;
    world_of(s, __ldg(body + j), wx, wy, wz);
    const unsigned o = perm ? perm[j] : (unsigned)j;
    const int cnt = __float_as_int(near_meta[j].x);
    int flag = 1;
    if (cnt > 0) {                                           // !Nearest_Points[i].empty() (:349)
        const float fsf = (float)fs;
        const float cx = fmul(fadd(floorf(fdiv(wx, fsf)), 0.5f), fsf), cy = fmul(fadd(floorf(fdiv(wy, fsf)), 0.5f), fsf),
                    cz = fmul(fadd(floorf(fdiv(wz, fsf)), 0.5f), fsf);
        const float4 n0 = near[(size_t)j * 5];
        const double half = 0.5 * fs;
        if ((double)fabsf(fsub(n0.x, cx)) > half && (double)fabsf(fsub(n0.y, cy)) > half && (double)fabsf(fsub(n0.z, cz)) > half) {
            flag = 2;                                        // nearest neighbour far from the voxel centre (:358-363)
        } else {
            const float dist = fsqrt(sqdist3_eigen(wx, wy, wz, cx, cy, cz));
            if (cnt >= 5) {
                for (int r = 0; r < 5; ++r) {
                    const float4 q = near[(size_t)j * 5 + r];
                    if ((double)fsqrt(sqdist3_eigen(q.x, q.y, q.z, cx, cy, cz)) < (double)dist + 1e-6) { flag = 0; break; }
                }
            }
        }
    }
    world_out[o] = make_float4(wx, wy, wz, 0.f);
    flag_add[o] = flag == 1;
    flag_noneed[o] = flag == 2;
}

// ---- host launchers ---------------------------------------------------------------------------------
void launch_fused(const FusedParams& P, const SolveParams& S, int grid, cudaStream_t stream) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(UNIT_POINTS); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = P.pdl_fused ? 1 : 0;
    cudaLaunchKernelEx(&cfg, k_fused, P, S);
}
void launch_search(const FusedParams& P, int search_grid, cudaStream_t stream, bool par_probes, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)search_grid); cfg.blockDim = dim3(SEARCH_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (P.map.ivox) cudaLaunchKernelEx(&cfg, k_search<true, false>, P);
    else if (par_probes) cudaLaunchKernelEx(&cfg, k_search<false, true>, P);
    else cudaLaunchKernelEx(&cfg, k_search<false, false>, P);
}
// CTAs of k_search that fit on one SM (its grid never needs to exceed the resident set: tiles are dealt dynamically)
int search_tile_points() { return SEARCH_THREADS; }
int search_ctas_per_sm(bool ivox) {
    int nb = 0;
    const cudaError_t e = ivox ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_search<true, false>, SEARCH_THREADS, 0)
                               : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_search<false, false>, SEARCH_THREADS, 0);
    if (e != cudaSuccess || nb < 1) nb = 4;
    return nb;
}
void launch_knn(const MapView& map, const float4* q, int nq, int k, uint32_t max_sq_bits, float* out_xyz, float* out_sqd,
                int* out_id, int* out_count, SearchStats* stats, int grid, cudaStream_t stream) {
    k_knn<<<grid, 128, 0, stream>>>(map, q, nq, k, max_sq_bits, out_xyz, out_sqd, out_id, out_count, stats);
}
void launch_knn_ivox(const MapView& map, const float4* q, int nq, int k, double max_range_sq, float* out_xyz, float* out_sqd,
                     int* out_id, int* out_count, SearchStats* stats, cudaStream_t stream) {
    k_knn_ivox<<<(nq + 127) / 128, 128, 0, stream>>>(map, q, nq, k, max_range_sq, out_xyz, out_sqd, out_id, out_count, stats);
}
void launch_ivox_map_filter(const float4* body, int n, const double* state, const float4* near, const float2* near_meta,
                            const unsigned int* perm, double fs, float4* world_out, unsigned char* flag_add,
                            unsigned char* flag_noneed, cudaStream_t stream) {
    k_ivox_map_filter<<<(n + 255) / 256, 256, 0, stream>>>(body, n, state, near, near_meta, perm, fs, world_out, flag_add, flag_noneed);
}
void launch_transform(const float4* body, int n, const double* state, float4* world, cudaStream_t stream) {
    k_transform<<<(n + 255) / 256, 256, 0, stream>>>(body, n, state, world);
}

}  // namespace flivo
