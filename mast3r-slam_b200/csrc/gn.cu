// Sim3 Gauss-Newton on the device: accumulate -> per-edge finalise -> assemble -> blocked fp64
// Cholesky -> back-substitution + convergence test + retraction.  No host synchronisation and no
// atomics anywhere; every reduction has a fixed order (bitwise reproducible, also across ranks).
//
// Reference: backends/mpsgraph/gauss_newton.py:23-280 (rays), gauss_newton_points.py:17-207,
// gauss_newton_calib.py:17-274, sim3_ops.py (relative pose, Huber, exp/retract).
//
// Algebra used: in all three modes Ji = -Jj, and
//     Jj = J' * G,   J' = P * [ s_i^-1 I | M | X' ],   G = blockdiag(R_i, R_i, 1),
// with X' = T_ij Xj, M = -[X']x as the reference writes it (gauss_newton.py:172-182), R_i = R(q_i),
// and P = I (rays / points) or the 3x3 projection Jacobian (calib).  G is constant per edge, so the
// per-point work accumulates S = sum J'^T W J' and v = sum J'^T W err in the pose-free basis (29
// sums for rays/points thanks to the sparsity of J', 35 for calib); the edge block is
// A = G^T S G, b = G^T v, and  H_ii += A, H_jj += A, H_ij -= A, H_ji -= A, g_i -= b, g_j += b.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <new>
#include <set>
#include <utility>
#include <vector>

#include "m3s_common.cuh"

namespace m3s {

constexpr int kAccThreads = 256;
constexpr int kNB = 32;            // Cholesky panel width
constexpr int kEdgeStride = M3S_GN_EDGE_STRIDE;
constexpr double kHuber = 1.345;   // sim3_ops.py:295
constexpr double kReg = 1e-6;      // gauss_newton.py:254

struct Status {
    int applied;    // iterations whose update was applied
    int stopped;    // ||dx|| < delta_thresh hit
    int failed;     // factorisation failed / non-finite
    int attempted;  // iterations attempted
    int chol_fail;  // scratch flag of the running factorisation
    int xch_timeout;  // peer exchange: a rank did not publish within the spin bound (=> failed)
    int pad[2];
    double last_norm;
};

struct EdgeMeta {
    int ix, jx;   // keyframe rows of Xs / poses
    int il, jl;   // local free index or -1
    int active;   // contributes to H, g
    int pad[3];
};

struct KernelParams {
    int mode;
    float sig_inv_f;              // f32(1/sigma_ray | 1/sigma_point)
    float c_th, q_th;             // f32 thresholds (numpy compares f32 arrays with weak scalars)
    float fx, fy, cx, cy;         // calib
    double z_eps, sig_px_inv, sig_z_inv;
    double u_lo, u_hi, v_lo, v_hi;  // border <= u < W - border
};

// ------------------------------------------------------------------------------------------------
// small fp64 Sim3 helpers
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline void quat_mul(const double* a, const double* b, double* o) {
    const double x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3];
    const double x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
    o[0] = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2;
    o[1] = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2;
    o[2] = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2;
    o[3] = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2;
}

// linear map of sim3_ops.quat_rotate: R = I + 2w[q]x + 2[q]x^2 (valid for non-unit q as well)
__host__ __device__ inline void quat_to_mat(const double* q, double* R) {
    const double x = q[0], y = q[1], z = q[2], w = q[3];
    R[0] = 1.0 - 2.0 * (y * y + z * z); R[1] = 2.0 * (x * y - z * w); R[2] = 2.0 * (x * z + y * w);
    R[3] = 2.0 * (x * y + z * w); R[4] = 1.0 - 2.0 * (x * x + z * z); R[5] = 2.0 * (y * z - x * w);
    R[6] = 2.0 * (x * z - y * w); R[7] = 2.0 * (y * z + x * w); R[8] = 1.0 - 2.0 * (x * x + y * y);
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------------------
// K3: accumulate.  grid = nchunks * E blocks; block (e, chunk) streams points
// [chunk*chunk_pts, ...) of directed edge e.
//   per point reads: idx 4 + valid 1 + Q 4 (edge streams), Xj 12 + Cj 4 (coalesced), Xi 12 + Ci 4
//   (gather through idx) = 41 B; compulsory traffic 9 B/(edge,pt) + 16 B/(kf,pt).
// Per-thread fp64 accumulators -> warp shuffle tree -> shared across warps -> one partial per
// block.  No atomics.
// ------------------------------------------------------------------------------------------------
template <int MODE> struct AccCount { static constexpr int value = 29; };
template <> struct AccCount<M3S_GN_CALIB> { static constexpr int value = 35; };

struct EdgeConst {
    double sR[9];   // s_ij * R(q_ij)
    double t[3];    // t_ij
    double s_inv;   // 1 / s_i
};

template <int MODE>
__device__ __forceinline__ void accumulate_point(const KernelParams& P, const EdgeConst& ec,
                                                 float xjx, float xjy, float xjz, float xix,
                                                 float xiy, float xiz, float qc, double* acc) {
    const double Xj0 = xjx, Xj1 = xjy, Xj2 = xjz;
    const double x = ec.sR[0] * Xj0 + ec.sR[1] * Xj1 + ec.sR[2] * Xj2 + ec.t[0];
    const double y = ec.sR[3] * Xj0 + ec.sR[4] * Xj1 + ec.sR[5] * Xj2 + ec.t[1];
    const double z = ec.sR[6] * Xj0 + ec.sR[7] * Xj1 + ec.sR[8] * Xj2 + ec.t[2];
    if (MODE == M3S_GN_CALIB) {
        // gauss_newton_calib.py:125-221.  Observed side (Xi, fp32 K) is evaluated in fp32 exactly as
        // numpy does (f32 arrays times np.float32 scalars), predicted side in fp64.
        if (!(z > P.z_eps) || !((double)xiz > P.z_eps)) return;
        const double zinv = 1.0 / z;
        const float zinv_i = __fdiv_rn(1.0f, xiz);
        const double pju = (double)P.fx * x * zinv + (double)P.cx;
        const double pjv = (double)P.fy * y * zinv + (double)P.cy;
        if (!(pju >= P.u_lo && pju < P.u_hi && pjv >= P.v_lo && pjv < P.v_hi)) return;
        const float piu = __fadd_rn(__fmul_rn(__fmul_rn(P.fx, xix), zinv_i), P.cx);
        const float piv = __fadd_rn(__fmul_rn(__fmul_rn(P.fy, xiy), zinv_i), P.cy);
        const double e0 = (pju - (double)piu) * P.sig_px_inv;
        const double e1 = (pjv - (double)piv) * P.sig_px_inv;
        const double e2 = (log(z) - (double)logf(xiz)) * P.sig_z_inv;
        const float sw = __fsqrt_rn(qc);
        const double sqrt_w = sw, w2 = (double)__fmul_rn(sw, sw);
        const double zinv2 = zinv * zinv;
        const double pa = (double)P.fx * zinv * P.sig_px_inv;
        const double pb = -(double)P.fx * x * zinv2 * P.sig_px_inv;
        const double pc = (double)P.fy * zinv * P.sig_px_inv;
        const double pd = -(double)P.fy * y * zinv2 * P.sig_px_inv;
        const double pe = zinv * P.sig_z_inv;
        const double si = ec.s_inv;
        // rows of J' = P [ si I | M | X' ],  M = [[0,z,-y],[-z,0,x],[y,-x,0]]
        const double J[3][7] = {
            {pa * si, 0.0, pb * si, pb * y, pa * z - pb * x, -pa * y, pa * x + pb * z},
            {0.0, pc * si, pd * si, -pc * z + pd * y, -pd * x, pc * x, pc * y + pd * z},
            {0.0, 0.0, pe * si, pe * y, -pe * x, 0.0, pe * z}};
        const double err[3] = {e0, e1, e2};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const double ra = fabs(sqrt_w * err[c]);
            const double w = (ra < kHuber) ? w2 : (kHuber * __drcp_rn(ra)) * w2;
            int k = 0;
#pragma unroll
            for (int a = 0; a < 7; ++a) {
                const double wj = w * J[c][a];
#pragma unroll
                for (int b = a; b < 7; ++b) acc[k++] += wj * J[c][b];
                acc[28 + a] += wj * err[c];
            }
        }
    } else {
        const double e0 = x - (double)xix, e1 = y - (double)xiy, e2 = z - (double)xiz;
        float sw = __fmul_rn(P.sig_inv_f, __fsqrt_rn(qc));  // gauss_newton.py:147 (f32)
        if (MODE == M3S_GN_POINTS) {
            // gauss_newton_points.py:115-119: f32 norm + 1e-6, f32 reciprocal
            float ss = __fmul_rn(xix, xix);
            ss = __fadd_rn(ss, __fmul_rn(xiy, xiy));
            ss = __fadd_rn(ss, __fmul_rn(xiz, xiz));
            const float dist = __fadd_rn(__fsqrt_rn(ss), 1e-6f);
            sw = __fmul_rn(sw, __fdiv_rn(1.0f, dist));
        }
        const double sqrt_w = sw, w2 = (double)__fmul_rn(sw, sw);
        const double r0 = fabs(sqrt_w * e0), r1 = fabs(sqrt_w * e1), r2 = fabs(sqrt_w * e2);
        // Huber (sim3_ops.py:295-306): k / |r| via reciprocal (within 1 ulp of the division)
        const double w0 = (r0 < kHuber) ? w2 : (kHuber * __drcp_rn(r0)) * w2;
        const double w1 = (r1 < kHuber) ? w2 : (kHuber * __drcp_rn(r1)) * w2;
        const double w2c = (r2 < kHuber) ? w2 : (kHuber * __drcp_rn(r2)) * w2;
        const double w0y = w0 * y, w0z = w0 * z, w1x = w1 * x, w1z = w1 * z, w2x = w2c * x,
                     w2y = w2c * y;
        const double w0x = w0 * x, w1y = w1 * y, w2z = w2c * z;
        acc[0] += w0; acc[1] += w1; acc[2] += w2c;
        acc[3] += w0z; acc[4] += w0y; acc[5] += w1z; acc[6] += w1x; acc[7] += w2y; acc[8] += w2x;
        acc[9] += w0x; acc[10] += w1y; acc[11] += w2z;
        acc[12] += w1z * z + w2y * y;
        acc[13] += w0z * z + w2x * x;
        acc[14] += w0y * y + w1x * x;
        acc[15] += w2x * y;
        acc[16] += w1x * z;
        acc[17] += w0y * z;
        acc[18] += (w2y - w1y) * z;   // yz (w2 - w1)
        acc[19] += (w0x - w2x) * z;   // xz (w0 - w2)
        acc[20] += (w1x - w0x) * y;   // xy (w1 - w0)
        acc[21] += w0x * x + w1y * y + w2z * z;
        const double u0 = w0 * e0, u1 = w1 * e1, u2 = w2c * e2;
        acc[22] += u0; acc[23] += u1; acc[24] += u2;
        acc[25] += y * u2 - z * u1;
        acc[26] += z * u0 - x * u2;
        acc[27] += x * u1 - y * u0;
        acc[28] += x * u0 + y * u1 + z * u2;
    }
}

template <int MODE>
__global__ void __launch_bounds__(kAccThreads, 2)
gn_accum_kernel(const Status* __restrict__ status, KernelParams P, const double* __restrict__ poses,
                const EdgeMeta* __restrict__ meta, const float* __restrict__ Xs,
                const float* __restrict__ Cs, const int32_t* __restrict__ idx,
                const uint8_t* __restrict__ valid, const float* __restrict__ Q, long long N,
                int edge_begin, int nchunks, long long chunk_pts, int vec_ok,
                double* __restrict__ partials) {
    if (status->stopped | status->failed) return;
    constexpr int NA = AccCount<MODE>::value;
    const int e = edge_begin + blockIdx.x / nchunks;
    const int chunk = blockIdx.x % nchunks;
    const EdgeMeta em = meta[e];
    double* out = partials + ((size_t)e * nchunks + chunk) * kEdgeStride;
    if (!em.active) return;  // slot is zeroed by the finalise kernel

    __shared__ EdgeConst ec_s;
    __shared__ double red[kAccThreads / 32][NA];
    if (threadIdx.x == 0) {
        // sim3_relative (sim3_ops.py:96-126)
        const double* Ti = poses + 8 * em.ix;
        const double* Tj = poses + 8 * em.jx;
        const double si_inv = 1.0 / Ti[7];
        const double sij = si_inv * Tj[7];
        const double qi_inv[4] = {-Ti[3], -Ti[4], -Ti[5], Ti[6]};
        double qij[4], R[9], Ri[9];
        quat_mul(qi_inv, Tj + 3, qij);
        quat_to_mat(qij, R);
        quat_to_mat(qi_inv, Ri);
        const double d0 = Tj[0] - Ti[0], d1 = Tj[1] - Ti[1], d2 = Tj[2] - Ti[2];
#pragma unroll
        for (int k = 0; k < 9; ++k) ec_s.sR[k] = R[k] * sij;
        ec_s.t[0] = (Ri[0] * d0 + Ri[1] * d1 + Ri[2] * d2) * si_inv;
        ec_s.t[1] = (Ri[3] * d0 + Ri[4] * d1 + Ri[5] * d2) * si_inv;
        ec_s.t[2] = (Ri[6] * d0 + Ri[7] * d1 + Ri[8] * d2) * si_inv;
        ec_s.s_inv = si_inv;
    }
    __syncthreads();
#ifdef M3S_GN_EC_REGS
    const EdgeConst ec = ec_s;
#else
    // the 13 per-edge constants stay in shared memory and are re-read every trip (LDS.64, shared by
    // the 4 points of the trip): 26 fewer live registers next to the 58 accumulator registers
    const EdgeConst& ec = ec_s;
#endif

    double acc[NA];
#pragma unroll
    for (int k = 0; k < NA; ++k) acc[k] = 0.0;

    const float* __restrict__ Xi = Xs + (size_t)em.ix * N * 3;
    const float* __restrict__ Ci = Cs + (size_t)em.ix * N;
    const float* __restrict__ Xj = Xs + (size_t)em.jx * N * 3;
    const float* __restrict__ Cj = Cs + (size_t)em.jx * N;
    const int32_t* __restrict__ ide = idx + (size_t)e * N;
    const uint8_t* __restrict__ ve = valid + (size_t)e * N;
    const float* __restrict__ Qe = Q + (size_t)e * N;

    const long long n0 = (long long)chunk * chunk_pts;
    const long long n1 = min(N, n0 + chunk_pts);
    if (vec_ok) {
        // Vector path (N % 4 == 0, chunk_pts % 4 == 0): each thread takes 4 consecutive points per
        // trip.  All streaming loads (valid uchar4, Q / Cj float4, idx int4, Xj 3 x float4) are
        // issued together, then the 4 gathers of (Ci, Xi) with clamped indices, and only then the
        // arithmetic: 4x the memory-level parallelism of a point-at-a-time loop and no load sits
        // behind a data-dependent branch (first version: 66 % of stall samples were long-scoreboard).
        const uchar4* __restrict__ v4p = reinterpret_cast<const uchar4*>(ve);
        const float4* __restrict__ q4p = reinterpret_cast<const float4*>(Qe);
        const int4* __restrict__ i4p = reinterpret_cast<const int4*>(ide);
        const float4* __restrict__ c4p = reinterpret_cast<const float4*>(Cj);
        const float4* __restrict__ x4p = reinterpret_cast<const float4*>(Xj);
        for (long long g = (n0 >> 2) + threadIdx.x; g < (n1 >> 2); g += kAccThreads) {
#ifndef M3S_GN_EC_REGS
            asm volatile("" ::: "memory");  // keep the per-edge constants out of loop-carried registers
#endif
            const uchar4 v4 = __ldg(v4p + g);
            const float4 q4 = __ldg(q4p + g);
            const int4 i4 = __ldg(i4p + g);
            const float4 c4 = __ldg(c4p + g);
            const float4 xa = __ldg(x4p + 3 * g), xb = __ldg(x4p + 3 * g + 1), xc = __ldg(x4p + 3 * g + 2);
            const unsigned char vv[4] = {v4.x, v4.y, v4.z, v4.w};
            const float qq[4] = {q4.x, q4.y, q4.z, q4.w};
            const int ii4[4] = {i4.x, i4.y, i4.z, i4.w};
            const float cc[4] = {c4.x, c4.y, c4.z, c4.w};
            const float xj[4][3] = {{xa.x, xa.y, xa.z}, {xa.w, xb.x, xb.y}, {xb.z, xb.w, xc.x}, {xc.y, xc.z, xc.w}};
            bool okp[4];
            float ci[4], xi[4][3];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                long long m = ii4[u];
                if (m < 0) m += N;  // numpy negative indexing (gauss_newton.py:125)
                okp[u] = vv[u] && (m >= 0) && (m < N) && (qq[u] > P.q_th) && (cc[u] > P.c_th);
                m = okp[u] ? m : 0;  // clamped gather: no branch in front of the load
                ci[u] = __ldg(Ci + m);
                xi[u][0] = __ldg(Xi + 3 * m); xi[u][1] = __ldg(Xi + 3 * m + 1); xi[u][2] = __ldg(Xi + 3 * m + 2);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (okp[u] && (ci[u] > P.c_th))
                    accumulate_point<MODE>(P, ec, xj[u][0], xj[u][1], xj[u][2], xi[u][0], xi[u][1],
                                           xi[u][2], qq[u], acc);
            }
        }
    } else {
        for (long long n = n0 + threadIdx.x; n < n1; n += kAccThreads) {
            if (!__ldg(ve + n)) continue;
            const float qc = __ldg(Qe + n);
            long long m = __ldg(ide + n);
            if (m < 0) m += N;  // numpy negative indexing (gauss_newton.py:125)
            if (m < 0 || m >= N) continue;
            if (!(qc > P.q_th) || !(__ldg(Cj + n) > P.c_th) || !(__ldg(Ci + m) > P.c_th)) continue;
            accumulate_point<MODE>(P, ec, __ldg(Xj + 3 * n), __ldg(Xj + 3 * n + 1), __ldg(Xj + 3 * n + 2),
                                   __ldg(Xi + 3 * m), __ldg(Xi + 3 * m + 1), __ldg(Xi + 3 * m + 2), qc,
                                   acc);
        }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NA; ++k) {
        const double v = warp_sum(acc[k]);
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NA) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kAccThreads / 32; ++w) s += red[w][threadIdx.x];
        out[threadIdx.x] = s;
    }
}

// ------------------------------------------------------------------------------------------------
// K3 (mixed precision, M3S_GN_PREC_MIXED): the same sums with the per-point arithmetic in fp32.
//
// What has to stay accurate is the *systematic* part of the residual e = T_ij Xj - Xi: the relative
// pose (s_ij R_ij, t_ij) is computed in fp64 and split into fp32 hi + lo parts, X' = hi-chain (3
// FMA) and e = (X'_hi - Xi) + lo-chain; the first difference is exact when the residual is small
// (Sterbenz), so e carries only the zero-mean rounding error of the hi chain (~2^-24 |X| per point).
// Zero-mean per-point errors average out over the ~1e5 points of an edge: the resulting pose
// perturbation is ~|e_round| / (sqrt(n) |J|) ~ 1e-9, three orders below the parity bar.  The same
// holds for the fp32 weights, Jacobian products and the per-thread fp32 partial sums (<= 32
// points); from the thread boundary upwards everything is fp64 with a fixed summation order.
//
// fp32 accumulators (29 / 35 registers instead of 58 / 70) also double the occupancy, and the block
// reduction goes through shared memory (one transposed pass) instead of 29 x 5 64-bit shuffles.
// ------------------------------------------------------------------------------------------------
struct EdgeConstF {
    float Rh[9], Rl[9];   // s_ij R_ij = Rh + Rl
    float th[3], tl[3];   // t_ij = th + tl
    float s_inv;          // 1 / s_i (Jacobian scale only)
};

constexpr float kHuberF = 1.345f;

// MUFU.SQRT / MUFU.RCP without the IEEE fix-up sequences (~1 ulp; mixed mode only)
__device__ __forceinline__ float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <int MODE>
__device__ __forceinline__ void accumulate_point_f32(const KernelParams& P, const EdgeConstF& ec,
                                                     float X0, float X1, float X2, float xix,
                                                     float xiy, float xiz, float qc, float* acc) {
    const float x = fmaf(ec.Rh[0], X0, fmaf(ec.Rh[1], X1, fmaf(ec.Rh[2], X2, ec.th[0])));
    const float y = fmaf(ec.Rh[3], X0, fmaf(ec.Rh[4], X1, fmaf(ec.Rh[5], X2, ec.th[1])));
    const float z = fmaf(ec.Rh[6], X0, fmaf(ec.Rh[7], X1, fmaf(ec.Rh[8], X2, ec.th[2])));
    const float xl = fmaf(ec.Rl[0], X0, fmaf(ec.Rl[1], X1, fmaf(ec.Rl[2], X2, ec.tl[0])));
    const float yl = fmaf(ec.Rl[3], X0, fmaf(ec.Rl[4], X1, fmaf(ec.Rl[5], X2, ec.tl[1])));
    const float zl = fmaf(ec.Rl[6], X0, fmaf(ec.Rl[7], X1, fmaf(ec.Rl[8], X2, ec.tl[2])));
    static_assert(MODE != M3S_GN_CALIB, "calib runs the fp64 pass (see m3s_gn_accumulate)");
    {
        const float e0 = (x - xix) + xl, e1 = (y - xiy) + yl, e2 = (z - xiz) + zl;
        float sw = P.sig_inv_f * sqrt_approx(qc);
        if (MODE == M3S_GN_POINTS) {
            const float dist = sqrt_approx(fmaf(xix, xix, fmaf(xiy, xiy, xiz * xiz))) + 1e-6f;
            sw *= rcp_approx(dist);
        }
        const float w2 = sw * sw;
        const float r0 = fabsf(sw * e0), r1 = fabsf(sw * e1), r2 = fabsf(sw * e2);
        const float hw = kHuberF * w2;
        const float w0 = (r0 < kHuberF) ? w2 : hw * rcp_approx(r0);
        const float w1 = (r1 < kHuberF) ? w2 : hw * rcp_approx(r1);
        const float w2c = (r2 < kHuberF) ? w2 : hw * rcp_approx(r2);
        const float w0y = w0 * y, w0z = w0 * z, w1x = w1 * x, w1z = w1 * z, w2x = w2c * x, w2y = w2c * y;
        const float w0x = w0 * x, w1y = w1 * y, w2z = w2c * z;
        acc[0] += w0; acc[1] += w1; acc[2] += w2c;
        acc[3] += w0z; acc[4] += w0y; acc[5] += w1z; acc[6] += w1x; acc[7] += w2y; acc[8] += w2x;
        acc[9] += w0x; acc[10] += w1y; acc[11] += w2z;
        acc[12] = fmaf(w1z, z, fmaf(w2y, y, acc[12]));
        acc[13] = fmaf(w0z, z, fmaf(w2x, x, acc[13]));
        acc[14] = fmaf(w0y, y, fmaf(w1x, x, acc[14]));
        acc[15] = fmaf(w2x, y, acc[15]);
        acc[16] = fmaf(w1x, z, acc[16]);
        acc[17] = fmaf(w0y, z, acc[17]);
        acc[18] = fmaf(w2y - w1y, z, acc[18]);
        acc[19] = fmaf(w0x - w2x, z, acc[19]);
        acc[20] = fmaf(w1x - w0x, y, acc[20]);
        acc[21] = fmaf(w0x, x, fmaf(w1y, y, fmaf(w2z, z, acc[21])));
        const float u0 = w0 * e0, u1 = w1 * e1, u2 = w2c * e2;
        acc[22] += u0; acc[23] += u1; acc[24] += u2;
        acc[25] += fmaf(y, u2, -z * u1);
        acc[26] += fmaf(z, u0, -x * u2);
        acc[27] += fmaf(x, u1, -y * u0);
        acc[28] = fmaf(x, u0, fmaf(y, u1, fmaf(z, u2, acc[28])));
    }
}

// ---- two points per instruction: the same arithmetic on the packed fp32x2 pipe ------------------
// Blackwell's FFMA2 / FMUL2 / FADD2 take a scalar register as a broadcast operand, so the edge
// constants cost nothing extra; lane .x / .y of every float2 is point u / u+1 of the trip.  A point
// that fails the gate is neutralised instead of branched around (coordinates -> 0, Q -> 0 => every
// weight is exactly 0 and every product finite), which also removes the per-point divergent branch.
// Accumulators are float2 (even / odd points), added once at the end: fixed order, reproducible.
#ifndef M3S_ACCF_SCALAR
#define M3S_ACCF_PACKED 1
#endif
__device__ __forceinline__ float2 bc2(float c) { return make_float2(c, c); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 neg2(float2 a) { return make_float2(-a.x, -a.y); }
__device__ __forceinline__ float2 sub2(float2 a, float2 b) { return __fadd2_rn(a, neg2(b)); }

template <int MODE>
__device__ __forceinline__ void accumulate_pair_f32(const KernelParams& P, const EdgeConstF& ec,
                                                    float2 X0, float2 X1, float2 X2, float2 xix,
                                                    float2 xiy, float2 xiz, float2 qc, float2* acc) {
    static_assert(MODE != M3S_GN_CALIB, "calib runs the fp64 pass (see m3s_gn_accumulate)");
    const float2 x = fma2(bc2(ec.Rh[0]), X0, fma2(bc2(ec.Rh[1]), X1, fma2(bc2(ec.Rh[2]), X2, bc2(ec.th[0]))));
    const float2 y = fma2(bc2(ec.Rh[3]), X0, fma2(bc2(ec.Rh[4]), X1, fma2(bc2(ec.Rh[5]), X2, bc2(ec.th[1]))));
    const float2 z = fma2(bc2(ec.Rh[6]), X0, fma2(bc2(ec.Rh[7]), X1, fma2(bc2(ec.Rh[8]), X2, bc2(ec.th[2]))));
    const float2 xl = fma2(bc2(ec.Rl[0]), X0, fma2(bc2(ec.Rl[1]), X1, fma2(bc2(ec.Rl[2]), X2, bc2(ec.tl[0]))));
    const float2 yl = fma2(bc2(ec.Rl[3]), X0, fma2(bc2(ec.Rl[4]), X1, fma2(bc2(ec.Rl[5]), X2, bc2(ec.tl[1]))));
    const float2 zl = fma2(bc2(ec.Rl[6]), X0, fma2(bc2(ec.Rl[7]), X1, fma2(bc2(ec.Rl[8]), X2, bc2(ec.tl[2]))));
    const float2 e0 = add2(sub2(x, xix), xl), e1 = add2(sub2(y, xiy), yl), e2 = add2(sub2(z, xiz), zl);
    float2 sw = mul2(bc2(P.sig_inv_f), make_float2(sqrt_approx(qc.x), sqrt_approx(qc.y)));
    if (MODE == M3S_GN_POINTS) {
        const float2 d2 = fma2(xix, xix, fma2(xiy, xiy, mul2(xiz, xiz)));
        const float2 dist = add2(make_float2(sqrt_approx(d2.x), sqrt_approx(d2.y)), bc2(1e-6f));
        sw = mul2(sw, make_float2(rcp_approx(dist.x), rcp_approx(dist.y)));
    }
    const float2 w2 = mul2(sw, sw);
    const float2 hw = mul2(bc2(kHuberF), w2);
    const float2 a0 = mul2(sw, e0), a1 = mul2(sw, e1), a2 = mul2(sw, e2);
    const float2 h0 = mul2(hw, make_float2(rcp_approx(fabsf(a0.x)), rcp_approx(fabsf(a0.y))));
    const float2 h1 = mul2(hw, make_float2(rcp_approx(fabsf(a1.x)), rcp_approx(fabsf(a1.y))));
    const float2 h2 = mul2(hw, make_float2(rcp_approx(fabsf(a2.x)), rcp_approx(fabsf(a2.y))));
    const float2 w0 = make_float2((fabsf(a0.x) < kHuberF) ? w2.x : h0.x, (fabsf(a0.y) < kHuberF) ? w2.y : h0.y);
    const float2 w1 = make_float2((fabsf(a1.x) < kHuberF) ? w2.x : h1.x, (fabsf(a1.y) < kHuberF) ? w2.y : h1.y);
    const float2 w2c = make_float2((fabsf(a2.x) < kHuberF) ? w2.x : h2.x, (fabsf(a2.y) < kHuberF) ? w2.y : h2.y);
    const float2 w0y = mul2(w0, y), w0z = mul2(w0, z), w1x = mul2(w1, x), w1z = mul2(w1, z);
    const float2 w2x = mul2(w2c, x), w2y = mul2(w2c, y);
    const float2 w0x = mul2(w0, x), w1y = mul2(w1, y), w2z = mul2(w2c, z);
    acc[0] = add2(acc[0], w0); acc[1] = add2(acc[1], w1); acc[2] = add2(acc[2], w2c);
    acc[3] = add2(acc[3], w0z); acc[4] = add2(acc[4], w0y); acc[5] = add2(acc[5], w1z);
    acc[6] = add2(acc[6], w1x); acc[7] = add2(acc[7], w2y); acc[8] = add2(acc[8], w2x);
    acc[9] = add2(acc[9], w0x); acc[10] = add2(acc[10], w1y); acc[11] = add2(acc[11], w2z);
    acc[12] = fma2(w1z, z, fma2(w2y, y, acc[12]));
    acc[13] = fma2(w0z, z, fma2(w2x, x, acc[13]));
    acc[14] = fma2(w0y, y, fma2(w1x, x, acc[14]));
    acc[15] = fma2(w2x, y, acc[15]);
    acc[16] = fma2(w1x, z, acc[16]);
    acc[17] = fma2(w0y, z, acc[17]);
    acc[18] = fma2(sub2(w2y, w1y), z, acc[18]);
    acc[19] = fma2(sub2(w0x, w2x), z, acc[19]);
    acc[20] = fma2(sub2(w1x, w0x), y, acc[20]);
    acc[21] = fma2(w0x, x, fma2(w1y, y, fma2(w2z, z, acc[21])));
    const float2 u0 = mul2(w0, e0), u1 = mul2(w1, e1), u2 = mul2(w2c, e2);
    acc[22] = add2(acc[22], u0); acc[23] = add2(acc[23], u1); acc[24] = add2(acc[24], u2);
    acc[25] = add2(acc[25], fma2(y, u2, neg2(mul2(z, u1))));
    acc[26] = add2(acc[26], fma2(z, u0, neg2(mul2(x, u2))));
    acc[27] = add2(acc[27], fma2(x, u1, neg2(mul2(y, u0))));
    acc[28] = fma2(x, u0, fma2(y, u1, fma2(z, u2, acc[28])));
}

#ifndef M3S_ACCF_MINBLOCKS
#define M3S_ACCF_MINBLOCKS 2
#endif

template <int MODE>
__global__ void __launch_bounds__(kAccThreads, M3S_ACCF_MINBLOCKS)
gn_accum_mixed_kernel(const Status* __restrict__ status, KernelParams P,
                      const double* __restrict__ poses, const EdgeMeta* __restrict__ meta,
                      const float* __restrict__ Xs, const float* __restrict__ Cs,
                      const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid,
                      const float* __restrict__ Q, long long N, int edge_begin, int nchunks,
                      long long chunk_pts, int vec_ok, double* __restrict__ partials) {
    if (status->stopped | status->failed) return;
    constexpr int NA = AccCount<MODE>::value;
    const int e = edge_begin + blockIdx.x / nchunks;
    const int chunk = blockIdx.x % nchunks;
    const EdgeMeta em = meta[e];
    double* out = partials + ((size_t)e * nchunks + chunk) * kEdgeStride;
    if (!em.active) return;

    __shared__ EdgeConstF ec_s;
    __shared__ float red[NA][kAccThreads];
    if (threadIdx.x == 0) {
        // sim3_relative (sim3_ops.py:96-126) in fp64, then split
        const double* Ti = poses + 8 * em.ix;
        const double* Tj = poses + 8 * em.jx;
        const double si_inv = 1.0 / Ti[7];
        const double sij = si_inv * Tj[7];
        const double qi_inv[4] = {-Ti[3], -Ti[4], -Ti[5], Ti[6]};
        double qij[4], R[9], Ri[9];
        quat_mul(qi_inv, Tj + 3, qij);
        quat_to_mat(qij, R);
        quat_to_mat(qi_inv, Ri);
        const double d0 = Tj[0] - Ti[0], d1 = Tj[1] - Ti[1], d2 = Tj[2] - Ti[2];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const double v = R[k] * sij;
            const float h = (float)v;
            ec_s.Rh[k] = h;
            ec_s.Rl[k] = (float)(v - (double)h);
        }
        const double t[3] = {(Ri[0] * d0 + Ri[1] * d1 + Ri[2] * d2) * si_inv,
                             (Ri[3] * d0 + Ri[4] * d1 + Ri[5] * d2) * si_inv,
                             (Ri[6] * d0 + Ri[7] * d1 + Ri[8] * d2) * si_inv};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float h = (float)t[k];
            ec_s.th[k] = h;
            ec_s.tl[k] = (float)(t[k] - (double)h);
        }
        ec_s.s_inv = (float)si_inv;
    }
    __syncthreads();
#ifdef M3S_ACCF_EC_SMEM
    const EdgeConstF& ec = ec_s;
#else
    const EdgeConstF ec = ec_s;  // 25 registers; re-reading them from shared memory every trip
                                 // cost 25 LDS per 4 points (l1tex data pipe 67 % busy, ncu s10a)
#endif

#ifdef M3S_ACCF_PACKED
    float2 acc2[NA];   // lane .x = even points of a trip, .y = odd points
#pragma unroll
    for (int k = 0; k < NA; ++k) acc2[k] = make_float2(0.0f, 0.0f);
#else
    float acc[NA];
#pragma unroll
    for (int k = 0; k < NA; ++k) acc[k] = 0.0f;
#endif

    const float* __restrict__ Xi = Xs + (size_t)em.ix * N * 3;
    const float* __restrict__ Ci = Cs + (size_t)em.ix * N;
    const float* __restrict__ Xj = Xs + (size_t)em.jx * N * 3;
    const float* __restrict__ Cj = Cs + (size_t)em.jx * N;
    const int32_t* __restrict__ ide = idx + (size_t)e * N;
    const uint8_t* __restrict__ ve = valid + (size_t)e * N;
    const float* __restrict__ Qe = Q + (size_t)e * N;

    const long long n0 = (long long)chunk * chunk_pts;
    const long long n1 = min(N, n0 + chunk_pts);
    if (vec_ok) {
        const uchar4* __restrict__ v4p = reinterpret_cast<const uchar4*>(ve);
        const float4* __restrict__ q4p = reinterpret_cast<const float4*>(Qe);
        const int4* __restrict__ i4p = reinterpret_cast<const int4*>(ide);
        const float4* __restrict__ c4p = reinterpret_cast<const float4*>(Cj);
        const float4* __restrict__ x4p = reinterpret_cast<const float4*>(Xj);
        const int Ni = (int)N;  // vec path is only taken for N < 2^31 (host check)
#ifndef M3S_ACCF_NO_PREFETCH  // default; the plain loop below is kept for A/B runs (B200: 1.38 -> 1.34 ms)
        // Two-stage software pipeline: the seven streaming vector loads of trip t+1 are issued before
        // the gathers and the arithmetic of trip t, so only the (idx-dependent) gather latency is
        // exposed per trip instead of stream + gather back to back.
        const int g_end = (int)(n1 >> 2);
        int g = (int)(n0 >> 2) + threadIdx.x;
        uchar4 v4 = make_uchar4(0, 0, 0, 0);
        float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f), c4 = q4, xa = q4, xb = q4, xc = q4;
        int4 i4 = make_int4(0, 0, 0, 0);
        if (g < g_end) {
            v4 = __ldg(v4p + g); q4 = __ldg(q4p + g); i4 = __ldg(i4p + g); c4 = __ldg(c4p + g);
            xa = __ldg(x4p + 3 * g); xb = __ldg(x4p + 3 * g + 1); xc = __ldg(x4p + 3 * g + 2);
        }
        for (; g < g_end; g += kAccThreads) {
            const unsigned char vv[4] = {v4.x, v4.y, v4.z, v4.w};
            const float qq[4] = {q4.x, q4.y, q4.z, q4.w};
            const int ii4[4] = {i4.x, i4.y, i4.z, i4.w};
            const float cc[4] = {c4.x, c4.y, c4.z, c4.w};
            const float xj[4][3] = {{xa.x, xa.y, xa.z}, {xa.w, xb.x, xb.y}, {xb.z, xb.w, xc.x}, {xc.y, xc.z, xc.w}};
            bool okp[4];
            float ci[4], xi[4][3];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int m = ii4[u];
                m += (m < 0) ? Ni : 0;  // numpy negative indexing (gauss_newton.py:125)
                okp[u] = vv[u] && ((unsigned)m < (unsigned)Ni) && (qq[u] > P.q_th) && (cc[u] > P.c_th);
                m = okp[u] ? m : 0;
                ci[u] = __ldg(Ci + m);
                const float* __restrict__ xp = Xi + 3 * (size_t)(unsigned)m;
                xi[u][0] = __ldg(xp); xi[u][1] = __ldg(xp + 1); xi[u][2] = __ldg(xp + 2);
            }
            const int gn = g + kAccThreads;
            if (gn < g_end) {
                v4 = __ldg(v4p + gn); q4 = __ldg(q4p + gn); i4 = __ldg(i4p + gn); c4 = __ldg(c4p + gn);
                xa = __ldg(x4p + 3 * gn); xb = __ldg(x4p + 3 * gn + 1); xc = __ldg(x4p + 3 * gn + 2);
            }
#ifdef M3S_ACCF_PACKED
            bool okf[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) okf[u] = okp[u] && (ci[u] > P.c_th);
#pragma unroll
            for (int u = 0; u < 4; u += 2) {
                if (okf[u] | okf[u + 1]) {
                    const bool a = okf[u], b = okf[u + 1];
                    accumulate_pair_f32<MODE>(
                        P, ec,
                        make_float2(a ? xj[u][0] : 0.f, b ? xj[u + 1][0] : 0.f),
                        make_float2(a ? xj[u][1] : 0.f, b ? xj[u + 1][1] : 0.f),
                        make_float2(a ? xj[u][2] : 0.f, b ? xj[u + 1][2] : 0.f),
                        make_float2(a ? xi[u][0] : 0.f, b ? xi[u + 1][0] : 0.f),
                        make_float2(a ? xi[u][1] : 0.f, b ? xi[u + 1][1] : 0.f),
                        make_float2(a ? xi[u][2] : 0.f, b ? xi[u + 1][2] : 0.f),
                        make_float2(a ? qq[u] : 0.f, b ? qq[u + 1] : 0.f), acc2);
                }
            }
#else
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (okp[u] && (ci[u] > P.c_th))
                    accumulate_point_f32<MODE>(P, ec, xj[u][0], xj[u][1], xj[u][2], xi[u][0], xi[u][1],
                                               xi[u][2], qq[u], acc);
            }
#endif
        }
#else
#ifndef M3S_ACCF_UNROLL
#define M3S_ACCF_UNROLL 1
#endif
        constexpr int kTripUnroll = M3S_ACCF_UNROLL;
#pragma unroll kTripUnroll
        for (int g = (int)(n0 >> 2) + threadIdx.x; g < (int)(n1 >> 2); g += kAccThreads) {
            const uchar4 v4 = __ldg(v4p + g);
            const float4 q4 = __ldg(q4p + g);
            const int4 i4 = __ldg(i4p + g);
            const float4 c4 = __ldg(c4p + g);
            const float4 xa = __ldg(x4p + 3 * g), xb = __ldg(x4p + 3 * g + 1), xc = __ldg(x4p + 3 * g + 2);
            const unsigned char vv[4] = {v4.x, v4.y, v4.z, v4.w};
            const float qq[4] = {q4.x, q4.y, q4.z, q4.w};
            const int ii4[4] = {i4.x, i4.y, i4.z, i4.w};
            const float cc[4] = {c4.x, c4.y, c4.z, c4.w};
            const float xj[4][3] = {{xa.x, xa.y, xa.z}, {xa.w, xb.x, xb.y}, {xb.z, xb.w, xc.x}, {xc.y, xc.z, xc.w}};
            bool okp[4];
            float ci[4], xi[4][3];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int m = ii4[u];
                m += (m < 0) ? Ni : 0;  // numpy negative indexing (gauss_newton.py:125)
                okp[u] = vv[u] && ((unsigned)m < (unsigned)Ni) && (qq[u] > P.q_th) && (cc[u] > P.c_th);
                m = okp[u] ? m : 0;
                ci[u] = __ldg(Ci + m);
                const float* __restrict__ xp = Xi + 3 * (size_t)(unsigned)m;
                xi[u][0] = __ldg(xp); xi[u][1] = __ldg(xp + 1); xi[u][2] = __ldg(xp + 2);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (okp[u] && (ci[u] > P.c_th))
                    accumulate_point_f32<MODE>(P, ec, xj[u][0], xj[u][1], xj[u][2], xi[u][0], xi[u][1],
                                               xi[u][2], qq[u], acc);
            }
        }
#endif
    } else {
        for (long long n = n0 + threadIdx.x; n < n1; n += kAccThreads) {
            if (!__ldg(ve + n)) continue;
            const float qc = __ldg(Qe + n);
            long long m = __ldg(ide + n);
            if (m < 0) m += N;
            if (m < 0 || m >= N) continue;
            if (!(qc > P.q_th) || !(__ldg(Cj + n) > P.c_th) || !(__ldg(Ci + m) > P.c_th)) continue;
#ifdef M3S_ACCF_PACKED   // one point in lane .x, a neutral point (Q = 0 => all weights 0) in lane .y
            accumulate_pair_f32<MODE>(P, ec, make_float2(__ldg(Xj + 3 * n), 0.f), make_float2(__ldg(Xj + 3 * n + 1), 0.f),
                                      make_float2(__ldg(Xj + 3 * n + 2), 0.f), make_float2(__ldg(Xi + 3 * m), 0.f),
                                      make_float2(__ldg(Xi + 3 * m + 1), 0.f), make_float2(__ldg(Xi + 3 * m + 2), 0.f),
                                      make_float2(qc, 0.f), acc2);
#else
            accumulate_point_f32<MODE>(P, ec, __ldg(Xj + 3 * n), __ldg(Xj + 3 * n + 1), __ldg(Xj + 3 * n + 2),
                                       __ldg(Xi + 3 * m), __ldg(Xi + 3 * m + 1), __ldg(Xi + 3 * m + 2),
                                       qc, acc);
#endif
        }
    }
    // block reduction in fp64 through one transposed shared-memory pass: thread (k, part) sums 32
    // of the 256 per-thread partials of accumulator k (rotated start: conflict-free), the 8 parts
    // are combined by a 3-step shuffle.  Fixed order => bitwise reproducible.
#ifdef M3S_ACCF_PACKED
#pragma unroll
    for (int k = 0; k < NA; ++k) red[k][threadIdx.x] = acc2[k].x + acc2[k].y;
#else
#pragma unroll
    for (int k = 0; k < NA; ++k) red[k][threadIdx.x] = acc[k];
#endif
    __syncthreads();
    const int lane = threadIdx.x & 31;
    constexpr int PASSES = (NA * 8 + kAccThreads - 1) / kAccThreads;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int id = ps * kAccThreads + threadIdx.x;
        const int k = min(id >> 3, NA - 1), part = id & 7;
        double s = 0.0;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) s += (double)red[k][part * 32 + ((i + lane) & 31)];
        s += __shfl_down_sync(0xffffffffu, s, 4);
        s += __shfl_down_sync(0xffffffffu, s, 2);
        s += __shfl_down_sync(0xffffffffu, s, 1);
        if (part == 0 && (id >> 3) < NA) out[id >> 3] = s;
    }
}

// ------------------------------------------------------------------------------------------------
// K3 (streamed): the mixed-precision pass with the five per-point streams (Xj 12 B, Q 4 B, idx 4 B,
// Cj 4 B, valid 1 B = 25 B per point) moved by the bulk-copy engine (cp.async.bulk, 1-D TMA) into
// a kStages-deep shared-memory ring, completion tracked by one mbarrier per stage.  ncu on the
// register-pipelined kernel above: 36 % of the stall samples are long-scoreboard waits on the
// FIRST use of the streamed registers -- one trip of lookahead does not cover the DRAM latency and
// deeper lookahead costs 25 registers per trip.  Here the lookahead is (kStages - 1) trips, costs
// no registers, and the freed registers hold the NEXT trip's (Ci, Xi) gathers while the current
// trip is being computed (two-stage software pipeline of the idx-dependent loads).  Arithmetic is
// accumulate_pair_f32 (packed fp32x2).  Needs N % 16 == 0, 16-B aligned arrays and chunks that are
// multiples of the stage size; anything else takes gn_accum_mixed_kernel.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok, spins = 0;
    do {
        asm volatile(
            "{\n.reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 24)) __trap();  // a copy that never lands must not hang the GPU
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// ring reads through 32-bit shared addresses (the stage base is a running register; a generic pointer made
// ptxas rebuild the shared-window base and the stage offset with ~20 uniform instructions per trip)
__device__ __forceinline__ float lds_f32(uint32_t a) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ int lds_s32(uint32_t a) {
    int v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ unsigned lds_u8(uint32_t a) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}

// wrapping shared-memory counter: old value, *a = (old >= wrap) ? 0 : old + 1.  (atom.add is turned into a
// warp-aggregated sequence of 12 instructions by ptxas even when one lane executes it; atom.inc is not, and the
// counter re-arms itself.)
__device__ __forceinline__ unsigned atoms_inc_u32(uint32_t a, unsigned wrap) {
    unsigned old;
    asm volatile("atom.shared.inc.u32 %0, [%1], %2;" : "=r"(old) : "r"(a), "r"(wrap) : "memory");
    return old;
}
__device__ __forceinline__ void l2_prefetch(const void* src, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

template <int TPB, int STAGES>
struct StreamCfg {
    static constexpr int kPts = TPB * 4;               // points per stage (4 per thread)
    static constexpr int kOffX = 0;                    // Xj  [kPts * 3] f32
    static constexpr int kOffQ = kPts * 12;            // Q   [kPts] f32
    static constexpr int kOffI = kPts * 16;            // idx [kPts] i32
    static constexpr int kOffC = kPts * 20;            // Cj  [kPts] f32
    static constexpr int kOffV = kPts * 24;            // valid [kPts] u8
    static constexpr int kStageBytes = kPts * 25;      // multiple of 16 (kPts % 16 == 0)
    static constexpr size_t ring_bytes = (size_t)kStageBytes * STAGES;
};
template <int MODE, int TPB, int STAGES>
constexpr size_t stream_smem_bytes() {
    const size_t ring = StreamCfg<TPB, STAGES>::ring_bytes;
    const size_t red = sizeof(float) * AccCount<MODE>::value * TPB;
    return ring > red ? ring : red;
}

// kSlim: Xj and Q of a trip are NOT carried in registers while its gathers are in flight; they are
// re-read from the ring when the trip is computed and the stage is handed back after that (one trip
// later).  The in-flight trip then costs 17 registers instead of 33, which is what separates 3 from 4
// resident CTAs per SM.
template <int MODE, int TPB, int STAGES, int MINB, bool kSlim>
__global__ void __launch_bounds__(TPB, MINB)
gn_accum_stream_kernel(const Status* __restrict__ status, KernelParams P,
                       const double* __restrict__ poses, const EdgeMeta* __restrict__ meta,
                       const float* __restrict__ Xs, const float* __restrict__ Cs,
                       const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid,
                       const float* __restrict__ Q, long long N, int edge_begin, int nchunks,
                       long long chunk_pts, double* __restrict__ partials) {
    if (status->stopped | status->failed) return;
    using Cfg = StreamCfg<TPB, STAGES>;
    constexpr int NA = AccCount<MODE>::value;
    const int e = edge_begin + blockIdx.x / nchunks;
    const int chunk = blockIdx.x % nchunks;
    const EdgeMeta em = meta[e];
    double* out = partials + ((size_t)e * nchunks + chunk) * kEdgeStride;
    if (!em.active) return;

    extern __shared__ __align__(128) unsigned char ring[];
    __shared__ EdgeConstF ec_s;
    __shared__ __align__(8) unsigned long long bars[STAGES];
    __shared__ unsigned int stage_cnt[STAGES];
    const int tid = threadIdx.x;
    if (tid < STAGES) stage_cnt[tid] = 0u;

    const float* __restrict__ Xi = Xs + (size_t)em.ix * N * 3;
    const float* __restrict__ Ci = Cs + (size_t)em.ix * N;
    const float* __restrict__ Xj = Xs + (size_t)em.jx * N * 3;
    const float* __restrict__ Cj = Cs + (size_t)em.jx * N;
    const int32_t* __restrict__ ide = idx + (size_t)e * N;
    const uint8_t* __restrict__ ve = valid + (size_t)e * N;
    const float* __restrict__ Qe = Q + (size_t)e * N;
    const long long n0 = (long long)chunk * chunk_pts;
    const long long n1 = min(N, n0 + chunk_pts);
    const int npts = (int)(n1 - n0);                       // multiple of 16
    const int trips = (npts + Cfg::kPts - 1) / Cfg::kPts;
    const uint32_t ring_s = smem_u32(ring), bars_s = smem_u32(bars), cnt_s = smem_u32(stage_cnt);

    auto issue = [&](int trip, int st) {  // one thread: arm the barrier of stage st (= trip % STAGES) and start its five copies
        const long long g0 = n0 + (long long)trip * Cfg::kPts;
        const uint32_t cnt = (uint32_t)min((long long)Cfg::kPts, n1 - g0);
        const uint32_t bar = bars_s + 8u * st, base = ring_s + (uint32_t)st * Cfg::kStageBytes;
        mbar_expect_tx(bar, cnt * 25u);
        bulk_g2s(base + Cfg::kOffX, Xj + 3 * g0, cnt * 12u, bar);
        bulk_g2s(base + Cfg::kOffQ, Qe + g0, cnt * 4u, bar);
        bulk_g2s(base + Cfg::kOffI, ide + g0, cnt * 4u, bar);
        bulk_g2s(base + Cfg::kOffC, Cj + g0, cnt * 4u, bar);
        bulk_g2s(base + Cfg::kOffV, ve + g0, cnt, bar);
#ifdef M3S_ACCS_GPF    // L2 prefetch of the gather targets assuming idx ~ identity + flow: measured SLOWER (1.208 vs 1.149 ms)
        l2_prefetch(Xi + 3 * g0, cnt * 12u);
        l2_prefetch(Ci + g0, cnt * 4u);
#endif
#ifdef M3S_ACCS_L2PF   // L2 prefetch of the stage M3S_ACCS_L2PF trips further on: measured SLOWER (8: 1.246, 16: 1.251 vs 1.201 ms)
        const long long gp = g0 + (long long)M3S_ACCS_L2PF * Cfg::kPts;
        if (gp < n1) {
            const uint32_t c2 = (uint32_t)min((long long)Cfg::kPts, n1 - gp);
            l2_prefetch(Xj + 3 * gp, c2 * 12u); l2_prefetch(Qe + gp, c2 * 4u);
            l2_prefetch(ide + gp, c2 * 4u); l2_prefetch(Cj + gp, c2 * 4u); l2_prefetch(ve + gp, c2);
        }
#endif
    };

    if (tid == 0) {
        for (int k = 0; k < STAGES; ++k) mbar_init(bars_s + 8u * k, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        for (int k = 0; k < STAGES && k < trips; ++k) issue(k, k);
        // sim3_relative (sim3_ops.py:96-126) in fp64, then split (as gn_accum_mixed_kernel)
        const double* Ti = poses + 8 * em.ix;
        const double* Tj = poses + 8 * em.jx;
        const double si_inv = 1.0 / Ti[7];
        const double sij = si_inv * Tj[7];
        const double qi_inv[4] = {-Ti[3], -Ti[4], -Ti[5], Ti[6]};
        double qij[4], R[9], Ri[9];
        quat_mul(qi_inv, Tj + 3, qij);
        quat_to_mat(qij, R);
        quat_to_mat(qi_inv, Ri);
        const double d0 = Tj[0] - Ti[0], d1 = Tj[1] - Ti[1], d2 = Tj[2] - Ti[2];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const double v = R[k] * sij;
            const float h = (float)v;
            ec_s.Rh[k] = h;
            ec_s.Rl[k] = (float)(v - (double)h);
        }
        const double t[3] = {(Ri[0] * d0 + Ri[1] * d1 + Ri[2] * d2) * si_inv,
                             (Ri[3] * d0 + Ri[4] * d1 + Ri[5] * d2) * si_inv,
                             (Ri[6] * d0 + Ri[7] * d1 + Ri[8] * d2) * si_inv};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float h = (float)t[k];
            ec_s.th[k] = h;
            ec_s.tl[k] = (float)(t[k] - (double)h);
        }
        ec_s.s_inv = (float)si_inv;
    }
    __syncthreads();
    const EdgeConstF ec = ec_s;
    const int Ni = (int)N;
    // gather bases held as register pairs (one IMAD.WIDE per address; ptxas otherwise re-derives
    // Xs + 3 N ix + 3 m from the kernel parameters in 64-bit pieces for every point: 3-4 instructions each)
    const float* Xi_r = Xi;
    const float* Ci_r = Ci;
    asm volatile("" : "+l"(Xi_r));
    asm volatile("" : "+l"(Ci_r));
    // fetches run in trip order: stage, wait parity and the points left are running values (no trip % STAGES)
    const int j0 = (tid >> 5) * 128 + (tid & 31);
    uint32_t f_st = 0, f_par = 0;
    int f_lim = npts - j0;

    float2 acc2[NA];
#pragma unroll
    for (int k = 0; k < NA; ++k) acc2[k] = make_float2(0.0f, 0.0f);

    // per-trip state of one thread: its 4 points of the stage
    struct Trip {
        float xj[kSlim ? 1 : 4][3];   // kSlim: unused
        float qq[kSlim ? 1 : 4];
        float ci[4], xi[4][3];
        bool ok[4];
    };
    auto release = [&](int trip, int st) {
        // Stage hand-back without anybody waiting: every warp counts itself out of the stage
        // (shared-memory atomic, ordered behind its ring reads in the LSU queue); the LAST warp to
        // leave re-arms the stage.  (A fixed producer warp that bar.sync'ed on the others spent half
        // of its time in that barrier: ncu s17, 12.6 % of all stall samples.)
        __syncwarp();
        if ((tid & 31) == 0) {
            const unsigned int old = atoms_inc_u32(cnt_s + 4u * (uint32_t)st, TPB / 32 - 1);
            if (old == TPB / 32 - 1) {
                if (trip + STAGES < trips) issue(trip + STAGES, st);
            }
        }
    };
    // wait for the stage, pull this thread's 4 points into registers, release the stage, start the
    // (Ci, Xi) gathers.  Returns with the gathers in flight.
    auto fetch = [&](int trip, Trip& T) {
        const int st = (int)f_st;
        mbar_wait(bars_s + 8u * f_st, f_par);
        const uint32_t sb = ring_s + f_st * (uint32_t)Cfg::kStageBytes;
        // Lane l of warp w takes points w*128 + l + 32*u (u = 0..3) of the stage: for a fixed u the
        // warp's 32 points are CONSECUTIVE, so with idx ~ identity + flow the (Ci, Xi) gathers of one
        // LDG touch 1-2 / 3-4 cache lines instead of 4 / 12 when a thread owns 4 adjacent points
        // (the L1 tag stage, one line per clock, was the limiter of the register-staged kernels).
        // The shared-memory reads are conflict-free either way (word stride 1, or 3 for Xj).
        const uint32_t a1 = sb + (uint32_t)j0, a4 = sb + 4u * (uint32_t)j0, a12 = sb + 12u * (uint32_t)j0;
        unsigned vv[4];
        int ii4[4];
        float cc[4], qg[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            // points past the end of a partial last stage read stale ring bytes: harmless, the gate
            // below fails on `in` and a failed point is neutralised before any arithmetic
            const bool in = 32 * u < f_lim;
            vv[u] = in ? lds_u8(a1 + Cfg::kOffV + 32 * u) : 0u;
            qg[u] = lds_f32(a4 + Cfg::kOffQ + 128 * u);
            ii4[u] = lds_s32(a4 + Cfg::kOffI + 128 * u);
            cc[u] = lds_f32(a4 + Cfg::kOffC + 128 * u);
            if constexpr (!kSlim) {
                T.qq[u] = qg[u];
                T.xj[u][0] = lds_f32(a12 + Cfg::kOffX + 384 * u);
                T.xj[u][1] = lds_f32(a12 + Cfg::kOffX + 384 * u + 4);
                T.xj[u][2] = lds_f32(a12 + Cfg::kOffX + 384 * u + 8);
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            int m = ii4[u];
            m += (m < 0) ? Ni : 0;  // numpy negative indexing (gauss_newton.py:125)
            T.ok[u] = vv[u] && ((unsigned)m < (unsigned)Ni) && (qg[u] > P.q_th) && (cc[u] > P.c_th);
            m = T.ok[u] ? m : 0;
            T.ci[u] = __ldg(Ci_r + (unsigned)m);
            const float* __restrict__ xp = Xi_r + 3 * (size_t)(unsigned)m;
            T.xi[u][0] = __ldg(xp); T.xi[u][1] = __ldg(xp + 1); T.xi[u][2] = __ldg(xp + 2);
        }
        // (placed after the gate, which consumes the shared-memory loads: they have landed)
        if constexpr (!kSlim) release(trip, st);
        if (++f_st == (uint32_t)STAGES) { f_st = 0; f_par ^= 1u; }
        f_lim -= Cfg::kPts;
    };
    auto compute = [&](int trip, const Trip& T) {
        bool okf[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) okf[u] = T.ok[u] && (T.ci[u] > P.c_th);
        float xj[4][3], qq[4];
        if constexpr (kSlim) {
            const unsigned char* base = ring + (size_t)(trip % STAGES) * Cfg::kStageBytes;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + 32 * u;
                const float* xp = reinterpret_cast<const float*>(base + Cfg::kOffX) + 3 * j;
                qq[u] = reinterpret_cast<const float*>(base + Cfg::kOffQ)[j];
                xj[u][0] = xp[0]; xj[u][1] = xp[1]; xj[u][2] = xp[2];
            }
        } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                qq[u] = T.qq[u];
                xj[u][0] = T.xj[u][0]; xj[u][1] = T.xj[u][1]; xj[u][2] = T.xj[u][2];
            }
        }
#ifdef M3S_ACCS_DUAL   // both pairs in one basic block (more ILP, 168 registers): measured 1.267 vs 1.257 ms
        const bool any4 = okf[0] | okf[1] | okf[2] | okf[3];
#endif
#ifdef M3S_ACCS_FASTPATH
        // every point of the warp's trip passes the gate (the common case): the same arithmetic on the raw
        // values, without the 14 selects per pair that neutralise a gated-out point (bit-identical).  Measured
        // SLOWER (1.176 vs 1.150 ms): the selects ride on the ALU pipe beside the packed FMA work, the vote and
        // the duplicated code do not
        if (__all_sync(0xffffffffu, okf[0] & okf[1] & okf[2] & okf[3])) {
#pragma unroll
            for (int u = 0; u < 4; u += 2)
                accumulate_pair_f32<MODE>(P, ec, make_float2(xj[u][0], xj[u + 1][0]), make_float2(xj[u][1], xj[u + 1][1]),
                                          make_float2(xj[u][2], xj[u + 1][2]), make_float2(T.xi[u][0], T.xi[u + 1][0]),
                                          make_float2(T.xi[u][1], T.xi[u + 1][1]), make_float2(T.xi[u][2], T.xi[u + 1][2]),
                                          make_float2(qq[u], qq[u + 1]), acc2);
        } else
#endif
#pragma unroll
        for (int u = 0; u < 4; u += 2) {
#ifdef M3S_ACCS_DUAL
            if (any4) {
#else
            if (okf[u] | okf[u + 1]) {
#endif
                const bool a = okf[u], b = okf[u + 1];
                accumulate_pair_f32<MODE>(
                    P, ec,
                    make_float2(a ? xj[u][0] : 0.f, b ? xj[u + 1][0] : 0.f),
                    make_float2(a ? xj[u][1] : 0.f, b ? xj[u + 1][1] : 0.f),
                    make_float2(a ? xj[u][2] : 0.f, b ? xj[u + 1][2] : 0.f),
                    make_float2(a ? T.xi[u][0] : 0.f, b ? T.xi[u + 1][0] : 0.f),
                    make_float2(a ? T.xi[u][1] : 0.f, b ? T.xi[u + 1][1] : 0.f),
                    make_float2(a ? T.xi[u][2] : 0.f, b ? T.xi[u + 1][2] : 0.f),
                    make_float2(a ? qq[u] : 0.f, b ? qq[u + 1] : 0.f), acc2);
            }
        }
        if constexpr (kSlim) release(trip, trip % STAGES);   // the arithmetic above consumed the ring reads
    };

    if (trips > 0) {
        Trip A, B;
        fetch(0, A);
        for (int t = 0; t < trips; t += 2) {   // A = trip t, its gathers in flight
            const bool has1 = t + 1 < trips;   // CTA-uniform
            if (has1) fetch(t + 1, B);
            compute(t, A);
            if (has1) {
                if (t + 2 < trips) fetch(t + 2, A);
                compute(t + 1, B);
            }
        }
    }

    // block reduction in fp64 (fixed order, bitwise reproducible): the ring is free once every
    // thread is past its last stage read; red[k][thread] aliases it
    __syncthreads();
    float* red = reinterpret_cast<float*>(ring);
#pragma unroll
    for (int k = 0; k < NA; ++k) red[k * TPB + tid] = acc2[k].x + acc2[k].y;
    __syncthreads();
    const int lane = tid & 31;
    constexpr int PARTS = TPB / 32;
    constexpr int PASSES = (NA * PARTS + TPB - 1) / TPB;
#pragma unroll
    for (int ps = 0; ps < PASSES; ++ps) {
        const int id = ps * TPB + tid;
        const int k = min(id / PARTS, NA - 1), part = id % PARTS;
        double sres = 0.0;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) sres += (double)red[k * TPB + part * 32 + ((i + lane) & 31)];
#pragma unroll
        for (int off = PARTS / 2; off > 0; off >>= 1) sres += __shfl_down_sync(0xffffffffu, sres, off);
        if (part == 0 && (id / PARTS) < NA) out[id / PARTS] = sres;
    }
}

// ------------------------------------------------------------------------------------------------
// Per-edge finalise: fixed-order sum over chunks, unpack to S (7x7) and v (7), transform by
// G = blockdiag(R_i, R_i, 1), write 28 upper + 7 (+1 spare) doubles per edge.  Edges outside
// [edge_begin, edge_end) and inactive edges get zeros (all-reduce friendly).  One 64-thread block
// per edge.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ int tri_index(int r, int c) {  // r <= c, 7x7 upper
    return r * 7 - (r * (r - 1)) / 2 + (c - r);
}

// Peer exchange (multi-GPU, edge-sharded): instead of writing its block into the local buffer and
// leaving the other ranks' slots to an NCCL all-reduce, the CTA of an OWNED edge stores the block
// straight into the receive buffer of every rank (its own included) through NVLink peer mappings
// (CUDA IPC); the LAST CTA of the launch then publishes the epoch number to every peer's flag word
// (threadfence.sys + ticket counter + st.release.sys).  gn_exchange_wait_kernel on the receiving
// side spins on its flags and copies the receive buffer (double-buffered by epoch parity, so a
// fast peer can never overwrite blocks that are still being read) into plan->edge_blocks.
constexpr int kMaxPeers = 8;
struct PeerTable {
    double* recv[kMaxPeers];                 // this epoch's receive buffer of rank r ([E, 36] f64)
    unsigned long long* flags[kMaxPeers];    // rank r's flag words [kMaxPeers]
    unsigned int* ticket;                    // local CTA counter
    unsigned long long epoch;
    int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

template <bool kPush>
__global__ void __launch_bounds__(64)
gn_edge_finalize_kernel(const Status* __restrict__ status, int mode,
                        const double* __restrict__ poses, const EdgeMeta* __restrict__ meta,
                        const double* __restrict__ partials, int nchunks, int edge_begin,
                        int edge_end, double* __restrict__ edge_blocks, const PeerTable peers) {
    // After the delta test has stopped the solve (identically on every rank) the remaining enqueued
    // iterations are no-ops -- but with the peer exchange attached they still run the push / publish
    // protocol (with zero blocks nobody reads), so the epoch sequence has no gaps and the
    // double-buffer argument (a rank can be at most one epoch ahead of a peer's copy) also holds
    // across consecutive solves.  A failed solve stops exchanging: the plan is unusable anyway.
    if (status->failed) return;
    const bool dead = status->stopped != 0;
    if (dead && !kPush) return;
    // kPush: the grid covers only the owned edges
    const int e = kPush ? edge_begin + (int)blockIdx.x : (int)blockIdx.x;
    const int tid = threadIdx.x;
    __shared__ double outb[kEdgeStride];
    double* out = kPush ? outb : edge_blocks + (size_t)e * kEdgeStride;
    const EdgeMeta em = meta[e];
    const bool skip = dead || (e < edge_begin || e >= edge_end || !em.active);
    if (skip) {
        if (tid < kEdgeStride) out[tid] = 0.0;
        if (!kPush) return;
    }
    if (!skip) {
    __shared__ double acc[kEdgeStride];
    __shared__ double S[7][7], v[7], G[7][7], T[7][7];
    const int NA = (mode == M3S_GN_CALIB) ? 35 : 29;
    if (tid < NA) {
        const double* p = partials + (size_t)e * nchunks * kEdgeStride + tid;
        double s = 0.0;
        for (int c = 0; c < nchunks; ++c) s += p[(size_t)c * kEdgeStride];
        acc[tid] = s;
    }
    if (tid < 49) G[tid / 7][tid % 7] = 0.0;
    __syncthreads();
    if (tid == 0) {
        const double* Ti = poses + 8 * em.ix;
        // R_i: Jj[c, :3] = s_i^-1 * R(q_i^-1) e_c = s_i^-1 * (row c of R(q_i)) etc.
        double Ri[9];
        quat_to_mat(Ti + 3, Ri);
        for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) { G[r][c] = Ri[r * 3 + c]; G[3 + r][3 + c] = Ri[r * 3 + c]; }
        G[6][6] = 1.0;
        if (mode == M3S_GN_CALIB) {
            int k = 0;
            for (int a = 0; a < 7; ++a)
                for (int b = a; b < 7; ++b) { S[a][b] = acc[k]; S[b][a] = acc[k]; ++k; }
            for (int a = 0; a < 7; ++a) v[a] = acc[28 + a];
        } else {
            const double si = 1.0 / Ti[7];
            const double* a = acc;
            for (int r = 0; r < 7; ++r)
                for (int c = 0; c < 7; ++c) S[r][c] = 0.0;
            S[0][0] = si * si * a[0]; S[1][1] = si * si * a[1]; S[2][2] = si * si * a[2];
            S[0][4] = si * a[3];  S[0][5] = -si * a[4];
            S[1][3] = -si * a[5]; S[1][5] = si * a[6];
            S[2][3] = si * a[7];  S[2][4] = -si * a[8];
            S[0][6] = si * a[9]; S[1][6] = si * a[10]; S[2][6] = si * a[11];
            S[3][3] = a[12]; S[4][4] = a[13]; S[5][5] = a[14];
            S[3][4] = -a[15]; S[3][5] = -a[16]; S[4][5] = -a[17];
            S[3][6] = a[18]; S[4][6] = a[19]; S[5][6] = a[20];
            S[6][6] = a[21];
            for (int r = 0; r < 7; ++r)
                for (int c = r + 1; c < 7; ++c) S[c][r] = S[r][c];
            v[0] = si * a[22]; v[1] = si * a[23]; v[2] = si * a[24];
            v[3] = a[25]; v[4] = a[26]; v[5] = a[27]; v[6] = a[28];
        }
    }
    __syncthreads();
    if (tid < 49) {  // T = S G
        const int r = tid / 7, c = tid % 7;
        double s = 0.0;
        for (int k = 0; k < 7; ++k) s += S[r][k] * G[k][c];
        T[r][c] = s;
    }
    __syncthreads();
    if (tid < 49) {  // A = G^T T (upper part stored)
        const int r = tid / 7, c = tid % 7;
        if (r <= c) {
            double s = 0.0;
            for (int k = 0; k < 7; ++k) s += G[k][r] * T[k][c];
            out[tri_index(r, c)] = s;
        }
    } else if (tid < 56) {  // b = G^T v
        const int r = tid - 49;
        double s = 0.0;
        for (int k = 0; k < 7; ++k) s += G[k][r] * v[k];
        out[28 + r] = s;
    } else if (tid == 56) {
        out[35] = 0.0;
    }
    }  // !skip
    if constexpr (kPush) {
        __shared__ unsigned int s_ticket;
        __syncthreads();
        if (tid < kEdgeStride) {
            const double val = outb[tid];
            for (int r = 0; r < peers.world; ++r) peers.recv[r][(size_t)e * kEdgeStride + tid] = val;
            __threadfence_system();
        }
        __syncthreads();
        if (tid == 0) s_ticket = atomicAdd(peers.ticket, 1u);
        __syncthreads();
        if (s_ticket == gridDim.x - 1) {  // every CTA's stores are fenced and counted: publish
            __threadfence_system();
            if (tid == 0) *peers.ticket = 0u;
            if (tid < peers.world) st_release_sys(peers.flags[tid] + peers.rank, peers.epoch);
        }
    }
}

// Wait until every rank has published `epoch`, then copy the receive buffer into the plan's block
// buffer.  A rank that owns no edges publishes from here (its finalize grid is empty).  Bounded
// spin (M3S_GN_XCH_TIMEOUT_MS, default 10 000; 0 = unbounded): a dead peer turns into status->failed +
// status->xch_timeout (reported as failed == 2 by m3s_gn_status) instead of a hung GPU.
__global__ void __launch_bounds__(256)
gn_exchange_wait_kernel(Status* __restrict__ status, const unsigned long long* __restrict__ my_flags,
                        const double* __restrict__ recv, double* __restrict__ edge_blocks,
                        long long count, const PeerTable peers, int publish_here,
                        unsigned long long timeout_ns) {
    if (status->failed) return;
    const bool dead = status->stopped != 0;  // see gn_edge_finalize_kernel: keep the epochs flowing
    __shared__ int s_ok;
    const int tid = threadIdx.x;
    if (tid == 0) s_ok = 1;
    __syncthreads();
    if (publish_here && blockIdx.x == 0 && tid < peers.world)
        st_release_sys(peers.flags[tid] + peers.rank, peers.epoch);
    if (tid < peers.world) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (ld_acquire_sys(my_flags + tid) < peers.epoch) {
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (timeout_ns != 0ull && t1 - t0 > timeout_ns) { s_ok = 0; break; }
            __nanosleep(100);
        }
    }
    __syncthreads();
    if (!s_ok) {
        if (tid == 0) { atomicExch(&status->xch_timeout, 1); atomicExch(&status->failed, 1); }
        return;
    }
    if (dead) return;
    for (long long i = (long long)blockIdx.x * blockDim.x + tid; i < count;
         i += (long long)gridDim.x * blockDim.x)
        edge_blocks[i] = recv[i];
}

// ------------------------------------------------------------------------------------------------
// Assemble dense H (lower + upper), g (as the extra row "-g" of the augmented matrix) from the
// per-edge blocks through CSR lists built on the host: deterministic, no atomics.
// Matrix storage: row-major [np + 1][np], np = dim rounded up to a multiple of 32; padded diagonal
// entries are 1 (so the padded factor is blockdiag(L, I)); row np holds -g (the right-hand side):
// factorising the augmented matrix performs the forward substitution on the fly.
// ------------------------------------------------------------------------------------------------
struct BlockRef {
    int row, col;     // local free ids; row == col for diagonal blocks; row > col otherwise
    int begin, end;   // range in the edge list
};

__global__ void __launch_bounds__(64)
gn_assemble_kernel(const Status* __restrict__ status, const BlockRef* __restrict__ blocks,
                   const int* __restrict__ list_edge, const int* __restrict__ list_sign,
                   const double* __restrict__ edge_blocks, double* __restrict__ Hm, int np, int dim,
                   double* __restrict__ g_out, int* __restrict__ bs_flags, int n_bs_flags) {
    if (status->stopped | status->failed) return;
    const BlockRef br = blocks[blockIdx.x];
    const int tid = threadIdx.x;
    // publication flags of the back-substitution: zeroed here instead of by a memset node of their own (the
    // previous solve's kernels are complete in stream order)
    if (blockIdx.x == 0) for (int i = tid; i < n_bs_flags; i += 64) bs_flags[i] = 0;
    if (tid < 49) {
        const int r = tid / 7, c = tid % 7;
        const int k = (r <= c) ? tri_index(r, c) : tri_index(c, r);
        double s = 0.0;
        for (int l = br.begin; l < br.end; ++l) s += edge_blocks[(size_t)list_edge[l] * kEdgeStride + k];
        if (br.row == br.col) {
            if (r == c) s += kReg;
            Hm[(size_t)(br.row * 7 + r) * np + br.col * 7 + c] = s;
        } else {
            s = -s;
            Hm[(size_t)(br.row * 7 + r) * np + br.col * 7 + c] = s;
            Hm[(size_t)(br.col * 7 + c) * np + br.row * 7 + r] = s;
        }
    } else if (tid < 56 && br.row == br.col) {
        const int r = tid - 49;
        double s = 0.0;
        for (int l = br.begin; l < br.end; ++l)
            s += (double)list_sign[l] * edge_blocks[(size_t)list_edge[l] * kEdgeStride + 28 + r];
        g_out[br.row * 7 + r] = s;
        Hm[(size_t)np * np + br.row * 7 + r] = -s;  // rhs row of the augmented matrix
    }
    (void)dim;
}

// identity on the padded diagonal (dim .. np-1); rhs row padding is zero from the memset
__global__ void gn_pad_diag_kernel(const Status* __restrict__ status, double* __restrict__ Hm,
                                   int np, int dim) {
    if (status->stopped | status->failed) return;
    const int k = dim + blockIdx.x * blockDim.x + threadIdx.x;
    if (k < np) Hm[(size_t)k * np + k] = 1.0;
}

// ------------------------------------------------------------------------------------------------
// Blocked right-looking Cholesky of the augmented matrix, NB = 32: ONE kernel per block step, a
// single dependent chain of 44 launches for the 200-keyframe graph (dim 1393).
//
// chol_step_kernel(k) holds three kinds of CTAs:
//   * update tiles (64x64 of the trailing lower triangle, incl. the rhs row): each tile solves the
//     64 + 64 panel rows it needs against L_k itself (one row per thread in registers, L_k from
//     Ldiag[k]) and applies the rank-32 update  A22 -= L21 L21^T  (4x4 register tiles).  The panel
//     solve is recomputed per tile (~1 us) instead of being a kernel of its own: every dependent
//     launch costs 2.5 us, a join of two graph branches 7 us (measured with the timeline below).
//     Tiles of tile-column 0 store their solved rows to a SEPARATE matrix Lm (the factor the
//     back-substitution reads), so no tile overwrites panel rows another tile still has to read.
//   * the diag CTA brings the NEXT diagonal block up to date on its own -- B = A[k+1,k] L_k^-T,
//     D = A[k+1,k+1] - B B^T -- and factorises it with one warp (chol32_warp), so the 32 dependent
//     column steps of block k+1 overlap the trailing update of step k instead of preceding step
//     k+1.  Update tiles leave block (k+1,k+1) alone.
//   * one CTA inverts L_k for the back-substitution, off the chain.
// chol_diag0_kernel factorises block 0 before the first step.
// ------------------------------------------------------------------------------------------------
constexpr int kStepThreads = 256;

// Optional timeline (M3S_GN_TIMELINE=1 at plan creation): %globaltimer stamps at entry / exit into
// tl[(kind * nblk + kb) * 2 + {0,1}]: kind 0 = diag CTA of block kb, 1 = update tile (0,0) of step
// kb until its panel rows are solved, 2 = the same tile until its update is stored, 3 = back-
// substitution CTA of block kb.  nullptr = disabled.
__device__ __forceinline__ void tl_stamp(long long* tl, int slot) {
    if (tl != nullptr) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        tl[slot] = t;
    }
}

// phase stamps inside the diag CTA (who = 0) / update tile (0,0) (who = 1) of block step kb: slot i of
// tl[(8 + 8 who) * nblk + 8 kb + i] (thread 0 only; call after a barrier)
__device__ __forceinline__ void tl_phase(long long* tl, int nblk, int who, int kb, int i) {
    if (tl != nullptr && threadIdx.x == 0) tl_stamp(tl, (8 + 8 * who) * nblk + 8 * kb + i);
}

// 1 / sqrt(d) for 1e-290 < d < 1e300: hardware seed (2^-22) + two Newton steps, no special cases
__device__ __forceinline__ double rsqrt_fast(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double h = 0.5 * d;
    double e = fma(-h * y, y, 0.5);
    y = fma(y, e, y);
    e = fma(-h * y, y, 0.5);
    return fma(y, e, y);
}

// 1 / d for 1e-290 < d < 1e300: hardware seed (~2^-20) + one third-order step (error e^3)
__device__ __forceinline__ double rcp_fast(double d) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double e = fma(-d, y, 1.0);
    const double t = fma(e, e, e);
    return fma(y, t, y);
}

// 32x32 Cholesky by ONE warp, row `lane` of the block in registers, no block barrier on the 32-step
// dependent chain.  Column j:
//   critical part (shuffles only): d_j and the entries a[j+1][j], a[j+2][j] are broadcast by three
//     independent shuffles, every lane derives rs = 1/sqrt(d_j) itself, scales its own entry
//     (l = a[j] rs) and updates the next TWO columns (a[j+1], a[j+2]) -- all the next two steps wait for;
//   bulk part: the scaled column goes to a double-buffered 32-double shared array and the updates
//     of columns >= j+3 are done one iteration LATER (all lanes read the same addresses: broadcast
//     LDS.128), in the shadow of the next column's shuffle -> rsqrt -> scale latency.  One
//     __syncwarp per column orders the shared-memory hand-over.
// Entries above the diagonal pick up finite garbage that nothing reads.
// (Block-wide version with a __syncthreads per column: 12 us per block; all-shuffle warp version:
//  ~5 us; measured with scripts/gn_solve_timeline.py.)
__device__ __forceinline__ bool chol32_warp(double (*W)[kNB + 1], double (*L)[kNB + 1], double* dinv,
                                            double (*col)[kNB], int lane) {
    double a[kNB];
#pragma unroll
    for (int c = 0; c < kNB; ++c) a[c] = (c <= lane) ? W[lane][c] : 0.0;
    bool bad = false;
    double lprev = 0.0;
#pragma unroll
    for (int j = 0; j < kNB; ++j) {
        double d = __shfl_sync(0xffffffffu, a[j], j);
        const double n1 = __shfl_sync(0xffffffffu, a[j], (j + 1) & 31);  // a[j+1][j]
        const double n2 = __shfl_sync(0xffffffffu, a[j], (j + 2) & 31);  // a[j+2][j]
        if (j > 0) {
            // bulk update of column j-1 (columns >= j+2 of it; j+0, j+1 were done in its critical part)
            const double2* cb = reinterpret_cast<const double2*>(col[(j - 1) & 1]);
#pragma unroll
            for (int c = j + 2; c < kNB; ++c) {
                const double2 v = cb[c >> 1];
                a[c] -= lprev * ((c & 1) ? v.y : v.x);
            }
        }
        if (!((d > 1e-290) && (d < 1e300))) { bad = true; d = 1.0; }  // warp-uniform
        const double rs = rsqrt_fast(d);
        const double l = (lane == j) ? d * rs : a[j] * rs;
        a[j] = l;
        if (lane == j) dinv[j] = rs;
        if (j + 1 < kNB) a[j + 1] -= l * (n1 * rs);
        if (j + 2 < kNB) a[j + 2] -= l * (n2 * rs);
        if (j + 3 < kNB) {
            col[j & 1][lane] = l;
            __syncwarp();
        }
        lprev = l;
    }
#pragma unroll
    for (int c = 0; c < kNB; ++c) L[lane][c] = (c <= lane) ? a[c] : 0.0;
    return !bad;
}

// 32x32 Cholesky by the whole CTA (>= 64 threads), blocked in 4 panels of 8 columns.  Warp 0
// factorises a panel with the rows in registers (8 entries per lane); per column the dependent chain
// is shuffle(d) -> reciprocal -> update of the <= 7 remaining panel columns, whose multipliers
// are shuffled before the reciprocal (independent of it) -- no shared-memory traffic and no
// bulk update on the chain.  The trailing update of the columns right of the panel (<= 300 entries
// x 8 FMAs) is spread over all threads between two barriers.  Measured against the one-warp
// left-looking version above (scripts/gn_solve_timeline.py, B200): diag CTA 8.84 vs 8.97 us per block
// step, solve 0.609 vs 0.616 ms -- the 32-column chain is not what the ~5 us of a block are made of.
constexpr int kCholPanel = 8;
__device__ __forceinline__ bool chol32_block(double (*W)[kNB + 1], double (*L)[kNB + 1], double* dinv,
                                             int* bad_flag) {
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31;
#pragma unroll 1
    for (int c0 = 0; c0 < kNB; c0 += kCholPanel) {
        if (tid < 32) {
            double a[kCholPanel];
#pragma unroll
            for (int j = 0; j < kCholPanel; ++j) a[j] = W[lane][c0 + j];  // rows < c0 + j: unused garbage-free zeros
            bool bad = false;
            double Dj[kCholPanel];
            // LDL^T-style inside the panel: columns stay UNSCALED (u = l * sqrt(D)) while the chain
            // runs, so a column step is shuffle -> reciprocal (MUFU + 3 fp64 ops) -> 2 fp64 ops
            // instead of shuffle -> rsqrt (MUFU + 6) -> 3: fp64 latency (~25 clk) is what the 32-step
            // chain is made of.  The square roots are taken afterwards, all 8 independent.
#pragma unroll
            for (int j = 0; j < kCholPanel; ++j) {
                double d = __shfl_sync(0xffffffffu, a[j], c0 + j);
                double nv[kCholPanel];
#pragma unroll
                for (int c = j + 1; c < kCholPanel; ++c) nv[c] = __shfl_sync(0xffffffffu, a[j], c0 + c);
                if (!((d > 1e-290) && (d < 1e300))) { bad = true; d = 1.0; }  // warp-uniform
                Dj[j] = d;
                const double t = a[j] * rcp_fast(d);
#pragma unroll
                for (int c = j + 1; c < kCholPanel; ++c) a[c] -= t * nv[c];
            }
#pragma unroll
            for (int j = 0; j < kCholPanel; ++j) {
                const double rs = rsqrt_fast(Dj[j]);
                const double l = (lane == c0 + j) ? Dj[j] * rs : a[j] * rs;
                if (lane >= c0 + j) W[lane][c0 + j] = l;
                if (lane == c0 + j) dinv[c0 + j] = rs;
            }
            if (bad && lane == 0) *bad_flag = 1;
        }
        __syncthreads();
        const int m0 = c0 + kCholPanel;          // first trailing row / column
        const int m = kNB - m0;                  // trailing size
        for (int i = tid; i < m * m; i += nt) {
            const int r = m0 + i / m, c = m0 + i % m;
            if (c <= r) {
                double sacc = 0.0;
#pragma unroll
                for (int j = 0; j < kCholPanel; ++j) sacc += W[r][c0 + j] * W[c][c0 + j];
                W[r][c] -= sacc;
            }
        }
        __syncthreads();
    }
    for (int i = tid; i < kNB * kNB; i += nt) {
        const int r = i / kNB, c = i % kNB;
        L[r][c] = (c <= r) ? W[r][c] : 0.0;
    }
    __syncthreads();
    return *bad_flag == 0;
}

// Column `lane` of the inverse of the lower-triangular 32x32 block L (shared memory), by forward
// substitution with the column in registers; written as row-major Linv[r][lane].
__device__ __forceinline__ void tri_inverse32(double (*L)[kNB + 1], const double* dinv,
                                              double* __restrict__ out, int lane) {
    double x[kNB];
#pragma unroll
    for (int r = 0; r < kNB; ++r) {
        double sacc = (r == lane) ? 1.0 : 0.0;
#pragma unroll
        for (int k = 0; k < r; ++k) sacc -= L[r][k] * x[k];
        x[r] = sacc * dinv[r];
    }
#pragma unroll
    for (int r = 0; r < kNB; ++r) out[r * kNB + lane] = x[r];
}

// Rows of X L^T = A for a 32-row block against the 32x32 factor L (shared memory), 8 threads per
// row: lane g of a row's 8-lane group owns columns g, g+8, g+16, g+24 of the row in registers.  Per
// column j the owner scales x_j by dinv[j] and broadcasts it inside the group (two 32-bit shuffles);
// every lane then updates its <= 4 columns right of j.  The dependent chain of a row is 32 x (DMUL +
// shuffle + DFMA) instead of 32 x (DMUL + up to 31 DFMAs issued by one thread).  Needs all 256
// threads of the CTA (32 rows x 8); X is overwritten in place.
__device__ __forceinline__ void trsm32_rows8(double (*X)[kNB + 1], double (*L)[kNB + 1],
                                             const double* dinv, int tid) {
    const int r = tid >> 3, g = tid & 7;
    double x[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) x[i] = X[r][g + 8 * i];
#pragma unroll
    for (int j = 0; j < kNB; ++j) {
        const int oi = j >> 3, og = j & 7;            // compile-time after unrolling
        double xj = x[oi] * dinv[j];                  // meaningful on the owner lane only
        xj = __shfl_sync(0xffffffffu, xj, og, 8);
        if (g == og) x[oi] = xj;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            // column c = g + 8 i is right of j  <=>  i > oi or (i == oi and g > og)
            if (i > oi) x[i] -= xj * L[g + 8 * i][j];
            else if (i == oi && g > og) x[i] -= xj * L[g + 8 * i][j];
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) X[r][g + 8 * i] = x[i];
}

// W -= B B^T on the lower triangle, 2 x 2 register tiles (136 threads busy): 4 LDS.64 per 4 DFMA
// instead of 2 per DFMA.
__device__ __forceinline__ void syrk32_lower(double (*Wm)[kNB + 1], double (*Bm)[kNB + 1], int tid) {
    if (tid < 136) {
        int ti = (int)((sqrtf(8.0f * (float)tid + 1.0f) - 1.0f) * 0.5f);
        while (ti * (ti + 1) / 2 > tid) --ti;
        while ((ti + 1) * (ti + 2) / 2 <= tid) ++ti;
        const int tj = tid - ti * (ti + 1) / 2;
        const int r = 2 * ti, c = 2 * tj;
        double s00 = 0.0, s01 = 0.0, s10 = 0.0, s11 = 0.0;
#pragma unroll 8
        for (int j = 0; j < kNB; ++j) {
            const double a0 = Bm[r][j], a1 = Bm[r + 1][j], b0 = Bm[c][j], b1 = Bm[c + 1][j];
            s00 += a0 * b0; s01 += a0 * b1; s10 += a1 * b0; s11 += a1 * b1;
        }
        Wm[r][c] -= s00;
        if (c + 1 <= r) Wm[r][c + 1] -= s01;
        Wm[r + 1][c] -= s10;
        Wm[r + 1][c + 1] -= s11;
    }
}

struct DiagSmem {
    double W[kNB][kNB + 1];   // block being factorised
    double L[kNB][kNB + 1];   // previous factor, then this block's factor
    double Bm[kNB][kNB + 1];  // B = L[kb, kb-1]
    double dinv[kNB];
    __align__(16) double col[2][kNB];  // chol32_warp's column hand-over
    int bad;
};

// Factorise diagonal block kb (whole CTA, any blockDim >= 32):
//   Ldiag[kb] <- chol(A[kb,kb] - B B^T),  B = A[kb,kb-1] L_{kb-1}^-T   (kb > 0),
// with A[kb,kb-1] read from Hm (final after step kb-2; step kb-1 does not touch it because solved
// panel rows go to Lm, not back into Hm).
__device__ __forceinline__ void chol_diag_block(DiagSmem& sm, Status* __restrict__ status,
                                                const double* __restrict__ Hm, double* __restrict__ Lm,
                                                double* __restrict__ Ldiag, double* __restrict__ Linv,
                                                int np, int kb, long long* __restrict__ tl) {
    const int tid = threadIdx.x, nt = blockDim.x;
    const int k0 = kb * kNB;
    const int nblk = np / kNB;
    if (tid == 0) tl_stamp(tl, (0 * nblk + kb) * 2);
    for (int i = tid; i < kNB * kNB; i += nt) {
        const int r = i / kNB, c = i % kNB;
        sm.W[r][c] = (c <= r) ? Hm[(size_t)(k0 + r) * np + k0 + c] : 0.0;
    }
    if (tid == 0) sm.bad = 0;
    if (kb > 0) {
        const double* Lp = Ldiag + (size_t)(kb - 1) * kNB * kNB;
        for (int i = tid; i < kNB * kNB; i += nt) {
            sm.L[i / kNB][i % kNB] = Lp[i];
            sm.Bm[i / kNB][i % kNB] = Hm[(size_t)(k0 + i / kNB) * np + (k0 - kNB) + (i % kNB)];
        }
        __syncthreads();
        tl_phase(tl, nblk, 0, kb, 0);   // loads landed
        if (tid < kNB) sm.dinv[tid] = 1.0 / sm.L[tid][tid];
        __syncthreads();
        if (nt == kStepThreads) {
            trsm32_rows8(sm.Bm, sm.L, sm.dinv, tid);
            __syncthreads();
            tl_phase(tl, nblk, 0, kb, 1);   // B solved
            syrk32_lower(sm.W, sm.Bm, tid);
        } else {
            if (tid < kNB) {
                // row tid of B: right-looking substitution against L_{kb-1}, row in registers
                double x[kNB];
#pragma unroll
                for (int c = 0; c < kNB; ++c) x[c] = sm.Bm[tid][c];
#pragma unroll
                for (int j = 0; j < kNB; ++j) {
                    const double xj = x[j] * sm.dinv[j];
                    x[j] = xj;
#pragma unroll
                    for (int c = j + 1; c < kNB; ++c) x[c] -= xj * sm.L[c][j];
                }
#pragma unroll
                for (int c = 0; c < kNB; ++c) sm.Bm[tid][c] = x[c];
            }
            __syncthreads();
            // W -= B B^T (lower part)
            for (int i = tid; i < kNB * kNB; i += nt) {
                const int r = i / kNB, c = i % kNB;
                if (c <= r) {
                    double s = 0.0;
#pragma unroll 8
                    for (int j = 0; j < kNB; ++j) s += sm.Bm[r][j] * sm.Bm[c][j];
                    sm.W[r][c] -= s;
                }
            }
        }
    }
    __syncthreads();
    tl_phase(tl, nblk, 0, kb, 2);   // W up to date
#ifdef M3S_CHOL_WARP   // the one-warp left-looking version (A/B)
    if (tid < 32) {
        if (!chol32_warp(sm.W, sm.L, sm.dinv, sm.col, tid) && tid == 0) sm.bad = 1;
    }
    __syncthreads();
#else
    chol32_block(sm.W, sm.L, sm.dinv, &sm.bad);
#endif
    tl_phase(tl, nblk, 0, kb, 3);   // factorised
    if (sm.bad) {
        if (tid == 0) status->chol_fail = 1;
        return;
    }
    double* Ld = Ldiag + (size_t)kb * kNB * kNB;
    for (int i = tid; i < kNB * kNB; i += nt) {
        const int r = i / kNB, c = i % kNB;
        Ld[i] = sm.L[r][c];
        if (c <= r) Lm[(size_t)(k0 + r) * np + k0 + c] = sm.L[r][c];
    }
    // the last block has no following step whose spare CTA could invert it
    if (kb == nblk - 1 && tid < 32) tri_inverse32(sm.L, sm.dinv, Linv + (size_t)kb * kNB * kNB, tid);
    __syncthreads();
    if (tid == 0) tl_stamp(tl, (0 * nblk + kb) * 2 + 1);
}

__global__ void __launch_bounds__(kStepThreads)
chol_diag0_kernel(Status* __restrict__ status, const double* __restrict__ Hm, double* __restrict__ Lm,
                  double* __restrict__ Ldiag, double* __restrict__ Linv, int np,
                  long long* __restrict__ tl) {
    if (status->stopped | status->failed | status->chol_fail) return;
    __shared__ DiagSmem sm;
    chol_diag_block(sm, status, Hm, Lm, Ldiag, Linv, np, 0, tl);
}

constexpr int kSyrkTile = 64;

struct TileSmem {
    double D[kNB][kNB + 1];        // L_k
    double dinv[kNB];
    double A[kSyrkTile][kNB + 1];  // panel rows of the tile's row range
    double B[kSyrkTile][kNB + 1];  // panel rows of the tile's column range
};

union StepSmem {
    TileSmem t;
    DiagSmem d;
};

// invert L_kb for the back-substitution (off the chain; the last block is inverted by its own diag CTA)
__device__ __forceinline__ void chol_invert_block(TileSmem& sm, const double* Ldiag,
                                                  double* __restrict__ Linv, int kb) {
    const int tid = threadIdx.x;
    const double* Ld = Ldiag + (size_t)kb * kNB * kNB;
    for (int i = tid; i < kNB * kNB; i += kStepThreads) sm.D[i / kNB][i % kNB] = __ldcg(Ld + i);
    __syncthreads();
    if (tid < kNB) sm.dinv[tid] = 1.0 / sm.D[tid][tid];
    __syncthreads();
    if (tid < 32) tri_inverse32(sm.D, sm.dinv, Linv + (size_t)kb * kNB * kNB, tid);
    __syncthreads();
}

// update tile (tr, tc) of block step kb (see chol_step_kernel).  Everything another CTA may have
// written earlier in the same launch (persistent kernel) is read with ld.global.cg.
__device__ __forceinline__ void chol_tile_body(TileSmem& sm, double* Hm, double* __restrict__ Lm,
                                               const double* Ldiag, int np, int kb, int tr, int tc,
                                               long long* __restrict__ tl) {
    const int tid = threadIdx.x;
    const int nblk = np / kNB;
    const int k0 = kb * kNB;
    const bool first = (tr == 0) && (tc == 0);
    const bool stamp = (tl != nullptr) && first && (tid == 0);
    if (stamp) tl_stamp(tl, (1 * nblk + kb) * 2);
    const int base = k0 + kNB;
    const int r0 = base + tr * kSyrkTile, c0 = base + tc * kSyrkTile;
    const int ty = tid / 16, tx = tid % 16;  // 16 x 16 threads, 4 x 4 outputs each (strided by 16)
    // the tile's current values first: their latency hides behind the panel solve
    double cval[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gr = r0 + ty + 16 * i;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gc = c0 + tx + 16 * j;
            cval[i][j] = (gr <= np && gc < np && gc <= gr) ? __ldcg(Hm + (size_t)gr * np + gc) : 0.0;
        }
    }
    {
        const double* Ld = Ldiag + (size_t)kb * kNB * kNB;
        for (int i = tid; i < kNB * kNB; i += kStepThreads) sm.D[i / kNB][i % kNB] = __ldcg(Ld + i);
    }
    for (int i = tid; i < kSyrkTile * kNB; i += kStepThreads) {
        const int r = i / kNB, c = i % kNB;
        const int gr = r0 + r, gc = c0 + r;
        sm.A[r][c] = (gr <= np) ? __ldcg(Hm + (size_t)gr * np + k0 + c) : 0.0;
        if (tc != tr) sm.B[r][c] = (gc < np) ? __ldcg(Hm + (size_t)gc * np + k0 + c) : 0.0;
    }
    __syncthreads();
    if (first) tl_phase(tl, nblk, 1, kb, 0);   // loads landed
    if (tid < kNB) sm.dinv[tid] = 1.0 / sm.D[tid][tid];
    __syncthreads();
    // panel solve: threads 0..63 take the rows of A, threads 64..127 the rows of B
    if (tid < kSyrkTile || (tid < 2 * kSyrkTile && tc != tr)) {
        double (*P)[kNB + 1] = (tid < kSyrkTile) ? sm.A : sm.B;
        const int r = tid & (kSyrkTile - 1);
        double x[kNB];
#pragma unroll
        for (int c = 0; c < kNB; ++c) x[c] = P[r][c];
#pragma unroll
        for (int j = 0; j < kNB; ++j) {
            const double xj = x[j] * sm.dinv[j];
            x[j] = xj;
#pragma unroll
            for (int c = j + 1; c < kNB; ++c) x[c] -= xj * sm.D[c][j];
        }
#pragma unroll
        for (int c = 0; c < kNB; ++c) P[r][c] = x[c];
    }
    __syncthreads();
    if (stamp) { tl_stamp(tl, (1 * nblk + kb) * 2 + 1); tl_stamp(tl, (2 * nblk + kb) * 2); }
    double (*Bp)[kNB + 1] = (tc == tr) ? sm.A : sm.B;
    if (tc == 0) {  // this tile column publishes the solved rows (the factor's block column kb)
        for (int i = tid; i < kSyrkTile * kNB; i += kStepThreads) {
            const int r = i / kNB, c = i % kNB;
            if (r0 + r <= np) Lm[(size_t)(r0 + r) * np + k0 + c] = sm.A[r][c];
        }
    }
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.0;
#pragma unroll 8
    for (int k = 0; k < kNB; ++k) {
        double a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { a[i] = sm.A[ty + 16 * i][k]; b[i] = Bp[tx + 16 * i][k]; }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] += a[i] * b[j];
    }
    if (first) tl_phase(tl, nblk, 1, kb, 1);   // rank-32 products done (thread 0)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gr = r0 + ty + 16 * i;
        if (gr > np) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gc = c0 + tx + 16 * j;
            if (!(gc < np && gc <= gr)) continue;
            // block (kb+1, kb+1) belongs to the diag CTA of this launch
            if (first && gr < base + kNB) continue;
            Hm[(size_t)gr * np + gc] = cval[i][j] - acc[i][j];
        }
    }
    if (stamp) tl_stamp(tl, (2 * nblk + kb) * 2 + 1);
    __syncthreads();  // persistent kernel: the shared panels are reloaded by this CTA's next tile
}

// 1-D grid: block 0 = diag CTA of block kb + 1 (first, so it is dispatched first: it is the longest
// CTA of the launch), block 1 inverts L_kb, blocks 2.. = the tiles of the trailing lower triangle,
// row by row: tile rows 0 .. ntc-1 hold tr + 1 tiles each, a last tile row `ntc` (present when the
// rhs row starts a new tile row) holds ntc tiles.
__global__ void __launch_bounds__(kStepThreads, 2)  // 2 CTAs / SM: the 255 tiles of step 0 in one wave
chol_step_kernel(Status* __restrict__ status, double* __restrict__ Hm, double* __restrict__ Lm,
                 double* __restrict__ Ldiag, double* __restrict__ Linv, int np, int kb, int ntc,
                 long long* __restrict__ tl) {
    // Programmatic dependent launch (when the launch carries the attribute; no-ops otherwise): let
    // the NEXT block step's CTAs be scheduled while this grid is still running, and do not touch
    // anything the previous grid wrote before it has completed and flushed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (status->stopped | status->failed | status->chol_fail) return;
    __shared__ StepSmem sm;
    const int nblk = np / kNB;
    if (blockIdx.x == 0) {
        if (kb + 1 < nblk) chol_diag_block(sm.d, status, Hm, Lm, Ldiag, Linv, np, kb + 1, tl);
        return;
    }
    if (blockIdx.x == 1) {
        if (kb != nblk - 1) chol_invert_block(sm.t, Ldiag, Linv, kb);
        return;
    }
    int tr, tc;
    {
        const int t = (int)blockIdx.x - 2;
        const int ntri = ntc * (ntc + 1) / 2;
        if (t >= ntri) {
            tr = ntc;
            tc = t - ntri;
        } else {
            tr = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
            while (tr * (tr + 1) / 2 > t) --tr;
            while ((tr + 1) * (tr + 2) / 2 <= t) ++tr;
            tc = t - tr * (tr + 1) / 2;
        }
    }
    chol_tile_body(sm.t, Hm, Lm, Ldiag, np, kb, tr, tc, tl);
}

// ------------------------------------------------------------------------------------------------
// Persistent factorisation (opt-in, M3S_GN_CHOL=fused): the same right-looking blocked Cholesky as
// above in ONE cooperative launch, as a dataflow of two roles:
//   * CTA 0 is the CHAIN: iteration c brings diagonal block c up to date with the last block step
//     (B = A[c,c-1] L_{c-1}^-T against the factor it still holds in shared memory, W = A[c,c] - B B^T),
//     factorises it and publishes L_c.  It waits for ONE tile only: tile (0,0) of step c-2.
//   * CTAs 1.. are WORKERS: the 64x64 tiles of the trailing update of step kb, dealt round-robin
//     (chol_tile_body, shared with the step kernel); a worker step needs L_kb (flag from the chain)
//     and all tiles of step kb-1 (barrier among the workers); the last worker inverts L_kb.
// Every CTA is co-resident by construction (cudaLaunchCooperativeKernel).  Flags and the barrier are
// monotonic words in global memory (zeroed by a memset before the launch): publish = fence +
// st.release, wait = ld.acquire poll with a bound, so a lost CTA turns into status->failed instead of
// a hung device.  Everything another CTA may have written during the launch is read with
// ld.global.cg (L1 is not coherent across SMs).
// ------------------------------------------------------------------------------------------------
constexpr int kFusedThreads = 256;
constexpr int kGridSpinLimit = 1 << 21;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

struct FusedChainSmem {
    double W[kNB][kNB + 1];    // block being factorised
    double L[kNB][kNB + 1];    // the most recent factor (kept across steps: next step's B-solve)
    double Bm[kNB][kNB + 1];   // A[kb,kb-1], then B = A[kb,kb-1] L_{kb-1}^-T
    double dinv[kNB];          // 1 / diag of the most recent factor
    int bad;
};
union FusedSmem {
    TileSmem t;
    FusedChainSmem c;
};

// chain CTA: factorise diagonal block kb; sm.L / sm.dinv hold L_{kb-1} on entry (kb > 0), L_kb on exit
__device__ __forceinline__ void fused_chain_block(FusedChainSmem& sm, Status* __restrict__ status,
                                                  const double* Hm, double* __restrict__ Lm,
                                                  double* __restrict__ Ldiag, double* __restrict__ Linv,
                                                  int np, int kb, long long* __restrict__ tl) {
    const int tid = threadIdx.x;
    const int k0 = kb * kNB;
    const int nblk = np / kNB;
    if (tid == 0) tl_stamp(tl, (0 * nblk + kb) * 2);
    for (int i = tid; i < kNB * kNB; i += kFusedThreads) {
        const int r = i / kNB, c = i % kNB;
        sm.W[r][c] = (c <= r) ? __ldcg(Hm + (size_t)(k0 + r) * np + k0 + c) : 0.0;
        if (kb > 0) sm.Bm[r][c] = __ldcg(Hm + (size_t)(k0 + r) * np + (k0 - kNB) + c);
    }
    if (tid == 0) sm.bad = 0;
    __syncthreads();
    if (kb > 0) {
        trsm32_rows8(sm.Bm, sm.L, sm.dinv, tid);
        __syncthreads();
        syrk32_lower(sm.W, sm.Bm, tid);
        __syncthreads();
    }
    chol32_block(sm.W, sm.L, sm.dinv, &sm.bad);
    if (sm.bad) {
        if (tid == 0) status->chol_fail = 1;
        return;
    }
    double* Ld = Ldiag + (size_t)kb * kNB * kNB;
    for (int i = tid; i < kNB * kNB; i += kFusedThreads) {
        const int r = i / kNB, c = i % kNB;
        const double l = sm.L[r][c];
        Ld[i] = l;
        if (c <= r) Lm[(size_t)(k0 + r) * np + k0 + c] = l;
    }
    // the last block has no following step whose spare worker could invert it
    if (kb == nblk - 1 && tid < 32) tri_inverse32(sm.L, sm.dinv, Linv + (size_t)kb * kNB * kNB, tid);
    if (tid == 0) tl_stamp(tl, (0 * nblk + kb) * 2 + 1);
}

// Control words of one launch (zeroed by a memset node before it):
//   ctl[0]  worker barrier counter (monotonic: W arrivals per step)
//   ctl[1]  chain_done   = number of diagonal blocks factorised and published by the chain CTA
//   ctl[2]  tile00_done  = 1 + the last step whose tile (0,0) is stored (the two blocks of the next-but-
//                          one diagonal block's input: A[kb+2, kb+1], A[kb+2, kb+2])
//   ctl[3]  fail_at      = 1 + index of the diagonal block whose factorisation failed (0 = none)
// Dependencies: chain iteration c (factorise block c) needs tile (0,0) of step c-2; workers' step s
// needs L_s (chain iteration s) and every tile of step s-1 (worker barrier).  The chain CTA therefore
// never waits for the bulk of a trailing update, only for one tile that the first worker does first.
__device__ __forceinline__ bool fused_wait_ge(const unsigned* p, unsigned v, const unsigned* fail_at,
                                              unsigned fail_le, int* s_flag) {
    // thread 0 polls; *s_flag: 1 = reached, 2 = a failed factorisation makes the wait moot, 0 = timeout
    if (threadIdx.x == 0) {
        int spins = 0, r = 0;
        for (;;) {
            if (ld_acquire_u32(p) >= v) { r = 1; break; }
            const unsigned f = ld_acquire_u32(fail_at);
            if (f != 0u && f <= fail_le) { r = 2; break; }
            if (++spins > kGridSpinLimit) break;
        }
        *s_flag = r;
    }
    __syncthreads();
    const int r = *s_flag;
    __syncthreads();
    return r == 1;
}

__device__ __forceinline__ void st_release_u32(unsigned* p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kFusedThreads, 2)
chol_fused_kernel(Status* __restrict__ status, double* Hm, double* __restrict__ Lm,
                  double* __restrict__ Ldiag, double* __restrict__ Linv, int np, unsigned* ctl,
                  long long* __restrict__ tl) {
    // uniform over the grid: both flags are only written by earlier kernels (or, on a timeout, long
    // after every CTA has passed this line)
    if (status->stopped | status->failed) return;
    __shared__ FusedSmem sm;
    __shared__ int s_flag;
    const int nblk = np / kNB;
    const int tid = threadIdx.x;
    if (blockIdx.x == 0) {
        // ---- chain ----
        for (int c = 0; c < nblk; ++c) {
            if (c >= 2) {
                if (!fused_wait_ge(ctl + 2, (unsigned)(c - 1), ctl + 3, 0u, &s_flag)) {
                    if (tid == 0) status->failed = 1;   // timeout: a worker is lost
                    return;
                }
            }
            fused_chain_block(sm.c, status, Hm, Lm, Ldiag, Linv, np, c, tl);
            __syncthreads();
            if (sm.c.bad) {
                if (tid == 0) { __threadfence(); st_release_u32(ctl + 3, (unsigned)(c + 1)); }
                return;
            }
            if (tid == 0) { __threadfence(); st_release_u32(ctl + 1, (unsigned)(c + 1)); }
        }
        return;
    }
    // ---- workers ----
    const int W = (int)gridDim.x - 1, w = (int)blockIdx.x - 1;
    for (int kb = 0; kb < nblk; ++kb) {
        // L_kb published?  (a failed factorisation of block <= kb ends every worker at this step)
        if (!fused_wait_ge(ctl + 1, (unsigned)(kb + 1), ctl + 3, (unsigned)(kb + 1), &s_flag)) {
            if (s_flag == 0 && tid == 0) status->failed = 1;
            return;
        }
        const int k0 = kb * kNB;
        const int rows_below = np + 1 - (k0 + kNB);  // incl. the rhs row
        const int trailing = np - (k0 + kNB);
        const int ntr = (rows_below + kSyrkTile - 1) / kSyrkTile;
        const int ntc0 = (trailing + kSyrkTile - 1) / kSyrkTile;
        const int ntc = ntc0 > 0 ? ntc0 : 1;   // trailing == 0 (last step): one tile, the rhs row
        const int ntri = ntc * (ntc + 1) / 2;
        const int ntiles = (ntc0 == 0) ? 1 : ntri + ((ntr > ntc0) ? ntc0 : 0);
        for (int t = w; t < ntiles; t += W) {
            int tr, tc;
            if (t >= ntri) {
                tr = ntc;
                tc = t - ntri;
            } else {
                tr = (int)((sqrtf(8.0f * (float)t + 1.0f) - 1.0f) * 0.5f);
                while (tr * (tr + 1) / 2 > t) --tr;
                while ((tr + 1) * (tr + 2) / 2 <= t) ++tr;
                tc = t - tr * (tr + 1) / 2;
            }
            chol_tile_body(sm.t, Hm, Lm, Ldiag, np, kb, tr, tc, tl);   // ends with __syncthreads
            if (t == 0 && tid == 0) { __threadfence(); st_release_u32(ctl + 2, (unsigned)(kb + 1)); }
        }
        // the least loaded worker inverts L_kb for the back-substitution, off every chain
        if (w == W - 1 && kb != nblk - 1) chol_invert_block(sm.t, Ldiag, Linv, kb);
        if (kb + 1 < nblk) {
            // worker barrier: every tile of step kb is stored before any tile of step kb + 1 starts
            __syncthreads();
            if (tid == 0) {
                __threadfence();
                atomicAdd(ctl, 1u);
                const unsigned target = (unsigned)(kb + 1) * (unsigned)W;
                int spins = 0, ok = 1;
                while (ld_acquire_u32(ctl) < target) {
                    if (++spins > kGridSpinLimit) { ok = 0; break; }
                }
                __threadfence();
                s_flag = ok;
            }
            __syncthreads();
            if (!s_flag) {
                // timeout, or the other workers left at a failed factorisation: leave as well
                if (tid == 0 && ld_acquire_u32(ctl + 3) == 0u) status->failed = 1;
                return;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Dataflow factorisation (default when every tile CTA is co-resident; M3S_GN_CHOL=steps|fused select the
// other paths): ONE cooperative launch in which every 64x64 tile of the lower triangle stays in the shared
// memory of the CTA that owns it for the whole factorisation; the only global traffic per block step
// is the 32-column panel and the only synchronisation are per-panel acquire/release flags (no grid
// barrier, no kernel boundary):
//   * 64x64 tiles (r, c), 32-column block steps k (tile column c holds steps 2c and 2c + 1).
//   * the factor is kept as M_k = G_kk^-1 with M_k A_kk M_k^T = I (ANY such M works: the panel is
//     P_rk = A_rk M_k^T, the trailing update A_rc -= P_rk P_ck^T, the right-hand-side row is one more
//     panel row, and the back-substitution needs G_kk^-T = M_k^T): the triangular solves of the step
//     kernels become 32^3 GEMMs and the Cholesky factor of the diagonal block is never formed.
//   * tl_elim32_w: M_k by a block LDL^T with 2x2 pivots on three warps (see there).
//   * roles: CTA d < nt owns the diagonal tile (d, d) and the 64 right-hand-side entries of tile column
//     d; every tile (r, c), r > c, has a CTA of its own.  The chain from one diagonal tile to the next:
//     M_{2d+1} -> [tile (d+1, d): panel = S M^T, publish] -> [CTA d+1: T -= P P^T, two eliminations].
//     The diagonal CTAs and the tiles (c + 1, c) take the blocks that are dispatched to SMs without a
//     second CTA (cooperative launch, 253 CTAs for 22 tile rows on 148 SMs, 2 CTAs per SM at most).
// Flags are monotonic words zeroed by a memset before the launch; a wait is one thread per flag polling
// with ld.acquire (bounded: a lost producer turns into status->failed and an abort word that releases
// every other wait); data written during the launch is read with ld.global.cg.
// MEASURED (B200, 200 keyframes = 44 block steps, scripts/gn_solve_timeline.py, profiles/r2_gn_timeline_*):
// factorisation 457 us (launch-per-step path: 495 us); per diagonal tile ~20.8 us = elimination 3.3 +
// [M stored, P10, T11 update, fence, publish] 2.9 + elimination 3.7 + publish 0.4 + [tile (d+1, d): hop,
// panel GEMM, store, fence, publish; CTA d+1: hop] 6 + panel load 1.2 + T update 2.3.  What the numbers
// say about this GPU (profiles/microbench/latency_b200.txt): dependent DFMA 8.7 clk but only ~53 DFMA
// lanes/clk/SM, L2 round trip ~300 clk, __threadfence ~410 clk, st.release.gpu ~790 clk -- a flag hop with
// its data costs ~1.5 us, and a 64x64x32 update ~1.3 us of a whole SM.
// ------------------------------------------------------------------------------------------------
constexpr int kTS = 64;          // tile size
constexpr int kTlThreads = 256;
constexpr int kTP = 66;          // tile pitch in doubles (16-B aligned rows, conflict-free column walks)
constexpr int kPP = 34;          // panel pitch in doubles

struct TilesSmem {
    double T[kTS][kTP];    // diagonal tile (d, d)                       (diagonal CTAs)
    double S[kTS][kTP];    // tile (d, d-1) / the off-diagonal CTA's tile
    double PA[kTS][kPP];   // panel rows of this tile row; scratch of elim32
    double PB[kTS][kPP];   // panel rows of the tile column
    double Mi[kNB][kPP];   // M_k
    double rhs[kTS];       // right-hand-side entries of tile column d
    double yk[kNB];
    double piv[16][4];     // elim32: (alpha, beta, det) of each 2x2 pivot
    int wres[2][4];        // results of the flag waits (double-buffered: one barrier per wait)
    int prog;              // elim32: pivots published by warp 0
    int hsflag;            // elim32: segments handed over by warp 2
    int bad;
};

// control words: [0] abort | lflag[2 nt] | p10flag[nt] | yflag[2 nt] | pflag[2 nt][nt]
__host__ __device__ inline int tl_ctl_words(int nt) { return 8 + 5 * nt + 2 * nt * nt; }
struct TilesCtl {
    unsigned* abort_w;
    unsigned* lflag;
    unsigned* p10flag;
    unsigned* yflag;
    unsigned* pflag;
    int nt;
    __device__ TilesCtl(unsigned* ctl, int nt_) : abort_w(ctl), lflag(ctl + 8), p10flag(ctl + 8 + 2 * nt_),
                                                   yflag(ctl + 8 + 3 * nt_), pflag(ctl + 8 + 5 * nt_), nt(nt_) {}
    __device__ unsigned* panel(int k, int r) const { return pflag + k * nt + r; }
};

// Wait for up to three flags (null = none): lane 0 of warps 0..2 poll one each.  One barrier per call.
__device__ __forceinline__ bool tl_wait(TilesSmem& sm, int& phase, const TilesCtl& ctl, const unsigned* f0,
                                        const unsigned* f1, const unsigned* f2) {
    const int tid = threadIdx.x;
    const int slot = phase & 1;
    phase ^= 1;
    if ((tid & 31) == 0 && tid < 96) {
        const unsigned* f = (tid == 0) ? f0 : (tid == 32 ? f1 : f2);
        int ok = 1;
        if (f != nullptr) {
            int spins = 0;
            while (ld_acquire_u32(f) == 0u) {
                ++spins;
                if (spins > kGridSpinLimit || ((spins & 63) == 0 && ld_acquire_u32(ctl.abort_w) != 0u)) { ok = 0; break; }
            }
            if (!ok) st_release_u32(ctl.abort_w, 1u);
        }
        sm.wres[slot][tid >> 5] = ok;
    }
    __syncthreads();
    return (sm.wres[slot][0] & sm.wres[slot][1] & sm.wres[slot][2]) != 0;
}

// after the data stores of all threads and a __syncthreads: one thread publishes
__device__ __forceinline__ void st_relaxed_u32(unsigned* p, unsigned v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void tl_publish(unsigned* flag) {
    __threadfence();
    st_relaxed_u32(flag, 1u);
}

// 64x64 tile of the assembled matrix (rows r0.., columns c0..); beyond np: identity
__device__ __forceinline__ void tl_load_tile(double (*dst)[kTP], const double* __restrict__ Hm, int np, int r0, int c0) {
    for (int i = threadIdx.x; i < kTS * kTS / 2; i += kTlThreads) {
        const int r = i >> 5, c = (i & 31) * 2;
        const int gr = r0 + r, gc = c0 + c;
        double2 v;
        if (gr < np && gc + 1 < np) v = *reinterpret_cast<const double2*>(Hm + (size_t)gr * np + gc);
        else {
            v.x = (gr < np && gc < np) ? Hm[(size_t)gr * np + gc] : ((gr == gc) ? 1.0 : 0.0);
            v.y = (gr < np && gc + 1 < np) ? Hm[(size_t)gr * np + gc + 1] : ((gr == gc + 1) ? 1.0 : 0.0);
        }
        *reinterpret_cast<double2*>(&dst[r][c]) = v;
    }
}

// `rows` panel rows (r0.., columns 32 k ..) of the factor written during this launch; rows >= np: 0
__device__ __forceinline__ void tl_load_panel(double (*dst)[kPP], const double* Lm, int np, int r0, int k, int rows) {
    for (int i = threadIdx.x; i < rows * (kNB / 2); i += kTlThreads) {
        const int r = i >> 4, c = (i & 15) * 2;
        double2 v = make_double2(0.0, 0.0);
        if (r0 + r < np) v = __ldcg(reinterpret_cast<const double2*>(Lm + (size_t)(r0 + r) * np + k * kNB + c));
        *reinterpret_cast<double2*>(&dst[r][c]) = v;
    }
}
__device__ __forceinline__ void tl_store_panel(const double (*src)[kPP], double* Lm, int np, int r0, int k, int rows) {
    for (int i = threadIdx.x; i < rows * (kNB / 2); i += kTlThreads) {
        const int r = i >> 4, c = (i & 15) * 2;
        if (r0 + r < np)
            *reinterpret_cast<double2*>(Lm + (size_t)(r0 + r) * np + k * kNB + c) = *reinterpret_cast<const double2*>(&src[r][c]);
    }
}
__device__ __forceinline__ void tl_load_block32(double (*dst)[kPP], const double* src) {  // Linv block, row-major 32x32
    for (int i = threadIdx.x; i < kNB * kNB / 2; i += kTlThreads) {
        const int r = i >> 4, c = (i & 15) * 2;
        *reinterpret_cast<double2*>(&dst[r][c]) = __ldcg(reinterpret_cast<const double2*>(src + r * kNB + c));
    }
}
__device__ __forceinline__ void tl_store_block32(const double (*src)[kPP], double* dst) {
    for (int i = threadIdx.x; i < kNB * kNB / 2; i += kTlThreads) {
        const int r = i >> 4, c = (i & 15) * 2;
        *reinterpret_cast<double2*>(dst + r * kNB + c) = *reinterpret_cast<const double2*>(&src[r][c]);
    }
}

// acc[i][j] = sum_k A[ty + 16 i][k] B[tx + 16 j][k], k < 32 (thread grid 16 x 16; lda, ldb even)
template <int MI, int NJ>
__device__ __forceinline__ void tl_gemm_nt(const double* A, int lda, const double* B, int ldb, double (&acc)[MI][NJ]) {
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) acc[i][j] = 0.0;
#pragma unroll 4
    for (int k = 0; k < kNB; k += 2) {
        double2 a[MI], b[NJ];
#pragma unroll
        for (int i = 0; i < MI; ++i) a[i] = *reinterpret_cast<const double2*>(A + (ty + 16 * i) * lda + k);
#pragma unroll
        for (int j = 0; j < NJ; ++j) b[j] = *reinterpret_cast<const double2*>(B + (tx + 16 * j) * ldb + k);
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) acc[i][j] = fma(a[i].y, b[j].y, fma(a[i].x, b[j].x, acc[i][j]));
    }
}
// the same for a SYMMETRIC product A A^T of which only the lower triangle is kept: register tiles with
// i < j lie above the diagonal (rows ty + 16 i < columns tx + 16 j) and are skipped -- 10 of 16 tiles
template <int MI>
__device__ __forceinline__ void tl_syrk_lower(const double* A, int lda, double (&acc)[MI][MI]) {
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < MI; ++j) acc[i][j] = 0.0;
#pragma unroll 4
    for (int k = 0; k < kNB; k += 2) {
        double2 a[MI], b[MI];
#pragma unroll
        for (int i = 0; i < MI; ++i) a[i] = *reinterpret_cast<const double2*>(A + (ty + 16 * i) * lda + k);
#pragma unroll
        for (int j = 0; j < MI; ++j) b[j] = *reinterpret_cast<const double2*>(A + (tx + 16 * j) * lda + k);
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j <= i; ++j) acc[i][j] = fma(a[i].y, b[j].y, fma(a[i].x, b[j].x, acc[i][j]));
    }
}
template <int MI>
__device__ __forceinline__ void tl_sub_lower(double* Cm, int ldc, const double (&acc)[MI][MI]) {
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) Cm[(ty + 16 * i) * ldc + tx + 16 * j] -= acc[i][j];
}
template <int MI, int NJ>
__device__ __forceinline__ void tl_sub(double* Cm, int ldc, const double (&acc)[MI][NJ]) {
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) Cm[(ty + 16 * i) * ldc + tx + 16 * j] -= acc[i][j];
}
template <int MI, int NJ>
__device__ __forceinline__ void tl_put(double* Cm, int ldc, const double (&acc)[MI][NJ]) {
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
#pragma unroll
    for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) Cm[(ty + 16 * i) * ldc + tx + 16 * j] = acc[i][j];
}

// out[i] (+)= sign * sum_m v[m] P[i][m], i < rows (4 threads per output)
__device__ __forceinline__ void tl_rowdot(const double (*P)[kPP], const double* v, double* out, int rows, bool subtract) {
    const int i = threadIdx.x >> 2, part = threadIdx.x & 3;
    double s = 0.0;
    if (i < rows) {
#pragma unroll
        for (int m = 0; m < 8; ++m) s = fma(v[part * 8 + m], P[i][part * 8 + m], s);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (i < rows && part == 0) out[i] = subtract ? out[i] - s : s;
}

// M (32 x 32, into Mo) with M W M^T = I for the symmetric block W = Tm[o .. o+31][o .. o+31]: block LDL^T
// with 2x2 pivots, W = L D L^T, M = D^-1/2 L^-1 (the dependent chain per PAIR of columns is
// det = a c - b^2 -> ONE reciprocal -> one FMA per updated entry, instead of shuffle -> rsqrt -> scale ->
// update per column).  Called by three warps (kElimWarp0 .. +2; no block barrier inside; sm.prog =
// sm.hsflag = 0 and a block barrier before the call, a block barrier after it):
//   chain   the dependent chain, 8 columns (one SEGMENT) in registers at a time; publishes the pivot
//           columns X, Y, the multipliers M1, M2 and the pivots of every pair through shared memory;
//   inverse N = L^-1 and M = D^-1/2 N from the published multipliers (lane = column, registers);
//   helper  owns the columns of the segments the chain has not reached yet, applies every published pair
//           to them and hands the next segment over through `hs` when the chain gets there.
// What shaped this code (clock stamps per pair via `dbg`, profiles/microbench/elim_bench.cu):
//   * with all 32 columns in the chain warp's registers ptxas scheduled ~60 trailing DFMAs in front of
//     each determinant (in-order issue: 450 clk per pair) -> trailing update on the helper warp;
//   * run by warps 0-2 of the factorisation kernel -- the warps whose lane 0 polls the flags, writes the
//     time stamps and does the other single-thread jobs -- every warp-synchronising instruction of the
//     chain (SHFL.IDX, then __syncwarp) took 200-700 clk for the first pairs of each call: 17 400 clk
//     per elimination against 6 200 on warps 4-6 (same code, same data; stand-alone: 6 200 on any warp).
//     Hence kElimWarp0 = 4 and shared-memory broadcasts instead of shuffles;
//   * GEMM work of other warps of the same CTA beside the elimination (tried: 128 workers doing the
//     T10 / T11 updates meanwhile) slows the chain 5x through the shared-memory pipeline -> the
//     elimination runs alone on its SM.
// sc = 4 x 16 x 32 doubles (X, Y, M1, M2), hs = 8 x 32 doubles.  ~6 500 clk = 3.4 us per call.
#ifdef M3S_TL_SPIN_SLEEP
#define TL_SPIN_PAUSE __nanosleep(M3S_TL_SPIN_SLEEP);
#else
#define TL_SPIN_PAUSE
#endif
#if defined(M3S_TL_NOFENCE)
#define TL_FENCE_BLOCK
#elif defined(M3S_TL_ACQREL)
#define TL_FENCE_BLOCK asm volatile("fence.acq_rel.cta;" ::: "memory");
#else
#define TL_FENCE_BLOCK __threadfence_block();
#endif
// the three elimination warps: NOT warp 0 (thread 0 polls flags, writes time stamps and carries other
// single-thread duties; see the measurements in DESIGN.md)
#ifndef M3S_TL_ELIM_WARP0
#define M3S_TL_ELIM_WARP0 4
#endif
constexpr int kElimWarp0 = M3S_TL_ELIM_WARP0;
__device__ __forceinline__ bool tl_is_elim_warp() {
    const int w = (int)(threadIdx.x >> 5) - kElimWarp0;
    return w >= 0 && w <= 2;
}
__device__ __noinline__ void tl_elim32_w(TilesSmem& sm, double (*Tm)[kTP], int o, double* sc, double* hs,
                                         double (*Mo)[kPP], long long* dbg) {
    const int lane = threadIdx.x & 31, warp = (int)(threadIdx.x >> 5) - kElimWarp0;
    double* X = sc;
    double* Y = sc + 512;
    double* M1 = sc + 1024;
    double* M2 = sc + 1536;
    volatile int* prog = &sm.prog;
    volatile int* hsf = &sm.hsflag;
    if (dbg != nullptr && warp == 0 && lane == 0) dbg[48] = clock64();
    if (warp == 0) {
        double a[8];   // the 8 columns of the current segment
        bool isbad = false;
#pragma unroll
        for (int c = 0; c < 8; ++c) a[c] = Tm[o + lane][o + c];   // row `lane`; only entries c <= lane (lower triangle) are ever used
#pragma unroll 1
        for (int s = 0; s < 4; ++s) {
            if (s > 0) {
                while (*hsf < s) { TL_SPIN_PAUSE }
                __syncwarp();
#pragma unroll
                for (int c = 0; c < 8; ++c) a[c] = hs[c * 32 + lane];
            }
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                const int q = 4 * s + p, j = 2 * q;
                // The pivot block and the pivot-column entries of the rows that follow go through SHARED
                // MEMORY (store own x, y -> __syncwarp -> broadcast loads), not through shuffles (30 SHFL per
                // pair cost more issue slots and, measured inside the factorisation kernel, far more time).
                const double x = a[2 * p], y = a[2 * p + 1];
                X[q * 32 + lane] = x;
                Y[q * 32 + lane] = y;
                __syncwarp();
                double al = X[q * 32 + j];
                double be = X[q * 32 + j + 1];
                double ga = Y[q * 32 + j + 1];
                double xc[8], yc[8];
#pragma unroll
                for (int c = 2 * p + 2; c < 8; c += 2) {
                    const double2 xv = *reinterpret_cast<const double2*>(X + q * 32 + 8 * s + c);
                    const double2 yv = *reinterpret_cast<const double2*>(Y + q * 32 + 8 * s + c);
                    xc[c] = xv.x; xc[c + 1] = xv.y;
                    yc[c] = yv.x; yc[c + 1] = yv.y;
                }
                double det = fma(al, ga, -(be * be));
                if (!(al > 1e-290) || !(det > 1e-300) || !(det < 1e300)) {  // warp-uniform
                    isbad = true;
                    al = 1.0; be = 0.0; ga = 1.0; det = 1.0;
                }
                const double r = rcp_fast(det);
                const double n1 = fma(x, ga, -(y * be)), n2 = fma(y, al, -(x * be));   // independent of r
#pragma unroll
                for (int c = 2 * p + 2; c < 8; ++c) {
                    const double z = fma(n1, xc[c], n2 * yc[c]);                        // independent of r
                    a[c] = fma(-r, z, a[c]);                                            // ONE op after the reciprocal
                }
                M1[q * 32 + lane] = n1 * r;
                M2[q * 32 + lane] = n2 * r;
                if (lane == 0) { sm.piv[q][0] = al; sm.piv[q][1] = be; sm.piv[q][2] = det; }
                __syncwarp();
                if (lane == 0) { TL_FENCE_BLOCK *prog = q + 1; if (dbg != nullptr) dbg[q] = clock64(); }
            }
        }
        if (isbad && lane == 0) sm.bad = 1;
    } else if (warp == 1) {
        // lane = column c of N in registers (loop over the pairs fully unrolled: static register indices)
        double n[kNB];
#pragma unroll
        for (int i = 0; i < kNB; ++i) n[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
            const int j = 2 * q;
            while (*prog < q + 1) { TL_SPIN_PAUSE }
            __syncwarp();
            const double nj = n[j], nj1 = n[j + 1];
            const double2* m1v = reinterpret_cast<const double2*>(M1 + q * 32);
            const double2* m2v = reinterpret_cast<const double2*>(M2 + q * 32);
#pragma unroll
            for (int i = j + 2; i < kNB; i += 2) {
                const double2 u = m1v[i >> 1], v = m2v[i >> 1];
                n[i] = fma(-v.x, nj1, fma(-u.x, nj, n[i]));
                n[i + 1] = fma(-v.y, nj1, fma(-u.y, nj, n[i + 1]));
            }
            const double al = sm.piv[q][0], be = sm.piv[q][1], det = sm.piv[q][2];
            // F with F D F^T = I for D = [[al, be], [be, ga]]: F = [[1/sqrt(al), 0], [-be / (al s), 1/s]], s^2 = det / al
            const double ra = rsqrt_fast(al);
            const double ia = ra * ra;
            const double rs = rsqrt_fast(det * ia);
            const double f10 = -(be * ia) * rs;
            Mo[j][lane] = ra * nj;
            Mo[j + 1][lane] = fma(f10, nj, rs * nj1);
            if (dbg != nullptr && lane == 0) dbg[16 + q] = clock64();
        }
    } else if (warp == 2) {
        double h[24];   // columns 8 .. 31 of row `lane`
#pragma unroll
        for (int c = 0; c < 24; ++c) h[c] = Tm[o + lane][o + 8 + c];
        // loop over the pairs fully unrolled (static register indices); in the pair that ends a segment the
        // 8 columns handed over next are updated and published first, the rest afterwards.  (Handing a
        // segment over two pairs early, with the chain applying the last two pairs itself, was measured
        // slower: 7 800 vs 6 200 clk stand-alone -- the helper's pace, not the hop, bounds the hand-over.)
#pragma unroll
        for (int q = 0; q < 12; ++q) {
            while (*prog < q + 1) { TL_SPIN_PAUSE }
            __syncwarp();
            const double m1 = M1[q * 32 + lane], m2 = M2[q * 32 + lane];
            const double2* Xq = reinterpret_cast<const double2*>(X + q * 32);
            const double2* Yq = reinterpret_cast<const double2*>(Y + q * 32);
            const int c0 = 8 * (q / 4 + 1);   // first column still owned here
#pragma unroll
            for (int c = c0; c < c0 + 8; c += 2) {
                const double2 xv = Xq[c >> 1], yv = Yq[c >> 1];
                h[c - 8] = fma(-m2, yv.x, fma(-m1, xv.x, h[c - 8]));
                h[c - 7] = fma(-m2, yv.y, fma(-m1, xv.y, h[c - 7]));
            }
            if ((q & 3) == 3) {
#pragma unroll
                for (int c = 0; c < 8; ++c) hs[c * 32 + lane] = h[c0 - 8 + c];
                __syncwarp();
                if (lane == 0) { TL_FENCE_BLOCK *hsf = q / 4 + 1; }
            }
#pragma unroll
            for (int c = c0 + 8; c < kNB; c += 2) {
                const double2 xv = Xq[c >> 1], yv = Yq[c >> 1];
                h[c - 8] = fma(-m2, yv.x, fma(-m1, xv.x, h[c - 8]));
                h[c - 7] = fma(-m2, yv.y, fma(-m1, xv.y, h[c - 7]));
            }
            if (dbg != nullptr && lane == 0) dbg[32 + q] = clock64();
        }
    }
}

__global__ void __launch_bounds__(kTlThreads, 2)
chol_tiles_kernel(Status* __restrict__ status, const double* __restrict__ Hm, double* Lm, double* Linv, int np,
                  unsigned* ctlw, int first_diag, long long* __restrict__ tl) {
    if (status->stopped | status->failed) return;   // uniform over the grid (written by earlier kernels only)
    extern __shared__ __align__(16) unsigned char tl_smem_raw[];
    TilesSmem& sm = *reinterpret_cast<TilesSmem*>(tl_smem_raw);
    const int tid = threadIdx.x;
    const int nblk = np / kNB;
    const int nt = (np + kTS - 1) / kTS;
    const TilesCtl ctl(ctlw, nt);
    int phase = 0;
    if (tid == 0) sm.bad = 0;
    const double* ybase = Lm + (size_t)np * np;   // row np of the factor: y
    double* ystore = Lm + (size_t)np * np;

    // Role of this CTA.  CTAs are dispatched breadth-first (block b < #SMs -> SM b, the rest fill the
    // second slot of SMs 0, 1, ...): the nt diagonal CTAs -- the dependent chain -- take the LAST nt
    // blocks of the first wave (first_diag = min(#SMs, grid) - nt), whose SMs get no second CTA as long as
    // grid <= 2 #SMs - nt, so no neighbour competes for their fp64 pipe.  Correct for any dispatch order.
    const int bid = (int)blockIdx.x;
    const bool is_diag = bid >= first_diag && bid < first_diag + nt;
    if (tid == 0 && tl != nullptr && bid < 232) {   // placement diagnostics (scripts/gn_solve_timeline.py)
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        tl[16 * nblk + 120 + bid] = (long long)smid + 1;
    }
    if (is_diag) {
        // ------------------------------ diagonal CTA d ------------------------------------------
        const int d = bid - first_diag;
        if (tid == 0 && tl != nullptr) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            tl[16 * nblk + d] = (long long)smid;   // who = 1 phase slots are unused by this kernel
        }
        const int r0 = d * kTS;
        tl_load_tile(sm.T, Hm, np, r0, r0);
        if (tid < kTS) sm.rhs[tid] = (r0 + tid < np) ? Hm[(size_t)np * np + r0 + tid] : 0.0;
        __syncthreads();
        // Scratch of the eliminations: S rows 0 .. 31 (this CTA has no second tile) = elim32's X / Y / M1 /
        // M2, PB rows 0 .. 31 = the segment hand-over buffer, PB rows 32 .. 63 = a second M buffer.
        double* esc = &sm.S[0][0];
        double* ehs = &sm.PB[0][0];
        double (*MiB)[kPP] = reinterpret_cast<double (*)[kPP]>(&sm.PB[kNB][0]);
        // block steps 0 .. 2d-1: the panels of tile row d (the last two come from tile (d, d-1), one flag hop
        // after M_{2d-2} / M_{2d-1}: that hop + this update is what separates two diagonal tiles on the chain)
        for (int k = 0; k < 2 * d; ++k) {
            const bool last = (k == 2 * d - 1);
            if (last && tid == 0) tl_stamp(tl, 8 * nblk + 8 * k + 3);
            if (!tl_wait(sm, phase, ctl, ctl.panel(k, d), ctl.yflag + k, nullptr)) goto lost;
            if (last && tid == 0) tl_stamp(tl, 8 * nblk + 8 * k + 4);
            tl_load_panel(sm.PA, Lm, np, r0, k, kTS);
            if (tid < kNB) sm.yk[tid] = __ldcg(ybase + k * kNB + tid);
            __syncthreads();
            if (last && tid == 0) tl_stamp(tl, 8 * nblk + 8 * k + 5);
            {   // only the lower triangle of the diagonal tile is kept up to date (and read)
                double acc[4][4];
                tl_syrk_lower<4>(&sm.PA[0][0], kPP, acc);
                tl_sub_lower<4>(&sm.T[0][0], kTP, acc);
            }
            tl_rowdot(sm.PA, sm.yk, sm.rhs, kTS, true);
            __syncthreads();
            if (last && tid == 0) tl_stamp(tl, 8 * nblk + 8 * k + 6);
        }
        // ---- block step 2d: first elimination
        if (tid == 0) { sm.prog = 0; sm.hsflag = 0; tl_stamp(tl, (0 * nblk + 2 * d) * 2); }
        __syncthreads();
        if (tl_is_elim_warp()) tl_elim32_w(sm, sm.T, 0, esc, ehs, MiB, (tl != nullptr && d == 3) ? tl + 16 * nblk + 64 : nullptr);
        __syncthreads();
        // ---- block step 2d: M_{2d} is in MiB
        {
            const int k = 2 * d;
            if (tid == 0) tl_stamp(tl, 8 * nblk + 8 * k + 0);
            tl_store_block32(MiB, Linv + (size_t)k * kNB * kNB);
            tl_rowdot(MiB, sm.rhs, sm.yk, kNB, false);                     // y_k = M_k rhs_k
            const bool second = k + 1 < nblk;
            if (second) {
                double acc[2][2];
                tl_gemm_nt<2, 2>(&sm.T[kNB][0], kTP, &MiB[0][0], kPP, acc);   // P10 = T10 M^T
                tl_put<2, 2>(&sm.PA[0][0], kPP, acc);
            }
            __syncthreads();
            if (tid < kNB) ystore[k * kNB + tid] = sm.yk[tid];
            if (second) {
                tl_store_panel(sm.PA, Lm, np, r0 + kNB, k, kNB);
                double acc[2][2];
                tl_syrk_lower<2>(&sm.PA[0][0], kPP, acc);
                tl_sub_lower<2>(&sm.T[kNB][kNB], kTP, acc);
                tl_rowdot(sm.PA, sm.yk, sm.rhs + kNB, kNB, true);
            }
            if (tid == 0) { sm.prog = 0; sm.hsflag = 0; }
            __syncthreads();
            if (tid == 96) {   // ONE fence for everything stored since the last publish, finished BEFORE the chain goes on
                __threadfence();
                st_relaxed_u32(ctl.lflag + k, 1u);
                st_relaxed_u32(ctl.yflag + k, 1u);
                st_relaxed_u32(ctl.p10flag + d, 1u);
            }
#ifdef M3S_TL_SYNC_PUBLISH   // A/B: finish the fence before the second elimination starts
            __syncthreads();
#endif
            if (tid == 0) tl_stamp(tl, (0 * nblk + k) * 2 + 1);
            if (second) {
                // ---- block step 2d+1
                if (tid == 0) tl_stamp(tl, (0 * nblk + k + 1) * 2);
                if (tl_is_elim_warp()) tl_elim32_w(sm, sm.T, kNB, esc, ehs, sm.Mi, nullptr);
                __syncthreads();
                if (tid == 0) tl_stamp(tl, 8 * nblk + 8 * (k + 1) + 0);
                tl_store_block32(sm.Mi, Linv + (size_t)(k + 1) * kNB * kNB);
                tl_rowdot(sm.Mi, sm.rhs + kNB, sm.yk, kNB, false);
                __syncthreads();
                if (tid < kNB) ystore[(k + 1) * kNB + tid] = sm.yk[tid];
                __syncthreads();
                if (tid == 0) tl_stamp(tl, (0 * nblk + k + 1) * 2 + 1);
                if (tid == 64) {
                    __threadfence();
                    st_relaxed_u32(ctl.lflag + k + 1, 1u);
                    st_relaxed_u32(ctl.yflag + k + 1, 1u);
                }
            }
        }
        if (sm.bad && tid == 0) status->chol_fail = 1;
        return;
    }
    {
        // ------------------------------ off-diagonal CTA (r, c), r > c ---------------------------
        int t = bid < first_diag ? bid : bid - nt, c = 0;
        // Enumeration: the R = min(nt - 1, first_diag) blocks right before the diagonal CTAs take the tiles
        // (c + 1, c) -- each one flag hop inside the chain -- so that they, too, land on SMs without a
        // second CTA; every other block takes the remaining tiles in column order.
        int r;
        {
            const int R = min(nt - 1, first_diag), sub0 = first_diag - R;
            if (t >= sub0 && t < first_diag) {
                c = t - sub0;
                r = c + 1;
            } else {
                int u = t < sub0 ? t : t - R;
                if (u < nt - 1 - R) {          // sub-diagonal tiles that found no reserved block
                    c = R + u;
                    r = c + 1;
                } else {
                    u -= nt - 1 - R;
                    c = 0;
                    while (u >= nt - 2 - c) { u -= nt - 2 - c; ++c; }
                    r = c + 2 + u;
                }
            }
        }
        const int r0 = r * kTS, c0 = c * kTS;
        tl_load_tile(sm.S, Hm, np, r0, c0);
        __syncthreads();
        for (int k = 0; k < 2 * c; ++k) {
            if (!tl_wait(sm, phase, ctl, ctl.panel(k, r), ctl.panel(k, c), nullptr)) goto lost;
            tl_load_panel(sm.PA, Lm, np, r0, k, kTS);
            tl_load_panel(sm.PB, Lm, np, c0, k, kTS);
            __syncthreads();
            double acc[4][4];
            tl_gemm_nt<4, 4>(&sm.PA[0][0], kPP, &sm.PB[0][0], kPP, acc);
            tl_sub<4, 4>(&sm.S[0][0], kTP, acc);
            __syncthreads();
        }
        for (int h = 0; h < 2; ++h) {
            const int k = 2 * c + h;
            if (!tl_wait(sm, phase, ctl, ctl.lflag + k, h == 0 ? ctl.p10flag + c : nullptr, nullptr)) goto lost;
            tl_load_block32(sm.Mi, Linv + (size_t)k * kNB * kNB);
            if (h == 0) tl_load_panel(sm.PB, Lm, np, c0 + kNB, k, kNB);
            __syncthreads();
            {
                double acc[4][2];
                tl_gemm_nt<4, 2>(&sm.S[0][kNB * h], kTP, &sm.Mi[0][0], kPP, acc);
                tl_put<4, 2>(&sm.PA[0][0], kPP, acc);
            }
            __syncthreads();
            tl_store_panel(sm.PA, Lm, np, r0, k, kTS);
            if (h == 0) {
                double acc[4][2];
                tl_gemm_nt<4, 2>(&sm.PA[0][0], kPP, &sm.PB[0][0], kPP, acc);
                tl_sub<4, 2>(&sm.S[0][kNB], kTP, acc);
            }
            __syncthreads();
            if (tid == 64) tl_publish(ctl.panel(k, r));
        }
        return;
    }
lost:
    // a producer did not publish within the bound (or another CTA aborted): never hang the device
    if (tid == 0) status->failed = 1;
}

// ------------------------------------------------------------------------------------------------
// Back-substitution L^T dx = y (y = rhs row of the factorised augmented matrix), ||dx|| test
// (gauss_newton.py:263-265: stop BEFORE applying) and Sim3 retraction (sim3_ops.py:129-251) in one
// kernel.
// ------------------------------------------------------------------------------------------------
__device__ void sim3_retract_dev(const double* xi, double* T) {
    const double EPS = 1e-6;
    const double* tau = xi;
    const double* om = xi + 3;
    const double sigma = xi[6];
    const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
    const double th = sqrt(th2 + EPS);  // biased on purpose, as the reference
    const bool small_t = th2 < EPS, small_s = fabs(sigma) < EPS;
    const double imag = small_t ? 0.5 - th2 / 48.0 : sin(0.5 * th) / th;
    const double real = small_t ? 1.0 - th2 / 8.0 : cos(0.5 * th);
    const double dq[4] = {imag * om[0], imag * om[1], imag * om[2], real};
    const double s = exp(sigma);
    const double C = small_s ? 1.0 : (s - 1.0) / sigma;
    double A, Bc;
    if (small_s) {
        A = small_t ? 0.5 : (1.0 - cos(th)) / th2;
        Bc = small_t ? 1.0 / 6.0 : (th - sin(th)) / (th2 * th);
    } else if (small_t) {
        A = ((sigma - 1.0) * s + 1.0) / (sigma * sigma);
        Bc = (s * 0.5 * sigma * sigma + s - 1.0 - sigma * s) / (sigma * sigma * sigma);
    } else {
        A = (s * sin(th) * sigma + (1.0 - s * cos(th)) * th) / (th * (th2 + sigma * sigma));
        Bc = (C - ((s * cos(th) - 1.0) * sigma + s * sin(th) * th) / (th2 + sigma * sigma)) / th2;
    }
    const double c1[3] = {om[1] * tau[2] - om[2] * tau[1], om[2] * tau[0] - om[0] * tau[2],
                          om[0] * tau[1] - om[1] * tau[0]};
    const double c2[3] = {om[1] * c1[2] - om[2] * c1[1], om[2] * c1[0] - om[0] * c1[2],
                          om[0] * c1[1] - om[1] * c1[0]};
    double qn[4], R[9];
    quat_mul(dq, T + 3, qn);
    quat_to_mat(dq, R);
    const double t0 = T[0], t1 = T[1], t2 = T[2];
    T[0] = (R[0] * t0 + R[1] * t1 + R[2] * t2) * s + (C * tau[0] + A * c1[0] + Bc * c2[0]);
    T[1] = (R[3] * t0 + R[4] * t1 + R[5] * t2) * s + (C * tau[1] + A * c1[1] + Bc * c2[1]);
    T[2] = (R[6] * t0 + R[7] * t1 + R[8] * t2) * s + (C * tau[2] + A * c1[2] + Bc * c2[2]);
    T[3] = qn[0]; T[4] = qn[1]; T[5] = qn[2]; T[6] = qn[3];
    T[7] = s * T[7];
}

// Metal-semantics retraction (params.semantics == M3S_SEM_METAL): pose_update_kernel of
// shaders/gauss_newton.metal:255-301 -- everything in fp32, the step is cast to fp32 first
// (gn_metal_runner.py:303), and the translation is t' = R(dq) t ds + tau: dt = tau (:286), without
// the Sim3 V-matrix coefficients of the numpy path.  The pose buffer of that path is fp32, so the
// fp64 pose state is rounded to fp32 on every update.
__device__ void sim3_retract_metal_dev(const double* xi, double* T) {
    const float EPS = 1e-6f;
    const float tau[3] = {(float)xi[0], (float)xi[1], (float)xi[2]};
    const float om[3] = {(float)xi[3], (float)xi[4], (float)xi[5]};
    const float sigma = (float)xi[6];
    const float t[3] = {(float)T[0], (float)T[1], (float)T[2]};
    const float q[4] = {(float)T[3], (float)T[4], (float)T[5], (float)T[6]};
    const float s = (float)T[7];
    const float th2 = (om[0] * om[0] + om[1] * om[1]) + om[2] * om[2];
    const float th = sqrtf(th2 + EPS);
    const float half = 0.5f * th;
    const float imag = th2 < EPS ? 0.5f - th2 / 48.0f : sinf(half) / th;
    const float real = th2 < EPS ? 1.0f - th2 / 8.0f : cosf(half);
    const float dq[4] = {imag * om[0], imag * om[1], imag * om[2], real};
    const float ds = expf(sigma);
    // quat_mul(dq, q), gauss_newton.metal:17-24
    const float qn[4] = {dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1],
                         dq[3] * q[1] - dq[0] * q[2] + dq[1] * q[3] + dq[2] * q[0],
                         dq[3] * q[2] + dq[0] * q[1] - dq[1] * q[0] + dq[2] * q[3],
                         dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2]};
    // quat_rotate(dq, t) = t + w uv + qv x uv, uv = 2 qv x t   (gauss_newton.metal:32-36)
    const float uv[3] = {2.0f * (dq[1] * t[2] - dq[2] * t[1]), 2.0f * (dq[2] * t[0] - dq[0] * t[2]),
                         2.0f * (dq[0] * t[1] - dq[1] * t[0])};
    const float c2[3] = {dq[1] * uv[2] - dq[2] * uv[1], dq[2] * uv[0] - dq[0] * uv[2],
                         dq[0] * uv[1] - dq[1] * uv[0]};
#pragma unroll
    for (int k = 0; k < 3; ++k) T[k] = (double)((t[k] + dq[3] * uv[k] + c2[k]) * ds + tau[k]);
    T[3] = (double)qn[0]; T[4] = (double)qn[1]; T[5] = (double)qn[2]; T[6] = (double)qn[3];
    T[7] = (double)(ds * s);
}

// Dataflow back-substitution.  One CTA per 32-column block; the CTA with blockIdx i owns block
// b = nblk-1-i, so every CTA waits only on CTAs with a smaller blockIdx (dispatched before it: no
// deadlock even if the grid were not fully resident).  CTA b accumulates
//     s_b = sum_{r > b} L[r, b]^T x_r
// block by block as the x_r are published (r descending): the 32x32 tile of L for the next r is
// already in registers when its flag arrives, so a step of the critical path is flag -> 256-B read of
// x_r -> 4 FMAs per thread -> (for r = b+1) shared-memory reduction -> x_b = Linv_b^T (y_b - s_b)
// -> publish.  Flags carry the launch epoch (no reset pass).  The CTA of block 0 finishes last by
// construction; it then does the ||dx|| test BEFORE applying (gauss_newton.py:263-265) and the Sim3
// retraction (sim3_ops.py:129-251).
constexpr int kBackThreads = 256;
constexpr int kSpinLimit = 1 << 24;

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(kBackThreads)
gn_backsolve_update_kernel(Status* __restrict__ status, const double* __restrict__ Hm,
                           const double* __restrict__ Linv, int np, int dim, int num_free,
                           const int* __restrict__ free_kf, double delta_thresh,
                           double* __restrict__ xsol, int* __restrict__ flags, int epoch,
                           double* __restrict__ dx_out, double* __restrict__ poses,
                           long long* __restrict__ tl, int metal) {
    if (status->stopped | status->failed) return;
    const int nblk = np / kNB;
    const int b = nblk - 1 - (int)blockIdx.x;
    const int tid = threadIdx.x;
    if (tid == 0) tl_stamp(tl, (3 * nblk + b) * 2);
    if (status->chol_fail) {
        // every CTA sees the same flag (nobody clears it here); the last one records the failure
        if (b == 0 && tid == 0) { status->failed = 1; status->attempted += 1; }
        return;
    }
    __shared__ double Li[kNB][kNB + 1];               // Linv_b
    __shared__ double xr[kNB];                         // x_r of the block being applied
    __shared__ double part[kBackThreads / 32][kNB];    // per-row-group partial sums
    __shared__ double rhs[kNB];
    __shared__ int ok_s;
    const int c = tid & 31, g = tid >> 5;              // column c of the block, row group g (4 rows)
    for (int i = tid; i < kNB * kNB; i += kBackThreads) Li[i / kNB][i % kNB] = Linv[(size_t)b * kNB * kNB + i];
    double acc = 0.0;
    // tile of L for block row r, column block b: thread (c, g) holds rows 4g .. 4g+3 of column c
    double t[4];
    int r = nblk - 1;
    if (r > b) {
#pragma unroll
        for (int i = 0; i < 4; ++i) t[i] = Hm[(size_t)(r * kNB + 4 * g + i) * np + b * kNB + c];
    }
    for (; r > b; --r) {
        if (tid == 0) {
            int spins = 0, okv = 1;
            while (ld_acquire(flags + r) != epoch) {
                if (++spins > kSpinLimit) { okv = 0; break; }
                __nanosleep(20);
            }
            ok_s = okv;
        }
        __syncthreads();
        if (!ok_s) {  // cannot happen with in-order dispatch; never hang the device
            if (tid == 0) status->failed = 1;
            return;
        }
        if (tid < kNB) xr[tid] = __ldcg(xsol + r * kNB + tid);
        __syncthreads();
        double cur[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) cur[i] = t[i];
        if (r - 1 > b) {  // prefetch the next tile before using this one
#pragma unroll
            for (int i = 0; i < 4; ++i) t[i] = Hm[(size_t)((r - 1) * kNB + 4 * g + i) * np + b * kNB + c];
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) acc += cur[i] * xr[4 * g + i];
        __syncthreads();  // xr is rewritten in the next round
    }
    part[g][c] = acc;
    __syncthreads();
    if (tid < kNB) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kBackThreads / 32; ++w) s += part[w][tid];
        rhs[tid] = Hm[(size_t)np * np + b * kNB + tid] - s;  // y_b - s_b
    }
    __syncthreads();
    if (tid < kNB) {
        // x_b[i] = sum_r Linv[r][i] rhs[r]   (Linv lower triangular: r >= i)
        double s = 0.0;
#pragma unroll
        for (int rr = 0; rr < kNB; ++rr) s += Li[rr][tid] * rhs[rr];
        __stcg(xsol + b * kNB + tid, s);
        __threadfence();
    }
    __syncthreads();
    if (tid == 0) { st_release(flags + b, epoch); tl_stamp(tl, (3 * nblk + b) * 2 + 1); }
    if (b != 0) return;

    // ---- block 0: all of x is published (x_0 depended on every other block) ----------------------
    __shared__ double red[kBackThreads / 32];
    __shared__ int decision;  // 0 apply, 1 stop, 2 fail
    double psum = 0.0;
    int nonfinite = 0;
    for (int i = tid; i < dim; i += kBackThreads) {
        const double v = __ldcg(xsol + i);
        dx_out[i] = v;
        psum += v * v;
        if (!(fabs(v) <= 1.7e308)) nonfinite = 1;
    }
    psum = warp_sum(psum);
    nonfinite = __any_sync(0xffffffffu, nonfinite);
    if ((tid & 31) == 0) red[tid >> 5] = nonfinite ? __longlong_as_double(0x7ff8000000000000LL) : psum;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < kBackThreads / 32; ++w) s += red[w];
        const double nrm = sqrt(s);
        status->last_norm = nrm;
        status->attempted += 1;
        int d;
        if (!(nrm == nrm)) d = 2;                 // non-finite step: treat as failed solve
        else if (nrm < delta_thresh) d = 1;       // converged: do not apply
        else d = 0;
        if (d == 1) status->stopped = 1;
        if (d == 2) status->failed = 1;
        if (d == 0) status->applied += 1;
        decision = d;
    }
    __syncthreads();
    if (decision != 0) return;
    for (int f = tid; f < num_free; f += kBackThreads) {
        double xi[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) xi[k] = __ldcg(xsol + 7 * f + k);
        if (metal) sim3_retract_metal_dev(xi, poses + 8 * free_kf[f]);
        else sim3_retract_dev(xi, poses + 8 * free_kf[f]);
    }
}

// Grouped variant (default): one CTA per kBackGroup = 4 consecutive 32-column blocks, so the chain has one
// flag hop per 128 columns instead of one per 32 (a hop = fence + flag + remote poll + 256-B read, ~1.5 us;
// the arithmetic between two hops is ~0.3 us).  CTA gi owns blocks b_hi = nblk-1-4 gi down to b_lo; it
// accumulates s_b = sum_{r > b_hi} L[r, b]^T x_r for its blocks as the groups above publish, then walks its
// own blocks top-down inside shared memory.  Same flags / epoch / decision / retraction as the kernel above.
constexpr int kBackGroup = 4;

__global__ void __launch_bounds__(kBackThreads)
gn_backsolve_group_kernel(Status* __restrict__ status, const double* __restrict__ Lm,
                          const double* __restrict__ Linv, int np, int dim, int num_free,
                          const int* __restrict__ free_kf, double delta_thresh,
                          double* __restrict__ xsol, int* __restrict__ flags, int epoch,
                          double* __restrict__ dx_out, double* __restrict__ poses,
                          long long* __restrict__ tl, int metal, int finish) {
    if (status->stopped | status->failed) return;
    const int nblk = np / kNB;
    const int gi = blockIdx.x;
    const int ngroups = (nblk + kBackGroup - 1) / kBackGroup;
    const int b_hi = nblk - 1 - kBackGroup * gi;
    const int b_lo = max(b_hi - kBackGroup + 1, 0);
    const int nb = b_hi - b_lo + 1;                  // blocks of this group (the last group may be short)
    const int tid = threadIdx.x;
    if (tid == 0) tl_stamp(tl, (3 * nblk + b_hi) * 2);
    if (status->chol_fail) {
        if (gi == ngroups - 1 && tid == 0) { status->failed = 1; status->attempted += 1; }
        return;
    }
    __shared__ double Li[kBackGroup][kNB][kNB + 1];        // M of the group's blocks (index 0 = b_hi)
    __shared__ double xr[kBackGroup][kNB];                 // x of the group being applied / of this group
    __shared__ double part[kBackThreads / 32][kNB];
    __shared__ double rhs[kNB];
    __shared__ int ok_s;
    const int c = tid & 31, g = tid >> 5;                  // column c of a block, row group g (4 rows)
    for (int i = tid; i < nb * kNB * kNB; i += kBackThreads) {
        const int q = i / (kNB * kNB), e = i % (kNB * kNB);
        Li[q][e / kNB][e % kNB] = Linv[(size_t)(b_hi - q) * kNB * kNB + e];
    }
    double acc[kBackGroup];
#pragma unroll
    for (int q = 0; q < kBackGroup; ++q) acc[q] = 0.0;
    // contributions of the groups above, as they are published
    for (int gr = 0; gr < gi; ++gr) {
        const int r_hi = nblk - 1 - kBackGroup * gr;
        // tiles L[r, b] of this group's columns for the 4 blocks r of group gr: rows 4g .. 4g+3 of column c
        double t[kBackGroup][kBackGroup][4];
#pragma unroll
        for (int rr = 0; rr < kBackGroup; ++rr)
#pragma unroll
            for (int q = 0; q < kBackGroup; ++q)
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    t[rr][q][i] = (q < nb) ? Lm[(size_t)((r_hi - rr) * kNB + 4 * g + i) * np + (b_hi - q) * kNB + c] : 0.0;
        if (tid == 0) {
            int spins = 0, okv = 1;
            while (ld_acquire(flags + gr) != epoch) {
                if (++spins > kSpinLimit) { okv = 0; break; }
            }
            ok_s = okv;
        }
        __syncthreads();
        if (!ok_s) {  // cannot happen with in-order dispatch; never hang the device
            if (tid == 0) status->failed = 1;
            return;
        }
        if (tid < kBackGroup * kNB) xr[tid >> 5][tid & 31] = __ldcg(xsol + (r_hi - (tid >> 5)) * kNB + (tid & 31));
        __syncthreads();
#pragma unroll
        for (int rr = 0; rr < kBackGroup; ++rr)
#pragma unroll
            for (int q = 0; q < kBackGroup; ++q)
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[q] = fma(t[rr][q][i], xr[rr][4 * g + i], acc[q]);
        __syncthreads();  // xr is rewritten in the next round
    }
    // inside the group, top-down (unrolled: acc[] stays in registers; nb is uniform over the CTA)
#pragma unroll
    for (int q = 0; q < kBackGroup; ++q) {
        if (q >= nb) break;
        const int b = b_hi - q;
        part[g][c] = acc[q];
        __syncthreads();
        if (tid < kNB) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kBackThreads / 32; ++w) s += part[w][tid];
            rhs[tid] = Lm[(size_t)np * np + b * kNB + tid] - s;  // y_b - s_b
        }
        __syncthreads();
        {   // x_b[i] = sum_r M[r][i] rhs[r]: 8 partial sums of 4 terms per output
            double s = 0.0;
#pragma unroll
            for (int m = 0; m < 4; ++m) s = fma(Li[q][4 * g + m][c], rhs[4 * g + m], s);
            part[g][c] = s;
        }
        __syncthreads();
        if (tid < kNB) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < kBackThreads / 32; ++w) s += part[w][tid];
            xr[q][tid] = s;
            __stcg(xsol + b * kNB + tid, s);
        }
        __syncthreads();
        // x_b into the sums of the group's lower blocks: tile L[b, b'] rows 4g .. 4g+3, column c
#pragma unroll
        for (int q2 = q + 1; q2 < kBackGroup; ++q2) {
            if (q2 >= nb) break;
            double s = acc[q2];
#pragma unroll
            for (int i = 0; i < 4; ++i)
                s = fma(Lm[(size_t)(b * kNB + 4 * g + i) * np + (b_hi - q2) * kNB + c], xr[q][4 * g + i], s);
            acc[q2] = s;
        }
    }
    __syncthreads();
    if (tid == 0) { __threadfence(); st_release(flags + gi, epoch); tl_stamp(tl, (3 * nblk + b_lo) * 2 + 1); }
    if (b_lo != 0 || !finish) return;   // finish == 0: the sparse-front kernels complete and apply the step

    // ---- block 0: all of x is published (x_0 depended on every other block) ----------------------
    __shared__ double red[kBackThreads / 32];
    __shared__ int decision;  // 0 apply, 1 stop, 2 fail
    double psum = 0.0;
    int nonfinite = 0;
    for (int i = tid; i < dim; i += kBackThreads) {
        const double v = __ldcg(xsol + i);
        dx_out[i] = v;
        psum += v * v;
        if (!(fabs(v) <= 1.7e308)) nonfinite = 1;
    }
    psum = warp_sum(psum);
    nonfinite = __any_sync(0xffffffffu, nonfinite);
    if ((tid & 31) == 0) red[tid >> 5] = nonfinite ? __longlong_as_double(0x7ff8000000000000LL) : psum;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < kBackThreads / 32; ++w) s += red[w];
        const double nrm = sqrt(s);
        status->last_norm = nrm;
        status->attempted += 1;
        int d;
        if (!(nrm == nrm)) d = 2;                 // non-finite step: treat as failed solve
        else if (nrm < delta_thresh) d = 1;       // converged: do not apply
        else d = 0;
        if (d == 1) status->stopped = 1;
        if (d == 2) status->failed = 1;
        if (d == 0) status->applied += 1;
        decision = d;
    }
    __syncthreads();
    if (decision != 0) return;
    for (int f = tid; f < num_free; f += kBackThreads) {
        double xi[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) xi[k] = __ldcg(xsol + 7 * f + k);
        if (metal) sim3_retract_metal_dev(xi, poses + 8 * free_kf[f]);
        else sim3_retract_dev(xi, poses + 8 * free_kf[f]);
    }
}

// ------------------------------------------------------------------------------------------------
// Sparse front (default when the pose graph leaves H sparse enough, M3S_GN_SPARSE=0 disables): keyframes
// that are mutually non-adjacent and have few neighbours are eliminated EXACTLY as 7x7 blocks, level by
// level (each level = an independent set of the graph that the previous levels left, fill included), and
// only the Schur complement on the remaining keyframes -- the "tail" -- goes through the dense blocked
// factorisation above.  For the window-2 chain + loop closures of config 5 (199 free keyframes) ~150
// keyframes leave in 7 levels and the dense chain shrinks from 22 tiles to 6.
//   level kernel (left-looking, one CTA per keyframe j of the level):
//     A~_aj = H_aj - sum_i W_ai W_ji^T   over the earlier-eliminated i with a, j in struct(i)   (a in {j} + struct(j))
//     b~_j  = -g_j - sum_i W_ji y_i
//     G = chol(A~_jj),  W_aj = A~_aj G^-T,  y_j = G^-1 b~_j         (W, G^-1, y kept for the back-substitution)
//   tail kernel: Hr_ab = H_ab - sum_i W_ai W_bi^T,  rhs_a = -g_a - sum_i W_ai y_i   (dense, tail order)
//   after the dense solve of the tail, the levels in reverse:  x_j = G^-T (y_j - sum_a W_aj^T x_a).
// Every sum runs over a host-built list in a fixed order: no atomics, bitwise reproducible, identical on every
// rank.  H (Hm) is only read.
// ------------------------------------------------------------------------------------------------
constexpr int kSpMaxRows = 25;      // 1 + the largest struct a level keyframe may have
constexpr int kSpThreads = 32 * (kSpMaxRows + 1);   // one warp per row block + one for the right-hand side

struct SpDev {
    const int* enode;      // [nE] free index of the eliminated keyframes, level-major
    const int* row_ptr;    // [nE + 1] rows of keyframe e: row 0 = itself, then struct(e)
    const int* row_node;   // [nrows] free index of the row's keyframe
    const int* con_ptr;    // [nrows + 1] contributor pairs of the row
    const int2* con;       // (W slot of (a, i), W slot of (j, i))
    const int* rc_ptr;     // [nE + 1] right-hand-side contributors of keyframe e
    const int2* rc;        // (W slot of (j, i), e index of i)
    double* W;             // [nrows][49]; the slot of row (e, s) is its row index (row 0 slots unused)
    double* Ginv;          // [nE][49]
    double* y;             // [nE][7]
};

// sum over a contributor list of the 7-term products  Wa[r, :] . Wb[c, :]  (rows of two 7x7 blocks), four list
// entries per round so that their index and block loads are in flight together (the lists are short and every
// load is an L2 round trip: taken one by one, a 25-entry list costs ~15k clk); fixed order, deterministic
__device__ __forceinline__ double sp_dot7(const double* __restrict__ a, const double* __restrict__ b) {
    double d = 0.0;
#pragma unroll
    for (int k = 0; k < 7; ++k) d = fma(a[k], b[k], d);
    return d;
}
__device__ __forceinline__ double sp_list_sum(const int2* __restrict__ lst, int q0, int q1,
                                              const double* __restrict__ Wa, int ra, const double* __restrict__ Wb,
                                              int rb, int sb) {
    double acc = 0.0;
    int q = q0;
    for (; q + 4 <= q1; q += 4) {
        const int2 w0 = lst[q], w1 = lst[q + 1], w2 = lst[q + 2], w3 = lst[q + 3];
        const double d0 = sp_dot7(Wa + (size_t)w0.x * 49 + ra, Wb + (size_t)w0.y * sb + rb);
        const double d1 = sp_dot7(Wa + (size_t)w1.x * 49 + ra, Wb + (size_t)w1.y * sb + rb);
        const double d2 = sp_dot7(Wa + (size_t)w2.x * 49 + ra, Wb + (size_t)w2.y * sb + rb);
        const double d3 = sp_dot7(Wa + (size_t)w3.x * 49 + ra, Wb + (size_t)w3.y * sb + rb);
        acc += (d0 + d1) + (d2 + d3);
    }
    for (; q < q1; ++q) {
        const int2 w = lst[q];
        acc += sp_dot7(Wa + (size_t)w.x * 49 + ra, Wb + (size_t)w.y * sb + rb);
    }
    return acc;
}

// One CTA per keyframe of the level, ONE WARP PER ROW BLOCK (rows <= 25) + one warp for the right-hand side.
// A row's contributor list is walked four entries at a time: the warp loads the 4 x (W_ai, W_ji) blocks -- 392
// doubles, 12-13 independent coalesced loads per lane, all in flight together -- into its slice of shared memory
// and every lane then forms its two entries of the 7x7 block from there.  (One thread per entry loading its own
// 14 doubles per contributor kept too few loads in flight: 10-18 us per level for a few MFLOP.)
constexpr int kSpBatch = 4;
constexpr int kSpStage = kSpBatch * 98;                       // doubles per warp
constexpr size_t kSpLevelSmem = sizeof(double) * kSpStage * kSpMaxRows;

__global__ void __launch_bounds__(kSpThreads)
sp_level_kernel(Status* __restrict__ status, SpDev sp, int e_begin, const double* __restrict__ Hm, int np) {
    if (status->stopped | status->failed) return;
    extern __shared__ __align__(16) double sp_stage[];
    __shared__ double At[kSpMaxRows][49];
    __shared__ double Gi[49];
    __shared__ double bt[7];
    const int e = e_begin + blockIdx.x;
    const int j = sp.enode[e];
    const int r0 = sp.row_ptr[e], nrows = sp.row_ptr[e + 1] - r0;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int grp = tid >> 6, t = tid & 63;
    const int r = t / 7, c = t % 7;
    const double* __restrict__ Wr = sp.W;   // (the slots read here were written by earlier launches, never by this one)
    if (warp < nrows) {
        const int s = warp;
        const int a = sp.row_node[r0 + s];
        double* st = sp_stage + (size_t)warp * kSpStage;
        const int t0 = lane, t1 = lane + 32;                  // this lane's entries of the block (t1 < 49: lanes 0 .. 16)
        const int ra0 = (t0 / 7) * 7, cb0 = 49 + (t0 % 7) * 7;
        const int ra1 = (t1 < 49 ? t1 / 7 : 0) * 7, cb1 = 49 + (t1 < 49 ? t1 % 7 : 0) * 7;
        double acc0 = 0.0, acc1 = 0.0;
        const int q0 = sp.con_ptr[r0 + s], q1 = sp.con_ptr[r0 + s + 1];
        for (int q = q0; q < q1; q += kSpBatch) {
            const int nb = min(kSpBatch, q1 - q);
            int2 w = make_int2(0, 0);
            if (lane < nb) w = sp.con[q + lane];
            double v[(kSpStage + 31) / 32];
#pragma unroll
            for (int k = 0; k < (kSpStage + 31) / 32; ++k) {
                const int idx = lane + 32 * k;
                const int cc = idx / 98, off = idx % 98;
                const int wx = __shfl_sync(0xffffffffu, w.x, cc & (kSpBatch - 1));
                const int wy = __shfl_sync(0xffffffffu, w.y, cc & (kSpBatch - 1));
                v[k] = 0.0;
                if (idx < kSpStage && cc < nb)
                    v[k] = (off < 49) ? Wr[(size_t)wx * 49 + off] : Wr[(size_t)wy * 49 + off - 49];
            }
            __syncwarp();                                     // the previous round's reads of the slice are done
#pragma unroll
            for (int k = 0; k < (kSpStage + 31) / 32; ++k)
                if (lane + 32 * k < kSpStage) st[lane + 32 * k] = v[k];
            __syncwarp();
            double d0[kSpBatch], d1[kSpBatch];
#pragma unroll
            for (int cc = 0; cc < kSpBatch; ++cc) {          // entries past the end of the list were staged as zeros
                d0[cc] = sp_dot7(st + cc * 98 + ra0, st + cc * 98 + cb0);
                d1[cc] = sp_dot7(st + cc * 98 + ra1, st + cc * 98 + cb1);
            }
            acc0 += (d0[0] + d0[1]) + (d0[2] + d0[3]);
            acc1 += (d1[0] + d1[1]) + (d1[2] + d1[3]);
        }
        At[s][t0] = Hm[(size_t)(7 * a + t0 / 7) * np + 7 * j + t0 % 7] - acc0;
        if (t1 < 49) At[s][t1] = Hm[(size_t)(7 * a + t1 / 7) * np + 7 * j + t1 % 7] - acc1;
    } else if (warp == kSpMaxRows && lane < 7) {
        bt[lane] = Hm[(size_t)np * np + 7 * j + lane] -
                   sp_list_sum(sp.rc, sp.rc_ptr[e], sp.rc_ptr[e + 1], sp.W, lane * 7, sp.y, 0, 7);
    }
    __syncthreads();
    if (tid == 0) {
        // 7x7 Cholesky and G^-1 by one thread, everything in registers (fully unrolled): the block is on the
        // critical path of the level, and 28 + 28 values through shared memory with IEEE sqrt / division took
        // longer than the rest of the kernel
        double g[7][7], gi[7], v[7][7];
        bool bad = false;
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
            for (int k = 0; k <= i; ++k) g[i][k] = At[0][i * 7 + k];
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            double d = g[k][k];
#pragma unroll
            for (int m = 0; m < k; ++m) d = fma(-g[k][m], g[k][m], d);
            if (!(d > 0.0) || !(d <= 1.7e308)) { bad = true; d = 1.0; }
            gi[k] = rsqrt_fast(d);
            g[k][k] = d * gi[k];
#pragma unroll
            for (int i = k + 1; i < 7; ++i) {
                double x = g[i][k];
#pragma unroll
                for (int m = 0; m < k; ++m) x = fma(-g[i][m], g[k][m], x);
                g[i][k] = x * gi[k];
            }
        }
        // V = G^-1 (lower): V_ii = 1 / G_ii, V_ij = -V_ii sum_{k = j .. i-1} G_ik V_kj
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            v[j][j] = gi[j];
#pragma unroll
            for (int i = j + 1; i < 7; ++i) {
                double x = 0.0;
#pragma unroll
                for (int k = j; k < i; ++k) x = fma(g[i][k], v[k][j], x);
                v[i][j] = -x * gi[i];
            }
        }
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
            for (int j = 0; j < 7; ++j) Gi[i * 7 + j] = (j <= i) ? v[i][j] : 0.0;
        if (bad) status->chol_fail = 1;
    }
    __syncthreads();
    for (int s = 1 + grp; s < nrows; s += kSpThreads / 64) {
        if (t < 49) {   // W_aj = A~_aj G^-T
            double d = 0.0;
#pragma unroll
            for (int k = 0; k < 7; ++k) d = fma(At[s][r * 7 + k], Gi[c * 7 + k], d);
            sp.W[(size_t)(r0 + s) * 49 + t] = d;
        }
    }
    if (tid < 49) sp.Ginv[(size_t)e * 49 + tid] = Gi[tid];
    if (tid >= 64 && tid < 71) {
        const int rr = tid - 64;
        double d = 0.0;
#pragma unroll
        for (int k = 0; k < 7; ++k) d = fma(Gi[rr * 7 + k], bt[k], d);
        sp.y[(size_t)e * 7 + rr] = d;
    }
}

// Schur complement on the tail, written as the dense augmented matrix the blocked factorisation reads
// (row-major [npR + 1][npR], both triangles, rhs in row npR; padding = identity, set once at plan creation).
struct SpTailDev {
    const int* rnode;      // [nR] free index of tail keyframe
    const int2* blk;       // [nblk] (ra, rb), ra >= rb, tail indices
    const int* con_ptr;    // [nblk + 1]
    const int2* con;       // (W slot of (a, i), W slot of (b, i))
    const int* rc_ptr;     // [nR + 1]
    const int2* rc;        // (W slot of (a, i), e index of i)
};

__global__ void __launch_bounds__(64)
sp_tail_kernel(const Status* __restrict__ status, SpDev sp, SpTailDev tl, const double* __restrict__ Hm, int np,
               double* __restrict__ Hr, int npR, unsigned* __restrict__ tctl, int n_tctl) {
    if (status->stopped | status->failed) return;
    // flag words of the dataflow factorisation that follows (instead of a memset node between the two)
    if (blockIdx.x == 0) for (int i = threadIdx.x; i < n_tctl; i += 64) tctl[i] = 0u;
    const int2 rb = tl.blk[blockIdx.x];
    const int a = tl.rnode[rb.x], b = tl.rnode[rb.y];
    const int t = threadIdx.x;
    if (t < 49) {
        const int r = t / 7, c = t % 7;
        const double acc = Hm[(size_t)(7 * a + r) * np + 7 * b + c] -
                           sp_list_sum(tl.con, tl.con_ptr[blockIdx.x], tl.con_ptr[blockIdx.x + 1], sp.W, r * 7, sp.W, c * 7, 49);
        Hr[(size_t)(7 * rb.x + r) * npR + 7 * rb.y + c] = acc;
        if (rb.x != rb.y) Hr[(size_t)(7 * rb.y + c) * npR + 7 * rb.x + r] = acc;
    } else if (rb.x == rb.y && t < 56) {
        const int rr = t - 49;
        Hr[(size_t)npR * npR + 7 * rb.x + rr] =
            Hm[(size_t)np * np + 7 * a + rr] -
            sp_list_sum(tl.rc, tl.rc_ptr[rb.x], tl.rc_ptr[rb.x + 1], sp.W, rr * 7, sp.y, 0, 7);
    }
}

// The whole back end of the sparse-front solve in ONE CTA: tail x into natural order, the levels in reverse
// (x_j = G^-T (y_j - sum_a W_aj^T x_a); 32 keyframes at a time, one warp each), the ||dx|| test before applying
// (gauss_newton.py:263-265) and the Sim3 retraction.  (As one launch per level + scatter + finish this was seven
// launches of 3-10 us each for ~50k FMAs of work.)
__global__ void __launch_bounds__(1024)
sp_back_finish_kernel(Status* __restrict__ status, SpDev sp, const int* __restrict__ level_begin, int levels,
                      const int* __restrict__ rnode, int nR, const double* __restrict__ xsol,
                      double* __restrict__ dx, int dim, int num_free, const int* __restrict__ free_kf,
                      double delta_thresh, double* __restrict__ poses, int metal) {
    if (status->stopped | status->failed) return;
    __shared__ double red[32];
    __shared__ int decision;  // 0 apply, 1 stop, 2 fail
    const int tid = threadIdx.x;
    for (int i = tid; i < 7 * nR; i += 1024) dx[7 * rnode[i / 7] + i % 7] = __ldcg(xsol + i);
    __syncthreads();
    // one WARP per keyframe (32 at a time), everything inside the warp: lane = 7 p + k holds partial p (rows
    // 1 + p, 5 + p, ...) of component k; the CTA only synchronises between levels
    const int warp = tid >> 5, lane = tid & 31;
    for (int l = levels - 1; l >= 0; --l) {
        const int b = level_begin[l], n = level_begin[l + 1] - b;
        for (int it = warp; it < n; it += 32) {
            const int e = b + it;
            const int r0 = sp.row_ptr[e], nrows = sp.row_ptr[e + 1] - r0;
            const int k = lane % 7, p = lane / 7;            // lanes 28 .. 31: p = 4, idle
            double acc = 0.0;
            if (lane < 28) {
                for (int s = 1 + p; s < nrows; s += 4) {
                    const double* w = sp.W + (size_t)(r0 + s) * 49 + k;
                    const double* xa = dx + 7 * sp.row_node[r0 + s];
#pragma unroll
                    for (int r = 0; r < 7; ++r) acc = fma(w[r * 7], xa[r], acc);
                }
            }
            acc += __shfl_down_sync(0xffffffffu, acc, 14);   // partials 2, 3 onto 0, 1
            acc += __shfl_down_sync(0xffffffffu, acc, 7);    // partial 1 onto 0: lanes 0 .. 6 hold the sums
            const double tk = (lane < 7) ? sp.y[(size_t)e * 7 + lane] - acc : 0.0;
            const double* gi = sp.Ginv + (size_t)e * 49;
            double d = 0.0;
#pragma unroll
            for (int kk = 0; kk < 7; ++kk) {
                const double t_kk = __shfl_sync(0xffffffffu, tk, kk);
                if (lane < 7) d = fma(gi[kk * 7 + lane], t_kk, d);
            }
            if (lane < 7) dx[7 * sp.enode[e] + lane] = d;
        }
        __syncthreads();   // the level's x is visible to the whole CTA before the next level reads it
    }
    double psum = 0.0;
    int nonfinite = 0;
    for (int i = tid; i < dim; i += 1024) {
        const double v = dx[i];
        psum += v * v;
        if (!(fabs(v) <= 1.7e308)) nonfinite = 1;
    }
    psum = warp_sum(psum);
    nonfinite = __any_sync(0xffffffffu, nonfinite);
    if ((tid & 31) == 0) red[tid >> 5] = nonfinite ? __longlong_as_double(0x7ff8000000000000LL) : psum;
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int w = 0; w < 32; ++w) s += red[w];
        const double nrm = sqrt(s);
        status->last_norm = nrm;
        status->attempted += 1;
        int d;
        if (!(nrm == nrm)) d = 2;
        else if (nrm < delta_thresh) d = 1;
        else d = 0;
        if (d == 1) status->stopped = 1;
        if (d == 2) status->failed = 1;
        if (d == 0) status->applied += 1;
        decision = d;
    }
    __syncthreads();
    if (decision != 0) return;
    for (int f = tid; f < num_free; f += 1024) {
        double xi[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) xi[k] = dx[7 * f + k];
        if (metal) sim3_retract_metal_dev(xi, poses + 8 * free_kf[f]);
        else sim3_retract_dev(xi, poses + 8 * free_kf[f]);
    }
}

__global__ void gn_poses_f32_to_f64(const float* __restrict__ in, double* __restrict__ out, int n,
                                    Status* status) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)in[i];
    if (i == 0 && status) {
        status->applied = 0; status->stopped = 0; status->failed = 0; status->attempted = 0;
        status->chol_fail = 0; status->last_norm = 0.0;
    }
}

__global__ void gn_poses_f64_to_f32(const double* __restrict__ in, float* __restrict__ out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (float)in[i];
}

}  // namespace m3s

using namespace m3s;

// ------------------------------------------------------------------------------------------------
// plan
// ------------------------------------------------------------------------------------------------
struct m3s_gn_plan {
    int device = 0;
    int E = 0, num_kf = 0, pin = 0;
    long long N = 0;
    int F = 0, dim = 0, np = 0;
    int nchunks = 1;
    long long chunk_pts = 0;
    int nblocks_H = 0;
    // device buffers
    double* poses = nullptr;        // [num_kf, 8]
    EdgeMeta* meta = nullptr;       // [E]
    double* partials = nullptr;     // [E, nchunks, 36]
    double* edge_blocks = nullptr;  // [E, 36]
    BlockRef* blocks = nullptr;
    int* list_edge = nullptr;
    int* list_sign = nullptr;
    int* free_kf = nullptr;         // [F]
    double* Hm = nullptr;           // [np + 1, np]
    double* Ldiag = nullptr;        // [np / 32, 32, 32] diagonal factor blocks
    double* Linv = nullptr;         // [np / 32, 32, 32] their inverses
    double* Lm = nullptr;           // [np + 1, np] the Cholesky factor (block columns below the
                                    // diagonal + diagonal blocks) and, in row np, y = L^-1 (-g)
    double* xsol = nullptr;         // [np] back-substitution result blocks
    int* flags = nullptr;           // [np / 32] publication flags of the back-substitution
    unsigned* gbar = nullptr;       // grid-barrier counter of the persistent factorisation
    int fused_grid = -1;            // CTAs of chol_fused_kernel (0 = not available, -1 = not probed)
    bool hm_clean = false;          // the structurally-zero entries of Hm are zero (see m3s_gn_assemble)
    unsigned* tctl = nullptr;       // flag words of the dataflow factorisation (chol_tiles_kernel)
    int tiles_grid = -1;            // its CTA count (0 = not available / not selected, -1 = not probed)
    long long* timeline = nullptr;  // [4, np / 32, 2] globaltimer stamps (M3S_GN_TIMELINE=1), else null
    // sparse front (sp_level_kernel ...): levels of exactly eliminated keyframes + a dense tail
    bool sp_on = false;
    int sp_nE = 0, sp_nR = 0, sp_dimR = 0, sp_npR = 0, sp_ntail_blocks = 0;
    std::vector<int> sp_level_begin;   // [levels + 1] into the level-major list of eliminated keyframes
    const int* sp_level_begin_d = nullptr;
    SpDev sp{};
    SpTailDev spt{};
    double* sp_Hr = nullptr;           // [npR + 1, npR] Schur complement on the tail (+ rhs row)
    int sp_tiles_grid = 0;             // CTAs of chol_tiles_kernel for the tail
    std::vector<void*> sp_allocs;
    double* g = nullptr;            // [dim]
    double* dx = nullptr;           // [dim]
    Status* status = nullptr;
    // assemble + factorisation (~90 launches) captured once per plan; all arguments are plan-constant
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    bool graph_failed = false;
    // peer exchange (m3s_gn_exchange_*): one IPC-exported allocation per rank
    //   [2][E,36] f64 receive buffers | kMaxPeers u64 flags | u32 ticket
    int xch_rank = 0, xch_world = 0;          // world > 1 once attached
    bool xch_attached = false;
    char* xch_local = nullptr;
    char* xch_peer[kMaxPeers] = {};            // peer mappings of the same region (own = local)
    unsigned long long xch_epoch = 0;
    unsigned long long xch_timeout_ns = 10000000000ull;
    bool xch_exported = false;   // an IPC handle of xch_local was handed out
    bool xch_released = true;    // every importer has unmapped it (m3s_gn_exchange_release)
};

static size_t xch_recv_bytes(const m3s_gn_plan* p) {
    return sizeof(double) * kEdgeStride * (size_t)std::max(p->E, 1);
}
static size_t xch_total_bytes(const m3s_gn_plan* p) {
    return 2 * xch_recv_bytes(p) + sizeof(unsigned long long) * kMaxPeers + 64;
}
static void xch_unmap(m3s_gn_plan* p) {  // close the peer mappings, keep the local region
    for (int r = 0; r < kMaxPeers; ++r) {
        if (p->xch_peer[r] && p->xch_peer[r] != p->xch_local) cudaIpcCloseMemHandle(p->xch_peer[r]);
        p->xch_peer[r] = nullptr;
    }
    p->xch_attached = false;
    cudaGetLastError();
}
static void xch_free(m3s_gn_plan* p) {
    xch_unmap(p);
    // Freeing an IPC-exported allocation that a peer process may still have mapped is undefined
    // behaviour.  Unless the caller confirmed (m3s_gn_exchange_release, after its own barrier) that
    // every importer has unmapped, the region is deliberately leaked (<= 2 x E x 288 B + 128 B).
    if (p->xch_local && (!p->xch_exported || p->xch_released)) cudaFree(p->xch_local);
    p->xch_local = nullptr;
    p->xch_exported = false;
    p->xch_released = true;
    p->xch_world = 0;
    cudaGetLastError();
}

static void plan_free(m3s_gn_plan* p);
static void sp_free(m3s_gn_plan* p) {
    for (void* q : p->sp_allocs) cudaFree(q);
    p->sp_allocs.clear();
    p->sp_on = false;
}
static void plan_free(m3s_gn_plan* p) {
    if (!p) return;
    xch_free(p);
    sp_free(p);
    if (p->graph_exec) cudaGraphExecDestroy(p->graph_exec);
    if (p->graph) cudaGraphDestroy(p->graph);
    cudaFree(p->Lm); cudaFree(p->xsol); cudaFree(p->flags); cudaFree(p->gbar); cudaFree(p->tctl); cudaFree(p->timeline);
    cudaFree(p->poses); cudaFree(p->meta); cudaFree(p->partials); cudaFree(p->edge_blocks);
    cudaFree(p->blocks); cudaFree(p->list_edge); cudaFree(p->list_sign); cudaFree(p->free_kf);
    cudaFree(p->Hm); cudaFree(p->Ldiag); cudaFree(p->Linv); cudaFree(p->g); cudaFree(p->dx); cudaFree(p->status);
    delete p;
}

template <typename T>
static bool dev_alloc_copy(T** dst, const std::vector<T>& src) {
    const size_t bytes = std::max<size_t>(src.size(), 1) * sizeof(T);
    if (cudaMalloc(reinterpret_cast<void**>(dst), bytes) != cudaSuccess) return false;
    if (!src.empty() && cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess)
        return false;
    return true;
}

// ------------------------------------------------------------------------------------------------
// Sparse front, host side: symbolic elimination on the keyframe graph at plan creation.
// Level l = a greedy independent set (lowest current degree first) of the keyframes with at most
// kSpMaxRows - 1 neighbours in the graph the earlier levels left (fill edges included).  Levels are added
// while one is worth its two launches (>= kSpMinLevel keyframes ~ 2/3 of a 64-row tile of the dense chain).
// Selected by default when it removes at least one tile from the dense chain; M3S_GN_SPARSE=0 or an
// explicit M3S_GN_CHOL=steps|fused|tiles keeps the whole system dense, M3S_GN_CHOL=sparse forces it.
// ------------------------------------------------------------------------------------------------
constexpr int kSpMinLevel = 6;
constexpr int kSpMaxLevels = 32;

template <typename T>
static bool sp_upload(m3s_gn_plan* p, const T** dst, const std::vector<T>& src) {
    T* d = nullptr;
    const size_t n = std::max<size_t>(src.size(), 1);
    if (cudaMalloc((void**)&d, sizeof(T) * n) != cudaSuccess) return false;
    p->sp_allocs.push_back(d);
    if (!src.empty() && cudaMemcpy(d, src.data(), sizeof(T) * src.size(), cudaMemcpyHostToDevice) != cudaSuccess) return false;
    *dst = d;
    return true;
}

static bool sp_env_allows(bool* forced) {
    *forced = false;
    if (const char* e = getenv("M3S_GN_SPARSE")) if (e[0] == '0') return false;
    if (const char* e = getenv("M3S_GN_CHOL")) {
        if (strcmp(e, "sparse") == 0) { *forced = true; return true; }
        if (e[0] != 0) return false;   // steps | fused | tiles: the dense paths
    }
    return true;
}

// off_pairs: the structurally non-zero off-diagonal blocks (hi, lo) of H in free-keyframe indices
static bool build_sparse_front(m3s_gn_plan* p, const std::vector<std::pair<int, int>>& off_pairs) {
    bool forced = false;
    const int F = p->F;
    if (!sp_env_allows(&forced) || F < 2) return true;
    std::vector<std::set<int>> adj(F);
    for (auto& pr : off_pairs) { adj[pr.first].insert(pr.second); adj[pr.second].insert(pr.first); }
    std::vector<char> alive(F, 1);
    std::vector<std::vector<int>> strct(F);
    std::vector<int> elim_pos(F, -1);        // e index (level-major) of an eliminated keyframe
    std::vector<int> enode, level_begin{0};
    int n_alive = F;
    int min_level = kSpMinLevel, max_levels = kSpMaxLevels;   // M3S_GN_SP_MINLEVEL / _MAXLEVELS: A/B knobs
    if (const char* e = getenv("M3S_GN_SP_MINLEVEL")) min_level = std::max(1, atoi(e));
    if (const char* e = getenv("M3S_GN_SP_MAXLEVELS")) max_levels = std::max(1, atoi(e));
    for (int l = 0; l < max_levels; ++l) {
        std::vector<std::pair<int, int>> cand;
        for (int v = 0; v < F; ++v)
            if (alive[v] && (int)adj[v].size() <= kSpMaxRows - 1) cand.push_back({(int)adj[v].size(), v});
        std::sort(cand.begin(), cand.end());
        std::vector<char> blocked(F, 0);
        std::vector<int> sel;
        for (auto& dv : cand) {
            const int v = dv.second;
            if (blocked[v] || (int)sel.size() >= n_alive - 1) continue;   // the tail keeps >= 1 keyframe
            sel.push_back(v);
            blocked[v] = 1;
            for (int n : adj[v]) blocked[n] = 1;
        }
        if ((int)sel.size() < (forced ? 1 : min_level)) break;
        std::sort(sel.begin(), sel.end());
        for (int v : sel) {
            strct[v].assign(adj[v].begin(), adj[v].end());
            for (int a : strct[v]) {
                adj[a].erase(v);
                for (int b : strct[v]) if (b != a) adj[a].insert(b);
            }
            adj[v].clear();
            alive[v] = 0;
            elim_pos[v] = (int)enode.size();
            enode.push_back(v);
            --n_alive;
        }
        level_begin.push_back((int)enode.size());
    }
    const int nE = (int)enode.size(), nR = F - nE;
    if (nE == 0) return true;
    const int dimR = 7 * nR, npR = ((dimR + kNB - 1) / kNB) * kNB;
    const int nt_full = (p->np + kTS - 1) / kTS, nt_tail = (npR + kTS - 1) / kTS;
    if (!forced && nt_tail >= nt_full) return true;   // nothing to win
    // rows, W slots (slot = global row index), contributor lists
    std::vector<int> row_ptr{0}, row_node, con_ptr{0}, rc_ptr{0};
    std::vector<int2> con, rc;
    std::map<std::pair<int, int>, int> wslot;   // (a, i) -> row index of a in keyframe i's rows
    for (int e = 0; e < nE; ++e) {
        const int j = enode[e];
        row_node.push_back(j);
        for (int a : strct[j]) { wslot[{a, j}] = (int)row_node.size(); row_node.push_back(a); }
        row_ptr.push_back((int)row_node.size());
    }
    // contributors of keyframe v: the earlier-eliminated i with v in struct(i), in elimination order
    std::vector<std::vector<int>> contrib_of(F);
    for (int e = 0; e < nE; ++e)
        for (int a : strct[enode[e]]) contrib_of[a].push_back(e);
    for (int e = 0; e < nE; ++e) {
        const int j = enode[e];
        for (int r = row_ptr[e]; r < row_ptr[e + 1]; ++r) {
            const int a = row_node[r];
            for (int ei : contrib_of[j]) {           // ei < e by construction (j was alive when ei left)
                const int i = enode[ei];
                auto ia = (a == j) ? wslot.find({j, i}) : wslot.find({a, i});
                if (ia == wslot.end()) continue;     // a is not in struct(i)
                con.push_back(make_int2(ia->second, wslot[{j, i}]));
            }
            con_ptr.push_back((int)con.size());
        }
        for (int ei : contrib_of[j]) rc.push_back(make_int2(wslot[{j, enode[ei]}], ei));
        rc_ptr.push_back((int)rc.size());
    }
    // tail
    std::vector<int> rnode;
    for (int v = 0; v < F; ++v) if (alive[v]) rnode.push_back(v);
    std::vector<int2> blk, tcon, trc;
    std::vector<int> tcon_ptr{0}, trc_ptr{0};
    for (int ra = 0; ra < nR; ++ra) {
        const int a = rnode[ra];
        for (int rb = 0; rb <= ra; ++rb) {
            const int b = rnode[rb];
            blk.push_back(make_int2(ra, rb));
            // contributors: i with a and b in struct(i) (both lists are in elimination order)
            const auto &ca = contrib_of[a], &cb = contrib_of[b];
            size_t x = 0, y = 0;
            while (x < ca.size() && y < cb.size()) {
                if (ca[x] < cb[y]) ++x;
                else if (ca[x] > cb[y]) ++y;
                else {
                    const int i = enode[ca[x]];
                    tcon.push_back(make_int2(wslot[{a, i}], wslot[{b, i}]));
                    ++x; ++y;
                }
            }
            tcon_ptr.push_back((int)tcon.size());
        }
        for (int ei : contrib_of[a]) trc.push_back(make_int2(wslot[{a, enode[ei]}], ei));
        trc_ptr.push_back((int)trc.size());
    }
    bool ok = true;
    ok = ok && sp_upload(p, &p->sp.enode, enode) && sp_upload(p, &p->sp.row_ptr, row_ptr) &&
         sp_upload(p, &p->sp.row_node, row_node) && sp_upload(p, &p->sp.con_ptr, con_ptr) &&
         sp_upload(p, &p->sp.con, con) && sp_upload(p, &p->sp.rc_ptr, rc_ptr) && sp_upload(p, &p->sp.rc, rc) &&
         sp_upload(p, &p->spt.rnode, rnode) && sp_upload(p, &p->spt.blk, blk) &&
         sp_upload(p, &p->spt.con_ptr, tcon_ptr) && sp_upload(p, &p->spt.con, tcon) &&
         sp_upload(p, &p->spt.rc_ptr, trc_ptr) && sp_upload(p, &p->spt.rc, trc) &&
         sp_upload(p, &p->sp_level_begin_d, level_begin);
    auto dalloc = [&](double** d, size_t n) {
        if (cudaMalloc((void**)d, sizeof(double) * std::max<size_t>(n, 1)) != cudaSuccess) return false;
        p->sp_allocs.push_back(*d);
        return cudaMemset(*d, 0, sizeof(double) * std::max<size_t>(n, 1)) == cudaSuccess;
    };
    ok = ok && dalloc(&p->sp.W, (size_t)row_node.size() * 49) && dalloc(&p->sp.Ginv, (size_t)nE * 49) &&
         dalloc(&p->sp.y, (size_t)nE * 7) && dalloc(&p->sp_Hr, (size_t)(npR + 1) * npR);
    if (ok && npR > dimR) {   // identity on the padded diagonal, once: sp_tail_kernel never writes there
        std::vector<double> ones(npR - dimR, 1.0);
        ok = cudaMemcpy2D(p->sp_Hr + (size_t)dimR * npR + dimR, sizeof(double) * (npR + 1), ones.data(),
                          sizeof(double), sizeof(double), npR - dimR, cudaMemcpyHostToDevice) == cudaSuccess;
    }
    if (!ok) { cudaGetLastError(); sp_free(p); return false; }
    p->sp_nE = nE; p->sp_nR = nR; p->sp_dimR = dimR; p->sp_npR = npR;
    p->sp_ntail_blocks = (int)blk.size();
    p->sp_level_begin = level_begin;
    p->sp_on = true;
    return true;
}

M3S_API m3s_gn_plan* m3s_gn_plan_create(const int32_t* ii, const int32_t* jj, int E, int num_kf,
                                        int64_t N, int pin, int device) {
    if (E < 0 || num_kf < 0 || N < 0 || pin < 0 || (E > 0 && (!ii || !jj))) {
        set_error("gn_plan_create: bad argument");
        return nullptr;
    }
    for (int e = 0; e < E; ++e) {
        if (ii[e] < 0 || ii[e] >= num_kf || jj[e] < 0 || jj[e] >= num_kf) {
            set_error("gn_plan_create: edge %d references keyframe outside [0, %d)", e, num_kf);
            return nullptr;
        }
    }
    int cur_dev = -1;
    cudaGetDevice(&cur_dev);
    if (cur_dev != device && cudaSetDevice(device) != cudaSuccess) {
        cuda_fail(cudaGetLastError(), "cudaSetDevice");
        return nullptr;
    }
    struct Restore { int d, want; ~Restore() { if (d >= 0 && d != want) cudaSetDevice(d); } } restore{cur_dev, device};
    m3s_gn_plan* p = new (std::nothrow) m3s_gn_plan();
    if (!p) { set_error("gn_plan_create: out of memory"); return nullptr; }
    p->device = device; p->E = E; p->num_kf = num_kf; p->pin = pin; p->N = N;

    // free-keyframe numbering: gauss_newton.py:66-81
    std::vector<int> uniq;
    if (E > 0 && num_kf > pin) {
        uniq.assign(ii, ii + E);
        uniq.insert(uniq.end(), jj, jj + E);
        std::sort(uniq.begin(), uniq.end());
        uniq.erase(std::unique(uniq.begin(), uniq.end()), uniq.end());
    }
    const int nu = (int)uniq.size();
    p->F = (nu > pin) ? nu - pin : 0;
    p->dim = 7 * p->F;
    p->np = ((p->dim + kNB - 1) / kNB) * kNB;
    std::vector<int> local(std::max(num_kf, 1), -1);
    std::vector<int> free_kf;
    for (int k = pin; k < nu; ++k) { local[uniq[k]] = k - pin; free_kf.push_back(uniq[k]); }

    std::vector<EdgeMeta> meta(E);
    int n_active = 0;
    for (int e = 0; e < E; ++e) {
        EdgeMeta m{};
        m.ix = ii[e]; m.jx = jj[e];
        m.il = p->F ? local[m.ix] : -1;
        m.jl = p->F ? local[m.jx] : -1;
        // both ends fixed: skipped by the reference (gauss_newton.py:113); self edge: its four
        // contributions cancel exactly (H_ii + H_jj - H_ij - H_ji = 0, g_i + g_j = 0)
        m.active = (m.il >= 0 || m.jl >= 0) && (m.ix != m.jx);
        n_active += m.active;
        meta[e] = m;
    }
    // assembly lists
    std::vector<BlockRef> blocks;
    std::vector<int> list_edge, list_sign;
    std::vector<std::pair<int, int>> off_pairs;
    if (p->F > 0) {
        std::vector<std::vector<std::pair<int, int>>> diag(p->F);
        std::map<std::pair<int, int>, std::vector<int>> off;
        for (int e = 0; e < E; ++e) {
            const EdgeMeta& m = meta[e];
            if (!m.active) continue;
            if (m.il >= 0) diag[m.il].push_back({e, -1});  // g_i -= b
            if (m.jl >= 0) diag[m.jl].push_back({e, +1});  // g_j += b
            if (m.il >= 0 && m.jl >= 0) {
                const int hi = std::max(m.il, m.jl), lo = std::min(m.il, m.jl);
                off[{hi, lo}].push_back(e);
            }
        }
        for (int f = 0; f < p->F; ++f) {
            BlockRef b{f, f, (int)list_edge.size(), 0};
            for (auto& pr : diag[f]) { list_edge.push_back(pr.first); list_sign.push_back(pr.second); }
            b.end = (int)list_edge.size();
            blocks.push_back(b);
        }
        for (auto& kv : off) {
            off_pairs.push_back(kv.first);
            BlockRef b{kv.first.first, kv.first.second, (int)list_edge.size(), 0};
            for (int e : kv.second) { list_edge.push_back(e); list_sign.push_back(0); }
            b.end = (int)list_edge.size();
            blocks.push_back(b);
        }
    }
    p->nblocks_H = (int)blocks.size();

    // chunking of the point dimension: aim for >= ~6 CTAs per SM in flight, 1024..32768 pts/CTA
    // (B200, 1000 edges x 196608 pts: 8192 -> 1.56 ms, 16384 -> 1.43, 32768 -> 1.38, 65536 -> 1.39:
    // the per-CTA prologue (relative pose in fp64) and the fp64 block reduction amortise)
    {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const long long target_blocks = 6LL * sms;
        long long nch = (n_active > 0) ? (target_blocks + n_active - 1) / n_active : 1;
        const long long max_ch = std::max<long long>(1, (N + 1023) / 1024);
#ifndef M3S_ACC_MAX_CHUNK
#define M3S_ACC_MAX_CHUNK 32768
#endif
        const long long min_ch = std::max<long long>(1, (N + M3S_ACC_MAX_CHUNK - 1) / M3S_ACC_MAX_CHUNK);
        nch = std::min(std::max(nch, min_ch), max_ch);
#ifndef M3S_ACC_NO_WAVE_FIT
        // Wave fit: (edge, chunk) CTAs of equal length run in waves of 3 per SM (the streamed kernel's residency),
        // so the pass costs ceil(CTAs / slots) chunk times: among the chunk counts near the target take the one
        // with the cheapest last wave (~400 points of fixed cost per CTA, from the 16 k / 32 k / 64 k A/B).
        // The plan does not know how many ranks will share its edges (and must chunk identically on all of them:
        // the chunk sums are part of the bitwise result), so the fit is over 1 / 2 / 4 / 8 equal shares at once.
        // Config 5 (1000 edges): 6 x 32 768 = 13.5 / 6.8 / 3.4 / 1.7 waves -> 7 x 28 160 = 15.8 / 7.9 / 3.9 / 1.97
        // (measured: 1 GPU 1.149 -> 1.152 ms, i.e. nothing; 8 GPUs see the A/B in DESIGN.md); config 3 (40 edges):
        // 23 chunks = 2.07 waves -> one full wave, 0.128 -> 0.107 ms per iteration.
        if (n_active > 0) {
            const long long slots = 3LL * sms;
            const long long hi = std::min(max_ch, std::max(nch + 6, min_ch + 6));
            long long best_cost = -1, best = nch;
            for (long long c = min_ch; c <= hi; ++c) {
                long long cpc = (N + c - 1) / c;
                cpc = std::max<long long>(((cpc + 511) / 512) * 512, 512);   // whole stages of the streamed kernel
                const long long real = std::max<long long>(1, (N + cpc - 1) / cpc);
                long long cost = 0;
                for (int w = 1; w <= 8; w *= 2) {
                    const long long share = (n_active + w - 1) / w;
                    if (w > 1 && share * real < slots) break;   // a share below one wave: not a sharding worth fitting
                    cost += w * ((share * real + slots - 1) / slots) * (cpc + 400);
                }
                if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = c; }
            }
            nch = best;
        }
#endif
        long long cp = (N + nch - 1) / std::max<long long>(nch, 1);
        cp = std::max<long long>(((cp + kAccThreads - 1) / kAccThreads) * kAccThreads, kAccThreads);
        if (N >= 4096) cp = ((cp + 511) / 512) * 512;   // whole 512-point stages (streamed kernel)
        p->chunk_pts = cp;
        p->nchunks = (int)std::max<long long>(1, (N + cp - 1) / cp);
    }

    bool ok = true;
    ok = ok && cudaMalloc((void**)&p->poses, sizeof(double) * 8 * std::max(num_kf, 1)) == cudaSuccess;
    ok = ok && dev_alloc_copy(&p->meta, meta);
    ok = ok && cudaMalloc((void**)&p->partials, sizeof(double) * kEdgeStride * (size_t)std::max(E, 1) * p->nchunks) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->edge_blocks, sizeof(double) * kEdgeStride * (size_t)std::max(E, 1)) == cudaSuccess;
    ok = ok && dev_alloc_copy(&p->blocks, blocks);
    ok = ok && dev_alloc_copy(&p->list_edge, list_edge);
    ok = ok && dev_alloc_copy(&p->list_sign, list_sign);
    ok = ok && dev_alloc_copy(&p->free_kf, free_kf);
    const size_t hm = (size_t)(p->np + 1) * std::max(p->np, 1);
    ok = ok && cudaMalloc((void**)&p->Hm, sizeof(double) * hm) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->Ldiag, sizeof(double) * kNB * std::max(p->np, kNB)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->Linv, sizeof(double) * kNB * std::max(p->np, kNB)) == cudaSuccess;
    const int nblk = std::max(p->np / kNB, 1);
    ok = ok && cudaMalloc((void**)&p->Lm, sizeof(double) * (size_t)(p->np + 1) * std::max(p->np, 1)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->xsol, sizeof(double) * std::max(p->np, kNB)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->flags, sizeof(int) * nblk) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->gbar, 64) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->tctl, sizeof(unsigned) * tl_ctl_words((p->np + kTS - 1) / kTS)) == cudaSuccess;
    if (const char* tlv = getenv("M3S_GN_TIMELINE")) {
        if (tlv[0] == '1') {
            ok = ok && cudaMalloc((void**)&p->timeline, sizeof(long long) * 24 * nblk) == cudaSuccess &&
                 cudaMemset(p->timeline, 0, sizeof(long long) * 24 * nblk) == cudaSuccess;
        }
    }
    ok = ok && cudaMalloc((void**)&p->g, sizeof(double) * std::max(p->dim, 1)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->dx, sizeof(double) * std::max(p->dim, 1)) == cudaSuccess;
    ok = ok && cudaMalloc((void**)&p->status, sizeof(Status)) == cudaSuccess;
    if (ok) {
        ok = cudaMemset(p->status, 0, sizeof(Status)) == cudaSuccess &&
             cudaMemset(p->edge_blocks, 0, sizeof(double) * kEdgeStride * (size_t)std::max(E, 1)) == cudaSuccess &&
             cudaMemset(p->dx, 0, sizeof(double) * std::max(p->dim, 1)) == cudaSuccess &&
             cudaMemset(p->g, 0, sizeof(double) * std::max(p->dim, 1)) == cudaSuccess;
    }
    if (ok && p->F > 0) ok = build_sparse_front(p, off_pairs);
    if (!ok) {
        cuda_fail(cudaGetLastError(), "gn_plan_create allocation");
        plan_free(p);
        return nullptr;
    }
    return p;
}

M3S_API void m3s_gn_plan_destroy(m3s_gn_plan* plan) { plan_free(plan); }
M3S_API int m3s_gn_plan_num_free(const m3s_gn_plan* plan) { return plan ? plan->F : 0; }
M3S_API int m3s_gn_plan_dim(const m3s_gn_plan* plan) { return plan ? plan->dim : 0; }
M3S_API int m3s_gn_plan_ld(const m3s_gn_plan* plan) { return plan ? plan->np : 0; }
M3S_API double* m3s_gn_plan_edge_blocks(m3s_gn_plan* plan, int64_t* count_out) {
    if (!plan) return nullptr;
    if (count_out) *count_out = (int64_t)plan->E * kEdgeStride;
    return plan->edge_blocks;
}
M3S_API double* m3s_gn_plan_H(m3s_gn_plan* plan) { return plan ? plan->Hm : nullptr; }
M3S_API double* m3s_gn_plan_g(m3s_gn_plan* plan) { return plan ? plan->g : nullptr; }
M3S_API double* m3s_gn_plan_dx(m3s_gn_plan* plan) { return plan ? plan->dx : nullptr; }
M3S_API double* m3s_gn_plan_poses(m3s_gn_plan* plan) { return plan ? plan->poses : nullptr; }
M3S_API long long* m3s_gn_plan_timeline(m3s_gn_plan* plan, int64_t* count_out) {
    if (count_out) *count_out = (plan && plan->timeline) ? 24LL * std::max(plan->np / kNB, 1) : 0;
    return plan ? plan->timeline : nullptr;
}

M3S_API int m3s_gn_set_poses(m3s_gn_plan* plan, const float* Twc, m3s_stream_t stream) {
    M3S_REQUIRE(plan && Twc, "gn_set_poses: null pointer");
    const int n = 8 * plan->num_kf;
    if (n == 0) return M3S_OK;
    gn_poses_f32_to_f64<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(Twc, plan->poses, n, plan->status);
    return launch_ok("gn_poses_f32_to_f64");
}

M3S_API int m3s_gn_get_poses(m3s_gn_plan* plan, float* Twc_out, m3s_stream_t stream) {
    M3S_REQUIRE(plan && Twc_out, "gn_get_poses: null pointer");
    const int n = 8 * plan->num_kf;
    if (n == 0) return M3S_OK;
    gn_poses_f64_to_f32<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(plan->poses, Twc_out, n);
    return launch_ok("gn_poses_f64_to_f32");
}

// M3S_GN_PREC_DEFAULT -> MIXED unless M3S_GN_PRECISION=f64 (read once)
// accumulate kernel of the mixed pass: M3S_GN_ACC_IMPL = stream (default: bulk-copy ring,
// gn_accum_stream_kernel) | regs (register-pipelined gn_accum_mixed_kernel).  Measured on B200,
// config 5: stream 1.256 ms, regs 1.38 ms (packed) / 1.34 ms (-DM3S_ACCF_SCALAR); a 256-thread /
// 2-CTA shape of the stream kernel spills at 128 registers and takes 1.82 ms.
static int acc_impl() {
    static const int v = [] {
        const char* e = getenv("M3S_GN_ACC_IMPL");
        if (e && !strcmp(e, "regs")) return 0;
        if (e && !strcmp(e, "slim")) return 2;
        return 1;
    }();
    return v;
}

static int resolve_precision(int requested) {
    if (requested == M3S_GN_PREC_F64 || requested == M3S_GN_PREC_MIXED) return requested;
    static const int env_default = [] {
        const char* v = getenv("M3S_GN_PRECISION");
        return (v && (!strcmp(v, "f64") || !strcmp(v, "F64"))) ? M3S_GN_PREC_F64 : M3S_GN_PREC_MIXED;
    }();
    return env_default;
}

static KernelParams make_kernel_params(const m3s_gn_params* gp) {
    KernelParams P{};
    P.mode = gp->mode;
    P.c_th = (float)gp->C_thresh;
    P.q_th = (float)gp->Q_thresh;
    if (gp->mode == M3S_GN_CALIB) {
        P.fx = gp->K4[0]; P.fy = gp->K4[1]; P.cx = gp->K4[2]; P.cy = gp->K4[3];
        P.z_eps = gp->z_eps;
        P.sig_px_inv = 1.0 / gp->sigma_a;
        P.sig_z_inv = 1.0 / gp->sigma_b;
        P.u_lo = gp->pixel_border; P.u_hi = gp->img_w - gp->pixel_border;
        P.v_lo = gp->pixel_border; P.v_hi = gp->img_h - gp->pixel_border;
    } else {
        P.sig_inv_f = (float)(1.0 / gp->sigma_a);
    }
    return P;
}

M3S_API int m3s_gn_accumulate(m3s_gn_plan* plan, const m3s_gn_params* gp, const float* Xs,
                              const float* Cs, const int32_t* idx, const uint8_t* valid,
                              const float* Q, int edge_begin, int edge_end, m3s_stream_t stream) {
    M3S_REQUIRE(plan && gp, "gn_accumulate: null pointer");
    M3S_REQUIRE(gp->mode >= 0 && gp->mode <= 2, "gn_accumulate: bad mode");
    M3S_REQUIRE(gp->precision >= 0 && gp->precision <= 2, "gn_accumulate: bad precision");
    M3S_REQUIRE(edge_begin >= 0 && edge_end <= plan->E && edge_begin <= edge_end, "gn_accumulate: bad edge range");
    if (plan->F == 0 || plan->E == 0) return M3S_OK;
    M3S_REQUIRE(Xs && Cs && idx && valid && Q, "gn_accumulate: null pointer");
    cudaStream_t st = as_stream(stream);
    const KernelParams P = make_kernel_params(gp);
    const int ne = edge_end - edge_begin;
    if (ne > 0 && plan->N > 0) {
        const unsigned grid = (unsigned)((long long)ne * plan->nchunks);
        const uintptr_t al = reinterpret_cast<uintptr_t>(Xs) | reinterpret_cast<uintptr_t>(Cs) |
                             reinterpret_cast<uintptr_t>(idx) | reinterpret_cast<uintptr_t>(Q);
        const int vec_ok = ((plan->N & 3) == 0) && ((al & 15) == 0) &&
                           ((reinterpret_cast<uintptr_t>(valid) & 3) == 0);
#define M3S_ACC(MODE)                                                                           \
    gn_accum_kernel<MODE><<<grid, kAccThreads, 0, st>>>(plan->status, P, plan->poses, plan->meta, \
                                                        Xs, Cs, idx, valid, Q, plan->N,          \
                                                        edge_begin, plan->nchunks,               \
                                                        plan->chunk_pts, vec_ok, plan->partials)
#define M3S_ACCF(MODE)                                                                          \
    gn_accum_mixed_kernel<MODE><<<grid, kAccThreads, 0, st>>>(plan->status, P, plan->poses,      \
                                                              plan->meta, Xs, Cs, idx, valid, Q, \
                                                              plan->N, edge_begin, plan->nchunks, \
                                                              plan->chunk_pts,                   \
                                                              vec_ok && plan->N < (1LL << 29),   \
                                                              plan->partials)
        // calib always takes the fp64 pass: its depth direction is weakly constrained (sigma_depth),
        // which amplifies the fp32 rounding of X' (4e-5 px) to a 1e-5 relative deviation of the
        // update (scripts/gn_precision_report.py), and its problems are 2-keyframe tracking solves
        // (config 4) where the pass takes microseconds either way
        // (Metal semantics: the per-point arithmetic of the reference's GPU kernel is fp32 -- the mixed pass)
        if ((resolve_precision(gp->precision) == M3S_GN_PREC_MIXED || gp->semantics == M3S_SEM_METAL) &&
            gp->mode != M3S_GN_CALIB) {
            // streamed kernel (bulk-copy ring): needs 16-point granularity everywhere
            const uintptr_t al_all = al | reinterpret_cast<uintptr_t>(valid);
            int impl = acc_impl();
            const bool stream_ok = ((plan->N & 15) == 0) && ((al_all & 15) == 0) &&
                                   plan->N < (1LL << 29) && (plan->chunk_pts % 512) == 0;   // whole 512-point stages
            if (!stream_ok) impl = 0;
#define M3S_ACCS(MODE, TPB, ST, MINB, SLIM)                                                          \
    do {                                                                                         \
        auto fn = gn_accum_stream_kernel<MODE, TPB, ST, MINB, SLIM>;                             \
        constexpr size_t sm = stream_smem_bytes<MODE, TPB, ST>();                                \
        /* per device, so not cached in a static: a process may drive several devices */        \
        M3S_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm)); \
        fn<<<grid, TPB, sm, st>>>(plan->status, P, plan->poses, plan->meta, Xs, Cs, idx, valid, Q, \
                                  plan->N, edge_begin, plan->nchunks, plan->chunk_pts,           \
                                  plan->partials);                                               \
    } while (0)
            if (impl == 1) {
                if (gp->mode == M3S_GN_RAYS) M3S_ACCS(M3S_GN_RAYS, 128, 5, 3, false);
                else M3S_ACCS(M3S_GN_POINTS, 128, 5, 3, false);
            } else if (impl == 2) {
                if (gp->mode == M3S_GN_RAYS) M3S_ACCS(M3S_GN_RAYS, 128, 4, 4, true);
                else M3S_ACCS(M3S_GN_POINTS, 128, 4, 4, true);
            } else if (gp->mode == M3S_GN_RAYS) M3S_ACCF(M3S_GN_RAYS);
            else M3S_ACCF(M3S_GN_POINTS);
#undef M3S_ACCS
        } else {
            if (gp->mode == M3S_GN_RAYS) M3S_ACC(M3S_GN_RAYS);
            else if (gp->mode == M3S_GN_POINTS) M3S_ACC(M3S_GN_POINTS);
            else M3S_ACC(M3S_GN_CALIB);
        }
#undef M3S_ACCF
#undef M3S_ACC
        if (int rc = launch_ok("gn_accum_kernel")) return rc;
    }
    PeerTable pt{};
    if (!plan->xch_attached) {
        gn_edge_finalize_kernel<false><<<plan->E, 64, 0, st>>>(
            plan->status, gp->mode, plan->poses, plan->meta, plan->partials,
            (plan->N > 0) ? plan->nchunks : 0, edge_begin, edge_end, plan->edge_blocks, pt);
        return launch_ok("gn_edge_finalize_kernel");
    }
    // peer exchange: push the owned blocks to every rank, then wait for everybody's
    const unsigned long long epoch = ++plan->xch_epoch;
    const size_t rb = xch_recv_bytes(plan);
    for (int r = 0; r < plan->xch_world; ++r) {
        pt.recv[r] = reinterpret_cast<double*>(plan->xch_peer[r] + (epoch & 1) * rb);
        pt.flags[r] = reinterpret_cast<unsigned long long*>(plan->xch_peer[r] + 2 * rb);
    }
    pt.ticket = reinterpret_cast<unsigned int*>(plan->xch_local + 2 * rb + sizeof(unsigned long long) * kMaxPeers);
    pt.epoch = epoch;
    pt.rank = plan->xch_rank;
    pt.world = plan->xch_world;
    if (ne > 0) {
        gn_edge_finalize_kernel<true><<<ne, 64, 0, st>>>(
            plan->status, gp->mode, plan->poses, plan->meta, plan->partials,
            (plan->N > 0) ? plan->nchunks : 0, edge_begin, edge_end, plan->edge_blocks, pt);
        if (int rc = launch_ok("gn_edge_finalize_kernel")) return rc;
    }
    const long long count = (long long)kEdgeStride * plan->E;
    const int wgrid = (int)std::min<long long>((count + 255) / 256, 64);
    gn_exchange_wait_kernel<<<wgrid, 256, 0, st>>>(
        plan->status, reinterpret_cast<const unsigned long long*>(plan->xch_local + 2 * rb),
        reinterpret_cast<const double*>(plan->xch_local + (epoch & 1) * rb), plan->edge_blocks, count,
        pt, ne == 0 ? 1 : 0, plan->xch_timeout_ns);
    return launch_ok("gn_exchange_wait_kernel");
}

// ---- peer exchange setup (CUDA IPC; host plumbing exchanges the 64-byte handles) ---------------
M3S_API int m3s_gn_exchange_create(m3s_gn_plan* plan, int rank, int world, void* ipc_handle_out) {
    M3S_REQUIRE(plan && ipc_handle_out, "gn_exchange_create: null pointer");
    M3S_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "gn_exchange_create: bad rank/world");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    xch_free(plan);
    M3S_CUDA(cudaSetDevice(plan->device));
    M3S_CUDA(cudaMalloc(reinterpret_cast<void**>(&plan->xch_local), xch_total_bytes(plan)));
    M3S_CUDA(cudaMemset(plan->xch_local, 0, xch_total_bytes(plan)));
    cudaIpcMemHandle_t h;
    M3S_CUDA(cudaIpcGetMemHandle(&h, plan->xch_local));
    memcpy(ipc_handle_out, &h, sizeof(h));
    plan->xch_exported = true;
    plan->xch_released = false;
    if (const char* e = getenv("M3S_GN_XCH_TIMEOUT_MS")) {
        const long long ms = atoll(e);
        plan->xch_timeout_ns = ms <= 0 ? 0ull : (unsigned long long)ms * 1000000ull;
    }
    plan->xch_rank = rank;
    plan->xch_world = world;
    plan->xch_epoch = 0;
    return M3S_OK;
}

M3S_API int m3s_gn_exchange_attach(m3s_gn_plan* plan, const void* ipc_handles) {
    M3S_REQUIRE(plan && ipc_handles, "gn_exchange_attach: null pointer");
    M3S_REQUIRE(plan->xch_local && plan->xch_world >= 1, "gn_exchange_attach: call m3s_gn_exchange_create first");
    M3S_CUDA(cudaSetDevice(plan->device));
    for (int r = 0; r < plan->xch_world; ++r) {
        if (r == plan->xch_rank) { plan->xch_peer[r] = plan->xch_local; continue; }
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char*>(ipc_handles) + (size_t)r * sizeof(h), sizeof(h));
        void* ptr = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("gn_exchange_attach: cudaIpcOpenMemHandle(rank %d): %s", r, cudaGetErrorString(e));
            for (int q = 0; q < r; ++q)
                if (q != plan->xch_rank && plan->xch_peer[q]) { cudaIpcCloseMemHandle(plan->xch_peer[q]); plan->xch_peer[q] = nullptr; }
            return M3S_ERR_CUDA;
        }
        plan->xch_peer[r] = static_cast<char*>(ptr);
    }
    plan->xch_attached = plan->xch_world > 1;
    return M3S_OK;
}

M3S_API int m3s_gn_exchange_detach(m3s_gn_plan* plan) {
    M3S_REQUIRE(plan, "gn_exchange_detach: null plan");
    M3S_CUDA(cudaSetDevice(plan->device));
    M3S_CUDA(cudaDeviceSynchronize());
    xch_unmap(plan);  // the local region stays allocated until the plan dies: peers may still map it
    return M3S_OK;
}

M3S_API int m3s_gn_exchange_release(m3s_gn_plan* plan) {
    M3S_REQUIRE(plan, "gn_exchange_release: null plan");
    M3S_REQUIRE(!plan->xch_attached, "gn_exchange_release: detach first");
    plan->xch_released = true;
    return M3S_OK;
}

M3S_API int m3s_gn_assemble(m3s_gn_plan* plan, m3s_stream_t stream) {
    M3S_REQUIRE(plan, "gn_assemble: null plan");
    if (plan->F == 0) return M3S_OK;
    cudaStream_t st = as_stream(stream);
    // gn_assemble_kernel REWRITES every structurally non-zero block (and the rhs row entries of the free
    // keyframes) on every call, so the 15.9 MB of Hm only have to be zeroed when something else wrote into
    // them: the launch-per-step and the persistent factorisations update Hm in place, the dataflow
    // factorisation keeps its tiles in shared memory and leaves Hm alone (hm_clean stays true).
    if (!plan->hm_clean) {
        M3S_CUDA(cudaMemsetAsync(plan->Hm, 0, sizeof(double) * (size_t)(plan->np + 1) * plan->np, st));
        plan->hm_clean = true;
    }
    gn_assemble_kernel<<<plan->nblocks_H, 64, 0, st>>>(plan->status, plan->blocks, plan->list_edge,
                                                       plan->list_sign, plan->edge_blocks, plan->Hm,
                                                       plan->np, plan->dim, plan->g, plan->flags,
                                                       std::max(plan->np / kNB, 1));
    if (int rc = launch_ok("gn_assemble_kernel")) return rc;
    if (plan->np > plan->dim && !plan->sp_on) {   // (the sparse front reads keyframe blocks only)
        gn_pad_diag_kernel<<<1, 32, 0, st>>>(plan->status, plan->Hm, plan->np, plan->dim);
        if (int rc = launch_ok("gn_pad_diag_kernel")) return rc;
    }
    return M3S_OK;
}

// M3S_GN_PDL=1 launches the block steps with programmatic stream serialisation (the kernel's
// griddepcontrol.wait orders them).  MEASURED on B200 (config 5): solve 0.635 ms with, 0.606 ms
// without -- the early-resident CTAs of later steps cost more than the launch gaps they hide; off by
// default, kept for A/B.
static bool use_pdl() {
    static const bool v = [] {
        const char* e = getenv("M3S_GN_PDL");
        return e && e[0] == '1';
    }();
    return v;
}

// Grid of the persistent factorisation: <= 2 CTAs per SM, never more than the chain CTA + the tiles of
// step 0; 0 = use the launch-per-step path (the default, see below; also when the device cannot
// launch cooperatively).  M3S_GN_CHOL_CTAS overrides the count (clamped to what is co-resident).
static int fused_grid_size(m3s_gn_plan* plan) {
    if (plan->fused_grid >= 0) return plan->fused_grid;
    plan->fused_grid = 0;
    // MEASURED (B200, config 5, dim 1393 = 44 block steps, profiles/r2_gn_timeline_*.txt): persistent
    // kernel 13.3 us per block step (585 us), launch-per-step 11.4 us (502 us).  The chain CTA is
    // not what bounds the persistent version: a worker step is tile 8 us (two CTAs share an SM) +
    // worker barrier / flag hops 5 us, and the kernel boundary of the step path is the cheaper
    // barrier (2.4 us).  Opt-in (M3S_GN_CHOL=fused) until the tiles are faster.
    const char* e = getenv("M3S_GN_CHOL");
    if (!(e && strcmp(e, "fused") == 0)) return 0;
    int coop = 0, sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, plan->device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, plan->device);
    if (!coop || sms <= 0 ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_fused_kernel, kFusedThreads, 0) != cudaSuccess ||
        per_sm < 1) {
        cudaGetLastError();
        return 0;
    }
    const int np = plan->np;
    const int trailing = np - kNB, rows_below = np + 1 - kNB;
    const int ntr = (rows_below + kSyrkTile - 1) / kSyrkTile, ntc = (trailing + kSyrkTile - 1) / kSyrkTile;
    const int ntiles0 = (ntc == 0) ? 1 : ntc * (ntc + 1) / 2 + ((ntr > ntc) ? ntc : 0);
    int g = std::min(sms * std::min(per_sm, 2), 1 + ntiles0);
    if (const char* e = getenv("M3S_GN_CHOL_CTAS")) {
        const int v = atoi(e);
        if (v >= 2) g = std::min(v, sms * per_sm);
    }
    plan->fused_grid = std::max(g, 2);
    return plan->fused_grid;
}

// Grid of the dataflow factorisation (chol_tiles_kernel): one CTA per diagonal tile + one per tile
// (r, c) with r - c >= 2; every CTA must be co-resident (cooperative launch, 2 CTAs per SM).  0 = use
// another path (M3S_GN_CHOL=steps|fused, no cooperative launch, or more tiles than resident CTAs).
static int tiles_grid_for(m3s_gn_plan* plan, int np) {
    int coop = 0, sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, plan->device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, plan->device);
    if (!coop || sms <= 0 ||
        cudaFuncSetAttribute(chol_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(TilesSmem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_tiles_kernel, kTlThreads,
                                                      sizeof(TilesSmem)) != cudaSuccess ||
        per_sm < 1) {
        cudaGetLastError();
        return 0;
    }
    const int nt = (np + kTS - 1) / kTS;
    const int grid = nt + nt * (nt - 1) / 2;
    return grid > sms * per_sm ? 0 : grid;
}

static int tiles_grid_size(m3s_gn_plan* plan) {
    if (plan->tiles_grid >= 0) return plan->tiles_grid;
    plan->tiles_grid = 0;
    const char* e = getenv("M3S_GN_CHOL");
    if (e && (strcmp(e, "steps") == 0 || strcmp(e, "fused") == 0)) return 0;
    int coop = 0, sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, plan->device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, plan->device);
    if (!coop || sms <= 0 ||
        cudaFuncSetAttribute(chol_tiles_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(TilesSmem)) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_tiles_kernel, kTlThreads,
                                                      sizeof(TilesSmem)) != cudaSuccess ||
        per_sm < 1) {
        cudaGetLastError();
        return 0;
    }
    const int nt = (plan->np + kTS - 1) / kTS;
    const int grid = nt + nt * (nt - 1) / 2;
    if (grid > sms * per_sm) return 0;
    plan->tiles_grid = grid;
    return grid;
}

// assemble + blocked factorisation of the augmented matrix, enqueued on `st`: one cooperative launch
// (chol_tiles_kernel, or chol_fused_kernel on request) or, as the fallback, one launch per block step
// sparse front: levels, Schur complement on the tail, dataflow factorisation of the tail.  Returns 1 when the
// tail's cooperative grid cannot be placed (the caller then factorises the whole system densely).
static int enqueue_sparse_front(m3s_gn_plan* plan, cudaStream_t st, bool capturing) {
    if (plan->sp_tiles_grid == 0) plan->sp_tiles_grid = tiles_grid_for(plan, plan->sp_npR);
    if (plan->sp_tiles_grid == 0 ||
        cudaFuncSetAttribute(sp_level_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSpLevelSmem) != cudaSuccess) {
        cudaGetLastError();
        plan->sp_on = false;
        return 1;
    }
    const int np = plan->np, npR = plan->sp_npR;
    const int levels = (int)plan->sp_level_begin.size() - 1;
    for (int l = 0; l < levels; ++l) {
        const int b = plan->sp_level_begin[l], n = plan->sp_level_begin[l + 1] - b;
        sp_level_kernel<<<n, kSpThreads, kSpLevelSmem, st>>>(plan->status, plan->sp, b, plan->Hm, np);
        if (int rc = launch_ok("sp_level_kernel")) return rc;
    }
    const int nt = (npR + kTS - 1) / kTS, grid = plan->sp_tiles_grid;
    sp_tail_kernel<<<plan->sp_ntail_blocks, 64, 0, st>>>(plan->status, plan->sp, plan->spt, plan->Hm, np,
                                                         plan->sp_Hr, npR, plan->tctl, tl_ctl_words(nt));
    if (int rc = launch_ok("sp_tail_kernel")) return rc;
    int first_diag = std::max(0, std::min(sm_count_current(), grid) - nt);
    void* args[] = {&plan->status, &plan->sp_Hr, &plan->Lm, &plan->Linv, const_cast<int*>(&npR),
                    &plan->tctl, &first_diag, &plan->timeline};
    const cudaError_t ce = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(chol_tiles_kernel), dim3(grid),
                                                       dim3(kTlThreads), args, sizeof(TilesSmem), st);
    if (ce == cudaSuccess) return launch_ok("chol_tiles_kernel");
    cudaGetLastError();
    if (capturing || (ce != cudaErrorCooperativeLaunchTooLarge && ce != cudaErrorLaunchOutOfResources))
        return cuda_fail(ce, "cudaLaunchCooperativeKernel(chol_tiles_kernel, tail)");
    plan->sp_on = false;   // the levels only wrote their own buffers: the dense paths start from Hm as assembled
    return 1;
}

static int enqueue_assemble_factor(m3s_gn_plan* plan, cudaStream_t st, bool capturing = false) {
    if (int rc = m3s_gn_assemble(plan, reinterpret_cast<m3s_stream_t>(st))) return rc;
    if (plan->sp_on) {
        const int rc = enqueue_sparse_front(plan, st, capturing);
        if (rc <= 0) return rc;
        if (plan->np > plan->dim) {   // dense fallback: m3s_gn_assemble skipped the padding for the sparse front
            gn_pad_diag_kernel<<<1, 32, 0, st>>>(plan->status, plan->Hm, plan->np, plan->dim);
            if (int rc2 = launch_ok("gn_pad_diag_kernel")) return rc2;
        }
    }
    const int np = plan->np;
    const int nblk = np / kNB;
    if (const int grid = tiles_grid_size(plan)) {
        const int nt = (np + kTS - 1) / kTS;
        M3S_CUDA(cudaMemsetAsync(plan->tctl, 0, sizeof(unsigned) * tl_ctl_words(nt), st));
        int first_diag = std::max(0, std::min(sm_count_current(), grid) - nt);
        void* args[] = {&plan->status, &plan->Hm, &plan->Lm, &plan->Linv, const_cast<int*>(&np),
                        &plan->tctl, &first_diag, &plan->timeline};
        const cudaError_t ce = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(chol_tiles_kernel), dim3(grid),
                                                           dim3(kTlThreads), args, sizeof(TilesSmem), st);
        if (ce == cudaSuccess) return launch_ok("chol_tiles_kernel");
        // the device cannot hold the grid right now (e.g. SMs partitioned away from this context): use the
        // launch-per-step path for this plan from here on -- another CUDA path, never a CPU one
        cudaGetLastError();
        // (while capturing, any failure just abandons the graph: the direct launch decides about the path)
        if (capturing || (ce != cudaErrorCooperativeLaunchTooLarge && ce != cudaErrorLaunchOutOfResources))
            return cuda_fail(ce, "cudaLaunchCooperativeKernel(chol_tiles_kernel)");
        plan->tiles_grid = 0;
    }
    plan->hm_clean = false;   // both paths below factorise Hm in place
    if (const int grid = fused_grid_size(plan)) {
        M3S_CUDA(cudaMemsetAsync(plan->gbar, 0, 64, st));
        void* args[] = {&plan->status, &plan->Hm, &plan->Lm, &plan->Ldiag, &plan->Linv,
                        const_cast<int*>(&np), &plan->gbar, &plan->timeline};
        M3S_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(chol_fused_kernel), dim3(grid),
                                             dim3(kFusedThreads), args, 0, st));
        return launch_ok("chol_fused_kernel");
    }
    chol_diag0_kernel<<<1, kStepThreads, 0, st>>>(plan->status, plan->Hm, plan->Lm, plan->Ldiag,
                                                   plan->Linv, np, plan->timeline);
    if (int rc = launch_ok("chol_diag0_kernel")) return rc;
    for (int kb = 0; kb < nblk; ++kb) {
        const int k0 = kb * kNB;
        const int rows_below = np + 1 - (k0 + kNB);  // incl. rhs row
        const int trailing = np - (k0 + kNB);
        const int ntr = (rows_below + kSyrkTile - 1) / kSyrkTile;
        const int ntc = (trailing + kSyrkTile - 1) / kSyrkTile;
        // trailing == 0 (last step): one tile that only solves the rhs row
        const int ntiles = (ntc == 0) ? 1 : ntc * (ntc + 1) / 2 + ((ntr > ntc) ? ntc : 0);
        if (use_pdl() && kb > 0) {
            // block step kb depends on step kb - 1 only: programmatic edge (also inside the captured
            // graph), the kernel's griddepcontrol.wait provides the ordering
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(2 + ntiles);
            cfg.blockDim = dim3(kStepThreads);
            cfg.dynamicSmemBytes = 0;
            cfg.stream = st;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at;
            cfg.numAttrs = 1;
            M3S_CUDA(cudaLaunchKernelEx(&cfg, chol_step_kernel, plan->status, plan->Hm, plan->Lm,
                                        plan->Ldiag, plan->Linv, np, kb, std::max(ntc, 1),
                                        plan->timeline));
        } else {
            chol_step_kernel<<<2 + ntiles, kStepThreads, 0, st>>>(plan->status, plan->Hm, plan->Lm,
                                                                 plan->Ldiag, plan->Linv, np, kb,
                                                                 std::max(ntc, 1), plan->timeline);
        }
        if (int rc = launch_ok("chol_step_kernel")) return rc;
    }
    return M3S_OK;
}

// Capture the sequence once on a private stream (so a caller that is itself capturing is not
// disturbed) and replay it with one cudaGraphLaunch per iteration: ~90 launches on the
// launch-per-step path, 5 stream operations incl. the cooperative launch on the dataflow path, whose
// host-side cost otherwise shows up as gaps on the device when the launching thread is slow (first
// process on a fresh box: 0.89 instead of 0.54 ms per solve).
// The zeroing of Hm must not depend on the state at capture time: on the dataflow path (Hm is never
// written by the factorisation) it is done once here, on the caller's stream, outside the graph; on
// the in-place paths it is forced into the graph.
static void build_solve_graph(m3s_gn_plan* plan, cudaStream_t st) {
    if (plan->sp_on || tiles_grid_size(plan) > 0) {   // neither path writes Hm
        if (!plan->hm_clean) {
            if (cudaMemsetAsync(plan->Hm, 0, sizeof(double) * (size_t)(plan->np + 1) * plan->np, st) != cudaSuccess) {
                plan->graph_failed = true; cudaGetLastError(); return;
            }
            plan->hm_clean = true;
        }
    } else {
        plan->hm_clean = false;
    }
    cudaStream_t cs = nullptr;
    if (cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking) != cudaSuccess) { plan->graph_failed = true; cudaGetLastError(); return; }
    bool ok = cudaStreamBeginCapture(cs, cudaStreamCaptureModeRelaxed) == cudaSuccess;
    if (ok) {
        const int rc = enqueue_assemble_factor(plan, cs, true);
        cudaGraph_t g = nullptr;
        const cudaError_t e = cudaStreamEndCapture(cs, &g);
        ok = (rc == M3S_OK) && (e == cudaSuccess) && g;
        if (ok) ok = cudaGraphInstantiate(&plan->graph_exec, g, 0) == cudaSuccess;
        if (ok) plan->graph = g;
        else if (g) cudaGraphDestroy(g);
    }
    if (!ok) { plan->graph_failed = true; plan->graph_exec = nullptr; cudaGetLastError(); }
    cudaStreamDestroy(cs);
}

M3S_API int m3s_gn_plan_factor_path(m3s_gn_plan* plan) {
    if (!plan || plan->F == 0) return 0;
    if (plan->sp_on) {   // 3 = sparse front + dataflow factorisation of the tail (if its grid can be placed)
        if (plan->sp_tiles_grid == 0) plan->sp_tiles_grid = tiles_grid_for(plan, plan->sp_npR);
        if (plan->sp_tiles_grid > 0) return 3;
        plan->sp_on = false;
    }
    if (tiles_grid_size(plan) > 0) return 2;
    return fused_grid_size(plan) > 0 ? 1 : 0;
}

M3S_API int m3s_gn_plan_sparse_info(const m3s_gn_plan* plan, int32_t* out) {
    M3S_REQUIRE(plan && out, "gn_plan_sparse_info: null pointer");
    out[0] = plan->sp_on ? (int)plan->sp_level_begin.size() - 1 : 0;
    out[1] = plan->sp_on ? plan->sp_nE : 0;
    out[2] = plan->sp_on ? plan->sp_nR : plan->F;
    out[3] = plan->sp_on ? plan->sp_npR : plan->np;
    return M3S_OK;
}

M3S_API int m3s_gn_solve_update(m3s_gn_plan* plan, const m3s_gn_params* gp, m3s_stream_t stream) {
    M3S_REQUIRE(plan && gp, "gn_solve_update: null pointer");
    if (plan->F == 0) return M3S_OK;
    cudaStream_t st = as_stream(stream);
    const int np = plan->np;
    // M3S_GN_GRAPH=0 launches directly (A/B, debugging)
    static const bool use_graph = [] { const char* e = getenv("M3S_GN_GRAPH"); return !(e && e[0] == '0'); }();
    if (use_graph && !plan->graph_exec && !plan->graph_failed &&
        (plan->sp_on || tiles_grid_size(plan) > 0 || (np > 2 * kNB && fused_grid_size(plan) == 0)))
        build_solve_graph(plan, st);
    bool launched = false;
    if (plan->graph_exec) {
        const cudaError_t ge = cudaGraphLaunch(plan->graph_exec, st);
        if (ge == cudaSuccess) {
            launched = true;
        } else {
            // e.g. the co-resident grid cannot be placed any more: drop the graph, the direct launch
            // below picks the path (and reports what it cannot recover from)
            cudaGetLastError();
            cudaGraphExecDestroy(plan->graph_exec);
            plan->graph_exec = nullptr;
            plan->graph_failed = true;
        }
    }
    if (!launched) {
        if (int rc = enqueue_assemble_factor(plan, st)) return rc;
    }
    const int metal = gp->semantics == M3S_SEM_METAL ? 1 : 0;
    if (plan->sp_on) {
        // dense back-substitution of the tail, then the levels in reverse, then the ||dx|| test + retraction
        const int npR = plan->sp_npR, nblkR = npR / kNB;
        gn_backsolve_group_kernel<<<(nblkR + kBackGroup - 1) / kBackGroup, kBackThreads, 0, st>>>(
            plan->status, plan->Lm, plan->Linv, npR, plan->sp_dimR, plan->F, plan->free_kf, gp->delta_thresh,
            plan->xsol, plan->flags, 1, plan->dx, plan->poses, plan->timeline, metal, 0);
        if (int rc = launch_ok("gn_backsolve_group_kernel")) return rc;
        sp_back_finish_kernel<<<1, 1024, 0, st>>>(plan->status, plan->sp, plan->sp_level_begin_d,
                                                  (int)plan->sp_level_begin.size() - 1, plan->spt.rnode, plan->sp_nR,
                                                  plan->xsol, plan->dx, plan->dim, plan->F, plan->free_kf,
                                                  gp->delta_thresh, plan->poses, metal);
        return launch_ok("sp_back_finish_kernel");
    }
    // M3S_GN_BACKSOLVE=block: one CTA per 32-column block (round 1); read per call (A/B, tests)
    const char* bse = getenv("M3S_GN_BACKSOLVE");
    const bool per_block = bse && strcmp(bse, "block") == 0;
    const int nblk = np / kNB;
    if (per_block)
        gn_backsolve_update_kernel<<<nblk, kBackThreads, 0, st>>>(
            plan->status, plan->Lm, plan->Linv, np, plan->dim, plan->F, plan->free_kf, gp->delta_thresh,
            plan->xsol, plan->flags, 1, plan->dx, plan->poses, plan->timeline,
            metal);
    else
        gn_backsolve_group_kernel<<<(nblk + kBackGroup - 1) / kBackGroup, kBackThreads, 0, st>>>(
            plan->status, plan->Lm, plan->Linv, np, plan->dim, plan->F, plan->free_kf, gp->delta_thresh,
            plan->xsol, plan->flags, 1, plan->dx, plan->poses, plan->timeline, metal, 1);
    return launch_ok("gn_backsolve_update_kernel");
}

M3S_API int m3s_gn_status(m3s_gn_plan* plan, int32_t* status_host, double* last_dx_norm_host,
                          m3s_stream_t stream) {
    M3S_REQUIRE(plan, "gn_status: null plan");
    Status s;
    M3S_CUDA(cudaMemcpyAsync(&s, plan->status, sizeof(Status), cudaMemcpyDeviceToHost, as_stream(stream)));
    M3S_CUDA(cudaStreamSynchronize(as_stream(stream)));
    if (status_host) {
        status_host[0] = s.applied; status_host[1] = s.stopped;
        status_host[2] = s.failed ? (s.xch_timeout ? 2 : 1) : 0;
        status_host[3] = s.attempted;
    }
    if (last_dx_norm_host) *last_dx_norm_host = s.last_norm;
    return M3S_OK;
}

M3S_API int m3s_gauss_newton(m3s_gn_plan* plan, const m3s_gn_params* gp, const float* Twc,
                             const float* Xs, const float* Cs, const int32_t* idx,
                             const uint8_t* valid, const float* Q, float* Twc_out,
                             m3s_stream_t stream) {
    M3S_REQUIRE(plan && gp && Twc && Twc_out, "gauss_newton: null pointer");
    if (int rc = m3s_gn_set_poses(plan, Twc, stream)) return rc;
    if (plan->F > 0) {
        for (int it = 0; it < gp->max_iter; ++it) {
            if (int rc = m3s_gn_accumulate(plan, gp, Xs, Cs, idx, valid, Q, 0, plan->E, stream)) return rc;
            if (int rc = m3s_gn_solve_update(plan, gp, stream)) return rc;
        }
    }
    return m3s_gn_get_poses(plan, Twc_out, stream);
}
