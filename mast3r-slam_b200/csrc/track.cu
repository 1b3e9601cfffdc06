// Tracker single-pose solves: the frame-to-keyframe Sim3 Gauss-Newton of
// FrameTracker (tracker.py:216-406), the direct caller of matching.match.
//
//   m3s_track_gate       tracker.py:86-113  Qk = sqrt(Qff[idx] * Qkf), valid_opt / valid_kf masks, counts
//   m3s_track_opt_pose   tracker.py:255-313 (_opt_pose_ray_dist_sim3: 4 residuals, ray + distance,
//                        geometry.py:46-92) and :315-406 (_opt_pose_calib_sim3: pixel + log-depth,
//                        geometry.py:155-227), each with _solve (:216-253), act_Sim3 Jacobian
//                        (geometry.py:95-137), Sim3.retr = T * exp(tau) (liegroups/sim3.py:106-154,
//                        253-262) and check_convergence (optimizer.py:11-46).
//
// One problem = one frame pair: N points, 7 unknowns, ~6 MB and ~30 MFLOP per iteration: the solve is
// latency-bound, so the WHOLE loop is ONE cooperative launch (`track_persistent_kernel`, at most
// 2 CTAs per SM, all co-resident) with two grid-wide barriers per iteration and no host
// synchronisation: every CTA accumulates its grid-stride share of the points (thread -> warp shuffle
// -> shared -> one 36-double partial per CTA: 28 upper-triangular entries of J^T W J, 7 of -J^T W r,
// cost); barrier; CTA 0 sums the partials in CTA order, factorises the 7x7 H + 1e-6 I (Cholesky),
// retracts, tests convergence and writes the outputs; barrier; a `done` flag in global memory
// replays the reference's `break`.
// The reference evaluates everything in MLX fp32; here the state, the per-point arithmetic and the
// reductions are fp64 (N is small: ~6 MB and ~30 MFLOP per iteration, the solve is latency-bound),
// without atomics, so results are bitwise reproducible.
#include <math.h>

#include <algorithm>

#include <cooperative_groups.h>

#include "m3s_common.cuh"

namespace cg = cooperative_groups;

namespace m3s {

constexpr int kTrkThreads = 256;
constexpr int kTrkAcc = 36;  // 28 (H upper) + 7 (g) + 1 (cost)

struct TrackState {
    double T[8];    // T_CkCf (t, q_xyzw, s)
    double Tk[8];   // T_WCk
    double old_cost, cost, dnorm;
    int iters, done, failed, pad;
};

struct TrackConst {
    int mode, max_iters, img_w, img_h, pixel_border, nblocks;
    double si_a, si_b;  // 1/sigma_ray, 1/sigma_dist   |  1/sigma_pixel, 1/sigma_depth
    double huber, rel_error, delta_norm, depth_eps;
    double fx, fy, cx, cy;
};

// ---- Sim3 helpers (liegroups/so3.py:125-167, sim3.py:195-231) ---------------------------------
__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
// SO3.act: p + qw * t + qxyz x t, t = 2 (qxyz x p)
__device__ __forceinline__ void quat_act(const double* q, const double* p, double* o) {
    double t[3], c[3];
    cross3(q, p, t);
    t[0] *= 2.0; t[1] *= 2.0; t[2] *= 2.0;
    cross3(q, t, c);
    o[0] = p[0] + q[3] * t[0] + c[0];
    o[1] = p[1] + q[3] * t[1] + c[1];
    o[2] = p[2] + q[3] * t[2] + c[2];
}
__device__ __forceinline__ void quat_mul(const double* a, const double* b, double* o) {
    const double x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3], x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
    o[0] = w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2;
    o[1] = w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2;
    o[2] = w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2;
    o[3] = w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2;
}
// Sim3.__mul__: (t1 + s1 R1 t2, q1 q2, s1 s2)
__device__ __forceinline__ void sim3_mul(const double* a, const double* b, double* o) {
    double r[3];
    quat_act(a + 3, b, r);
    o[0] = a[0] + a[7] * r[0];
    o[1] = a[1] + a[7] * r[1];
    o[2] = a[2] + a[7] * r[2];
    quat_mul(a + 3, b + 3, o + 3);
    o[7] = a[7] * b[7];
}
// Sim3.inv: s_inv = 1 / (s + 1e-10), t_inv = -s_inv R^-1 t
__device__ __forceinline__ void sim3_inv(const double* a, double* o) {
    const double qi[4] = {-a[3], -a[4], -a[5], a[6]};
    const double si = 1.0 / (a[7] + 1e-10);
    double r[3];
    quat_act(qi, a, r);
    o[0] = -si * r[0]; o[1] = -si * r[1]; o[2] = -si * r[2];
    o[3] = qi[0]; o[4] = qi[1]; o[5] = qi[2]; o[6] = qi[3];
    o[7] = si;
}
// Sim3.exp (sim3.py:106-154: SE3-like translation, scale = exp(sigma)) with SO3.exp (so3.py:65-96)
__device__ __forceinline__ void sim3_exp(const double* tau, double* o) {
    const double* v = tau;
    const double* w = tau + 3;
    const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    const double th = sqrt(th2 + 1e-10);
    const bool small = th2 < 1e-8;
    const double A = small ? 1.0 - th2 / 6.0 : sin(th) / th;
    const double B = small ? 0.5 - th2 / 24.0 : (1.0 - cos(th)) / th2;
    const double Cc = small ? 1.0 / 6.0 - th2 / 120.0 : (1.0 - A) / th2;
    double wv[3], wwv[3];
    cross3(w, v, wv);
    cross3(w, wv, wwv);
    for (int k = 0; k < 3; ++k) o[k] = v[k] + B * wv[k] + Cc * wwv[k];
    const double half = 0.5 * th;
    const double sh = small ? 0.5 - th2 / 48.0 : sin(half) / th;
    const double ch = small ? 1.0 - th2 / 8.0 : cos(half);
    o[3] = sh * w[0]; o[4] = sh * w[1]; o[5] = sh * w[2]; o[6] = ch;
    o[7] = exp(tau[6]);
}

__device__ __forceinline__ void write_outputs(const TrackState* st, float* T_WCf_out,
                                              float* T_CkCf_out) {
    double Tw[8];
    sim3_mul(st->Tk, st->T, Tw);  // T_WCf = T_WCk * T_CkCf (tracker.py:312)
    for (int k = 0; k < 8; ++k) {
        if (T_WCf_out) T_WCf_out[k] = (float)Tw[k];
        if (T_CkCf_out) T_CkCf_out[k] = (float)st->T[k];
    }
}

__device__ __forceinline__ void track_init_body(TrackState* st, const float* __restrict__ T_WCf,
                                                const float* __restrict__ T_WCk, float* T_WCf_out,
                                                float* T_CkCf_out, double* stats, int max_iters) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double Tf[8], Ti[8];
    for (int k = 0; k < 8; ++k) { Tf[k] = (double)T_WCf[k]; st->Tk[k] = (double)T_WCk[k]; }
    sim3_inv(st->Tk, Ti);
    sim3_mul(Ti, Tf, st->T);  // T_CkCf = T_WCk^-1 * T_WCf (tracker.py:280)
    st->old_cost = INFINITY;
    st->cost = 0.0;
    st->dnorm = 0.0;
    st->iters = 0;
    st->failed = 0;
    st->done = (max_iters <= 0);
    write_outputs(st, T_WCf_out, T_CkCf_out);
    if (stats) { stats[0] = 0.0; stats[1] = 0.0; stats[2] = 0.0; stats[3] = 0.0; }
}

// one weighted residual row: a = robust * J_c (7), b = robust * r_c  (tracker.py:232-243)
__device__ __forceinline__ void add_row(double (&acc)[kTrkAcc], const double (&J)[7], double r,
                                        double si, double huber) {
    const double wr = fabs(si * r);
    const double hw = (wr < huber) ? 1.0 : huber / wr;  // optimizer.py:49-62
    const double rs = si * sqrt(hw);
    double a[7];
#pragma unroll
    for (int i = 0; i < 7; ++i) a[i] = rs * J[i];
    const double b = rs * r;
    int k = 0;
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
        for (int j = i; j < 7; ++j) acc[k++] += a[i] * a[j];
#pragma unroll
    for (int i = 0; i < 7; ++i) acc[28 + i] -= a[i] * b;  // g = -A^T b
    acc[35] += 0.5 * b * b;
}

// J_c = -(d[0..2] . [I | -skew(pW) | pW]) for one row d of the 4x3 (3x3) measurement Jacobian
__device__ __forceinline__ void chain_row(const double* d, const double* p, double (&J)[7]) {
    J[0] = -d[0];
    J[1] = -d[1];
    J[2] = -d[2];
    // -skew(p) = [[0, pz, -py], [-pz, 0, px], [py, -px, 0]]  (geometry.py:124-126)
    J[3] = -(d[1] * -p[2] + d[2] * p[1]);
    J[4] = -(d[0] * p[2] + d[2] * -p[0]);
    J[5] = -(d[0] * -p[1] + d[1] * p[0]);
    J[6] = -(d[0] * p[0] + d[1] * p[1] + d[2] * p[2]);
}

// (state and partials are re-read through volatile / L2 loads: they change inside the launch)
template <int MODE>
__device__ __forceinline__ void
track_accum_body(const TrackState* st, const TrackConst& c, const float* __restrict__ Xf,
                 const int32_t* __restrict__ idx, const float* __restrict__ Xk,
                 const float* __restrict__ Qk, const uint8_t* __restrict__ valid, int N,
                 double* partial) {
    double T[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) T[k] = reinterpret_cast<const volatile double*>(st->T)[k];
    double acc[kTrkAcc];
#pragma unroll
    for (int k = 0; k < kTrkAcc; ++k) acc[k] = 0.0;

    // four points per trip (same per-thread order as a plain grid-stride loop): all the index /
    // weight loads first, then the eight gathers, then the arithmetic -- with ~5 points per thread
    // the loop is pure load latency otherwise
    const int stride = gridDim.x * kTrkThreads;
    for (int n0 = blockIdx.x * kTrkThreads + threadIdx.x; n0 < N; n0 += 4 * stride) {
      bool okp[4];
      float qf[4];
      int mi[4];
      float xf[4][3], xk[4][3];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
          const int n = n0 + u * stride;
          const bool in = n < N;
          const int nn = in ? n : 0;
          qf[u] = __ldg(Qk + nn);
          mi[u] = idx ? __ldg(idx + nn) : nn;
          okp[u] = in && (__ldg(valid + nn) != 0);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
          const int n = n0 + u * stride;
          okp[u] = okp[u] && (qf[u] > 0.0f) && (mi[u] >= 0) && (mi[u] < N);  // sqrt_info = 0: nothing
          const size_t m = okp[u] ? (size_t)mi[u] : 0, nn = okp[u] ? (size_t)n : 0;
#pragma unroll
          for (int k = 0; k < 3; ++k) { xf[u][k] = __ldg(Xf + 3 * m + k); xk[u][k] = __ldg(Xk + 3 * nn + k); }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (!okp[u]) continue;
        const int n = n0 + u * stride;
        const double sq = sqrt((double)qf[u]);
        const double X[3] = {(double)xf[u][0], (double)xf[u][1], (double)xf[u][2]};
        const double Y[3] = {(double)xk[u][0], (double)xk[u][1], (double)xk[u][2]};
        double p[3];
        quat_act(T + 3, X, p);
#pragma unroll
        for (int k = 0; k < 3; ++k) p[k] = T[7] * p[k] + T[k];  // Sim3.act
        double J[7];
        if (MODE == M3S_TRACK_RAY_DIST) {
            // point_to_ray_dist (geometry.py:46-92)
            const double d = sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2] + 1e-10);
            const double di = 1.0 / d, di2 = di * di;
            const double rr[3] = {di * p[0], di * p[1], di * p[2]};
            const double dk = sqrt(Y[0] * Y[0] + Y[1] * Y[1] + Y[2] * Y[2] + 1e-10);
            const double dki = 1.0 / dk;
            const double si_ray = c.si_a * sq, si_dist = c.si_b * sq;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                double drow[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) drow[k] = di * ((r == k ? 1.0 : 0.0) - di2 * p[r] * p[k]);
                chain_row(drow, p, J);
                add_row(acc, J, dki * Y[r] - rr[r], si_ray, c.huber);
            }
            chain_row(rr, p, J);
            add_row(acc, J, dk - d, si_dist, c.huber);
        } else {
            // project_calib (geometry.py:155-227); img_size = (h, w)
            const double zden = p[2] + 1e-10;
            const double u = (c.fx * p[0] + c.cx * p[2]) / zden, v = (c.fy * p[1] + c.cy * p[2]) / zden;
            const bool vproj = (u > c.pixel_border) && (u < c.img_w - 1 - c.pixel_border) &&
                               (v > c.pixel_border) && (v < c.img_h - 1 - c.pixel_border) &&
                               (p[2] > c.depth_eps);
            const bool vmeas = Y[2] > c.depth_eps;  // tracker.py:207
            if (!(vproj && vmeas)) continue;
            const double logz = log(p[2] + 1e-10);
            const double zi = 1.0 / zden;
            const double mu = (double)(n % c.img_w), mv = (double)(n / c.img_w);
            const double mz = log(Y[2] + 1e-10);
            const double si_px = c.si_a * sq, si_z = c.si_b * sq;
            const double d0[3] = {c.fx * zi, 0.0, -c.fx * p[0] * zi * zi};
            const double d1[3] = {0.0, c.fy * zi, -c.fy * p[1] * zi * zi};
            const double d2[3] = {0.0, 0.0, zi};
            chain_row(d0, p, J);
            add_row(acc, J, mu - u, si_px, c.huber);
            chain_row(d1, p, J);
            add_row(acc, J, mv - v, si_px, c.huber);
            chain_row(d2, p, J);
            add_row(acc, J, mz - logz, si_z, c.huber);
        }
      }
    }
    // thread -> warp (shuffle tree) -> block (fixed order over warps)
    __shared__ double sh[kTrkThreads / 32][kTrkAcc];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kTrkAcc; ++k) {
        double v = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) sh[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < kTrkAcc) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < kTrkThreads / 32; ++w) s += sh[w][threadIdx.x];
        partial[(size_t)blockIdx.x * kTrkAcc + threadIdx.x] = s;
    }
}

// One CTA.  Threads 0..35 sum the block partials in block order; thread 0 solves and updates.
__device__ __forceinline__ void
track_solve_body(TrackState* st, const TrackConst& c, const double* partial, float* T_WCf_out,
                 float* T_CkCf_out, double* stats) {
    // fixed-order sum of the CTA partials: thread (k, part) adds CTAs part, part + 7, ...; the 7
    // parts are then added in order (independent loads in flight instead of one serial chain)
    __shared__ double s[kTrkAcc];
    __shared__ double sp[kTrkAcc][7];
    if (threadIdx.x < kTrkAcc * 7) {
        const int k = threadIdx.x / 7, part = threadIdx.x % 7;
        double v = 0.0;
#pragma unroll 4
        for (int b = part; b < c.nblocks; b += 7) v += __ldcg(partial + (size_t)b * kTrkAcc + k);
        sp[k][part] = v;
    }
    __syncthreads();
    if (threadIdx.x < kTrkAcc) {
        double v = 0.0;
#pragma unroll
        for (int part = 0; part < 7; ++part) v += sp[threadIdx.x][part];
        s[threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x != 0) return;
    // H + 1e-6 I (linalg.py:43), lower Cholesky, two triangular solves: tau = H^-1 g.  Fully
    // unrolled (L stays in registers, one reciprocal per pivot): this serial stretch sits between two
    // grid barriers and was 60 % of an iteration with L in local memory.
    double L[7][7], tau[7], rinv[7];
    {
        int k = 0;
#pragma unroll
        for (int i = 0; i < 7; ++i)
#pragma unroll
            for (int j = i; j < 7; ++j) { L[j][i] = s[k]; ++k; }
    }
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 7; ++j) {
        double d = L[j][j] + 1e-6;
#pragma unroll
        for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
        ok = ok && (d > 0.0) && isfinite(d);
        d = sqrt(d);
        L[j][j] = d;
        rinv[j] = 1.0 / d;
#pragma unroll
        for (int i = j + 1; i < 7; ++i) {
            double v = L[i][j];
#pragma unroll
            for (int k = 0; k < j; ++k) v -= L[i][k] * L[j][k];
            L[i][j] = v * rinv[j];
        }
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) {
        double v = s[28 + i];
#pragma unroll
        for (int k = 0; k < i; ++k) v -= L[i][k] * tau[k];
        tau[i] = v * rinv[i];
    }
#pragma unroll
    for (int i = 6; i >= 0; --i) {
        double v = tau[i];
#pragma unroll
        for (int k = i + 1; k < 7; ++k) v -= L[k][i] * tau[k];
        tau[i] = v * rinv[i];
    }
#pragma unroll
    for (int i = 0; i < 7; ++i) ok = ok && isfinite(tau[i]);
    if (!ok) {  // reference: exception -> "Optimization failed", pose left as it was (tracker.py:140-142)
        st->failed = 1;
        st->done = 1;
        if (stats) stats[3] = 1.0;
        return;
    }
    double E[8], Tn[8];
    sim3_exp(tau, E);
    sim3_mul(st->T, E, Tn);  // retr: T * exp(tau) (sim3.py:253-262), applied BEFORE the test
    for (int k = 0; k < 8; ++k) st->T[k] = Tn[k];
    const double cost = s[35];
    double dn = 0.0;
    for (int i = 0; i < 7; ++i) dn += tau[i] * tau[i];
    dn = sqrt(dn);
    // check_convergence (optimizer.py:33-37); first step: inf/inf = NaN -> not converged
    const double rel = fabs((st->old_cost - cost) / (st->old_cost + 1e-10));
    const bool conv = (rel < c.rel_error) || (dn < c.delta_norm);
    st->old_cost = cost;
    st->cost = cost;
    st->dnorm = dn;
    st->iters += 1;
    if (conv || st->iters >= c.max_iters) st->done = 1;
    write_outputs(st, T_WCf_out, T_CkCf_out);
    if (stats) { stats[0] = (double)st->iters; stats[1] = cost; stats[2] = dn; stats[3] = 0.0; }
}

template <int MODE>
__global__ void __launch_bounds__(kTrkThreads)
track_persistent_kernel(TrackState* st, TrackConst c, const float* __restrict__ Xf,
                        const int32_t* __restrict__ idx, const float* __restrict__ Xk,
                        const float* __restrict__ Qk, const uint8_t* __restrict__ valid, int N,
                        double* partial, const float* __restrict__ T_WCf,
                        const float* __restrict__ T_WCk, float* T_WCf_out, float* T_CkCf_out,
                        double* stats) {
    cg::grid_group grid = cg::this_grid();
    track_init_body(st, T_WCf, T_WCk, T_WCf_out, T_CkCf_out, stats, c.max_iters);
    grid.sync();
    for (int it = 0; it < c.max_iters; ++it) {
        if (*reinterpret_cast<volatile int*>(&st->done)) break;  // same value in every CTA
        track_accum_body<MODE>(st, c, Xf, idx, Xk, Qk, valid, N, partial);
        grid.sync();
        if (blockIdx.x == 0) track_solve_body(st, c, partial, T_WCf_out, T_CkCf_out, stats);
        grid.sync();
    }
}

// tracker.py:86-113: combined match confidence and the two validity masks, with their counts.
__global__ void __launch_bounds__(256)
track_gate_kernel(const float* __restrict__ Cf, const float* __restrict__ Ck,
                  const float* __restrict__ Qff, const float* __restrict__ Qkf,
                  const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid_match,
                  float c_conf, float q_conf, float* __restrict__ Qk, uint8_t* __restrict__ valid_opt,
                  uint8_t* __restrict__ valid_kf, int32_t* __restrict__ counts, int N) {
    int n_opt = 0, n_kf = 0;
    for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < N; n += gridDim.x * blockDim.x) {
        const int m = __ldg(idx + n);
        const bool in = (m >= 0) && (m < N);
        const float q = __fsqrt_rn(__fmul_rn(in ? __ldg(Qff + m) : 0.0f, __ldg(Qkf + n)));
        const bool vm = in && (__ldg(valid_match + n) != 0);
        const bool vq = q > q_conf;
        const bool kf = vm && vq;
        const bool opt = kf && (__ldg(Cf + m) > c_conf) && (__ldg(Ck + n) > c_conf);
        Qk[n] = q;
        valid_opt[n] = opt;
        valid_kf[n] = kf;
        n_opt += opt;
        n_kf += kf;
    }
    n_opt = __reduce_add_sync(0xffffffffu, n_opt);
    n_kf = __reduce_add_sync(0xffffffffu, n_kf);
    if ((threadIdx.x & 31) == 0) {  // integer sums: order-independent
        if (n_opt) atomicAdd(counts, n_opt);
        if (n_kf) atomicAdd(counts + 1, n_kf);
    }
}

}  // namespace m3s

using namespace m3s;

static int track_blocks(int N) {
    const int want = (N + kTrkThreads - 1) / kTrkThreads;
    const int cap = 2 * sm_count_current();
    return want < 1 ? 1 : (want < cap ? want : cap);
}

M3S_API size_t m3s_track_workspace_bytes(int N) {
    if (N < 0) return 0;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    return 256 + (size_t)track_blocks(N) * kTrkAcc * sizeof(double);
}

M3S_API int m3s_track_gate(const float* Cf, const float* Ck, const float* Qff, const float* Qkf,
                           const int32_t* idx_f2k, const uint8_t* valid_match, float C_conf,
                           float Q_conf, float* Qk_out, uint8_t* valid_opt_out,
                           uint8_t* valid_kf_out, int32_t* counts_out, int N, m3s_stream_t stream) {
    M3S_REQUIRE(N >= 0, "track_gate: bad N");
    M3S_REQUIRE(counts_out, "track_gate: null counts");
    cudaStream_t sst = as_stream(stream);
    M3S_CUDA(cudaMemsetAsync(counts_out, 0, 2 * sizeof(int32_t), sst));
    if (N == 0) return M3S_OK;
    M3S_REQUIRE(Cf && Ck && Qff && Qkf && idx_f2k && valid_match && Qk_out && valid_opt_out && valid_kf_out,
                "track_gate: null pointer");
    const int grid = (N + 255) / 256 < 4 * sm_count_current() ? (N + 255) / 256 : 4 * sm_count_current();
    track_gate_kernel<<<grid, 256, 0, sst>>>(Cf, Ck, Qff, Qkf, idx_f2k, valid_match, C_conf, Q_conf,
                                            Qk_out, valid_opt_out, valid_kf_out, counts_out, N);
    return launch_ok("track_gate_kernel");
}

M3S_API int m3s_track_opt_pose(const m3s_track_params* prm, const float* Xf, const int32_t* idx_f2k,
                               const float* Xk, const float* Qk, const uint8_t* valid, int N,
                               const float* T_WCf, const float* T_WCk, float* T_WCf_out,
                               float* T_CkCf_out, double* stats_out, void* workspace,
                               size_t workspace_bytes, m3s_stream_t stream) {
    M3S_REQUIRE(prm, "track_opt_pose: null params");
    M3S_REQUIRE(prm->mode == M3S_TRACK_RAY_DIST || prm->mode == M3S_TRACK_CALIB, "track_opt_pose: bad mode");
    M3S_REQUIRE(N >= 0, "track_opt_pose: bad N");
    M3S_REQUIRE(T_WCf && T_WCk && T_WCf_out && T_CkCf_out && workspace, "track_opt_pose: null pointer");
    M3S_REQUIRE(N == 0 || (Xf && Xk && Qk && valid), "track_opt_pose: null point arrays");
    if (prm->mode == M3S_TRACK_CALIB)
        M3S_REQUIRE(prm->img_w > 0 && prm->img_h > 0 && (long long)prm->img_w * prm->img_h == N,
                    "track_opt_pose: calib mode needs img_w * img_h == N");
    const size_t need = m3s_track_workspace_bytes(N);
    if (need == 0) return cuda_fail(cudaGetLastError(), "cudaGetDevice");
    if (workspace_bytes < need) {
        set_error("track_opt_pose: workspace %zu < %zu bytes", workspace_bytes, need);
        return M3S_ERR_WORKSPACE;
    }
    cudaStream_t sst = as_stream(stream);
    TrackState* st = reinterpret_cast<TrackState*>(workspace);
    double* partial = reinterpret_cast<double*>(reinterpret_cast<char*>(workspace) + 256);
    TrackConst c;
    c.mode = prm->mode;
    c.max_iters = prm->max_iters;
    c.img_w = prm->img_w; c.img_h = prm->img_h; c.pixel_border = prm->pixel_border;
    c.nblocks = track_blocks(N);
    const bool rd = prm->mode == M3S_TRACK_RAY_DIST;
    c.si_a = 1.0 / (rd ? prm->sigma_ray : prm->sigma_pixel);
    c.si_b = 1.0 / (rd ? prm->sigma_dist : prm->sigma_depth);
    c.huber = prm->huber; c.rel_error = prm->rel_error; c.delta_norm = prm->delta_norm;
    c.depth_eps = prm->depth_eps;
    c.fx = prm->K4[0]; c.fy = prm->K4[1]; c.cx = prm->K4[2]; c.cy = prm->K4[3];
    // one cooperative launch: every CTA must be resident (<= 2 per SM by construction; clamp to what
    // the occupancy calculator allows for this kernel)
    const void* fn = rd ? (const void*)track_persistent_kernel<M3S_TRACK_RAY_DIST>
                        : (const void*)track_persistent_kernel<M3S_TRACK_CALIB>;
    int per_sm = 0;
    M3S_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kTrkThreads, 0));
    M3S_REQUIRE(per_sm >= 1, "track_opt_pose: kernel does not fit an SM");
    c.nblocks = std::min(c.nblocks, per_sm * sm_count_current());
    void* args[] = {(void*)&st, (void*)&c, (void*)&Xf, (void*)&idx_f2k, (void*)&Xk, (void*)&Qk,
                    (void*)&valid, (void*)&N, (void*)&partial, (void*)&T_WCf, (void*)&T_WCk,
                    (void*)&T_WCf_out, (void*)&T_CkCf_out, (void*)&stats_out};
    M3S_CUDA(cudaLaunchCooperativeKernel(fn, dim3(c.nblocks), dim3(kTrkThreads), args, 0, sst));
    return launch_ok("track_opt_pose");
}
