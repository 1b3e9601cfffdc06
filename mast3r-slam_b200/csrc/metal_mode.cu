// Metal-semantics mode: the behaviour of the reference's GPU path -- its Metal
// compute shaders -- instead of its numpy path, for callers that want to reproduce what the fork
// computes on Apple hardware:
//   * iter_proj      shaders/iter_proj.metal:9-165: all-fp32 bilinear blend, PER-THREAD early exit
//                    (:149-152), |det| < 1e-10 => zero step (:68-78), clamp bound f32(W-1) - 0.001f;
//   * refine_matches shaders/refine_matches.metal:160-215 driven by refine_metal_runner.py:136-164: the
//                    window is walked with step `dilation`, coarse to fine (dilation_max .. 1), IN PLACE
//                    (each level starts from the previous level's winner), sequential fp32 score sum,
//                    best score starts at -1e10f;
//   * Gauss-Newton   retraction of shaders/gauss_newton.metal:255-301 (fp32, dt = tau) -- in gn.cu.
// Metal leaves contraction / reassociation to its compiler; the semantics mirrored here are the ones
// the parity tests fix by compiling the unmodified shaders as C++: every operation
// separately rounded, dot = ((x*x' + y*y') + z*z'), clamp = min(max()).  Results are BIT-EXACT
// against that build.  This translation unit is built with -fmad=false on top of the explicit
// round-to-nearest intrinsics.
#include "m3s_common.cuh"

namespace m3s {

namespace {

__device__ __forceinline__ float mmax(float v, float lo) { return (v < lo) ? lo : v; }   // shim max
__device__ __forceinline__ float mmin(float a, float b) { return (b < a) ? b : a; }      // shim min
__device__ __forceinline__ float mclamp(float v, float lo, float hi) { return mmin(mmax(v, lo), hi); }
__device__ __forceinline__ float mdot3(float ax, float ay, float az, float bx, float by, float bz) {
    return __fadd_rn(__fadd_rn(__fmul_rn(ax, bx), __fmul_rn(ay, by)), __fmul_rn(az, bz));
}
// w00*a + w01*b + w10*c + w11*d, left to right (iter_proj.metal:42-67)
__device__ __forceinline__ float mblend(float w00, float w01, float w10, float w11, float a, float b,
                                        float c, float d) {
    return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w00, a), __fmul_rn(w01, b)), __fmul_rn(w10, c)),
                     __fmul_rn(w11, d));
}

constexpr int kMetalThreads = 128;

__global__ void __launch_bounds__(kMetalThreads)
iter_proj_metal_kernel(const float* __restrict__ rays, const float* __restrict__ pts,
                       const float* __restrict__ p_init, float* __restrict__ p_out,
                       uint8_t* __restrict__ valid_out, int H, int W, int N, long long total,
                       int max_iter, float lambda, float thresh) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / N);
    const float* __restrict__ img = rays + (size_t)b * H * W * 9;
    float u = __ldg(p_init + 2 * t), v = __ldg(p_init + 2 * t + 1);
    const float t0 = __ldg(pts + 3 * t), t1 = __ldg(pts + 3 * t + 1), t2 = __ldg(pts + 3 * t + 2);
    const float uhi = __fsub_rn((float)(W - 1), 0.001f), vhi = __fsub_rn((float)(H - 1), 0.001f);
    // A NaN position is undefined behaviour in the shader (int(floor(NaN))); here it ends the loop.
    bool alive = (u == u) && (v == v);
    for (int it = 0; it < max_iter && alive; ++it) {
        const float cu = mclamp(u, 0.0f, uhi), cv = mclamp(v, 0.0f, vhi);
        const float fu = floorf(cu), fv = floorf(cv);
        const int x0 = (int)fu, y0 = (int)fv;
        const int x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
        const float fx = __fsub_rn(cu, (float)x0), fy = __fsub_rn(cv, (float)y0);
        const float ofx = __fsub_rn(1.0f, fx), ofy = __fsub_rn(1.0f, fy);
        const float w00 = __fmul_rn(ofx, ofy), w01 = __fmul_rn(ofx, fy);
        const float w10 = __fmul_rn(fx, ofy), w11 = __fmul_rn(fx, fy);
        const float* __restrict__ r00 = img + ((size_t)y0 * W + x0) * 9;
        const float* __restrict__ r01 = img + ((size_t)y1 * W + x0) * 9;
        const float* __restrict__ r10 = img + ((size_t)y0 * W + x1) * 9;
        const float* __restrict__ r11 = img + ((size_t)y1 * W + x1) * 9;
        float s[9];
#pragma unroll
        for (int ch = 0; ch < 9; ++ch)
            s[ch] = mblend(w00, w01, w10, w11, __ldg(r00 + ch), __ldg(r01 + ch), __ldg(r10 + ch), __ldg(r11 + ch));
        const float e0 = __fsub_rn(s[0], t0), e1 = __fsub_rn(s[1], t1), e2 = __fsub_rn(s[2], t2);
        const float a = __fadd_rn(mdot3(s[3], s[4], s[5], s[3], s[4], s[5]), lambda);
        const float bq = mdot3(s[3], s[4], s[5], s[6], s[7], s[8]);
        const float d = __fadd_rn(mdot3(s[6], s[7], s[8], s[6], s[7], s[8]), lambda);
        const float j0 = mdot3(s[3], s[4], s[5], e0, e1, e2);
        const float j1 = mdot3(s[6], s[7], s[8], e0, e1, e2);
        // solve_2x2(a, b, b, d, -j0, -j1) (iter_proj.metal:68-78)
        const float det = __fsub_rn(__fmul_rn(a, d), __fmul_rn(bq, bq));
        float dx = 0.0f, dy = 0.0f;
        if (!(fabsf(det) < 1e-10f)) {
            const float inv = __fdiv_rn(1.0f, det);
            const float e = -j0, f = -j1;
            dx = __fmul_rn(__fsub_rn(__fmul_rn(d, e), __fmul_rn(bq, f)), inv);
            dy = __fmul_rn(__fadd_rn(__fmul_rn(-bq, e), __fmul_rn(a, f)), inv);
        }
        u = __fadd_rn(u, dx);
        v = __fadd_rn(v, dy);
        const float dn = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        if (dn < thresh) break;
        alive = (u == u) && (v == v);
    }
    float2 o;
    o.x = mclamp(u, 0.0f, (float)(W - 1));
    o.y = mclamp(v, 0.0f, (float)(H - 1));
    reinterpret_cast<float2*>(p_out)[t] = o;
    valid_out[t] = (u >= 0.0f) && (u < (float)W) && (v >= 0.0f) && (v < (float)H);
}

// one thread per query; the query descriptor sits in registers for D <= 32 (D4 = D / 4 float4)
template <int D4>
__global__ void __launch_bounds__(kMetalThreads)
refine_metal_kernel(const float* __restrict__ D11, const float* __restrict__ D21,
                    const int32_t* __restrict__ p1, int32_t* __restrict__ p1_out, int H, int W, int D,
                    int N, long long total, int radius, int dilation_max) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / N);
    const float* __restrict__ img = D11 + (size_t)b * H * W * D;
    const float* __restrict__ qp = D21 + (size_t)t * D;
    [[maybe_unused]] float4 q[D4 > 0 ? D4 : 1];
    if constexpr (D4 > 0) {
#pragma unroll
        for (int k = 0; k < D4; ++k) q[k] = __ldg(reinterpret_cast<const float4*>(qp) + k);
    }
    const int2 c0 = __ldg(reinterpret_cast<const int2*>(p1) + t);
    int cx = c0.x, cy = c0.y;
    for (int dil = max(dilation_max, 1); dil >= 1; --dil) {
        float best = -1e10f;
        int bx = cx, by = cy;
        for (int dy = -radius; dy <= radius; ++dy) {
            const int ny = cy + dy * dil;
            if (ny < 0 || ny >= H) continue;
            for (int dx = -radius; dx <= radius; ++dx) {
                const int nx = cx + dx * dil;
                if (nx < 0 || nx >= W) continue;
                const float* __restrict__ cd = img + ((size_t)ny * W + nx) * D;
                float sc = 0.0f;
                if constexpr (D4 > 0) {
#pragma unroll
                    for (int k = 0; k < D4; ++k) {
                        const float4 cv = __ldg(reinterpret_cast<const float4*>(cd) + k);
                        sc = __fadd_rn(sc, __fmul_rn(q[k].x, cv.x));
                        sc = __fadd_rn(sc, __fmul_rn(q[k].y, cv.y));
                        sc = __fadd_rn(sc, __fmul_rn(q[k].z, cv.z));
                        sc = __fadd_rn(sc, __fmul_rn(q[k].w, cv.w));
                    }
                } else {
                    for (int k = 0; k < D; ++k) sc = __fadd_rn(sc, __fmul_rn(__ldg(qp + k), __ldg(cd + k)));
                }
                if (sc > best) { best = sc; bx = nx; by = ny; }
            }
        }
        cx = bx;
        cy = by;
    }
    reinterpret_cast<int2*>(p1_out)[t] = make_int2(cx, cy);
}

}  // namespace

}  // namespace m3s

using namespace m3s;

M3S_API int m3s_iter_proj_metal(const float* rays, const float* pts, const float* p_init, float* p_out,
                                uint8_t* valid_out, int B, int H, int W, int N, int max_iter,
                                float lambda_init, float convergence_thresh, m3s_stream_t stream) {
    M3S_REQUIRE(rays && pts && p_init && p_out && valid_out, "iter_proj_metal: null pointer");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0 && N >= 0 && max_iter >= 0, "iter_proj_metal: bad shape");
    const long long total = (long long)B * N;
    if (total == 0) return M3S_OK;
    const unsigned grid = (unsigned)((total + kMetalThreads - 1) / kMetalThreads);
    iter_proj_metal_kernel<<<grid, kMetalThreads, 0, as_stream(stream)>>>(
        rays, pts, p_init, p_out, valid_out, H, W, N, total, max_iter, lambda_init, convergence_thresh);
    return launch_ok("iter_proj_metal_kernel");
}

M3S_API int m3s_refine_matches_metal(const float* D11, const float* D21, const int32_t* p1,
                                     int32_t* p1_out, int B, int H, int W, int D, int N, int radius,
                                     int dilation_max, m3s_stream_t stream) {
    M3S_REQUIRE(D11 && D21 && p1 && p1_out, "refine_matches_metal: null pointer");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0 && D > 0 && N >= 0 && radius >= 0, "refine_matches_metal: bad shape");
    const long long total = (long long)B * N;
    if (total == 0) return M3S_OK;
    const unsigned grid = (unsigned)((total + kMetalThreads - 1) / kMetalThreads);
    cudaStream_t st = as_stream(stream);
    const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21)) & 15) == 0;
#define M3S_RM_CASE(DD)                                                                              \
    case DD:                                                                                         \
        refine_metal_kernel<DD / 4><<<grid, kMetalThreads, 0, st>>>(D11, D21, p1, p1_out, H, W, D, N, \
                                                                    total, radius, dilation_max);    \
        return launch_ok("refine_metal_kernel")
    if (aligned) {
        switch (D) {
            M3S_RM_CASE(16);
            M3S_RM_CASE(24);
            M3S_RM_CASE(32);
            default: break;
        }
    }
#undef M3S_RM_CASE
    refine_metal_kernel<0><<<grid, kMetalThreads, 0, st>>>(D11, D21, p1, p1_out, H, W, D, N, total, radius,
                                                           dilation_max);
    return launch_ok("refine_metal_kernel");
}
