// refine_matches: local argmax of descriptor dot products (backends/mpsgraph/kernels.py:496-537).
//
// Semantics (numpy path): window (2r+1)^2 around trunc(p1), step 1 (dilation_max has no effect in
// the reference's numpy implementation: every level restarts from the original p1 and the last
// level, dilation 1, overwrites the result), scan dy outer / dx inner, strict '>', out-of-image
// candidates skipped, centre kept when no candidate is inside the image.
//
// Scores are fp32 sums over D products.  The reference's np.dot is BLAS sdot whose summation order
// is unspecified, so FMA accumulation (k ascending) is used here; for descriptors on a dyadic grid
// every order is exact and the argmax is bit-exact unconditionally.
#include "m3s_common.cuh"

namespace m3s {

constexpr int kRefineThreads = 128;

__device__ __forceinline__ int2 load_center(const int32_t* p1, long long t) {
    return __ldg(reinterpret_cast<const int2*>(p1) + t);
}
__device__ __forceinline__ int2 load_center(const float* p1, long long t) {
    const float2 f = __ldg(reinterpret_cast<const float2*>(p1) + t);
    return make_int2((int)f.x, (int)f.y);  // astype(int32): truncation toward zero
}

// One thread per query; query descriptor in registers (D4 float4), candidates through L1.
template <int D4, typename PT>
__global__ void __launch_bounds__(kRefineThreads)
refine_kernel_vec(const float* __restrict__ D11, const float* __restrict__ D21,
                  const PT* __restrict__ p1, int32_t* __restrict__ p1_out, int H, int W, int N,
                  long long total, int radius) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / N);
    constexpr int D = D4 * 4;
    const float4* __restrict__ img = reinterpret_cast<const float4*>(D11 + (size_t)b * H * W * D);
    const float4* __restrict__ qp = reinterpret_cast<const float4*>(D21 + (size_t)t * D);
    float4 q[D4];
#pragma unroll
    for (int k = 0; k < D4; ++k) q[k] = __ldg(qp + k);
    const int2 c = load_center(p1, t);
    float best = -INFINITY;
    int bx = c.x, by = c.y;
    // clip the window to the image once: equivalent to skipping OOB candidates
    const int y_lo = max(c.y - radius, 0), y_hi = min(c.y + radius, H - 1);
    const int x_lo = max(c.x - radius, 0), x_hi = min(c.x + radius, W - 1);
    for (int ny = y_lo; ny <= y_hi; ++ny) {
        const float4* __restrict__ row = img + ((size_t)ny * W) * D4;
        for (int nx = x_lo; nx <= x_hi; ++nx) {
            const float4* __restrict__ cd = row + (size_t)nx * D4;
            float sc = 0.0f;
#pragma unroll
            for (int k = 0; k < D4; ++k) {
                const float4 v = __ldg(cd + k);
                sc = fmaf(q[k].x, v.x, sc);
                sc = fmaf(q[k].y, v.y, sc);
                sc = fmaf(q[k].z, v.z, sc);
                sc = fmaf(q[k].w, v.w, sc);
            }
            if (sc > best) { best = sc; bx = nx; by = ny; }
        }
    }
    reinterpret_cast<int2*>(p1_out)[t] = make_int2(bx, by);
}

// Any descriptor dimension (scalar loads).
template <typename PT>
__global__ void __launch_bounds__(kRefineThreads)
refine_kernel_generic(const float* __restrict__ D11, const float* __restrict__ D21,
                      const PT* __restrict__ p1, int32_t* __restrict__ p1_out, int H, int W, int D,
                      int N, long long total, int radius) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const int b = (int)(t / N);
    const float* __restrict__ img = D11 + (size_t)b * H * W * D;
    const float* __restrict__ qp = D21 + (size_t)t * D;
    const int2 c = load_center(p1, t);
    float best = -INFINITY;
    int bx = c.x, by = c.y;
    const int y_lo = max(c.y - radius, 0), y_hi = min(c.y + radius, H - 1);
    const int x_lo = max(c.x - radius, 0), x_hi = min(c.x + radius, W - 1);
    for (int ny = y_lo; ny <= y_hi; ++ny) {
        for (int nx = x_lo; nx <= x_hi; ++nx) {
            const float* __restrict__ cd = img + ((size_t)ny * W + nx) * D;
            float sc = 0.0f;
            for (int k = 0; k < D; ++k) sc = fmaf(__ldg(qp + k), __ldg(cd + k), sc);
            if (sc > best) { best = sc; bx = nx; by = ny; }
        }
    }
    reinterpret_cast<int2*>(p1_out)[t] = make_int2(bx, by);
}

template <typename PT>
int refine_dispatch(const float* D11, const float* D21, const PT* p1, int32_t* p1_out, int B,
                    int H, int W, int D, int N, int radius, cudaStream_t st) {
    const long long total = (long long)B * N;
    if (total == 0) return M3S_OK;
    const unsigned grid = (unsigned)((total + kRefineThreads - 1) / kRefineThreads);
    const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21)) & 15) == 0;
#define M3S_REFINE_CASE(DD)                                                                     \
    case DD:                                                                                    \
        refine_kernel_vec<DD / 4, PT><<<grid, kRefineThreads, 0, st>>>(D11, D21, p1, p1_out, H, \
                                                                       W, N, total, radius);    \
        return launch_ok("refine_kernel_vec")
    if (aligned) {
        switch (D) {
            M3S_REFINE_CASE(16);
            M3S_REFINE_CASE(24);
            M3S_REFINE_CASE(32);
            M3S_REFINE_CASE(64);
            default: break;
        }
    }
#undef M3S_REFINE_CASE
    refine_kernel_generic<PT><<<grid, kRefineThreads, 0, st>>>(D11, D21, p1, p1_out, H, W, D, N,
                                                                total, radius);
    return launch_ok("refine_kernel_generic");
}

}  // namespace m3s

using namespace m3s;

M3S_API int m3s_refine_matches(const float* D11, const float* D21, const int32_t* p1,
                               int32_t* p1_out, int B, int H, int W, int D, int N, int radius,
                               int dilation_max, m3s_stream_t stream) {
    (void)dilation_max;  // no effect in the reference's numpy path (kernels.py:510-535)
    M3S_REQUIRE(D11 && D21 && p1 && p1_out, "refine_matches: null pointer");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0 && D > 0 && N >= 0 && radius >= 0, "refine_matches: bad shape");
    return refine_dispatch<int32_t>(D11, D21, p1, p1_out, B, H, W, D, N, radius, as_stream(stream));
}

M3S_API int m3s_refine_matches_f32(const float* D11, const float* D21, const float* p1,
                                   int32_t* p1_out, int B, int H, int W, int D, int N, int radius,
                                   int dilation_max, m3s_stream_t stream) {
    (void)dilation_max;
    M3S_REQUIRE(D11 && D21 && p1 && p1_out, "refine_matches_f32: null pointer");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0 && D > 0 && N >= 0 && radius >= 0, "refine_matches_f32: bad shape");
    return refine_dispatch<float>(D11, D21, p1, p1_out, B, H, W, D, N, radius, as_stream(stream));
}
