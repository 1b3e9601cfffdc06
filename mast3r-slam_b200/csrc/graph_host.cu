// Array-level kernels of the factor-graph host: the two elementwise /
// gather steps that sit between match() and gauss_newton_*() in the reference's FactorGraph.
//
//   m3s_constrain_points_to_ray  geometry.py:273-302 (constrain_points_to_ray) + :229-246 (backproject)
//   m3s_edge_confidence          global_opt.py:86-104 (add_factors: Qj = sqrt(Qii[idx] * Qji),
//                                valid = valid_match & (Qj > Q_conf), per-edge valid count)
//
// Both are HBM-streaming (24 B/px and 13-17 B/px); fp32 with explicit round-to-nearest ops in the
// reference's evaluation order so the results are bit-identical to a numpy float32 restatement.
#include "m3s_common.cuh"

namespace m3s {

__global__ void __launch_bounds__(256)
constrain_points_kernel(const float* __restrict__ Xs, float fx, float fy, float cx, float cy,
                        float* __restrict__ out, int W, long long NP, long long total) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    const long long n = t % NP;
    const float u = (float)(n % W), v = (float)(n / W);
    const float z = __ldg(Xs + 3 * t + 2);
    // backproject: x = (u - cx) / fx * z   (left to right, every op rounded)
    out[3 * t] = __fmul_rn(__fdiv_rn(__fsub_rn(u, cx), fx), z);
    out[3 * t + 1] = __fmul_rn(__fdiv_rn(__fsub_rn(v, cy), fy), z);
    out[3 * t + 2] = z;
}

// one block per (edge, chunk of 4096 points); per-edge valid count by warp ballot + one atomicAdd
constexpr int kConfChunk = 4096;
__global__ void __launch_bounds__(256)
edge_confidence_kernel(const float* __restrict__ Qii, const float* __restrict__ Qji,
                       const int32_t* __restrict__ idx, const uint8_t* __restrict__ valid_match,
                       float q_conf, float* __restrict__ Qj_out, uint8_t* __restrict__ valid_out,
                       int32_t* __restrict__ count, long long N, int nchunks) {
    const int e = blockIdx.x / nchunks, chunk = blockIdx.x % nchunks;
    const long long base = (long long)e * N;
    const long long n0 = (long long)chunk * kConfChunk, n1 = min(N, n0 + kConfChunk);
    int cnt = 0;
    for (long long n = n0 + threadIdx.x; n < n1; n += blockDim.x) {
        const long long m = __ldg(idx + base + n);
        const bool in = (m >= 0) && (m < N);
        const float qi = in ? __ldg(Qii + base + m) : 0.0f;
        const float q = __fsqrt_rn(__fmul_rn(qi, __ldg(Qji + base + n)));
        const bool ok = in && (__ldg(valid_match + base + n) != 0) && (q > q_conf);
        Qj_out[base + n] = q;
        if (valid_out) valid_out[base + n] = ok;
        cnt += ok;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    __shared__ int s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += s[w];
        if (tot) atomicAdd(count + e, tot);  // integer: order-independent
    }
}

}  // namespace m3s

using namespace m3s;

M3S_API int m3s_constrain_points_to_ray(const float* Xs, const float* K4_host, float* out, int B,
                                        int H, int W, m3s_stream_t stream) {
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0, "constrain_points_to_ray: bad shape");
    const long long NP = (long long)H * W, total = (long long)B * NP;
    if (total == 0) return M3S_OK;
    M3S_REQUIRE(Xs && K4_host && out, "constrain_points_to_ray: null pointer");
    const unsigned grid = (unsigned)((total + 255) / 256);
    constrain_points_kernel<<<grid, 256, 0, as_stream(stream)>>>(Xs, K4_host[0], K4_host[1], K4_host[2],
                                                                 K4_host[3], out, W, NP, total);
    return launch_ok("constrain_points_kernel");
}

M3S_API int m3s_edge_confidence(const float* Qii, const float* Qji, const int32_t* idx,
                                const uint8_t* valid_match, float q_conf, float* Qj_out,
                                uint8_t* valid_out, int32_t* valid_count, int E, long long N,
                                m3s_stream_t stream) {
    M3S_REQUIRE(E >= 0 && N >= 0, "edge_confidence: bad shape");
    if (E == 0) return M3S_OK;  // empty tensors carry null pointers
    M3S_REQUIRE(Qii && Qji && idx && valid_match && Qj_out && valid_count, "edge_confidence: null pointer");
    cudaStream_t st = as_stream(stream);
    M3S_CUDA(cudaMemsetAsync(valid_count, 0, sizeof(int32_t) * E, st));
    if (N == 0) return M3S_OK;
    const int nchunks = (int)((N + kConfChunk - 1) / kConfChunk);
    edge_confidence_kernel<<<(unsigned)((long long)E * nchunks), 256, 0, st>>>(
        Qii, Qji, idx, valid_match, q_conf, Qj_out, valid_out, valid_count, N, nchunks);
    return launch_ok("edge_confidence_kernel");
}
