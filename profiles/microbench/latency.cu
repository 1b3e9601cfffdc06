// Dependent-issue latencies that bound the Gauss-Newton factorisation chain (one warp, one SM):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o latency latency.cu && ./latency
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ double rcp_fast(double d) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double e = fma(-d, y, 1.0);
    const double t = fma(e, e, e);
    return fma(y, t, y);
}
__device__ __forceinline__ double rsqrt_fast(double d) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
    const double h = 0.5 * d;
    double e = fma(-h * y, y, 0.5);
    y = fma(y, e, y);
    e = fma(-h * y, y, 0.5);
    return fma(y, e, y);
}

constexpr int N = 512;
__global__ void lat(double* out, long long* cyc, const int* chase, unsigned* flag, double seed) {
    const int lane = threadIdx.x;
    double x = seed + lane * 1e-3;
    long long t0, t1;
    // 0: dependent DFMA
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = fma(x, 0.999999, 1e-9);
    t1 = clock64(); if (lane == 0) cyc[0] = t1 - t0;
    // 1: dependent DMUL
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = x * 1.0000001;
    t1 = clock64(); if (lane == 0) cyc[1] = t1 - t0;
    // 2: dependent DADD
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = x + 1e-9;
    t1 = clock64(); if (lane == 0) cyc[2] = t1 - t0;
    // 3: rcp.approx.f64 seed alone (dependent)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) { double y; asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x)); x = y; }
    t1 = clock64(); if (lane == 0) cyc[3] = t1 - t0;
    // 4: rcp_fast (seed + 3 fma)
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N; ++i) x = rcp_fast(x);
    t1 = clock64(); if (lane == 0) cyc[4] = t1 - t0;
    // 5: rsqrt_fast
    x = fabs(x) + 1.0;
    t0 = clock64();
#pragma unroll 8
    for (int i = 0; i < N; ++i) x = rsqrt_fast(x) + 1.0;
    t1 = clock64(); if (lane == 0) cyc[5] = t1 - t0;
    // 6: 64-bit shuffle (dependent)
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);
    t1 = clock64(); if (lane == 0) cyc[6] = t1 - t0;
    // 7: dependent FFMA (reference)
    float f = (float)x;
    t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N; ++i) f = fmaf(f, 0.99999f, 1e-6f);
    t1 = clock64(); if (lane == 0) cyc[7] = t1 - t0;
    x += f;
    // 8: ld.global.cg pointer chase (L2 latency)
    int p = lane == 0 ? 0 : 0;
    t0 = clock64();
    for (int i = 0; i < 64; ++i) p = __ldcg(chase + p);
    t1 = clock64(); if (lane == 0) cyc[8] = (t1 - t0);
    x += p;
    // 9: ld.acquire.gpu of one word, dependent
    unsigned v = 0;
    t0 = clock64();
    for (int i = 0; i < 64; ++i) { unsigned w; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(w) : "l"(flag + (v & 1)) : "memory"); v += w; }
    t1 = clock64(); if (lane == 0) cyc[9] = (t1 - t0);
    // 10: store + __threadfence
    t0 = clock64();
    for (int i = 0; i < 64; ++i) { if (lane == 0) out[64 + i] = x; __threadfence(); }
    t1 = clock64(); if (lane == 0) cyc[10] = (t1 - t0);
    // 11: store + st.release.gpu
    t0 = clock64();
    for (int i = 0; i < 64; ++i) { if (lane == 0) { out[128 + i] = x; asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag + 2), "r"(v + i) : "memory"); } }
    t1 = clock64(); if (lane == 0) cyc[11] = (t1 - t0);
    // 12: __threadfence_block
    t0 = clock64();
    for (int i = 0; i < 64; ++i) { if (lane == 0) out[192 + i] = x; __threadfence_block(); }
    t1 = clock64(); if (lane == 0) cyc[12] = (t1 - t0);
    // 13: __syncthreads (1 warp)
    t0 = clock64();
    for (int i = 0; i < 64; ++i) __syncthreads();
    t1 = clock64(); if (lane == 0) cyc[13] = (t1 - t0);
    out[lane] = x + v;
}

__global__ void bar256(long long* cyc) {
    long long t0 = clock64();
    for (int i = 0; i < 64; ++i) __syncthreads();
    long long t1 = clock64();
    if (threadIdx.x == 0) cyc[14] = t1 - t0;
}

int main() {
    double* out; long long* cyc; int* chase; unsigned* flag;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 256); cudaMalloc(&flag, 64);
    const int M = 1 << 22;   // 16 MB of ints: stays in L2, misses L1 (ld.cg)
    cudaMalloc(&chase, sizeof(int) * M);
    int* h = new int[M];
    for (int i = 0; i < M; ++i) h[i] = (int)(((long long)i + 1048583) % M);
    cudaMemcpy(chase, h, sizeof(int) * M, cudaMemcpyHostToDevice);
    cudaMemset(flag, 0, 64);
    for (int rep = 0; rep < 2; ++rep) { lat<<<1, 32>>>(out, cyc, chase, flag, 1.2345); bar256<<<1, 256>>>(cyc); }
    cudaDeviceSynchronize();
    long long c[16];
    cudaMemcpy(c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    const char* names[] = {"DFMA dependent", "DMUL dependent", "DADD dependent", "rcp.approx.f64 seed", "rcp_fast (seed + 3 DFMA)",
                           "rsqrt_fast (+1 DADD)", "SHFL 64-bit dependent", "FFMA dependent", "ld.global.cg chase (L2 hit)",
                           "ld.acquire.gpu dependent", "st + __threadfence", "st + st.release.gpu", "st + __threadfence_block",
                           "__syncthreads (1 warp)", "__syncthreads (8 warps)"};
    const int div[] = {N, N, N, N, N, N, N, N, 64, 64, 64, 64, 64, 64, 64};
    for (int i = 0; i < 15; ++i) printf("%-30s %8.1f clk\n", names[i], (double)c[i] / div[i]);
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
