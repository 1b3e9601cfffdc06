// Instruction-fetch cost of straight-line code on one SM (one warp):
//   cold  = first execution of a 4-KB block of independent FFMAs
//   warm  = immediate re-execution
//   after k other 4-KB blocks = re-execution after k x 4 KB of different code ran in between
//   nvcc -gencode arch=compute_100a,code=sm_100a -O1 -o icache icache.cu && ./icache
#include <cstdio>
#include <cuda_runtime.h>

template <int ID>
__device__ __noinline__ float blk(float x) {
    float a = x, b = x + 1.f, c = x + 2.f, d = x + 3.f;
    // 256 FFMA (4 chains), distinct immediates per ID so the compiler cannot merge blocks
#pragma unroll
    for (int i = 0; i < 64; ++i) {
        a = fmaf(a, 1.0001f + ID * 1e-6f, 0.5f + i);
        b = fmaf(b, 0.9999f + ID * 1e-6f, 0.25f + i);
        c = fmaf(c, 1.0002f + ID * 1e-6f, 0.125f + i);
        d = fmaf(d, 0.9998f + ID * 1e-6f, 0.0625f + i);
    }
    return (a + b) + (c + d);
}

template <int N> struct Run {
    __device__ static float go(float x) { return Run<N - 1>::go(blk<N>(x)); }
};
template <> struct Run<0> { __device__ static float go(float x) { return x; } };

__global__ void k(float* out, long long* cyc, float seed) {
    float x = seed;
    long long t0, t1;
    int n = 0;
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // cold
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // warm
    x = Run<1>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 4 KB
    x = Run<2>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 8 KB (blocks 1,2: block 1 now warm)
    x = Run<4>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 16 KB
    x = Run<8>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 32 KB
    x = Run<16>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 64 KB
    x = Run<32>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 128 KB
    x = Run<48>::go(x);
    t0 = clock64(); x = blk<100>(x); t1 = clock64(); cyc[n++] = t1 - t0;   // after 192 KB
    t0 = clock64(); x = Run<48>::go(x); t1 = clock64(); cyc[n++] = t1 - t0;  // 192 KB, second pass
    t0 = clock64(); x = blk<99>(x); t1 = clock64(); cyc[n++] = t1 - t0;    // another cold block
    out[threadIdx.x] = x;
}

int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 256);
    k<<<1, 32>>>(out, cyc, 1.0f);
    cudaDeviceSynchronize();
    long long c[16];
    cudaMemcpy(c, cyc, sizeof(c), cudaMemcpyDeviceToHost);
    const char* names[] = {"cold 4-KB block", "warm (immediately again)", "after 4 KB of other code", "after 8 KB", "after 16 KB",
                           "after 32 KB", "after 64 KB", "after 128 KB", "after 192 KB", "192 KB second pass (48 blocks)", "another cold block"};
    for (int i = 0; i < 11; ++i) printf("%-34s %8lld clk\n", names[i], c[i]);
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
