// Pipe-rate microbenchmark for B200 (sm_100a): lanes/clk/SM of the instruction classes the matching
// kernels are bound by.  One CTA of 1024 threads per SM; every thread runs ITER x UNROLL independent
// chains; rate = threads * ops / (clock64 delta).   nvcc -arch=sm_100a -O3 -o pipes pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 512
template <int OP> __global__ void __launch_bounds__(1024) k(float* out, long long* clk, float seed) {
    __shared__ __align__(16) double sm[4096];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = seed * i;
    __syncthreads();
    float a[8]; double d[8]; float2 f2[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed + i + threadIdx.x; d[i] = a[i]; f2[i] = make_float2(a[i], a[i] + 1); }
    float m0 = seed * 1.5f, m1 = seed * 0.5f; double dm0 = m0, dm1 = m1;
    float2 m2 = make_float2(m0, m1), m3 = make_float2(m1, m0);
    int lane16 = (threadIdx.x * 2) & 4095;
    int lane8 = threadIdx.x & 4095;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == 0) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(m0), "f"(m1));          // FFMA, 2 shared srcs
            if (OP == 1) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(a[(i + 1) & 7]), "f"(m1)); // FFMA, distinct regs
            if (OP == 2) { unsigned long long x, y, z; asm volatile("mov.b64 %0, {%3,%4}; mov.b64 %1, {%5,%6}; mov.b64 %2, {%7,%8}; fma.rn.f32x2 %0, %1, %2, %0; mov.b64 {%3,%4}, %0;" : "=l"(x), "=l"(y), "=l"(z), "+f"(f2[i].x), "+f"(f2[i].y) : "f"(m2.x), "f"(m2.y), "f"(m3.x), "f"(m3.y)); }
            if (OP == 14) { unsigned long long x, y, z; asm volatile("mov.b64 %0, {%3,%4}; mov.b64 %1, {%5,%6}; mov.b64 %2, {%7,%8}; fma.rn.f32x2 %0, %1, %2, %0; mov.b64 {%3,%4}, %0;" : "=l"(x), "=l"(y), "=l"(z), "+f"(f2[i].x), "+f"(f2[i].y) : "f"(f2[(i + 1) & 7].x), "f"(f2[(i + 1) & 7].y), "f"(m3.x), "f"(m3.y)); }
            if (OP == 3) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(dm0));
            if (OP == 4) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(dm0));
            if (OP == 5) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(d[i]) : "d"(dm0), "d"(dm1));
            if (OP == 6) asm volatile("cvt.f64.f32 %0, %1;" : "=d"(d[i]) : "f"(a[i]));
            if (OP == 7) asm volatile("cvt.rn.f32.f64 %0, %1;" : "=f"(a[i]) : "d"(d[i]));
            if (OP == 8) { double2 v; asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"((unsigned)__cvta_generic_to_shared(&sm[(lane16 + i * 2 * 32) & 4094]))); d[i] += v.x; }
            if (OP == 9) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"((unsigned)__cvta_generic_to_shared(&sm[(lane8 + i * 32) & 4095]))); d[i] += v; }
            if (OP == 10) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (OP == 11) { int x = __float_as_int(a[i]); asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x) : "r"(it), "r"(i)); a[i] = __int_as_float(x); }
            if (OP == 12) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m0));
            if (OP == 13) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m0));
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + (float)d[i] + f2[i].x + f2[i].y;
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int threads, float* out, long long* clk) {
    k<OP><<<148, threads>>>(out, clk, 1.0f); cudaDeviceSynchronize();
    k<OP><<<148, threads>>>(out, clk, 1.0f); cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    printf("%-28s threads=%4d  cycles=%9.0f  thread-ops/clk/SM=%7.2f\n", name, threads, avg, (double)threads * ITER * 8 / avg);
}
int main() {
    float* out; long long* clk; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 148 * 8);
    for (int th : {256, 512, 1024}) {
        run<0>("FFMA shared-src", th, out, clk); run<1>("FFMA distinct", th, out, clk); run<2>("FFMA2 (x2 flops)", th, out, clk); run<14>("FFMA2 distinct", th, out, clk);
        run<12>("FMUL", th, out, clk); run<13>("FADD", th, out, clk);
        run<3>("DADD", th, out, clk); run<4>("DMUL", th, out, clk); run<5>("DFMA", th, out, clk);
        run<6>("F2F.F64.F32", th, out, clk); run<7>("F2F.F32.F64", th, out, clk);
        run<8>("LDS.128 (+DADD)", th, out, clk); run<9>("LDS.64 (+DADD)", th, out, clk);
        run<10>("MUFU.RCP", th, out, clk); run<11>("LOP3", th, out, clk);
    }
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
