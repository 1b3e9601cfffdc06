// Pipe-rate microbenchmark for B200 (sm_100a): lanes/clk/SM of the instruction classes the matching
// kernels are bound by.  One CTA per SM; every thread runs ITER x 8 (or 16) independent chains;
// rate = threads * ops / (clock64 delta).   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu
//
// Every measured instruction has a LOOP-CARRIED data dependency and its result reaches the final
// store, so ptxas cannot drop it (round 1's F2F rows were dead code: the conversion result was
// overwritten every iteration).  Where the dependency needs a second instruction it sits on another
// pipe with a higher rate (LOP3 on the ALU pipe), noted in the row name; check the SASS with
//   cuobjdump -sass pipes | grep -c F2F
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 512
enum Op { FFMA_SH, FFMA_DI, FFMA2_SH, FFMA2_DI, FMUL, FADD, DADD, DMUL, DFMA, F2F_64_32, F2F_32_64, F2F_RT,
          LDS128, LDS64, LDS32, RCP, LOP3, PRMT, SHF, IMAD, FSETP_FSEL, WIDEN_ALU };

template <int OP> __global__ void __launch_bounds__(1024) k(float* out, long long* clk, float seed) {
    __shared__ __align__(16) unsigned sm[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = (unsigned)(seed * i);
    __syncthreads();
    float a[8]; double d[8]; float2 f2[8]; unsigned u[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        a[i] = seed + i + threadIdx.x; d[i] = a[i]; f2[i] = make_float2(a[i], a[i] + 1);
        u[i] = threadIdx.x * 16 + i * 4;
    }
    float m0 = seed * 1.0000001f, m1 = seed * 1e-9f; double dm0 = m0, dm1 = m1;
    float2 m2 = make_float2(m0, m1), m3 = make_float2(m1, m0);
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (OP == FFMA_SH) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(m0), "f"(m1));
            if (OP == FFMA_DI) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(a[i]) : "f"(a[(i + 1) & 7]), "f"(m1));
            if (OP == FFMA2_SH) { unsigned long long x, y, z; asm volatile("mov.b64 %0, {%3,%4}; mov.b64 %1, {%5,%6}; mov.b64 %2, {%7,%8}; fma.rn.f32x2 %0, %1, %2, %0; mov.b64 {%3,%4}, %0;" : "=l"(x), "=l"(y), "=l"(z), "+f"(f2[i].x), "+f"(f2[i].y) : "f"(m2.x), "f"(m2.y), "f"(m3.x), "f"(m3.y)); }
            if (OP == FFMA2_DI) { unsigned long long x, y, z; asm volatile("mov.b64 %0, {%3,%4}; mov.b64 %1, {%5,%6}; mov.b64 %2, {%7,%8}; fma.rn.f32x2 %0, %1, %2, %0; mov.b64 {%3,%4}, %0;" : "=l"(x), "=l"(y), "=l"(z), "+f"(f2[i].x), "+f"(f2[i].y) : "f"(f2[(i + 1) & 7].x), "f"(f2[(i + 1) & 7].y), "f"(m3.x), "f"(m3.y)); }
            if (OP == FMUL) asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m0));
            if (OP == FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(m1));
            if (OP == DADD) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(dm1));
            if (OP == DMUL) asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(d[i]) : "d"(dm0));
            if (OP == DFMA) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(d[i]) : "d"(dm0), "d"(dm1));
            // f32 -> f64 widening; the next input is the low word of the result xor the old input (ALU)
            if (OP == F2F_64_32) asm volatile("{ .reg .b32 lo, hi; cvt.f64.f32 %1, %0; mov.b64 {lo, hi}, %1; xor.b32 lo, lo, hi; and.b32 lo, lo, 0x3fffffff; mov.b32 %0, lo; }" : "+f"(a[i]), "+d"(d[i]));
            // f64 -> f32 rounding; the next input gets the result or-ed into its low word (ALU)
            if (OP == F2F_32_64) asm volatile("{ .reg .b32 lo, hi, r; cvt.rn.f32.f64 %0, %1; mov.b32 r, %0; mov.b64 {lo, hi}, %1; xor.b32 lo, lo, r; mov.b64 %1, {lo, hi}; }" : "+f"(a[i]), "+d"(d[i]));
            // both directions chained (2 XU instructions per counted op pair; reported per instruction)
            if (OP == F2F_RT) asm volatile("cvt.f64.f32 %1, %0; cvt.rn.f32.f64 %0, %1;" : "+f"(a[i]), "+d"(d[i]));
            // shared-memory loads: the loaded word feeds the next address (pointer chase inside the CTA's array)
            if (OP == LDS128) { unsigned x, y, z, w; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(x), "=r"(y), "=r"(z), "=r"(w) : "r"(sbase + (u[i] & 0x7ff0))); u[i] = u[i] + 16 * 1024 + ((x ^ y ^ z ^ w) & 0x10); }
            if (OP == LDS64) { unsigned x, y; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(x), "=r"(y) : "r"(sbase + ((u[i] >> 1) & 0x7ff8))); u[i] = u[i] + 16 * 1024 + ((x ^ y) & 0x10); }
            if (OP == LDS32) { unsigned x; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x) : "r"(sbase + ((u[i] >> 2) & 0x7ffc))); u[i] = u[i] + 16 * 1024 + (x & 0x10); }
            if (OP == RCP) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
            if (OP == LOP3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(u[i]) : "r"(it), "r"(i));
            if (OP == PRMT) asm volatile("prmt.b32 %0, %0, %1, 0x2103;" : "+r"(u[i]) : "r"(it));
            if (OP == SHF) asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(u[i]) : "r"(it));
            if (OP == IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(u[i]) : "r"(it | 1), "r"(i));
            if (OP == FSETP_FSEL) asm volatile("{ .reg .pred p; setp.gt.f32 p, %0, %1; selp.f32 %0, %1, %0, p; }" : "+f"(a[i]) : "f"(m0 + it));
            // exact f32 -> (scaled) f64 widening on the ALU pipe: hi = sign | (|b| >> 3), lo = b << 29
            // (4 ALU instructions per value; counted per VALUE)
            if (OP == WIDEN_ALU) asm volatile("{ .reg .b32 t, hi, lo; and.b32 t, %0, 0x7fffffff; shr.u32 t, t, 3; lop3.b32 hi, t, %0, 0x80000000, 0xf8; shl.b32 lo, %0, 29; xor.b32 %0, hi, lo; }" : "+r"(u[i]));
        }
    }
    long long t1 = clock64();
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + (float)d[i] + f2[i].x + f2[i].y + (float)u[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}
template <int OP> void run(const char* name, int threads, float* out, long long* clk, double per = 1.0) {
    k<OP><<<148, threads>>>(out, clk, 1.0f); cudaDeviceSynchronize();
    k<OP><<<148, threads>>>(out, clk, 1.0f); cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    printf("%-44s threads=%4d  cycles=%9.0f  thread-ops/clk/SM=%7.2f\n", name, threads, avg, per * threads * ITER * 8 / avg);
}
int main() {
    float* out; long long* clk; cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 148 * 8);
    for (int th : {512, 1024}) {
        run<FFMA_SH>("FFMA (2 shared srcs)", th, out, clk); run<FFMA_DI>("FFMA (distinct regs)", th, out, clk);
        run<FFMA2_SH>("FFMA2 (x2 flops per op)", th, out, clk); run<FFMA2_DI>("FFMA2 distinct", th, out, clk);
        run<FMUL>("FMUL", th, out, clk); run<FADD>("FADD", th, out, clk);
        run<DADD>("DADD", th, out, clk); run<DMUL>("DMUL", th, out, clk); run<DFMA>("DFMA", th, out, clk);
        run<F2F_64_32>("F2F.F64.F32 (+2 LOP3)", th, out, clk); run<F2F_32_64>("F2F.F32.F64 (+1 LOP3)", th, out, clk);
        run<F2F_RT>("F2F round trip (per instruction)", th, out, clk, 2.0);
        run<LDS128>("LDS.128 dependent (+ALU)", th, out, clk); run<LDS64>("LDS.64 dependent (+ALU)", th, out, clk);
        run<LDS32>("LDS.32 dependent (+ALU)", th, out, clk);
        run<RCP>("MUFU.RCP", th, out, clk); run<LOP3>("LOP3", th, out, clk); run<PRMT>("PRMT", th, out, clk);
        run<SHF>("SHF", th, out, clk); run<IMAD>("IMAD", th, out, clk); run<FSETP_FSEL>("FSETP+FSEL (per pair)", th, out, clk);
        run<WIDEN_ALU>("f32->f64 widening by ALU ops (per value)", th, out, clk);
    }
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
