// Stand-alone timing of tl_elim32_w (gn.cu): the same 32x32 elimination four times back to back in one
// CTA, optionally after a spin-wait (idle SM) or a burst of GEMM work, with per-pair clock stamps.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../include -I../../mast3r-slam_b200/csrc \
//        -o elim_bench elim_bench.cu && ./elim_bench
#include "../../mast3r-slam_b200/csrc/gn.cu"
#include <cstdarg>
namespace m3s {   // the few host helpers gn.cu takes from api.cu
void set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap); }
int cuda_fail(cudaError_t e, const char* what) { fprintf(stderr, "%s: %s\n", what, cudaGetErrorString(e)); return M3S_ERR_CUDA; }
int sm_count_current() { return 148; }
unsigned long long* path_counters() { return nullptr; }
}

__global__ void __launch_bounds__(256, 2) bench(const double* W, long long* dbg, double* out, int mode, volatile int* gate) {
    extern __shared__ __align__(16) unsigned char raw[];
    TilesSmem& sm = *reinterpret_cast<TilesSmem*>(raw);
    const int tid = threadIdx.x;
    for (int rep = 0; rep < 4; ++rep) {
        for (int i = tid; i < 64 * 64; i += 256) {
            const int r = i >> 6, c = i & 63;
            sm.T[r][c] = (r < 32 && c < 32) ? W[r * 32 + c] : (r == c ? 1.0 : 0.0);
        }
        if (tid == 0) { sm.prog = 0; sm.hsflag = 0; sm.bad = 0; }
        __syncthreads();
        if (mode == 1 && rep == 2) {
            // idle wait: one thread polls a global word for ~20 us, everybody else sits at the barrier
            if (tid == 0) { long long t0 = clock64(); while (clock64() - t0 < 40000) { (void)*gate; } }
            __syncthreads();
        }
        if (mode == 2 && rep == 2) {
            // burst of fp64 GEMM work by all warps right before
            for (int it = 0; it < 4; ++it) {
                double acc[4][4];
                tl_gemm_nt<4, 4>(&sm.PA[0][0], kPP, &sm.PB[0][0], kPP, acc);
                tl_sub<4, 4>(&sm.S[0][0], kTP, acc);
                __syncthreads();
            }
        }
        if (tl_is_elim_warp()) tl_elim32_w(sm, sm.T, 0, &sm.S[0][0], &sm.PB[0][0], sm.Mi, dbg + 64 * rep);
        __syncthreads();
    }
    if (tid < 32) out[tid] = sm.Mi[tid][0];
}

// block 0 runs the elimination; every other block of the grid keeps its SM busy with shared-memory GEMMs
// (mode 0), shared-memory GEMMs + global loads/stores (mode 1) until block 0 is done
__global__ void __launch_bounds__(256, 2) bench_neighbours(const double* W, long long* dbg, double* out, int mode, volatile int* gate,
                                                         double* scratch) {
    extern __shared__ __align__(16) unsigned char raw[];
    TilesSmem& sm = *reinterpret_cast<TilesSmem*>(raw);
    const int tid = threadIdx.x;
    if (tid == 0) { unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); dbg[1000 + blockIdx.x] = smid; }
    if (blockIdx.x != 0) {
        for (int i = tid; i < 64 * kPP; i += 256) { (&sm.PA[0][0])[i] = 1e-3 * i; (&sm.PB[0][0])[i] = 1e-3; }
        __syncthreads();
        while (*gate == 0) {
            double acc[4][4];
            tl_gemm_nt<4, 4>(&sm.PA[0][0], kPP, &sm.PB[0][0], kPP, acc);
            tl_sub<4, 4>(&sm.S[0][0], kTP, acc);
            if (mode == 1) {
                double* g = scratch + (size_t)blockIdx.x * 8192;
                for (int i = tid; i < 4096; i += 256) g[i] = __ldcg(g + 4096 + i) + acc[0][0];
            }
            __syncthreads();
        }
        return;
    }
    for (int rep = 0; rep < 4; ++rep) {
        for (int i = tid; i < 64 * 64; i += 256) {
            const int r = i >> 6, c = i & 63;
            sm.T[r][c] = (r < 32 && c < 32) ? W[r * 32 + c] : (r == c ? 1.0 : 0.0);
        }
        if (tid == 0) { sm.prog = 0; sm.hsflag = 0; sm.bad = 0; }
        __syncthreads();
        if (tl_is_elim_warp()) tl_elim32_w(sm, sm.T, 0, &sm.S[0][0], &sm.PB[0][0], sm.Mi, dbg + 64 * rep);
        __syncthreads();
    }
    if (tid < 32) out[tid] = sm.Mi[tid][0];
    __syncthreads();
    if (tid == 0) { __threadfence(); *gate = 1; }
}

int main() {
    double hW[1024];
    for (int r = 0; r < 32; ++r)
        for (int c = 0; c < 32; ++c) hW[r * 32 + c] = (r == c ? 40.0 : 0.0) + 1.0 / (1.0 + abs(r - c));
    double *W, *out; long long* dbg; int* gate;
    cudaMalloc(&W, sizeof(hW)); cudaMalloc(&out, 4096); cudaMalloc(&dbg, 8 * 256 * 3); cudaMalloc(&gate, 4);
    cudaMemcpy(W, hW, sizeof(hW), cudaMemcpyHostToDevice);
    cudaMemset(gate, 0, 4);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TilesSmem));
    for (int mode = 0; mode < 3; ++mode) {
        cudaMemset(dbg, 0, 8 * 256 * 3);
        bench<<<1, 256, sizeof(TilesSmem)>>>(W, dbg, out, mode, gate);
        cudaDeviceSynchronize();
        long long h[256];
        cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost);
        printf("mode %d (%s before the third call): %s\n", mode, mode == 0 ? "nothing" : mode == 1 ? "20-us idle wait" : "GEMM burst",
               cudaGetErrorString(cudaGetLastError()));
        for (int rep = 0; rep < 4; ++rep) {
            printf("  call %d chain  :", rep);
            for (int q = 0; q < 16; ++q) printf(" %5lld", h[64 * rep + q] - h[64 * rep + 48]);
            printf("\n  call %d inverse:", rep);
            for (int q = 0; q < 16; ++q) printf(" %5lld", h[64 * rep + 16 + q] - h[64 * rep + 48]);
            printf("\n");
        }
    }
    double* scratch;
    cudaMalloc(&scratch, sizeof(double) * 8192 * 300);
    cudaMemset(scratch, 0, sizeof(double) * 8192 * 300);
    cudaFuncSetAttribute(bench_neighbours, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TilesSmem));
    const int grids[] = {2, 4, 148, 232};
    for (int gi = 0; gi < 4; ++gi)
        for (int mode = 0; mode < 2; ++mode) {
            cudaMemset(dbg, 0, 8 * 256 * 3); cudaMemset(gate, 0, 4);
            long long* dbg2; cudaMalloc(&dbg2, 8 * 2048); cudaMemset(dbg2, 0, 8 * 2048);
            bench_neighbours<<<grids[gi], 256, sizeof(TilesSmem)>>>(W, dbg2, out, mode, gate, scratch);
            cudaDeviceSynchronize();
            static long long h[2048];
            cudaMemcpy(h, dbg2, sizeof(h), cudaMemcpyDeviceToHost);
            printf("grid %d, neighbours %s: %s; SM of block 0..3: %lld %lld %lld %lld\n", grids[gi],
                   mode == 0 ? "shared-memory GEMMs" : "GEMMs + global traffic", cudaGetErrorString(cudaGetLastError()),
                   h[1000], h[1001], h[1002], h[1003]);
            for (int rep = 1; rep < 3; ++rep) {
                printf("  call %d chain  :", rep);
                for (int q = 0; q < 16; ++q) printf(" %5lld", h[64 * rep + q] - h[64 * rep + 48]);
                printf("\n");
            }
            cudaFree(dbg2);
        }
    return 0;
}
