// Shared host/device helpers for the m3s-b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "m3s.h"

#define M3S_API extern "C" __attribute__((visibility("default")))

namespace m3s {

// thread-local error message behind m3s_last_error()
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define M3S_CUDA(call)                                              \
    do {                                                            \
        cudaError_t e__ = (call);                                   \
        if (e__ != cudaSuccess) return ::m3s::cuda_fail(e__, #call); \
    } while (0)

#define M3S_REQUIRE(cond, msg)              \
    do {                                    \
        if (!(cond)) {                      \
            ::m3s::set_error("%s", msg);    \
            return M3S_ERR_INVALID;         \
        }                                   \
    } while (0)

inline int launch_ok(const char* name) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, name);
    return M3S_OK;
}

// number of SMs of the current device (cached per device)
int sm_count_current();

static inline cudaStream_t as_stream(m3s_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

// optional device counters of the tile kernels' fallback paths (m3s_set_path_counters; null = off):
//   [0] refine_matches queries scanned from global memory instead of the staged window
//   [1] iter_proj points of tiles that staged no window (bounding box of the starts too large)
//   [2] iter_proj points of windowed tiles that left the window at least once
unsigned long long* path_counters();

// prep_for_iter_proj + iter_proj fused into one kernel (match_exact.cu); device pointers, p_out
// [B,H*W,2] f32, valid_out [B,H*W] u8.  M3S_ERR_UNSUPPORTED => use the separate entry points.
int iter_proj_fused(const float* X11, const float* X21, const int32_t* idx_init, float* p_out,
                    uint8_t* valid_out, int B, int H, int W, int max_iter, float lambda_init,
                    float convergence_thresh, void* workspace, size_t workspace_bytes,
                    m3s_stream_t stream);

}  // namespace m3s
