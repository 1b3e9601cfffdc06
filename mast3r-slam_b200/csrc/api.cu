// Library plumbing (errors, device queries) and the composite match_iterative_proj entry point.
#include <stdarg.h>

#include "m3s_common.cuh"

namespace m3s {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return M3S_ERR_CUDA;
}

int sm_count_current() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

static unsigned long long* g_path_counters = nullptr;
unsigned long long* path_counters() { return g_path_counters; }

}  // namespace m3s

using namespace m3s;

M3S_API int m3s_version(void) { return 1; }

M3S_API int m3s_set_path_counters(unsigned long long* dev_counters) {
    g_path_counters = dev_counters;
    return M3S_OK;
}

M3S_API const char* m3s_last_error(void) { return g_err; }

M3S_API int m3s_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

M3S_API int m3s_sm_count(int device) {
    int n = 0;
    M3S_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
    return n;
}

// ------------------------------------------------------------------------------------------------
// match_iterative_proj (matching.py:339-461) as one enqueue: prep -> iter_proj -> refine -> tail.
// Workspace layout (256-B aligned sections): rays [T,9] f32 | pts [T,3] | p_init [T,2] |
// p [T,2] f32 | p1 [T,2] i32 | valid_proj [T] u8 | iter_proj scratch.      T = B*H*W
// ------------------------------------------------------------------------------------------------
static inline size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

M3S_API size_t m3s_match_workspace_bytes(int B, int H, int W, int max_iter) {
    const size_t T = (size_t)B * H * W;
    return align_up(T * 9 * 4) + align_up(T * 3 * 4) + align_up(T * 2 * 4) + align_up(T * 2 * 4) +
           align_up(T * 2 * 4) + align_up(T) + align_up(m3s_iter_proj_workspace_bytes(max_iter));
}

M3S_API int m3s_match_iterative_proj(const float* X11, const float* X21, const float* D11,
                                     const float* D21, const int32_t* idx_init, int32_t* idx_out,
                                     uint8_t* valid_out, int B, int H, int W, int D, int max_iter,
                                     float lambda_init, float convergence_thresh, int radius,
                                     int dilation_max, float dist_thresh, void* workspace,
                                     size_t workspace_bytes, m3s_stream_t stream) {
    M3S_REQUIRE(X11 && X21 && idx_out && valid_out, "match_iterative_proj: null pointer");
    M3S_REQUIRE(radius <= 0 || (D11 && D21), "match_iterative_proj: descriptors required");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0, "match_iterative_proj: bad shape");
    if (B == 0) return M3S_OK;
    if (!workspace || workspace_bytes < m3s_match_workspace_bytes(B, H, W, max_iter)) {
        set_error("match_iterative_proj: workspace too small (%zu < %zu)", workspace_bytes,
                  m3s_match_workspace_bytes(B, H, W, max_iter));
        return M3S_ERR_WORKSPACE;
    }
    const size_t T = (size_t)B * H * W;
    const int N = H * W;
    char* w = reinterpret_cast<char*>(workspace);
    float* rays = reinterpret_cast<float*>(w); w += align_up(T * 9 * 4);
    float* pts = reinterpret_cast<float*>(w); w += align_up(T * 3 * 4);
    float* p_init = reinterpret_cast<float*>(w); w += align_up(T * 2 * 4);
    float* p = reinterpret_cast<float*>(w); w += align_up(T * 2 * 4);
    int32_t* p1 = reinterpret_cast<int32_t*>(w); w += align_up(T * 2 * 4);
    uint8_t* vproj = reinterpret_cast<uint8_t*>(w); w += align_up(T);
    void* ip_ws = w;
    // prep fused into the iter_proj tile kernel (the ray image never reaches HBM) is
    // opt-in (M3S_MATCH_FUSED=1: measured slower on B200, see match_exact.cu); otherwise, and for the
    // shapes the fused kernel does not take, the two separate kernels run
    int rc = m3s::iter_proj_fused(X11, X21, idx_init, p, vproj, B, H, W, max_iter, lambda_init,
                                  convergence_thresh, ip_ws, m3s_iter_proj_workspace_bytes(max_iter),
                                  stream);
    if (rc == M3S_ERR_UNSUPPORTED) {
        rc = m3s_prep_iter_proj(X11, X21, idx_init, rays, pts, p_init, B, H, W, stream);
        if (rc) return rc;
        rc = m3s_iter_proj(rays, pts, p_init, p, vproj, B, H, W, N, max_iter, lambda_init,
                           convergence_thresh, ip_ws, m3s_iter_proj_workspace_bytes(max_iter), stream);
    }
    if (rc) return rc;
    if (radius > 0) {
        rc = m3s_refine_matches_f32(D11, D21, p, p1, B, H, W, D, N, radius, dilation_max, stream);
    } else {
        // use_refine == False / radius 0: p1 = trunc(p) (matching.py:436); radius-0 refine with
        // a null-free descriptor path would need D; a truncating copy is the same result.
        rc = m3s_refine_matches_f32(X11, X11, p, p1, B, H, W, 3, N, -0, dilation_max, stream);
        // NOTE: radius 0 on an in-image centre returns the centre; an out-of-image centre is also
        // returned unchanged, i.e. exactly trunc(p).
    }
    if (rc) return rc;
    return m3s_match_tail(X11, X21, p1, vproj, idx_out, valid_out, B, H, W, dist_thresh, stream);
}

// ------------------------------------------------------------------------------------------------
// m3s_match_iterative_proj_static: the same enqueue with its intermediates in a file-scope device
// array instead of a caller workspace, for callers that may not allocate (stream capture, serving
// loops).  No counterpart in the reference.
//   * limit: B*H*W <= M3S_MATCH_STATIC_MAX_PIXELS (one 512 x 512 pair); the array is 73 B per pixel
//     + section padding = 18.3 MiB, held by every GPU that loads the library until exit
//   * no cudaMalloc / cudaFree / allocating library call on this path
//   * every call on a device shares the array: calls that overlap on the GPU (two streams) race;
//     what must start at zero (the iter_proj early-exit slots) is cleared by m3s_iter_proj on the
//     caller's stream at each call, all other sections are written before they are read
// In host code `g_match_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
// ------------------------------------------------------------------------------------------------
static constexpr size_t kStaticMaxPixels = M3S_MATCH_STATIC_MAX_PIXELS;
static constexpr size_t kStaticScratchBytes = kStaticMaxPixels * 73 + 4096;  // 7 sections, <= 255 B pad each + iter_proj scratch
static __device__ __align__(256) unsigned char g_match_scratch[kStaticScratchBytes];

static int match_scratch_address(void** out) {
    static void* cached[64] = {nullptr};  // per device; a benign race writes the same value twice
    int dev = 0;
    M3S_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        set_error("match_iterative_proj_static: device ordinal %d out of range", dev);
        return M3S_ERR_INVALID;
    }
    void* ptr = __atomic_load_n(&cached[dev], __ATOMIC_ACQUIRE);
    if (!ptr) {
        M3S_CUDA(cudaGetSymbolAddress(&ptr, g_match_scratch));
        if (!ptr) {
            set_error("match_iterative_proj_static: cudaGetSymbolAddress returned a null address");
            return M3S_ERR_CUDA;
        }
        __atomic_store_n(&cached[dev], ptr, __ATOMIC_RELEASE);
    }
    *out = ptr;
    return M3S_OK;
}

M3S_API int m3s_match_iterative_proj_static(const float* X11, const float* X21, const float* D11,
                                            const float* D21, const int32_t* idx_init,
                                            int32_t* idx_out, uint8_t* valid_out, int B, int H, int W,
                                            int D, int max_iter, float lambda_init,
                                            float convergence_thresh, int radius, int dilation_max,
                                            float dist_thresh, m3s_stream_t stream) {
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0, "match_iterative_proj_static: bad shape");
    const unsigned long long T = (unsigned long long)B * (unsigned long long)H * (unsigned long long)W;
    if (T > kStaticMaxPixels || m3s_match_workspace_bytes(B, H, W, max_iter) > kStaticScratchBytes) {
        set_error("match_iterative_proj_static: B*H*W = %llu pixels exceeds the static scratch limit "
                  "of %zu (M3S_MATCH_STATIC_MAX_PIXELS); use m3s_match_iterative_proj with a "
                  "caller workspace", T, kStaticMaxPixels);
        return M3S_ERR_WORKSPACE;
    }
    void* scratch = nullptr;
    if (int rc = match_scratch_address(&scratch)) return rc;
    return m3s_match_iterative_proj(X11, X21, D11, D21, idx_init, idx_out, valid_out, B, H, W, D,
                                    max_iter, lambda_init, convergence_thresh, radius, dilation_max,
                                    dist_thresh, scratch, kStaticScratchBytes, stream);
}
