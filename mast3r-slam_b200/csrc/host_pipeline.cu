// Host-buffer entry point of match_iterative_proj: what the reference's caller actually holds are
// HOST arrays (mast3r_utils.py:490 hands numpy-convertible arrays to matching.match), so the end to
// end cost of the path is dominated by the PCIe transfer of the inputs (216 B/px in, 5 B/px out),
// not by the kernels.  This entry point hides everything else behind that transfer:
//
//   the batch is cut into chunks of `chunk_pairs` frame pairs; chunk c runs on slot c % kSlots, each
//   slot owning a CUDA stream, device buffers for one chunk and the kernels' workspace.  Per chunk:
//   H2D of the four inputs (async, copy engine 0) -> prep / iter_proj / refine / tail (SMs) ->
//   D2H of idx + valid (async, copy engine 1).  Consecutive chunks are on different streams, so
//   chunk c+1 is uploading while chunk c computes and chunk c-1 downloads.  Stream order inside a
//   slot makes the buffer reuse safe without extra events.
//
// Host pointers should be page-locked (cudaHostAlloc / cudaHostRegister / torch pin_memory) for
// the copies to be asynchronous; pageable memory still gives correct results (the runtime stages
// it), just without the overlap.  The call returns after the last D2H completed.
#include <mutex>

#include "m3s_common.cuh"

namespace m3s {

constexpr int kSlots = 3;

struct Slot {
    cudaStream_t stream = nullptr;
    char* buf = nullptr;     // X11 | X21 | D11 | D21 | idx_init | idx | valid | workspace
    size_t bytes = 0;
};

struct HostPipe {
    std::mutex mu;
    int device = -1;
    Slot slots[kSlots];
};

static HostPipe g_pipes[16];

static inline size_t up256(size_t v) { return (v + 255) & ~(size_t)255; }

static void release_pipe(HostPipe& hp) {
    for (Slot& s : hp.slots) {
        if (s.buf) cudaFree(s.buf);
        if (s.stream) cudaStreamDestroy(s.stream);
        s = Slot();
    }
    hp.device = -1;
}

}  // namespace m3s

using namespace m3s;

M3S_API int m3s_match_iterative_proj_host(const float* X11_host, const float* X21_host,
                                          const float* D11_host, const float* D21_host,
                                          const int32_t* idx_init_host, int32_t* idx_out_host,
                                          uint8_t* valid_out_host, int B, int H, int W, int D,
                                          int max_iter, float lambda_init, float convergence_thresh,
                                          int radius, int dilation_max, float dist_thresh,
                                          int chunk_pairs, int device) {
    M3S_REQUIRE(X11_host && X21_host && idx_out_host && valid_out_host, "match_host: null pointer");
    M3S_REQUIRE(radius <= 0 || (D11_host && D21_host), "match_host: descriptors required");
    M3S_REQUIRE(B >= 0 && H > 0 && W > 0 && D >= 0, "match_host: bad shape");
    M3S_REQUIRE(device >= 0 && device < 16, "match_host: bad device ordinal");
    if (B == 0) return M3S_OK;
    if (chunk_pairs <= 0) chunk_pairs = 4;
    if (chunk_pairs > B) chunk_pairs = B;

    int cur = -1;
    M3S_CUDA(cudaGetDevice(&cur));
    if (cur != device) M3S_CUDA(cudaSetDevice(device));
    struct Restore { int d, want; ~Restore() { if (d >= 0 && d != want) cudaSetDevice(d); } } restore{cur, device};

    HostPipe& hp = g_pipes[device];
    std::lock_guard<std::mutex> lock(hp.mu);
    hp.device = device;

    const size_t npx = (size_t)H * W;
    const bool use_desc = radius > 0;
    // per-chunk section sizes
    const size_t cX = up256((size_t)chunk_pairs * npx * 3 * 4);
    const size_t cD = use_desc ? up256((size_t)chunk_pairs * npx * D * 4) : 0;
    const size_t cI = up256((size_t)chunk_pairs * npx * 4);
    const size_t cV = up256((size_t)chunk_pairs * npx);
    const size_t cW = up256(m3s_match_workspace_bytes(chunk_pairs, H, W, max_iter));
    const size_t need = 2 * cX + 2 * cD + 2 * cI + cV + cW;

    const int nchunks = (B + chunk_pairs - 1) / chunk_pairs;
    const int nslots = nchunks < kSlots ? nchunks : kSlots;
    for (int s = 0; s < nslots; ++s) {
        Slot& sl = hp.slots[s];
        if (!sl.stream) M3S_CUDA(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
        if (sl.bytes < need) {
            if (sl.buf) { M3S_CUDA(cudaStreamSynchronize(sl.stream)); cudaFree(sl.buf); sl.buf = nullptr; sl.bytes = 0; }
            M3S_CUDA(cudaMalloc(reinterpret_cast<void**>(&sl.buf), need));
            sl.bytes = need;
        }
    }

    int rc = M3S_OK;
    for (int c = 0; c < nchunks && rc == M3S_OK; ++c) {
        Slot& sl = hp.slots[c % kSlots];
        const int b0 = c * chunk_pairs;
        const int nb = (b0 + chunk_pairs <= B) ? chunk_pairs : B - b0;
        const size_t px0 = (size_t)b0 * npx, npc = (size_t)nb * npx;
        char* w = sl.buf;
        float* dX11 = reinterpret_cast<float*>(w); w += cX;
        float* dX21 = reinterpret_cast<float*>(w); w += cX;
        float* dD11 = reinterpret_cast<float*>(w); w += cD;
        float* dD21 = reinterpret_cast<float*>(w); w += cD;
        int32_t* dInit = reinterpret_cast<int32_t*>(w); w += cI;
        int32_t* dIdx = reinterpret_cast<int32_t*>(w); w += cI;
        uint8_t* dValid = reinterpret_cast<uint8_t*>(w); w += cV;
        void* ws = w;
        cudaStream_t st = sl.stream;
        M3S_CUDA(cudaMemcpyAsync(dX11, X11_host + px0 * 3, npc * 3 * 4, cudaMemcpyHostToDevice, st));
        M3S_CUDA(cudaMemcpyAsync(dX21, X21_host + px0 * 3, npc * 3 * 4, cudaMemcpyHostToDevice, st));
        if (use_desc) {
            M3S_CUDA(cudaMemcpyAsync(dD11, D11_host + px0 * D, npc * D * 4, cudaMemcpyHostToDevice, st));
            M3S_CUDA(cudaMemcpyAsync(dD21, D21_host + px0 * D, npc * D * 4, cudaMemcpyHostToDevice, st));
        }
        if (idx_init_host)
            M3S_CUDA(cudaMemcpyAsync(dInit, idx_init_host + px0, npc * 4, cudaMemcpyHostToDevice, st));
        rc = m3s_match_iterative_proj(dX11, dX21, use_desc ? dD11 : nullptr, use_desc ? dD21 : nullptr,
                                      idx_init_host ? dInit : nullptr, dIdx, dValid, nb, H, W, D,
                                      max_iter, lambda_init, convergence_thresh, radius, dilation_max,
                                      dist_thresh, ws, cW, reinterpret_cast<m3s_stream_t>(st));
        if (rc) break;
        M3S_CUDA(cudaMemcpyAsync(idx_out_host + px0, dIdx, npc * 4, cudaMemcpyDeviceToHost, st));
        M3S_CUDA(cudaMemcpyAsync(valid_out_host + px0, dValid, npc, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < nslots; ++s) {
        const cudaError_t e = cudaStreamSynchronize(hp.slots[s].stream);
        if (e != cudaSuccess && rc == M3S_OK) rc = cuda_fail(e, "match_host: stream synchronize");
    }
    return rc;
}

M3S_API int m3s_match_host_release(int device) {
    M3S_REQUIRE(device >= 0 && device < 16, "match_host_release: bad device ordinal");
    HostPipe& hp = g_pipes[device];
    std::lock_guard<std::mutex> lock(hp.mu);
    if (hp.device < 0) return M3S_OK;
    int cur = -1;
    cudaGetDevice(&cur);
    if (cur != device) cudaSetDevice(device);
    release_pipe(hp);
    if (cur >= 0 && cur != device) cudaSetDevice(cur);
    return M3S_OK;
}
