// refine_matches: local argmax of descriptor dot products (backends/mpsgraph/kernels.py:496-537).
//
// Semantics (numpy path): window (2r+1)^2 around trunc(p1), step 1 (dilation_max has no effect in
// the reference's numpy implementation: every level restarts from the original p1 and the last
// level, dilation 1, overwrites the result), scan dy outer / dx inner, strict '>', out-of-image
// candidates skipped, centre kept when no candidate is inside the image.
//
// Scores are fp32 sums over D products.  The reference's np.dot is BLAS sdot whose summation order
// is unspecified; every kernel here uses ONE canonical order -- four component-wise FMA chains over
// the float4 chunks, combined as (x + z) + (y + w) -- so all code paths agree bit for bit with each
// other; for descriptors on a dyadic grid every order is exact and the argmax equals the
// reference's unconditionally.
//
// Two implementations:
//  * refine_tiled_kernel (full-frame matching, N == H*W, D == 24, r <= 3): the hot one.
//    CTA = 32 x 16 queries (128 threads, 4 queries per thread stacked in y, lanes along x).  The
//    candidate neighbourhood of the whole tile (40 x 24 pixels of D11) is staged once in shared
//    memory (96-B pixel pitch, chunk index XOR-swizzled => conflict-free LDS.128 across lanes); each
//    candidate descriptor a thread loads is reused by up to 4 queries from registers (FFMA2 packed
//    FMAs).  A query whose window does not fit the staged tile (incoherent flow) takes the generic
//    global-memory path, so the result never depends on the tiling.
//  * refine_kernel_vec / _generic: one thread per query straight from global memory/L1 (any N, D, r).
#include <limits.h>

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

struct Acc4 {
    float2 xy, zw;
    __device__ __forceinline__ Acc4() : xy(make_float2(0.f, 0.f)), zw(make_float2(0.f, 0.f)) {}
    __device__ __forceinline__ void fma4(const float4& a, const float4& b) {
#if defined(M3S_REFINE_SCALAR_FMA)  // A/B experiment: same arithmetic with four scalar FFMA
        xy.x = fmaf(a.x, b.x, xy.x); xy.y = fmaf(a.y, b.y, xy.y);
        zw.x = fmaf(a.z, b.z, zw.x); zw.y = fmaf(a.w, b.w, zw.y);
#elif defined(M3S_REFINE_ASM_FMA)  // A/B experiment: volatile asm keeps the interleaved chain order
        unsigned long long axy, azw, bxy, bzw, cxy, czw;
        asm("mov.b64 %0, {%1, %2};" : "=l"(axy) : "f"(a.x), "f"(a.y));
        asm("mov.b64 %0, {%1, %2};" : "=l"(azw) : "f"(a.z), "f"(a.w));
        asm("mov.b64 %0, {%1, %2};" : "=l"(bxy) : "f"(b.x), "f"(b.y));
        asm("mov.b64 %0, {%1, %2};" : "=l"(bzw) : "f"(b.z), "f"(b.w));
        asm("mov.b64 %0, {%1, %2};" : "=l"(cxy) : "f"(xy.x), "f"(xy.y));
        asm("mov.b64 %0, {%1, %2};" : "=l"(czw) : "f"(zw.x), "f"(zw.y));
        asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(cxy) : "l"(axy), "l"(bxy));
        asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(czw) : "l"(azw), "l"(bzw));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(xy.x), "=f"(xy.y) : "l"(cxy));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(zw.x), "=f"(zw.y) : "l"(czw));
#else
        xy = __ffma2_rn(make_float2(a.x, a.y), make_float2(b.x, b.y), xy);
        zw = __ffma2_rn(make_float2(a.z, a.w), make_float2(b.z, b.w), zw);
#endif
    }
    __device__ __forceinline__ float sum() const {
        const float2 t = __fadd2_rn(xy, zw);
        return t.x + t.y;
    }
};

// generic per-query scan from global memory (canonical summation order)
template <int D4>
__device__ __forceinline__ void scan_global(const float4* __restrict__ img, const float4 (&q)[D4],
                                            int cx, int cy, int radius, int H, int W, int& bx,
                                            int& by) {
    float best = -INFINITY;
    bx = cx;
    by = cy;
    const int y_lo = max(cy - radius, 0), y_hi = min(cy + radius, H - 1);
    const int x_lo = max(cx - radius, 0), x_hi = min(cx + radius, W - 1);
    for (int ny = y_lo; ny <= y_hi; ++ny) {
        const float4* __restrict__ row = img + ((size_t)ny * W) * D4;
        for (int nx = x_lo; nx <= x_hi; ++nx) {
            const float4* __restrict__ cd = row + (size_t)nx * D4;
            Acc4 a;
#pragma unroll
            for (int k = 0; k < D4; ++k) a.fma4(q[k], __ldg(cd + k));
            const float sc = a.sum();
            if (sc > best) { best = sc; bx = nx; by = ny; }
        }
    }
}

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
    int bx, by;
    scan_global<D4>(img, q, c.x, c.y, radius, H, W, bx, by);
    reinterpret_cast<int2*>(p1_out)[t] = make_int2(bx, by);
}

// Any descriptor dimension (scalar loads; same canonical order: component j = k % 4).
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
            float a[4] = {0.f, 0.f, 0.f, 0.f};
            for (int k = 0; k < D; ++k) a[k & 3] = fmaf(__ldg(qp + k), __ldg(cd + k), a[k & 3]);
            const float sc = (a[0] + a[2]) + (a[1] + a[3]);
            if (sc > best) { best = sc; bx = nx; by = ny; }
        }
    }
    reinterpret_cast<int2*>(p1_out)[t] = make_int2(bx, by);
}

// ------------------------------------------------------------------------------------------------
// Tiled kernel.
// ------------------------------------------------------------------------------------------------
// Tile shape: 32 x 16 queries, 40 x 24 window, 4 warps, 2 CTAs/SM.  Pixel pitch in shared memory:
//   default              96 B, the chunk index XOR-swizzled by bit 2 of the window column
//   -DM3S_REFINE_PITCH112 112 B (6 chunks + 16 B pad), no swizzle (round 1)
// (A 32 x 12 tile with 3 CTAs/SM does not work: 9 warps put 3 on one SM sub-partition, whose 16 K
// registers then cap a thread at 168 -- the scan needs ~200 and spills: 2.45 ms.)
// -DM3S_REFINE_TILEY=20 -DM3S_REFINE_WINY=28: 5 warps per CTA, 2 CTAs/SM = 10 warps (A/B, see profiles/)
#ifndef M3S_REFINE_TILEY
#define M3S_REFINE_TILEY 16
#endif
#ifndef M3S_REFINE_WINY
#define M3S_REFINE_WINY (M3S_REFINE_TILEY + 8)
#endif
constexpr int kTileX = 32, kTileY = M3S_REFINE_TILEY, kQ = 4;  // queries per CTA: 32 x 16, 4 per thread (in y)
constexpr int kWinX = 40, kWinY = M3S_REFINE_WINY;            // staged candidate window (pixels)
static_assert(kTileY % kQ == 0 && kWinY >= kTileY + 6, "refine tile shape");
#if defined(M3S_REFINE_PITCH112)
constexpr bool kSwz = false;
#else
constexpr bool kSwz = true;
#endif
#ifndef M3S_REFINE_MINBLOCKS
#define M3S_REFINE_MINBLOCKS 2
#endif
constexpr int kTiledMinBlocks = M3S_REFINE_MINBLOCKS;
constexpr int kTiledThreads = 32 * kTileY / kQ;
// TIMING-ONLY experiments (results are wrong): how long is the scan without the staging / query traffic?
#ifdef M3S_REFINE_TIMING_NOSTAGE
constexpr bool kTimingNoStage = true;
#else
constexpr bool kTimingNoStage = false;
#endif
#ifdef M3S_REFINE_TIMING_NOQ
constexpr bool kTimingNoQ = true;
#else
constexpr bool kTimingNoQ = false;
#endif
constexpr int kMaxTiledRadius = 3;                   // 32 + 2*3 <= 40, tile rows + 2*3 <= window rows

template <int D4>
struct TiledCfg {
    // float4 per pixel: + 16 B pad (lane stride 7 chunks: conflict-free LDS.128 across lanes), or no pad
    // and chunk d of window column c stored in slot d ^ ((c >> 2) & 1): lanes c..c+7 then hit the
    // eight 16-B bank groups {0,6,4,2} + d and {1,7,5,3} + d -- also conflict-free, 14 % less smem
    static constexpr int pitch4 = kSwz ? D4 : D4 + 1;
    // + one padded row: lanes walk the warp-wide column count from their own base without a clamp,
    // so reads may run past the end of the last window row (values discarded by the column mask)
    // (a lane's base column is <= kWinX - 7, the warp-wide column count <= 32, the software pipeline
    // looks one pair ahead: <= 28 pixels past the end)
    static constexpr int pad_px = kWinX + 4;
    static constexpr size_t smem_bytes = (size_t)(kWinX * kWinY + pad_px) * pitch4 * 16;
};

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory");
}

// One (warp-uniform) window row for the queries selected by the compile-time mask KMASK:
// straight-line code, no per-candidate branches.  Every lane walks NC columns of ITS OWN window
// (lane-relative column i -> smem column min(xb + i, kWinX - 1)); a candidate descriptor is D4
// conflict-free LDS.128 and is reused from registers by every active query.  Lanes / queries for
// which the candidate is outside their own (2r+1)^2 window compute a dot product that the compare
// predicate discards.  Position packed as (row << 16) | col in lane-relative window coordinates.
//
// Candidates are evaluated TWO AT A TIME (columns i, i+1): with 8 warps per SM the FMA pipe only
// stays busy if every warp carries enough independent FFMA2 chains, and the top / bottom rows of
// the union window serve a single query (2 chains).  A pair gives >= 4 chains for every mask.  The
// pair after the current one is fetched chunk by chunk into the registers the current pair has
// just finished with (c[d] is dead after chunk d), so its LDS latency is covered by the remaining
// chunks and the score epilogue -- a software pipeline without a second register buffer.
// pointer(s) to the D4 chunks of window column `col` of a row: chunk d is e[d] for even d, o[d] for odd d
// (e == o without the swizzle)
template <int D4>
struct Cand {
    const float4* __restrict__ e;
    const float4* __restrict__ o;
    __device__ __forceinline__ float4 operator[](int d) const { return (d & 1) ? o[d] : e[d]; }
};
template <int D4>
__device__ __forceinline__ Cand<D4> cand_ptr(const float4* __restrict__ row, int col) {
    const float4* p = row + col * TiledCfg<D4>::pitch4;
    if constexpr (kSwz) {
        const int s = (col >> 2) & 1;
        return Cand<D4>{p + s, p - s};
    } else {
        return Cand<D4>{p, p};
    }
}

template <unsigned KMASK>
__device__ __forceinline__ void cand_pick(const Acc4 (&a)[kQ], int pos, unsigned bit,
                                          const unsigned (&rmask)[kQ], float (&best)[kQ],
                                          int (&bpos)[kQ]) {
#pragma unroll
    for (int k = 0; k < kQ; ++k) {
        if ((KMASK >> k) & 1u) {
            const float sc = a[k].sum();
            // rmask[k]: columns of this row inside query k's own window (0 when the row is not);
            // bitwise & (no short-circuit): keeps this a select, not a branch
            const bool take = ((rmask[k] & bit) != 0u) & (sc > best[k]);
            best[k] = take ? sc : best[k];
            bpos[k] = take ? pos : bpos[k];
        }
    }
}

// evaluate the pair (c0 = column i, c1 = column i+1; c1 skipped when !SECOND) and refill c0 / c1
// from n0 / n1 (the next pair) as their chunks retire
template <int D4, unsigned KMASK, bool SECOND>
__device__ __forceinline__ void pair_eval(float4 (&c0)[D4], float4 (&c1)[D4],
                                          const Cand<D4> n0, const Cand<D4> n1,
                                          int i, int jbits, const float4 (&q)[kQ][D4],
                                          const unsigned (&rmask)[kQ], float (&best)[kQ],
                                          int (&bpos)[kQ]) {
    Acc4 a0[kQ], a1[kQ];
#pragma unroll
    for (int d = 0; d < D4; ++d) {
#pragma unroll
        for (int k = 0; k < kQ; ++k) {
            if ((KMASK >> k) & 1u) {
                a0[k].fma4(q[k][d], c0[d]);
                if (SECOND) a1[k].fma4(q[k][d], c1[d]);
            }
        }
        c0[d] = n0[d];
        c1[d] = n1[d];
    }
    const unsigned bit = 1u << i;
    cand_pick<KMASK>(a0, jbits | i, bit, rmask, best, bpos);  // scan order: i before i+1
    if (SECOND) cand_pick<KMASK>(a1, jbits | (i + 1), bit << 1, rmask, best, bpos);
}

// `row` = this window row, `nrow` = the next one (its first pair is prefetched by the last pair here)
template <int D4, unsigned KMASK>
__device__ __forceinline__ void row_scan(const float4* __restrict__ row, const float4* __restrict__ nrow,
                                         int xb, int NC, int j, float4 (&c0)[D4], float4 (&c1)[D4],
                                         const float4 (&q)[kQ][D4], const unsigned (&rmask)[kQ],
                                         float (&best)[kQ], int (&bpos)[kQ]) {
    const int jbits = j << 16;
    int i = 0;
    for (; i + 2 < NC; i += 2)
        pair_eval<D4, KMASK, true>(c0, c1, cand_ptr<D4>(row, xb + i + 2), cand_ptr<D4>(row, xb + i + 3),
                                   i, jbits, q, rmask, best, bpos);
    const Cand<D4> n0 = cand_ptr<D4>(nrow, xb);
    const Cand<D4> n1 = cand_ptr<D4>(nrow, xb + 1);
    if (i + 1 < NC)
        pair_eval<D4, KMASK, true>(c0, c1, n0, n1, i, jbits, q, rmask, best, bpos);
    else
        pair_eval<D4, KMASK, false>(c0, c1, n0, n1, i, jbits, q, rmask, best, bpos);
}

// Default scan: ONE candidate per evaluation, two register buffers (cA, cB) in ping-pong: the loads
// of the candidate after next are issued before the current one is evaluated.  Measured on B200 (64
// pairs, profiles/r2_refine_ab.txt): 1.459 ms against 1.529 ms for the pair evaluation of round 1
// (row_scan, -DM3S_REFINE_PAIR) -- 169 instead of 185 instructions per two candidates and ~200
// instead of 220 registers.  Invariant on entry and exit: cA holds column 0 of the row to scan.
template <int D4, unsigned KMASK>
__device__ __forceinline__ void cand_eval(const float4 (&c)[D4], int i, int jbits, const float4 (&q)[kQ][D4],
                                          const unsigned (&rmask)[kQ], float (&best)[kQ], int (&bpos)[kQ]) {
    Acc4 a[kQ];
#pragma unroll
    for (int d = 0; d < D4; ++d) {
#pragma unroll
        for (int k = 0; k < kQ; ++k)
            if ((KMASK >> k) & 1u) a[k].fma4(q[k][d], c[d]);
    }
    cand_pick<KMASK>(a, jbits | i, 1u << i, rmask, best, bpos);
}
template <int D4>
__device__ __forceinline__ void cand_load(float4 (&c)[D4], const Cand<D4> p) {
#pragma unroll
    for (int d = 0; d < D4; ++d) c[d] = p[d];
}
template <int D4, unsigned KMASK>
__device__ __forceinline__ void row_scan_pp(const float4* __restrict__ row, const float4* __restrict__ nrow,
                                            int xb, int NC, int j, float4 (&cA)[D4], float4 (&cB)[D4],
                                            const float4 (&q)[kQ][D4], const unsigned (&rmask)[kQ],
                                            float (&best)[kQ], int (&bpos)[kQ]) {
    const int jbits = j << 16;
    int i = 0;
    for (; i + 2 < NC; i += 2) {
        cand_load<D4>(cB, cand_ptr<D4>(row, xb + i + 1));
        cand_eval<D4, KMASK>(cA, i, jbits, q, rmask, best, bpos);
        cand_load<D4>(cA, cand_ptr<D4>(row, xb + i + 2));
        cand_eval<D4, KMASK>(cB, i + 1, jbits, q, rmask, best, bpos);
    }
    if (i + 1 < NC) {
        cand_load<D4>(cB, cand_ptr<D4>(row, xb + i + 1));
        cand_eval<D4, KMASK>(cA, i, jbits, q, rmask, best, bpos);
        cand_load<D4>(cA, cand_ptr<D4>(nrow, xb));
        cand_eval<D4, KMASK>(cB, i + 1, jbits, q, rmask, best, bpos);
    } else {
        cand_load<D4>(cB, cand_ptr<D4>(nrow, xb));
        cand_eval<D4, KMASK>(cA, i, jbits, q, rmask, best, bpos);
#pragma unroll
        for (int d = 0; d < D4; ++d) cA[d] = cB[d];
    }
}

template <int D4, typename PT>
__global__ void __launch_bounds__(kTiledThreads, kTiledMinBlocks)
refine_tiled_kernel(const float* __restrict__ D11, const float* __restrict__ D21,
                    const PT* __restrict__ p1, int32_t* __restrict__ p1_out, int H, int W,
                    int radius, unsigned long long* __restrict__ counters) {
    extern __shared__ float4 tile[];  // [kWinY][kWinX][pitch4]
    __shared__ int s_min[2], s_max[2], s_fit;
    constexpr int P4 = TiledCfg<D4>::pitch4;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.z;
    const int qx = blockIdx.x * kTileX + lane;
    const int qy0 = blockIdx.y * kTileY + warp * kQ;
    const size_t N = (size_t)H * W;
    const float4* __restrict__ img = reinterpret_cast<const float4*>(D11) + (size_t)b * N * D4;

    if (threadIdx.x == 0) {
        s_min[0] = s_min[1] = INT_MAX;
        s_max[0] = s_max[1] = INT_MIN;
        s_fit = 0;
    }
    // centres first (they place the window); query descriptors are fetched while the tile streams in
    int cx[kQ], cy[kQ];
    bool ok[kQ];
    int mnx = INT_MAX, mny = INT_MAX, mxx = INT_MIN, mxy = INT_MIN;
#pragma unroll
    for (int k = 0; k < kQ; ++k) {
        const int qy = qy0 + k;
        ok[k] = (qx < W) && (qy < H);
        cx[k] = cy[k] = 0;
        if (ok[k]) {
            const int2 c = load_center(p1, (long long)((size_t)b * N + (size_t)qy * W + qx));
            cx[k] = c.x; cy[k] = c.y;
            mnx = min(mnx, c.x); mxx = max(mxx, c.x);
            mny = min(mny, c.y); mxy = max(mxy, c.y);
        }
    }
    __syncthreads();
    mnx = __reduce_min_sync(0xffffffffu, mnx); mny = __reduce_min_sync(0xffffffffu, mny);
    mxx = __reduce_max_sync(0xffffffffu, mxx); mxy = __reduce_max_sync(0xffffffffu, mxy);
    if (lane == 0) {
        atomicMin(&s_min[0], mnx); atomicMin(&s_min[1], mny);
        atomicMax(&s_max[0], mxx); atomicMax(&s_max[1], mxy);
    }
    __syncthreads();
    if (s_min[0] == INT_MAX) return;  // tile entirely outside the image
    // window origin: centre the slack around the bounding box of the centres
    const int ext_x = s_max[0] - s_min[0] + 1 + 2 * radius, ext_y = s_max[1] - s_min[1] + 1 + 2 * radius;
    const int wx0 = s_min[0] - radius - max((kWinX - ext_x) / 2, 0);
    const int wy0 = s_min[1] - radius - max((kWinY - ext_y) / 2, 0);
    bool fit[kQ];
    int nfit = 0, nok = 0;
#pragma unroll
    for (int k = 0; k < kQ; ++k) {
        fit[k] = ok[k] && (cx[k] - radius >= wx0) && (cx[k] + radius < wx0 + kWinX) &&
                 (cy[k] - radius >= wy0) && (cy[k] + radius < wy0 + kWinY);
        nfit += fit[k];
        nok += ok[k];
    }
    {
        const int v = __reduce_add_sync(0xffffffffu, nfit * 4096 + nok);  // <= 128 each per warp, <= 1536 per CTA
        if (lane == 0) atomicAdd(&s_fit, v);
    }
    __syncthreads();
    const bool use_tile = (s_fit >> 12) * 4 >= (s_fit & 4095);  // >= 25 % of the queries fit
    if (use_tile) {
        // asynchronous staging (LDGSTS, no register round trip): a window row is kWinX * D4
        // consecutive float4 in global memory; thread t copies float4 t and t + 128 of every row.
        constexpr int ROW4 = kWinX * D4;  // 240
        constexpr int NCP = (ROW4 + kTiledThreads - 1) / kTiledThreads;  // copies per thread and row
        bool in[NCP];
        int f[NCP];
        unsigned doff[NCP];
#pragma unroll
        for (int c = 0; c < NCP; ++c) {
            f[c] = threadIdx.x + c * kTiledThreads;
            const int px = f[c] / D4, d = f[c] - px * D4;
            in[c] = (f[c] < ROW4) && (wx0 + px >= 0) && (wx0 + px < W);
            doff[c] = (unsigned)(px * P4 + (kSwz ? (d ^ ((px >> 2) & 1)) : d)) * 16u;
        }
        // hoisted addressing: one 64-bit source pointer and one 32-bit shared address per row,
        // advanced by constant strides; only rows inside the image are visited (the per-row
        // address arithmetic of the first version was 11 % of all executed instructions, ncu s15)
        const int wy_lo = max(0, -wy0), wy_hi = min(kWinY, H - wy0);
        const size_t grow = (size_t)W * D4;  // float4 per image row
        // `src` may point before the row start when wx0 < 0; only in-image elements are touched
        const float4* __restrict__ src = img + ((long long)(wy0 + wy_lo) * W + wx0) * D4;
        unsigned d0 = (unsigned)__cvta_generic_to_shared(tile + wy_lo * kWinX * P4);
        constexpr unsigned srow = kWinX * P4 * 16;
#pragma unroll 4
        for (int wy = wy_lo; wy < wy_hi; ++wy) {
#pragma unroll
            for (int c = 0; c < NCP; ++c)
                if (in[c] && !kTimingNoStage)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d0 + doff[c]), "l"(src + f[c]) : "memory");
            src += grow;
            d0 += srow;
        }
    }
    // Query descriptors are fetched HERE, after the window's cp.async are in flight.  Measured on B200
    // (64 pairs, profiles/r2_refine_ab.txt): fetching them before the first barrier costs +2 %
    // (1.567 vs 1.537 ms -- 24 registers live across the staging loop), and warming the L2 for the CTA
    // one wave ahead with prefetch.global.L2 costs +11 % (1.70 ms): the prologue is not DRAM-latency bound.
    float4 q[kQ][D4];
#pragma unroll
    for (int k = 0; k < kQ; ++k) {
        if (ok[k]) {
            const float4* __restrict__ qp =
                reinterpret_cast<const float4*>(D21) + ((size_t)b * N + (size_t)(qy0 + k) * W + qx) * D4;
#pragma unroll
            for (int d = 0; d < D4; ++d)
                q[k][d] = kTimingNoQ ? make_float4((float)lane, (float)k, (float)d, 1.f) : __ldg(qp + d);
        } else {
#pragma unroll
            for (int d = 0; d < D4; ++d) q[k][d] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    cp_async_wait_all();
    __syncthreads();

    float best[kQ];
    int bpos[kQ];
#pragma unroll
    for (int k = 0; k < kQ; ++k) { best[k] = -INFINITY; bpos[k] = -1; }

    int xb = 0, yb = 0;  // this lane's window base inside the staged tile (tile-relative)
    if (use_tile) {
        // lane-relative window geometry
        int X0 = INT_MAX, Y0 = INT_MAX;
#pragma unroll
        for (int k = 0; k < kQ; ++k)
            if (fit[k]) { X0 = min(X0, cx[k] - radius); Y0 = min(Y0, cy[k] - radius); }
        if (X0 == INT_MAX) { X0 = wx0; Y0 = wy0; }  // no fitting query: walk a dummy window
        xb = X0 - wx0;
        yb = Y0 - wy0;
        int nr = 0, nc = 0;
        int ylo[kQ];
        unsigned yspan[kQ], colmask[kQ];
#pragma unroll
        for (int k = 0; k < kQ; ++k) {
            // candidate range of query k clipped to the image, in lane-relative coordinates: rows
            // as (lo, span) (an empty range is lo = 1 << 20), columns as a bit mask (bit i = column
            // X0 + i).  A query whose columns do not fit 32 bits (its centre is > 25 px right of a
            // stacked neighbour's: incoherent flow) goes to the global path.
            const int ya = max(cy[k] - radius, 0), yz = min(cy[k] + radius, H - 1);
            const int xa = max(cx[k] - radius, 0), xz = min(cx[k] + radius, W - 1);
            if (fit[k] && xz - X0 >= 32) fit[k] = false;
            const bool any = fit[k] && ya <= yz && xa <= xz;
            ylo[k] = any ? ya - Y0 : (1 << 20);
            yspan[k] = any ? (unsigned)(yz - ya) : 0u;
            colmask[k] = any ? ((0xffffffffu >> (31 - (xz - xa))) << (xa - X0)) : 0u;
            if (any) { nr = max(nr, yz - Y0 + 1); nc = max(nc, xz - X0 + 1); }
        }
        const int NR = __reduce_max_sync(0xffffffffu, nr), NC = __reduce_max_sync(0xffffffffu, nc);
        float4 c0[D4], c1[D4];  // the candidate pair in flight (columns i, i+1 of the current row)
        {
            const float4* __restrict__ r0 = tile + (min(yb, kWinY - 1) * kWinX) * P4;
            const Cand<D4> p0 = cand_ptr<D4>(r0, xb);
            const Cand<D4> p1 = cand_ptr<D4>(r0, xb + 1);
#pragma unroll
            for (int d = 0; d < D4; ++d) { c0[d] = p0[d]; c1[d] = p1[d]; }
        }
        for (int j = 0; j < NR; ++j) {
            unsigned rmask[kQ];
            unsigned kmask = 0;
#pragma unroll
            for (int k = 0; k < kQ; ++k) {
                const bool rowok = (unsigned)(j - ylo[k]) <= yspan[k];
                rmask[k] = rowok ? colmask[k] : 0u;
                if (__any_sync(0xffffffffu, rowok)) kmask |= 1u << k;
            }
            const float4* __restrict__ row = tile + (min(yb + j, kWinY - 1) * kWinX) * P4;
            const float4* __restrict__ nrow = tile + (min(yb + j + 1, kWinY - 1) * kWinX) * P4;
#ifdef M3S_REFINE_PAIR
#define M3S_ROW(MASK) case MASK: row_scan<D4, MASK>(row, nrow, xb, NC, j, c0, c1, q, rmask, best, bpos); break
#else
#define M3S_ROW(MASK) case MASK: row_scan_pp<D4, MASK>(row, nrow, xb, NC, j, c0, c1, q, rmask, best, bpos); break
#endif
            switch (kmask) {  // warp-uniform
                case 0: {  // no query uses this row: just advance the pipeline to the next row
                    const Cand<D4> p0 = cand_ptr<D4>(nrow, xb);
                    const Cand<D4> p1 = cand_ptr<D4>(nrow, xb + 1);
#pragma unroll
                    for (int d = 0; d < D4; ++d) { c0[d] = p0[d]; c1[d] = p1[d]; }
                    break;
                }
                M3S_ROW(1); M3S_ROW(2); M3S_ROW(3); M3S_ROW(4); M3S_ROW(5); M3S_ROW(6); M3S_ROW(7);
                M3S_ROW(8); M3S_ROW(9); M3S_ROW(10); M3S_ROW(11); M3S_ROW(12); M3S_ROW(13);
                M3S_ROW(14); M3S_ROW(15);
            }
#undef M3S_ROW
        }
    }
    // write-back; queries whose window is not covered by the staged tile take the global path
    if (counters != nullptr) {
        int ng = 0;
#pragma unroll
        for (int k = 0; k < kQ; ++k) ng += (ok[k] && !(use_tile && fit[k])) ? 1 : 0;
        ng = __reduce_add_sync(0xffffffffu, ng);
        if (lane == 0 && ng) atomicAdd(counters, (unsigned long long)ng);
    }
#pragma unroll
    for (int k = 0; k < kQ; ++k) {
        if (!ok[k]) continue;
        int bx, by;
        if (use_tile && fit[k]) {
            bx = (bpos[k] < 0) ? cx[k] : (wx0 + xb + (bpos[k] & 0xffff));
            by = (bpos[k] < 0) ? cy[k] : (wy0 + yb + (bpos[k] >> 16));
        } else {
            scan_global<D4>(img, q[k], cx[k], cy[k], radius, H, W, bx, by);
        }
        const size_t t = (size_t)b * N + (size_t)(qy0 + k) * W + qx;
        reinterpret_cast<int2*>(p1_out)[t] = make_int2(bx, by);
    }
}

template <typename PT>
int refine_dispatch(const float* D11, const float* D21, const PT* p1, int32_t* p1_out, int B,
                    int H, int W, int D, int N, int radius, cudaStream_t st) {
    const long long total = (long long)B * N;
    if (total == 0) return M3S_OK;
    const bool aligned = ((reinterpret_cast<uintptr_t>(D11) | reinterpret_cast<uintptr_t>(D21)) & 15) == 0;
    if (aligned && D == 24 && N == H * W && radius >= 1 && radius <= kMaxTiledRadius && B <= 65535) {
        constexpr size_t smem = TiledCfg<6>::smem_bytes;
        M3S_CUDA(cudaFuncSetAttribute(refine_tiled_kernel<6, PT>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const dim3 grid((W + kTileX - 1) / kTileX, (H + kTileY - 1) / kTileY, B);
        refine_tiled_kernel<6, PT><<<grid, kTiledThreads, smem, st>>>(D11, D21, p1, p1_out, H, W, radius,
                                                                       path_counters());
        return launch_ok("refine_tiled_kernel");
    }
    const unsigned grid = (unsigned)((total + kRefineThreads - 1) / kRefineThreads);
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
