/*
 * m3s.h -- C ABI of the B200-native (sm_100a) MASt3R-SLAM backend hot path:
 *   iterative projective matching (iter_proj) -> descriptor refine_matches -> match tail, and
 *   Sim3 Gauss-Newton (rays / points / calib): accumulate -> assemble -> damped solve -> retract.
 *
 * This is the drop-in boundary.  Every entry point replaces one function of the reference's
 * operator layer (paths relative to /root/reference/src/mlx_mast3r_slam/), cited per function.
 * The reference is Python: its "FFI" for this path is the numpy-in / numpy-out operator module
 * backends/mpsgraph/kernels.py, which a maintainer would re-point at this library with the ctypes
 * stub shown in INTEGRATION.md (our own copy of that stub is mast3r-slam_b200/_lib.py).
 *
 * Conventions
 *   - plain pointers and sizes only; no torch / numpy types.  Unless a parameter name ends in
 *     `_host`, pointers are DEVICE pointers (cudaMalloc / torch CUDA tensors), densely packed,
 *     row-major, with the layouts of the reference's numpy arrays.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  All work is
 *     enqueued asynchronously on it; no entry point synchronises unless documented.
 *   - return value: 0 = ok, <0 = error (M3S_ERR_*); m3s_last_error() returns a thread-local
 *     message.  There is NO CPU fallback: without a CUDA device every compute entry point fails
 *     with M3S_ERR_CUDA.
 *   - inputs are never modified; outputs are caller-allocated.
 */
#ifndef M3S_H_
#define M3S_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define M3S_OK 0
#define M3S_ERR_INVALID (-1)   /* bad argument (null pointer, non-positive size, ...) */
#define M3S_ERR_CUDA (-2)      /* CUDA runtime error or no device */
#define M3S_ERR_WORKSPACE (-3) /* workspace too small */
#define M3S_ERR_UNSUPPORTED (-4)

#define M3S_GN_RAYS 0   /* gauss_newton_rays   (backends/mpsgraph/gauss_newton.py:23)        */
#define M3S_GN_POINTS 1 /* gauss_newton_points (backends/mpsgraph/gauss_newton_points.py:17) */
#define M3S_GN_CALIB 2  /* gauss_newton_calib  (backends/mpsgraph/gauss_newton_calib.py:17)  */

typedef void* m3s_stream_t;

/* ---- library ------------------------------------------------------------------------------ */
int m3s_version(void);              /* 100 * major + minor */
const char* m3s_last_error(void);   /* thread-local, never NULL */
int m3s_device_count(void);         /* CUDA devices visible to this process (0 if none) */
int m3s_sm_count(int device);       /* multiprocessor count of `device` (<0 on error) */

/* ---- matching ------------------------------------------------------------------------------ */

/* Replaces prep_for_iter_proj (matching.py:134-175) incl. normalize_rays (:121-131),
 * img_gradient (image.py:9-34) and lin_to_pixel (matching.py:106-118).
 *   X11, X21 [B,H,W,3] f32; idx_init [B,H*W] i32 or NULL (identity)
 *   -> rays_out [B,H,W,9] f32 = (ray, d/dx, d/dy), pts_out [B,H*W,3], p_init_out [B,H*W,2]. */
int m3s_prep_iter_proj(const float* X11, const float* X21, const int32_t* idx_init,
                       float* rays_out, float* pts_out, float* p_init_out, int B, int H, int W,
                       m3s_stream_t stream);

/* Scratch bytes m3s_iter_proj needs (per-iteration max-step slots for the global early exit). */
size_t m3s_iter_proj_workspace_bytes(int max_iter);

/* Replaces iter_proj (backends/mpsgraph/kernels.py:107-148), numpy semantics of
 * _iter_proj_numpy (:151-254): <=max_iter Levenberg-Marquardt steps per point on the bilinearly
 * sampled 9-channel ray image, fp64 blend weights, GLOBAL early exit when the largest step of an
 * iteration is < convergence_thresh (resolved on the device: no host synchronisation).
 *   rays [B,H,W,9], pts [B,N,3], p_init [B,N,2]  ->  p_out [B,N,2] f32 (clamped), valid_out [B,N] u8
 * Bit-exact against the reference's numpy path for finite inputs.  NaN positions yield valid=0. */
int m3s_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_out,
                  uint8_t* valid_out, int B, int H, int W, int N, int max_iter, float lambda_init,
                  float convergence_thresh, void* workspace, size_t workspace_bytes,
                  m3s_stream_t stream);

/* Replaces refine_matches (backends/mpsgraph/kernels.py:463-493), numpy semantics of
 * _refine_matches_numpy (:496-537): argmax over the (2*radius+1)^2 window (step 1; dilation_max
 * is accepted and, exactly as in the reference's numpy path, has no effect) of <D21[n], D11[y,x]>,
 * scan dy outer / dx inner, strict '>', out-of-image candidates skipped.
 *   D11 [B,H,W,D] f32, D21 [B,N,D] f32, p1 [B,N,2] i32 -> p1_out [B,N,2] i32 */
int m3s_refine_matches(const float* D11, const float* D21, const int32_t* p1, int32_t* p1_out,
                       int B, int H, int W, int D, int N, int radius, int dilation_max,
                       m3s_stream_t stream);

/* Same, taking the float positions iter_proj produced; the `p1.astype(int32)` truncation of
 * matching.py:410 is fused into the load. */
int m3s_refine_matches_f32(const float* D11, const float* D21, const float* p1, int32_t* p1_out,
                           int B, int H, int W, int D, int N, int radius, int dilation_max,
                           m3s_stream_t stream);

/* Metal-semantics variants: what the reference computes on its GPU path.
 * iter_proj: shaders/iter_proj.metal:9-165 via metal_runner.py:118-196 -- all-fp32 bilinear blend,
 *   PER-POINT early exit when its own step is < convergence_thresh, zero step when |det| < 1e-10,
 *   clamp bound f32(W-1) - 0.001f.  No workspace (nothing global to resolve).
 * refine_matches: shaders/refine_matches.metal:160-215 via refine_metal_runner.py:136-164 -- window
 *   step = dilation, levels dilation_max .. 1 applied IN PLACE (each level restarts from the previous
 *   winner), sequential fp32 score sum, initial best score -1e10f.
 * Bit-exact against the shaders compiled as C++ with separately rounded operations
 * (DESIGN.md s7); a NaN start position (undefined behaviour in the shader) yields valid = 0. */
int m3s_iter_proj_metal(const float* rays, const float* pts, const float* p_init, float* p_out,
                        uint8_t* valid_out, int B, int H, int W, int N, int max_iter,
                        float lambda_init, float convergence_thresh, m3s_stream_t stream);
int m3s_refine_matches_metal(const float* D11, const float* D21, const int32_t* p1, int32_t* p1_out,
                             int B, int H, int W, int D, int N, int radius, int dilation_max,
                             m3s_stream_t stream);

/* Replaces the tail of match_iterative_proj (matching.py:438-461):
 *   gather X11 at clip(p1), valid = valid_proj & (||X11[p1] - X21|| < dist_thresh),
 *   idx = p1.x + W * p1.y.     X11, X21 [B,H,W,3]; p1 [B,H*W,2] i32; valid_proj [B,H*W] u8
 *   -> idx_out [B,H*W] i32, valid_out [B,H*W] u8 (the caller views it as [B,H*W,1]). */
int m3s_match_tail(const float* X11, const float* X21, const int32_t* p1,
                   const uint8_t* valid_proj, int32_t* idx_out, uint8_t* valid_out, int B, int H,
                   int W, float dist_thresh, m3s_stream_t stream);

/* Diagnostics (bench / tests): dev_counters = 3 x uint64 in device memory, or NULL to switch off
 * (default).  While set, the tile kernels add: [0] refine_matches queries that took the global-memory
 * scan instead of the staged window, [1] iter_proj points of tiles without a staged window, [2]
 * iter_proj points of windowed tiles that left the window at least once.  Process-global, not
 * thread-safe; results never depend on it. */
int m3s_set_path_counters(unsigned long long* dev_counters);

/* Replaces match_simple (matching.py:41-90), the route matching.match takes under the reference's
 * default configuration (config.py:65 use_simple=True): idx = idx_init (identity when NULL),
 * valid = ||X11[idx] - X21|| < dist_thresh.  X11, X21 [B,H,W,3]; idx_init [B,H*W] i32 or NULL
 *   -> idx_out [B,H*W] i32, valid_out [B,H*W] u8 (viewed as [B,H*W,1] by the caller).
 * An index outside [0, H*W) gives valid = 0 (the reference's gather is undefined there). */
int m3s_match_simple(const float* X11, const float* X21, const int32_t* idx_init, int32_t* idx_out,
                     uint8_t* valid_out, int B, int H, int W, float dist_thresh, m3s_stream_t stream);

/* Whole match_iterative_proj (matching.py:339-461) on device-resident inputs:
 * prep -> iter_proj -> trunc -> refine_matches -> tail, intermediates in `workspace`.
 *   X11, X21 [B,H,W,3]; D11, D21 [B,H,W,D]; idx_init [B,H*W] i32 or NULL
 *   -> idx_out [B,H*W] i32, valid_out [B,H*W] u8.   radius <= 0 skips the refinement. */
size_t m3s_match_workspace_bytes(int B, int H, int W, int max_iter);
int m3s_match_iterative_proj(const float* X11, const float* X21, const float* D11,
                             const float* D21, const int32_t* idx_init, int32_t* idx_out,
                             uint8_t* valid_out, int B, int H, int W, int D, int max_iter,
                             float lambda_init, float convergence_thresh, int radius,
                             int dilation_max, float dist_thresh, void* workspace,
                             size_t workspace_bytes, m3s_stream_t stream);

/* m3s_match_iterative_proj without a caller workspace and without any allocation: the
 * intermediates live in a file-scope device array of the library.  No counterpart in the
 * reference; added for callers that cannot call
 * cudaMalloc (stream capture, tight serving loops, harnesses that forbid it).
 *   - same arguments, results and error codes as m3s_match_iterative_proj
 *   - B*H*W > M3S_MATCH_STATIC_MAX_PIXELS => M3S_ERR_WORKSPACE with a message naming the limit
 *   - the array (73 B per pixel of the limit, 18.3 MiB) is held by every GPU that loads the
 *     library, until exit; it keeps the previous call's contents (the path clears what must
 *     start at zero itself)
 *   - all calls on one device share it: calls that overlap on the GPU (different streams, no
 *     ordering between them) race.  Calls on one stream, or ordered by events, are safe. */
#define M3S_MATCH_STATIC_MAX_PIXELS 262144 /* one 512 x 512 pair, or e.g. 4 pairs of 256 x 256 */
int m3s_match_iterative_proj_static(const float* X11, const float* X21, const float* D11,
                                    const float* D21, const int32_t* idx_init, int32_t* idx_out,
                                    uint8_t* valid_out, int B, int H, int W, int D, int max_iter,
                                    float lambda_init, float convergence_thresh, int radius,
                                    int dilation_max, float dist_thresh, m3s_stream_t stream);

/* Same computation on HOST arrays (what the reference's caller holds: mast3r_utils.py:490 passes
 * host arrays to matching.match): the batch is cut into chunks of `chunk_pairs` pairs (<= 0: 4)
 * that are uploaded, matched and downloaded on three rotating streams, so the upload of chunk
 * c+1 overlaps the kernels of chunk c and the download of chunk c-1.  Device buffers for one
 * chunk per stream are kept between calls (m3s_match_host_release frees them).  Host pointers
 * should be page-locked for the overlap; pageable memory is still correct.  `device` = CUDA
 * ordinal.  SYNCHRONOUS: returns when idx_out_host / valid_out_host are complete. */
int m3s_match_iterative_proj_host(const float* X11_host, const float* X21_host,
                                  const float* D11_host, const float* D21_host,
                                  const int32_t* idx_init_host, int32_t* idx_out_host,
                                  uint8_t* valid_out_host, int B, int H, int W, int D,
                                  int max_iter, float lambda_init, float convergence_thresh,
                                  int radius, int dilation_max, float dist_thresh,
                                  int chunk_pairs, int device);
int m3s_match_host_release(int device);

/* ---- factor-graph host steps (between match() and gauss_newton_*()) --------------------------- */

/* Replaces constrain_points_to_ray (geometry.py:273-302) = backproject (geometry.py:229-246) of
 * the pixel grid with each point's own depth:  x = (u - cx) / fx * z,  y = (v - cy) / fy * z.
 *   Xs [B,H*W,3] f32 (device), K4_host = (fx, fy, cx, cy) on the HOST -> out [B,H*W,3] (may alias Xs) */
int m3s_constrain_points_to_ray(const float* Xs, const float* K4_host, float* out, int B, int H,
                                int W, m3s_stream_t stream);

/* Replaces the confidence / match-fraction step of FactorGraph.add_factors
 * (global_opt.py:86-104):  Qj = sqrt(Qii[e, idx[e,n]] * Qji[e,n]),
 * valid = valid_match & (Qj > Q_conf), valid_count[e] = sum_n valid.
 *   Qii, Qji [E,N] f32; idx [E,N] i32; valid_match [E,N] u8
 *   -> Qj_out [E,N] f32, valid_out [E,N] u8 (or NULL), valid_count [E] i32.
 * idx outside [0,N) (an error in the reference) gives Qj = 0, valid = 0. */
int m3s_edge_confidence(const float* Qii, const float* Qji, const int32_t* idx,
                        const uint8_t* valid_match, float q_conf, float* Qj_out,
                        uint8_t* valid_out, int32_t* valid_count, int E, long long N,
                        m3s_stream_t stream);

/* ---- Gauss-Newton -------------------------------------------------------------------------- */

/* Per-point arithmetic of m3s_gn_accumulate.
 *   F64:   every per-point quantity in fp64, as the reference's numpy path (gauss_newton.py:92-220).
 *   MIXED: per-point residual / Jacobian products in fp32 (what the reference's own GPU kernel does,
 *          gauss_newton.metal:66-252), with the relative pose split into fp32 hi + lo parts so the
 *          residual keeps its systematic accuracy; per-thread fp32 partial sums over <= 32 points,
 *          everything above that in fp64.  Pose updates agree with the fp64 pass to 1e-7..5e-7
 *          relative (bar: 1e-5; scripts/gn_precision_report.py, tests/test_gpu_gn.py).  Rays and
 *          points only: calib always runs the F64 pass (weakly constrained depth direction).
 *   DEFAULT: MIXED unless the environment says M3S_GN_PRECISION=f64. */
#define M3S_GN_PREC_DEFAULT 0
#define M3S_GN_PREC_F64 1
#define M3S_GN_PREC_MIXED 2

/* Which implementation of the reference is reproduced.
 *   NUMPY (default): its numpy path (backends/mpsgraph/{kernels,gauss_newton*}.py) -- what its dispatcher
 *          runs without Metal, and what the parity bar of this library is stated against.
 *   METAL: its Metal compute shaders (backends/mpsgraph/shaders/*.metal + *_metal_runner.py): for
 *          Gauss-Newton fp32 per-point arithmetic (the MIXED pass; calib keeps F64), the step cast to
 *          fp32 and the fp32 retraction with dt = tau of gauss_newton.metal:255-301 on fp32-rounded
 *          poses; for matching see m3s_iter_proj_metal / m3s_refine_matches_metal. */
#define M3S_SEM_NUMPY 0
#define M3S_SEM_METAL 1

/* Scalars of gauss_newton_{rays,points,calib} (backends/mpsgraph/kernels.py:262-279, 325-346,
 * 396-412).  sigma_a = sigma_ray | sigma_point | sigma_pixel; sigma_b = sigma_depth (calib). */
typedef struct m3s_gn_params {
    int32_t mode;         /* M3S_GN_* */
    int32_t max_iter;
    int32_t pin;
    int32_t pixel_border; /* calib */
    int32_t img_w, img_h; /* calib: img_size = (width, height) */
    double sigma_a, sigma_b;
    double C_thresh, Q_thresh;
    double delta_thresh;
    double z_eps;         /* calib */
    float K4[4];          /* calib: fx, fy, cx, cy (the reference passes K as fp32) */
    int32_t precision;    /* M3S_GN_PREC_*: arithmetic of the per-point pass (reductions, H, the
                             solve and the poses are fp64 in every mode) */
    int32_t semantics;    /* M3S_SEM_*: which of the reference's two code paths is reproduced */
} m3s_gn_params;

/* Per-graph plan: the free-keyframe numbering of gauss_newton.py:70-78 (sorted unique(ii U jj),
 * first `pin` fixed), the deterministic H/g assembly lists (gauss_newton.py:222-251) and all
 * device workspaces (fp64 poses, per-edge blocks, dense H, g, dx).  Topology arrays are HOST
 * pointers.  `device` = CUDA device ordinal.  Returns NULL on error (see m3s_last_error). */
typedef struct m3s_gn_plan m3s_gn_plan;
m3s_gn_plan* m3s_gn_plan_create(const int32_t* ii_host, const int32_t* jj_host, int E, int num_kf,
                                int64_t N, int pin, int device);
void m3s_gn_plan_destroy(m3s_gn_plan* plan);
int m3s_gn_plan_num_free(const m3s_gn_plan* plan);      /* F; 0 => solver is a no-op (early out) */
int m3s_gn_plan_dim(const m3s_gn_plan* plan);           /* 7 F */
int m3s_gn_plan_ld(const m3s_gn_plan* plan);            /* row stride of H: dim rounded up to 32 */
/* factorisation path m3s_gn_solve_update will take for this plan on the current device:
 * 0 = one launch per block step, 1 = persistent chain + workers (M3S_GN_CHOL=fused),
 * 2 = dataflow over shared-memory-resident tiles (default when every tile CTA is co-resident),
 * 3 = sparse front (keyframes with few neighbours eliminated exactly as 7x7 blocks, level by level) +
 *     path 2 on the Schur complement of the remaining keyframes (default when the graph is sparse
 *     enough to shorten the dense chain; M3S_GN_SPARSE=0 or M3S_GN_CHOL=steps|fused|tiles keep it dense).
 * Same system as gauss_newton.py:254-265 (H + 1e-6 I, Cholesky); only the elimination order differs. */
int m3s_gn_plan_factor_path(m3s_gn_plan* plan);
/* sparse front of the plan: out[0] = levels (0 = not used), out[1] = keyframes eliminated in the levels,
 * out[2] = keyframes of the dense tail, out[3] = padded dimension of the tail system.  Diagnostic; no
 * counterpart in the reference (its solve is one dense Cholesky, gauss_newton.py:254-262). */
int m3s_gn_plan_sparse_info(const m3s_gn_plan* plan, int32_t* out);
/* device pointer + element count of the packed per-edge blocks [E][M3S_GN_EDGE_STRIDE] f64 that
 * m3s_gn_accumulate writes and m3s_gn_solve_update reads: the buffer a multi-GPU caller
 * all-reduces (sum) between the two when edges are sharded across ranks. */
#define M3S_GN_EDGE_STRIDE 36
double* m3s_gn_plan_edge_blocks(m3s_gn_plan* plan, int64_t* count_out);
double* m3s_gn_plan_H(m3s_gn_plan* plan);   /* [dim rows, ld] f64 (device): H + 1e-6 I after
                                               m3s_gn_assemble; used as the factorisation's work
                                               matrix (destroyed) by m3s_gn_solve_update */
double* m3s_gn_plan_g(m3s_gn_plan* plan);   /* [dim] */
double* m3s_gn_plan_dx(m3s_gn_plan* plan);  /* [dim] last solved step */
/* Development aid: with M3S_GN_TIMELINE=1 in the environment when the plan is created, the solve
 * kernels stamp %globaltimer (ns) at entry / exit: [4 kinds (diag, panel, syrk, backsolve)]
 * [dim/32 blocks][2] int64 on the device; NULL / 0 when disabled. */
long long* m3s_gn_plan_timeline(m3s_gn_plan* plan, int64_t* count_out);
double* m3s_gn_plan_poses(m3s_gn_plan* plan); /* [num_kf, 8] f64: the solver's pose state
                                                 (gauss_newton.py:92, Twc.astype(float64)) */

/* Load poses Twc [num_kf,8] f32 (tx,ty,tz,qx,qy,qz,qw,s) into the plan's fp64 state
 * (gauss_newton.py:92-94) and reset the iteration status. */
int m3s_gn_set_poses(m3s_gn_plan* plan, const float* Twc, m3s_stream_t stream);
/* Write the fp64 state back as f32 (gauss_newton.py:279). */
int m3s_gn_get_poses(m3s_gn_plan* plan, float* Twc_out, m3s_stream_t stream);

/* One residual/Jacobian/reduction pass over directed edges [edge_begin, edge_end)
 * (gauss_newton.py:106-220; Metal gn_jacobian_kernel gauss_newton.metal:66-252):
 * per edge the symmetric 7x7 block and 7-vector in the pose-free basis, reduced warp -> block ->
 * fixed-order chunk sum (no atomics); slots of edges outside the range are zero-filled so a
 * sum all-reduce over ranks reproduces the single-GPU buffer bit for bit.
 *   Xs [K,N,3] f32, Cs [K,N] f32, idx [E,N] i32, valid [E,N] u8, Q [E,N] f32. */
int m3s_gn_accumulate(m3s_gn_plan* plan, const m3s_gn_params* params, const float* Xs,
                      const float* Cs, const int32_t* idx, const uint8_t* valid, const float* Q,
                      int edge_begin, int edge_end, m3s_stream_t stream);

/* Peer exchange for edge-sharded multi-GPU solves (replaces the NCCL all-reduce of
 * the packed blocks, which the reference has no counterpart for -- it is single-device).
 * Once attached, m3s_gn_accumulate stores every owned edge block straight into the receive buffer
 * of EVERY rank through NVLink peer mappings (CUDA IPC) from inside gn_edge_finalize_kernel, the
 * last CTA publishes an epoch flag to all peers (st.release.sys), and a wait kernel on each rank
 * acquires all flags and copies the (epoch-parity double-buffered) receive buffer into the plan's
 * block buffer -- all enqueued on `stream`, no host synchronisation, no collective library call.
 * Every rank must issue the same sequence of m3s_gn_accumulate calls (it is a collective).
 *   create: allocates this rank's region, writes its 64-byte cudaIpcMemHandle to ipc_handle_out;
 *   attach: ipc_handles = world x 64 bytes, rank-ordered (exchanged by the host, e.g.
 *           torch.distributed.all_gather); opens the peers' regions.  The caller must barrier
 *           after attach and before the first accumulate, and before detach / plan destroy.
 *   detach: unmaps the peers (the rank's own region lives until the plan is destroyed; barrier
 *           after detach on all ranks before destroying plans).
 *   release: tells the plan that every importer has unmapped this rank's region (call it after the
 *           post-detach barrier).  A plan destroyed, or re-created, without it LEAKS the exported
 *           region on purpose: freeing memory a peer process may still have mapped is undefined.
 * After the delta test stops a solve, the remaining enqueued iterations keep running the (empty)
 * exchange so that epochs stay consecutive across solves.  A peer that does not publish within
 * M3S_GN_XCH_TIMEOUT_MS (environment, read at create; default 10000, 0 = wait for ever) marks the
 * solve failed with the distinct status 2 (m3s_gn_status) instead of hanging the device.  world <= 8. */
int m3s_gn_exchange_create(m3s_gn_plan* plan, int rank, int world, void* ipc_handle_out);
int m3s_gn_exchange_attach(m3s_gn_plan* plan, const void* ipc_handles);
int m3s_gn_exchange_detach(m3s_gn_plan* plan);
int m3s_gn_exchange_release(m3s_gn_plan* plan);

/* Assemble only: dense H (+ 1e-6 I, gauss_newton.py:254) and g from the per-edge blocks through
 * fixed-order lists (gauss_newton.py:222-251).  m3s_gn_solve_update calls it itself; exposed for
 * inspection and tests. */
int m3s_gn_assemble(m3s_gn_plan* plan, m3s_stream_t stream);

/* Assemble dense H, g from the per-edge blocks (gauss_newton.py:222-251), H += 1e-6 I, solve
 * H dx = -g by Cholesky (:254-260), test ||dx|| < delta_thresh BEFORE applying (:263-265),
 * retract the free poses T <- exp(dx) o T (sim3_ops.py:229-251).  Entirely on the device. */
int m3s_gn_solve_update(m3s_gn_plan* plan, const m3s_gn_params* params, m3s_stream_t stream);

/* status_host[0] = iterations applied, [1] = 1 if stopped by the delta test, [2] = 1 if the solve
 * failed numerically (non-finite / not positive definite) or 2 if the peer exchange timed out,
 * [3] = iterations attempted.  Synchronises. */
int m3s_gn_status(m3s_gn_plan* plan, int32_t* status_host, double* last_dx_norm_host,
                  m3s_stream_t stream);

/* Whole solver: gauss_newton_rays / _points / _calib of backends/mpsgraph/kernels.py:262,396,325.
 *   Twc [K,8] f32 device -> Twc_out [K,8] f32 device.  No host synchronisation. */
int m3s_gauss_newton(m3s_gn_plan* plan, const m3s_gn_params* params, const float* Twc,
                     const float* Xs, const float* Cs, const int32_t* idx, const uint8_t* valid,
                     const float* Q, float* Twc_out, m3s_stream_t stream);

/* ---- tracker single-pose solves ------------------------------------------ */

#define M3S_TRACK_RAY_DIST 0 /* FrameTracker._opt_pose_ray_dist_sim3 (tracker.py:255-313) */
#define M3S_TRACK_CALIB 1    /* FrameTracker._opt_pose_calib_sim3    (tracker.py:315-406) */

/* config "tracking" section (config.py:74-91) as it reaches the solver */
typedef struct m3s_track_params {
    int32_t mode;          /* M3S_TRACK_* */
    int32_t max_iters;     /* 10 */
    int32_t img_w, img_h;  /* calib only; img_w * img_h == N (pixel n = (n % img_w, n / img_w)) */
    int32_t pixel_border;  /* calib only */
    int32_t reserved;
    double sigma_ray, sigma_dist;    /* 0.003, 10 */
    double sigma_pixel, sigma_depth; /* 1, 10 */
    double huber;                    /* 1.345 */
    double rel_error, delta_norm;    /* 1e-3, 1e-3 (optimizer.py:11-46) */
    double depth_eps;                /* calib only */
    float K4[4];                     /* fx, fy, cx, cy (calib only) */
} m3s_track_params;

/* Replaces the confidence / mask block of FrameTracker.track (tracker.py:86-113):
 *   Qk = sqrt(Qff[idx] * Qkf); valid_kf = valid_match & (Qk > Q_conf);
 *   valid_opt = valid_kf & (Cf[idx] > C_conf) & (Ck > C_conf); counts_out = {sum valid_opt, sum valid_kf}.
 * All arrays [N] on the device (Cf, Qff indexed through idx_f2k). */
int m3s_track_gate(const float* Cf, const float* Ck, const float* Qff, const float* Qkf,
                   const int32_t* idx_f2k, const uint8_t* valid_match, float C_conf, float Q_conf,
                   float* Qk_out, uint8_t* valid_opt_out, uint8_t* valid_kf_out,
                   int32_t* counts_out, int N, m3s_stream_t stream);

size_t m3s_track_workspace_bytes(int N);

/* Replaces FrameTracker._opt_pose_ray_dist_sim3 / _opt_pose_calib_sim3 incl. _solve (tracker.py:216-253),
 * point_to_ray_dist / act_Sim3 / project_calib Jacobians (geometry.py:46-227), Sim3.retr and
 * check_convergence (optimizer.py:11-46).  The whole <= max_iters loop is enqueued without host
 * synchronisation (a device flag replays the reference's `break`).
 *   Xf [N,3] frame points; idx_f2k [N] i32 gathers them (Xf[idx_f2k], tracker.py:214) or NULL if the
 *   caller already gathered; Xk [N,3] keyframe points; Qk [N]; valid [N] u8 (valid_opt; in calib
 *   mode the measurement / projection masks are derived from Xk and the projected point);
 *   T_WCf, T_WCk [8] f32 device (t, q_xyzw, s) -> T_WCf_out, T_CkCf_out [8] f32 device;
 *   stats_out (device, may be NULL) = {iterations done, last cost, last |tau|, 1 if the solve failed}. */
int m3s_track_opt_pose(const m3s_track_params* params, const float* Xf, const int32_t* idx_f2k,
                       const float* Xk, const float* Qk, const uint8_t* valid, int N,
                       const float* T_WCf, const float* T_WCk, float* T_WCf_out, float* T_CkCf_out,
                       double* stats_out, void* workspace, size_t workspace_bytes,
                       m3s_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* M3S_H_ */
