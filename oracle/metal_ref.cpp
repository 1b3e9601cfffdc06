// CPU execution of the reference's UNMODIFIED Metal compute shaders (compiled as C++ through
// oracle/metal_shim/metal_stdlib, from where they lie under $M3S_REFERENCE_SHADERS =
// /root/reference/src/mlx_mast3r_slam/backends/mpsgraph/shaders), one "thread" per loop iteration.
// Oracle of the Metal-semantics mode.  TEST INFRASTRUCTURE ONLY: built by
// oracle/Makefile into oracle/_ref/libm3s_metal_ref.so (git-ignored, travels with gpurun); nothing
// under mast3r-slam_b200/ loads it.
//
// Host-side dispatch restated from the reference's runners:
//   iter_proj           metal_runner.py:118-196   (one dispatch, thread per point)
//   refine_matches      refine_metal_runner.py:136-164 (multiscale kernel, in place, dilation_max..1)
//   gn jacobian / pose update  gn_metal_runner.py:166-214, 302-330 (per-point outputs; the CPU
//                       accumulation of :216-300 is restated in oracle/metal_oracle.py)
#include <metal_stdlib>

namespace mip {
#include "iter_proj.metal"
}
namespace mrf {
#include "refine_matches.metal"
}
namespace mgr {
#include "gauss_newton.metal"
}
namespace mgp {
#include "gauss_newton_points.metal"
}
namespace mgc {
#include "gauss_newton_calib.metal"
}

#define MREF_API extern "C" __attribute__((visibility("default")))

MREF_API void mref_iter_proj(const float* rays, const float* pts, const float* p_init, float* p_out,
                             unsigned char* valid_out, int B, int H, int W, int N, int max_iter,
                             float lambda_init, float thresh) {
    static_assert(sizeof(bool) == 1, "bool buffers are byte arrays");
    const long long total = (long long)B * N;
#pragma omp parallel for schedule(static)
    for (long long t = 0; t < total; ++t)
        mip::iter_proj_kernel(rays, pts, p_init, p_out, reinterpret_cast<bool*>(valid_out), B, H, W, N,
                              max_iter, lambda_init, thresh, (uint)t);
}

// p1 [B,N,2] int32 is updated in place, coarse to fine (refine_metal_runner.py:136-164)
MREF_API void mref_refine_matches(const float* D11, const float* D21, int* p1, int B, int H, int W, int D,
                                  int N, int radius, int dilation_max) {
    const long long total = (long long)B * N;
    for (int dil = dilation_max > 1 ? dilation_max : 1; dil >= 1; --dil) {
#pragma omp parallel for schedule(static)
        for (long long t = 0; t < total; ++t)
            mrf::refine_matches_multiscale_kernel(D11, D21, p1, B, H, W, D, N, radius, dil, (uint)t);
    }
}

// mode 0 rays, 1 points, 2 calib.  Outputs per (edge, point): JtJ_i/j [28], Jtr_i/j [7], JtJ_ij [49],
// valid [1] (entries of invalid points are left untouched, as on the GPU).
MREF_API void mref_gn_jacobian(int mode, const float* Twc, const float* Xs, const float* Cs, const int* ii,
                               const int* jj, const int* idx, const unsigned char* valid, const float* Q,
                               const float* K4, float* JtJ_i, float* Jtr_i, float* JtJ_j, float* Jtr_j,
                               float* JtJ_ij, unsigned char* valid_out, int num_kf, int num_pts,
                               int num_edges, float sigma_a_inv, float sigma_b_inv, float C_thresh,
                               float Q_thresh, int img_w, int img_h, int pixel_border, float z_eps) {
    const long long total = (long long)num_edges * num_pts;
    const bool* v = reinterpret_cast<const bool*>(valid);
    bool* vo = reinterpret_cast<bool*>(valid_out);
#pragma omp parallel for schedule(static)
    for (long long t = 0; t < total; ++t) {
        if (mode == 0)
            mgr::gn_jacobian_kernel(Twc, Xs, Cs, ii, jj, idx, v, Q, JtJ_i, Jtr_i, JtJ_j, Jtr_j, JtJ_ij, vo,
                                    num_kf, num_pts, num_edges, sigma_a_inv, C_thresh, Q_thresh, (uint)t);
        else if (mode == 1)
            mgp::gn_points_jacobian_kernel(Twc, Xs, Cs, ii, jj, idx, v, Q, JtJ_i, Jtr_i, JtJ_j, Jtr_j, JtJ_ij,
                                           vo, num_kf, num_pts, num_edges, sigma_a_inv, C_thresh, Q_thresh,
                                           (uint)t);
        else
            mgc::gn_calib_jacobian_kernel(Twc, Xs, Cs, ii, jj, idx, v, Q, K4, JtJ_i, Jtr_i, JtJ_j, Jtr_j,
                                          JtJ_ij, vo, num_kf, num_pts, num_edges, img_w, img_h, pixel_border,
                                          z_eps, sigma_a_inv, sigma_b_inv, C_thresh, Q_thresh, (uint)t);
    }
}

MREF_API void mref_pose_update(float* Twc, const float* dx, const int* free_kf_idx, int num_free) {
    for (int t = 0; t < num_free; ++t) mgr::pose_update_kernel(Twc, dx, free_kf_idx, num_free, (uint)t);
}
