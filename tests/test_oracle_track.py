"""CPU checks of the tracker restatement in oracle/oracle.py.  The reference's
tracker is MLX code and cannot run here, so these tests pin the restatement to properties the
reference's formulas imply: group axioms of the Sim3 helpers, Jacobians against finite differences,
convergence to the ground truth, float32 (reference precision) vs float64 agreement."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def synth():
    from mast3r_slam_b200 import synth as s
    return s


def _rand_sim3(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    return np.concatenate([rng.normal(size=3), q, [np.exp(0.2 * rng.normal())]])


def test_sim3_helpers_group_axioms(oracle):
    rng = np.random.default_rng(0)
    a, b = _rand_sim3(rng), _rand_sim3(rng)
    p = rng.normal(size=(5, 3))
    act = lambda T, x: T[7] * oracle._so3_act(T[3:7][None], x) + T[:3]  # noqa: E731
    ab = oracle._sim3_mul(a, b)
    assert np.allclose(act(ab, p), act(a, act(b, p)), atol=1e-12)
    e = oracle._sim3_mul(a, oracle._sim3_inv(a))
    assert np.allclose(e, [0, 0, 0, 0, 0, 0, 1, 1], atol=1e-9)
    # exp: pure scale / pure translation / small-angle branch continuity
    assert np.allclose(oracle._sim3_exp(np.array([0, 0, 0, 0, 0, 0, 0.3]))[7], np.exp(0.3))
    assert np.allclose(oracle._sim3_exp(np.array([1.0, 2, 3, 0, 0, 0, 0]))[:3], [1, 2, 3], atol=1e-4)
    w = np.array([1.0, -2.0, 0.5])
    w /= np.linalg.norm(w)
    lo, hi = oracle._sim3_exp(np.r_[0.1, 0.2, 0.3, 0.99e-4 * w, 0.0]), oracle._sim3_exp(np.r_[0.1, 0.2, 0.3, 1.01e-4 * w, 0.0])
    assert np.abs(lo - hi).max() < 1e-5
    q = oracle._sim3_exp(np.r_[0, 0, 0, 0.7 * w, 0.0])[3:7]
    assert np.allclose(q, np.r_[np.sin(0.35) * w, np.cos(0.35)], atol=1e-5)  # theta biased by 1e-10


def test_jacobians_finite_difference(oracle):
    rng = np.random.default_rng(1)
    X = rng.normal(size=(6, 3)) + np.array([0, 0, 3.0])
    rd, J = oracle._ray_dist(X, True)
    for k in range(3):
        d = np.zeros(3)
        d[k] = 1e-6
        num = (oracle._ray_dist(X + d, False) - oracle._ray_dist(X - d, False)) / 2e-6
        assert np.allclose(J[:, :, k], num, atol=1e-7)
    # act_Sim3 Jacobian = derivative of exp(tau) o T acting on X at tau = 0 (left perturbation)
    T = _rand_sim3(rng)
    pW, dX = oracle._act_jac(T, X)
    for k in range(7):
        tau = np.zeros(7)
        tau[k] = 1e-6
        Tp = oracle._sim3_mul(oracle._sim3_exp(tau), T)
        Tm = oracle._sim3_mul(oracle._sim3_exp(-tau), T)
        act = lambda A: A[7] * oracle._so3_act(A[3:7][None], X) + A[:3]  # noqa: E731
        num = (act(Tp) - act(Tm)) / 2e-6
        assert np.allclose(dX[:, :, k], num, atol=1e-5), k


@pytest.mark.parametrize("seed", [0, 1])
def test_tracker_oracle_converges(oracle, synth, seed):
    d = synth.make_track_pair(seed, 96, 128)
    Qk, v_opt, v_kf, (n_opt, n_kf) = oracle.track_gate_numpy(d["Cf"], d["Ck"], d["Qff"], d["Qkf"],
                                                            d["idx_f2k"], d["valid_match"])
    assert n_opt == int(v_opt.sum()) and 0 < n_opt <= n_kf
    # gate against a scalar loop
    for n in range(0, len(Qk), 997):
        m = d["idx_f2k"][n]
        q = np.sqrt(np.float32(d["Qff"][m] * d["Qkf"][n]))
        assert Qk[n, 0] == q
        assert v_opt[n, 0] == bool(d["valid_match"][n, 0] and q > 1.5 and d["Cf"][m] > 0 and d["Ck"][n] > 0)
    Xf = d["Xf"][d["idx_f2k"]]
    e0 = np.abs(d["T_WCf"] - d["T_WCf_gt"]).max()
    Tw, Tr, info = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt)
    assert 2 <= info["iterations"] <= 10 and np.abs(Tw - d["T_WCf_gt"]).max() < 0.02 * e0
    assert np.allclose(oracle._sim3_mul(np.asarray(d["T_WCk"], np.float64), Tr), Tw)
    Tw32, _, i32 = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt,
                                                      dtype=np.float32)
    assert np.abs(Tw32 - Tw).max() < 1e-3 * e0
    Twc, _, ic = oracle.opt_pose_calib_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt, d["K4"],
                                                  d["img_size"])
    assert np.abs(Twc - d["T_WCf_gt"]).max() < 0.02 * e0
