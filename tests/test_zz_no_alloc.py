"""m3s_match_iterative_proj_static: the allocation-free variant of the whole-match entry point
(scratch in a file-scope device array of the library) against m3s_match_iterative_proj with a
caller workspace.  Called through the C ABI (ctypes); torch only provides the device buffers.

The bar is the one the matching tests use: indices and validity are integer work, so bit-exact.
Kept in a file of its own that sorts last, so that under `pytest -x` a failure here hides nothing.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ARGS = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, radius=3, dilation_max=2,
            dist_thresh=0.1)


@pytest.fixture(scope="module")
def env():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import _lib, kernels
    assert kernels.is_available(), "CUDA extension not loaded"
    return torch, _lib, _lib.load()


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _inputs(torch, seeds, h, w, d):
    from mast3r_slam_b200 import synth
    ps = synth.make_pairs(seeds, h, w, d, "dyadic")
    dev = torch.device("cuda", 0)
    return {k: torch.from_numpy(np.ascontiguousarray(v, dtype=np.float32)).to(dev) for k, v in ps.items()}


def _call(torch, lib, check, t, static, idx_init=None, **over):
    a = dict(ARGS, **over)
    b, h, w, _ = t["X11"].shape
    d = t["D11"].shape[-1]
    dev = t["X11"].device
    # poison the outputs so that a call that writes nothing cannot pass
    idx = torch.full((b, h * w), -7, dtype=torch.int32, device=dev)
    valid = torch.full((b, h * w), 3, dtype=torch.uint8, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    head = (_ptr(t["X11"]), _ptr(t["X21"]), _ptr(t["D11"]), _ptr(t["D21"]), _ptr(idx_init), _ptr(idx),
            _ptr(valid), b, h, w, d, a["max_iter"], a["lambda_init"], a["convergence_thresh"],
            a["radius"], a["dilation_max"], a["dist_thresh"])
    if static:
        rc = lib.m3s_match_iterative_proj_static(*head, stream)
    else:
        nb = lib.m3s_match_workspace_bytes(b, h, w, a["max_iter"])
        ws = torch.empty(nb, dtype=torch.uint8, device=dev)
        rc = lib.m3s_match_iterative_proj(*head, _ptr(ws), nb, stream)
    check(rc, "match_iterative_proj_static" if static else "match_iterative_proj")
    torch.cuda.synchronize(dev)
    return idx.cpu().numpy(), valid.cpu().numpy()


def test_static_matches_workspace_variant_twice(env):
    """two calls in a row on different inputs and shapes: the second one runs over whatever the
    first left in the shared array and must still agree with the workspace variant"""
    torch, _lib, lib = env
    first = _inputs(torch, [11, 12], 72, 96, 24)
    second = _inputs(torch, [13], 120, 88, 16)
    for t in (first, second, first):
        idx_w, valid_w = _call(torch, lib, _lib.check, t, static=False)
        idx_s, valid_s = _call(torch, lib, _lib.check, t, static=True)
        assert idx_s.min() >= 0 and valid_s.max() <= 1
        assert np.array_equal(idx_s, idx_w)
        assert np.array_equal(valid_s, valid_w)
        assert valid_s.mean() > 0.5  # the scene matches: not a trivially empty answer
    # with an initial correspondence, no refinement, and a single iteration (no early-exit scratch)
    idx0 = torch.from_numpy(idx_w.copy()).to(first["X11"].device)
    for over in (dict(radius=0), dict(max_iter=1)):
        a = _call(torch, lib, _lib.check, first, static=False, idx_init=idx0, **over)
        b = _call(torch, lib, _lib.check, first, static=True, idx_init=idx0, **over)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_static_full_limit(env):
    """exactly M3S_MATCH_STATIC_MAX_PIXELS pixels (one 512 x 512 pair) fits"""
    torch, _lib, lib = env
    t = _inputs(torch, [21], 512, 512, 24)
    idx_w, valid_w = _call(torch, lib, _lib.check, t, static=False)
    idx_s, valid_s = _call(torch, lib, _lib.check, t, static=True)
    assert np.array_equal(idx_s, idx_w) and np.array_equal(valid_s, valid_w)


def test_static_rejects_oversize(env):
    torch, _lib, lib = env
    t = _inputs(torch, [1], 8, 8, 4)
    dev = t["X11"].device
    idx = torch.empty((1, 64), dtype=torch.int32, device=dev)
    valid = torch.empty((1, 64), dtype=torch.uint8, device=dev)
    # the shape check comes before any pointer is used: claim a batch beyond the limit
    for b, h, w in ((2, 512, 512), (1, 513, 512), (65536, 65536, 2)):
        rc = lib.m3s_match_iterative_proj_static(
            _ptr(t["X11"]), _ptr(t["X21"]), _ptr(t["D11"]), _ptr(t["D21"]), None, _ptr(idx), _ptr(valid),
            b, h, w, 4, 10, 1e-8, 1e-6, 3, 2, 0.1, None)
        assert rc == -3, rc  # M3S_ERR_WORKSPACE
        msg = lib.m3s_last_error().decode()
        assert "262144" in msg and "limit" in msg, msg
    torch.cuda.synchronize(dev)
