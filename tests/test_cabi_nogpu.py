"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol include/m3s.h
declares, and fails loudly (no fallback) when there is no CUDA device."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from mast3r_slam_b200 import _build, _lib
    _build.build_native()
    return _lib.load()


def test_header_symbols_exported(lib):
    from mast3r_slam_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "m3s.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(m3s_[A-Za-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/m3s.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_library_has_no_torch_dependency():
    import subprocess
    from mast3r_slam_b200 import _lib
    out = subprocess.run(["ldd", _lib.SO_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "python" not in out


def test_version_and_errors(lib):
    assert lib.m3s_version() >= 1
    assert isinstance(lib.m3s_last_error(), bytes)
    # argument validation happens before any CUDA call
    rc = lib.m3s_iter_proj(None, None, None, None, None, 1, 4, 4, 4, 10, 1e-8, 1e-6, None, 0, None)
    assert rc == -1 and b"null" in lib.m3s_last_error()
    assert lib.m3s_iter_proj_workspace_bytes(10) > 0
    assert lib.m3s_match_workspace_bytes(1, 384, 512, 10) > 196608 * 9 * 4


def test_no_silent_cpu_fallback(lib):
    """without a device the product raises; it never routes through the oracle or numpy"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mast3r_slam_b200 import kernels
    from mast3r_slam_b200._lib import M3SError
    assert kernels.is_available() is False
    rays = np.zeros((1, 4, 4, 9), np.float32)
    with pytest.raises((M3SError, RuntimeError)):
        kernels.iter_proj(rays, np.zeros((1, 3, 3), np.float32), np.zeros((1, 3, 2), np.float32))
    with pytest.raises((M3SError, RuntimeError)):
        kernels.GNPlan([0], [1], 2, 16, 1, torch.device("cuda", 0))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "mast3r-slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("no CPU", ""), f"{f} mentions the oracle"


def test_config_defaults_match_reference():
    from mast3r_slam_b200.config import DEFAULT_CONFIG
    m, l = DEFAULT_CONFIG["matching"], DEFAULT_CONFIG["local_opt"]
    assert (m["max_iter"], m["lambda_init"], m["convergence_thresh"], m["dist_thresh"]) == (10, 1e-8, 1e-6, 0.1)
    assert (m["refine_radius"], m["refine_dilation"], m["use_refine"]) == (3, 2, True)
    assert (l["pin"], l["max_iters"], l["Q_conf"], l["sigma_ray"], l["delta_norm"]) == (1, 10, 1.5, 0.003, 1e-3)
