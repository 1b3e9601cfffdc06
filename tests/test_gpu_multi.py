"""Multi-GPU parity, collected by pytest: spawns tests/mgpu_check.py on 2 ranks (and
on every GPU of the box when there are more) through torch.distributed.run.  Skips on a 1-GPU box.

mgpu_check asserts, on every rank: edge-sharded Gauss-Newton with the fused peer-store exchange and
with the NCCL exchange == the single-GPU solve bit for bit (first solve, 10 repeated solves, 6
consecutive early-stopped solves per exchange, a rank that owns no edges), all ranks agree, the result
is within 1e-5 of the oracle, and pair-sharded matching == un-sharded matching."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worlds():
    try:
        import torch
        n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        n = 0
    return n


@pytest.mark.parametrize("world", [2, 8])
def test_mgpu_check(world):
    n = _worlds()
    if n < world:
        pytest.skip(f"needs {world} GPUs, box has {n}")
    env = dict(os.environ)
    env.pop("M3S_GN_EXCHANGE", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "mgpu_check.py")]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    tail = (r.stdout + r.stderr)[-3000:]
    assert r.returncode == 0, tail
    assert "mgpu_check ok" in r.stdout, tail
