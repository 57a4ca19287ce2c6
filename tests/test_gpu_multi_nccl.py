"""The NCCL path of variant 3 (cells of one mesh split over GPUs): one process per GPU
under torchrun, compared with the single-rank run and with the oracle's scipy loop
(tests/nccl_partitioned_check.py).  Needs at least two GPUs; skipped otherwise (the same code with
gloo ranks sharing one GPU is tests/test_gpu_partitioned.py::test_cells_split_over_ranks)."""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return sorted({w for w in (2, 3, n) if 2 <= w <= min(n, 8)})


@pytest.mark.skipif(not _worlds(), reason="needs at least two GPUs")
@pytest.mark.parametrize("world", _worlds() or [2])
def test_partitioned_solver_over_nccl(world):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "nccl_partitioned_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and lines, res.stdout[-2000:] + res.stderr[-2000:]
    rep = json.loads(lines[-1])
    assert rep["ok"] and rep["backend"] == "nccl" and rep["world"] == world
    assert rep["max_abs_vs_single_rank"] < 1e-13 and rep["max_rel_vs_oracle"] < 1e-10
