"""Multi-GPU identity gate: the domain-decomposed XPBD pass loop (one exchange per
(energy, colour) pass, xpbd_partition.h) and the domain-decomposed VBD sweep with per-colour
halo exchange (NCCL point-to-point, and the fused peer-memory exchange) on k GPUs reproduces the single-GPU result bit for bit (every vertex update is a deterministic
function of its stencil neighbours and the per-vertex accumulation order is kept).  Needs >= 2 GPUs; skipped otherwise."""
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _ngpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("scene", ["cloth", "beam", "cloth_edit", "xpbd_cloth", "xpbd_cloth_lambda", "xpbd_beam"])
def test_k_gpu_equals_one_gpu(tmp_path, scene):
    n = _ngpus()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(n, 4)
    out = tmp_path / "res.npz"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(ROOT / "tests" / "multi_worker.py"), scene, str(out)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    d = np.load(out)
    for mode in ("nccl", "p2p"):  # ncclSend / ncclRecv schedule, fused peer-memory stores + event counters
        for graph in (1, 0):      # CUDA-graph replay, plain launches
            assert np.array_equal(d[f"x_{mode}_{graph}"], d["x1"]), (mode, graph)
            assert np.array_equal(d[f"v_{mode}_{graph}"], d["v1"]), (mode, graph)
