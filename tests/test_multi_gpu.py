"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): tests/multi_gpu_check.py
under torchrun -- 2-rank x-slab run == 1-rank run (pair set, numneighs bit-exact; forces,
energies, positions 1e-10; FD lattice to the TTM tolerances), across rebuilds with migration."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("ttm", [False, True])
def test_two_ranks_equal_one_rank(ttm):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533" if ttm else "29532", os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    if ttm:
        cmd.append("--ttm")
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "MULTI_GPU_CHECK PASSED" in res.stdout
