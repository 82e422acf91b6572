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


@pytest.mark.parametrize("mode", ["nve", "ttm", "nrb", "ttm3d", "nve_cuts", "ttm_cuts"])
def test_two_ranks_equal_one_rank(mode):
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    port = {"nve": "29532", "ttm": "29533", "ttm3d": "29534", "nrb": "29535", "nve_cuts": "29536", "ttm_cuts": "29537"}[mode]
    cuts = mode.endswith("_cuts")                 # slabs of uneven width / equal atom count (imdcu_comm_set_cuts)
    mode = mode.replace("_cuts", "")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", port, os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    if mode != "nve":
        cmd.append("--" + mode)
    if cuts:
        cmd.append("--cuts")
    res = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "MULTI_GPU_CHECK PASSED" in res.stdout
