"""3-D TTM sweep: the numpy restatement (oracle/ttm3d_np.py) against the reference's own do_DIFF +
ttm_fill_ghost_layers (imd_ttm_3D.c:1603-1829, 3314-3504): committed golden outputs of
oracle/_ref/ref3d_diff, and -- where that binary exists -- live runs on further seeds.
The comparison is bit-exact: same IEEE operations in the same order."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import ttm3d_np as t3  # noqa: E402
from ttm3d_util import GOLDEN, load_golden  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref3d_diff")


def run_oracle(case):
    f = t3.make_case(case["shape"], seed=case["seed"], hole=case["hole"], pbc=case["pbc"], fd_min_atoms=case["fd_min_atoms"])
    e = t3.run(f, case["h"], case["tau"], case["fd_min_atoms"], case["diff_substeps"], case["laser_active"],
               case["nsweeps"], case["pbc"])
    return e, f


def test_fixtures_present():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_equals_reference_golden(path):
    case, ref = load_golden(path)
    e, f = run_oracle(case)
    for k in ("natoms", "temp", "U", "xi"):
        assert np.array_equal(f[k], ref[k]), k                      # whole extended lattice, bit for bit
    assert e == pytest.approx(ref["eabs"], rel=1e-13)                # sum order of the cumulative Eabs
    act = ref["natoms"][1:-1, 1:-1, 1:-1] >= case["fd_min_atoms"]
    assert act.any() and (~act).any()                                # the case has active and inactive cells


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/ref3d_diff not built (needs /root/reference)")
@pytest.mark.parametrize("seed,pbc,shape", [(11, (0, 1, 1), (5, 7, 6)), (12, (1, 1, 1), (1, 4, 3)), (13, (0, 1, 1), (14, 3, 33))])
def test_oracle_equals_reference_live(seed, pbc, shape):
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from make_golden_ttm3d import run_reference
    case = dict(shape=shape, seed=seed, hole=min(shape) >= 3, pbc=pbc, fd_min_atoms=20, h=(15.0, 17.0, 16.0),
                tau=0.1 / 300, diff_substeps=300, laser_active=1, nsweeps=5)
    f0 = t3.make_case(shape, seed=seed, hole=case["hole"], pbc=pbc, fd_min_atoms=20)
    eabs, nat, temp, U, xi = run_reference(f0, case)
    e, f = run_oracle(case)
    assert np.array_equal(f["natoms"], nat) and np.array_equal(f["temp"], temp)
    assert np.array_equal(f["U"], U) and np.array_equal(f["xi"], xi)
    assert e == pytest.approx(eabs, rel=1e-13)


def test_conservation_and_fixed_point():
    """Size-independent properties the GPU tests rely on at full size: without coupling and source
    the sweep conserves sum(Ce*T) over the active cells (conservative form, :1715-1733), and a
    uniform temperature is a fixed point."""
    shape = (10, 9, 8)
    f = t3.make_case(shape, seed=2, hole=True, pbc=(1, 1, 1))
    f["fd_g"][:] = 0.0
    f["source"][:] = 0.0
    f["Ce"][:] = 2.5e-3
    t3.fill_ghost_layers(f, (1, 1, 1))
    c = (slice(1, -1),) * 3
    act = f["natoms"][c] >= 20
    s0 = (f["temp"][c] * act).sum()
    t3.run(f, (16.0, 16.0, 16.0), 0.1 / 250, 20, 250, 0, 6, (1, 1, 1))
    assert (f["temp"][c] * act).sum() == pytest.approx(s0, rel=1e-13)
    g = t3.make_case(shape, seed=2, hole=True, pbc=(1, 1, 1))
    g["fd_g"][:] = 0.0
    g["source"][:] = 0.0
    g["temp"][:] = 0.3
    t3.run(g, (16.0, 16.0, 16.0), 0.1 / 250, 20, 250, 0, 3, (1, 1, 1))
    assert np.array_equal(g["temp"], np.full_like(g["temp"], 0.3))
