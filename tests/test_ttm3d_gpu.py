"""3-D TTM sweep on the GPU (imd_b200/csrc/ttm3d.cu, through the C ABI) against the reference's
golden outputs and the numpy restatement: bit-exact for temp, U, xi, natoms on every cell the
path defines (interior + ghost faces); Eabs to 1e-12 (different summation order).  Full-size
(BASELINE.json configs[4] is 512^3) checks use size-independent properties."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import ttm3d_np as t3  # noqa: E402
from ttm3d_util import GOLDEN, faces_and_interior, load_golden  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from imd_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


def gpu_run(ctx, f, case):
    ctx.ttm3d_init(case["shape"], case["h"], case["pbc"], case["fd_min_atoms"])
    ctx.ttm3d_upload(**f)
    e = ctx.ttm3d_diff(case["tau"], case["diff_substeps"], case["laser_active"], case["nsweeps"])
    return e, ctx.ttm3d_download()


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_gpu_equals_reference_golden(ctx, path):
    case, ref = load_golden(path)
    f = t3.make_case(case["shape"], seed=case["seed"], hole=case["hole"], pbc=case["pbc"], fd_min_atoms=case["fd_min_atoms"])
    e, out = gpu_run(ctx, f, case)
    m = faces_and_interior(ref["temp"].shape)
    for k in ("natoms", "temp", "U", "xi"):
        assert np.array_equal(out[k][m], ref[k][m]), k
    assert e == pytest.approx(ref["eabs"], rel=1e-12)


@pytest.mark.parametrize("shape,pbc", [((1, 1, 1), (1, 1, 1)), ((1, 40, 3), (0, 1, 1)), ((33, 17, 65), (0, 0, 1)),
                                       ((64, 64, 64), (0, 1, 1)), ((130, 47, 96), (1, 1, 0))])
def test_gpu_equals_oracle(ctx, shape, pbc):
    """ragged tiles (not multiples of 32 x 16), single cells, every open/periodic mix, x chunks."""
    case = dict(shape=shape, seed=21, hole=min(shape) >= 4, pbc=pbc, fd_min_atoms=20, h=(15.0, 16.0, 17.0),
                tau=0.1 / 250, diff_substeps=250, laser_active=1, nsweeps=4)
    f = t3.make_case(shape, seed=21, hole=case["hole"], pbc=pbc)
    e, out = gpu_run(ctx, {k: v.copy() for k, v in f.items()}, case)
    e_o = t3.run(f, case["h"], case["tau"], 20, 250, 1, 4, pbc)
    m = faces_and_interior(f["temp"].shape)
    for k in ("natoms", "temp", "U", "xi"):
        assert np.array_equal(out[k][m], f[k][m]), k
    assert e == pytest.approx(e_o, rel=1e-12)


def test_sweeps_compose(ctx):
    """5 sweeps in one call == 2 + 3 sweeps in two calls (lattice swap state is kept)."""
    case = dict(shape=(40, 33, 34), seed=4, hole=True, pbc=(0, 1, 1), fd_min_atoms=20, h=(16.0, 16.0, 16.0),
                tau=0.1 / 250, diff_substeps=250, laser_active=1, nsweeps=5)
    f = t3.make_case(case["shape"], seed=4, pbc=case["pbc"])
    e1, a = gpu_run(ctx, f, case)
    ctx.ttm3d_init(case["shape"], case["h"], case["pbc"], 20)
    ctx.ttm3d_upload(**f)
    ctx.ttm3d_diff(case["tau"], 250, 1, 2)
    e2 = ctx.ttm3d_diff(case["tau"], 250, 1, 3)
    b = ctx.ttm3d_download()
    for k in ("temp", "U", "xi"):
        assert np.array_equal(a[k], b[k])
    assert e1 == pytest.approx(e2, rel=1e-14)


def test_full_size_properties(ctx):
    """256 x 512 x 512 lattice (half of configs[4], 67 M cells): (i) a uniform temperature is a fixed
    point; (ii) without coupling/source and with uniform Ce the sweep conserves sum(T) over the
    active cells; (iii) inactive cells never change."""
    shape = (256, 512, 512)
    e = tuple(s + 2 for s in shape)
    rng = np.random.default_rng(0)
    nat = np.full(e, 60, np.int32)
    nat[: shape[0] // 8] = 3                                            # vacuum in front of the film
    temp = (0.05 + 0.5 * rng.random(e, dtype=np.float32)).astype(np.float64)
    fd_k = (10.0 + 20.0 * rng.random(e, dtype=np.float32)).astype(np.float64)
    const = lambda v: np.full(e, v)                                     # noqa: E731
    ctx.ttm3d_init(shape, (16.0, 16.0, 16.0), (0, 1, 1), 20)
    ctx.ttm3d_upload(natoms=nat, temp=const(0.3), fd_k=fd_k, fd_g=const(0.0), Ce=const(2.5e-3), md_temp=const(0.03),
                     source=const(0.0), U=const(1.0), xi=const(0.0), dens=const(2700.0))
    ctx.ttm3d_diff(0.1 / 250, 250, 0, 3, want_eabs=False)
    out = ctx.ttm3d_download(("temp",))
    assert np.array_equal(out["temp"][1:-1, 1:-1, 1:-1], np.full(shape, 0.3))
    ctx.ttm3d_upload(temp=temp)
    ctx.ttm3d_fill_ghost_layers()                                       # uploaded ghosts are not periodic images
    c = (slice(1, -1),) * 3
    act = nat[c] >= 20
    s0 = temp[c][act].sum()
    ctx.ttm3d_diff(0.1 / 250, 250, 0, 4, want_eabs=False)
    out = ctx.ttm3d_download(("temp", "U"))
    assert out["temp"][c][act].sum() == pytest.approx(s0, rel=1e-12)
    assert np.array_equal(out["temp"][c][~act], temp[c][~act])
    assert np.abs(out["temp"][c][act] - temp[c][act]).max() > 1e-6      # it did diffuse


def test_errors(ctx):
    from imd_b200 import capi
    c2 = capi.Context(0)
    with pytest.raises(capi.ImdcuError):
        c2.ttm3d_diff(1e-3, 1, 0, 1)                                    # not initialised
    with pytest.raises(capi.ImdcuError):
        c2.ttm3d_init((0, 4, 4), (1.0, 1.0, 1.0), (1, 1, 1), 1)
    c2.close()
