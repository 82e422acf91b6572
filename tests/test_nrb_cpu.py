"""CPU side of the non-reflecting boundary: the numpy restatement of nrb_forces (oracle/nrb_np.py) and
the host mirror of init_nrb (imd_b200/nrb.py) against dumps of the UNMODIFIED reference built with
-DNRB -DREFPOS (tests/golden/ttm_nrb_*.npz, made by tests/golden/make_golden_ttm.py)."""
import numpy as np
import pytest

from imd_b200 import nrb
from oracle import nrb_np
from tests.ttm_util import NRB_CASES, TtmGolden

PBC = (0, 1, 1)


@pytest.mark.parametrize("name", NRB_CASES)
def test_oracle_nrb_forces_matches_reference(name):
    g = TtmGolden(name)
    st = g.nrb_state()
    box = np.diag(g.box)
    assert (st.bnd > 0).sum() > 0
    for s in g.steps:
        a = g.atoms(s)
        if s <= 1:
            continue
        got = nrb_np.nrb_forces(a["id"], a["pos"], g.pre_mom(s), a["mass"], st.bnd, st.nrbi, st.refpos, box, PBC,
                                g.nrb["nrb_k"])
        scale = np.abs(a["mom"][st.bnd > 0]).max()
        assert np.abs(got - a["mom"]).max() < 1e-13 * max(scale, 1.0), (name, s)
        assert np.array_equal(got[st.bnd == 0], g.pre_mom(s)[st.bnd == 0])


@pytest.mark.parametrize("name", NRB_CASES)
def test_oracle_startup_images_matches_reference(name):
    """steps <= 1: the reference's send_forces multiplies boundary momenta by the image count (exactly)."""
    g = TtmGolden(name)
    st = g.nrb_state()
    box = np.diag(g.box)
    gdim = np.floor(box / np.sqrt((g.spec.r_c + g.margin) ** 2)).astype(int)
    seen = set()
    for s in [t for t in g.steps if t <= 1]:
        a = g.atoms(s)
        got = nrb_np.startup_images(a["pos"], g.pre_mom(s), st.bnd, box, PBC, gdim)
        assert np.array_equal(got, a["mom"]), (name, s)
        seen |= set(np.round(a["mom"][st.bnd > 0, 0] / g.pre_mom(s)[st.bnd > 0, 0]).astype(int))
    assert seen == {1, 2, 4}


@pytest.mark.parametrize("name", NRB_CASES)
def test_host_init_nrb_matches_reference(name):
    g = TtmGolden(name)
    ref = g.nrb_state()
    a = g.atoms(0)
    st = nrb.init_nrb(a["pos"], a["id"], g.box, PBC, g.nrb["nrb_alat"], g.nrb["nrb_eps"])
    assert np.array_equal(st.bnd, ref.bnd)
    assert np.array_equal(st.nrbi, ref.nrbi)
    assert np.array_equal(st.neigh, ref.neigh)
    assert np.array_equal(st.refpos, ref.refpos)
    ids, bnd, nrbi, refpos = st.entries()
    assert ids.shape[0] == (ref.bnd > 0).sum() + (ref.neigh > 0).sum()     # the +x layer and the one below
    assert set(np.unique(nrbi[bnd > 0])) - {0} <= set(ids)


def test_init_nrb_open_y_has_three_types():
    """open y: -y and +y faces become types 2 and 3, x keeps priority at the edges (imd_nrb.c:252-268)."""
    from imd_b200 import synth
    pos, box = synth.film_lattice((6, 5, 4), 4.05, 5.0, 5.0)
    ids = np.arange(1, pos.shape[0] + 1, dtype=np.int32)
    st = nrb.init_nrb(pos, ids, box, (0, 0, 1), 4.05, 0.3)
    assert set(np.unique(st.bnd)) == {0, 1, 2, 3}
    x1 = pos[:, 0] > pos[:, 0].max() - 1.0
    assert (st.bnd[x1] == 1).all()
    # a type-2 atom in the bulk of its face lists sites 0, 3, 6, 7 and nothing else
    m = (st.bnd == 2) & (pos[:, 0] > pos[:, 0].min() + 3.0) & (pos[:, 0] < pos[:, 0].max() - 3.0)
    assert m.any()
    assert ((st.nrbi[m] != 0) == np.isin(np.arange(12), (0, 3, 6, 7))[None, :]).all()
    with pytest.raises(ValueError):
        nrb.init_nrb(pos, ids - 1, box, (0, 0, 1), 4.05, 0.3)
