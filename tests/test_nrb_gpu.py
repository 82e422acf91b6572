"""GPU parity of the non-reflecting boundary (imd_b200/csrc/nrb.cu: nrb_forces, the integrator hooks)
through the C ABI, against dumps of the unmodified reference built with -DNRB -DREFPOS
(tests/golden/ttm_nrb_*.npz) and against the CPU oracle (oracle/nrb_np.py) on a seeded film.

Tolerances: a boundary momentum is a sum of 1 + 4 terms; the device adds the neighbour terms in the order
of its own list, the reference in the order of its half list: 1e-12 relative to the largest momentum.
steps <= 1 (image multiplication, exact doublings): bit-exact.  Whole runs: the TTM tolerances of
tests/ttm_util.py."""
import numpy as np
import pytest

from imd_b200 import host, nrb, synth
from tests.ttm_util import NRB_CASES, TtmGolden, compare_cells

pytestmark = pytest.mark.gpu
PBC = (0, 1, 1)


def make_imd(g, s, with_ttm):
    a = g.atoms(s)
    imd = host.Imd(0)
    imd.setup_potentials(g.tables)
    imd.make_box(g.box, PBC, g.margin)
    imd.set_params(1, g.timestep)
    imd.set_atoms(a["pos"], g.pre_mom(s), a["mass"], a["type"], a["id"])
    if with_ttm:
        imd.init_ttm(g.tp, g.eos, g.nranks)
    imd.set_nrb(g.nrb["nrb_k"], g.nrb_state(), startup_images=True)
    return imd


@pytest.mark.parametrize("name", NRB_CASES)
def test_nrb_forces_matches_reference(name):
    """One calc_forces(steps) from the reference's state before it, at every dumped step."""
    g = TtmGolden(name)
    bnd = g.nrb_state().bnd > 0
    for s in g.steps:
        imd = make_imd(g, s, False)
        imd.calc_forces(s)
        got = imd.atoms(fields=("id", "mom", "force"))
        a = g.atoms(s)
        assert np.array_equal(got["id"], a["id"])
        assert np.array_equal(got["mom"][~bnd], g.pre_mom(s)[~bnd])
        if s <= 1:
            assert np.array_equal(got["mom"], a["mom"]), (name, s)
        else:
            scale = np.abs(a["mom"]).max()
            assert np.abs(got["mom"] - a["mom"]).max() < 1e-12 * scale, (name, s)
            assert np.abs(got["mom"][bnd] - g.pre_mom(s)[bnd]).max() > 1e-3 * scale      # it did something
        imd.close()


@pytest.mark.parametrize("name", NRB_CASES)
def test_full_loop_with_nrb_matches_reference(name):
    """calc_forces (with nrb_forces) -> calc_ttm -> move_atoms_ttm for the whole reference run."""
    g = TtmGolden(name)
    imd = make_imd(g, 0, True)
    box = np.diag(g.box)
    bnd = g.nrb_state().bnd > 0
    s = 0
    for target in g.steps:
        if target > s:
            imd.main_loop_ttm(target - s)
            s = target
        imd.calc_forces(s)
        sc = imd.calc_ttm(s)
        got = imd.atoms(fields=("id", "pos", "mom", "force", "numneighs"))
        a, ref = g.atoms(s), g.ttm(s)
        d = got["pos"] - a["pos"]
        d[:, 1:] -= np.round(d[:, 1:] / box[1:]) * box[1:]
        assert np.abs(d).max() < 1e-9, (name, s)
        assert np.array_equal(got["numneighs"], a["numneighs"])
        assert np.abs(got["force"] - a["force"]).max() < 1e-9 * np.abs(a["force"]).max()
        assert np.abs(got["mom"] - a["mom"]).max() < 1e-9 * np.abs(a["mom"]).max(), (name, s)
        hdr = ref["hdr"][0]
        assert int(sc["diff_substeps"]) == int(hdr[4])
        compare_cells(imd.ttm_lattice()["fields"], ref["fields"], s, int(hdr[6]), name)
        p_before = got["mom"][bnd]
        imd.move_atoms_ttm()
        after = imd.atoms(fields=("id", "mom"))
        assert np.array_equal(after["mom"][bnd], p_before)              # boundary atoms: no kick, no heating
        s += 1
        imd.steps = s
    imd.close()


def test_nve_with_nrb_against_oracle():
    """A film the reference never saw, NVE integrator, NRB state from the host mirror of init_nrb:
    device vs EAM oracle + numpy nrb_forces + the oracle's NVE step (boundary atoms as a virtual type with
    restriction 0 0 0, which is what 'no kick' amounts to without PDECAY)."""
    from oracle import nrb_np
    from oracle.binding import EamOracle
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    pos, box = synth.film_lattice((14, 5, 6), spec.a, 12.0, 14.0)
    n = pos.shape[0]
    ids = np.arange(1, n + 1, dtype=np.int32)
    st = nrb.init_nrb(pos, ids, box, PBC, spec.a, 0.3)
    assert (st.bnd == 1).sum() == 2 * 5 * 6 and set(np.unique(st.bnd)) == {0, 1}
    rng = np.random.default_rng(3)
    pos = pos + rng.normal(0, 0.04, pos.shape)
    mom = synth.maxwell_momenta(n, spec.mass, 0.06, seed=5)
    mass = np.full(n, spec.mass)
    types = np.zeros(n, np.int32)
    nrb_k = 1.1
    imd = host.Imd(0)
    imd.setup_potentials(tables)
    imd.make_box(box, PBC, 0.4)
    imd.set_params(1, 0.1)
    imd.set_atoms(pos, mom, mass, types, ids)
    imd.set_nrb(nrb_k, st)                                # no startup images: steps <= 1 leave momenta alone
    orc = EamOracle(box, PBC, tables, 0.4)
    op, om = pos.copy(), mom.copy()
    L = np.diag(box)
    vt = (st.bnd > 0).astype(np.int32)
    rs = np.array([[1.0, 1, 1], [0.0, 0, 0]])
    ref0 = op.copy()
    valid = False
    for s in range(6):
        if not valid:
            orc.fix_cells(op); orc.make_nblist(); ref0 = op.copy()
        f = orc.calc_forces(op, types)
        if s > 1:
            om = nrb_np.nrb_forces(ids, op, om, mass, st.bnd, st.nrbi, st.refpos, L, PBC, nrb_k)
        imd.calc_forces(s)
        got = imd.atoms(fields=("id", "mom"))
        o = np.argsort(got["id"])
        assert np.abs(got["mom"][o] - om).max() < 1e-11 * np.abs(om).max(), s
        orc.move_atoms_nve(op, om, f["force"], mass, 0.1, vtype=vt, restrictions=rs)
        valid = orc.max_displacement2(op, ref0) <= (0.5 * 0.4) ** 2
        imd.move_atoms()
        got = imd.atoms(fields=("id", "pos", "mom"))
        o = np.argsort(got["id"])
        d = got["pos"][o] - op
        d[:, 1:] -= np.round(d[:, 1:] / L[1:]) * L[1:]
        assert np.abs(d).max() < 1e-11, s
        assert np.abs(got["mom"][o] - om).max() < 1e-11 * np.abs(om).max(), s
    imd.set_nrb(nrb_k, None)                              # off again: the next step kicks every atom
    before = imd.atoms(fields=("id", "mom"))
    imd.calc_forces(6); imd.move_atoms()
    after = imd.atoms(fields=("id", "mom"))
    ob, oa = np.argsort(before["id"]), np.argsort(after["id"])
    assert (np.abs(after["mom"][oa] - before["mom"][ob]).max(axis=1)[st.bnd > 0] > 0).all()
    imd.close(); orc.close()


def test_set_nrb_errors():
    g = TtmGolden(NRB_CASES[0])
    a = g.atoms(0)
    imd = host.Imd(0)
    imd.setup_potentials(g.tables)
    imd.make_box(g.box, PBC, g.margin)
    imd.set_params(1, g.timestep)
    st = g.nrb_state()
    with pytest.raises(RuntimeError):
        imd.set_nrb(1.2, st)                              # atoms first
    imd.set_atoms(a["pos"], a["mom"], a["mass"], a["type"], a["id"])
    ids, bnd, nrbi, refpos = st.entries()
    with pytest.raises(RuntimeError):                     # a listed neighbour without an entry
        m = bnd > 0
        imd.ctx.set_nrb(1.2, ids[m], bnd[m], nrbi[m], refpos[m])
    with pytest.raises(RuntimeError):
        imd.ctx.set_nrb(-1.0, ids, bnd, nrbi, refpos)
    with pytest.raises(RuntimeError):
        imd.ctx.set_nrb(1.2, np.concatenate([ids, ids[:1]]), np.concatenate([bnd, bnd[:1]]),
                        np.concatenate([nrbi, nrbi[:1]]), np.concatenate([refpos, refpos[:1]]))
    imd.ctx.set_nrb(1.2, ids, bnd, nrbi, refpos)
    imd.close()
