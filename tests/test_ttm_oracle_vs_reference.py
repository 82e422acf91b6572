"""Pins the CPU restatement of the 1-D TTM (oracle/ttm_oracle.c) against dumps of the
UNMODIFIED reference's -DMPI -DTTM build run over shmpi/ with 2 and 4 ranks
(tests/golden/ttm_*.npz, made by tests/golden/make_golden_ttm.py).  CPU only."""
import numpy as np
import pytest

from oracle.binding import EamOracle, Eos, TtmOracle
from tests.ttm_util import (REL_LATTICE, TE_ABS_K, TMM_CASES, TTM_CASES, TtmGolden, compare_cells)


@pytest.mark.parametrize("name", TTM_CASES[:1])
@pytest.mark.parametrize("key", ["cve", "ee"])
def test_eos_interpolation_matches_libnn(name, key):
    """orc_eos_interp == lpi_interpolate_point of the reference's own library (nn 1.86.2)."""
    g = TtmGolden(name)
    e = Eos(g.eos[key])
    pr, ref = g.z["eos_%s_probe_xy" % key], g.z["eos_%s_probe_z" % key]
    got = e.interp(pr[:, 0], pr[:, 1])
    assert not np.isnan(ref).any()
    np.testing.assert_allclose(got, ref, rtol=1e-13, atol=0)


def make_oracle(g):
    return TtmOracle(g.tp, g.eos, g.box, float(g.tables["pair"].end[0]), g.timestep, g.nranks)


@pytest.mark.parametrize("name", TTM_CASES + TMM_CASES[:1])
def test_calc_ttm_on_reference_atoms(name):
    """calc_ttm fed with the reference's own atoms at consecutive steps 0,1[,2]."""
    g = TtmGolden(name)
    orc = make_oracle(g)
    n = g.atoms(0)["id"].shape[0]
    fdi = np.zeros(n, np.int32)
    consecutive = [s for i, s in enumerate(g.steps) if s == i]
    assert len(consecutive) >= 2
    for s in consecutive:
        a, ref = g.atoms(s), g.ttm(s)
        orc.step(s, a["pos"], a["mom"], a["mass"], a["type"], a["numneighs"], fdi)
        sc = orc.scalars()
        hdr = ref["hdr"][0]
        assert int(sc["diff_substeps"]) == int(hdr[4])
        assert int(sc["last_active_cell_global"]) == int(hdr[6])
        assert int(sc["cur_vlattice_proc"]) == int(hdr[7])
        np.testing.assert_allclose(sc["max_dt_ttm"], hdr[5], rtol=1e-7 if s else 1e-14)
        np.testing.assert_allclose(sc["Eabs"], float(ref["Eabs"]), rtol=1e-9 if s else 1e-12)   # the source follows Te (do_tmm)
        np.testing.assert_allclose(sc["tot_elec_energy"], float(ref["tot_elec_energy"]), rtol=1e-7 if s else 1e-12)
        compare_cells(orc.cells(), ref["fields"], s, int(hdr[6]), name)
        ar = orc.arrays()
        assert np.array_equal(ar["natomsglobal"], ref["natomsglobal"])
        assert np.array_equal(ar["fluxl"], ref["fluxl"]) and np.array_equal(ar["fluxr"], ref["fluxr"])
        pscale = np.sqrt(g.spec.mass * 0.0258) / g.spec.mass
        np.testing.assert_allclose(ar["vcom"], ref["vcom"], rtol=0, atol=1e-12 * pscale)
        np.testing.assert_allclose(ar["xiarr"], ref["xiarr"], rtol=1e-10 if s == 0 else 2e-7,
                                   atol=1e-18 if s == 0 else 2e-7 * np.abs(ref["xiarr"]).max())
        v, vref = orc.vlattice(), ref["vlat"]
        np.testing.assert_allclose(v[:, :5], vref[:, :5], rtol=1e-10 if s == 0 else 2e-7)
    orc.close()


def oracle_main_loop(g, nsteps, on_step):
    """The reference's main loop (imd_main_3d.c:151-916) with oracle pieces only:
    calc_forces (cells + list rebuilt every step; the list content does not change forces),
    calc_ttm, move_atoms_ttm."""
    a0 = g.atoms(0)
    pos, mom = a0["pos"].copy(), a0["mom"].copy()
    mass, types = a0["mass"].copy(), a0["type"].copy()
    n = pos.shape[0]
    eam = EamOracle(g.box, (0, 1, 1), g.tables, g.margin)
    ttm = make_oracle(g)
    fdi = np.zeros(n, np.int32)
    for s in range(nsteps + 1):
        eam.fix_cells(pos)
        eam.make_nblist()
        f = eam.calc_forces(pos, types)
        ttm.step(s, pos, mom, mass, types, f["numneighs"], fdi)
        on_step(s, pos, mom, f, ttm)
        ttm.move_atoms(pos, mom, f["force"], mass, types, pdecay=g.pdecay)
    eam.close(); ttm.close()


@pytest.mark.parametrize("name", TTM_CASES + TMM_CASES)
def test_full_loop_matches_reference(name):
    """EAM forces -> calc_ttm -> move_atoms_ttm for the whole run, oracle only, compared with
    the reference at every dumped step (atoms and FD lattice)."""
    g = TtmGolden(name)
    seen = []

    def check(s, pos, mom, f, ttm):
        if s not in g.steps:
            return
        seen.append(s)
        a, ref = g.atoms(s), g.ttm(s)
        box = np.diag(g.box)
        d = pos - a["pos"]
        d[:, 1:] -= np.round(d[:, 1:] / box[1:]) * box[1:]          # y,z are periodic; wraps happen at rebuilds
        assert np.abs(d).max() < 1e-9
        pscale = np.sqrt(g.spec.mass * 0.0258)
        assert np.abs(mom - a["mom"]).max() < 1e-9 * pscale * 10
        assert np.array_equal(f["numneighs"], a["numneighs"])
        fs = np.abs(a["force"]).max()
        assert np.abs(f["force"] - a["force"]).max() < 1e-9 * fs
        hdr = ref["hdr"][0]
        compare_cells(ttm.cells(), ref["fields"], s, int(hdr[6]), name)
        sc = ttm.scalars()
        assert int(sc["diff_substeps"]) == int(hdr[4])
        np.testing.assert_allclose(sc["Eabs"], float(ref["Eabs"]), rtol=1e-9)

    oracle_main_loop(g, g.steps[-1], check)
    assert seen == g.steps


def test_move_atoms_ttm_one_step():
    """move_atoms_ttm alone: reference atoms of step 0 -> reference atoms of step 1."""
    g = TtmGolden(TTM_CASES[2])                         # the case with the PDECAY absorber
    a0, a1 = g.atoms(0), g.atoms(1)
    orc = make_oracle(g)
    fdi = np.zeros(a0["id"].shape[0], np.int32)
    orc.step(0, a0["pos"], a0["mom"], a0["mass"], a0["type"], a0["numneighs"], fdi)
    pos, mom = a0["pos"].copy(), a0["mom"].copy()
    orc.move_atoms(pos, mom, a0["force"], a0["mass"], a0["type"], pdecay=g.pdecay)
    box = np.diag(g.box)
    d = pos - a1["pos"]
    d[:, 1:] -= np.round(d[:, 1:] / box[1:]) * box[1:]
    assert np.abs(d).max() < 1e-12
    np.testing.assert_allclose(mom, a1["mom"], rtol=1e-11, atol=1e-13)
    assert REL_LATTICE < 1e-8 and TE_ABS_K < 1e-3
