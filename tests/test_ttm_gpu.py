"""GPU parity of the 1-D TTM path (calc_ttm, move_atoms_ttm) through the C ABI, against
(a) dumps of the unmodified reference (tests/golden/ttm_*.npz) and (b) the CPU oracle on
seeded inputs.  Tolerances: tests/ttm_util.py."""
import numpy as np
import pytest

from imd_b200 import host, synth
from tests.ttm_util import (REL_STEP0, REL_TE, TE_ABS_K, TMM_CASES, TTM_CASES, TTM_FIELDS, TtmGolden, compare_cells)

pytestmark = pytest.mark.gpu


def make_imd(g, atoms=None, nranks=None):
    a = g.atoms(0) if atoms is None else atoms
    imd = host.Imd(0)
    imd.setup_potentials(g.tables)
    imd.make_box(g.box, (0, 1, 1), g.margin)
    imd.set_params(1, g.timestep)
    if g.pdecay:
        imd.ctx.set_pdecay(*g.pdecay)
    imd.set_atoms(a["pos"], a["mom"], a["mass"], a["type"], a["id"])
    imd.init_ttm(g.tp, g.eos, g.nranks if nranks is None else nranks)
    return imd


@pytest.mark.parametrize("name", TTM_CASES)
def test_calc_ttm_step0_matches_reference(name):
    """Step 0 has no root finding on the path: every field to 1e-10 (we see ~1e-15)."""
    g = TtmGolden(name)
    imd = make_imd(g)
    imd.calc_forces(0)
    got = imd.atoms(fields=("id", "numneighs", "force"))
    a = g.atoms(0)
    assert np.array_equal(got["numneighs"], a["numneighs"])
    sc = imd.calc_ttm(0)
    ref = g.ttm(0)
    hdr = ref["hdr"][0]
    assert int(sc["diff_substeps"]) == int(hdr[4])
    assert int(sc["last_active_cell_global"]) == int(hdr[6])
    assert int(sc["cur_vlattice_proc"]) == int(hdr[7])
    np.testing.assert_allclose(sc["max_dt_ttm"], hdr[5], rtol=1e-14)
    np.testing.assert_allclose(sc["Eabs"], float(ref["Eabs"]), rtol=1e-11)
    np.testing.assert_allclose(sc["tot_elec_energy"], float(ref["tot_elec_energy"]), rtol=1e-11)
    lat = imd.ttm_lattice()
    compare_cells(lat["fields"], ref["fields"], 0, int(hdr[6]), name)
    pscale = np.sqrt(g.spec.mass * 0.0258) / g.spec.mass
    np.testing.assert_allclose(lat["vcom"], ref["vcom"], rtol=0, atol=1e-11 * pscale)
    np.testing.assert_allclose(lat["xiarr"], ref["xiarr"], rtol=REL_STEP0, atol=1e-18)
    np.testing.assert_allclose(lat["vlat"][:, :5], ref["vlat"][:, :5], rtol=REL_STEP0)
    imd.close()


@pytest.mark.parametrize("name", TTM_CASES + TMM_CASES)
def test_full_loop_matches_reference(name):
    """calc_forces -> calc_ttm -> move_atoms_ttm on the device for the whole reference run,
    compared at every dumped step: atoms, numneighs (bit exact) and the FD lattice."""
    g = TtmGolden(name)
    imd = make_imd(g)
    box = np.diag(g.box)
    s = 0
    for target in g.steps:
        # run up to (not including) `target`, then do its force + ttm part step by step
        if target > s:
            imd.main_loop_ttm(target - s)
            s = target
        imd.calc_forces(s)
        sc = imd.calc_ttm(s)
        got = imd.atoms(fields=("id", "pos", "mom", "force", "numneighs"))
        a, ref = g.atoms(s), g.ttm(s)
        d = got["pos"] - a["pos"]
        d[:, 1:] -= np.round(d[:, 1:] / box[1:]) * box[1:]
        assert np.abs(d).max() < 1e-9
        assert np.array_equal(got["numneighs"], a["numneighs"])
        fs = np.abs(a["force"]).max()
        assert np.abs(got["force"] - a["force"]).max() < 1e-9 * fs
        hdr = ref["hdr"][0]
        assert int(sc["diff_substeps"]) == int(hdr[4])
        np.testing.assert_allclose(sc["Eabs"], float(ref["Eabs"]), rtol=1e-9)
        compare_cells(imd.ttm_lattice()["fields"], ref["fields"], s, int(hdr[6]), name)
        imd.move_atoms_ttm()
        s += 1
        imd.steps = s
    imd.close()


def test_host_buffer_steps_equal_resident_steps():
    """imdcu_update_atoms uploads the momenta on a second stream beside the force passes; every reader of
    the momenta (cell sort, update_fd, NRB, integrators, downloads) waits for them.  A run that hands
    positions and momenta through host buffers every step must therefore equal the resident run bit for
    bit -- atoms and FD lattice -- across neighbour-list rebuilds."""
    g = TtmGolden("ttm_al_48x5x5_p2_hot")
    nsteps = 30

    def run(through_host):
        imd = make_imd(g)
        n = imd.natoms
        hp, hm = np.empty((n, 3)), np.empty((n, 3))
        imd.ctx.download_into(hp, hm)
        for s in range(nsteps):
            if through_host:
                imd.ctx.update_atoms(hp, hm)
            imd.calc_forces(s)
            imd.calc_ttm(s)
            imd.move_atoms_ttm()
            if through_host:
                imd.ctx.download_into(hp, hm)
        out = imd.atoms(fields=("id", "pos", "mom")), {k: np.array(v) for k, v in imd.ttm_lattice()["fields"].items()}, imd.nbl_count
        imd.close()
        return out

    (a, lat_a, nb_a), (b, lat_b, nb_b) = run(False), run(True)
    assert nb_a == nb_b and nb_a >= 2, (nb_a, nb_b)
    assert np.array_equal(a["pos"], b["pos"]) and np.array_equal(a["mom"], b["mom"])
    for k in lat_a:
        assert np.array_equal(lat_a[k], lat_b[k], equal_nan=True), k


def test_against_oracle_seeded_film():
    """A film the reference never saw (different size, seed, slab count, emulated rank count):
    device vs CPU oracle, step by step for a few steps."""
    from oracle.binding import EamOracle, TtmOracle
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    g0 = TtmGolden(TTM_CASES[0])                       # only for the EOS diagonal flags
    eos = g0.eos
    pos, box = synth.film_lattice((40, 5, 5), spec.a, 18.0, 22.0)
    rng = np.random.default_rng(5)
    pos = pos + rng.normal(0, 0.03, pos.shape)
    n = pos.shape[0]
    mom = synth.maxwell_momenta(n, spec.mass, 0.04, seed=11)
    mass = np.full(n, spec.mass); types = np.zeros(n, np.int32); ids = np.arange(n, dtype=np.int32)
    tp = synth.TtmParams(ttmdimx=12, fd_n_timesteps=25, fd_min_atoms=15, vlatdim=6, vlatbuffer=1, I0=3.0e16)
    P = 3
    imd = host.Imd(0)
    imd.setup_potentials(tables)
    imd.make_box(box, (0, 1, 1), 0.4)
    imd.set_params(1, 0.1)
    imd.set_atoms(pos, mom, mass, types, ids)
    imd.init_ttm(tp, eos, P)
    eam = EamOracle(box, (0, 1, 1), tables, 0.4)
    ttm = TtmOracle(tp, eos, box, float(tables["pair"].end[0]), 0.1, P)
    opos, omom = pos.copy(), mom.copy()
    fdi = np.zeros(n, np.int32)
    for s in range(4):
        eam.fix_cells(opos); eam.make_nblist()
        f = eam.calc_forces(opos, types)
        ttm.step(s, opos, omom, mass, types, f["numneighs"], fdi)
        imd.calc_forces(s)
        sc = imd.calc_ttm(s)
        osc = ttm.scalars()
        assert int(sc["diff_substeps"]) == int(osc["diff_substeps"])
        assert int(sc["last_active_cell_global"]) == int(osc["last_active_cell_global"])
        lat, oc = imd.ttm_lattice(), ttm.cells()
        compare_cells(lat["fields"], oc, s, int(osc["last_active_cell_global"]), "seeded")
        np.testing.assert_allclose(sc["Eabs"], osc["Eabs"], rtol=1e-9)
        ttm.move_atoms(opos, omom, f["force"], mass, types)
        imd.move_atoms_ttm()
        got = imd.atoms(fields=("id", "pos", "mom"))
        d = got["pos"] - opos
        bx = np.diag(box)
        d[:, 1:] -= np.round(d[:, 1:] / bx[1:]) * bx[1:]
        assert np.abs(d).max() < 1e-10
        np.testing.assert_allclose(got["mom"], omom, rtol=0, atol=1e-10 * np.sqrt(spec.mass * 0.04))
    assert TE_ABS_K > 0 and len(TTM_FIELDS) == 13
    imd.close(); eam.close(); ttm.close()


def test_ttm_error_paths():
    g = TtmGolden(TTM_CASES[0])
    imd = host.Imd(0)
    imd.setup_potentials(g.tables)
    imd.make_box(g.box, (0, 1, 1), g.margin)
    imd.set_params(1, g.timestep)
    from imd_b200 import capi
    with pytest.raises(capi.ImdcuError):                 # EOS tables not set
        imd.ctx.ttm_init(g.tp, 2)
    imd.ctx.ttm_set_eos(0, g.eos["cve"]); imd.ctx.ttm_set_eos(1, g.eos["ee"])
    with pytest.raises(capi.ImdcuError):                 # global_fd_dim.x mod num_cpus != 0
        imd.ctx.ttm_init(g.tp, 5)
    with pytest.raises(capi.ImdcuError):                 # calc_ttm before init
        imd.ctx.calc_ttm(0)
    # do_tmm: K12 needs its 1500 rows, wavelengths must ascend, the phase table must be there, lambda inside eps_bb
    tabs = synth.make_tmm_tables()
    with pytest.raises(capi.ImdcuError):
        imd.ctx.ttm_set_tmm(5.0, tabs["eps_bb"], tabs["K12"][:100])
    with pytest.raises(capi.ImdcuError):
        imd.ctx.ttm_set_tmm(5.0, tabs["eps_bb"][::-1], tabs["K12"])
    imd.ctx.ttm_set_tmm(5.0, tabs["eps_bb"], tabs["K12"])
    with pytest.raises(capi.ImdcuError):                 # IMDCU_EOS_PHASE missing
        imd.ctx.ttm_init(g.tp, 2)
    imd.ctx.ttm_set_eos(2, g.eos["phase"])
    import dataclasses
    with pytest.raises(capi.ImdcuError):                 # 5 um is outside the interband table
        imd.ctx.ttm_init(dataclasses.replace(g.tp, lam=5e-6), 2)
    imd.ctx.ttm_init(g.tp, 2)
    imd.close()


def test_tmm_sample_too_short():
    """The reference stops with "Sample is too short for TMM" when the field never drops below
    exp(-tmm_threshold) inside the FD lattice (imd_tmm.c:335-345); the library reports the same at calc_ttm."""
    import dataclasses
    from imd_b200 import capi
    g = TtmGolden(TMM_CASES[0])
    g.tp = dataclasses.replace(g.tp, tmm_threshold=40.0)
    imd = make_imd(g)
    with pytest.raises(capi.ImdcuError, match="too short"):
        for s in range(4):
            imd.calc_forces(s)
            imd.calc_ttm(s)
            imd.move_atoms_ttm()
    imd.close()


def test_device_math():
    """log/exp/pow/div/sqrt/rcp of ttm_math.cuh as compiled for the device (hardware approximations
    refined by Newton steps) against libm / IEEE over the magnitudes the wide-range models produce."""
    from imd_b200 import capi
    c = capi.Context(0)
    rng = np.random.default_rng(5)
    n = 300_000
    ulps = lambda got, ref: np.abs(got - ref) / np.spacing(np.abs(ref))        # noqa: E731
    x = np.exp(rng.uniform(-160, 160, n)) * rng.choice([1.0, 1.0], n)
    y = np.exp(rng.uniform(-160, 160, n)) * rng.choice([-1.0, 1.0], n)
    assert ulps(c.eval_ttm_math("div", x, y), x / y).max() <= 1.0
    assert ulps(c.eval_ttm_math("rcp", y), 1.0 / y).max() <= 1.0
    assert ulps(c.eval_ttm_math("sqrt", x), np.sqrt(x)).max() <= 1.0
    assert ulps(c.eval_ttm_math("log", x), np.log(x)).max() <= 2.0
    e = rng.uniform(-700, 700, n)
    assert ulps(c.eval_ttm_math("exp", e), np.exp(e)).max() <= 1.0
    xb = np.exp(rng.uniform(-60, 60, n))
    yb = rng.uniform(-4, 4, n)
    ref = np.power(xb, yb)
    assert (np.abs(c.eval_ttm_math("pow", xb, yb) - ref) / ref <= 2.0 ** -51 * (1 + np.abs(yb * np.log(xb)))).all()
    c.close()


def test_full_size_fused_schedule_equals_sequential_calls():
    """BASELINE.json configs[2] at full size (Al film, 100^3 cubic cells = 4 000 000 atoms, 1-D TTM): the
    main-loop entry point (calc_ttm forked between the EAM passes, lattice kernel beside pass 2) gives what
    the three separate entry points give one after the other.
    The atom->slab sums end in floating-point atomics (one set per CTA), so lattice fields agree to round-off,
    not bit for bit."""
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    pos, box = synth.film_lattice((100, 100, 100), spec.a, 10 * spec.a, 10 * spec.a)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=3)
    eos = synth.make_eos_tables()
    gx = int(box[0, 0] / 16.0)
    tp = synth.TtmParams(ttmdimx=gx, fd_n_timesteps=250, fd_min_atoms=100, vlatdim=30, vlatbuffer=8,
                         I0=5.0e15, laser_t_0=100e-15, laser_sigma_t=42.5e-15)
    nsteps = 4

    def run(fused):
        imd = host.Imd(0)
        imd.setup_potentials(tables)
        imd.make_box(box, (0, 1, 1), 0.4)
        imd.set_params(1, 0.1)
        imd.set_atoms(pos, mom, spec.mass)
        imd.init_ttm(tp, eos, 1)
        if fused:
            imd.main_loop_ttm(nsteps)
        else:
            for s in range(nsteps):
                imd.calc_forces(s)
                imd.calc_ttm(s)
                imd.move_atoms_ttm()
        at = imd.atoms(fields=("id", "pos", "mom"))
        lat = {k: np.array(v) for k, v in imd.ttm_lattice()["fields"].items()}
        imd.close()
        return at, lat

    (a, la), (b, lb) = run(True), run(False)
    assert np.array_equal(la["natoms"], lb["natoms"]) and int(la["natoms"].max()) > 100000
    assert np.abs(a["pos"] - b["pos"]).max() < 1e-10
    assert np.abs(a["mom"] - b["mom"]).max() < 1e-9 * np.abs(a["mom"]).max()
    act = la["natoms"] >= tp.fd_min_atoms
    # Te comes out of the reference's Brent minimiser (1e-5 K, iteration path follows the last bits of its
    # inputs): fields downstream of Te to 2e-7, xi ~ (Te - Ti) to 1e-5, lattice quantities 1e-9
    for k, tol in (("md_temp", 1e-9), ("dens", 1e-9), ("temp", 2e-7), ("U", 2e-7), ("xi", 1e-5)):
        ref = np.abs(lb[k][act]).max() + 1e-300
        assert np.abs(la[k][act] - lb[k][act]).max() < tol * ref, k


def test_full_size_calc_ttm_vs_oracle():
    """BASELINE.json configs[2] at full size (Al film, 100^3 cubic cells = 4 000 000 atoms, 250 sub-steps, pulse on):
    calc_ttm on the device against the CPU restatement of imd_ttm.c (update_fd, do_ADV, FILLMESH, CFL, do_DIFF) fed
    with the SAME atom state every step -- the device's positions, momenta and numneighs (the latter checked
    bit for bit against the oracle's make_nblist/calc_forces at ~1 M atoms in test_eam_gpu.py).  This is the
    lattice kernel and the atom->slab kernels at the size bench.py times, not a self-comparison."""
    from oracle.binding import TtmOracle
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    pos, box = synth.film_lattice((100, 100, 100), spec.a, 10 * spec.a, 10 * spec.a)
    n = pos.shape[0]
    mom = synth.maxwell_momenta(n, spec.mass, 0.0258, seed=3)
    eos = synth.make_eos_tables()
    gx = int(box[0, 0] / 16.0)
    tp = synth.TtmParams(ttmdimx=gx, fd_n_timesteps=250, fd_min_atoms=100, vlatdim=30, vlatbuffer=8,
                         I0=5.0e15, laser_t_0=3e-15, laser_sigma_t=42.5e-15)       # pulse peak inside the 3 steps
    P = 2
    imd = host.Imd(0)
    imd.setup_potentials(tables)
    imd.make_box(box, (0, 1, 1), 0.4)
    imd.set_params(1, 0.1)
    imd.set_atoms(pos, mom, spec.mass)
    imd.init_ttm(tp, eos, P)
    ttm = TtmOracle(tp, eos, box, float(tables["pair"].end[0]), 0.1, P)
    mass = np.full(n, spec.mass); types = np.zeros(n, np.int32)
    fdi = np.zeros(n, np.int32)
    for s in range(3):
        imd.calc_forces(s)
        a = imd.atoms(fields=("id", "pos", "mom", "numneighs"))
        ttm.step(s, a["pos"], a["mom"], mass, types, a["numneighs"], fdi)
        sc, osc = imd.calc_ttm(s), ttm.scalars()
        assert int(sc["diff_substeps"]) == int(osc["diff_substeps"])
        assert int(sc["last_active_cell_global"]) == int(osc["last_active_cell_global"])
        assert int(sc["laser_active"]) == int(osc["laser_active"])
        np.testing.assert_allclose(sc["Eabs"], osc["Eabs"], rtol=1e-9)
        lat, oc, oa = imd.ttm_lattice(), ttm.cells(), ttm.arrays()
        compare_cells(lat["fields"], oc, s, int(osc["last_active_cell_global"]), "full size 4M")
        np.testing.assert_allclose(lat["xiarr"], oa["xiarr"], rtol=REL_TE if s else REL_STEP0,
                                   atol=(REL_TE if s else 1e-14) * np.abs(oa["xiarr"]).max())
        np.testing.assert_allclose(lat["vcom"], oa["vcom"], rtol=0, atol=1e-9 * np.abs(oa["vcom"]).max())
        imd.move_atoms_ttm()
    assert int(lat["fields"]["natoms"].max()) > 100000
    imd.close(); ttm.close()


def test_fused_forces_ttm_call_equals_the_two_calls():
    """imdcu_calc_forces_ttm (calc_forces + calc_ttm in one call, lattice kernel beside EAM pass 2: what the IMD stub
    uses) against imdcu_calc_forces followed by imdcu_calc_ttm, step by step on the same golden start: scalars, atoms
    and FD lattice bit for bit."""
    g = TtmGolden(TTM_CASES[0])
    runs = []
    for fused in (False, True):
        a = g.atoms(0)
        imd = host.Imd(0)
        imd.setup_potentials(g.tables)
        imd.make_box(g.box, (0, 1, 1), g.margin)
        imd.set_params(1, g.timestep)
        if g.pdecay:
            imd.ctx.set_pdecay(*g.pdecay)
        imd.set_atoms(a["pos"], a["mom"], a["mass"], a["type"], a["id"])
        imd.init_ttm(g.tp, g.eos, g.nranks)
        hist = []
        for step in range(8):
            if fused:
                o8, sc = imd.ctx.calc_forces_ttm(step)
            else:
                o8 = imd.ctx.calc_forces(step)
                sc = imd.ctx.calc_ttm(step)
            imd.ctx.move_atoms_ttm()
            hist.append((o8.copy(), dict(sc)))
        runs.append((hist, imd.atoms(("id", "pos", "mom", "force")), imd.ttm_lattice()))
        imd.close()
    (h0, a0, l0), (h1, a1, l1) = runs
    for (o0, s0), (o1, s1) in zip(h0, h1):
        assert np.array_equal(o0, o1)
        assert all(s0[k] == s1[k] or (np.isnan(s0[k]) and np.isnan(s1[k])) for k in s0), (s0, s1)
    for k in ("pos", "mom", "force"):
        assert np.array_equal(a0[k], a1[k]), k
    def same_bits(x0, x1, key):                                   # FD cells, xi, v_com, virtual lattice: the same bits
        if isinstance(x0, dict):
            assert set(x0) == set(x1), key
            for k in x0:
                same_bits(x0[k], x1[k], "%s.%s" % (key, k))
        else:
            x0, x1 = np.asarray(x0), np.asarray(x1)
            assert x0.shape == x1.shape and x0.tobytes() == x1.tobytes(), key
    same_bits(l0, l1, "lattice")
