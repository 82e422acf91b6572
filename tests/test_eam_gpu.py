"""Parity tests proper: the CUDA path, called through the C ABI, against
(1) the golden dumps of the UNMODIFIED reference (tests/golden) and
(2) the oracle on seeded inputs.
Integer results (cell ids, neighbour index sets, numneighs) bit exact; FP64 results within
1e-10 relative (north_star), tolerance written below."""
import numpy as np
import pytest

from imd_b200 import host, synth
from tests.util import GOLDEN_CASES, INTERP_CASES, Golden, canon_pairs

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def rel_err(a, b, floor=1e-2):
    """max |a-b| relative to the scale of b; `floor` keeps a perfect lattice (all forces
    ~1e-15 by symmetry) from turning round-off into a relative error."""
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), floor)


# kernel instantiations (imdcu_set_tuning): lanes per atom of the EAM passes x interleaved-row pass 1.
# lpa = 1 with fused = 2 is what bench.py times (k_eam_pass1_rep4<1,896>: raw table rows in four bank-staggered
# copies; k_eam_pass2<1,...>); fused = 1: k_eam_pass1_fused (48-byte rows), 0: the general kernel; lpa 0 = by system size
VARIANTS = [(0, 1), (1, 1), (1, 0), (2, 1), (4, 1), (4, 0), (8, 0), (1, 2), (4, 2), (0, 2)]


def make_imd(box, pbc, tables, margin, timestep=0.1, ntypes=1, lpa=0, fused=2):
    imd = host.Imd(0)
    imd.ctx.set_tuning(lpa=lpa, fused_table=fused)
    imd.setup_potentials(tables)
    imd.make_box(box, pbc, margin)
    imd.set_params(ntypes, timestep)
    return imd


@pytest.mark.parametrize("interp", ["3point", "4point", "spline"])
def test_table_access_matches_oracle(interp):
    from oracle.binding import Table
    tables = synth.make_eam_tables(synth.CU, interp=interp)
    imd = make_imd(np.diag([30.0, 30, 30]), (1, 1, 1), tables, 0.5)
    rng = np.random.default_rng(0)
    for which, key in ((0, "pair"), (1, "rho"), (2, "emb")):
        pt = tables[key]
        # inside, below begin (is_short clamp), above end (clamped to end), exact knots
        arg = np.concatenate([rng.uniform(pt.begin[0] - 0.5, pt.end[0] + 1.0, 20000),
                              pt.begin[0] + pt.step[0] * np.arange(0, pt.len[0], 37),
                              [pt.begin[0], pt.end[0]]])
        val, grad = imd.ctx.eval_table(which, arg)
        oval, ograd = Table(pt).pair_int(arg)
        # same operations; only FMA contraction may differ -> 1e-13 relative to the table scale
        assert np.max(np.abs(val - oval)) <= 1e-13 * np.max(np.abs(oval))
        assert np.max(np.abs(grad - ograd)) <= 1e-12 * np.max(np.abs(ograd))
    imd.close()


@pytest.mark.parametrize("lpa,fused", VARIANTS)
@pytest.mark.parametrize("case", GOLDEN_CASES + INTERP_CASES)
def test_golden_cells_nblist_forces(case, lpa, fused):
    g = Golden(case)
    for s in g.steps:
        if "s%d_nblpairs" % s not in g.z:
            continue
        imd = make_imd(g.box, g.pbc, g.tables, g.margin, g.timestep, g.ntypes, lpa, fused)
        sc = g.step(s, "scalars")
        assert [int(x) for x in imd.global_cell_dim] == [int(x) for x in sc[4:7]]
        assert imd.cellsz == sc[7]
        imd.set_atoms(g.step(s, "pre_pos"), g.step(s, "pre_mom"), g.step(s, "pre_mass"),
                      g.step(s, "pre_type"), g.step(s, "pre_id"))
        imd.calc_forces(s)
        got = imd.atoms()
        assert np.array_equal(got["id"], g.ids_sorted(s))
        # bit exact: wrapped positions, cell ids, neighbour index set
        assert np.array_equal(got["pos"], g.by_id(s, "post_pos"))
        assert np.array_equal(got["cell"], g.by_id(s, "cell"))
        mine, ref = canon_pairs(imd.nblist_pairs()), canon_pairs(g.step(s, "nblpairs"))
        assert mine.shape == ref.shape and np.array_equal(mine, ref)
        # FP64 within 1e-10 relative
        # (perfect lattice at step 0: the net force is round-off of ~50 pair forces of ~0.3 eV/A each; the
        # cubic weights of the four-point mode cancel more, so its floor is a pair force, not 1e-2)
        floor = 1e-2 if g.interp == "3point" else 0.1
        for key in ("force", "epot", "rho", "dF"):
            assert rel_err(got[key], g.by_id(s, key), floor) < RTOL, (case, s, key)
        assert abs(imd.tot_pot_energy - sc[0]) < RTOL * abs(sc[0])
        assert abs(imd.virial - sc[1]) < 1e-9 * max(1.0, abs(sc[1]))
        imd.close()


@pytest.mark.parametrize("case", ["cu_6x6x6_pbc", "ni3al_6x6x6_pbc", "cu_6x6x6_pbc_4point", "cu_6x6x6_pbc_spline",
                                  "ni3al_6x5x5_free_x_spline"])
@pytest.mark.parametrize("lpa", [0, 1])
def test_golden_trajectory(case, lpa):
    """NVE from the reference's step-0 state: calc_forces -> move_atoms -> check_nblist on the
    device reproduces the reference's dumped states and its rebuild schedule (one and two species:
    the second case runs the MULTI kernels, per-column table grids, two masses)."""
    g = Golden(case)
    imd = make_imd(g.box, g.pbc, g.tables, g.margin, g.timestep, g.ntypes, lpa, 1)
    imd.set_atoms(g.step(0, "pre_pos"), g.step(0, "pre_mom"), g.step(0, "pre_mass"),
                  g.step(0, "pre_type"), g.step(0, "pre_id"))
    # the reference's .eng (write_eng_file, imd_io.c:2290-2420, built with -DHPO: %.16e): one row per step,
    # columns time, Epot = tot_pot_energy / natoms, temperature = 2 tot_kin_energy / nactive, pressure, volume
    eng = g.z["eng"]
    n = g.step(0, "pre_id").shape[0]
    nsteps = max(g.steps) + 1
    assert eng.shape[0] == nsteps
    mine = np.zeros((nsteps, 2))
    for step in range(0, nsteps):
        count0 = imd.nbl_count
        imd.calc_forces(step)
        if step in g.steps:
            got = imd.atoms()
            assert rel_err(got["force"], g.by_id(step, "force")) < 1e-8, step
            assert (imd.nbl_count != count0) == ("s%d_nblpairs" % step in g.z), step
            assert abs(imd.tot_pot_energy - g.step(step, "scalars")[0]) < 1e-9 * abs(g.step(step, "scalars")[0])
        imd.move_atoms()
        mine[step] = imd.tot_pot_energy / n, 2.0 * imd.tot_kin_energy / (3 * n)
    # NVE energy drift equal to the reference's (north_star): Etot per atom = Epot + 3/2 T, step by step.
    # A trajectory is chaotic, so the bound grows with the step; observed <= 2e-12 after 40 steps.
    etot_ref = eng[:, 1] + 1.5 * eng[:, 2]
    etot = mine[:, 0] + 1.5 * mine[:, 1]
    err_pot = np.abs(mine[:, 0] - eng[:, 1]).max() / np.abs(eng[:, 1]).max()
    err_kin = np.abs(mine[:, 1] - eng[:, 2]).max() / np.abs(eng[:, 2]).max()
    drift_ref, drift = etot_ref - etot_ref[0], etot - etot[0]
    err_drift = np.abs(drift - drift_ref).max()
    print("%s lpa=%d: Epot/atom rel err %.2e, T rel err %.2e, drift ref %.3e mine %.3e, |d drift| %.2e eV/atom"
          % (case, lpa, err_pot, err_kin, np.abs(drift_ref).max(), np.abs(drift).max(), err_drift))
    assert err_pot < 1e-9 and err_kin < 1e-8
    assert err_drift < 1e-9 * np.abs(etot_ref[0]), (err_drift, np.abs(drift_ref).max())
    imd.close()


@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_filtered_nblist_equals_exact_build(case):
    """The list build with the single-precision classifier (k_nbl_build2: float r^2 against cellsz -+ a measured error
    margin, the exact test in the shell between) gives the list of the exact-only build (k_nbl_build) -- same pairs in
    the same order, hence bit-identical forces -- on periodic images, free surfaces and two species."""
    g = Golden(case)
    s = g.steps[0]
    out = {}
    for filt in (0, 1):
        imd = make_imd(g.box, g.pbc, g.tables, g.margin, g.timestep, g.ntypes)
        imd.ctx.set_tuning(nbl_filter=filt)
        imd.set_atoms(g.step(s, "pre_pos"), g.step(s, "pre_mom"), g.step(s, "pre_mass"),
                      g.step(s, "pre_type"), g.step(s, "pre_id"))
        imd.calc_forces(s)
        out[filt] = (canon_pairs(imd.nblist_pairs()), imd.atoms(("id", "force", "epot", "rho", "numneighs")))
        imd.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert np.array_equal(out[1][0], canon_pairs(g.step(s, "nblpairs")))             # and the reference's
    for key in ("force", "epot", "rho", "numneighs"):
        assert np.array_equal(out[0][1][key], out[1][1][key]), key


def tables_on_different_grids(spec):
    """Single-species tables whose rho column is tabulated on another r^2 grid than the pair column:
    the interleaved-row kernel (k_eam_pass1_fused) does not apply, the general one runs."""
    pair, rho, emb = synth.eam_columns(spec)
    r2 = np.linspace(1.1, 0.97 * spec.r_c**2, 1777)
    r = np.sqrt(r2)
    f = spec.f_e * np.exp(-spec.beta * (r / spec.r_e - 1.0)) * synth._smooth(r, np.sqrt(0.97) * spec.r_c)
    rho = (1.1, 0.97 * spec.r_c**2, (0.97 * spec.r_c**2 - 1.1) / 1776, f)
    mk = synth._table_from_columns
    return {"pair": mk([pair[0]], [pair[1]], [pair[2]], [pair[3]], True),
            "rho": mk([rho[0]], [rho[1]], [rho[2]], [rho[3]], True),
            "emb": mk([emb[0]], [emb[1]], [emb[2]], [emb[3]], False)}


@pytest.mark.parametrize("ncell,pbc,spec,same_grid", [((9, 7, 8), (1, 1, 1), synth.CU, True),
                                                      ((12, 6, 6), (0, 1, 1), synth.AL, True),
                                                      ((6, 6, 6), (0, 0, 0), synth.CU, True),
                                                      ((8, 7, 6), (1, 1, 1), synth.CU, False),
                                                      ((10, 5, 6), (0, 1, 1), synth.AL, False)])
@pytest.mark.parametrize("lpa", [0, 1, 2])
def test_oracle_parity_seeded(ncell, pbc, spec, same_grid, lpa):
    from oracle.binding import EamOracle
    tables = synth.make_eam_tables(spec) if same_grid else tables_on_different_grids(spec)
    pos, box = synth.fcc_lattice(ncell, spec.a)
    rng = np.random.default_rng(7)
    pos = pos + rng.normal(0, 0.08, pos.shape)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.05, seed=3)
    margin = 0.4
    imd = make_imd(box, pbc, tables, margin, lpa=lpa)
    imd.set_atoms(pos, mom, spec.mass)
    orc = EamOracle(box, pbc, tables, margin)
    p = pos.copy()
    cells = orc.fix_cells(p)
    pairs = orc.make_nblist()
    ref = orc.calc_forces(p, np.zeros(pos.shape[0], np.int32))
    imd.calc_forces(0)
    got = imd.atoms()
    assert np.array_equal(got["pos"], p)
    assert np.array_equal(got["cell"], cells)
    assert np.array_equal(canon_pairs(imd.nblist_pairs()), canon_pairs(pairs))
    assert np.array_equal(got["numneighs"], ref["numneighs"])
    for key in ("force", "epot", "rho", "dF"):
        assert rel_err(got[key], ref[key]) < RTOL, key
    assert abs(imd.tot_pot_energy - ref["tot_pot_energy"]) < RTOL * abs(ref["tot_pot_energy"])
    # integrator + displacement check against the oracle's
    f = ref["force"]
    mass = np.full(pos.shape[0], spec.mass)
    p2, m2 = p.copy(), mom.copy()
    ekin = orc.move_atoms_nve(p2, m2, f, mass, 0.1)
    imd.move_atoms()
    got2 = imd.atoms(("id", "pos", "mom"))
    assert rel_err(got2["pos"], p2) < 1e-14 and rel_err(got2["mom"], m2) < 1e-13
    assert abs(imd.tot_kin_energy - ekin) < 1e-12 * ekin
    imd.close()
    orc.close()


def test_energy_conservation_and_momentum():
    """Size-independent properties on a larger box: Newton's third law (sum F = 0),
    NVE energy drift small, total momentum conserved."""
    spec = synth.CU
    tables = synth.make_eam_tables(spec)
    pos, box = synth.fcc_lattice((20, 20, 20), spec.a)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=5)
    imd = make_imd(box, (1, 1, 1), tables, 0.5)
    imd.set_atoms(pos, mom, spec.mass)
    e = imd.main_loop(60)
    got = imd.atoms(("id", "force", "mom"))
    fsum = np.abs(got["force"].sum(axis=0)).max()
    assert fsum < 1e-9 * np.abs(got["force"]).sum()
    assert np.abs(got["mom"].sum(axis=0)).max() < 1e-8
    etot = e[1:, 0] + e[1:, 1]        # leap-frog: first step has no half-step history
    drift = np.abs(etot - etot[0]).max() / pos.shape[0]
    assert drift < 2e-5, drift
    assert imd.nbl_count >= 1
    imd.close()


def test_full_size_properties_and_kernel_variants():
    """BASELINE.json configs[1] at full size (Al, 80^3 cubic cells = 2 048 000 atoms): the hot instantiations
    (one lane per atom, interleaved-row pass 1, pipelined neighbour loop) against (a) size-independent
    properties -- sum F = 0, momentum and NVE energy conserved -- and (b) the four-lanes-per-atom general
    instantiations that the golden tests exercise at small sizes, on the same resident atoms."""
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    pos, box = synth.fcc_lattice((80, 80, 80), spec.a)
    rng = np.random.default_rng(11)
    pos = pos + rng.normal(0.0, 0.04, pos.shape)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=7)
    imd = make_imd(box, (1, 1, 1), tables, 0.4)
    imd.set_atoms(pos, mom, spec.mass)
    imd.calc_forces(0)
    a = imd.atoms(("id", "force", "epot", "rho", "dF"))
    e_hot = imd.tot_pot_energy
    fscale = np.abs(a["force"]).max()
    assert np.abs(a["force"].sum(axis=0)).max() < 1e-9 * np.abs(a["force"]).sum()
    # k_eam_pass1_rep4 (raw rows, dv / d2v formed in the kernel with the macro's operations) at both CTA sizes:
    # the same operations in the same order as the interleaved-row kernel
    for blk in (896, 768):
        imd.ctx.set_tuning(lpa=0, fused_table=2, rep4_block=blk)
        imd.calc_forces(0)
        r = imd.atoms(("id", "force", "epot", "rho", "dF"))
        same = all(np.array_equal(a[k], r[k]) for k in ("force", "epot", "rho", "dF"))
        print("k_eam_pass1_rep4, %d threads: bit-identical to k_eam_pass1_fused: %s" % (blk, same))
        assert np.abs(a["force"] - r["force"]).max() < 1e-13 * fscale
        for key in ("epot", "rho", "dF"):
            assert rel_err(a[key], r[key]) < 1e-14, (key, blk)
        assert abs(imd.tot_pot_energy - e_hot) < 1e-13 * abs(e_hot)
    # the exact-only list build on the same atoms: same list, order included -> bit-identical results
    imd.ctx.set_tuning(lpa=0, fused_table=2, rep4_block=896, nbl_filter=0)
    imd.ctx.invalidate_nblist()
    imd.calc_forces(0)
    r = imd.atoms(("id", "force", "epot", "rho", "dF"))
    assert all(np.array_equal(a[k], r[k]) for k in ("force", "epot", "rho", "dF")), "the filtered list differs from the exact build"
    imd.ctx.set_tuning(nbl_filter=1)
    imd.ctx.invalidate_nblist()
    imd.ctx.set_tuning(lpa=4, fused_table=0)
    imd.calc_forces(0)
    b = imd.atoms(("id", "force", "epot", "rho", "dF"))
    e_lpa = imd.tot_pot_energy
    imd.ctx.set_tuning(lpa=0, fused_table=2)
    assert np.abs(a["force"] - b["force"]).max() < 1e-11 * fscale
    for key in ("epot", "rho", "dF"):
        assert rel_err(a[key], b[key]) < 1e-12, key
    assert abs(e_hot - e_lpa) < 1e-12 * abs(e_hot)
    e = imd.main_loop(25)
    got = imd.atoms(("id", "mom"))
    assert np.abs(got["mom"].sum(axis=0) - mom.sum(axis=0)).max() < 1e-6
    etot = e[1:, 0] + e[1:, 1]
    assert np.abs(etot - etot[0]).max() / pos.shape[0] < 2e-5
    assert imd.nbl_count >= 1
    imd.close()


@pytest.mark.parametrize("ncell,pbc", [((64, 64, 64), (1, 1, 1)), ((100, 48, 48), (0, 1, 1))])
def test_full_size_hot_kernels_vs_oracle(ncell, pbc):
    """The instantiations bench.py times -- one lane per atom, k_eam_pass1_fused<1> + k_eam_pass2<1,...>, chosen by
    system size (n >= 148 x 2048) -- against the CPU restatement of make_nblist / calc_forces
    (imd_forces_nbl.c:136-287, 419-985) on the WHOLE system: ~1 M Al atoms, bulk and film."""
    from oracle.binding import EamOracle
    spec = synth.AL
    tables = synth.make_eam_tables(spec)
    pos, box = synth.fcc_lattice(ncell, spec.a)
    if not pbc[0]:
        pos[:, 0] += 5 * spec.a
        box[0, 0] += 10 * spec.a
    rng = np.random.default_rng(21)
    pos = pos + rng.normal(0.0, 0.06, pos.shape)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=9)
    assert pos.shape[0] >= 148 * 2048
    imd = make_imd(box, pbc, tables, 0.4)
    imd.set_atoms(pos, mom, spec.mass)
    imd.calc_forces(0)
    got = imd.atoms(("id", "pos", "cell", "force", "epot", "rho", "dF", "numneighs"))
    orc = EamOracle(box, pbc, tables, 0.4)
    p = pos.copy()
    cells = orc.fix_cells(p)
    orc.make_nblist()
    ref = orc.calc_forces(p, np.zeros(pos.shape[0], np.int32))
    assert np.array_equal(got["pos"], p)
    assert np.array_equal(got["cell"], cells)
    assert np.array_equal(got["numneighs"], ref["numneighs"])          # every pair inside the cutoff, both sides
    errs = {k: rel_err(got[k], ref[k]) for k in ("force", "epot", "rho", "dF")}
    e_rel = abs(imd.tot_pot_energy - ref["tot_pot_energy"]) / abs(ref["tot_pot_energy"])
    print("full size %s: %d atoms, rel err %s, Epot %.2e" % (ncell, pos.shape[0], errs, e_rel))
    for k, v in errs.items():
        assert v < RTOL, (k, v)
    assert e_rel < RTOL
    imd.close()
    orc.close()
