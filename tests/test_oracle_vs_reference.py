"""Pins the CPU restatement (oracle/) against the UNMODIFIED reference: the fixtures in
tests/golden were dumped from the reference's own calc_forces (see make_golden.py).
Integer results (cell ids, neighbour index sets) must agree bit for bit; FP64 results to
1e-10 relative (summation order differs only where the reference's intra-cell atom order
is history dependent, imd_alloc.c:111-128)."""
import numpy as np
import pytest

from oracle.binding import EamOracle
from tests.util import GOLDEN_CASES, INTERP_CASES, Golden, canon_pairs

RTOL = 1e-10


def rel_err(a, b):
    scale = np.max(np.abs(b)) + 1e-300
    return np.max(np.abs(a - b)) / scale


@pytest.mark.parametrize("case", GOLDEN_CASES + INTERP_CASES)
def test_cells_nblist_forces(case):
    g = Golden(case)
    for s in g.steps:
        if "s%d_nblpairs" % s not in g.z:
            continue      # only steps where the reference rebuilt its list fix the cell state
        ids = g.step(s, "pre_id")
        pos = np.ascontiguousarray(g.step(s, "pre_pos"))
        orc = EamOracle(g.box, g.pbc, g.tables, g.margin, ntypes=g.ntypes)
        st = orc.state
        assert [st.gdim[0], st.gdim[1], st.gdim[2]] == [int(x) for x in g.step(s, "scalars")[4:7]]
        assert st.cellsz == g.step(s, "scalars")[7]
        cells = orc.fix_cells(pos)
        # wrapped positions and cell ids: bit exact (same atom order pre/post in the dump?)
        post_ids = g.step(s, "post_id")
        o_pre, o_post = np.argsort(ids), np.argsort(post_ids)
        assert np.array_equal(pos[o_pre], g.step(s, "post_pos")[o_post])
        assert np.array_equal(cells[o_pre], g.step(s, "cell")[o_post])
        pairs = orc.make_nblist()
        mine = canon_pairs(ids[pairs])
        ref = canon_pairs(g.step(s, "nblpairs"))
        assert mine.shape == ref.shape and np.array_equal(mine, ref)
        out = orc.calc_forces(pos, g.step(s, "pre_type"))
        for key in ("force", "epot", "rho", "dF"):
            assert rel_err(out[key][o_pre], g.step(s, key)[o_post]) < RTOL, key
        assert abs(out["tot_pot_energy"] - g.step(s, "scalars")[0]) < RTOL * abs(g.step(s, "scalars")[0])
        assert abs(out["virial"] - g.step(s, "scalars")[1]) < 1e-9 * max(1.0, abs(g.step(s, "scalars")[1]))
        orc.close()


def test_trajectory_matches_reference():
    """Oracle-driven NVE (calc_forces -> move_atoms_nve -> check_nblist) from the reference's
    step-0 state reproduces the reference's state at its later dump steps."""
    g = Golden("cu_6x6x6_pbc")
    ids = g.step(0, "pre_id")
    pos = np.ascontiguousarray(g.step(0, "pre_pos")).copy()
    mom = np.ascontiguousarray(g.step(0, "pre_mom")).copy()
    mass, types = g.step(0, "pre_mass"), g.step(0, "pre_type")
    orc = EamOracle(g.box, g.pbc, g.tables, g.margin)
    valid, ref_pos = False, None
    last = max(g.steps)
    for step in range(0, last + 1):
        if step in g.steps:
            o_me, o_ref = np.argsort(ids), np.argsort(g.step(step, "pre_id"))
            # positions may differ by a box vector between rebuilds only after wrapping;
            # pre_* is taken before calc_forces wraps, so they must agree directly
            assert rel_err(pos[o_me], g.step(step, "pre_pos")[o_ref]) < 1e-9, step
            assert rel_err(mom[o_me], g.step(step, "pre_mom")[o_ref]) < 1e-8, step
        if not valid:
            orc.fix_cells(pos)
            orc.make_nblist()
            ref_pos = pos.copy()
            valid = True
        out = orc.calc_forces(pos, types)
        if step in g.steps:
            o_me, o_ref = np.argsort(ids), np.argsort(g.step(step, "post_id"))
            assert rel_err(out["force"][o_me], g.step(step, "force")[o_ref]) < 1e-8, step
            rebuilt_ref = "s%d_nblpairs" % step in g.z
            assert rebuilt_ref == np.array_equal(ref_pos, pos), step
        orc.move_atoms_nve(pos, mom, out["force"], mass, g.timestep)
        if orc.max_displacement2(pos, ref_pos) > (0.5 * g.margin) ** 2:
            valid = False
    orc.close()


@pytest.mark.parametrize("interp", ["3point", "4point", "spline"])
def test_table_finish_host_mirror_equals_oracle(interp):
    """imd_b200/synth.py finishes the tables the way read_pot_table does for the mode (init_threepoint /
    init_fourpoint / init_spline, imd_potential.c:1167-1276); the C restatement does the same from the raw
    columns.  Both must give the same bits (extra rows and, for splines, the second-derivative table)."""
    import ctypes as C
    from oracle import binding
    from imd_b200 import synth
    L = binding.lib()
    for tables in (synth.make_eam_tables(synth.CU, interp=interp), synth.make_eam_tables_binary(synth.NI3AL, interp=interp)):
        for key, radial in (("pair", 1), ("rho", 1), ("emb", 0)):
            pt = tables[key]
            t = binding.Table(pt)
            raw = np.array(pt.table, copy=True)
            for c in range(pt.ncols):
                raw[pt.len[c]:, c] = 0.0                 # what read_pot_table2 leaves behind the data
            t._tab[...] = raw
            if interp == "spline":
                t._tab2[...] = 0.0
                L.orc_init_spline(C.byref(t.c), radial)
                assert np.array_equal(t._tab2, pt.table2)
            elif interp == "4point":
                L.orc_init_fourpoint(C.byref(t.c))
            else:
                L.orc_init_threepoint(C.byref(t.c))
            assert np.array_equal(t._tab, pt.table), (interp, key)
