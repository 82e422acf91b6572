"""Device-side spatial distributions (imdcu_distrib, csrc/distrib.cu) through the C ABI against the numpy oracle
(oracle/distrib_np.py, pinned on the reference's files) and against those files themselves (write_distrib,
imd_distrib.c:33-218; fixtures tests/golden/distrib_*.npz written by the unmodified reference)."""
import os

import numpy as np
import pytest

from imd_b200 import host, synth
from oracle import distrib_np as dn

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["distrib_cu_7x6x6_box", "distrib_cu_8x6x6_sub1d"]


def make_imd(box, pbc, n_types=1):
    imd = host.Imd(0)
    imd.setup_potentials(synth.make_eam_tables(synth.CU))
    imd.make_box(np.diag(box), tuple(int(x) for x in pbc), 0.5)
    imd.set_params(n_types, 0.1)
    return imd


@pytest.mark.parametrize("case", CASES)
def test_device_distrib_equals_oracle_and_reference_files(case):
    z = np.load(os.path.join(GOLD, case + ".npz"))
    dim, ll, ur = z["dim"], z["ll"], z["ur"]
    for n in (1, 2):
        # the reference's sequence of that step from the state entering it (imd_main_3d.c:406,564,719):
        # calc_forces -> move_atoms -> write_distrib, so POTENG belongs to the positions before the move
        imd = make_imd(z["box"], z["pbc"])
        imd.set_atoms(z["n%d_pre_pos" % n], z["n%d_pre_mom" % n], z["n%d_pre_mass" % n], z["n%d_pre_type" % n],
                      z["n%d_pre_id" % n])
        imd.calc_forces(10 * n)
        imd.move_atoms()
        got = imd.ctx.distrib(dim, ll, ur)
        a = imd.atoms(fields=("id", "pos", "mom", "epot", "mass"))   # the device's own positions, momenta, Epot
        s = dn.sums(a["pos"], a["mom"], a["mass"], a["epot"], dim, ll, ur)
        assert np.array_equal(got["num"], s["num"].astype(np.float32)), case          # binning: exact
        scale = {k: max(np.abs(s[k]).max(), 1e-300) for k in s}
        for k in ("ekin", "epot", "vxavg"):                      # double sums, one float rounding at the end
            err = np.abs(got[k] - s[k]).max() / scale[k]
            assert err <= 2.0 ** -23, (case, n, k, err)
        # against the files the reference wrote for this state
        f = dn.reference_files({k: v.astype(np.float64) for k, v in got.items()}, dim, ll, ur)
        assert np.array_equal(f["dens"], z["n%d_file_dens" % n]), case
        tol = 2.0 ** -23 * max(1.0, s["num"].max())              # the reference sums a bin's atoms in float
        for k, fk in (("ekin", "Ekin"), ("epot", "Epot")):
            ref = z["n%d_file_%s" % (n, fk)]
            err = np.abs(f[k] - ref).max() / np.abs(ref).max()
            print("%s file %d %s: device vs reference file, max err / max %.2e (tolerance %.1e)" % (case, n, fk, err, tol))
            assert err <= tol, (case, n, k, err)
        imd.close()


def test_full_size_distrib_conserves_the_global_sums():
    """1 M atoms, 1-D slabs (the dist_dim of a film run, shared-memory histogram) and a 3-D grid too large for it:
    the bins add up to the number of atoms, Ekin and Epot of the whole system; both code paths agree bin by bin
    after summing the 3-D grid over y and z."""
    spec = synth.CU
    pos, box = synth.fcc_lattice((64, 64, 64), spec.a)
    rng = np.random.default_rng(3)
    pos = pos + rng.normal(0, 0.05, pos.shape)
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.05)
    imd = make_imd(np.diag(box), (1, 1, 1))
    imd.set_atoms(pos, mom, spec.mass)
    imd.calc_forces(0)
    n = pos.shape[0]
    ll, ur = (-1.0, -1.0, -1.0), tuple(np.diag(box) + 1.0)      # a margin: the noise moves atoms across the faces
    d1 = imd.ctx.distrib((40, 1, 1), ll, ur)
    d3 = imd.ctx.distrib((40, 24, 20), ll, ur)
    ekin = float(np.sum(mom * mom) / (2 * spec.mass))
    for d in (d1, d3):
        assert d["num"].astype(np.float64).sum() == n
        assert abs(d["ekin"].astype(np.float64).sum() - ekin) <= 1e-6 * ekin
        assert abs(d["epot"].astype(np.float64).sum() - imd.tot_pot_energy) <= 1e-6 * abs(imd.tot_pot_energy)
    for k in ("num", "ekin", "epot", "vxavg"):
        a3 = d3[k].astype(np.float64).reshape(40, 24 * 20).sum(axis=1)
        sc = max(np.abs(d1[k]).max(), 1e-30)
        assert np.abs(a3 - d1[k]).max() <= 1e-5 * sc, k
    imd.close()


def test_error_returns_of_the_round_2_entry_points():
    """Bad arguments and wrong call order come back as IMDCU_EINVAL with a message, nothing exits or hangs:
    imdcu_distrib, imdcu_comm_set_cuts, imdcu_set_tuning, imdcu_calc_forces_ttm."""
    from imd_b200 import capi
    imd = host.Imd(0)
    imd.setup_potentials(synth.make_eam_tables(synth.CU))
    with pytest.raises(capi.ImdcuError) as e:                    # cuts before the box exists
        imd.ctx.comm_set_cuts([0, 4])
    assert "imdcu_set_box first" in str(e.value)
    imd.make_box(np.diag([30.0, 30.0, 30.0]), (1, 1, 1), 0.5)
    imd.set_params(1, 0.1)
    g = int(imd.global_cell_dim[0])
    imd.ctx.comm_set_cuts([0, g])                                # one rank: the only valid cut array
    with pytest.raises(capi.ImdcuError) as e:
        imd.ctx.comm_set_cuts([0, g - 1])
    assert "xcut[nranks] = global_cell_dim.x" in str(e.value)
    pos, _ = synth.fcc_lattice((8, 8, 8), synth.CU.a)
    imd.set_atoms(pos, synth.maxwell_momenta(pos.shape[0], synth.CU.mass, 0.05), synth.CU.mass)
    with pytest.raises(capi.ImdcuError) as e:                    # and not once the atoms are there
        imd.ctx.comm_set_cuts([0, g])
    assert "already uploaded" in str(e.value)
    imd.calc_forces(0)
    with pytest.raises(capi.ImdcuError) as e:                    # empty or inverted histogram box
        imd.ctx.distrib((4, 0, 1), (0, 0, 0), (30, 30, 30))
    assert "dist_dim >= 1" in str(e.value)
    with pytest.raises(capi.ImdcuError):
        imd.ctx.distrib((4, 1, 1), (0, 0, 0), (30, 0, 30))
    for key, val in (("fused_table", 3), ("rep4_block", 512), ("lpa", 3), ("no_such_knob", 1)):
        with pytest.raises(capi.ImdcuError):
            imd.ctx.set_tuning(**{key: val})
    with pytest.raises(capi.ImdcuError) as e:                    # the fused TTM call without a TTM
        imd.ctx.calc_forces_ttm(0)
    assert "imdcu_ttm_init" in str(e.value)
    d = imd.ctx.distrib((2, 2, 2), (0, 0, 0), (30, 30, 30))     # still usable after the rejected calls
    assert d["num"].sum() == pos.shape[0]
    imd.close()
