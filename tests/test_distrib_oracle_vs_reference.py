"""The numpy restatement of the reference's spatial distributions (oracle/distrib_np.py) against files the
unmodified reference wrote (tests/golden/distrib_*.npz, tests/golden/make_golden_distrib.py): pins the oracle."""
import os

import numpy as np
import pytest

from oracle import distrib_np as dn

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["distrib_cu_7x6x6_box", "distrib_cu_8x6x6_sub1d"]


@pytest.mark.parametrize("case", CASES)
def test_oracle_equals_reference_files(case):
    z = np.load(os.path.join(GOLD, case + ".npz"))
    dim, ll, ur = z["dim"], z["ll"], z["ur"]
    for n in (1, 2):
        pos, mom, mass, epot = z["n%d_pos" % n], z["n%d_mom" % n], z["n%d_mass" % n], z["n%d_epot" % n]
        s = dn.sums(pos, mom, mass, epot, dim, ll, ur)
        f = dn.reference_files(s, dim, ll, ur)
        inside = int(s["num"].sum())
        assert 0 < inside <= pos.shape[0]
        # the count is a sum of ones: exact in float, so the density file pins the binning bit for bit
        assert np.array_equal(f["dens"], z["n%d_file_dens" % n]), case
        # averages: the reference sums in float in its cell order, the oracle in double -> float rounding of a sum
        per_bin = max(1.0, s["num"].max())
        tol = 2.0 ** -24 * per_bin
        for key, fk in (("ekin", "Ekin"), ("epot", "Epot")):
            ref = z["n%d_file_%s" % (n, fk)]
            err = np.abs(f[key] - ref).max() / np.abs(ref).max()
            assert err <= tol, (case, n, key, err, tol)
        if case.endswith("sub1d"):
            assert inside < pos.shape[0]                        # the sub-box rejects atoms
        # the reference's own arithmetic (float accumulators, atom by atom) lands on its files within the
        # reordering of the atoms between its cells and the checkpoint's order
        q = dn.reference_files({k: v.astype(np.float64) for k, v in
                                dn.sums_float_sequential(pos, mom, mass, epot, dim, ll, ur).items()}, dim, ll, ur)
        assert np.array_equal(q["dens"], z["n%d_file_dens" % n])
        assert np.abs(q["ekin"] - z["n%d_file_Ekin" % n]).max() <= tol * np.abs(z["n%d_file_Ekin" % n]).max()


def test_truncation_toward_zero_puts_atoms_just_below_ll_into_bin_zero():
    pos = np.array([[-0.4, 1.0, 1.0], [-1.2, 1.0, 1.0], [9.99, 1.0, 1.0], [10.0, 1.0, 1.0]])
    b = dn.bin_index(pos, (10, 1, 1), (0.0, 0.0, 0.0), (10.0, 2.0, 2.0))
    assert list(b) == [0, -1, 9, -1]                             # (int)(-0.4) = 0 as in C; -1.2 -> -1 rejected
