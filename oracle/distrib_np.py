"""TEST INFRASTRUCTURE (oracle), not product code: numpy restatement of the reference's spatial distributions.

Follows /root/reference/imd_distrib.c:
  make_distrib_density        :377-411   number of atoms per bin (float)
  make_write_distrib_select   :678-716   per-bin sum of a per-atom quantity (float), :738-746 division by the count
  dist_Ekin_fun :219-222  p.p / (2 m)      dist_Epot_fun :230-238  POTENG      dist_vxavg_fun :364-367  p_x / m
  write_distrib_density       :434-460   density = count * dist_size / volume (float)

Bins are orthogonal boxes: num = (int)(scale * (x - ll)) per direction with C truncation toward zero -- an atom up to
one bin width BELOW ll lands in bin 0, as in the reference -- rejected when < 0 or >= dim; the flat index is
(numx * dim.y + numy) * dim.z + numz.  The reference accumulates in float, atom by atom in its cell order; `sums`
accumulates in double (what the device does) and `reference_files` converts to what the reference writes, so the
comparison with its files is to float rounding of a sum of the bin's atoms.

Pinned against files the unmodified reference wrote (tests/golden/distrib_*.npz, made by
tests/golden/make_golden_distrib.py): tests/test_distrib_oracle_vs_reference.py.
"""
import numpy as np


def bin_index(pos, dim, ll, ur):
    """Flat bin of every atom, -1 when outside (imd_distrib.c:385-408)."""
    dim = np.asarray(dim, np.int64)
    ll, ur = np.asarray(ll, np.float64), np.asarray(ur, np.float64)
    scale = dim / (ur - ll)
    ok = np.ones(pos.shape[0], bool)
    num = np.zeros(pos.shape[0], np.int64)
    for d in range(3):
        nd = np.trunc(scale[d] * (pos[:, d] - ll[d])).astype(np.int64)      # (int) conversion: toward zero
        ok &= (nd >= 0) & (nd < dim[d])
        num = num * dim[d] + np.where(ok, nd, 0)
    return np.where(ok, num, -1)


def sums(pos, mom, mass, epot, dim, ll, ur):
    """Per-bin sums in double: count, Ekin, Epot, v_x (the device's output before its conversion to float)."""
    size = int(np.prod(dim))
    b = bin_index(pos, dim, ll, ur)
    m = b >= 0
    mass = np.broadcast_to(np.asarray(mass, np.float64), (pos.shape[0],))
    ekin = (mom[:, 0] * mom[:, 0] + mom[:, 1] * mom[:, 1] + mom[:, 2] * mom[:, 2]) / (2 * mass)
    out = {}
    for key, val in (("num", np.ones(pos.shape[0])), ("ekin", ekin), ("epot", epot), ("vxavg", mom[:, 0] / mass)):
        out[key] = np.bincount(b[m], weights=np.asarray(val, np.float64)[m], minlength=size)
    return out


def reference_files(s, dim, ll, ur):
    """What the reference writes from the sums: per-atom averages of bins that hold atoms (others keep the sum, 0),
    and the number density (imd_distrib.c:738-746, 447-455), as float32."""
    num = s["num"].astype(np.float32)
    out = {}
    for key in ("ekin", "epot", "vxavg"):
        d = s[key].astype(np.float32)
        out[key] = np.where(num > 0, d / np.where(num > 0, num, 1), d).astype(np.float32)
    vol = float(np.prod(np.asarray(ur, np.float64) - np.asarray(ll, np.float64)))
    out["dens"] = (num * np.float32(num.shape[0] / vol)).astype(np.float32)
    return out


def sums_float_sequential(pos, mom, mass, epot, dim, ll, ur):
    """The reference's own arithmetic: float accumulators, `*dat += <double expression>` atom by atom in the given
    order (small cases only: a Python loop)."""
    size = int(np.prod(dim))
    b = bin_index(pos, dim, ll, ur)
    mass = np.broadcast_to(np.asarray(mass, np.float64), (pos.shape[0],))
    acc = {k: np.zeros(size, np.float32) for k in ("num", "ekin", "epot", "vxavg")}
    for i in range(pos.shape[0]):
        k = b[i]
        if k < 0:
            continue
        acc["num"][k] = np.float32(np.float64(acc["num"][k]) + 1.0)
        e = (mom[i, 0] * mom[i, 0] + mom[i, 1] * mom[i, 1] + mom[i, 2] * mom[i, 2]) / (2 * mass[i])
        acc["ekin"][k] = np.float32(np.float64(acc["ekin"][k]) + e)
        acc["epot"][k] = np.float32(np.float64(acc["epot"][k]) + epot[i])
        acc["vxavg"][k] = np.float32(np.float64(acc["vxavg"][k]) + mom[i, 0] / mass[i])
    return acc
