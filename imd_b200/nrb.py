"""imd_b200/nrb.py -- host side of the non-reflecting boundary: the one-time setup the reference does in
init_nrb (imd_nrb.c:125-768), mirrored in numpy so that the Python host can switch NRB on without IMD.
The per-step part (nrb_forces, the integrator hooks) is device code: imd_b200/csrc/nrb.cu.

init_nrb in words: reference positions = current positions (:203-216); boundary region = the outermost
atomic layer, x > max(x) - alat/4 -> type 1, else y < min(y) + alat/4 -> type 2, else
y > max(y) - alat/4 -> type 3 (:222-268; with periodic y only type 1 exists, :226-230); for every
boundary atom the atoms that sit at the fcc (100) nearest-neighbour offsets listed in SITES, each
component within nrb_eps (:327-700), are recorded by atom number in NRBI[site].
"""
from __future__ import annotations

import numpy as np

# site -> (offset in units of alat/2 relative to the boundary atom, boundary types that use it)
SITES = {
    0: ((-1, +1, 0), (1, 2)),      # imd_nrb.c:327-343
    1: ((-1, -1, 0), (1, 3)),      # :346-362
    2: ((+1, -1, 0), (3,)),        # :364-379
    3: ((+1, +1, 0), (2,)),        # :382-397
    4: ((0, -1, +1), (3,)),        # :440-455
    5: ((0, -1, -1), (3,)),        # :457-471
    6: ((0, +1, -1), (2,)),        # :486-500
    7: ((0, +1, +1), (2,)),        # :473-484
    8: ((-1, 0, +1), (1,)),        # :421-436
    9: ((-1, 0, -1), (1,)),        # :403-418
}


class NrbState:
    """What init_nrb leaves per atom: bnd (NRBBND), neigh (NRBNEIGH), nrbi (NRBI, 0 = unset), refpos."""

    def __init__(self, ids, bnd, neigh, nrbi, refpos):
        self.ids, self.bnd, self.neigh, self.nrbi, self.refpos = ids, bnd, neigh, nrbi, refpos

    def entries(self):
        """Rows imdcu_set_nrb wants: the boundary atoms and every atom one of them lists."""
        listed = np.unique(self.nrbi[self.bnd > 0])
        listed = listed[listed != 0]
        m = (self.bnd > 0) | np.isin(self.ids, listed)
        return self.ids[m], self.bnd[m], self.nrbi[m], self.refpos[m]


def boundary_types(pos, pbc_dirs, nrb_alat):
    """NRBBND per atom (imd_nrb.c:203-268)."""
    xhi = pos[:, 0].max() - nrb_alat / 4
    yhi = pos[:, 1].max() - nrb_alat / 4
    ylo = pos[:, 1].min() + nrb_alat / 4
    if pbc_dirs[1] == 1:
        yhi, ylo = 10e9, -10e9
    bnd = np.zeros(pos.shape[0], np.int32)
    bnd[pos[:, 1] > yhi] = 3
    bnd[pos[:, 1] < ylo] = 2
    bnd[pos[:, 0] > xhi] = 1                         # x has priority, then -y, then +y
    return bnd


def init_nrb(pos, ids, box, pbc_dirs, nrb_alat, nrb_eps):
    """NrbState for atoms at `pos` (numbers `ids`, none of them 0) in the orthorhombic `box`."""
    from scipy.spatial import cKDTree
    pos = np.asarray(pos, np.float64)
    ids = np.asarray(ids, np.int32)
    if (ids == 0).any():
        raise ValueError("atom number 0 marks an unset NRB site in the reference; number the atoms from 1")
    L = np.diag(np.asarray(box, np.float64).reshape(3, 3))
    n = pos.shape[0]
    nd = nrb_alat / 2.0
    bnd = boundary_types(pos, pbc_dirs, nrb_alat)
    nrbi = np.zeros((n, 12), np.int32)
    neigh = np.zeros(n, np.int32)
    tree = cKDTree(pos)
    # images of a target across the periodic directions (the reference sees them as buffer-cell copies)
    shifts = [np.zeros(3)]
    for ax in (0, 1, 2):
        if pbc_dirs[ax] == 1:
            shifts = [s + k * np.eye(3)[ax] * L[ax] for s in shifts for k in (-1, 0, 1)]
    for site, (off, types) in SITES.items():
        who = np.where(np.isin(bnd, types))[0]
        if who.size == 0:
            continue
        target = pos[who] + nd * np.asarray(off, np.float64)
        for s in shifts:
            hits = tree.query_ball_point(target + s, r=nrb_eps, p=np.inf)
            for w, h in zip(who, hits):
                for q in h:
                    if q != w:
                        nrbi[w, site] = ids[q]
                        neigh[q] = 1
    return NrbState(ids, bnd, neigh, nrbi, pos.copy())
