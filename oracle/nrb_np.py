"""oracle/nrb_np.py -- TEST INFRASTRUCTURE: numpy restatement of the reference's nrb_forces
(imd_nrb.c:792-1129) and of what its send_forces does to boundary momenta while nrb_forces is not yet
called (steps <= 1; imd_comm_force_3d.c:1046-1052,1157-1168).  Pinned against dumps of the unmodified
reference (oracle/_ref/imd_ref_mpi_eam_ttm_nrb, tests/golden/ttm_nrb_*.npz) in
tests/test_nrb_cpu.py.  Not restated: nrb_test_forces (imd_forces_nbl.c:1297-1345), a left-over test
that adds drand48() forces to atoms around x = 1100 A for steps < 300.
"""
import numpy as np

MASS = 26.9815                    # imd_nrb.c:805 (hard-coded for the neighbour terms)
SQRT2HALF = 0.70710678118655      # globals.h:426
DBLSQRT2 = 2.82842712474619       # globals.h:427


def _nint(x):
    """makros.h:103: (int)(x < 0 ? x - 0.5 : x + 0.5)"""
    return np.trunc(np.where(x < 0, x - 0.5, x + 0.5))


def _minimg(u, box, pbc_dirs):
    """MINIMGZ always, MINIMGY if pbc_dirs.y (:849-852, :926-928); each is a no-op without pbc."""
    u = u.copy()
    for ax in (2, 1):
        if pbc_dirs[ax] == 1:
            u[..., ax] = u[..., ax] - box[ax] * _nint(u[..., ax] * (1.0 / box[ax]))
    return u


def nrb_forces(ids, pos, mom, mass, bnd, nrbi, refpos, box, pbc_dirs, nrb_k, in_list=None):
    """Momenta after nrb_forces.  box = (Lx, Ly, Lz); nrb_k = the parameter (init_nrb takes
    sqrt(nrb_k / 26.9815), :149).  in_list(id_i, id_j) -> bool tells whether the pair is in the
    neighbour list (None: every listed site is, true for nearest neighbours a/sqrt(2) apart)."""
    k = np.sqrt(nrb_k / MASS)
    index = {int(a): i for i, a in enumerate(ids)}
    out = mom.copy()
    for i in np.where(bnd > 0)[0]:
        u = _minimg(pos[i] - refpos[i], box, pbc_dirs)
        if bnd[i] == 1:
            ud = np.array([-k * 4.0 * u[0], -k * DBLSQRT2 * u[1], -k * DBLSQRT2 * u[2]])
        else:
            ud = np.array([-k * DBLSQRT2 * u[0], -k * u[1] * 4, -k * DBLSQRT2 * u[2]])
        p = np.zeros(3) + ud * mass[i]
        for r in range(12):
            q = int(nrbi[i, r])
            if q == 0 or q not in index:
                continue
            if in_list is not None and not in_list(int(ids[i]), q):
                continue
            j = index[q]
            U = _minimg(pos[j] - refpos[j], box, pbc_dirs)
            V = mom[j] / MASS                                  # IMPULS_ALT: momenta of before the call
            if bnd[i] == 1:
                d = np.array([k * U[0], k * SQRT2HALF * U[1], k * SQRT2HALF * U[2]])
            else:
                d = np.array([k * SQRT2HALF * U[0], k * U[1], k * SQRT2HALF * U[2]])
            d = d - 0.25 * V
            p = p + d * MASS
        out[i] = p
    return out


def startup_images(pos, mom, bnd, box, pbc_dirs, global_cell_dim):
    """steps <= 1: every boundary atom's momentum times the number of periodic y/z images of its cell
    (cpu_dim = (P,1,1): a cell in the first or last y/z layer has one buffer-cell copy per such layer)."""
    out = mom.copy()
    mult = np.ones(pos.shape[0])
    for ax in (1, 2):
        if pbc_dirs[ax] == 1:
            g = int(global_cell_dim[ax])
            c = np.floor(pos[:, ax] / (box[ax] / g)).astype(int)
            mult = mult * (1 + (c == 0) + (c == g - 1))
    m = bnd > 0
    out[m] = mom[m] * mult[m, None]
    return out
