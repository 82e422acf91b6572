"""Shared helpers of the 3-D TTM sweep tests."""
import glob
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "ttm3d_*.npz")))


def load_golden(path):
    z = np.load(path)
    case = {k[5:]: z[k] for k in z.files if k.startswith("case_")}
    case = dict(shape=tuple(int(v) for v in case["shape"]), seed=int(case["seed"]), hole=bool(case["hole"]),
                pbc=tuple(int(v) for v in case["pbc"]), fd_min_atoms=int(case["fd_min_atoms"]),
                h=tuple(float(v) for v in case["h"]), tau=float(case["tau"]), diff_substeps=int(case["diff_substeps"]),
                laser_active=int(case["laser_active"]), nsweeps=int(case["nsweeps"]))
    out = dict(eabs=float(z["eabs"]), natoms=z["natoms"], temp=z["temp"], U=z["U"], xi=z["xi"])
    return case, out


def faces_and_interior(shape_ext):
    """Mask of the cells the path defines: interior + the six ghost faces (edges and corners of the
    ghost layer are never written by ttm_fill_ghost_layers nor read by do_DIFF)."""
    m = np.zeros(shape_ext, bool)
    m[1:-1, 1:-1, 1:-1] = True
    m[0, 1:-1, 1:-1] = m[-1, 1:-1, 1:-1] = True
    m[1:-1, 0, 1:-1] = m[1:-1, -1, 1:-1] = True
    m[1:-1, 1:-1, 0] = m[1:-1, 1:-1, -1] = True
    return m
