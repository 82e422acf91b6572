#!/usr/bin/env python
"""Generates tests/golden/ttm3d_*.npz: outputs of the UNMODIFIED reference functions do_DIFF +
ttm_fill_ghost_layers (imd_ttm_3D.c) run through oracle/_ref/ref3d_diff (oracle/ref3d_driver.c)
on seeded synthetic lattices.  Needs `make -C oracle ref` (i.e. /root/reference); the fixtures are
committed so that the tests run anywhere.  Inputs are regenerated from the seed by
oracle.ttm3d_np.make_case, the fixture stores the case parameters and the reference's outputs.

    python tests/golden/make_golden_ttm3d.py
"""
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ttm3d_np as t3  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref3d_diff")

CASES = {
    "ttm3d_film_12x10x8": dict(shape=(12, 10, 8), seed=3, hole=True, pbc=(0, 1, 1), fd_min_atoms=20,
                               h=(17.0, 16.0, 15.0), tau=0.1 / 250, diff_substeps=250, laser_active=1, nsweeps=7),
    "ttm3d_bulk_9x16x11": dict(shape=(9, 16, 11), seed=5, hole=True, pbc=(1, 1, 1), fd_min_atoms=25,
                               h=(12.0, 13.0, 14.0), tau=0.1 / 400, diff_substeps=400, laser_active=0, nsweeps=4),
    "ttm3d_ragged_35x19x37": dict(shape=(35, 19, 37), seed=7, hole=True, pbc=(0, 1, 1), fd_min_atoms=20,
                                  h=(16.0, 16.0, 16.0), tau=0.1 / 250, diff_substeps=250, laser_active=1, nsweeps=3),
}


def run_reference(f, case, exe=REF):
    """f: dict of extended-lattice fields -> (Eabs, natoms, temp, U, xi) of the reference."""
    d = tempfile.mkdtemp(prefix="ref3d_")
    nx, ny, nz = case["shape"]
    with open(os.path.join(d, "in.bin"), "wb") as o:
        np.array([nx, ny, nz, case["fd_min_atoms"], case["diff_substeps"], case["laser_active"], case["nsweeps"],
                  case["pbc"][0], case["pbc"][1], case["pbc"][2], 0, 0], np.int32).tofile(o)
        np.array(list(case["h"]) + [case["tau"]], np.float64).tofile(o)
        for k in t3.FIELDS:
            np.ascontiguousarray(f[k], dtype=np.float64).tofile(o)
    r = subprocess.run([exe, os.path.join(d, "in.bin"), os.path.join(d, "out.bin")],
                       env=dict(os.environ, SHMPI_NP="1", SHMPI_RING_MB="64"), capture_output=True, text=True, timeout=300)
    if r.returncode != 0:
        raise RuntimeError("ref3d_diff failed: %s %s" % (r.stdout, r.stderr))
    raw = np.fromfile(os.path.join(d, "out.bin"))
    arr = raw[1:].reshape(4, *f["temp"].shape)
    return float(raw[0]), arr[0].astype(np.int32), arr[1], arr[2], arr[3]


def main():
    if not os.path.exists(REF):
        raise SystemExit("build oracle/_ref/ref3d_diff first (make -C oracle ref)")
    for name, case in CASES.items():
        f = t3.make_case(case["shape"], seed=case["seed"], hole=case["hole"], pbc=case["pbc"],
                         fd_min_atoms=case["fd_min_atoms"])
        eabs, nat, temp, U, xi = run_reference(f, case)
        out = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(out, eabs=eabs, natoms=nat, temp=temp, U=U, xi=xi,
                            **{"case_" + k: np.asarray(v) for k, v in case.items()})
        print(name, "Eabs", eabs, "->", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
