#!/usr/bin/env python
"""Golden fixtures for the spatial distributions (imd_distrib.c), made by RUNNING THE UNMODIFIED REFERENCE.

oracle/_ref/imd_ref_eam_nve_nbl runs with dist_int = checkpt_int, so that for a step both the reference's own
distribution files (<outfiles>.<n>.Ekin / .Epot / .dens, binary float32, write_distrib imd_distrib.c:33-218) and
its checkpoint of the same state (write_config, same block of the main loop, imd_main_3d.c:700-722) exist.  The
fixture keeps the checkpoint's atoms (number, mass, position, velocity, Epot) and the files' arrays, and -- from the
dump hooks of oracle/ref_hooks.c -- the atoms as they ENTER calc_forces of that step: POTENG in the files belongs to
those positions, the checkpoint's positions are the ones after move_atoms, so a device run reproduces the reference's
sequence calc_forces -> move_atoms -> write_distrib from the entering state.

    python tests/golden/make_golden_distrib.py        # needs `make -C oracle ref` (where /root/reference exists)
"""
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from imd_b200 import synth  # noqa: E402
from tests.golden.make_golden import load_dump  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl")

CASES = {
    # name: (box_param, pbc, starttemp, dist_dim, dist_ll, dist_ur)   dist_ur None: the reference's default, the box
    # free x surfaces: atoms drift below x = 0 (truncation toward zero puts them into bin 0) and beyond box_x
    "distrib_cu_7x6x6_box": ((7, 6, 6), (0, 1, 1), 0.30, (6, 4, 3), None, None),
    # 1-D slabs along x on a sub-box: atoms outside in every direction are rejected
    "distrib_cu_8x6x6_sub1d": ((8, 6, 6), (1, 1, 1), 0.10, (9, 1, 1), (3.0, 2.0, 1.5), (25.0, 19.0, 20.5)),
}


def read_dist(path):
    raw = open(path, "rb").read()
    i = raw.index(b"#E\n") + 3
    hdr = raw[:i].decode()
    assert hdr.startswith("#F L"), hdr                         # little endian, binary
    return np.frombuffer(raw[i:], dtype="<f4").copy()


def run_case(name, box_param, pbc, starttemp, dim, ll, ur):
    work = tempfile.mkdtemp(prefix="imdgold_dist_")
    try:
        synth.make_eam_tables(synth.CU, work)
        os.makedirs(os.path.join(work, "out"))
        ex = "checkpt_int 10\ndist_int 10\ndist_dim %d %d %d\ndist_Ekin_flag 1\ndist_Epot_flag 1\ndist_dens_flag 1\n" % dim
        if ll is not None:
            ex += "dist_ll %.17g %.17g %.17g\ndist_ur %.17g %.17g %.17g\n" % (tuple(ll) + tuple(ur))
        synth.write_param_file(os.path.join(work, "run.param"), synth.CU, box_param, 20, "out/run", pbc=pbc,
                               starttemp=starttemp, nbl_margin=0.5, eng_int=1, extra=ex)
        os.makedirs(os.path.join(work, "dump"))
        env = dict(os.environ, IMD_DUMP_DIR=os.path.join(work, "dump"), IMD_DUMP_STEPS="10,20")
        res = subprocess.run([REF_BIN, "-p", "run.param"], cwd=work, env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True)
        if res.returncode != 0:
            print(res.stdout[-3000:])
            raise RuntimeError("reference run failed for %s" % name)
        box = np.array(box_param, np.float64) * synth.CU.a
        arrays = {"dim": np.array(dim, np.int32), "ll": np.array(ll if ll is not None else (0.0, 0.0, 0.0)),
                  "ur": np.array(ur if ur is not None else box), "box": box, "pbc": np.array(pbc, np.int32),
                  "box_param": np.array(box_param, np.int32)}
        for n in (1, 2):                                       # steps 10 and 20
            rows = [ln.split() for ln in open(os.path.join(work, "out", "run.%05d.chkpt" % n)) if not ln.startswith("#")]
            a = np.array(rows, dtype=np.float64)               # number type mass x y z vx vy vz Epot eam_rho
            arrays["n%d_id" % n] = a[:, 0].astype(np.int32)
            arrays["n%d_mass" % n] = a[:, 2]
            arrays["n%d_pos" % n] = a[:, 3:6]
            arrays["n%d_mom" % n] = a[:, 6:9] * a[:, 2:3]
            arrays["n%d_epot" % n] = a[:, 9]
            d = load_dump(os.path.join(work, "dump"), 10 * n)      # state entering calc_forces(10 n)
            for key in ("pre_id", "pre_pos", "pre_mom", "pre_mass", "pre_type"):
                arrays["n%d_%s" % (n, key)] = d[key]
            for key in ("Ekin", "Epot", "dens"):
                arrays["n%d_file_%s" % (n, key)] = read_dist(os.path.join(work, "out", "run.%d.%s" % (n, key)))
        out = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(out, **arrays)
        print("%-26s atoms %5d bins %s -> %s (%.0f kB)" % (name, arrays["n1_id"].shape[0], dim, out,
                                                           os.path.getsize(out) / 1e3))
    finally:
        shutil.rmtree(work, ignore_errors=True)


def main():
    if not os.path.exists(REF_BIN):
        sys.exit("missing %s: run `make -C oracle ref` where /root/reference exists" % REF_BIN)
    for name, args in CASES.items():
        if len(sys.argv) < 2 or name in sys.argv[1:]:
            run_case(name, *args)


if __name__ == "__main__":
    main()
