#!/usr/bin/env python
"""Generate the golden fixtures under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE.

Needs oracle/_ref/imd_ref_eam_nve_nbl (``make -C oracle ref``; only possible where
/root/reference exists).  The binary is the reference's serial EAM2/NBL/NVE build linked
with oracle/ref_hooks.c, which dumps the inputs and outputs of the reference's own
calc_forces (imd_forces_nbl.c:295) at chosen steps.  The fixtures are committed, so the
tests and the GPU box never need the reference.

    python tests/golden/make_golden.py            # regenerates all *.npz
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from imd_b200 import synth  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl")

CASES = {
    # name: (spec, box_param, pbc, starttemp, steps, dump_steps, nbl_margin)
    "cu_6x6x6_pbc": (synth.CU, (6, 6, 6), (1, 1, 1), 0.10, 40, (0, 1, 20, 40), 0.5),
    "cu_8x6x7_pbc": (synth.CU, (8, 6, 7), (1, 1, 1), 0.05, 12, (0, 12), 0.5),
    "cu_7x6x6_free_x": (synth.CU, (7, 6, 6), (0, 1, 1), 0.05, 12, (0, 12), 0.5),
    "al_5x5x5_pbc": (synth.AL, (5, 5, 5), (1, 1, 1), 0.05, 10, (0, 10), 0.4),
    # two species (L1_2), every table column on its own r^2 grid, asymmetric rho columns
    "ni3al_6x6x6_pbc": (synth.NI3AL, (6, 6, 6), (1, 1, 1), 0.08, 30, (0, 1, 15, 30), 0.5),
    "ni3al_7x5x6_free_x": (synth.NI3AL, (7, 5, 6), (0, 1, 1), 0.05, 12, (0, 12), 0.4),
    # the other two table-interpolation modes of potaccess.h (reference built with -DFOURPOINT / -DSPLINE)
    "cu_6x6x6_pbc_4point": (synth.CU, (6, 6, 6), (1, 1, 1), 0.10, 20, (0, 1, 20), 0.5),
    "cu_6x6x6_pbc_spline": (synth.CU, (6, 6, 6), (1, 1, 1), 0.10, 20, (0, 1, 20), 0.5),
    "ni3al_6x5x5_free_x_4point": (synth.NI3AL, (6, 5, 5), (0, 1, 1), 0.08, 12, (0, 12), 0.5),
    "ni3al_6x5x5_free_x_spline": (synth.NI3AL, (6, 5, 5), (0, 1, 1), 0.08, 12, (0, 12), 0.5),
}


def interp_of(name):
    return "4point" if name.endswith("_4point") else "spline" if name.endswith("_spline") else "3point"


def load_dump(d, step):
    def rd(name, dtype, cols=None):
        a = np.fromfile(os.path.join(d, "s%d.%s.bin" % (step, name)), dtype=dtype)
        return a.reshape(-1, cols) if cols else a
    out = {}
    for tag in ("pre", "post"):
        out[tag + "_id"] = rd(tag + "_id", np.int32)
        out[tag + "_pos"] = rd(tag + "_pos", np.float64, 3)
        out[tag + "_mom"] = rd(tag + "_mom", np.float64, 3)
        out[tag + "_type"] = rd(tag + "_type", np.int32)
        out[tag + "_mass"] = rd(tag + "_mass", np.float64)
    out["force"] = rd("post_force", np.float64, 3)
    out["epot"] = rd("post_epot", np.float64)
    out["rho"] = rd("post_rho", np.float64)
    out["dF"] = rd("post_dF", np.float64)
    out["cell"] = rd("post_cell", np.int32, 3)
    out["scalars"] = rd("post_scalars", np.float64)
    p = os.path.join(d, "s%d.post_nblpairs.bin" % step)
    if os.path.exists(p):
        out["nblpairs"] = np.fromfile(p, dtype=np.int32).reshape(-1, 2)
    return out


def run_case(name, spec, box_param, pbc, starttemp, steps, dump_steps, margin, keep=None):
    work = tempfile.mkdtemp(prefix="imdgold_")
    try:
        binary = isinstance(spec, synth.BinarySpec)
        if binary:
            synth.make_eam_tables_binary(spec, work)
        else:
            synth.make_eam_tables(spec, work)
        os.makedirs(os.path.join(work, "out"))
        os.makedirs(os.path.join(work, "dump"))
        synth.write_param_file(os.path.join(work, "run.param"), spec, box_param, steps, "out/run",
                               pbc=pbc, starttemp=starttemp, nbl_margin=margin, eng_int=1,
                               coordname="_l12" if binary else "_fcc", masses=spec.masses if binary else None)
        env = dict(os.environ, IMD_DUMP_DIR=os.path.join(work, "dump"),
                   IMD_DUMP_STEPS=",".join(str(s) for s in dump_steps))
        interp = interp_of(name)
        ref_bin = REF_BIN if interp == "3point" else REF_BIN + "_" + interp
        res = subprocess.run([ref_bin, "-p", "run.param"], cwd=work, env=env,
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if res.returncode != 0:
            print(res.stdout[-3000:])
            raise RuntimeError("reference run failed for %s" % name)
        arrays = {}
        for s in dump_steps:
            for k, v in load_dump(os.path.join(work, "dump"), s).items():
                arrays["s%d_%s" % (s, k)] = v
        # .eng: time, Epot/atom, Ekin-ish columns (imd_io.c:2290+); kept verbatim as numbers
        eng = np.loadtxt(os.path.join(work, "out", "run.eng"), comments="#")
        arrays["eng"] = eng
        arrays["box_param"] = np.array(box_param, dtype=np.int32)
        arrays["pbc"] = np.array(pbc, dtype=np.int32)
        arrays["dump_steps"] = np.array(dump_steps, dtype=np.int32)
        arrays["nbl_margin"] = np.array(margin)
        arrays["timestep"] = np.array(0.1)
        arrays["spec_name"] = np.array(spec.name)
        arrays["interp"] = np.array(interp)
        out = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(out, **arrays)
        print("%-18s atoms %6d  -> %s (%.0f kB)" % (name, arrays["s0_pre_id"].shape[0], out,
                                                   os.path.getsize(out) / 1e3))
        if keep:
            shutil.copytree(work, keep, dirs_exist_ok=True)
    finally:
        shutil.rmtree(work, ignore_errors=True)


def main():
    if not os.path.exists(REF_BIN):
        sys.exit("missing %s: run `make -C oracle ref` where /root/reference exists" % REF_BIN)
    only = sys.argv[1:]
    for name, args in CASES.items():
        if not only or name in only:
            run_case(name, *args)


if __name__ == "__main__":
    main()
