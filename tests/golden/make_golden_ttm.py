#!/usr/bin/env python
"""Generate the TTM golden fixtures (tests/golden/ttm_*.npz) by RUNNING THE UNMODIFIED
REFERENCE's MPI + TTM build over shmpi/.

Needs oracle/_ref/imd_ref_mpi_eam_ttm and oracle/_ref/nn_probe (``make -C oracle ref``).
The binary is the reference compiled with -DMPI -DNBL -DEAM2 -DTTM -DTTM1D -DTMM -DPDECAY,
linked with oracle/ref_hooks.c (calc_forces dumps) and oracle/ref_hooks_ttm.c (analytic
source instead of the data-file-dependent transfer-matrix solver; calc_ttm dumps).
nn_probe runs the reference's interpolation library (libnn.a) on the same EOS tables to
record the diagonal choice of its Delaunay triangulation and probe values.

    python tests/golden/make_golden_ttm.py
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

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm")
REF_BIN_NRB = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_nrb")   # + -DNRB -DREFPOS, imd_nrb.c
REF_BIN_TMM = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_tmm")   # do_tmm / tmm_init NOT replaced
NN_PROBE = os.path.join(ROOT, "oracle", "_ref", "nn_probe")

CASES = {
    # name: (box_param, nranks, ttm params, steps, dump_steps, shiftx_front, shiftx_rear, starttemp, extra)
    "ttm_al_56x6x6_p2": ((56, 6, 6), 2, synth.TtmParams(ttmdimx=16, fd_n_timesteps=20), 12,
                          (0, 1, 2, 5, 12), 20.0, 20.0, 0.0258, ""),
    "ttm_al_56x6x6_p4": ((56, 6, 6), 4, synth.TtmParams(ttmdimx=16, fd_n_timesteps=20, I0=2.0e16), 8,
                          (0, 1, 4, 8), 20.0, 20.0, 0.0258, ""),
    # hot lattice + strong pulse: atoms cross FD slabs (advection fluxes), rear absorber on
    "ttm_al_48x5x5_p2_hot": ((48, 5, 5), 2, synth.TtmParams(ttmdimx=12, fd_n_timesteps=30, I0=4.0e16,
                                                           fd_min_atoms=20, vlatbuffer=1), 30,
                              (0, 1, 10, 20, 30), 25.0, 15.0, 0.12,
                              "xipdecay 0.5\nramp_start 180.0\nramp_end 200.0\n"),
    # non-reflecting rear boundary (imd_nrb.c): warm lattice + pulse, so that the boundary plane and
    # its four neighbours per atom move; steps 0 and 1 run without nrb_forces (steps > 1 only)
    "ttm_nrb_al_40x5x5_p2": ((40, 5, 5), 2, synth.TtmParams(ttmdimx=10, fd_n_timesteps=20, I0=2.0e16), 24,
                              (0, 1, 2, 3, 12, 24), 20.0, 20.0, 0.05,
                              "nrb_alat 4.05\nnrb_eps 0.3\nnrb_k 1.2\n"),
    # the reference's own transfer-matrix source (imd_tmm.c) on synthetic alu_eps_bb.dat / K12.dat
    "ttm_tmm_al_96x5x5_p2": ((96, 5, 5), 2, synth.TtmParams(ttmdimx=26, fd_n_timesteps=20, I0=2.0e16, tmm_threshold=5.0,
                                                           use_tmm=1), 14,
                              (0, 1, 2, 3, 7, 14), 20.0, 20.0, 0.0258, ""),
    "ttm_tmm_al_96x5x5_p4": ((96, 5, 5), 4, synth.TtmParams(ttmdimx=28, fd_n_timesteps=25, I0=6.0e16, tmm_threshold=6.0,
                                                           use_tmm=1, laser_t_0=20e-15), 12,
                              (0, 2, 4, 8, 12), 20.0, 20.0, 0.05, ""),
    "ttm_nrb_al_56x5x6_p4": ((56, 5, 6), 4, synth.TtmParams(ttmdimx=16, fd_n_timesteps=20, I0=3.0e16), 20,
                              (0, 2, 3, 10, 20), 20.0, 20.0, 0.08,
                              "nrb_alat 4.05\nnrb_eps 0.3\nnrb_k 0.9\n"),
}


def rd(d, step, name, rank, dtype, cols=None):
    a = np.fromfile(os.path.join(d, "s%d.%s.r%d.bin" % (step, name, rank)), dtype=dtype)
    return a.reshape(-1, cols) if cols else a


def load_atoms(d, step, nranks):
    out = {}
    cat = lambda name, dt, cols=None: np.concatenate([rd(d, step, name, r, dt, cols) for r in range(nranks)])  # noqa: E731
    for tag in ("post",):                      # calc_forces does not change positions/momenta
        out[tag + "_id"] = cat(tag + "_id", np.int32)
        out[tag + "_pos"] = cat(tag + "_pos", np.float64, 3)
        out[tag + "_mom"] = cat(tag + "_mom", np.float64, 3)
    out["type"] = cat("post_type", np.int32)
    out["mass"] = cat("post_mass", np.float64)
    out["force"] = cat("post_force", np.float64, 3)
    out["numneighs"] = cat("post_numneighs", np.int32)
    out["scalars"] = rd(d, step, "post_scalars", 0, np.float64)
    nrb = os.path.exists(os.path.join(d, "s%d.post_nrbbnd.r0.bin" % step))
    if nrb:
        out["pre_id"] = cat("pre_id", np.int32)
        out["pre_mom"] = cat("pre_mom", np.float64, 3)[np.argsort(out["pre_id"], kind="stable")]
        del out["pre_id"]
        out["nrbbnd"] = cat("post_nrbbnd", np.int32)
        out["nrbneigh"] = cat("post_nrbneigh", np.int32)
        out["nrbi"] = cat("post_nrbi", np.int32, 12)
        out["refpos"] = cat("post_refpos", np.float64, 3)
    # sort everything by atom id so that the fixture does not depend on the decomposition
    for tag in ("post",):
        o = np.argsort(out[tag + "_id"], kind="stable")
        for k in ("_id", "_pos", "_mom"):
            out[tag + k] = out[tag + k][o]
        if tag == "post":
            for k in ("type", "mass", "force", "numneighs") + (("nrbbnd", "nrbneigh", "nrbi", "refpos") if nrb else ()):
                out[k] = out[k][o]
    return out


def load_ttm(d, step, nranks):
    cells, hdrs = [], []
    extra = None
    for r in range(nranks):
        a = np.fromfile(os.path.join(d, "s%d.ttm.r%d.bin" % (step, r)), dtype=np.float64)
        hdr = a[:16]
        G, L, vlatdim = int(hdr[0]), int(hdr[1]), int(hdr[15])
        p = 16
        cells.append(a[p:p + 13 * L].reshape(L, 13)); p += 13 * L
        glob = a[p:p + 7 * G].reshape(7, G); p += 7 * G
        vlat = a[p:p + 6 * vlatdim].reshape(vlatdim, 6)
        hdrs.append(hdr)
        if r == int(hdr[7]) or extra is None:          # the virtual lattice of its owner rank
            if r == int(hdr[7]):
                extra = (glob, vlat)
            elif extra is None:
                extra = (glob, vlat)
    hdrs = np.array(hdrs)
    glob, vlat = extra
    return {"cells": np.concatenate(cells), "hdr": hdrs, "xiarr": glob[0], "vcom": glob[1:4], "natomsglobal": glob[4],
            "fluxl": glob[5], "fluxr": glob[6], "vlat": vlat,
            "Eabs": np.array(hdrs[:, 8].sum()), "tot_elec_energy": np.array(hdrs[:, 9].sum())}


def probe_eos(tab, path, rng, nprobe=400):
    """diag flags + lpi_interpolate_point values from the reference's own library."""
    x = rng.uniform(tab.x[0], tab.x[-1], nprobe)
    y = rng.uniform(tab.y[0], tab.y[-1], nprobe)
    # a few points exactly on grid lines and nodes
    x[:20] = tab.x[rng.integers(0, tab.x.size, 20)]
    y[10:30] = tab.y[rng.integers(0, tab.y.size, 20)]
    pr = np.stack([x, y], axis=1)
    pr.tofile(path + ".probes")
    subprocess.check_call([NN_PROBE, path, path + ".probes", path + ".out"])
    raw = open(path + ".out", "rb").read()
    hdr = np.frombuffer(raw[:16], dtype=np.int32)
    nx, ny, npr = int(hdr[0]), int(hdr[1]), int(hdr[2])
    nd = (nx - 1) * (ny - 1)
    diag = np.frombuffer(raw[16:16 + nd], dtype=np.uint8).reshape(nx - 1, ny - 1).copy()
    off = 16 + nd + (8 - nd % 8) % 8
    z = np.frombuffer(raw[off:off + 8 * npr], dtype=np.float64).copy()
    assert diag.max() <= 1, "triangulation is not a split of the grid rectangles"
    return diag, pr, z


def run_case(name, box_param, nranks, tp, steps, dump_steps, sfront, srear, starttemp, extra):
    spec = synth.AL
    work = tempfile.mkdtemp(prefix="imdgold_ttm_")
    try:
        synth.make_eam_tables(spec, work)
        eos = synth.make_eos_tables()
        synth.write_eos_files(eos, work)
        if tp.use_tmm:
            synth.write_tmm_files(synth.make_tmm_tables(), work)
        rng = np.random.default_rng(7)
        arrays = {}
        for key, fn in (("cve", "EOS_cve_from_r_te.txt"), ("ee", "EOS_ee_from_r_tesqrt.txt")):
            diag, pr, z = probe_eos(eos[key], os.path.join(work, fn), rng)
            arrays["eos_%s_diag" % key] = diag
            arrays["eos_%s_probe_xy" % key] = pr
            arrays["eos_%s_probe_z" % key] = z
        run = os.path.join(work, "w")
        os.makedirs(os.path.join(run, "out"))
        os.makedirs(os.path.join(run, "dump"))
        ex = tp.param_text() + "shiftx_front %.17g\nshiftx_rear %.17g\n" % (sfront, srear) + extra
        synth.write_param_file(os.path.join(run, "run.param"), spec, box_param, steps, "out/run", potdir="..",
                               pbc=(0, 1, 1), starttemp=starttemp, nbl_margin=0.4, eng_int=1, extra=ex,
                               ensemble="ttm", cpu_dim=(nranks, 1, 1))
        env = dict(os.environ, SHMPI_NP=str(nranks), IMD_DUMP_DIR="dump",
                   IMD_DUMP_STEPS=",".join(str(s) for s in dump_steps),
                   IMD_SRC_MU="%.17g" % tp.src_mu, IMD_SRC_ABS="%.17g" % tp.src_abs)
        res = subprocess.run([REF_BIN_TMM if tp.use_tmm else REF_BIN_NRB if "nrb_alat" in extra else REF_BIN, "-p", "run.param"], cwd=run, env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, timeout=600)
        if res.returncode != 0:
            print(res.stdout[-3000:])
            raise RuntimeError("reference TTM run failed for %s" % name)
        for s in dump_steps:
            for k, v in load_atoms(os.path.join(run, "dump"), s, nranks).items():
                if k in ("nrbbnd", "nrbneigh", "nrbi", "refpos"):      # static after init_nrb: kept once
                    if "nrb_" + k in arrays:
                        assert np.array_equal(arrays["nrb_" + k], v), "NRB state changed at step %d" % s
                    arrays["nrb_" + k] = v
                    continue
                arrays["s%d_%s" % (s, k)] = v
            for k, v in load_ttm(os.path.join(run, "dump"), s, nranks).items():
                arrays["s%d_ttm_%s" % (s, k)] = v
        arrays["eng"] = np.loadtxt(os.path.join(run, "out", "run.eng"), comments="#")
        sc = arrays["s%d_scalars" % dump_steps[0]]
        arrays["box"] = np.array([sc[8], sc[9], sc[10]])
        arrays["box_param"] = np.array(box_param, dtype=np.int32)
        arrays["nranks"] = np.array(nranks)
        arrays["dump_steps"] = np.array(dump_steps, dtype=np.int32)
        arrays["nsteps"] = np.array(steps)
        arrays["timestep"] = np.array(0.1)
        arrays["nbl_margin"] = np.array(0.4)
        arrays["starttemp"] = np.array(starttemp)
        arrays["extra"] = np.array(extra)
        for k, v in tp.__dict__.items():
            arrays["tp_" + k] = np.array(v)
        out = os.path.join(ROOT, "tests", "golden", name + ".npz")
        np.savez_compressed(out, **arrays)
        print("%-24s atoms %6d ranks %d -> %s (%.0f kB)" % (name, arrays["s0_post_id"].shape[0], nranks, out,
                                                          os.path.getsize(out) / 1e3))
        print("    " + [ln for ln in res.stdout.splitlines() if ln.startswith("step:")][-1])
    finally:
        shutil.rmtree(work, ignore_errors=True)


def main():
    for b in (REF_BIN, NN_PROBE):
        if not os.path.exists(b):
            sys.exit("missing %s: run `make -C oracle ref` where /root/reference exists" % b)
    only = sys.argv[1:]
    for name, args in CASES.items():
        if only and name not in only:
            continue
        run_case(name, *args)


if __name__ == "__main__":
    main()
