"""Time calc_ttm's lattice kernel alone (k_ttm_grid: one CTA, a latency chain) on a small Al film with the
bench's TTM parameters (FD cell ~16 A, 250 sub-steps, vlatdim 30).  Its duration does not depend on the
atom count, so a 100x8x8-cell film (25 600 atoms) stands for the 4 M-atom bench workload.
usage: python profiles/ttm_grid_probe.py [ncx] [substeps] [reps]   -> one JSON line"""
import json
import sys
import time

import numpy as np

from imd_b200 import host, synth

ncx = int(sys.argv[1]) if len(sys.argv) > 1 else 100
sub = int(sys.argv[2]) if len(sys.argv) > 2 else 250
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 40
spec = synth.AL
pos, box = synth.fcc_lattice((ncx, 8, 8), spec.a)
pos[:, 0] += 10 * spec.a
box[0, 0] += 20 * spec.a
imd = host.Imd(0)
imd.setup_potentials(synth.make_eam_tables(spec))
imd.make_box(box, (0, 1, 1), 0.4)
imd.set_params(1, 0.1)
mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=7)
imd.set_atoms(pos, mom, spec.mass)
gx = int(box[0, 0] / 16.0)
tp = synth.TtmParams(ttmdimx=gx, fd_n_timesteps=sub, fd_min_atoms=100, vlatdim=30, vlatbuffer=8, I0=5.0e15,
                     laser_t_0=100e-15, laser_sigma_t=42.5e-15)
imd.init_ttm(tp, synth.make_eos_tables(), 1)
ctx = imd.ctx
ctx.calc_forces(0)
ctx.calc_ttm(0)
for s in range(1, 4):
    ctx.calc_ttm(s, False)
ctx.synchronize()
t0 = time.perf_counter()
for s in range(4, 4 + reps):
    ctx.calc_ttm(s, False)
ctx.synchronize()
ms = (time.perf_counter() - t0) / reps * 1e3
sc = ctx.calc_ttm(4 + reps)
lat = imd.ttm_lattice()
te = lat["fields"]["temp"]
print(json.dumps({"G": gx, "substeps": int(sc["diff_substeps"]), "ms_per_calc_ttm": round(ms, 4),
                  "te_sum": float(np.sum(te)), "te_max": float(np.max(te)), "xi_sum": float(np.sum(lat["fields"]["xi"]))}))
