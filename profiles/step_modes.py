#!/usr/bin/env python
"""Ways of driving one MD step of the bench film through the C ABI, ms per step (100 steps each after 20 warm-up):
  run_ttm            one library call for all steps (what bench.py's `value` times)
  fused + move       imdcu_calc_forces_ttm, imdcu_move_atoms_ttm per step (what integration/imd_cuda.c does)
  three calls        imdcu_calc_forces, imdcu_calc_ttm, imdcu_move_atoms_ttm per step
    python profiles/step_modes.py [ncx ncy ncz]"""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from imd_b200 import host, synth

cells = tuple(int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (100, 100, 100)
spec, pos, box, ids = bench.build_system(cells, 0, 1, True)
imd = host.Imd(0)
imd.setup_potentials(synth.make_eam_tables(spec))
imd.make_box(box, (0, 1, 1), 0.4)
imd.set_params(1, 0.1)
imd.set_atoms(pos, synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=1234), spec.mass, None, ids)
_, _, gx = bench.film_geometry(cells, 1)
imd.init_ttm(bench.bench_ttm_params(gx, 250, False), synth.make_eos_tables(), 1)
c = imd.ctx
step = 0
def timed(fn, n):
    global step
    c.synchronize(); t0 = time.perf_counter(); r0 = c.nbl_count
    for _ in range(n):
        fn(step); step += 1
    c.synchronize()
    return 1e3 * (time.perf_counter() - t0) / n, c.nbl_count - r0
c.run_ttm(step, 20, False); step += 20
t0 = time.perf_counter(); r0 = c.nbl_count; c.run_ttm(step, 100, False); c.synchronize(); step += 100
print("run_ttm (one call)        %.3f ms per step, %d rebuilds" % (1e3 * (time.perf_counter() - t0) / 100, c.nbl_count - r0))
def fused(s):
    c.calc_forces_ttm(s); c.move_atoms_ttm()
def three(s):
    c.calc_forces(s); c.calc_ttm(s); c.move_atoms_ttm()
print("fused + move per step     %.3f ms per step, %d rebuilds" % timed(fused, 100))
print("three calls per step      %.3f ms per step, %d rebuilds" % timed(three, 100))
print("fused + move per step     %.3f ms per step, %d rebuilds" % timed(fused, 100))
