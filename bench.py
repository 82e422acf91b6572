#!/usr/bin/env python
"""bench.py -- EAM(+TTM) atom-steps/s of the B200-native hot path (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W            # our arm
    python bench.py --impl reference --gpus N --steps K ...   # the reference's CPU code

A "step" is one MD step of the main-loop segment calc_forces -> calc_ttm -> move_atoms_ttm ->
check_nblist (imd_main_3d.c:406,534,564,812) over the whole synthetic system, neighbour-list
rebuilds included whenever check_nblist asks for them.  Default workload (the one BASELINE.json's
metric is quoted on): N=1 is configs[2], "Al slab 4M atoms + TTM coupling on 1 B200" (film of
100^3 cubic cells = 4 000 000 atoms, 1-D FD lattice along the laser axis, 250 sub-steps, pulse source);
N>1 is configs[3]: the film grows with N (4 M atoms per GPU, 32 M on 8), x slabs over NCCL, non-reflecting
rear boundary.  --workload nve gives configs[1] (Al 80^3 = 2 048 000 atoms NVE), --workload ttm3d configs[4].

value : whole-job atom-steps/s with atoms resident in HBM (device-timed, CUDA events).
e2e   : the same step driven through the C ABI with HOST buffers: every step uploads
        positions+momenta from pinned host memory (imdcu_update_atoms) and downloads the
        new positions+momenta (the next step's inputs) and the energy sums.
roofline : dominant kernel (k_eam_pass1_rep4), algorithmic bytes (64 + 4*N_h per atom,
        SURVEY.md 8d) / its mean launch time measured live with CUDA events.
cpu_baseline : the UNMODIFIED reference (oracle/_ref, author's flags), its own MPI build run
        over shmpi/ with one rank per host core (cpu_dim P 1 1), a bounded sample of the
        same workload; falls back to P concurrent serial runs if the MPI run fails.
N > 1  : weak scaling -- every GPU owns an x slab of --cells cubic cells (cpu_dim = N 1 1), halo
        planes over NCCL; --ttm adds calc_ttm + move_atoms_ttm (BASELINE.json configs[2]).
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "EAM+TTM atom-steps/s"          # --workload ttm (default); --workload nve reports "EAM NVE atom-steps/s"
UNIT = "atom-steps/s"


def metric_name(args):
    return METRIC if args.ttm else "EAM NVE atom-steps/s"
N_H_AL = 39                 # half-list neighbours per atom, Al r_c 6.3 + margin 0.4
REF_FAST = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl_fast")
REF_IEEE = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl")
REF_MPI = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_nve_nbl_fast")
REF_TTM = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_fast")          # -DMPI -DTTM -DTTM1D -DTMM -DPDECAY
REF_TTM_NRB = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_nrb_fast")  # + -DNRB -DREFPOS
VAC_CELLS = 10              # vacuum in front of and behind the film, in lattice constants
FD_CELL = 16.0              # Angstrom: target width of an FD cell (ttmdimx = box_x / FD_CELL, rounded to the rank count)


def measured_traffic(kernel, units):
    """DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json),
    scaled to this run's unit count; None when no capture exists."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]
        return float(t["bytes_per_unit"]) * units, t["source"]
    except Exception:
        return None, None


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ----------------------------------------------------------------------------- clocks

def nvml_handle(local_rank):
    """NVML handle of the GPU torch calls `local_rank` (by UUID: CUDA_VISIBLE_DEVICES may renumber), or None."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(local_rank).uuid)
        for cand in (uuid, "GPU-" + uuid):
            try:
                return pynvml.nvmlDeviceGetHandleByUUID(cand.encode() if hasattr(cand, "encode") else cand)
            except Exception:
                continue
        return pynvml.nvmlDeviceGetHandleByIndex(local_rank)
    except Exception:
        return None


class ClockSampler:
    """SM clock and throttle reasons while the GPU is under load: NVML queries from a thread every 10 ms (the timed
    region of the default run lasts ~140 ms); one-shot nvidia-smi queries every 0.2 s where NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NVML_REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index, period=0.2):
        self.idx, self.period = gpu_index, period
        self.lines, self.stop_flag, self.t = [], threading.Event(), None
        self.h = nvml_handle(gpu_index)

    def _poll_nvml(self):
        import pynvml
        mx = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        while not self.stop_flag.is_set():
            sm = pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            f = ["Active" if bits & b else "Not Active" for b, _ in self.NVML_REASONS]
            self.lines.append("%d, %d, %s" % (sm, mx, ", ".join(f)))
            self.stop_flag.wait(0.01)

    def _poll(self):
        if self.h is not None:
            try:
                self._poll_nvml()
                return
            except Exception:
                pass                                 # fall back to nvidia-smi below
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5)
                self.lines.extend(out.stdout.strip().splitlines())
            except Exception:
                return
            self.stop_flag.wait(self.period)

    def start(self):
        self.t = threading.Thread(target=self._poll, daemon=True)
        self.t.start()

    def stop(self):
        self.stop_flag.set()
        if self.t:
            self.t.join(timeout=6)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------- reference arm

def run_reference_sample(box_cells, steps, nproc, use_fast=True):
    """P concurrent runs of the UNMODIFIED serial reference binary (the reference's only
    parallel mode is MPI, which this image does not have): each process runs the same Al
    NVE sample; throughput = sum over processes of atoms*steps / its main-loop time."""
    from imd_b200 import synth
    binary = REF_FAST if (use_fast and os.path.exists(REF_FAST)) else REF_IEEE
    if not os.path.exists(binary):
        return None
    spec = synth.AL
    work = tempfile.mkdtemp(prefix="imdbench_")
    try:
        synth.make_eam_tables(spec, work)
        procs = []
        for p in range(nproc):
            d = os.path.join(work, "p%d" % p)
            os.makedirs(os.path.join(d, "out"))
            synth.write_param_file(os.path.join(d, "run.param"), spec, box_cells, steps, "out/run",
                                   potdir="..", pbc=(1, 1, 1), starttemp=0.0258, nbl_margin=0.4,
                                   eng_int=0, seed=4711 + p)
            procs.append(subprocess.Popen([binary, "-p", "run.param"], cwd=d, stdout=subprocess.PIPE,
                                          stderr=subprocess.STDOUT, text=True))
        t0 = time.perf_counter()
        outs = [p.communicate()[0] for p in procs]
        wall = time.perf_counter() - t0
        natoms = 4 * box_cells[0] * box_cells[1] * box_cells[2]
        rate, tmain = 0.0, []
        for o, p in zip(outs, procs):
            m = re.search(r"([0-9.eE+-]+) seconds excluding setup time", o)
            if p.returncode != 0 or not m:
                raise RuntimeError("reference run failed:\n" + o[-2000:])
            tm = float(m.group(1))
            tmain.append(tm)
            rate += natoms * (steps + 1) / tm          # imd.c:388-392 counts steps_max+1 steps
        return {"value": rate, "natoms": natoms, "steps": steps + 1, "nproc": nproc,
                "main_loop_s": float(np.max(tmain)), "wall_s": wall,
                "binary": os.path.basename(binary)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def run_reference_mpi(global_cells, steps, cpu_dim):
    """The reference's own MPI build (cpu_dim = x slabs like the GPU decomposition, or IMD's 3-D process grid) over
    shmpi/; throughput from the reference's own main-loop timer (imd.c:387-392, MPI_Wtime)."""
    from imd_b200 import synth
    if not os.path.exists(REF_MPI):
        return None
    spec = synth.AL
    nproc = cpu_dim[0] * cpu_dim[1] * cpu_dim[2]
    work = tempfile.mkdtemp(prefix="imdbench_mpi_")
    try:
        synth.make_eam_tables(spec, work)
        os.makedirs(os.path.join(work, "out"))
        synth.write_param_file(os.path.join(work, "run.param"), spec, tuple(global_cells), steps, "out/run",
                               potdir=".", pbc=(1, 1, 1), starttemp=0.0258, nbl_margin=0.4, eng_int=0,
                               cpu_dim=tuple(cpu_dim))
        ring_mb = max(4, int(4 * global_cells[1] * global_cells[2] * 2 * 80 * 2 / 1e6) + 1)
        env = dict(os.environ, SHMPI_NP=str(nproc), SHMPI_RING_MB=str(ring_mb))
        t0 = time.perf_counter()
        res = subprocess.run([REF_MPI, "-p", "run.param"], cwd=work, env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, timeout=900)
        wall = time.perf_counter() - t0
        m = re.search(r"([0-9.eE+-]+) seconds excluding setup time", res.stdout)
        n = re.search(r"Did (\d+) steps with (\d+) atoms", res.stdout)
        if res.returncode != 0 or not m or not n:
            raise RuntimeError("reference MPI run failed:\n" + res.stdout[-1500:])
        tm, nsteps, natoms = float(m.group(1)), int(n.group(1)), int(n.group(2))
        return {"value": natoms * nsteps / tm, "natoms": natoms, "steps": nsteps, "nproc": nproc,
                "main_loop_s": tm, "wall_s": wall, "binary": os.path.basename(REF_MPI), "mode": "mpi",
                "cpu_dim": list(cpu_dim)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def grid3(nproc, cells):
    """IMD-style 3-D process grid for nproc ranks: the factorisation (px, py, pz) with px*py*pz = nproc that divides
    the lattice and minimises the surface of a rank's sub-box (what imd's users pick by hand for cpu_dim)."""
    best = None
    for px in range(1, nproc + 1):
        if nproc % px or cells[0] % px:
            continue
        for py in range(1, nproc // px + 1):
            if (nproc // px) % py or cells[1] % py:
                continue
            pz = nproc // px // py
            if cells[2] % pz:
                continue
            lx, ly, lz = cells[0] / px, cells[1] / py, cells[2] / pz
            if min(lx, ly, lz) < 4:                      # >= 4 lattice constants = 16 A > one MD cell per rank
                continue
            surf = lx * ly + ly * lz + lx * lz
            if best is None or surf < best[0]:
                best = (surf, (px, py, pz))
    return best[1] if best else None


def film_geometry(cells, world):
    """Global film of the EAM+TTM workload: (cells.x * world) x cells.y x cells.z lattice cells, VAC_CELLS of vacuum
    in front and behind; ttmdimx ~ box_x / 16 A rounded to a multiple of 8 * world (so that the reference's
    P = 2, 4, 8 ... ranks and the GPUs' x slabs all divide it)."""
    from imd_b200 import synth
    a = synth.AL.a
    ncx = cells[0] * world
    lx = (ncx + 2 * VAC_CELLS) * a
    unit = 8 * world
    gx = max(unit, int(round(lx / FD_CELL / unit)) * unit)
    return ncx, lx, gx


def run_reference_ttm(cells, world, steps, nproc, nrb, substeps, binary=None, env_extra=None, timeout=1500):
    """The reference's own EAM + TTM(1-D) MPI build (imd_ttm.c:75-156 per step, move_atoms_ttm) over shmpi/ on
    the bench film: ensemble ttm, pbc 0 1 1, shiftx_front/rear vacuum, the same FD lattice, sub-step count and pulse
    as the GPU arm (analytic Beer-Lambert source on both sides, ref_hooks_ttm.c).  TTM1D talks to ranks myid +- 1
    (imd_ttm.c:884-939, 3290-3333): cpu_dim must be P 1 1, P >= 2."""
    from imd_b200 import synth
    binary = binary or (REF_TTM_NRB if nrb else REF_TTM)
    if not os.path.exists(binary):
        return None
    spec = synth.AL
    ncx, lx, gx = film_geometry(cells, world)
    work = tempfile.mkdtemp(prefix="imdbench_ttm_")
    try:
        synth.make_eam_tables(spec, work)
        synth.write_eos_files(synth.make_eos_tables(), work)
        run = os.path.join(work, "w")
        os.makedirs(os.path.join(run, "out"))
        tp = bench_ttm_params(gx, substeps, False)
        ex = tp.param_text() + "shiftx_front %.17g\nshiftx_rear %.17g\n" % (VAC_CELLS * spec.a, VAC_CELLS * spec.a)
        if nrb:
            ex += "nrb_alat %.17g\nnrb_eps 0.3\nnrb_k 1.2\n" % spec.a
        synth.write_param_file(os.path.join(run, "run.param"), spec, (ncx, cells[1], cells[2]), steps, "out/run",
                               potdir="..", pbc=(0, 1, 1), starttemp=0.0258, nbl_margin=0.4, eng_int=0, extra=ex,
                               ensemble="ttm", cpu_dim=(nproc, 1, 1))
        # shmpi ring per rank pair (virtual memory, touched where used): two cell layers of a yz plane (~4 atoms per cubic
        # cell and layer) at up to 240 bytes per atom (the -DNRB -DREFPOS build sends 18 doubles and more), x2 headroom
        ring_mb = max(8, int(4 * cells[1] * cells[2] * 2 * 240 * 2 / 1e6) + 1)
        env = dict(os.environ, SHMPI_NP=str(nproc), SHMPI_RING_MB=str(ring_mb),
                   IMD_SRC_MU="%.17g" % tp.src_mu, IMD_SRC_ABS="%.17g" % tp.src_abs)
        env.pop("IMD_DUMP_DIR", None)
        env.update(env_extra or {})
        t0 = time.perf_counter()
        res = subprocess.run([binary, "-p", "run.param"], cwd=run, env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, timeout=timeout)
        wall = time.perf_counter() - t0
        m = re.search(r"([0-9.eE+-]+) seconds excluding setup time", res.stdout)
        n = re.search(r"Did (\d+) steps with (\d+) atoms", res.stdout)
        if res.returncode != 0 or not m or not n:
            raise RuntimeError("reference TTM run failed:\n" + res.stdout[-1500:])
        tm, nsteps, natoms = float(m.group(1)), int(n.group(1)), int(n.group(2))
        return {"value": natoms * nsteps / tm, "natoms": natoms, "steps": nsteps, "nproc": nproc,
                "main_loop_s": tm, "wall_s": wall, "binary": os.path.basename(binary), "mode": "mpi",
                "cpu_dim": [nproc, 1, 1], "ttmdimx": gx,
                "stdout_tail": "\n".join(ln for ln in res.stdout.splitlines() if "IMD_CUDA_TIMING" in ln or "imd_cuda:" in ln)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


IMD_CUDA_TTM = os.path.join(ROOT, "integration", "_bin", "imd_mpi_eam_ttm_tmm_nbl_cuda")
IMD_CUDA_TTM_NRB = os.path.join(ROOT, "integration", "_bin", "imd_mpi_eam_ttm_tmm_nrb_nbl_cuda")


def run_imd_cuda_binary(cells, steps, substeps, world=1, nrb=False):
    """IMD ITSELF with the CUDA hot path (integration/imd_cuda.c: the reference's own host program, parameter parser,
    generator and main loop, linked against libimdcu.so) on the bench film, one rank per GPU (cpu_dim world 1 1 over
    shmpi/, x slabs over NCCL; with the non-reflecting boundary the -DNRB -DREFPOS binary): the same parameter
    file the reference arm gets except cpu_dim.  Atoms stay on the device; per step only the global sums come
    back.  Timed by IMD's own main-loop timer (MPI_Wtime: wall clock), like the reference arm.  That timer also
    holds what happens ONCE inside the first calc_forces (CUDA context, cell_array packed and uploaded, the count pass
    of the first list build: ~1.2 s, and it varies between runs); the stub reports the duration of that first call
    (IMD_CUDA_TIMING=1) and the rate is the other steps over the rest of the timer."""
    r = run_reference_ttm(cells, world, steps, world, nrb, substeps, binary=IMD_CUDA_TTM_NRB if nrb else IMD_CUDA_TTM,
                          env_extra={"IMD_CUDA_TIMING": "1"}, timeout=300)
    if not r:
        return None
    m = re.search(r"first calc_forces ([0-9.eE+-]+) s; per step calc_forces ([0-9.eE+-]+) ms, calc_ttm ([0-9.eE+-]+) ms, "
                  r"move_atoms ([0-9.eE+-]+) ms", r.get("stdout_tail", ""))
    if not m:
        raise RuntimeError("IMD-with-CUDA run printed no IMD_CUDA_TIMING line: %s" % r.get("stdout_tail", "")[-300:])
    first, in_lib = float(m.group(1)), float(m.group(2)) + float(m.group(3)) + float(m.group(4))
    rest, nrest = r["main_loop_s"] - first, r["steps"] - 1
    return {"value": r["natoms"] * nrest / rest, "steps": nrest, "ms_per_step": 1e3 * rest / nrest, "first_call_s": first,
            "ms_per_step_in_library": in_lib, "binary": r["binary"],
            "sample": ("integration/_bin/%s (IMD's own host + libimdcu.so), %d rank(s) on %d GPU(s)%s, the bench film "
                       "(%d atoms; the host's equal-width x slabs), %d steps: "
                       "IMD's main-loop timer %.3f s, of which the first calc_forces %.3f s (CUDA context, cell_array "
                       "packed and uploaded, first list build; not part of the rate); the other %d steps %.3f ms each, "
                       "%.3f ms of that inside libimdcu.so on rank 0"
                       % (r["binary"], world, world, ", non-reflecting rear boundary" if nrb else "", r["natoms"], r["steps"],
                          r["main_loop_s"], first, nrest, 1e3 * rest / nrest, in_lib))}


def bench_ttm_params(gx, substeps, tmm):
    from imd_b200 import synth
    return synth.TtmParams(ttmdimx=gx, fd_n_timesteps=substeps, fd_min_atoms=100, vlatdim=30, vlatbuffer=8, I0=5.0e15,
                           laser_t_0=100e-15, laser_sigma_t=42.5e-15, use_tmm=1 if tmm else 0, tmm_threshold=4.0)


def reference_sample(args, world, steps):
    """The reference on the bench workload itself, with all the host cores it can use.
    EAM+TTM film: the -DTTM MPI build on P x-slab ranks, P = the largest divisor of ttmdimx <= cores (TTM1D needs
    cpu_dim P 1 1).  NVE slab: the -DNVE MPI build, x slabs (cpu_dim P 1 1) AND IMD's 3-D process grid, the faster
    of the two is reported; concurrent serial runs of a 24^3 sample as the fallback."""
    cores = max(1, host_cores())
    cells = args.cells
    if args.ttm:
        ncx, lx, gx = film_geometry(cells, world)
        nproc = max([p for p in range(2, min(cores, 64) + 1) if gx % p == 0] or [2])
        r = run_reference_ttm(cells, world, steps, nproc, args.nrb, args.ttm_substeps)
        if r:
            r["sample"] = ("the reference's own EAM+TTM MPI build (%s: -DMPI -DNBL -DEAM2 -DTTM -DTTM1D -DTMM -DPDECAY%s, author's "
                           "flags -O3 -mavx2 -ffast-math) over shmpi/, %d ranks (cpu_dim %d 1 1, the only layout TTM1D "
                           "supports) on %d host cores, Al film %dx%dx%d cells = %d atoms, ttmdimx %d, fd_n_timesteps %d, "
                           "%d steps (main loop %.1f s, wall %.1f s)"
                           % (r["binary"], " -DNRB -DREFPOS" if args.nrb else "", nproc, nproc, cores, ncx, cells[1], cells[2],
                              r["natoms"], gx, args.ttm_substeps, r["steps"], r["main_loop_s"], r["wall_s"]))
        return r
    gx = cells[0] * world
    gcells = (gx, cells[1], cells[2])
    nproc = max([p for p in range(1, min(cores, 64) + 1) if gx % p == 0 and gx // p >= 4] or [1])
    tried = []
    try:
        cands = [(nproc, 1, 1)] if nproc >= 2 else []
        n3 = max([p for p in range(2, min(cores, 64) + 1) if grid3(p, gcells)] or [0])
        if n3 >= 2 and grid3(n3, gcells) not in cands:
            cands.append(grid3(n3, gcells))
        best = None
        for cd in cands:
            r = run_reference_mpi(gcells, steps, cd)
            if r:
                tried.append("cpu_dim %d %d %d: %.3g atom-steps/s" % (cd[0], cd[1], cd[2], r["value"]))
                if best is None or r["value"] > best["value"]:
                    best = r
        if best:
            cd = best["cpu_dim"]
            best["sample"] = ("the reference's MPI build (%s, author's flags) over shmpi/ on %d host cores, Al fcc %dx%dx%d "
                              "cells = %d atoms, %d NVE steps, 300 K; tried %s; reported cpu_dim %d %d %d = %d ranks "
                              "(main loop %.1f s)" % (best["binary"], cores, gx, cells[1], cells[2], best["natoms"], best["steps"],
                                                      "; ".join(tried), cd[0], cd[1], cd[2], best["nproc"], best["main_loop_s"]))
            return best
    except Exception as ex:                               # fall through to the serial form
        sys.stderr.write("bench: MPI reference failed (%s); using concurrent serial runs\n" % ex)
    r = run_reference_sample((24, 24, 24), steps, cores)
    if r:
        r["mode"] = "serial x%d" % cores
        r["sample"] = ("%d concurrent serial runs of the unmodified reference (%s), each Al fcc 24^3 cells = %d atoms, "
                       "%d NVE steps (%.1f s wall)" % (r["nproc"], r["binary"], r["natoms"], r["steps"], r["wall_s"]))
    return r


def reference_steps(cells, world, want):
    """Bound the CPU sample: about 3e8 atom-steps (10-30 s on 16 cores), never fewer than 5 steps."""
    natoms = 4 * cells[0] * cells[1] * cells[2] * world
    return int(max(5, min(want, 3.0e8 / natoms)))


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def reference_arm(args, rank, world):
    if rank != 0:
        return 0
    steps = reference_steps(args.cells, max(1, args.gpus), args.steps + args.warmup)
    failure = None
    try:
        r = reference_sample(args, max(1, args.gpus), steps)
    except Exception as ex:                               # the arm must always print its line
        r, failure = None, "reference run failed: %s" % " ".join(str(ex).split())[-400:]
    base = {"metric": metric_name(args), "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args)}
    if r is None:
        base["unavailable"] = failure or "oracle/_ref binaries not built (make -C oracle ref needs /root/reference)"
        print(json.dumps(base))
        return 0
    sample = r["sample"]
    base.update({"value": r["value"], "ms_per_step": 1e3 * r["main_loop_s"] / r["steps"],
                 "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["nproc"], "kind": "reference",
                                  "sample": sample},
                 "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(base))
    return 0


def workload_config(args):
    c = args.cells
    n = 4 * c[0] * c[1] * c[2]
    if getattr(args, "ttm", False):
        w = ("Al fcc EAM2 + 1-D TTM film (calc_forces, calc_ttm, move_atoms_ttm), %dx%dx%d cubic cells = %d atoms per "
             "GPU, vacuum 10 a front and rear, pbc 0 1 1, 300 K, timestep 0.1, nbl_margin 0.4, FD cell ~16 A, "
             "fd_n_timesteps %d, 100 fs pulse source (BASELINE.json configs[%d])"
             % (c[0], c[1], c[2], n, args.ttm_substeps, 2 if args.gpus <= 1 else 3))
        if getattr(args, "tmm", False):
            w += ", laser source by do_tmm (transfer-matrix Helmholtz solver per sub-step, synthetic eps_bb/K12 tables)"
        if getattr(args, "nrb", False):
            w += (", non-reflecting rear boundary (nrb_forces on the last atomic layer, nrb_k 1.2; "
                  "%d atoms in all)" % (n * max(1, args.gpus)))
    else:
        w = ("Al fcc EAM2 NVE slab, %dx%dx%d cubic cells = %d atoms per GPU, 300 K, timestep 0.1 (1.018 fs), "
             "nbl_margin 0.4 (BASELINE.json configs[1])" % (c[0], c[1], c[2], n))
    return {"workload": w, "atoms_per_gpu": n,
            "l2_policy": "inputs larger than L2 (neighbour list alone ~0.8 GB per 2M atoms)",
            "parallelism": "1 GPU" if args.gpus == 1 else "x-slab domain decomposition, %d GPUs" % args.gpus}


# ----------------------------------------------------------------------------- our arm

def system_box(cells, world=1, ttm=False):
    """Global box of the weak-scaling run: (cells.x * world) x cells.y x cells.z cubic cells, with --ttm vacuum in
    front of and behind the film along x."""
    from imd_b200 import synth
    a = synth.AL.a
    vac = VAC_CELLS * a if ttm else 0.0
    return np.diag([cells[0] * world * a + 2 * vac, cells[1] * a, cells[2] * a]).astype(np.float64)


def layer_histogram(cells, world, ttm, gdimx):
    """Atoms per global x cell layer of the lattice build_system makes, without making it: an fcc cell has two atomic
    planes along x, at (i + 0.25) a and (i + 0.75) a, of 2 cells.y cells.z atoms each (input of imdcu_balance_cuts)."""
    from imd_b200 import synth
    a = synth.AL.a
    vac = VAC_CELLS * a if ttm else 0.0
    ncx = cells[0] * world
    lx = ncx * a + 2 * vac
    x = (np.arange(ncx)[:, None] + np.array([0.25, 0.75])[None, :]).reshape(-1) * a + vac
    layer = np.clip((gdimx * (x / lx)).astype(np.int64), 0, gdimx - 1)
    return np.bincount(layer, minlength=gdimx) * (2 * cells[1] * cells[2])


def build_system(cells, rank=0, world=1, ttm=False, x_lo=None, x_hi=None):
    """This rank's part of the global lattice: the cubic cells that touch [x_lo, x_hi) (default: the rank's
    equal-width slab) plus one cell either side; the caller still filters with Imd.owned_mask, because the cuts of
    the library lie on cell-layer boundaries, not on lattice planes.  Returns ids that are global indices."""
    from imd_b200 import synth
    spec = synth.AL
    a = spec.a
    ncx = cells[0] * world
    vac = VAC_CELLS * a if ttm else 0.0
    lx = ncx * a + 2 * vac
    lo = rank * lx / world if x_lo is None else x_lo
    hi = (rank + 1) * lx / world if x_hi is None else x_hi
    i0 = max(0, int(np.floor((lo - vac) / a)) - 1)
    i1 = min(ncx, int(np.ceil((hi - vac) / a)) + 1)
    pos, box = synth.fcc_lattice((max(i1 - i0, 0), cells[1], cells[2]), a)
    pos[:, 0] += i0 * a + vac
    box[0, 0] = lx
    ids = (np.arange(pos.shape[0], dtype=np.int64) + 4 * i0 * cells[1] * cells[2] + 1).astype(np.int32)
    return spec, pos, box, ids


def rear_boundary_state(cells, world):
    """NRB state of the film's rear (+x) surface: init_nrb (host mirror, imd_b200/nrb.py) on the last
    lattice cell layer, which holds the boundary layer and the layer of its listed neighbours.  Every
    rank builds the same table; only the rank that owns those atoms finds them."""
    from imd_b200 import nrb, synth
    spec = synth.AL
    ncx = cells[0] * world
    pos, box = synth.fcc_lattice((1, cells[1], cells[2]), spec.a)
    pos[:, 0] += (ncx - 1) * spec.a + VAC_CELLS * spec.a
    box[0, 0] = ncx * spec.a + 2 * VAC_CELLS * spec.a
    ids = (np.arange(pos.shape[0], dtype=np.int64) + 4 * (ncx - 1) * cells[1] * cells[2] + 1).astype(np.int32)
    return nrb.init_nrb(pos, ids, box, (0, 1, 1), spec.a, 0.3)


def bind_to_gpu_cpus(local_rank):
    """Several ranks on one node: keep this rank's host thread -- and with it the pinned buffers it is about to allocate
    -- on the CPUs NVML names as closest to its GPU (nvmlDeviceSetCpuAffinity).  Matters for the host-buffer (e2e)
    path only: eight processes moving 384 MB per step each should not cross sockets.  Never fatal."""
    try:
        import pynvml
        h = nvml_handle(local_rank)
        if h is None:
            return "not bound (NVML unavailable)"
        before = len(os.sched_getaffinity(0))
        pynvml.nvmlDeviceSetCpuAffinity(h)
        cpus = sorted(os.sched_getaffinity(0))
        return "bound to %d of %d CPUs (%d..%d) near GPU %d" % (len(cpus), before, cpus[0], cpus[-1], local_rank)
    except Exception as ex:
        return "not bound (%s)" % str(ex)[:80]


def our_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from imd_b200 import host, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    host_binding = bind_to_gpu_cpus(local_rank) if world > 1 and not args.no_cpu_binding else "one rank: not bound"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    spec = synth.AL
    box = system_box(args.cells, world, args.ttm)
    natoms = 4 * args.cells[0] * args.cells[1] * args.cells[2]      # per GPU on average
    tables = synth.make_eam_tables(spec)
    comm = None
    if world > 1:
        from imd_b200 import capi
        obj = [capi.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        comm = (obj[0], rank, world)
    imd = host.Imd(local_rank, comm)
    imd.setup_potentials(tables)
    imd.make_box(box, (0, 1, 1) if args.ttm else (1, 1, 1), 0.4)
    imd.set_params(1, 0.1)
    xcut = None
    if world > 1 and not args.equal_slabs:
        # slabs of equal atom count on cell-layer boundaries: the film's vacuum would leave the end ranks
        # of equal-width slabs short of atoms and the step time is the slowest rank's
        from imd_b200 import capi
        g = int(imd.global_cell_dim[0])
        xcut = capi.balance_cuts(world, layer_histogram(args.cells, world, args.ttm, g), 2)
        imd.set_cuts(xcut)
    w = box[0, 0] / int(imd.global_cell_dim[0])
    spec, pos, box, ids = build_system(args.cells, rank, world, args.ttm, imd.slab["xoff"] * w,
                                       (imd.slab["xoff"] + imd.slab["ldimx"]) * w)
    mask = imd.owned_mask(pos)
    pos, ids = pos[mask], ids[mask]
    mom = synth.maxwell_momenta(pos.shape[0], spec.mass, 0.0258, seed=1234 + rank)
    imd.set_atoms(pos, mom, spec.mass, None, ids)
    cnt = torch.tensor([pos.shape[0]], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(cnt)
    if int(cnt.item()) != natoms * world:
        raise SystemExit("bench.py: slab generation lost atoms (%d != %d)" % (int(cnt.item()), natoms * world))
    slabs = None
    if world > 1:
        per = [torch.zeros(1, dtype=torch.int64, device="cuda") for _ in range(world)]
        dist.all_gather(per, torch.tensor([pos.shape[0]], dtype=torch.int64, device="cuda"))
        per = [int(t.item()) for t in per]
        slabs = {"kind": "equal width" if xcut is None else "equal atom count (imdcu_balance_cuts)",
                 "x_cuts": None if xcut is None else [int(x) for x in xcut], "cell_layers": int(imd.global_cell_dim[0]),
                 "atoms_per_rank_min": min(per), "atoms_per_rank_max": max(per)}
    ctx = imd.ctx
    if args.ttm:
        eos = synth.make_eos_tables()          # diag = None: all rectangles split the same way
        _, _, gx = film_geometry(args.cells, world)
        tp = bench_ttm_params(gx, args.ttm_substeps, args.tmm)
        imd.init_ttm(tp, eos, world)
    if args.nrb:
        st = rear_boundary_state(args.cells, world)
        if int((st.bnd == 1).sum()) != 2 * args.cells[1] * args.cells[2]:
            raise SystemExit("bench.py: NRB setup did not find the rear layer")
        imd.set_nrb(1.2, st)
    run = (lambda first, n: ctx.run_ttm(first, n, False)) if args.ttm else (lambda first, n: ctx.run_nve(first, n, False))
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ---- device-resident throughput
    sampler = ClockSampler(local_rank)
    sampler.start()                              # clocks under load: warm-up + timed region
    run(0, args.warmup)
    barrier()
    l0 = ctx.launch_count
    nbl0 = ctx.nbl_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    run(args.warmup, args.steps)
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    launches = ctx.launch_count - l0
    rebuilds = ctx.nbl_count - nbl0
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = natoms * world * args.steps / (ms * 1e-3)

    # ---- end to end through the C ABI with host buffers (pinned)
    ke = max(1, min(args.steps, args.e2e_steps))
    cap = int(max(natoms, pos.shape[0]) * 1.25) + 1024   # the slab population changes when atoms migrate
    hp = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
    hm = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
    np_p, np_m = hp.numpy(), hm.numpy()
    ctx.download_into(np_p, np_m)

    def e2e_step(s):
        ctx.update_atoms(np_p, np_m)             # H2D: this step's inputs (first ctx.n rows)
        out = ctx.calc_forces(s)                 # D2H: energy sums
        if args.ttm:
            ctx.calc_ttm(args.warmup + args.steps + 2 + s)
            out2, _ = ctx.move_atoms_ttm()
        else:
            out2, _ = ctx.move_atoms_nve()
        ctx.download_into(np_p, np_m)            # D2H: the step's results = the next step's inputs
        return out[0] + out2[0]

    for s in range(2):
        e2e_step(s)
    barrier()
    t0 = time.perf_counter()
    for s in range(ke):
        e2e_step(s)
    barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = natoms * world * ke / float(t.item())
    h2d = natoms * 48
    d2h = natoms * 48 + 2 * 64

    # ---- diagnostic: wall time of every step, one library call per step (each call ends with the integrator's
    # synchronise), with the list-rebuild counter: shows what a rebuild costs and how even the other steps are
    if args.per_step > 0:
        rows = []
        first = args.warmup + args.steps + 1000
        for s in range(args.per_step):
            barrier()
            t0 = time.perf_counter()
            run(first + s, 1)
            torch.cuda.synchronize()
            rows.append((round((time.perf_counter() - t0) * 1e3, 3), int(ctx.nbl_count)))
        if rank == 0:
            sys.stderr.write("PER_STEP_MS " + json.dumps(rows) + "\n")

    # ---- per-kernel times (CUDA events inside the library, separate short segment)
    roofline = None
    phases = None
    # every rank runs the segment (the steps exchange halos); rank 0 reports its own kernel times
    ctx.set_profiling(True)
    run(args.warmup + args.steps + 100, args.profile_steps)
    ms_ph, ln_ph = ctx.phase_times()
    ctx.set_profiling(False)
    barrier()
    if rank == 0:
        phases = {k: {"ms_total": ms_ph[k], "launches": ln_ph[k]} for k in ms_ph if ln_ph[k]}
        p1_ms = ms_ph["eam_pass1"] / max(1, ln_ph["eam_pass1"])
        p2_ms = ms_ph["eam_pass2"] / max(1, ln_ph["eam_pass2"])
        n_loc = ctx.n                                  # atoms this rank owns (= natoms on one GPU)
        alg_bytes = (64 + 4 * N_H_AL) * n_loc          # SURVEY 8d, pass 1: 64 + 4 N_h bytes per atom
        alg_flops = 2 * N_H_AL * 75 * n_loc            # ~75 FP64 flops per pair evaluation in pass 1, both sides
        peak, peak_src = measured_peak_gbs()
        achieved = alg_bytes / (p1_ms * 1e-3) / 1e9
        fp64_peak = ctx.probe_fp64()
        fp64_ach = alg_flops / (p1_ms * 1e-3) / 1e9
        traffic, traffic_src = measured_traffic("k_eam_pass1", n_loc)
        p1_kernel = {"0": "k_eam_pass1", "1": "k_eam_pass1_fused"}.get(os.environ.get("IMDCU_FUSED_TABLE", "2"), "k_eam_pass1_rep4")
        roofline = {"kernel": p1_kernel, "bound": "l1", "achieved": achieved, "peak": peak,
                    "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src, "alg_bytes_per_launch": alg_bytes, "ms_per_launch": p1_ms,
                    "fp64": {"achieved": fp64_ach, "peak": fp64_peak, "unit": "GFLOP/s", "frac": fp64_ach / fp64_peak,
                             "alg_flops_per_launch": alg_flops,
                             "peak_source": "measured in this run (imdcu_probe_fp64: independent DFMA chains on every SM)"},
                    "pass2": {"kernel": "k_eam_pass2", "ms_per_launch": p2_ms,
                              "achieved": (80 + 4 * N_H_AL) * n_loc / (p2_ms * 1e-3) / 1e9,
                              "frac": (80 + 4 * N_H_AL) * n_loc / (p2_ms * 1e-3) / 1e9 / peak},
                    "note": "binding unit by ncu: the L1 data pipe (l1tex__data_pipe_lsu_wavefronts 92 % of peak: per pair three "
                            "16-byte table rows from shared memory and one 32-byte gather), with issue slots 53 % and the FP64 "
                            "pipe 44 % busy; HBM and FP64 fractions are reported against their measured peaks, the north star's "
                            ">= 0.60 of HBM is NOT met by the force kernels (profiles/r02f_rep4_ncu.txt, DESIGN.md section 3)"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            r = reference_sample(args, 1, reference_steps(args.cells, 1, 30))
            if r:
                cpu = {"value": r["value"], "unit": UNIT, "cores": r["nproc"], "kind": "reference",
                       "sample": r["sample"]}
        except Exception as ex:                    # the baseline must not take the bench down
            cpu = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "failed: %s" % ex}

    if rank == 0:
        line = {"metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(args), "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": ke},
                "gpu_launches": int(launches), "nbl_rebuilds": int(rebuilds),
                "roofline": roofline, "cpu_baseline": cpu, "phases": phases}
        if slabs:
            line["slabs"] = slabs
        line["e2e"]["host_binding_rank0"] = host_binding
        if world > 1:
            line["e2e"]["bound_by"] = ("every rank moves its atoms through host buffers each step (%.0f MB up, %.0f MB down per rank): "
                                       "%d ranks share this node's PCIe root complexes and host memory" % (h2d / 1e6, d2h / 1e6, world))
    imd.close()
    if rank == 0:
        # second end-to-end figure: the drop-in binary itself (the number a user of the reference sees)
        if args.ttm and not args.tmm and not args.no_cpu_baseline and args.imd_binary_steps > 0:
            try:
                torch.cuda.synchronize()
                rb = run_imd_cuda_binary(args.cells, args.imd_binary_steps, args.ttm_substeps, world, bool(args.nrb))
                if rb:
                    line["e2e_imd_binary"] = {"value": rb["value"], "unit": UNIT, "ms_per_step": rb["ms_per_step"],
                                              "ms_per_step_in_library": rb["ms_per_step_in_library"],
                                              "first_call_s": rb["first_call_s"], "sample": rb["sample"]}
            except Exception as ex:
                line["e2e_imd_binary"] = {"value": None, "unit": UNIT, "sample": "failed: %s" % str(ex)[-300:]}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


# ----------------------------------------------------------------------------- 3-D TTM sweep (configs[4])

def ttm3d_reference_sample(n, sweeps, nproc):
    """nproc concurrent runs of oracle/_ref/ref3d_diff (the reference's do_DIFF + ttm_fill_ghost_layers,
    one rank each; imd_ttm_3D.c has no parallel driver that runs here) on an n^3 lattice."""
    from oracle import ttm3d_np as t3
    exe = os.path.join(ROOT, "oracle", "_ref", "ref3d_diff")
    if not os.path.exists(exe):
        return None
    work = tempfile.mkdtemp(prefix="imdbench_t3_")
    try:
        f = t3.make_case((n, n, n), seed=1, hole=False, pbc=(0, 1, 1))
        with open(os.path.join(work, "in.bin"), "wb") as o:
            np.array([n, n, n, 20, 250, 1, sweeps, 0, 1, 1, 0, 0], np.int32).tofile(o)
            np.array([16.0, 16.0, 16.0, 0.1 / 250]).tofile(o)
            for k in t3.FIELDS:
                np.ascontiguousarray(f[k], dtype=np.float64).tofile(o)
        env = dict(os.environ, SHMPI_NP="1", SHMPI_RING_MB="64")
        t0 = time.perf_counter()
        procs = [subprocess.Popen([exe, "in.bin", "out%d.bin" % p], cwd=work, env=env, stdout=subprocess.PIPE,
                                  stderr=subprocess.STDOUT, text=True) for p in range(nproc)]
        outs = [p.communicate()[0] for p in procs]
        wall = time.perf_counter() - t0
        rate = 0.0
        for o, p in zip(outs, procs):
            m = re.search(r"ref3d: (\d+) sweeps of (\d+) cells in ([0-9.eE+-]+) s", o)
            if p.returncode != 0 or not m:
                raise RuntimeError("ref3d_diff failed: " + o[-500:])
            rate += int(m.group(1)) * int(m.group(2)) / float(m.group(3))
        return {"value": rate, "nproc": nproc, "wall_s": wall,
                "sample": "%d concurrent single-rank runs of the reference's do_DIFF + ttm_fill_ghost_layers "
                          "(oracle/_ref/ref3d_diff, imd_ttm_3D.c compiled -O2), %d^3 cells, %d sweeps each (%.1f s wall)"
                          % (nproc, n, sweeps, wall)}
    finally:
        shutil.rmtree(work, ignore_errors=True)


def ttm3d_arm(args, rank, local_rank, world):
    """BASELINE.json configs[4]: standalone 3-D TTM electron-temperature diffusion sweep, cell-updates/s.
    One step = one sweep { do_DIFF; ttm_fill_ghost_layers } over the lattice; N GPUs = x slabs of a lattice
    N times as long (weak scaling), ghost planes {natoms, temp, fd_k} over NCCL after every sweep."""
    metric, unit = "TTM-3D cell-updates/s", "cell-updates/s"
    nx, ny, nz = args.grid
    cfg = {"workload": "3-D TTM diffusion sweep (do_DIFF + ttm_fill_ghost_layers), %dx%dx%d FD cells, film with "
                       "vacuum front (1/8 inactive), pbc 0 1 1 (BASELINE.json configs[4])" % (nx, ny, nz),
           "cells": nx * ny * nz, "l2_policy": "inputs larger than L2 (11 fields x %.1f GB)" % (nx * ny * nz * 8 / 1e9),
           "parallelism": "1 GPU" if world == 1 else "x slabs of a %dx%dx%d lattice on %d GPUs, NCCL ghost planes"
                          % (nx * world, ny, nz, world)}
    if args.impl == "reference":
        if rank != 0:
            return 0
        r = ttm3d_reference_sample(128, max(2, min(args.steps, 10)), max(1, host_cores()))
        base = {"metric": metric, "unit": unit, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": cfg}
        if r is None:
            base["unavailable"] = "oracle/_ref/ref3d_diff not built"
        else:
            base.update({"value": r["value"], "ms_per_step": None,
                         "cpu_baseline": {"value": r["value"], "unit": unit, "cores": r["nproc"], "kind": "reference",
                                          "sample": r["sample"]},
                         "e2e": {"value": r["value"], "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                         "gpu_launches": 0})
        print(json.dumps(base))
        return 0
    import torch
    import torch.distributed as dist
    from imd_b200 import capi
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = capi.Context(local_rank)
    if world > 1:
        obj = [capi.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        ctx.comm_init(obj[0], rank, world)
    e = (nx + 2, ny + 2, nz + 2)
    rng = np.random.default_rng(7 + rank)
    nat = np.full(e, 60, np.int32)
    if rank == 0:
        nat[: max(1, nx // 8)] = 3                     # vacuum in front of the film (first slab only)
    temp = (0.05 + 0.5 * rng.random(e, dtype=np.float32)).astype(np.float64)
    fd_k = (10.0 + 20.0 * rng.random(e, dtype=np.float32)).astype(np.float64)
    const = lambda v: np.full(e, v)                                                    # noqa: E731
    ctx.ttm3d_init((nx, ny, nz), (16.0, 16.0, 16.0), (0, 1, 1), 20)
    ctx.ttm3d_upload(natoms=nat, temp=temp, fd_k=fd_k, fd_g=const(3e-4), Ce=const(2.5e-3), md_temp=const(0.03),
                     source=const(1e-4), U=const(1.0), xi=const(0.0), dens=const(2700.0))
    del fd_k, nat
    ctx.ttm3d_fill_ghost_layers()
    tau = 0.1 / 250
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)

    def barrier():
        ctx.synchronize()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    sampler = ClockSampler(local_rank)
    sampler.start()
    ctx.ttm3d_diff(tau, 250, 1, args.warmup, want_eabs=False)
    barrier()
    l0 = ctx.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    ctx.ttm3d_diff(tau, 250, 1, args.steps, want_eabs=False)
    ev1.record(stream)
    barrier()
    clocks = sampler.stop()
    launches = ctx.launch_count - l0
    t = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    cells = nx * ny * nz
    value = cells * world * args.steps / (ms * 1e-3)
    # end to end with host buffers: upload Te, one sweep, download Te (pinned)
    ht = torch.from_numpy(temp).pin_memory()
    hn = ht.numpy()
    ke = max(1, min(args.steps, args.e2e_steps, 5))
    L = ctx.L
    for _ in range(1):
        ctx.ttm3d_upload(temp=hn); ctx.ttm3d_diff(tau, 250, 1, 1); L.imdcu_ttm3d_download(ctx.h, None, capi._dp(hn), None, None)
    barrier()
    t0 = time.perf_counter()
    for _ in range(ke):
        ctx.ttm3d_upload(temp=hn)
        ctx.ttm3d_diff(tau, 250, 1, 1)
        L.imdcu_ttm3d_download(ctx.h, None, capi._dp(hn), None, None)
    barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = cells * world * ke / float(t.item())
    # kernel-only time of k_ttm3d_diff (phase 7)
    ctx.set_profiling(True)
    ctx.ttm3d_diff(tau, 250, 1, args.profile_steps, want_eabs=False)
    ms_ph, ln_ph = ctx.phase_times()
    ctx.set_profiling(False)
    k_ms = ms_ph["ttm"] / max(1, ln_ph["ttm"])
    peak, peak_src = measured_peak_gbs()
    alg = 108 * cells
    ach = alg / (k_ms * 1e-3) / 1e9
    traffic, traffic_src = measured_traffic("k_ttm3d_diff", cells)
    roofline = {"kernel": "k_ttm3d_diff_direct", "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "alg_bytes_per_launch": alg, "ms_per_launch": k_ms,
                "note": "108 B per cell update: 4 + 10 x 8 read, 3 x 8 written (SURVEY 8d)"}
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            r = ttm3d_reference_sample(128, 5, max(1, host_cores()))
            if r:
                cpu = {"value": r["value"], "unit": unit, "cores": r["nproc"], "kind": "reference", "sample": r["sample"]}
        except Exception as ex:
            cpu = {"value": None, "unit": unit, "cores": 0, "kind": "reference", "sample": "failed: %s" % ex}
    if rank == 0:
        print(json.dumps({"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg, "clocks": clocks,
                          "e2e": {"value": e2e_value, "unit": unit, "h2d_bytes_per_step": int(np.prod(e)) * 8,
                                  "d2h_bytes_per_step": int(np.prod(e)) * 8, "steps": ke},
                          "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu}))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cells", type=int, nargs=3, default=None,
                    help="cubic fcc cells per GPU (default 100 100 100 for the EAM+TTM film, 80 80 80 for --workload nve)")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--profile-steps", type=int, default=5)
    ap.add_argument("--imd-binary-steps", type=int, default=400,
                    help="N = 1: steps of the extra run of IMD-with-CUDA (integration/_bin), 0 = skip")
    ap.add_argument("--no-cpu-binding", action="store_true", help="N > 1: do not bind the rank's host thread to the CPUs near its GPU")
    ap.add_argument("--per-step", type=int, default=0, help="diagnostic: also time this many single steps, printed to stderr")
    ap.add_argument("--equal-slabs", action="store_true", help="N > 1: the reference's equal-width x slabs instead of slabs of equal atom count")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ttm", action="store_true", help="same as --workload ttm (kept for older command lines)")
    ap.add_argument("--ttm-substeps", type=int, default=250, help="fd_n_timesteps (manual example: 250)")
    ap.add_argument("--tmm", action="store_true",
                    help="with --ttm: the reference's transfer-matrix laser source (do_tmm) instead of the analytic one")
    ap.add_argument("--nrb", action="store_true",
                    help="EAM+TTM film: non-reflecting rear boundary (default on several GPUs = BASELINE.json configs[3])")
    ap.add_argument("--no-nrb", action="store_true", help="EAM+TTM film on several GPUs without the NRB")
    ap.add_argument("--workload", default="ttm", choices=["ttm", "nve", "md", "ttm3d"],
                    help="ttm (default): EAM+TTM film, the workload BASELINE.json's metric names (configs[2] on 1 GPU, "
                         "configs[3] on several); nve: Al NVE slab (configs[1]); md: nve unless --ttm is given; "
                         "ttm3d: standalone 3-D TTM sweep (configs[4])")
    ap.add_argument("--grid", type=int, nargs=3, default=[512, 512, 512], help="FD cells for --workload ttm3d")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.workload == "ttm":
        args.ttm = True
    elif args.workload == "nve":
        args.ttm = False
    if args.cells is None:
        args.cells = [100, 100, 100] if args.ttm else [80, 80, 80]
    if args.ttm and args.gpus > 1 and not args.no_nrb:
        args.nrb = True
    if not args.ttm:
        args.nrb = args.tmm = False
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # stdout carries the ONE JSON line and nothing else: library chatter (NCCL prints its version to
    # stdout when NCCL_DEBUG is set) goes to stderr while the run lasts
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            if args.workload == "ttm3d":
                rc = ttm3d_arm(args, rank, local_rank, world)
            elif args.impl == "reference":
                rc = reference_arm(args, rank, world)
            else:
                rc = our_arm(args, rank, local_rank, world)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
        sys.stdout.write(buf.getvalue())
        sys.stdout.flush()
    return rc


if __name__ == "__main__":
    sys.exit(main())
