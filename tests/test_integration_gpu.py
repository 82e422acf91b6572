"""IMD itself, built with the CUDA hot path (integration/_bin/*_cuda), against the unmodified reference.

Both programs are IMD's own host (parameter parser imd_param.c:4128, potential reader
imd_potential.c:161-466, generator, writers, main loop imd_main_3d.c:151-916) compiled in place from the
reference tree; they differ only in who provides calc_forces / move_atoms_nve / check_nblist
(imd_forces_nbl.c + imd_integrate.c there, integration/imd_cuda.c -> libimdcu.so here).  Each test writes ONE
parameter file and one set of potential files in IMD's formats and hands the same files to both binaries, so
what is compared is the reference's own output files (.eng from write_eng_file imd_io.c:2290, .chkpt from
write_config imd_io_3d.c:721-979).  The reference's own example inputs cannot travel to the GPU box and may
not be copied, hence imd_b200.synth writes the files; the readers are the reference's.

Binaries are built where /root/reference exists (`make -C integration`, `make -C oracle ref`) and shipped.
"""
import os
import subprocess

import numpy as np
import pytest

from imd_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl")
CUDA_BIN = os.path.join(ROOT, "integration", "_bin", "imd_eam_nve_nbl_cuda")

needs_bins = pytest.mark.skipif(not (os.path.exists(REF_BIN) and os.path.exists(CUDA_BIN)),
                                reason="integration/_bin or oracle/_ref not built (needs the reference sources)")


def write_inputs(work, steps, box=(8, 8, 8), extra=""):
    synth.make_eam_tables(synth.CU, work)
    os.makedirs(os.path.join(work, "out"), exist_ok=True)
    synth.write_param_file(os.path.join(work, "run.param"), synth.CU, box, steps, "out/run", pbc=(1, 1, 1),
                           starttemp=0.08, nbl_margin=0.5, eng_int=1, extra=extra)


def run(binary, work, *args):
    env = dict(os.environ)
    env.pop("IMD_DUMP_DIR", None)
    res = subprocess.run([binary, "-p", "run.param", *args], cwd=work, env=env, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True, timeout=600)
    return res


def eng_of(work):
    return np.atleast_2d(np.loadtxt(os.path.join(work, "out", "run.eng"), comments="#"))


def chkpt_path(work, number):
    """<outfiles>.<number>.chkpt; the writer pads the number to five digits (the restart reader takes both)."""
    for name in ("run.%05d.chkpt" % number, "run.%d.chkpt" % number):
        p = os.path.join(work, "out", name)
        if os.path.exists(p):
            return p
    raise FileNotFoundError("no checkpoint %d under %s/out: %s" % (number, work, os.listdir(os.path.join(work, "out"))))


def read_chkpt(path):
    """Atom lines of a .chkpt (ASCII, header lines start with '#'): number type mass x y z vx vy vz Epot ..."""
    rows = [ln.split() for ln in open(path) if ln.strip() and not ln.startswith("#")]
    a = np.array(rows, dtype=np.float64)
    return a[np.argsort(a[:, 0])]


@needs_bins
def test_cuda_binary_fails_loudly_without_a_gpu(tmp_path):
    """No CPU fallback: without a CUDA device the _cuda binary stops in imderror with the library's message."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    work = str(tmp_path)
    write_inputs(work, 2, box=(4, 4, 4))
    res = run(CUDA_BIN, work)
    assert res.returncode != 0, res.stdout[-2000:]
    assert "imdcu" in res.stdout, res.stdout[-2000:]


@needs_bins
@pytest.mark.gpu
def test_cuda_binary_eng_equals_reference(tmp_path):
    """40 NVE steps of 2048 Cu atoms: .eng of the CUDA build equals the reference's row by row."""
    steps = 40
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps)
    ra, rb = run(REF_BIN, wa), run(CUDA_BIN, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "hot path on CUDA device" in rb.stdout
    ea, eb = eng_of(wa), eng_of(wb)
    assert ea.shape == eb.shape and ea.shape[0] >= steps
    assert np.array_equal(ea[:, 0], eb[:, 0])                       # time column
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    eta, etb = ea[:, 1] + 1.5 * ea[:, 2], eb[:, 1] + 1.5 * eb[:, 2]
    err_drift = np.abs((eta - eta[0]) - (etb - etb[0])).max()
    print("imd_eam_nve_nbl_cuda vs reference: Epot/atom rel %.2e, T rel %.2e, |d drift| %.2e eV/atom"
          % (err_pot, err_kin, err_drift))
    assert err_pot < 1e-9 and err_kin < 1e-8
    assert err_drift < 1e-9 * abs(eta[0])


@needs_bins
@pytest.mark.gpu
def test_chkpt_written_by_cuda_build_restarts_in_reference(tmp_path):
    """sync_to_host: the CUDA build writes run.1.chkpt at step 20 through IMD's own write_config;
    the file equals the reference's own checkpoint of that step, and the PLAIN reference restarted from it
    (-r 1) lands on the uninterrupted reference run."""
    steps, at = 40, 20
    extra = "checkpt_int  %d\n" % at
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps, extra=extra)
    ra, rb = run(REF_BIN, wa), run(CUDA_BIN, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    ca, cb = read_chkpt(chkpt_path(wa, 1)), read_chkpt(chkpt_path(wb, 1))
    assert ca.shape == cb.shape
    assert np.array_equal(ca[:, :3], cb[:, :3])                     # number, type, mass
    scale = np.abs(ca[:, 3:]).max(axis=0) + 1e-300
    err = (np.abs(ca[:, 3:] - cb[:, 3:]).max(axis=0) / scale).max()
    print("run.1.chkpt (step %d), CUDA build vs reference: max rel err over columns %.2e" % (at, err))
    assert err < 1e-8
    eng_full = eng_of(wa)
    rc = run(REF_BIN, wb, "-r", "1")                                # the reference continues the CUDA build's run
    assert rc.returncode == 0, rc.stdout[-3000:]
    assert "Restarting from" in rc.stdout
    eng_re = eng_of(wb)
    tail_ref = eng_full[eng_full[:, 0] > at * 0.1 + 1e-9]
    tail_re = eng_re[-tail_ref.shape[0]:]
    assert tail_ref.shape[0] >= steps - at - 1
    assert np.allclose(tail_re[:, 0], tail_ref[:, 0], rtol=0, atol=1e-9)
    err_pot = np.abs(tail_re[:, 1] - tail_ref[:, 1]).max() / np.abs(tail_ref[:, 1]).max()
    err_kin = np.abs(tail_re[:, 2] - tail_ref[:, 2]).max() / np.abs(tail_ref[:, 2]).max()
    print("reference restarted from the CUDA build's checkpoint: Epot rel %.2e, T rel %.2e" % (err_pot, err_kin))
    assert err_pot < 1e-9 and err_kin < 1e-8
