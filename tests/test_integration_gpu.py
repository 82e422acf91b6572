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


def run(binary, work, *args, env_extra=None):
    env = dict(os.environ, **(env_extra or {}))
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


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["4point", "spline"])
def test_cuda_binary_other_interpolation_modes_equal_reference(tmp_path, mode):
    """The reference's compile-time FOURPOINT / SPLINE table access (make-target substrings, potaccess.h:365-457,
    503-582) in the drop-in: imd_eam_nve_nbl_<mode>_cuda hands the mode (and, for the spline, the second-derivative
    tables read_pot_table built, pot_table_t.table2) to the library; .eng against the reference built the same way."""
    ref = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl_" + mode)
    cuda = os.path.join(ROOT, "integration", "_bin", "imd_eam_nve_nbl_%s_cuda" % mode)
    if not (os.path.exists(ref) and os.path.exists(cuda)):
        pytest.skip("the %s builds are missing (needs the reference sources)" % mode)
    steps = 40
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps)
    ra, rb = run(ref, wa), run(cuda, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    ea, eb = eng_of(wa), eng_of(wb)
    assert ea.shape == eb.shape and ea.shape[0] >= steps
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    e3 = eng_of_3point(tmp_path, steps)
    differs = np.abs(ea[:, 1] - e3[:, 1]).max() / np.abs(ea[:, 1]).max()
    print("imd_eam_nve_nbl_%s_cuda vs the reference built with the same mode: Epot rel %.2e, T rel %.2e "
          "(the 3-point reference differs by %.1e)" % (mode, err_pot, err_kin, differs))
    assert err_pot < 1e-11 and err_kin < 1e-13                      # observed 7e-13 / 3e-13 and 2e-15
    assert differs > 100 * max(err_pot, 1e-13)                     # the mode matters: not the 3-point path by accident


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["3point", "4point", "spline"])
def test_cuda_binary_two_species_equals_reference(tmp_path, mode):
    """ntypes = 2 through the files: Ni3Al on the L1_2 structure (IMD's own generator, coordname _l12),
    4/4/2-column potential files with every column on its own r^2 grid and asymmetric rho columns; free x surfaces.
    The drop-in hands pot_table_t with its columns to the library as the reference's readers left it."""
    suffix = "" if mode == "3point" else "_" + mode
    ref = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl" + suffix)
    cuda = os.path.join(ROOT, "integration", "_bin", "imd_eam_nve_nbl%s_cuda" % suffix)
    if not (os.path.exists(ref) and os.path.exists(cuda)):
        pytest.skip("the %s builds are missing (needs the reference sources)" % mode)
    steps = 30
    spec = synth.NI3AL
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(os.path.join(w, "out"))
        synth.make_eam_tables_binary(spec, w)
        synth.write_param_file(os.path.join(w, "run.param"), spec, (7, 5, 6), steps, "out/run", pbc=(0, 1, 1),
                               starttemp=0.08, nbl_margin=0.4, eng_int=1, coordname="_l12", masses=spec.masses,
                               extra="checkpt_int %d\n" % steps)
    ra, rb = run(ref, wa), run(cuda, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    ea, eb = eng_of(wa), eng_of(wb)
    assert ea.shape == eb.shape and ea.shape[0] >= steps
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    ca, cb = read_chkpt(chkpt_path(wa, 1)), read_chkpt(chkpt_path(wb, 1))
    assert ca.shape == cb.shape and np.array_equal(ca[:, :3], cb[:, :3])
    assert set(np.unique(ca[:, 1])) == {0.0, 1.0}                   # both species present
    scale = np.abs(ca[:, 3:]).max(axis=0) + 1e-300
    err_chk = (np.abs(ca[:, 3:] - cb[:, 3:]).max(axis=0) / scale).max()
    print("two species, %s: Epot rel %.2e, T rel %.2e, .chkpt of step %d worst column %.2e"
          % (mode, err_pot, err_kin, steps, err_chk))
    assert err_pot < 1e-12 and err_kin < 1e-13 and err_chk < 1e-11   # observed 5e-14, 4e-15, 4e-13


def eng_of_3point(tmp_path, steps):
    w = str(tmp_path / "ref3")
    os.makedirs(w)
    write_inputs(w, steps)
    r = run(REF_BIN, w)
    assert r.returncode == 0, r.stdout[-3000:]
    return eng_of(w)


@needs_bins
@pytest.mark.gpu
def test_chkpt_written_by_cuda_build_restarts_in_reference(tmp_path):
    """sync_to_host: the CUDA build writes run.1.chkpt at step 20 through IMD's own write_config;
    the file equals the reference's own checkpoint of that step, and the PLAIN reference restarted from it
    (-r 1) lands on the uninterrupted reference run."""
    steps, at = 40, 20
    extra = ("checkpt_int  %d\ndist_int %d\ndist_dim 5 4 3\ndist_Ekin_flag 1\ndist_Epot_flag 1\ndist_dens_flag 1\n"
             % (at, at))
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps, extra=extra)
    ra, rb = run(REF_BIN, wa), run(CUDA_BIN, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "distributions (60 bins) are summed on the device" in rb.stdout
    ca, cb = read_chkpt(chkpt_path(wa, 1)), read_chkpt(chkpt_path(wb, 1))
    assert ca.shape == cb.shape
    assert np.array_equal(ca[:, :3], cb[:, :3])                     # number, type, mass
    scale = np.abs(ca[:, 3:]).max(axis=0) + 1e-300
    err = (np.abs(ca[:, 3:] - cb[:, 3:]).max(axis=0) / scale).max()
    print("run.1.chkpt (step %d), CUDA build vs reference: max rel err over columns %.2e" % (at, err))
    assert err < 1e-8
    # write_distrib (imd_distrib.c:33-218) of the same steps.  wb: per-bin sums from the device (imdcu_distrib) in
    # the reference's file layout; wc: the same binary with IMD_CUDA_HOST_DISTRIB=1, i.e. the reference's own
    # write_distrib behind sync_to_host.  Binary float32 after the header; the count (density) is exact, the
    # averages are float sums in cell order in the reference and double sums rounded once on the device.
    wc = str(tmp_path / "cuda_hostdist")
    os.makedirs(wc)
    write_inputs(wc, steps, extra=extra)
    rc = run(CUDA_BIN, wc, env_extra={"IMD_CUDA_HOST_DISTRIB": "1"})
    assert rc.returncode == 0, rc.stdout[-3000:]
    assert "summed on the device" not in rc.stdout
    worst = 0.0
    for n in (0, 1, 2):
        for key, tol in (("dens", 0.0), ("Ekin", 1e-5), ("Epot", 1e-5)):
            fa, fb, fc = (np.frombuffer(open(os.path.join(w, "out", "run.%d.%s" % (n, key)), "rb").read().split(b"#E\n", 1)[1],
                                        dtype="<f4") for w in (wa, wb, wc))
            assert fa.shape == fb.shape == fc.shape == (60,)
            for f in (fb, fc):
                assert np.abs(fa - f).max() <= tol * np.abs(fa).max(), (n, key, np.abs(fa - f).max())
            worst = max(worst, np.abs(fa - fb).max() / np.abs(fa).max())
    for key in ("dens", "Ekin", "Epot"):                            # <outfiles>.minmax.<key>: one line per file
        ma, mb = (np.loadtxt(os.path.join(w, "out", "run.minmax.%s" % key)) for w in (wa, wb))
        assert ma.shape == mb.shape == (3, 3)
        assert np.allclose(ma, mb, rtol=1e-5, atol=0)
    print("device-side distributions vs the reference's files: worst rel err %.2e" % worst)
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


REF_FORCE_BIN = os.path.join(ROOT, "oracle", "_ref", "imd_ref_eam_nve_nbl_force")
CUDA_FORCE_BIN = os.path.join(ROOT, "integration", "_bin", "imd_eam_nve_nbl_force_cuda")


def read_force_file(path):
    """<outfiles>.<n>.force (write_atoms_force, imd_io.c:1950-1973): type x y z Fx Fy Fz per atom, in cell order."""
    rows = np.loadtxt(path, comments="#")
    return rows[:, 1:4], rows[:, 4:7]


@pytest.mark.skipif(not (os.path.exists(REF_FORCE_BIN) and os.path.exists(CUDA_FORCE_BIN)),
                    reason="the -DFORCE builds are missing (needs the reference sources)")
@pytest.mark.gpu
def test_configs0_eng_and_force_files_equal_reference(tmp_path):
    """BASELINE configs[0] through IMD's own files: Cu fcc, 32^3 cells = 131 072 atoms, 100 NVE steps, the
    reference's -DFORCE build against the same build with the CUDA hot path.  .eng row by row, and the .force files
    of steps 0, 50 and 100: write_atoms_force runs right after calc_forces (imd_main_3d.c:429-434), so the files
    show KRAFT as imdcu_sync_to_host stages it back.  The files carry no atom numbers and the two programs keep
    their cells in different orders: rows are matched by position (periodic nearest neighbour)."""
    from scipy.spatial import cKDTree
    steps = 100
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps, box=(32, 32, 32), extra="force_int 50\n")
    ra, rb = run(REF_FORCE_BIN, wa), run(CUDA_FORCE_BIN, wb)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "131072 atoms" in ra.stdout and "hot path on CUDA device" in rb.stdout
    ea, eb = eng_of(wa), eng_of(wb)
    assert ea.shape == eb.shape and ea.shape[0] >= steps
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    box = 32 * synth.CU.a
    worst_x = worst_f = fmax = 0.0                                   # step 0 is the perfect lattice: forces ~1e-15
    for n in (0, 1, 2):
        xa, fa = read_force_file(os.path.join(wa, "out", "run.%05d.force" % n))
        xb, fb = read_force_file(os.path.join(wb, "out", "run.%05d.force" % n))
        assert xa.shape == xb.shape == (131072, 3)
        dist, perm = cKDTree(np.mod(xb, box), boxsize=box).query(np.mod(xa, box))
        assert np.array_equal(np.sort(perm), np.arange(131072))          # a one-to-one match
        worst_x = max(worst_x, dist.max())
        worst_f, fmax = max(worst_f, np.abs(fa - fb[perm]).max()), max(fmax, np.abs(fa).max())
    worst_f /= fmax
    print("configs[0], 131 072 Cu atoms, 100 steps: .eng Epot rel %.2e, T rel %.2e; .force files: positions %.2e A, "
          "forces %.2e of the largest component (%.2f eV/A)" % (err_pot, err_kin, worst_x, worst_f, fmax))
    assert err_pot < 3e-10 and err_kin < 1e-12                       # observed 2.3e-11, 2.7e-14
    assert worst_x < 1e-12 and worst_f < 2e-12                       # observed 5.7e-14 A, 1.4e-13


# ----------------------------------------------------------------------------- MPI hosts (one rank per GPU)
REF_MPI_TTM = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm")
CUDA_MPI_NVE = os.path.join(ROOT, "integration", "_bin", "imd_mpi_eam_nve_nbl_cuda")
CUDA_MPI_TTM = os.path.join(ROOT, "integration", "_bin", "imd_mpi_eam_ttm_tmm_nbl_cuda")
needs_mpi_bins = pytest.mark.skipif(not all(os.path.exists(b) for b in (REF_BIN, REF_MPI_TTM, CUDA_MPI_NVE, CUDA_MPI_TTM)),
                                    reason="integration/_bin or oracle/_ref not built (needs the reference sources)")


def run_mpi(binary, work, nranks, env_extra=None, *args):
    """The MPI builds run over oracle/shmpi here (this image has no MPI): SHMPI_NP ranks are forked by MPI_Init."""
    env = dict(os.environ, SHMPI_NP=str(nranks))
    env.pop("IMD_DUMP_DIR", None)
    env.update(env_extra or {})
    return subprocess.run([binary, "-p", "run.param", *args], cwd=work, env=env, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=600)


@needs_mpi_bins
@pytest.mark.gpu
def test_mpi_cuda_binary_on_one_rank_equals_serial_reference(tmp_path):
    """imd_mpi_eam_nve_nbl_cuda (MPI host: imd_mpi_util.c, cpugrid, the NCCL hand-shake skipped for one rank)
    on one rank against the serial reference: same .eng."""
    steps = 30
    wa, wb = str(tmp_path / "ref"), str(tmp_path / "cuda")
    for w in (wa, wb):
        os.makedirs(w)
        write_inputs(w, steps, box=(7, 6, 6))
    ra, rb = run(REF_BIN, wa), run_mpi(CUDA_MPI_NVE, wb, 1)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    ea, eb = eng_of(wa), eng_of(wb)
    assert ea.shape == eb.shape
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    print("imd_mpi_eam_nve_nbl_cuda (1 rank) vs serial reference: Epot rel %.2e, T rel %.2e" % (err_pot, err_kin))
    assert err_pot < 1e-9 and err_kin < 1e-8


def read_dist_text(path):
    """A distribution file in one of the ASCII formats (write_distrib_header, imd_distrib.c:805): the header lines
    start with '#'; rows are [ix iy iz] value."""
    rows = [ln.split() for ln in open(path) if ln.strip() and not ln.startswith("#")]
    return np.array(rows, dtype=np.float64)


def compare_text_dists(wa, wb, numbers, bins):
    """dist_vxavg_flag 2 (ASCII with bin coordinates) and dist_dens_flag 3 (plain ASCII), printed with %e."""
    worst = 0.0
    for n in numbers:
        va, vb = (read_dist_text(os.path.join(w, "out", "run.%d.vxavg" % n)) for w in (wa, wb))
        assert va.shape == vb.shape == (bins, 4)
        assert np.array_equal(va[:, :3], vb[:, :3])
        assert np.abs(va[:, 3] - vb[:, 3]).max() <= 1e-5 * np.abs(va[:, 3]).max(), n
        worst = max(worst, np.abs(va[:, 3] - vb[:, 3]).max() / np.abs(va[:, 3]).max())
        da, db = (read_dist_text(os.path.join(w, "out", "run.%d.dens" % n)) for w in (wa, wb))
        assert da.shape == db.shape == (bins, 1)
        assert np.array_equal(da, db) and da.sum() > 0
    return worst


DIST_TTM = "dist_int 6\ndist_dim 14 2 1\ndist_vxavg_flag 2\ndist_dens_flag 3\n"


def read_ttm(path):
    """<outfiles>.<n>.ttm (ttm_writeout, imd_ttm.c:1683-1969): one line per FD cell, then the virtual lattice."""
    rows = [ln.split() for ln in open(path) if ln.strip() and not ln.startswith("#")]
    return np.array(rows, dtype=np.float64)


@needs_mpi_bins
@pytest.mark.gpu
def test_ttm_cuda_binary_equals_reference_ttm_build(tmp_path):
    """imd_mpi_eam_ttm_tmm_nbl_cuda against the reference's own -DTTM -DTTM1D -DTMM -DPDECAY build (2 ranks over
    shmpi): the parameter files differ in cpu_dim only (2 1 1 there, 1 1 1 on one GPU with IMDCU_TTM_REF_RANKS=2,
    because the reference's 1-D lattice result depends on its rank count).  Both read the SAME
    start configuration (coordname file, box_from_header): the reference's generator draws the initial velocities
    per rank, so a run from `starttemp` on two ranks starts 2e-3 away in temperature from one on one rank (checked
    with the reference alone).  Compared: .eng with
    its E_el column (calc_ttm -> tot_elec_energy, imd_io.c:2345) and the .ttm files (ttm_writeout), which the
    reference prints with 7 digits."""
    from imd_b200 import synth as sy
    steps, ttm_int = 12, 6
    tp = sy.TtmParams(ttmdimx=16, fd_n_timesteps=20)
    work = str(tmp_path)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    pos, box = sy.fcc_lattice((56, 6, 6), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.0258),
                         sy.AL.mass, box)
    runs = {}
    for name, ranks in (("ref", 2), ("cuda", 1)):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        ex = tp.param_text() + "box_from_header 1\nshiftx_front 20\nshiftx_rear 20\nttm_int %d\n" % ttm_int + DIST_TTM
        sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (56, 6, 6), steps, "out/run", potdir="..",
                            pbc=(0, 1, 1), starttemp=0.0258, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                            cpu_dim=(ranks, 1, 1), coordname="../start.conf")
        runs[name] = w
    src = {"IMD_SRC_MU": "%.17g" % tp.src_mu, "IMD_SRC_ABS": "%.17g" % tp.src_abs}      # analytic source on both sides
    ra = run_mpi(REF_MPI_TTM, runs["ref"], 2, src)
    rb = run_mpi(CUDA_MPI_TTM, runs["cuda"], 1, dict(src, IMDCU_TTM_REF_RANKS="2"))
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "hot path on CUDA device" in rb.stdout
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape and ea.shape[1] >= 6
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    err_el = (np.abs(ea[:, 5] - eb[:, 5]) / np.abs(ea[:, 5])).max()
    print("imd_mpi_eam_ttm_tmm_nbl_cuda vs reference TTM build: Epot rel %.2e, T rel %.2e, E_el rel %.2e (7 digits printed)"
          % (err_pot, err_kin, err_el))
    assert err_pot < 1e-8 and err_kin < 1e-7 and err_el < 2e-6
    cols = "x y z natoms temp md_temp U xi source dens vx vy vz fd_k fd_g Z proc Ce".split()
    worst = np.zeros(len(cols))
    where = [""] * len(cols)
    for n in range(steps // ttm_int + 1):
        ta = read_ttm(os.path.join(runs["ref"], "out", "run.%d.ttm" % n))
        tb = read_ttm(os.path.join(runs["cuda"], "out", "run.%d.ttm" % n))
        assert ta.shape == tb.shape, (n, ta.shape, tb.shape)
        assert np.array_equal(ta[:, [0, 1, 2, 3, 16]], tb[:, [0, 1, 2, 3, 16]]), n        # x y z natoms proc
        if n == 0:
            # The reference sends the virtual lattice from its owner to rank 0 only `if(steps>0)` (imd_ttm.c:1828):
            # in the step-0 file rank 0 prints its own copy, whose fd_k, fd_g, Z, Ce it never computes when another
            # rank owns the virtual lattice (here rank 1) -- uninitialised memory (fd_k 1.6e-153 was seen).  Rows of
            # the virtual lattice (natoms = -2) of file 0 are therefore not comparable; the FD cells are.
            keep = ta[:, 3] != -2
            assert keep.sum() == tp.ttmdimx
            ta, tb = ta[keep], tb[keep]
        num = np.abs(ta - tb)
        rel = np.where(np.abs(ta) > 1e-12, num / np.maximum(np.abs(ta), 1e-30), num)
        # v_com of a cell is a difference of O(1e-3) momenta sums: compare it on the scale of the column
        for col in (10, 11, 12):
            rel[:, col] = num[:, col] / max(np.abs(ta[:, col]).max(), 1e-30)
        for col in range(len(cols)):
            r = int(np.argmax(rel[:, col]))
            if rel[r, col] > worst[col]:
                worst[col] = rel[r, col]
                where[col] = "file %d row %d: reference %.6e, CUDA build %.6e" % (n, r, ta[r, col], tb[r, col])
    for col in range(len(cols)):
        if worst[col] > 0:
            print("  .ttm column %-8s worst relative difference %.2e  (%s)" % (cols[col], worst[col], where[col]))
    print(".ttm files 0..%d: worst relative difference %.2e" % (steps // ttm_int, worst.max()))
    assert worst.max() < 5e-6
    assert "distributions (28 bins) are summed on the device" in rb.stdout
    wd = compare_text_dists(runs["ref"], runs["cuda"], (0, 1, 2), 28)
    print(".vxavg / .dens files (ASCII formats) vs the reference's: worst relative difference %.2e" % wd)


@needs_mpi_bins
@pytest.mark.gpu
def test_ttm_cuda_binary_hot_film_with_pdecay_equals_reference(tmp_path):
    """Hot lattice + strong pulse + the -DPDECAY rear absorber (xipdecay, ramp_start, ramp_end: imd_integrate.c:
    690-705 in move_atoms_ttm) through the files: atoms cross FD slabs (advection fluxes), cells deactivate.  The
    film of the library-level golden ttm_al_48x5x5_p2_hot; tolerances are those of that golden's comparison."""
    from imd_b200 import synth as sy
    steps, ttm_int = 30, 15
    tp = sy.TtmParams(ttmdimx=12, fd_n_timesteps=30, I0=4.0e16, fd_min_atoms=20, vlatbuffer=1)
    work = str(tmp_path)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    pos, box = sy.fcc_lattice((48, 5, 5), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.12),
                         sy.AL.mass, box)
    runs = {}
    for name, ranks in (("ref", 2), ("cuda", 1)):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        ex = (tp.param_text() + "box_from_header 1\nshiftx_front 25\nshiftx_rear 15\nttm_int %d\n"
              "xipdecay 0.5\nramp_start 180.0\nramp_end 200.0\n" % ttm_int)
        sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (48, 5, 5), steps, "out/run", potdir="..",
                            pbc=(0, 1, 1), starttemp=0.12, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                            cpu_dim=(ranks, 1, 1), coordname="../start.conf")
        runs[name] = w
    src = {"IMD_SRC_MU": "%.17g" % tp.src_mu, "IMD_SRC_ABS": "%.17g" % tp.src_abs}
    ra = run_mpi(REF_MPI_TTM, runs["ref"], 2, src)
    rb = run_mpi(CUDA_MPI_TTM, runs["cuda"], 1, dict(src, IMDCU_TTM_REF_RANKS="2"))
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape and ea.shape[1] >= 6
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    err_el = (np.abs(ea[:, 5] - eb[:, 5]) / np.abs(ea[:, 5])).max()
    worst = 0.0
    for n in range(1, steps // ttm_int + 1):
        ta = read_ttm(os.path.join(runs["ref"], "out", "run.%d.ttm" % n))
        tb = read_ttm(os.path.join(runs["cuda"], "out", "run.%d.ttm" % n))
        assert ta.shape == tb.shape, (n, ta.shape, tb.shape)
        assert np.array_equal(ta[:, [0, 1, 2, 3, 16]], tb[:, [0, 1, 2, 3, 16]]), n
        num = np.abs(ta - tb)
        rel = np.where(np.abs(ta) > 1e-12, num / np.maximum(np.abs(ta), 1e-30), num)
        for col in (10, 11, 12):
            rel[:, col] = num[:, col] / max(np.abs(ta[:, col]).max(), 1e-30)
        worst = max(worst, rel.max())
    print("hot film with the rear absorber through the files: Epot rel %.2e, T rel %.2e, E_el rel %.2e, .ttm worst %.2e"
          % (err_pot, err_kin, err_el, worst))
    # observed 2e-13, 4e-12, 1.4e-7 (E_el is printed with seven digits) and equal .ttm files
    assert err_pot < 1e-11 and err_kin < 1e-10 and err_el < 2e-6 and worst < 2e-6


REF_MPI_TMM = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_tmm")


@pytest.mark.skipif(not (os.path.exists(REF_MPI_TMM) and os.path.exists(CUDA_MPI_TTM)),
                    reason="integration/_bin or oracle/_ref not built (needs the reference sources)")
@pytest.mark.gpu
def test_ttm_cuda_binary_with_do_tmm_equals_reference_tmm_build(tmp_path):
    """The laser source by the transfer-matrix solver, through the files: the reference build that keeps its OWN
    tmm_init / do_tmm (imd_tmm.c:98-696; 2 ranks over shmpi) against imd_mpi_eam_ttm_tmm_nbl_cuda, whose tmm_init
    reads ../alu_eps_bb.dat and ../K12.dat itself and hands them to imdcu_ttm_set_tmm (do_tmm then runs inside the
    device lattice kernel).  Same film as the library-level golden ttm_tmm_al_96x5x5_p2; .eng with its E_el column
    and the .ttm files (source column included) are compared."""
    from imd_b200 import synth as sy
    steps, ttm_int = 14, 7
    tp = sy.TtmParams(ttmdimx=26, fd_n_timesteps=20, I0=2.0e16, tmm_threshold=5.0, use_tmm=1)
    work = str(tmp_path)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    sy.write_tmm_files(sy.make_tmm_tables(), work)
    pos, box = sy.fcc_lattice((96, 5, 5), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.0258),
                         sy.AL.mass, box)
    runs = {}
    for name, ranks in (("ref", 2), ("cuda", 1)):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        ex = tp.param_text() + "box_from_header 1\nshiftx_front 20\nshiftx_rear 20\nttm_int %d\n" % ttm_int
        sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (96, 5, 5), steps, "out/run", potdir="..",
                            pbc=(0, 1, 1), starttemp=0.0258, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                            cpu_dim=(ranks, 1, 1), coordname="../start.conf")
        runs[name] = w
    ra = run_mpi(REF_MPI_TMM, runs["ref"], 2)
    rb = run_mpi(CUDA_MPI_TTM, runs["cuda"], 1, {"IMDCU_TTM_REF_RANKS": "2"})
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "transfer-matrix method (do_tmm on the device)" in rb.stdout
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape and ea.shape[1] >= 6
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    err_el = (np.abs(ea[:, 5] - eb[:, 5]) / np.abs(ea[:, 5])).max()
    assert ea[-1, 5] > 1.5 * ea[0, 5]                               # the pulse has heated the electrons
    worst, src_max = 0.0, 0.0
    for n in range(1, steps // ttm_int + 1):                        # file 0: rank 0's uninitialised virtual lattice
        ta = read_ttm(os.path.join(runs["ref"], "out", "run.%d.ttm" % n))
        tb = read_ttm(os.path.join(runs["cuda"], "out", "run.%d.ttm" % n))
        assert ta.shape == tb.shape, (n, ta.shape, tb.shape)
        assert np.array_equal(ta[:, [0, 1, 2, 3, 16]], tb[:, [0, 1, 2, 3, 16]]), n
        num = np.abs(ta - tb)
        rel = np.where(np.abs(ta) > 1e-12, num / np.maximum(np.abs(ta), 1e-30), num)
        for col in (8, 10, 11, 12):                                 # source, vx vy vz: against the column's largest entry
            rel[:, col] = num[:, col] / max(np.abs(ta[:, col]).max(), 1e-30)
        worst, src_max = max(worst, rel.max()), max(src_max, np.abs(ta[:, 8]).max())
    print("do_tmm through the files: Epot rel %.2e, T rel %.2e, E_el rel %.2e; .ttm files 1..%d worst relative "
          "difference %.2e (largest source term %.3e)" % (err_pot, err_kin, err_el, steps // ttm_int, worst, src_max))
    assert src_max > 0
    # observed 2e-14, 2e-12, 0 and 3.6e-7 (the .ttm files carry seven digits: one unit of the last is up to 1e-6)
    assert err_pot < 1e-11 and err_kin < 1e-10 and err_el < 2e-6 and worst < 2e-6


REF_MPI_NRB = os.path.join(ROOT, "oracle", "_ref", "imd_ref_mpi_eam_ttm_nrb")
CUDA_MPI_NRB = os.path.join(ROOT, "integration", "_bin", "imd_mpi_eam_ttm_tmm_nrb_nbl_cuda")
needs_nrb_bins = pytest.mark.skipif(not (os.path.exists(REF_MPI_NRB) and os.path.exists(CUDA_MPI_NRB)),
                                    reason="the -DNRB builds are missing (needs the reference sources)")


def _nrb_film_runs(work, ranks_cuda):
    """BASELINE configs[3] in small: Al film + TTM + non-reflecting rear boundary (nrb_k 1.2), warm lattice and a
    pulse so that the boundary plane moves; the film of the library-level golden ttm_nrb_al_40x5x5_p2."""
    from imd_b200 import synth as sy
    steps, ttm_int = 24, 12
    tp = sy.TtmParams(ttmdimx=10, fd_n_timesteps=20, I0=2.0e16)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    pos, box = sy.fcc_lattice((40, 5, 5), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.05),
                         sy.AL.mass, box)
    runs = {}
    for name, ranks in (("ref", 2), ("cuda", ranks_cuda)):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        ex = (tp.param_text() + "box_from_header 1\nshiftx_front 20\nshiftx_rear 20\nttm_int %d\ncheckpt_int %d\n"
              "nrb_alat 4.05\nnrb_eps 0.3\nnrb_k 1.2\n" % (ttm_int, ttm_int))
        sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (40, 5, 5), steps, "out/run", potdir="..",
                            pbc=(0, 1, 1), starttemp=0.05, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                            cpu_dim=(ranks, 1, 1), coordname="../start.conf")
        runs[name] = w
    src = {"IMD_SRC_MU": "%.17g" % tp.src_mu, "IMD_SRC_ABS": "%.17g" % tp.src_abs}
    ra = run_mpi(REF_MPI_NRB, runs["ref"], 2, src)
    rb = run_mpi(CUDA_MPI_NRB, runs["cuda"], ranks_cuda, dict(src, IMDCU_TTM_REF_RANKS="2") if ranks_cuda == 1 else src)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "non-reflecting boundary on the device" in rb.stdout
    return runs, steps, ttm_int, pos.shape[0]


def _compare_nrb_runs(runs, steps, ttm_int, natoms, label):
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape and ea.shape[1] >= 6
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    err_el = (np.abs(ea[:, 5] - eb[:, 5]) / np.abs(ea[:, 5])).max()
    worst_ttm = worst_chk = 0.0
    for n in range(1, steps // ttm_int + 1):
        ta = read_ttm(os.path.join(runs["ref"], "out", "run.%d.ttm" % n))
        tb = read_ttm(os.path.join(runs["cuda"], "out", "run.%d.ttm" % n))
        assert ta.shape == tb.shape, (n, ta.shape, tb.shape)
        assert np.array_equal(ta[:, [0, 1, 2, 3, 16]], tb[:, [0, 1, 2, 3, 16]]), n
        num = np.abs(ta - tb)
        rel = np.where(np.abs(ta) > 1e-12, num / np.maximum(np.abs(ta), 1e-30), num)
        for col in (10, 11, 12):
            rel[:, col] = num[:, col] / max(np.abs(ta[:, col]).max(), 1e-30)
        worst_ttm = max(worst_ttm, rel.max())
        # .chkpt through imdcu_sync_to_host and the .nrb restart file beside it (nrb_writerestart, imd_nrb.c:2143-2214:
        # number, NRBBND, NRBNEIGH, position, NRBI[12], REF_POS of every boundary and neighbour atom)
        ca, cb = read_chkpt(chkpt_path(runs["ref"], n)), read_chkpt(chkpt_path(runs["cuda"], n))
        assert ca.shape == cb.shape and ca.shape[0] == natoms
        assert np.array_equal(ca[:, :3], cb[:, :3])
        scale = np.abs(ca[:, 3:]).max(axis=0) + 1e-300
        worst_chk = max(worst_chk, (np.abs(ca[:, 3:] - cb[:, 3:]).max(axis=0) / scale).max())
        na, nb = (np.loadtxt(os.path.join(runs[k], "out", "run.%05d.nrb" % n)) for k in ("ref", "cuda"))
        na, nb = na[np.argsort(na[:, 0])], nb[np.argsort(nb[:, 0])]
        assert na.shape == nb.shape and na.shape[0] > 0
        assert np.array_equal(na[:, [0, 1, 2] + list(range(6, 18))], nb[:, [0, 1, 2] + list(range(6, 18))]), n
        assert np.abs(na[:, 3:6] - nb[:, 3:6]).max() < 2e-6 and np.array_equal(na[:, 18:], nb[:, 18:]), n   # %lf: 6 decimals
    print("%s: Epot rel %.2e, T rel %.2e, E_el rel %.2e; .ttm worst %.2e; .chkpt worst %.2e; .nrb files equal"
          % (label, err_pot, err_kin, err_el, worst_ttm, worst_chk))
    assert err_pot < 1e-8 and err_kin < 1e-7 and err_el < 2e-6 and worst_ttm < 5e-6 and worst_chk < 1e-8


@needs_nrb_bins
@pytest.mark.gpu
def test_nrb_cuda_binary_equals_reference_nrb_build(tmp_path):
    """The -DNRB -DREFPOS build with the CUDA hot path (init_nrb on the host once, nrb_forces and the integrator hooks
    on the device) against the reference's own NRB build on two ranks: .eng, .ttm, .chkpt and .nrb files."""
    runs, steps, ttm_int, natoms = _nrb_film_runs(str(tmp_path), 1)
    _compare_nrb_runs(runs, steps, ttm_int, natoms, "imd_mpi_eam_ttm_tmm_nrb_nbl_cuda vs the reference NRB build")


def _ngpu():
    import torch
    return torch.cuda.device_count()


@needs_mpi_bins
@pytest.mark.gpu
def test_mpi_cuda_binary_on_two_gpus_equals_serial_reference(tmp_path):
    """imd_mpi_eam_nve_nbl_cuda on TWO ranks, one GPU each (cpu_dim 2 1 1: IMD distributes the atoms, the stub hands
    each rank's cell_array to its device, x slabs over NCCL with the unique id broadcast by MPI_Bcast) against the
    serial reference started from the same configuration file."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    steps = 30
    work = str(tmp_path)
    synth.make_eam_tables(synth.CU, work)
    pos, box = synth.fcc_lattice((12, 6, 6), synth.CU.a)
    synth.write_config_file(os.path.join(work, "start.conf"), pos, synth.maxwell_momenta(pos.shape[0], synth.CU.mass, 0.1),
                            synth.CU.mass, box)
    runs = {}
    for name, ranks in (("ref", 1), ("cuda", 2)):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        synth.write_param_file(os.path.join(w, "run.param"), synth.CU, (12, 6, 6), steps, "out/run", potdir="..",
                               pbc=(1, 1, 1), starttemp=0.1, nbl_margin=0.5, eng_int=1,
                               extra="box_from_header 1\ncheckpt_int 15\n", cpu_dim=(ranks, 1, 1), coordname="../start.conf")
        runs[name] = w
    ra, rb = run(REF_BIN, runs["ref"]), run_mpi(CUDA_MPI_NVE, runs["cuda"], 2)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "x slabs over NCCL" in rb.stdout
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    print("imd_mpi_eam_nve_nbl_cuda on 2 GPUs vs serial reference: Epot rel %.2e, T rel %.2e" % (err_pot, err_kin))
    assert err_pot < 1e-9 and err_kin < 1e-8
    # .chkpt of steps 15 and 30: every rank refills its cell_array (imdcu_sync_to_host) and the reference's write_config
    # collects the ranks' atoms (imd_io_3d.c:721-979); compared by atom number with the serial reference's files
    worst = 0.0
    for n in (1, 2):
        ca, cb = read_chkpt(chkpt_path(runs["ref"], n)), read_chkpt(chkpt_path(runs["cuda"], n))
        assert ca.shape == cb.shape and ca.shape[0] == pos.shape[0]
        assert np.array_equal(ca[:, :3], cb[:, :3])
        d = np.abs(ca[:, 3:6] - cb[:, 3:6])
        d = np.minimum(d, np.abs(d - np.diag(box)[None, :]))             # a position on the periodic boundary
        scale = np.abs(ca[:, 6:]).max(axis=0) + 1e-300
        worst = max(worst, d.max() / box.max(), (np.abs(ca[:, 6:] - cb[:, 6:]).max(axis=0) / scale).max())
    print(".chkpt files of two ranks / two GPUs vs the serial reference's: worst rel err %.2e" % worst)
    assert worst < 1e-12                                                 # observed 1.2e-14


@needs_mpi_bins
@pytest.mark.gpu
def test_ttm_cuda_binary_on_two_gpus_equals_reference_ttm_build(tmp_path):
    """imd_mpi_eam_ttm_tmm_nbl_cuda on two ranks / two GPUs against the reference's TTM build on two ranks: the SAME
    parameter file for both (cpu_dim 2 1 1), .eng and .ttm files compared as in the one-GPU test."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    from imd_b200 import synth as sy
    steps, ttm_int = 12, 6
    tp = sy.TtmParams(ttmdimx=16, fd_n_timesteps=20)
    work = str(tmp_path)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    pos, box = sy.fcc_lattice((56, 6, 6), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.0258),
                         sy.AL.mass, box)
    runs = {}
    for name in ("ref", "cuda"):
        w = os.path.join(work, name)
        os.makedirs(os.path.join(w, "out"))
        ex = tp.param_text() + "box_from_header 1\nshiftx_front 20\nshiftx_rear 20\nttm_int %d\n" % ttm_int + DIST_TTM
        sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (56, 6, 6), steps, "out/run", potdir="..",
                            pbc=(0, 1, 1), starttemp=0.0258, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                            cpu_dim=(2, 1, 1), coordname="../start.conf")
        runs[name] = w
    src = {"IMD_SRC_MU": "%.17g" % tp.src_mu, "IMD_SRC_ABS": "%.17g" % tp.src_abs}
    ra, rb = run_mpi(REF_MPI_TTM, runs["ref"], 2, src), run_mpi(CUDA_MPI_TTM, runs["cuda"], 2, src)
    assert ra.returncode == 0, ra.stdout[-3000:]
    assert rb.returncode == 0, rb.stdout[-3000:]
    assert "x slabs over NCCL" in rb.stdout
    ea, eb = eng_of(runs["ref"]), eng_of(runs["cuda"])
    assert ea.shape == eb.shape and ea.shape[1] >= 6
    err_pot = np.abs(ea[:, 1] - eb[:, 1]).max() / np.abs(ea[:, 1]).max()
    err_kin = np.abs(ea[:, 2] - eb[:, 2]).max() / np.abs(ea[:, 2]).max()
    err_el = (np.abs(ea[:, 5] - eb[:, 5]) / np.abs(ea[:, 5])).max()
    print("imd_mpi_eam_ttm_tmm_nbl_cuda on 2 GPUs vs reference TTM build: Epot rel %.2e, T rel %.2e, E_el rel %.2e"
          % (err_pot, err_kin, err_el))
    assert err_pot < 1e-8 and err_kin < 1e-7 and err_el < 2e-6
    worst = 0.0
    for n in range(1, steps // ttm_int + 1):                     # file 0: rank 0's uninitialised virtual lattice (see above)
        ta = read_ttm(os.path.join(runs["ref"], "out", "run.%d.ttm" % n))
        tb = read_ttm(os.path.join(runs["cuda"], "out", "run.%d.ttm" % n))
        assert ta.shape == tb.shape, (n, ta.shape, tb.shape)
        assert np.array_equal(ta[:, [0, 1, 2, 3, 16]], tb[:, [0, 1, 2, 3, 16]]), n
        num = np.abs(ta - tb)
        rel = np.where(np.abs(ta) > 1e-12, num / np.maximum(np.abs(ta), 1e-30), num)
        for col in (10, 11, 12):
            rel[:, col] = num[:, col] / max(np.abs(ta[:, col]).max(), 1e-30)
        worst = max(worst, rel.max())
    print(".ttm files 1..%d on 2 GPUs: worst relative difference %.2e" % (steps // ttm_int, worst))
    assert worst < 5e-6
    assert "distributions (28 bins) are summed on the device" in rb.stdout
    wd = compare_text_dists(runs["ref"], runs["cuda"], (0, 1, 2), 28)          # per-rank device sums + MPI_Reduce
    print(".vxavg / .dens files on 2 GPUs vs the reference's: worst relative difference %.2e" % wd)


@needs_nrb_bins
@pytest.mark.gpu
def test_nrb_cuda_binary_on_two_gpus_equals_reference_nrb_build(tmp_path):
    """The same film on two ranks / two GPUs (the NRB table is gathered over the ranks; the boundary lives on the
    last rank): BASELINE configs[3] in small, through IMD's own files."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    runs, steps, ttm_int, natoms = _nrb_film_runs(str(tmp_path), 2)
    _compare_nrb_runs(runs, steps, ttm_int, natoms, "imd_mpi_eam_ttm_tmm_nrb_nbl_cuda on 2 GPUs vs the reference NRB build")


@needs_nrb_bins
@pytest.mark.gpu
def test_nrb_reference_positions_follow_atoms_across_ranks(tmp_path):
    """-DREFPOS puts REF_POS of every atom into the checkpoint.  The CUDA build keeps them on the host of the rank
    that owned the atom when init_nrb set them; an atom that has moved to the other slab since is asked for over MPI
    (imd_cuda.c: cu_nrb_resolve).  Film shifted so that an atomic plane lies ON the cut between the two slabs: half
    of its atoms change rank at the first list rebuilds.  Expected refpos: the start configuration, shifted."""
    if _ngpu() < 2:
        pytest.skip("needs 2 GPUs")
    from imd_b200 import synth as sy
    steps = 40
    front = 101.0 - 1.0125 - 39 * 2.025                   # plane 39 at x = 101.0 = 15 of 30 cell layers of the 202 A box
    tp = sy.TtmParams(ttmdimx=10, fd_n_timesteps=20, I0=2.0e16)
    work = str(tmp_path)
    sy.make_eam_tables(sy.AL, work)
    sy.write_eos_files(sy.make_eos_tables(), work)
    pos, box = sy.fcc_lattice((40, 5, 5), sy.AL.a)
    sy.write_config_file(os.path.join(work, "start.conf"), pos, sy.maxwell_momenta(pos.shape[0], sy.AL.mass, 0.05),
                         sy.AL.mass, box)
    w = os.path.join(work, "cuda")
    os.makedirs(os.path.join(w, "out"))
    ex = (tp.param_text() + "box_from_header 1\nshiftx_front %.17g\nshiftx_rear %.17g\nttm_int 0\ncheckpt_int %d\n"
          "nrb_alat 4.05\nnrb_eps 0.3\nnrb_k 1.2\n" % (front, 40.0 - front, steps))
    sy.write_param_file(os.path.join(w, "run.param"), sy.AL, (40, 5, 5), steps, "out/run", potdir="..",
                        pbc=(0, 1, 1), starttemp=0.05, nbl_margin=0.4, eng_int=1, extra=ex, ensemble="ttm",
                        cpu_dim=(2, 1, 1), coordname="../start.conf")
    r = run_mpi(CUDA_MPI_NRB, w, 2, {"IMD_SRC_MU": "%.17g" % tp.src_mu, "IMD_SRC_ABS": "%.17g" % tp.src_abs})
    assert r.returncode == 0, r.stdout[-3000:]
    assert "changed their rank are fetched from their first owner" in r.stdout, r.stdout[-3000:]
    c = read_chkpt(chkpt_path(w, 1))                       # number type mass x y z vx vy vz Epot refpos_x refpos_y refpos_z eam_rho
    assert c.shape[0] == pos.shape[0] and np.array_equal(c[:, 0], np.arange(1, pos.shape[0] + 1))
    want = pos + np.array([front, 0.0, 0.0])
    err = np.abs(c[:, 10:13] - want).max()
    moved = np.abs(c[:, 3:6] - want).max()
    print("refpos of %d atoms after %d steps on two ranks: max |refpos - start| %.2e A (atoms moved up to %.2f A)"
          % (c.shape[0], steps, err, moved))
    assert err < 1e-12 and moved > 0.1
