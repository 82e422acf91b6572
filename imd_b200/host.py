"""Host-side mirror of the reference's call sites for the hot path.

The reference drives the path through global-state functions:
``calc_forces(steps)`` (imd_forces_nbl.c:295), ``move_atoms()`` -> ``move_atoms_nve``
(imd_integrate.c:35), ``check_nblist()`` (imd_forces_nbl.c:1062), with results in the
globals ``tot_pot_energy, virial, tot_kin_energy, have_valid_nbl, nbl_count``.  ``Imd``
keeps that vocabulary over one C-ABI context so the parity tests read like a main loop
(imd_main_3d.c:151-916).  All arithmetic happens in libimdcu.so on the GPU.
"""
from __future__ import annotations

import numpy as np

from . import capi, synth

TAB_PAIR, TAB_RHO, TAB_EMBED = 0, 1, 2
capi_EOS_CVE, capi_EOS_EE = 0, 1


def x_layer_of_atoms(pos, box, pbc_x, gdim_x):
    """Global x cell layer of every atom by cell_coord (imd_geom_3d.c:1077-1097)."""
    tbox_x = np.linalg.inv(np.asarray(box, np.float64).reshape(3, 3))[:, 0]
    s = np.asarray(pos, np.float64) @ tbox_x
    if pbc_x:
        s = s - np.floor(s)
    g = int(gdim_x)
    return np.clip((g * s).astype(np.int64), 0, g - 1)


def slab_rank_of_atoms(pos, box, pbc_x, gdim_x, nranks, xcut=None):
    """Owner rank of every atom for the x-slab decomposition cpu_dim = nranks 1 1: rank r owns the global cell
    layers [xcut[r], xcut[r+1]); without cuts the reference's equal slabs [r*l, (r+1)*l), l = gdim_x/nranks
    (imd_geom_mpi_3d.c:119-128).  Host-side helper: which atoms a rank uploads."""
    g = int(gdim_x)
    if xcut is None:
        if g % nranks != 0:
            raise ValueError("global_cell_dim.x %d is not a multiple of the number of ranks %d" % (g, nranks))
        xcut = np.arange(nranks + 1) * (g // nranks)
    cx = x_layer_of_atoms(pos, box, pbc_x, g)
    return (np.searchsorted(np.asarray(xcut), cx, side="right") - 1).astype(np.int32)


class Imd:
    def __init__(self, device: int = 0, comm=None):
        """comm = (ncclUniqueId bytes, rank, nranks): x-slab decomposition over nranks GPUs, the
        reference's cpu_dim = nranks 1 1 (one process per GPU, like one MPI rank per core)."""
        self.ctx = capi.Context(device)
        self.rank, self.nranks = 0, 1
        if comm is not None:
            uid, self.rank, self.nranks = comm
            self.ctx.comm_init(uid, self.rank, self.nranks)
        self.tot_pot_energy = 0.0
        self.virial = 0.0
        self.tot_kin_energy = 0.0
        self.have_valid_nbl = 0
        self.steps = 0

    # setup_potentials (imd_potential.c:43): pair_pot, rho_h_tab, embed_pot
    def setup_potentials(self, tables):
        # the reference's compile-time FOURPOINT / SPLINE (read_pot_table finishes the tables accordingly,
        # imd_potential.c:268-275) is a property of the tables handed in
        self.ctx.set_interp(getattr(tables["pair"], "interp", "3point"))
        self.ctx.set_table(TAB_PAIR, tables["pair"])
        self.ctx.set_table(TAB_RHO, tables["rho"])
        self.ctx.set_table(TAB_EMBED, tables["emb"])

    # make_box + init_cells (imd_geom_3d.c:52, 118)
    def make_box(self, box, pbc_dirs, nbl_margin, nbl_size=1.3, cell_divisor=None):
        self.ctx.set_box(box, pbc_dirs, nbl_margin, nbl_size, cell_divisor)
        self.global_cell_dim, self.cellsz = self.ctx.cell_grid()
        self.box = np.asarray(box, np.float64).reshape(3, 3)
        self.pbc_dirs = tuple(int(p) for p in pbc_dirs)
        self.slab = self.ctx.comm_rank()

    def owned_mask(self, pos):
        """Which atoms lie in this rank's x slab (cell_coord, imd_geom_3d.c:1077-1097; an atom the
        host puts one slab off migrates at the first rebuild)."""
        if self.nranks == 1:
            return np.ones(pos.shape[0], bool)
        cx = x_layer_of_atoms(pos, self.box, self.pbc_dirs[0], self.global_cell_dim[0])
        return (cx >= self.slab["xoff"]) & (cx < self.slab["xoff"] + self.slab["ldimx"])

    def set_cuts(self, xcut):
        """Slabs of equal atom count (imdcu_comm_set_cuts): after make_box, before set_atoms."""
        self.ctx.comm_set_cuts(xcut)
        self.slab = self.ctx.comm_rank()

    def set_params(self, ntypes=1, timestep=0.1, restrictions=None):
        self.ctx.set_params(ntypes, timestep, restrictions)

    # read_atoms / generate_atoms hand-over (imd.c:131-142)
    def set_atoms(self, pos, mom, mass, types=None, ids=None, vtypes=None):
        self.ctx.upload_atoms(pos, mom, mass, types, ids, vtypes)
        self.have_valid_nbl = 0

    @property
    def natoms(self):
        return self.ctx.n

    @property
    def nbl_count(self):
        return self.ctx.nbl_count

    def calc_forces(self, steps=None):
        out = self.ctx.calc_forces(self.steps if steps is None else steps)
        self.tot_pot_energy, self.virial = float(out[0]), float(out[1])
        self.have_valid_nbl = 1
        return out

    def move_atoms(self):
        """move_atoms_nve followed by check_nblist, as the main loop does (:564, :812)."""
        out, valid = self.ctx.move_atoms_nve()
        self.tot_kin_energy = float(out[0])
        self.have_valid_nbl = int(valid)
        return out

    def main_loop(self, nsteps):
        """calc_forces -> move_atoms -> check_nblist for nsteps; returns [nsteps,2] energies."""
        e = self.ctx.run_nve(self.steps, nsteps, True)
        self.steps += nsteps
        if nsteps:
            self.tot_pot_energy, self.tot_kin_energy = float(e[-1, 0]), float(e[-1, 1])
        return e

    # init_nrb (imd_forces_nbl.c:281-282, first calc_forces): NRB state to the device
    def set_nrb(self, nrb_k, state=None, startup_images=False):
        """state: imd_b200.nrb.NrbState (from nrb.init_nrb, or filled from the reference's .nrb restart
        data); None switches NRB off.  startup_images: see imdcu_set_nrb."""
        if state is None:
            e = np.zeros(0, np.int32)
            self.ctx.set_nrb(1.0, e, e, np.zeros((0, 12), np.int32), np.zeros((0, 3)))
            return
        ids, bnd, nrbi, refpos = state.entries()
        self.ctx.set_nrb(nrb_k, ids, bnd, nrbi, refpos, startup_images)

    # init_ttm + tmm_init (imd.c:237,245): EOS tables + TTM parameters
    def init_ttm(self, tp, eos, nranks_ref=1, tmm_tables=None):
        self.ctx.ttm_set_eos(capi_EOS_CVE, eos["cve"])
        self.ctx.ttm_set_eos(capi_EOS_EE, eos["ee"])
        if getattr(tp, "use_tmm", 0):
            # tmm_init (imd.c:245): the two data files of do_tmm and the phase table it looks up
            tabs = tmm_tables if tmm_tables is not None else synth.make_tmm_tables()
            self.ctx.ttm_set_eos(2, eos["phase"])
            self.ctx.ttm_set_tmm(tp.tmm_threshold, tabs["eps_bb"], tabs["K12"])
        self.ctx.ttm_init(tp, nranks_ref)

    # calc_ttm (imd_main_3d.c:534)
    def calc_ttm(self, steps=None):
        return self.ctx.calc_ttm(self.steps if steps is None else steps)

    def move_atoms_ttm(self):
        out, valid = self.ctx.move_atoms_ttm()
        self.tot_kin_energy = float(out[0])
        self.have_valid_nbl = int(valid)
        return out

    def main_loop_ttm(self, nsteps):
        """calc_forces -> calc_ttm -> move_atoms_ttm -> check_nblist for nsteps."""
        e = self.ctx.run_ttm(self.steps, nsteps, True)
        self.steps += nsteps
        if nsteps:
            self.tot_pot_energy, self.tot_kin_energy = float(e[-1, 0]), float(e[-1, 1])
        return e

    def ttm_lattice(self):
        return self.ctx.download_ttm()

    def atoms(self, fields=("id", "pos", "mom", "force", "epot", "rho", "dF", "numneighs", "cell"),
              sort_by_id=True):
        f = tuple(fields)
        if sort_by_id and "id" not in f:
            f = ("id",) + f
        d = self.ctx.download(f)
        if sort_by_id:
            order = np.argsort(d["id"], kind="stable")
            d = {k: v[order] for k, v in d.items()}
        return d

    def nblist_pairs(self):
        return self.ctx.nblist_pairs()

    def close(self):
        self.ctx.close()
