"""ctypes binding of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs, never by the product package ``imd_b200``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


class OrcTable(C.Structure):
    _fields_ = [("ncols", C.c_int), ("maxsteps", C.c_int),
                ("begin", c_dp), ("end", c_dp), ("step", c_dp), ("invstep", c_dp),
                ("len", c_ip), ("table", c_dp), ("table2", c_dp)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("eam_oracle.c", "ttm_oracle.c", "oracle.h", "ttm_oracle.h")]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_create.restype = C.c_void_p
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_setup.argtypes = [C.c_void_p, c_dp, c_ip, C.c_double, C.c_double, c_ip]
        L.orc_fix_cells.argtypes = [C.c_void_p, C.c_long, c_dp]
        L.orc_make_nblist.argtypes = [C.c_void_p]
        L.orc_calc_forces.argtypes = [C.c_void_p, c_dp, c_ip, C.c_int,
                                      C.POINTER(OrcTable), C.POINTER(OrcTable), C.POINTER(OrcTable),
                                      c_dp, c_dp, c_dp, c_dp, c_ip, c_dp]
        L.orc_move_atoms_nve.restype = C.c_double
        L.orc_move_atoms_nve.argtypes = [C.c_long, c_dp, c_dp, c_dp, c_dp, c_ip, c_dp, C.c_double,
                                         C.c_int, C.c_double, C.c_double, C.c_double]
        L.orc_max_displacement2.restype = C.c_double
        L.orc_max_displacement2.argtypes = [C.c_long, c_dp, c_dp]
        L.orc_export_pairs.restype = C.c_long
        L.orc_export_pairs.argtypes = [C.c_void_p, c_ip]
        L.orc_pair_int2.argtypes = [C.POINTER(OrcTable), C.c_int, C.c_int, C.c_double, c_dp, c_dp, c_ip]
        L.orc_pair_int.argtypes = [C.POINTER(OrcTable), C.c_int, C.c_int, C.c_double, c_dp, c_dp, c_ip]
        L.orc_set_interp.argtypes = [C.c_int]
        L.orc_init_fourpoint.argtypes = [C.POINTER(OrcTable)]
        L.orc_init_spline.argtypes = [C.POINTER(OrcTable), C.c_int]
        L.orc_init_threepoint.argtypes = [C.POINTER(OrcTable)]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


INTERP_MODES = {"3point": 0, "4point": 1, "spline": 2}


class Table:
    """Keeps the numpy buffers alive next to the C struct."""

    def __init__(self, pt):
        self.pt = pt
        self._keep = [np.ascontiguousarray(x, dtype=np.float64) for x in (pt.begin, pt.end, pt.step, pt.invstep)]
        self._len = np.ascontiguousarray(pt.len, dtype=np.int32)
        self._tab = np.ascontiguousarray(pt.table, dtype=np.float64)
        self._tab2 = None if getattr(pt, "table2", None) is None else np.ascontiguousarray(pt.table2, dtype=np.float64)
        self.c = OrcTable(pt.ncols, pt.maxsteps, _dp(self._keep[0]), _dp(self._keep[1]),
                          _dp(self._keep[2]), _dp(self._keep[3]), _ip(self._len), _dp(self._tab),
                          None if self._tab2 is None else _dp(self._tab2))
        self.interp = INTERP_MODES[getattr(pt, "interp", "3point")]

    def pair_int(self, r2, col=0):
        L = lib()
        r2 = np.atleast_1d(np.asarray(r2, dtype=np.float64))
        pot = np.empty_like(r2)
        grad = np.empty_like(r2)
        p, g, s = C.c_double(), C.c_double(), C.c_int(0)
        L.orc_set_interp(self.interp)
        for i, x in enumerate(r2):
            L.orc_pair_int(C.byref(self.c), col, self.pt.ncols, float(x), C.byref(p), C.byref(g), C.byref(s))
            pot[i], grad[i] = p.value, g.value
        return pot, grad


class EamOracle:
    """Serial CPU restatement of fix_cells / make_nblist / calc_forces / move_atoms_nve."""

    def __init__(self, box, pbc, tables, nbl_margin, ntypes=1, divisor=(1, 1, 1)):
        self.L = lib()
        self.s = self.L.orc_create()
        self.box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        self.pbc = np.ascontiguousarray(pbc, dtype=np.int32)
        self.ntypes = ntypes
        self.pair, self.rho, self.emb = Table(tables["pair"]), Table(tables["rho"]), Table(tables["emb"])
        r2_cut = max(float(tables["pair"].end.max()), float(tables["rho"].end.max()))
        div = np.ascontiguousarray(divisor, dtype=np.int32)
        rc = self.L.orc_setup(self.s, _dp(self.box), _ip(self.pbc), r2_cut, float(nbl_margin), _ip(div))
        if rc != 0:
            raise RuntimeError("orc_setup failed: %d" % rc)
        self.n = 0

    def close(self):
        if self.s:
            self.L.orc_destroy(self.s)
            self.s = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def state(self):
        class _S(C.Structure):
            _fields_ = [("box", C.c_double * 9), ("tbox", C.c_double * 9), ("height", C.c_double * 3),
                        ("pbc", C.c_int * 3), ("cellsz", C.c_double), ("gdim", C.c_int * 3),
                        ("cdim", C.c_int * 3), ("n", C.c_long), ("cl_off", c_ip), ("cell_atom", c_ip),
                        ("flat_pos", c_dp), ("n_flat", C.c_int), ("atom_cell", c_ip)]
        return C.cast(self.s, C.POINTER(_S)).contents

    def fix_cells(self, pos):
        """Wraps ``pos`` in place (do_boundaries) and bins; returns global cell coords [n,3]."""
        assert pos.dtype == np.float64 and pos.flags.c_contiguous
        self.n = pos.shape[0]
        self.L.orc_fix_cells(self.s, self.n, _dp(pos))
        st = self.state
        return np.ctypeslib.as_array(st.atom_cell, shape=(self.n, 3)).copy()

    def make_nblist(self):
        self.L.orc_make_nblist(self.s)
        tn = self.L.orc_export_pairs(self.s, None)
        pairs = np.empty((tn, 2), dtype=np.int32)
        self.L.orc_export_pairs(self.s, _ip(pairs))
        return pairs

    def calc_forces(self, pos, types):
        n = pos.shape[0]
        types = np.ascontiguousarray(types, dtype=np.int32)
        force = np.empty((n, 3)); epot = np.empty(n); rho = np.empty(n); dF = np.empty(n)
        nn = np.empty(n, dtype=np.int32); out8 = np.zeros(8)
        self.L.orc_set_interp(self.pair.interp)      # the mode of the tables handed in (a global of the C library)
        self.L.orc_calc_forces(self.s, _dp(pos), _ip(types), self.ntypes,
                               C.byref(self.pair.c), C.byref(self.rho.c), C.byref(self.emb.c),
                               _dp(force), _dp(epot), _dp(rho), _dp(dF), _ip(nn), _dp(out8))
        return dict(force=force, epot=epot, rho=rho, dF=dF, numneighs=nn,
                    tot_pot_energy=out8[0], virial=out8[1])

    def move_atoms_nve(self, pos, mom, force, mass, timestep, vtype=None, restrictions=None,
                       pdecay=None):
        n = pos.shape[0]
        vt = np.zeros(n, dtype=np.int32) if vtype is None else np.ascontiguousarray(vtype, dtype=np.int32)
        rs = np.ones((int(vt.max()) + 1, 3)) if restrictions is None else np.ascontiguousarray(restrictions, dtype=np.float64)
        force = np.ascontiguousarray(force, dtype=np.float64)
        mass = np.ascontiguousarray(mass, dtype=np.float64)
        pd = pdecay or (0, 0.0, 0.0, 1.0)
        return self.L.orc_move_atoms_nve(n, _dp(pos), _dp(mom), _dp(force), _dp(mass), _ip(vt), _dp(rs),
                                         float(timestep), int(pd[0]), float(pd[1]), float(pd[2]), float(pd[3]))

    def max_displacement2(self, pos, ref):
        return self.L.orc_max_displacement2(pos.shape[0], _dp(pos), _dp(ref))


# ----------------------------------------------------------------------------- TTM

class OrcEos(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("x", c_dp), ("y", c_dp), ("z", c_dp),
                ("diag", C.POINTER(C.c_ubyte))]


class OrcTmmParams(C.Structure):
    _fields_ = [("tmm_threshold", C.c_double), ("laser_t_1", C.c_double), ("laser_sigma_t1", C.c_double),
                ("nbb", C.c_int), ("nk", C.c_int), ("eps_bb", c_dp), ("K12", c_dp), ("phase", OrcEos)]


class OrcTtmParams(C.Structure):
    _fields_ = [("G", C.c_int), ("P", C.c_int), ("fd_n_timesteps", C.c_int), ("fd_min_atoms", C.c_int),
                ("vlatdim", C.c_int), ("vlatbuffer", C.c_int),
                ("vlatdens", C.c_double), ("atomic_weight", C.c_double), ("atomic_charge", C.c_double),
                ("Temin", C.c_double), ("timestep", C.c_double),
                ("box_x", C.c_double), ("box_y", C.c_double), ("box_z", C.c_double),
                ("r2_cut_pair", C.c_double),
                ("I0", C.c_double), ("lam", C.c_double), ("laser_t_0", C.c_double),
                ("laser_sigma_t", C.c_double), ("tmm_laser_threshold", C.c_double),
                ("src_mu", C.c_double), ("src_abs", C.c_double)]


TTM_FIELDS = ["natoms", "natoms_old", "temp", "md_temp", "xi", "dens", "fd_k", "fd_g", "Z", "ne", "Ce", "U",
              "source"]
_TTM_DECL = False


def _ttm_lib():
    global _TTM_DECL
    L = lib()
    if not _TTM_DECL:
        L.orc_ttm_create.restype = C.c_void_p
        L.orc_ttm_create.argtypes = [C.POINTER(OrcTtmParams), C.POINTER(OrcEos), C.POINTER(OrcEos)]
        L.orc_ttm_destroy.argtypes = [C.c_void_p]
        L.orc_ttm_step.argtypes = [C.c_void_p, C.c_int, C.c_long, c_dp, c_dp, c_dp, c_ip, c_ip, c_ip]
        L.orc_ttm_error.restype = C.c_char_p
        L.orc_ttm_error.argtypes = [C.c_void_p]
        L.orc_ttm_get_cells.argtypes = [C.c_void_p, c_dp]
        L.orc_ttm_get_arrays.argtypes = [C.c_void_p, c_dp, c_dp, c_dp, c_dp, c_dp]
        L.orc_ttm_get_scalars.argtypes = [C.c_void_p, c_dp]
        L.orc_ttm_get_vlattice.argtypes = [C.c_void_p, c_dp]
        L.orc_move_atoms_ttm.restype = C.c_double
        L.orc_move_atoms_ttm.argtypes = [C.c_void_p, C.c_long, c_dp, c_dp, c_dp, c_dp, c_ip, c_ip, c_dp,
                                         C.c_int, C.c_double, C.c_double, C.c_double]
        L.orc_eos_interp.restype = C.c_double
        L.orc_eos_interp.argtypes = [C.POINTER(OrcEos), C.c_double, C.c_double]
        L.orc_eos_te_from_r_ee.restype = C.c_double
        L.orc_eos_te_from_r_ee.argtypes = [C.POINTER(OrcEos), C.c_double, C.c_double]
        L.orc_ttm_mean_charge.restype = C.c_double
        L.orc_ttm_mean_charge.argtypes = [C.c_double] * 4
        for f in (L.orc_ttm_kappa, L.orc_ttm_gamma):
            f.restype = C.c_double
            f.argtypes = [C.c_double] * 5 + [C.c_int]
        _TTM_DECL = True
    return L


class Eos:
    """Keeps the numpy buffers of an EOS table (imd_b200.synth.EosTable) alive next to the C struct."""

    def __init__(self, tab):
        self.x = np.ascontiguousarray(tab.x, np.float64)
        self.y = np.ascontiguousarray(tab.y, np.float64)
        self.z = np.ascontiguousarray(tab.z, np.float64)
        d = tab.diag if tab.diag is not None else np.zeros((self.x.size - 1, self.y.size - 1), np.uint8)
        self.diag = np.ascontiguousarray(d, np.uint8)
        self.c = OrcEos(self.x.size, self.y.size, _dp(self.x), _dp(self.y), _dp(self.z),
                        self.diag.ctypes.data_as(C.POINTER(C.c_ubyte)))

    def interp(self, x, y):
        L = _ttm_lib()
        return np.array([L.orc_eos_interp(C.byref(self.c), float(a), float(b)) for a, b in zip(x, y)])

    def te_from_r_ee(self, r, e):
        L = _ttm_lib()
        return np.array([L.orc_eos_te_from_r_ee(C.byref(self.c), float(a), float(b)) for a, b in zip(r, e)])


class TtmOracle:
    """calc_ttm + move_atoms_ttm of the reference (1-D TTM), emulating P MPI ranks."""

    def __init__(self, tp, eos, box, r2_cut_pair, timestep, nranks):
        self.L = _ttm_lib()
        self.cve, self.ee = Eos(eos["cve"]), Eos(eos["ee"])
        box = np.asarray(box, np.float64).reshape(3, 3)
        self.par = OrcTtmParams(tp.ttmdimx, nranks, tp.fd_n_timesteps, tp.fd_min_atoms, tp.vlatdim, tp.vlatbuffer,
                                tp.vlatdens, tp.atomic_weight, tp.atomic_charge, tp.Temin, timestep,
                                box[0, 0], box[1, 1], box[2, 2], r2_cut_pair, tp.I0, tp.lam, tp.laser_t_0,
                                tp.laser_sigma_t, tp.tmm_laser_threshold, tp.src_mu, tp.src_abs)
        self.G = tp.ttmdimx
        self.vlatdim = tp.vlatdim
        self.t = self.L.orc_ttm_create(C.byref(self.par), C.byref(self.cve.c), C.byref(self.ee.c))
        if not self.t:
            raise RuntimeError("orc_ttm_create failed (ttmdimx % nranks != 0?)")
        if getattr(tp, "use_tmm", 0):
            # the reference's own transfer-matrix source (do_tmm) on the synthetic data files
            from imd_b200 import synth
            tabs = synth.make_tmm_tables()
            self.phase = Eos(eos["phase"])
            self._bb = np.ascontiguousarray(tabs["eps_bb"], np.float64)
            self._k12 = np.ascontiguousarray(tabs["K12"], np.float64)
            self.tmm = OrcTmmParams(tp.tmm_threshold, 0.0, 0.0, self._bb.shape[0], self._k12.shape[0],
                                    _dp(self._bb), _dp(self._k12), self.phase.c)
            self.L.orc_ttm_set_tmm.argtypes = [C.c_void_p, C.POINTER(OrcTmmParams)]
            self.L.orc_ttm_set_tmm(self.t, C.byref(self.tmm))

    def close(self):
        if getattr(self, "t", None):
            self.L.orc_ttm_destroy(self.t)
            self.t = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def step(self, steps, pos, mom, mass, types, numneighs, fdi):
        """fdi: int32 [n] in/out (slab index of the previous call)."""
        n = pos.shape[0]
        pos = np.ascontiguousarray(pos, np.float64); mom = np.ascontiguousarray(mom, np.float64)
        mass = np.ascontiguousarray(np.broadcast_to(np.asarray(mass, np.float64), (n,)))
        types = np.ascontiguousarray(types, np.int32); nn = np.ascontiguousarray(numneighs, np.int32)
        assert fdi.dtype == np.int32 and fdi.flags.c_contiguous
        rc = self.L.orc_ttm_step(self.t, int(steps), n, _dp(pos), _dp(mom), _dp(mass), _ip(types), _ip(nn), _ip(fdi))
        if rc != 0:
            raise RuntimeError("orc_ttm_step: " + self.L.orc_ttm_error(self.t).decode())

    def cells(self):
        f = np.zeros((self.G, len(TTM_FIELDS)))
        self.L.orc_ttm_get_cells(self.t, _dp(f))
        return {k: f[:, i].copy() for i, k in enumerate(TTM_FIELDS)}

    def arrays(self):
        xi = np.zeros(self.G); vcom = np.zeros((3, self.G)); na = np.zeros(self.G)
        fl = np.zeros(self.G); fr = np.zeros(self.G)
        self.L.orc_ttm_get_arrays(self.t, _dp(xi), _dp(vcom), _dp(na), _dp(fl), _dp(fr))
        return dict(xiarr=xi, vcom=vcom, natomsglobal=na, fluxl=fl, fluxr=fr)

    def scalars(self):
        s = np.zeros(16)
        self.L.orc_ttm_get_scalars(self.t, _dp(s))
        keys = ["diff_substeps", "max_dt_ttm", "last_active_cell_global", "cur_vlattice_proc", "Eabs",
                "tot_elec_energy", "fd_hx", "fd_vol", "neighvol", "laser_active", "tmm_time"]
        return dict(zip(keys, s))

    def vlattice(self):
        v = np.zeros((max(self.vlatdim, 1), 6))
        self.L.orc_ttm_get_vlattice(self.t, _dp(v))
        return v[: self.vlatdim]

    def move_atoms(self, pos, mom, force, mass, types, vtypes=None, restrictions=None, pdecay=None):
        n = pos.shape[0]
        assert pos.flags.c_contiguous and mom.flags.c_contiguous
        force = np.ascontiguousarray(force, np.float64)
        mass = np.ascontiguousarray(np.broadcast_to(np.asarray(mass, np.float64), (n,)))
        types = np.ascontiguousarray(types, np.int32)
        vt = types if vtypes is None else np.ascontiguousarray(vtypes, np.int32)
        rs = np.ones((int(vt.max()) + 1, 3)) if restrictions is None else np.ascontiguousarray(restrictions, np.float64)
        pd = pdecay or (0, 0.0, 0.0, 1.0)
        return self.L.orc_move_atoms_ttm(self.t, n, _dp(pos), _dp(mom), _dp(force), _dp(mass), _ip(types), _ip(vt),
                                         _dp(rs), int(pd[0]), float(pd[1]), float(pd[2]), float(pd[3]))
