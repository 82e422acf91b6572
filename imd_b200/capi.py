"""ctypes binding of the C ABI in include/imdcu.h (imd_b200/libimdcu.so).

This is the same boundary a C host (the IMD stubs of INTEGRATION.md) binds; nothing
here computes anything.  There is no CPU fallback: if the library is missing, or there
is no CUDA device, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libimdcu.so")

NPHASE = 8
INTERP_MODES = {"3point": 0, "4point": 1, "spline": 2}
PHASE_NAMES = ["bin_sort", "nbl_build", "ghost_refresh", "eam_pass1", "embed", "eam_pass2",
               "integrate", "ttm"]

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)
c_lp = C.POINTER(C.c_int64)

_lib = None


class TtmParamsC(C.Structure):
    """imdcu_ttm_params (include/imdcu.h)."""
    _fields_ = [("ttmdimx", C.c_int), ("nranks_ref", C.c_int), ("fd_n_timesteps", C.c_int),
                ("fd_min_atoms", C.c_int), ("vlatdim", C.c_int), ("vlatbuffer", C.c_int),
                ("vlatdens", C.c_double), ("atomic_weight", C.c_double), ("atomic_charge", C.c_double),
                ("Temin", C.c_double), ("I0", C.c_double), ("lam", C.c_double), ("laser_t_0", C.c_double),
                ("laser_sigma_t", C.c_double), ("tmm_laser_threshold", C.c_double), ("src_mu", C.c_double),
                ("src_abs", C.c_double)]


TTM_FIELDS = ["natoms", "natoms_old", "temp", "md_temp", "xi", "dens", "fd_k", "fd_g", "Z", "ne", "Ce", "U",
              "source"]
TTM_SCALARS = ["diff_substeps", "max_dt_ttm", "last_active_cell_global", "cur_vlattice_proc", "Eabs",
               "tot_elec_energy", "fd_hx", "fd_vol", "neighvol", "laser_active", "tmm_time", "err"]


class ImdcuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("imdcu error %d: %s" % (code, msg))
        self.code = code


def load():
    """dlopen libimdcu.so and declare every symbol of include/imdcu.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImdcuError(-100, "%s not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    sig = {
        "imdcu_abi_version": (C.c_int, []),
        "imdcu_device_count": (C.c_int, [c_ip]),
        "imdcu_create": (C.c_int, [C.POINTER(vp), C.c_int]),
        "imdcu_destroy": (C.c_int, [vp]),
        "imdcu_last_error": (C.c_char_p, [vp]),
        "imdcu_set_box": (C.c_int, [vp, c_dp, c_ip, C.c_double, C.c_double, c_ip]),
        "imdcu_set_table": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_dp, c_dp, c_ip, c_dp]),
        "imdcu_set_interp": (C.c_int, [vp, C.c_int]),
        "imdcu_set_table2": (C.c_int, [vp, C.c_int, c_dp]),
        "imdcu_set_params": (C.c_int, [vp, C.c_int, C.c_double, C.c_int, c_dp]),
        "imdcu_set_pdecay": (C.c_int, [vp, C.c_int, C.c_double, C.c_double, C.c_double]),
        "imdcu_set_tuning": (C.c_int, [vp, C.c_char_p, C.c_int]),
        "imdcu_upload_atoms": (C.c_int, [vp, C.c_int64, c_ip, c_ip, c_ip, c_dp, c_dp, c_dp]),
        "imdcu_num_atoms": (C.c_int64, [vp]),
        "imdcu_update_atoms": (C.c_int, [vp, c_dp, c_dp]),
        "imdcu_download_atoms": (C.c_int, [vp, c_ip, c_ip, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_ip, c_ip]),
        "imdcu_download_vtypes": (C.c_int, [vp, c_ip]),
        "imdcu_calc_forces": (C.c_int, [vp, C.c_int, c_dp]),
        "imdcu_move_atoms_nve": (C.c_int, [vp, c_dp, c_ip]),
        "imdcu_run_nve": (C.c_int, [vp, C.c_int, C.c_int, c_dp]),
        "imdcu_comm_unique_id": (C.c_int, [C.c_char_p]),
        "imdcu_comm_init": (C.c_int, [vp, C.c_char_p, C.c_int, C.c_int]),
        "imdcu_comm_rank": (C.c_int, [vp, c_ip, c_ip, c_ip, c_ip]),
        "imdcu_balance_cuts": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_int64), C.c_int, c_ip]),
        "imdcu_comm_set_cuts": (C.c_int, [vp, c_ip]),
        "imdcu_ttm_set_eos": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, c_dp, c_dp, c_dp, C.POINTER(C.c_ubyte)]),
        "imdcu_ttm_init": (C.c_int, [vp, C.POINTER(TtmParamsC)]),
        "imdcu_ttm_set_tmm": (C.c_int, [vp, C.c_double, C.c_double, C.c_double, C.c_int, c_dp, C.c_int, c_dp]),
        "imdcu_calc_forces_ttm": (C.c_int, [vp, C.c_int, c_dp, c_dp]),
        "imdcu_calc_ttm": (C.c_int, [vp, C.c_int, c_dp]),
        "imdcu_move_atoms_ttm": (C.c_int, [vp, c_dp, c_ip]),
        "imdcu_run_ttm": (C.c_int, [vp, C.c_int, C.c_int, c_dp]),
        "imdcu_download_ttm": (C.c_int, [vp, c_dp, c_dp, c_dp, c_dp]),
        "imdcu_set_nrb": (C.c_int, [vp, C.c_double, C.c_int64, c_ip, c_ip, c_ip, c_dp, C.c_int]),
        "imdcu_ttm3d_init": (C.c_int, [vp, C.c_int, C.c_int, C.c_int, c_dp, c_ip, C.c_int]),
        "imdcu_ttm3d_upload": (C.c_int, [vp, c_ip, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp, c_dp]),
        "imdcu_ttm3d_diff": (C.c_int, [vp, C.c_double, C.c_int, C.c_int, C.c_int, c_dp]),
        "imdcu_ttm3d_fill_ghost_layers": (C.c_int, [vp]),
        "imdcu_ttm3d_download": (C.c_int, [vp, c_ip, c_dp, c_dp, c_dp]),
        "imdcu_get_cell_grid": (C.c_int, [vp, c_ip, c_dp]),
        "imdcu_get_nblist_pairs": (C.c_int, [vp, c_lp, c_ip]),
        "imdcu_nbl_count": (C.c_int, [vp]),
        "imdcu_short_distance": (C.c_int, [vp]),
        "imdcu_invalidate_nblist": (C.c_int, [vp]),
        "imdcu_launch_count": (C.c_int64, [vp]),
        "imdcu_set_profiling": (C.c_int, [vp, C.c_int]),
        "imdcu_get_phase_times": (C.c_int, [vp, c_dp, c_lp]),
        "imdcu_stream": (vp, [vp]),
        "imdcu_synchronize": (C.c_int, [vp]),
        "imdcu_probe_fp64": (C.c_int, [vp, c_dp]),
        "imdcu_eval_table": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, c_dp, c_dp, c_dp]),
        "imdcu_distrib": (C.c_int, [vp, C.POINTER(C.c_int), c_dp, c_dp, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                    C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "imdcu_eval_table_noalloc": (C.c_int, [vp, C.c_int, C.c_int, C.c_int64, c_dp, c_dp, c_dp]),
        "imdcu_eval_ttm_math": (C.c_int, [vp, C.c_int, C.c_int64, c_dp, c_dp, c_dp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    L._declared = sorted(sig)
    _lib = L
    return L


EVAL_NOALLOC_MAX = 65536      # IMDCU_EVAL_NOALLOC_MAX (include/imdcu.h)


def balance_cuts(nranks, atoms_per_layer, min_layers=1):
    """imdcu_balance_cuts (host arithmetic, no GPU): slab boundaries of equal atom count on cell-layer boundaries."""
    h = np.ascontiguousarray(atoms_per_layer, np.int64)
    x = np.zeros(nranks + 1, np.int32)
    rc = load().imdcu_balance_cuts(int(nranks), int(h.shape[0]), h.ctypes.data_as(C.POINTER(C.c_int64)), int(min_layers),
                                   x.ctypes.data_as(C.POINTER(C.c_int32)))
    if rc != 0:
        raise ValueError("imdcu_balance_cuts: bad argument (rc %d)" % rc)
    return x


def _dp(a):
    return None if a is None else a.ctypes.data_as(c_dp)


def _ip(a):
    return None if a is None else a.ctypes.data_as(c_ip)


class Context:
    """One imdcu_ctx (one GPU).  Thin: argument marshalling and error translation only."""

    def __init__(self, device: int = 0):
        self.L = load()
        h = C.c_void_p()
        rc = self.L.imdcu_create(C.byref(h), device)
        if rc != 0:
            raise ImdcuError(rc, (self.L.imdcu_last_error(None) or b"").decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.L.imdcu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise ImdcuError(rc, (self.L.imdcu_last_error(self.h) or b"").decode())

    # ---- multi-GPU
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        rc = load().imdcu_comm_unique_id(buf)
        if rc != 0:
            raise ImdcuError(rc, "ncclGetUniqueId failed")
        return buf.raw

    def comm_init(self, uid: bytes, rank: int, nranks: int):
        self._ck(self.L.imdcu_comm_init(self.h, uid, int(rank), int(nranks)))

    def comm_set_cuts(self, xcut):
        """imdcu_comm_set_cuts: slab boundaries (global x cell layers), after set_box, before upload_atoms."""
        x = np.ascontiguousarray(xcut, np.int32)
        self._ck(self.L.imdcu_comm_set_cuts(self.h, _ip(x)))

    def comm_rank(self):
        v = [C.c_int32(0) for _ in range(4)]
        self._ck(self.L.imdcu_comm_rank(self.h, *[C.byref(x) for x in v]))
        return dict(rank=v[0].value, nranks=v[1].value, xoff=v[2].value, ldimx=v[3].value)

    # ---- set-up
    def set_interp(self, mode):
        """Table interpolation of the run: "3point" (default), "4point", "spline" (potaccess.h:24-36)."""
        self._ck(self.L.imdcu_set_interp(self.h, INTERP_MODES[mode] if isinstance(mode, str) else int(mode)))

    def set_table(self, which: int, pt):
        begin = np.ascontiguousarray(pt.begin, np.float64); end = np.ascontiguousarray(pt.end, np.float64)
        step = np.ascontiguousarray(pt.step, np.float64); inv = np.ascontiguousarray(pt.invstep, np.float64)
        ln = np.ascontiguousarray(pt.len, np.int32); tab = np.ascontiguousarray(pt.table, np.float64)
        self._ck(self.L.imdcu_set_table(self.h, which, pt.ncols, pt.maxsteps, _dp(begin), _dp(end), _dp(step),
                                        _dp(inv), _ip(ln), _dp(tab)))
        if getattr(pt, "table2", None) is not None:
            tab2 = np.ascontiguousarray(pt.table2, np.float64)
            self._ck(self.L.imdcu_set_table2(self.h, which, _dp(tab2)))

    def set_box(self, box, pbc, nbl_margin, nbl_size=1.3, divisor=None):
        box = np.ascontiguousarray(box, np.float64).reshape(9)
        pbc = np.ascontiguousarray(pbc, np.int32)
        div = None if divisor is None else np.ascontiguousarray(divisor, np.int32)
        self._ck(self.L.imdcu_set_box(self.h, _dp(box), _ip(pbc), float(nbl_margin), float(nbl_size), _ip(div)))

    def set_params(self, ntypes, timestep, restrictions=None):
        rs = np.ones((1, 3)) if restrictions is None else np.ascontiguousarray(restrictions, np.float64).reshape(-1, 3)
        self._ck(self.L.imdcu_set_params(self.h, int(ntypes), float(timestep), rs.shape[0], _dp(rs)))

    def set_pdecay(self, enabled, xipdecay=0.0, ramp_start=0.0, ramp_end=1.0):
        self._ck(self.L.imdcu_set_pdecay(self.h, int(enabled), xipdecay, ramp_start, ramp_end))

    def set_tuning(self, **knobs):
        """imdcu_set_tuning: lpa=0|1|2|4|8, fused_table=0|1, ttm_overlap=0|1, ..."""
        for k, v in knobs.items():
            self._ck(self.L.imdcu_set_tuning(self.h, k.encode(), int(v)))

    def set_nrb(self, nrb_k, ids, bnd, nrbi, refpos, startup_images=False):
        """imdcu_set_nrb: entries for the boundary atoms and their listed neighbours (ids empty: off)."""
        ids = np.ascontiguousarray(ids, np.int32); bnd = np.ascontiguousarray(bnd, np.int32)
        nrbi = np.ascontiguousarray(nrbi, np.int32).reshape(-1, 12); refpos = np.ascontiguousarray(refpos, np.float64).reshape(-1, 3)
        n = ids.shape[0]
        if not (bnd.shape[0] == n and nrbi.shape[0] == n and refpos.shape[0] == n):
            raise ValueError("set_nrb: ids, bnd, nrbi, refpos must have one row per entry")
        self._ck(self.L.imdcu_set_nrb(self.h, float(nrb_k), n, _ip(ids), _ip(bnd), _ip(nrbi), _dp(refpos), int(bool(startup_images))))

    # ---- atoms
    def upload_atoms(self, pos, mom, mass, types=None, ids=None, vtypes=None):
        pos = np.ascontiguousarray(pos, np.float64); mom = np.ascontiguousarray(mom, np.float64)
        n = pos.shape[0]
        mass = np.ascontiguousarray(np.broadcast_to(np.asarray(mass, np.float64), (n,)))
        ty = None if types is None else np.ascontiguousarray(types, np.int32)
        vt = None if vtypes is None else np.ascontiguousarray(vtypes, np.int32)
        idd = None if ids is None else np.ascontiguousarray(ids, np.int32)
        self._ck(self.L.imdcu_upload_atoms(self.h, n, _ip(idd), _ip(ty), _ip(vt), _dp(mass), _dp(pos), _dp(mom)))

    def update_atoms(self, pos, mom):
        """pos/mom [n,3] float64 C-contiguous in device order (pinned memory for full speed)."""
        self._ck(self.L.imdcu_update_atoms(self.h, _dp(pos), _dp(mom)))

    def download_into(self, pos=None, mom=None, force=None, epot=None):
        """Raw download into caller buffers (e.g. pinned), device order."""
        self._ck(self.L.imdcu_download_atoms(self.h, None, None, None, _dp(pos), _dp(mom), _dp(force), _dp(epot),
                                             None, None, None, None))

    @property
    def n(self):
        return int(self.L.imdcu_num_atoms(self.h))

    def download(self, fields=("id", "pos", "mom", "force", "epot", "rho", "dF", "numneighs", "cell")):
        n = self.n
        buf = {
            "id": np.empty(n, np.int32), "type": np.empty(n, np.int32), "mass": np.empty(n),
            "pos": np.empty((n, 3)), "mom": np.empty((n, 3)), "force": np.empty((n, 3)),
            "epot": np.empty(n), "rho": np.empty(n), "dF": np.empty(n),
            "numneighs": np.empty(n, np.int32), "cell": np.empty((n, 3), np.int32),
        }
        g = lambda k: buf[k] if k in fields else None  # noqa: E731
        self._ck(self.L.imdcu_download_atoms(self.h, _ip(g("id")), _ip(g("type")), _dp(g("mass")), _dp(g("pos")),
                                             _dp(g("mom")), _dp(g("force")), _dp(g("epot")), _dp(g("rho")),
                                             _dp(g("dF")), _ip(g("numneighs")), _ip(g("cell"))))
        return {k: buf[k] for k in fields}

    # ---- hot path
    def calc_forces(self, steps=0):
        out = np.zeros(8)
        self._ck(self.L.imdcu_calc_forces(self.h, steps, _dp(out)))
        return out

    def move_atoms_nve(self):
        out = np.zeros(8); valid = C.c_int32(0)
        self._ck(self.L.imdcu_move_atoms_nve(self.h, _dp(out), C.byref(valid)))
        return out, bool(valid.value)

    def run_nve(self, first_step, nsteps, want_energies=True):
        e = np.zeros((nsteps, 2)) if want_energies else None
        self._ck(self.L.imdcu_run_nve(self.h, first_step, nsteps, _dp(e)))
        return e

    # ---- two-temperature model
    def ttm_set_eos(self, which, tab):
        """tab: imd_b200.synth.EosTable (x, y, z[nx,ny], diag[nx-1,ny-1] or None)."""
        x = np.ascontiguousarray(tab.x, np.float64); y = np.ascontiguousarray(tab.y, np.float64)
        z = np.ascontiguousarray(tab.z, np.float64)
        d = None if tab.diag is None else np.ascontiguousarray(tab.diag, np.uint8)
        dp = None if d is None else d.ctypes.data_as(C.POINTER(C.c_ubyte))
        self._ck(self.L.imdcu_ttm_set_eos(self.h, which, x.size, y.size, _dp(x), _dp(y), _dp(z), dp))

    def ttm_set_tmm(self, tmm_threshold, eps_bb, K12, laser_t_1=0.0, laser_sigma_t1=0.0):
        """The reference's transfer-matrix source (do_tmm): eps_bb [nbb,3], K12 [>=1500,2] as tmm_init reads them."""
        bb = np.ascontiguousarray(eps_bb, np.float64); k12 = np.ascontiguousarray(K12, np.float64)
        self._ck(self.L.imdcu_ttm_set_tmm(self.h, float(tmm_threshold), float(laser_t_1), float(laser_sigma_t1),
                                          bb.shape[0], _dp(bb), k12.shape[0], _dp(k12)))

    def ttm_init(self, tp, nranks_ref=1):
        """tp: imd_b200.synth.TtmParams."""
        self._ttm_G, self._ttm_vlatdim = tp.ttmdimx, tp.vlatdim
        par = TtmParamsC(tp.ttmdimx, nranks_ref, tp.fd_n_timesteps, tp.fd_min_atoms, tp.vlatdim, tp.vlatbuffer,
                         tp.vlatdens, tp.atomic_weight, tp.atomic_charge, tp.Temin, tp.I0, tp.lam, tp.laser_t_0,
                         tp.laser_sigma_t, tp.tmm_laser_threshold, tp.src_mu, tp.src_abs)
        self._ck(self.L.imdcu_ttm_init(self.h, C.byref(par)))

    def calc_forces_ttm(self, steps):
        """imdcu_calc_forces_ttm: calc_forces + calc_ttm in one call (lattice kernel beside EAM pass 2)."""
        o8 = np.zeros(8); o16 = np.zeros(16)
        self._ck(self.L.imdcu_calc_forces_ttm(self.h, int(steps), _dp(o8), _dp(o16)))
        return o8, dict(zip(TTM_SCALARS, o16))

    def calc_ttm(self, steps, want_scalars=True):
        out = np.zeros(16) if want_scalars else None
        self._ck(self.L.imdcu_calc_ttm(self.h, int(steps), _dp(out)))
        return None if out is None else dict(zip(TTM_SCALARS, out))

    def move_atoms_ttm(self):
        out = np.zeros(8); valid = C.c_int32(0)
        self._ck(self.L.imdcu_move_atoms_ttm(self.h, _dp(out), C.byref(valid)))
        return out, bool(valid.value)

    def run_ttm(self, first_step, nsteps, want_energies=True):
        e = np.zeros((nsteps, 2)) if want_energies else None
        self._ck(self.L.imdcu_run_ttm(self.h, first_step, nsteps, _dp(e)))
        return e

    def download_ttm(self):
        G, nv = self._ttm_G, self._ttm_vlatdim
        cells = np.zeros((G, len(TTM_FIELDS))); xi = np.zeros(G); vcom = np.zeros((3, G))
        vlat = np.zeros((max(nv, 1), 6))
        self._ck(self.L.imdcu_download_ttm(self.h, _dp(cells), _dp(xi), _dp(vcom), _dp(vlat)))
        d = {k: cells[:, i].copy() for i, k in enumerate(TTM_FIELDS)}
        return dict(fields=d, xiarr=xi, vcom=vcom, vlat=vlat[:nv])

    # ---- 3-D TTM lattice sweep (imd_ttm_3D.c); arrays are the extended lattice (nx+2, ny+2, nz+2)
    T3_FIELDS = ("natoms", "temp", "fd_k", "fd_g", "Ce", "md_temp", "source", "U", "xi", "dens")

    def ttm3d_init(self, shape, h, pbc, fd_min_atoms):
        self._t3_shape = tuple(int(v) + 2 for v in shape)
        hh = np.ascontiguousarray(h, np.float64); pp = np.ascontiguousarray(pbc, np.int32)
        self._ck(self.L.imdcu_ttm3d_init(self.h, int(shape[0]), int(shape[1]), int(shape[2]), _dp(hh), _ip(pp),
                                         int(fd_min_atoms)))

    def ttm3d_upload(self, **fields):
        args = []
        for k in self.T3_FIELDS:
            a = fields.get(k)
            if a is not None:
                a = np.ascontiguousarray(a, np.int32 if k == "natoms" else np.float64)
                if a.shape != self._t3_shape:
                    raise ValueError("ttm3d field %s has shape %s, expected the extended lattice %s" % (k, a.shape, self._t3_shape))
            args.append(a)
        self._ck(self.L.imdcu_ttm3d_upload(self.h, _ip(args[0]), *[_dp(a) for a in args[1:]]))

    def ttm3d_diff(self, tau, diff_substeps, laser_active, nsweeps, want_eabs=True):
        e = np.zeros(1)
        self._ck(self.L.imdcu_ttm3d_diff(self.h, float(tau), int(diff_substeps), int(bool(laser_active)), int(nsweeps),
                                         _dp(e) if want_eabs else None))
        return float(e[0])

    def ttm3d_fill_ghost_layers(self):
        self._ck(self.L.imdcu_ttm3d_fill_ghost_layers(self.h))

    def ttm3d_download(self, fields=("natoms", "temp", "U", "xi")):
        sh = self._t3_shape
        out = {k: np.empty(sh, np.int32 if k == "natoms" else np.float64) for k in fields}
        g = out.get
        self._ck(self.L.imdcu_ttm3d_download(self.h, _ip(g("natoms")), _dp(g("temp")), _dp(g("U")), _dp(g("xi"))))
        return out

    def eval_ttm_math(self, op, a, b=None):
        """op: 'log', 'exp', 'pow', 'div', 'sqrt', 'rcp' -- the device versions of ttm_math.cuh."""
        code = {"log": 0, "exp": 1, "pow": 2, "div": 3, "sqrt": 4, "rcp": 5}[op]
        a = np.ascontiguousarray(a, np.float64)
        b = None if b is None else np.ascontiguousarray(b, np.float64)
        out = np.empty_like(a)
        self._ck(self.L.imdcu_eval_ttm_math(self.h, code, a.size, _dp(a), _dp(b), _dp(out)))
        return out

    # ---- inspection
    def cell_grid(self):
        g = np.zeros(3, np.int32); cs = C.c_double()
        self._ck(self.L.imdcu_get_cell_grid(self.h, _ip(g), C.byref(cs)))
        return g, cs.value

    def nblist_pairs(self):
        npairs = C.c_int64(0)
        self._ck(self.L.imdcu_get_nblist_pairs(self.h, C.byref(npairs), None))
        pairs = np.empty((npairs.value, 2), np.int32)
        self._ck(self.L.imdcu_get_nblist_pairs(self.h, C.byref(npairs), _ip(pairs)))
        return pairs

    @property
    def nbl_count(self):
        return int(self.L.imdcu_nbl_count(self.h))

    def short_distance(self):
        """True if a table was evaluated below its first argument since the last call (is_short)."""
        rc = self.L.imdcu_short_distance(self.h)
        if rc < 0:
            self._ck(rc)
        return rc > 0

    def invalidate_nblist(self):
        self._ck(self.L.imdcu_invalidate_nblist(self.h))

    @property
    def launch_count(self):
        return int(self.L.imdcu_launch_count(self.h))

    def set_profiling(self, on=True):
        self._ck(self.L.imdcu_set_profiling(self.h, int(on)))

    def phase_times(self):
        ms = np.zeros(NPHASE); ln = np.zeros(NPHASE, np.int64)
        self._ck(self.L.imdcu_get_phase_times(self.h, _dp(ms), ln.ctypes.data_as(c_lp)))
        return dict(zip(PHASE_NAMES, ms)), dict(zip(PHASE_NAMES, ln.tolist()))

    @property
    def stream(self):
        return int(self.L.imdcu_stream(self.h) or 0)

    def synchronize(self):
        self._ck(self.L.imdcu_synchronize(self.h))

    def probe_fp64(self):
        """measured FP64-FMA rate of this GPU in GFLOP/s"""
        v = C.c_double(0.0)
        self._ck(self.L.imdcu_probe_fp64(self.h, C.byref(v)))
        return v.value

    def distrib(self, dim, ll, ur):
        """imdcu_distrib: per-bin sums (float32) of this rank's atoms: num, ekin, epot, vxavg."""
        d = (C.c_int * 3)(*[int(x) for x in dim])
        lo = np.ascontiguousarray(ll, np.float64); hi = np.ascontiguousarray(ur, np.float64)
        size = int(dim[0]) * int(dim[1]) * int(dim[2])
        out = {k: np.zeros(size, np.float32) for k in ("num", "ekin", "epot", "vxavg")}
        fp = lambda a: a.ctypes.data_as(C.POINTER(C.c_float))  # noqa: E731
        self._ck(self.L.imdcu_distrib(self.h, d, _dp(lo), _dp(hi), fp(out["num"]), fp(out["ekin"]), fp(out["epot"]),
                                      fp(out["vxavg"])))
        return out

    def eval_table(self, which, arg, col=0):
        arg = np.ascontiguousarray(arg, np.float64)
        val = np.empty_like(arg); grad = np.empty_like(arg)
        self._ck(self.L.imdcu_eval_table(self.h, which, col, arg.shape[0], _dp(arg), _dp(val), _dp(grad)))
        return val, grad

    def eval_table_noalloc(self, which, arg, col=0):
        """imdcu_eval_table_noalloc: same result, device staging in the library's static array (no cudaMalloc);
        at most EVAL_NOALLOC_MAX arguments per call."""
        arg = np.ascontiguousarray(arg, np.float64)
        val = np.empty_like(arg); grad = np.empty_like(arg)
        self._ck(self.L.imdcu_eval_table_noalloc(self.h, which, col, arg.shape[0], _dp(arg), _dp(val), _dp(grad)))
        return val, grad
