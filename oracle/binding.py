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
                ("len", c_ip), ("table", c_dp)]


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("eam_oracle.c", "ttm_oracle.c", "oracle.h")]
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
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(c_dp)


def _ip(a):
    return a.ctypes.data_as(c_ip)


class Table:
    """Keeps the numpy buffers alive next to the C struct."""

    def __init__(self, pt):
        self.pt = pt
        self._keep = [np.ascontiguousarray(x, dtype=np.float64) for x in (pt.begin, pt.end, pt.step, pt.invstep)]
        self._len = np.ascontiguousarray(pt.len, dtype=np.int32)
        self._tab = np.ascontiguousarray(pt.table, dtype=np.float64)
        self.c = OrcTable(pt.ncols, pt.maxsteps, _dp(self._keep[0]), _dp(self._keep[1]),
                          _dp(self._keep[2]), _dp(self._keep[3]), _ip(self._len), _dp(self._tab))

    def pair_int(self, r2, col=0):
        L = lib()
        r2 = np.atleast_1d(np.asarray(r2, dtype=np.float64))
        pot = np.empty_like(r2)
        grad = np.empty_like(r2)
        p, g, s = C.c_double(), C.c_double(), C.c_int(0)
        for i, x in enumerate(r2):
            L.orc_pair_int2(C.byref(self.c), col, self.pt.ncols, float(x), C.byref(p), C.byref(g), C.byref(s))
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
