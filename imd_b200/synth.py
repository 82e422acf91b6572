"""Synthetic inputs for the EAM + TTM hot path.

Everything here is host-side set-up, not the hot path:

* analytic Johnson-style nearest-neighbour EAM functions tabulated in IMD potential
  format 2 (reader: reference imd_potential.c:394-466) -- ``write_pot_table2`` /
  ``read_pot_table2``; the in-memory form mirrors ``pot_table_t`` (types.h:537-549),
  including the shift-to-zero of radial tables (imd_potential.c:441-452) and the two
  extrapolated rows of ``init_threepoint`` (imd_potential.c:1260-1276);
* fcc lattices in the order and placement of the reference generator ``init_cubic``
  (imd_generate.c:291-586) is NOT reproduced bit for bit -- parity inputs come from the
  reference's own dumps (tests/golden); these lattices feed bench.py and property tests;
* IMD parameter files for the reference binary (keywords: imd_param.c:314-3722).
"""
from __future__ import annotations

import dataclasses
import io
import os
from typing import Dict, Optional, Sequence, Tuple

import numpy as np


# ----------------------------------------------------------------------------- tables

@dataclasses.dataclass
class PotTable:
    """Mirror of the reference's ``pot_table_t`` (types.h:537-549), ncols columns."""
    begin: np.ndarray      # [ncols]
    end: np.ndarray        # [ncols]
    step: np.ndarray       # [ncols]
    invstep: np.ndarray    # [ncols]
    len: np.ndarray        # [ncols] int32
    maxsteps: int
    table: np.ndarray      # [(maxsteps+2), ncols] float64, C order
    table2: Optional[np.ndarray] = None   # spline mode only: second derivatives, same shape (init_spline)
    interp: str = "3point"                # "3point" | "4point" | "spline" (potaccess.h:24-36)

    @property
    def ncols(self) -> int:
        return int(self.begin.shape[0])


def init_spline(tab: np.ndarray, ln, step, radial: bool) -> np.ndarray:
    """Second-derivative table of the spline mode and the extra row of ``tab`` (in place), the
    recurrences of init_spline (imd_potential.c:1201-1248) in the reference's operation order."""
    nrows, ncols = tab.shape
    tab2 = np.zeros_like(tab)
    for col in range(ncols):
        y, y2 = tab[:, col], tab2[:, col]
        n, st = int(ln[col]), float(step[col])
        u = np.zeros(nrows)
        y2[0] = u[0] = 0.0
        for i in range(1, n - 1):
            p = 0.5 * y2[i - 1] + 2.0
            y2[i] = -0.5 / p
            u[i] = (y[i + 1] - 2 * y[i] + y[i - 1]) / st
            u[i] = (6.0 * u[i] / (2 * st) - 0.5 * u[i - 1]) / p
        if radial:
            qn = 0.5
            un = (3.0 / st) * (y[n - 2] - y[n - 1]) / st
        else:
            qn = un = 0.0
        y2[n - 1] = (un - qn * u[n - 2]) / (qn * y2[n - 2] + 1.0)
        for k in range(n - 2, -1, -1):
            y2[k] = y2[k] * y2[k + 1] + u[k]
        y[n] = 2 * y[n - 1] - y[n - 2] + (st * st) * y2[n - 1]
        y2[n] = 2 * y2[n - 1] - y2[n - 2]
    return tab2


def _table_from_columns(begin, end, step, cols, radial: bool, interp: str = "3point") -> PotTable:
    """Assemble a PotTable from per-column value arrays exactly as read_pot_table2 +
    init_threepoint / init_fourpoint / init_spline do (imd_potential.c:394-466, 1167-1276)."""
    ncols = len(cols)
    begin = np.asarray(begin, dtype=np.float64)
    end = np.asarray(end, dtype=np.float64)
    step = np.asarray(step, dtype=np.float64)
    invstep = 1.0 / step
    numstep = 1 + (end - begin) / step
    ln = (numstep + 0.49).astype(np.int32)
    maxsteps = int(ln.max())
    tab = np.zeros((maxsteps + 2, ncols), dtype=np.float64)
    for i, c in enumerate(cols):
        c = np.asarray(c, dtype=np.float64)
        assert c.shape[0] == ln[i], (c.shape, ln[i])
        tab[: ln[i], i] = c
    if radial:
        for i in range(ncols):
            delta = tab[ln[i] - 1, i]
            if delta != 0.0:
                tab[: ln[i], i] -= delta
    tab2 = None
    if interp == "spline":
        tab2 = init_spline(tab, ln, step, radial)
    else:
        for i in range(ncols):
            n = int(ln[i])
            y = tab[:, i]
            if interp == "4point":         # init_fourpoint, imd_potential.c:1175-1191
                y[n] = 4 * y[n - 1] - 6 * y[n - 2] + 4 * y[n - 3] - y[n - 4]
                y[n + 1] = 10 * y[n - 1] - 20 * y[n - 2] + 15 * y[n - 3] - 4 * y[n - 4]
            else:                          # init_threepoint, imd_potential.c:1260-1276
                assert interp == "3point", interp
                y[n] = 3 * y[n - 1] - 3 * y[n - 2] + y[n - 3]
                y[n + 1] = 6 * y[n - 1] - 8 * y[n - 2] + 3 * y[n - 3]
    return PotTable(begin, end, step, invstep, ln, maxsteps, np.ascontiguousarray(tab),
                    None if tab2 is None else np.ascontiguousarray(tab2), interp)


def write_pot_table2(path: str, begin, end, step, cols) -> None:
    """IMD potential format 2: header '#F 2 ncols' / '#E', one info line per column
    (begin end step), then the values column after column (imd_potential.c:385-393)."""
    with open(path, "w") as f:
        f.write("#F 2 %d\n#E\n" % len(cols))
        for b, e, s in zip(begin, end, step):
            f.write("%.16e %.16e %.16e\n" % (b, e, s))
        for c in cols:
            f.write("\n")
            for v in c:
                f.write("%.16e\n" % v)


def read_pot_table2(path: str, ncols: int, radial: bool, interp: str = "3point") -> PotTable:
    """Reader for format 2 with the semantics of read_pot_table2 (imd_potential.c:394)."""
    with open(path) as f:
        lines = [ln for ln in f if not ln.startswith("#")]
    toks = " ".join(lines).split()
    vals = np.array(toks, dtype=np.float64)
    info = vals[: 3 * ncols].reshape(ncols, 3)
    begin, end, step = info[:, 0], info[:, 1], info[:, 2]
    numstep = 1 + (end - begin) / step
    ln = (numstep + 0.49).astype(np.int64)
    cols, p = [], 3 * ncols
    for i in range(ncols):
        cols.append(vals[p: p + ln[i]])
        p += ln[i]
    return _table_from_columns(begin, end, step, cols, radial, interp)


@dataclasses.dataclass
class EamSpec:
    """Analytic Johnson-style EAM: f(r)=f_e exp(-beta(r/r_e-1)) s(r),
    phi(r)=phi_e exp(-gamma(r/r_e-1)) s(r), s = quintic smoothstep to 0 on [0.9 r_c, r_c],
    F(rho) = -E_c (1-(alpha/beta) ln x) x^(alpha/beta) - 6 phi_e x^(gamma/beta), x=rho/rho_e."""
    name: str
    a: float          # lattice constant (Angstrom)
    r_c: float
    E_c: float
    phi_e: float
    alpha: float
    beta: float
    gamma: float
    mass: float       # IMD mass unit: amu * 1.0364e-4 is applied by the caller if wanted
    f_e: float = 1.0
    rho_e: float = 12.0
    rows: int = 2000
    r2_begin: float = 1.0

    @property
    def r_e(self) -> float:
        return self.a / np.sqrt(2.0)


CU = EamSpec("cu", a=3.615, r_c=4.95, E_c=3.54, phi_e=0.59, alpha=5.09, beta=5.85,
             gamma=8.0, mass=63.546)
AL = EamSpec("al", a=4.05, r_c=6.3, E_c=3.36, phi_e=0.30, alpha=4.60, beta=5.40,
             gamma=7.6, mass=26.9815)


def _smooth(r, r_c):
    t = np.clip((r - 0.9 * r_c) / (0.1 * r_c), 0.0, 1.0)
    return 1.0 - (10 * t**3 - 15 * t**4 + 6 * t**5)


def eam_columns(spec: EamSpec):
    """Return ((begin,end,step,col) for pair, rho) on an r^2 grid and embed on a rho grid."""
    r2 = np.linspace(spec.r2_begin, spec.r_c**2, spec.rows)
    step = (spec.r_c**2 - spec.r2_begin) / (spec.rows - 1)
    r = np.sqrt(r2)
    s = _smooth(r, spec.r_c)
    f = spec.f_e * np.exp(-spec.beta * (r / spec.r_e - 1.0)) * s
    phi = spec.phi_e * np.exp(-spec.gamma * (r / spec.r_e - 1.0)) * s
    rho_end = 3.0 * spec.rho_e
    rho = np.linspace(0.0, rho_end, spec.rows)
    rstep = rho_end / (spec.rows - 1)
    x = np.maximum(rho / spec.rho_e, 1e-300)
    ab, gb = spec.alpha / spec.beta, spec.gamma / spec.beta
    F = -spec.E_c * (1.0 - ab * np.log(x)) * x**ab - 6.0 * spec.phi_e * x**gb
    F[0] = 0.0
    return ((spec.r2_begin, spec.r_c**2, step, phi),
            (spec.r2_begin, spec.r_c**2, step, f),
            (0.0, rho_end, rstep, F))


def make_eam_tables(spec: EamSpec, outdir: Optional[str] = None, interp: str = "3point") -> Dict[str, PotTable]:
    """Tables for ntypes=1.  When ``outdir`` is given the three IMD format-2 files
    ``<name>.pair/.rho/.emb`` are written and the tables are re-read from the files, so
    that the in-memory values are exactly what the reference parses from the same text."""
    pair, rho, emb = eam_columns(spec)
    if outdir is not None:
        os.makedirs(outdir, exist_ok=True)
        out = {}
        for key, (b, e, s, c), radial in (("pair", pair, True), ("rho", rho, True), ("emb", emb, False)):
            path = os.path.join(outdir, "%s.%s" % (spec.name, key))
            write_pot_table2(path, [b], [e], [s], [c])
            out[key] = read_pot_table2(path, 1, radial, interp)
        return out
    return {
        "pair": _table_from_columns([pair[0]], [pair[1]], [pair[2]], [pair[3]], True, interp),
        "rho": _table_from_columns([rho[0]], [rho[1]], [rho[2]], [rho[3]], True, interp),
        "emb": _table_from_columns([emb[0]], [emb[1]], [emb[2]], [emb[3]], False, interp),
    }


@dataclasses.dataclass
class BinarySpec:
    """Two-species EAM (ntypes = 2) on the L1_2 structure (Cu3Au type, generate_fcc(6),
    imd_generate.c:483-490): every table column has its OWN r^2 grid (begin, end, step, length),
    the rho columns are asymmetric (col it*ntypes+jt != col jt*ntypes+it), as potaccess.h and
    imd_forces_nbl.c:557-572, 845-874 allow."""
    name: str
    a: float
    sa: EamSpec
    sb: EamSpec

    @property
    def masses(self):
        return (self.sa.mass, self.sb.mass)


NI3AL = BinarySpec("ni3al", a=3.57,
                   sa=EamSpec("ni", a=3.52, r_c=4.80, E_c=4.45, phi_e=0.74, alpha=4.98, beta=6.41, gamma=8.86,
                              mass=58.6934, rows=1700, r2_begin=1.0),
                   sb=EamSpec("alb", a=4.05, r_c=5.40, E_c=3.36, phi_e=0.30, alpha=4.60, beta=5.40, gamma=7.6,
                              mass=26.9815, rows=1900, r2_begin=1.2, rho_e=10.0))


def binary_columns(b: BinarySpec):
    """pair[4], rho[4], embed[2] column tuples (begin, end, step, values); column = it*2 + jt."""
    def radial(spec, fn, r_c, rows, r2_begin):
        r2 = np.linspace(r2_begin, r_c**2, rows)
        step = (r_c**2 - r2_begin) / (rows - 1)
        r = np.sqrt(r2)
        return (r2_begin, r_c**2, step, fn(r) * _smooth(r, r_c))
    sa, sb = b.sa, b.sb
    phi = lambda s: (lambda r: s.phi_e * np.exp(-s.gamma * (r / s.r_e - 1.0)))          # noqa: E731
    dens = lambda s, w: (lambda r: w * s.f_e * np.exp(-s.beta * (r / s.r_e - 1.0)))      # noqa: E731
    rc_x = 0.5 * (sa.r_c + sb.r_c)
    cross = lambda r: 0.5 * (phi(sa)(r) * 0.9 + phi(sb)(r) * 1.1)                        # noqa: E731
    pair = [radial(sa, phi(sa), sa.r_c, sa.rows, sa.r2_begin),
            radial(sa, cross, rc_x, 1800, 1.1), radial(sa, cross, rc_x, 1800, 1.1),
            radial(sb, phi(sb), sb.r_c, sb.rows, sb.r2_begin)]
    # rho column it*2+jt: density at an atom of type it from a neighbour of type jt
    rho = [radial(sa, dens(sa, 1.0), sa.r_c, sa.rows, sa.r2_begin),
           radial(sb, dens(sb, 0.85), 5.1, 1750, 1.15),
           radial(sa, dens(sa, 1.2), 4.9, 1650, 1.05),
           radial(sb, dens(sb, 1.0), sb.r_c, sb.rows, sb.r2_begin)]
    emb = []
    for s in (sa, sb):
        rho_end = 3.0 * s.rho_e
        x = np.linspace(0.0, rho_end, s.rows)
        xx = np.maximum(x / s.rho_e, 1e-300)
        ab, gb = s.alpha / s.beta, s.gamma / s.beta
        F = -s.E_c * (1.0 - ab * np.log(xx)) * xx**ab - 6.0 * s.phi_e * xx**gb
        F[0] = 0.0
        emb.append((0.0, rho_end, rho_end / (s.rows - 1), F))
    return pair, rho, emb


def make_eam_tables_binary(b: BinarySpec, outdir: Optional[str] = None, interp: str = "3point") -> Dict[str, PotTable]:
    """Like make_eam_tables for ntypes = 2 (files <name>.pair/.rho/.emb with 4/4/2 columns)."""
    pair, rho, emb = binary_columns(b)
    out = {}
    for key, cols, radial in (("pair", pair, True), ("rho", rho, True), ("emb", emb, False)):
        bg, en, st, vals = [c[0] for c in cols], [c[1] for c in cols], [c[2] for c in cols], [c[3] for c in cols]
        if outdir is not None:
            os.makedirs(outdir, exist_ok=True)
            path = os.path.join(outdir, "%s.%s" % (b.name, key))
            write_pot_table2(path, bg, en, st, vals)
            out[key] = read_pot_table2(path, len(cols), radial, interp)
        else:
            out[key] = _table_from_columns(bg, en, st, vals, radial, interp)
    return out


def l12_lattice(ncell: Sequence[int], a: float):
    """L1_2 positions and types as generate_fcc(6) makes them (imd_generate.c:440-560): fcc sites,
    corner atoms type 1, face atoms type 0."""
    nx, ny, nz = ncell
    pos, typ = [], []
    bu = a / 2
    for x in range(2 * nx):
        for y in range(2 * ny):
            for z in range(2 * nz):
                if (x + y + z) % 2 == 1:
                    continue
                typ.append(0 if (x % 2 == 1 or y % 2 == 1 or z % 2 == 1) else 1)
                pos.append(((x + 0.5) * bu, (y + 0.5) * bu, (z + 0.5) * bu))
    return np.array(pos), np.array(typ, np.int32), np.diag([nx * a, ny * a, nz * a]).astype(np.float64)


# ----------------------------------------------------------------------------- lattices

def fcc_lattice(ncell: Sequence[int], a: float, dtype=np.float64) -> Tuple[np.ndarray, np.ndarray]:
    """fcc positions (4 per cubic cell, basis offset a/4 as a generic interior point) and
    the orthorhombic box matrix (rows box_x, box_y, box_z)."""
    nx, ny, nz = ncell
    basis = np.array([[0, 0, 0], [0.5, 0.5, 0], [0.5, 0, 0.5], [0, 0.5, 0.5]], dtype=np.float64) + 0.25
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    cells = np.stack([ix, iy, iz], axis=-1).reshape(-1, 1, 3).astype(np.float64)
    pos = ((cells + basis[None, :, :]) * a).reshape(-1, 3).astype(dtype)
    box = np.diag([nx * a, ny * a, nz * a]).astype(np.float64)
    return pos, box


def maxwell_momenta(n: int, mass: float, temp_ev: float, seed: int = 4711) -> np.ndarray:
    """Gaussian momenta with <p^2/2m> = 3/2 kT per atom and zero total momentum
    (role of imd_maxwell.c; different RNG stream than drand48)."""
    rng = np.random.default_rng(seed)
    p = rng.normal(0.0, np.sqrt(mass * temp_ev), size=(n, 3))
    p -= p.mean(axis=0, keepdims=True)
    return p


def film_lattice(ncell: Sequence[int], a: float, vac_front: float, vac_rear: float):
    """fcc film: ncell cubic cells of material, vacuum of the given widths (Angstrom)
    before and behind it along x (the laser axis); box non-periodic in x."""
    pos, box = fcc_lattice(ncell, a)
    pos[:, 0] += vac_front
    box[0, 0] += vac_front + vac_rear
    return pos, box


# ----------------------------------------------------------------------------- IMD input

def write_param_file(path: str, spec: EamSpec, box_param: Sequence[int], maxsteps: int,
                     outfiles: str, potdir: str = ".", pbc=(1, 1, 1), starttemp: float = 0.0258,
                     timestep: float = 0.1, nbl_margin: float = 0.5, seed: int = 4711,
                     eng_int: int = 1, extra: str = "", ensemble: str = "nve",
                     cpu_dim=(1, 1, 1), coordname: str = "_fcc", masses=None) -> None:
    """IMD parameter file for the serial reference binary (keywords of imd_param.c)."""
    txt = io.StringIO()
    txt.write("coordname    %s\n" % coordname)
    txt.write("outfiles     %s\n" % outfiles)
    txt.write("ensemble     %s\n" % ensemble)
    txt.write("maxsteps     %d\n" % maxsteps)
    txt.write("timestep     %.17g\n" % timestep)
    masses = [spec.mass] if masses is None else list(masses)
    nt = len(masses)
    txt.write("ntypes       %d\ntotal_types  %d\n" % (nt, nt))
    txt.write("masses       %s\n" % " ".join("%.17g" % m for m in masses))
    txt.write("types        %s\n" % " ".join(str(t) for t in range(nt)))
    txt.write("box_param    %d %d %d\n" % tuple(box_param))
    txt.write("box_unit     %.17g\n" % spec.a)
    txt.write("pbc_dirs     %d %d %d\n" % tuple(pbc))
    txt.write("starttemp    %.17g\n" % starttemp)
    txt.write("eng_int      %d\ncheckpt_int  0\n" % eng_int)
    txt.write("seed         %d\n" % seed)
    txt.write("core_potential_file   %s\n" % os.path.join(potdir, spec.name + ".pair"))
    txt.write("embedding_energy_file %s\n" % os.path.join(potdir, spec.name + ".emb"))
    txt.write("atomic_e-density_file %s\n" % os.path.join(potdir, spec.name + ".rho"))
    txt.write("nbl_margin   %.17g\nnbl_size     1.3\n" % nbl_margin)
    txt.write("cpu_dim      %d %d %d\n" % tuple(cpu_dim))
    txt.write(extra)
    with open(path, "w") as f:
        f.write(txt.getvalue())


def write_config_file(path: str, pos: np.ndarray, mom: np.ndarray, mass, box: np.ndarray, types=None) -> None:
    """IMD ASCII configuration as read_atoms parses it (imd_io_3d.c:52-600; header as write_config writes it,
    imd_io.c:3357-3460): number type mass x y z vx vy vz.  The reader multiplies the velocities by the mass, so
    v = p / m is written; %.17g keeps every double.  With `box_from_header 1` the box comes from the #X #Y #Z
    lines (read_atoms then adds shiftx_front + shiftx_rear to box_x.x and shiftx_front to every x)."""
    n = pos.shape[0]
    mass = np.broadcast_to(np.asarray(mass, np.float64), (n,))
    types = np.zeros(n, np.int32) if types is None else np.asarray(types, np.int32)
    with open(path, "w") as f:
        f.write("#F A 1 1 1 3 3 0\n#C number type mass x y z vx vy vz\n")
        for tag, row in zip("XYZ", np.asarray(box, np.float64)):
            f.write("#%s %.17g %.17g %.17g\n" % (tag, row[0], row[1], row[2]))
        f.write("#E\n")
        for i in range(n):
            f.write("%d %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n"
                    % (i + 1, types[i], mass[i], pos[i, 0], pos[i, 1], pos[i, 2],
                       mom[i, 0] / mass[i], mom[i, 1] / mass[i], mom[i, 2] / mass[i]))


# ----------------------------------------------------------------------------- TTM inputs

AMU = 1.66053907e-27
KB = 1.380649e-23
HBAR = 1.054571817e-34
EMASS = 9.1093837015e-31


@dataclasses.dataclass
class EosTable:
    """One EOS interpolation table as nn_read_table parses it (imd_interpol.c:1170-1271):
    ``is`` x ``js`` scattered points (x, y, z) with x the outer index.  Ours are rectilinear:
    x[i] (density, kg/m^3), y[j], z[i, j].  ``diag`` (uint8 [is-1, js-1]) says which diagonal
    of each grid rectangle the reference's Delaunay triangulation (libnn, via Triangle) uses:
    0 = (i,j)-(i+1,j+1), 1 = (i+1,j)-(i,j+1); it is an INPUT of the device path because a
    triangulation of co-circular points is not unique ("EOS tables")."""
    x: np.ndarray
    y: np.ndarray
    z: np.ndarray
    diag: Optional[np.ndarray] = None

    def write(self, path: str) -> None:
        with open(path, "w") as f:
            f.write("%d %d\n" % (self.x.shape[0], self.y.shape[0]))
            f.write("%.17g %.17g %.17g %.17g\n" % (self.x[0], self.x[-1], self.y[0], self.y[-1]))
            for i in range(self.x.shape[0]):
                for j in range(self.y.shape[0]):
                    f.write("%.17g %.17g %.17g\n" % (self.x[i], self.y[j], self.z[i, j]))

    @staticmethod
    def read(path: str) -> "EosTable":
        with open(path) as f:
            toks = f.read().split()
        ni, nj = int(toks[0]), int(toks[1])
        v = np.array(toks[6:6 + 3 * ni * nj], dtype=np.float64).reshape(ni, nj, 3)
        return EosTable(np.ascontiguousarray(v[:, 0, 0]), np.ascontiguousarray(v[0, :, 1]),
                        np.ascontiguousarray(v[:, :, 2]))


def make_eos_tables(nrho: int = 48, nt: int = 96, z_free: float = 3.0, a_weight: float = 26.9815,
                    rho_range=(30.0, 4000.0), t_range=(30.0, 2.9e6)) -> Dict[str, EosTable]:
    """Synthetic free-electron-like Al EOS (the reference's EOS_*.txt are missing blobs):
    cve(rho,T) = c0 T / sqrt(T^2 + a(rho)^2)  [J/(K kg)],  ee = int_0^T cve = c0 (sqrt(T^2+a^2) - a)
    [J/kg], a = (3/pi^2) T_F(rho), c0 = 3/2 k_B n_e/rho -- Sommerfeld at low T, classical at high T,
    and ee consistent with cve as imd_ttm.c:1577-1579 requires.  Axes: density uniform; the cve table
    is over T (geometric), the ee table over sqrt(T) (uniform), the phase table over T_i."""
    rho = np.linspace(rho_range[0], rho_range[1], nrho)
    ne = z_free * rho / (a_weight * AMU)
    t_f = HBAR**2 * (3 * np.pi**2 * ne) ** (2.0 / 3.0) / (2 * EMASS) / KB
    a = (3.0 / np.pi**2) * t_f
    c0 = 1.5 * KB * z_free / (a_weight * AMU)
    t_geo = np.geomspace(t_range[0], t_range[1], nt)
    cve = c0 * t_geo[None, :] / np.sqrt(t_geo[None, :] ** 2 + a[:, None] ** 2)
    ts = np.linspace(np.sqrt(t_range[0]), np.sqrt(t_range[1]), nt)
    ee = c0 * (np.sqrt(ts[None, :] ** 4 + a[:, None] ** 2) - a[:, None])
    ti = np.linspace(t_range[0], 2.0e4, 8)
    phase = np.ones((nrho, ti.shape[0]))
    return {"cve": EosTable(rho, t_geo, cve), "ee": EosTable(rho, ts, ee), "phase": EosTable(rho, ti, phase)}


def make_tmm_tables(nbb: int = 64):
    """Synthetic stand-ins for the two data files do_tmm reads (imd_tmm.c:117-118; missing blobs upstream):
    eps_bb [nbb,3] = lambda (micrometres, ascending), Re, Im of the interband permittivity -- a Lorentzian
    bump around 0.8 um like aluminium's parallel-band absorption; K12 [1500,2] = the collision integrals
    K1, K2 at nu = (i+1)/30, smooth curves that join the asymptotes coded in tmm_K1/tmm_K2 (:667-696)."""
    lam = np.linspace(0.2, 2.0, nbb)
    w, w0, g = 1.0 / lam, 1.0 / 0.8, 0.25
    lor = 3.0 / (w0 * w0 - w * w - 1j * g * w)
    eps_bb = np.stack([lam, lor.real, lor.imag], axis=1)
    nu = (np.arange(1500) + 1.0) / 30.0
    k1 = 1.0 / (1.0 + 8.0 * nu * nu / 315.0)
    k2 = 1.0 / (1.0 + nu * nu / 6.0)
    return {"eps_bb": eps_bb, "K12": np.stack([k1, k2], axis=1)}


def write_tmm_files(tabs, outdir: str) -> None:
    """alu_eps_bb.dat (3 columns) and K12.dat (2 columns), one row per line (tmm_read_arr counts newlines)."""
    for fn, key in (("alu_eps_bb.dat", "eps_bb"), ("K12.dat", "K12")):
        with open(os.path.join(outdir, fn), "w") as f:
            for row in tabs[key]:
                f.write(" ".join("%.17g" % v for v in row) + "\n")


def write_eos_files(tabs: Dict[str, EosTable], outdir: str) -> None:
    """The file names init_ttm opens relative to the run directory's parent (imd_ttm.c:1130-1132)."""
    tabs["cve"].write(os.path.join(outdir, "EOS_cve_from_r_te.txt"))
    tabs["ee"].write(os.path.join(outdir, "EOS_ee_from_r_tesqrt.txt"))
    tabs["phase"].write(os.path.join(outdir, "EOS_phase_from_r_ti.txt"))


@dataclasses.dataclass
class TtmParams:
    """Run-time TTM parameters (keywords of imd_param.c:1979-2092) + the analytic source."""
    ttmdimx: int = 32
    fd_n_timesteps: int = 20
    fd_min_atoms: int = 10
    vlatdim: int = 8
    vlatbuffer: int = 2
    vlatdens: float = 2.665655433e+03       # globals.h:1226
    atomic_weight: float = 26.9815
    atomic_charge: float = 13.0
    I0: float = 5.0e15                      # W/m^2
    lam: float = 800e-9                     # m
    laser_t_0: float = 30e-15               # s
    laser_sigma_t: float = 20e-15           # s
    tmm_laser_threshold: float = 1e-5       # globals.h:479
    src_mu: float = 1.0 / 80.0              # 1/Angstrom, analytic Beer-Lambert source
    src_abs: float = 0.1
    Temin: float = 1e-3                     # globals.h:1196
    tmm_threshold: float = 10.0             # tmm_absorption_threshold: exp(-x) of the field is "absorbed"
    use_tmm: int = 0                        # 1: the reference's transfer-matrix source (do_tmm), 0: analytic

    def param_text(self) -> str:
        return ("fd_ext 1 1 1\nfd_n_timesteps %d\nttm_int 100000000\nfd_min_atoms %d\nttmdimx %d\n"
                "vlatdim %d\nvlatbuffer %d\natomic_weight %.17g\natomic_charge %.17g\nI0 %.17g\n"
                "lambda %.17g\nlaser_t_0 %.17g\nlaser_sigma_t %.17g\ntmm_threshold %.17g\n"
                % (self.fd_n_timesteps, self.fd_min_atoms, self.ttmdimx, self.vlatdim, self.vlatbuffer,
                   self.atomic_weight, self.atomic_charge, self.I0, self.lam, self.laser_t_0,
                   self.laser_sigma_t, self.tmm_threshold))
