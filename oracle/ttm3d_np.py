"""oracle/ttm3d_np.py -- TEST INFRASTRUCTURE: numpy restatement of the reference's 3-D TTM
diffusion sweep, do_DIFF (imd_ttm_3D.c:1603-1829) followed by ttm_fill_ghost_layers
(:3314-3504), single rank.  Pinned against the reference's own functions through
oracle/_ref/ref3d_diff (ref3d_driver.c) in tests/test_ttm3d_oracle_vs_reference.py.

Fields live on the extended lattice (nx+2, ny+2, nz+2), z fastest, like l1[i][j][k].
Only cells with natoms >= fd_min_atoms are updated (:1660); a neighbour with fewer atoms is
replaced by the cell itself (:1661-1666), which makes the face flux vanish.  DIRICHLET, FDTD
and COLRAD variants of the loop are compile-time options of the reference that its TTM target
does not set; they are not restated.
"""
import numpy as np

FIELDS = ("natoms", "temp", "fd_k", "fd_g", "Ce", "md_temp", "source", "U", "xi", "dens")
AMU = 1.660539040200000e-27          # constants.h


def fill_ghost_layers(f, pbc):
    """ttm_fill_ghost_layers (:3314-3504) on one rank.  Every plane type (mpi_xplane_block,
    mpi_yplane_block, mpi_zplane_block, :3148-3312) spans only the INTERIOR of the other two
    directions: faces are filled, edges and corners never (the 7-point stencil does not read
    them).  Transferred members: natoms, temp, fd_k (:2990-3006).  Non-periodic direction: ghost
    natoms = 0 over the face (:3384-3391) -- what the reference does on the surface ranks; its
    single-rank y branch does nothing and its single-rank z branch blocks in MPI_Recv
    (:3398-3451, :3467-3500), neither of which is restated."""
    nat, temp, k = f["natoms"], f["temp"], f["fd_k"]
    for ax in range(3):
        def sl(i):
            s = [slice(1, -1)] * 3
            s[ax] = i
            return tuple(s)
        if pbc[ax]:
            for a in (nat, temp, k):
                a[sl(-1)] = a[sl(1)]
                a[sl(0)] = a[sl(-2)]
        else:
            nat[sl(0)] = 0
            nat[sl(-1)] = 0


def diff_sweep(f, h, tau, fd_min_atoms, diff_substeps, laser_active):
    """One do_DIFF(tau) (:1603-1829).  Updates f in place (temp, U, xi); returns the Eabs increment."""
    nat, T, K = f["natoms"], f["temp"], f["fd_k"]
    c = (slice(1, -1),) * 3
    fd_vol = h[0] * h[1] * h[2]
    xi_fac = 26.9815 * AMU / 3.0 * 1e30 / float(diff_substeps)
    inv = [1.0 / (h[0] * h[0]), 1.0 / (h[1] * h[1]), 1.0 / (h[2] * h[2])]
    eabs = 0.0
    if laser_active:
        # the reference accumulates cell by cell in i,j,k order (:1651); same order here
        eabs = float(np.cumsum((f["source"][c] * fd_vol * tau).ravel())[-1]) if nat[c].size else 0.0
    act = nat[c] >= fd_min_atoms
    Tc, Kc = T[c], K[c]

    def nb(a, ax, d):
        s = [slice(1, -1)] * 3
        s[ax] = slice(2, None) if d > 0 else slice(0, -2)
        return a[tuple(s)]
    terms = None
    for ax in range(3):
        hi_ok = nb(nat, ax, +1) >= fd_min_atoms
        lo_ok = nb(nat, ax, -1) >= fd_min_atoms
        Thi = np.where(hi_ok, nb(T, ax, +1), Tc)
        Tlo = np.where(lo_ok, nb(T, ax, -1), Tc)
        Khi = np.where(hi_ok, nb(K, ax, +1), Kc)
        Klo = np.where(lo_ok, nb(K, ax, -1), Kc)
        a = ((Kc + Khi) / 2 * (Thi - Tc) * inv[ax])
        b = ((Kc + Klo) / 2 * (Tc - Tlo) * inv[ax])
        terms = (a - b) if terms is None else ((terms + a) - b)
    with np.errstate(divide="ignore", invalid="ignore"):
        Ce = f["Ce"][c]
        Tn = tau / Ce * (terms - f["fd_g"][c] * (Tc - f["md_temp"][c]) + f["source"][c]) + Tc
        Un = f["U"][c] + (Tn - Tc) * Ce * fd_vol / nat[c].astype(np.float64)
        dxi = (Tn - f["md_temp"][c]) * xi_fac * f["fd_g"][c] / f["md_temp"][c] / f["dens"][c]
    newT = np.where(act, Tn, Tc)
    f["U"][c] = np.where(act, Un, f["U"][c])
    f["xi"][c] = np.where(act, f["xi"][c] + dxi, f["xi"][c])
    T[c] = newT
    return eabs


def run(f, h, tau, fd_min_atoms, diff_substeps, laser_active, nsweeps, pbc):
    """nsweeps x { do_DIFF; ttm_fill_ghost_layers }, as the sub-step loop of calc_ttm (:193-200)
    with the wide-range properties frozen (the standalone sweep of BASELINE.json configs[4])."""
    eabs = 0.0
    for _ in range(nsweeps):
        eabs += diff_sweep(f, h, tau, fd_min_atoms, diff_substeps, laser_active)
        fill_ghost_layers(f, pbc)
    return eabs


def make_case(shape, seed=0, hole=True, pbc=(0, 1, 1), fd_min_atoms=20):
    """Synthetic lattice: hot spot in a metal film, a vacuum region (inactive cells) when `hole`."""
    nx, ny, nz = shape
    rng = np.random.default_rng(seed)
    e = (nx + 2, ny + 2, nz + 2)
    f = {}
    nat = rng.integers(40, 80, e)
    if hole:
        x = np.arange(e[0])[:, None, None]
        y = np.arange(e[1])[None, :, None]
        z = np.arange(e[2])[None, None, :]
        nat = np.where(x < 2 + nx // 6, rng.integers(0, fd_min_atoms, e), nat)            # sparse front
        nat = np.where((x - nx // 2) ** 2 + (y - ny // 2) ** 2 + (z - nz // 2) ** 2 < (min(shape) // 4) ** 2, 0, nat)
    f["natoms"] = nat.astype(np.int32)
    f["temp"] = 0.05 + 0.5 * rng.random(e)
    f["fd_k"] = 10.0 + 20.0 * rng.random(e)
    f["fd_g"] = 2e-4 + 4e-4 * rng.random(e)
    f["Ce"] = 2e-3 + 1e-3 * rng.random(e)
    f["md_temp"] = 0.025 + 0.01 * rng.random(e)
    f["source"] = 1e-3 * rng.random(e)
    f["U"] = rng.random(e)
    f["xi"] = np.zeros(e)
    f["dens"] = 2000.0 + 700.0 * rng.random(e)
    return f
