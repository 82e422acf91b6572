/* oracle/ttm_oracle.h -- TEST INFRASTRUCTURE ONLY (see oracle.h for the rules).
 *
 * CPU restatement of the reference's 1-D two-temperature model, imd_ttm.c (calc_ttm and
 * everything it calls) + move_atoms_ttm (imd_integrate.c:581-811), written from scratch.
 * The reference's result depends on its MPI decomposition (per-rank ghost cells, the
 * virtual-lattice owner rank, imd_ttm.c:1548-1552), so the restatement keeps the
 * reference's layout: P "ranks" each holding L = G/P local cells plus two ghost cells in
 * two lattices l1/l2 that are swapped by pointer, and performs the reference's exchanges
 * as plain copies.  P is a parameter (num_cpus of the reference run being matched).
 *
 * Parity status: PINNED against dumps of the unmodified reference built with -DMPI over
 * shmpi/ (oracle/_ref/imd_ref_mpi_eam_ttm; tests/golden/ttm_*.npz,
 * tests/test_ttm_oracle_vs_reference.py).  Two substitutions are shared by both sides and
 * documented in DESIGN.md: the laser source is the analytic Beer-Lambert term of
 * oracle/ref_hooks_ttm.c (the reference's transfer-matrix solver needs data files that are
 * not in its tree), and the EOS tables are synthetic.  Where the reference reads outside
 * an array (natomsglobal[-1], [G]; flux arrays likewise: imd_ttm.c:453,462,1246,1249) the
 * restatement reads 0.
 */
#ifndef IMD_TTM_ORACLE_H
#define IMD_TTM_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* rectilinear EOS table + the diagonal choice of the reference's Delaunay triangulation */
typedef struct {
  int nx, ny;
  const double *x, *y, *z;          /* z[i*ny + j] */
  const unsigned char *diag;        /* [(nx-1)*(ny-1)]: 0 = (i,j)-(i+1,j+1), 1 = (i+1,j)-(i,j+1) */
} orc_eos;

typedef struct {
  int    G, P;                      /* ttmdimx, num_cpus */
  int    fd_n_timesteps, fd_min_atoms, vlatdim, vlatbuffer;
  double vlatdens, atomic_weight, atomic_charge, Temin;
  double timestep;
  double box_x, box_y, box_z;       /* orthorhombic box edge lengths */
  double r2_cut_pair;               /* pair_pot.end[0] (neighvol, imd_ttm.c:981) */
  double I0, lambda, laser_t_0, laser_sigma_t, tmm_laser_threshold, src_mu, src_abs;
} orc_ttm_params;

typedef struct orc_ttm orc_ttm;

/* The reference's own laser source: transfer-matrix solution of the Helmholtz equation over the FD
 * cells with a wide-range permittivity (do_tmm, tmm_get_epsilon, Runge5, EE, tmm_K1/K2; imd_tmm.c:37-98,
 * 193-445, 503-570, 667-696).  eps_bb[nbb][3] = rows of alu_eps_bb.dat, K12[nk][2] = rows of K12.dat,
 * phase = EOS_phase_from_r_ti table.  Without this call the analytic source of ref_hooks_ttm.c is used. */
typedef struct {
  double tmm_threshold;             /* tmm_absorption_threshold */
  double laser_t_1, laser_sigma_t1; /* second pulse (globals.h:491-495), 0 0 = none */
  int    nbb, nk;
  const double *eps_bb, *K12;
  orc_eos phase;
} orc_tmm_params;
void orc_ttm_set_tmm(orc_ttm *t, const orc_tmm_params *q);

orc_ttm *orc_ttm_create(const orc_ttm_params *p, const orc_eos *cve, const orc_eos *ee);
void     orc_ttm_destroy(orc_ttm *t);

/* calc_ttm (imd_ttm.c:75-156) for MD step `steps`.  Atom arrays in caller order;
 * fdi[n] is the per-atom slab index of the previous call (p->fdi, in/out).
 * Returns 0, or <0 when the reference would call error() (message via orc_ttm_error). */
int  orc_ttm_step(orc_ttm *t, int steps, long n, const double *pos, const double *mom,
                  const double *mass, const int *type, const int *numneighs, int *fdi);
const char *orc_ttm_error(const orc_ttm *t);

/* results: global arrays of length G assembled from the per-rank lattices (l1 after the step) */
#define ORC_TTM_NFIELD 13   /* natoms natoms_old temp md_temp xi dens fd_k fd_g Z ne Ce U source */
void orc_ttm_get_cells(const orc_ttm *t, double *fields /* [G][13] */);
void orc_ttm_get_arrays(const orc_ttm *t, double *xiarr /*[G]*/, double *vcom /*[3][G]*/,
                        double *natoms /*[G]*/, double *fluxl /*[G]*/, double *fluxr /*[G]*/);
/* {diff_substeps, max_dt_ttm, last_active_cell_global, cur_vlattice_proc, Eabs (all ranks),
 *  tot_elec_energy (all ranks), fd_h.x, fd_vol, neighvol, laser_active, tmm_time} */
void orc_ttm_get_scalars(const orc_ttm *t, double out[16]);
void orc_ttm_get_vlattice(const orc_ttm *t, double *v /* [vlatdim][6]: temp md_temp fd_k fd_g Ce Z */);

/* move_atoms_ttm (imd_integrate.c:581-811, TTM1D + PDECAY branch); returns tot_kin_energy */
double orc_move_atoms_ttm(const orc_ttm *t, long n, double *pos, double *mom, const double *force,
                          const double *mass, const int *type, const int *vtype,
                          const double *restrictions, int pdecay, double xipdecay,
                          double ramp_start, double ramp_end);

/* single-function probes */
double orc_eos_interp(const orc_eos *e, double x, double y);                 /* lpi_interpolate_point */
double orc_eos_te_from_r_ee(const orc_eos *ee, double r, double e);          /* imd_ttm.c:2699-2724 */
double orc_ttm_mean_charge(double temp_K, double rho, double z0, double am); /* imd_ttm.c:2241-2280 */
double orc_ttm_kappa(double Te, double Ti, double Ne, double Z, double omega_laser, int laser_active);
double orc_ttm_gamma(double Te, double Ti, double Ne, double Z, double omega_laser, int laser_active);

/* 3-D diffusion sweep (imd_ttm_3D.c:1603-1829), standalone: see ttm3d_oracle.c */

#ifdef __cplusplus
}
#endif
#endif
