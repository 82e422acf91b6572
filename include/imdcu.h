/* include/imdcu.h -- C ABI of the B200-native EAM + TTM hot path (libimdcu.so).
 *
 * This is the drop-in boundary: plain C, plain pointers and sizes, no CUDA or torch
 * types.  The reference (fmqeisfeld/IMD, "lasMD") has no plugin/FFI interface; its
 * "operator API" for this path is the set of global-state C functions the main loop
 * calls.  Each entry point below names the reference function it
 * replaces (file:line in the reference tree); INTEGRATION.md shows the thin C stubs a
 * maintainer adds to IMD (a new `cuda` make-target substring) that forward
 * calc_forces / move_atoms_nve / check_nblist / calc_ttm to these calls.
 *
 * Conventions: every function returns 0 on success, a negative IMDCU_E* code on error
 * and never calls exit(); imdcu_last_error() returns the message (the IMD-side stub
 * passes it to imderror(), imd_misc.c:78-98).  All host pointers are borrowed for the
 * duration of the call only.  One context drives one GPU from one host thread (IMD is
 * MPI_THREAD_FUNNELED, imd.c:62).  There is NO CPU fallback: without a CUDA device
 * imdcu_create fails.
 */
#ifndef IMDCU_H
#define IMDCU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IMDCU_ABI_VERSION 3

#define IMDCU_OK            0
#define IMDCU_EINVAL       -1   /* bad argument / call order                       */
#define IMDCU_ECUDA        -2   /* CUDA runtime error (message has the details)    */
#define IMDCU_ENODEVICE    -3   /* no usable CUDA device                            */
#define IMDCU_EOVERFLOW    -4   /* "neighbor table full - increase nbl_size"       */
#define IMDCU_ENCCL        -5   /* NCCL error                                      */
#define IMDCU_ESHORT       -6   /* reserved                                         */

typedef struct imdcu_ctx imdcu_ctx;

/* which tabulated function set (globals pair_pot, rho_h_tab, embed_pot; types.h:537) */
#define IMDCU_TAB_PAIR  0
#define IMDCU_TAB_RHO   1
#define IMDCU_TAB_EMBED 2

/* ------------------------------------------------------------------ life cycle */
int  imdcu_abi_version(void);
/* number of CUDA devices this process sees (a host with one rank per GPU picks rank % count) */
int  imdcu_device_count(int *n);
/* device: CUDA ordinal.  Replaces nothing in the reference (there is no device). */
int  imdcu_create(imdcu_ctx **ctx, int device);
int  imdcu_destroy(imdcu_ctx *ctx);
const char *imdcu_last_error(const imdcu_ctx *ctx);   /* ctx may be NULL: create errors */

/* ------------------------------------------------------------------ multi-GPU (one context per rank) */
/* Domain decomposition in x slabs, cpu_dim = nranks 1 1 (setup_mpi_topology, imd_geom_mpi_3d.c;
 * the halo exchange of imd_comm_force_3d.c and the MPI_Allreduce calls of the path become NCCL
 * send/recv/allreduce).  id: an ncclUniqueId as 128 opaque bytes, created on one rank with
 * imdcu_comm_unique_id and distributed by the host (MPI_Bcast in IMD, torch.distributed in
 * bench.py).  Must be called before imdcu_set_box; nranks == 1 is a no-op.  Every rank then
 * uploads the atoms of its own slab (imdcu_comm_rank gives the owned global x cell layers
 * [xoff, xoff+ldimx)); atoms that cross a slab boundary migrate at neighbour-list rebuilds
 * (send_atoms of fix_cells, imd_fix_cells_3d.c:43-306). */
int  imdcu_comm_unique_id(char id[128]);
int  imdcu_comm_init(imdcu_ctx *ctx, const char id[128], int rank, int nranks);
int  imdcu_comm_rank(const imdcu_ctx *ctx, int *rank, int *nranks, int *xoff, int *ldimx);
/* Slabs of equal ATOM count instead of equal width (the static part of the reference's load balancing
 * along x, imd_main_3d.c:864-914 -- a film has vacuum in front and behind, equal-width slabs leave the end ranks
 * short of atoms).  imdcu_balance_cuts is host arithmetic only (no context, no GPU): from the number of atoms per
 * global x cell layer (the host's histogram, summed over the ranks) it puts cut r on the layer boundary where the
 * running count is closest to r/nranks of the total, every slab keeping >= min_layers layers; xcut has nranks + 1
 * entries, xcut[0] = 0, xcut[nranks] = gdimx.  imdcu_comm_set_cuts installs such boundaries: after imdcu_set_box,
 * before imdcu_upload_atoms, the same array on every rank; imdcu_comm_rank then reports the new [xoff, xoff+ldimx). */
int  imdcu_balance_cuts(int nranks, int gdimx, const int64_t *atoms_per_layer, int min_layers, int *xcut);
int  imdcu_comm_set_cuts(imdcu_ctx *ctx, const int *xcut);

/* ------------------------------------------------------------------ set-up */
/* Box and cell grid.  Replaces make_box (imd_geom_3d.c:52-108) and init_cells
 * (imd_geom_3d.c:118-330): box = rows box_x, box_y, box_z; tbox and heights are
 * derived with the reference's operation order.  Must be called after the radial
 * tables are set (cellsz comes from their r2 end, imd_potential.c:406).
 * cell_divisor = "cd" of init_cells (cpu_dim * fd_ext under TTM), NULL for 1 1 1. */
int  imdcu_set_box(imdcu_ctx *ctx, const double box[9], const int pbc_dirs[3],
                   double nbl_margin, double nbl_size, const int cell_divisor[3]);
/* One tabulated function set, layout of pot_table_t AFTER read_pot_table +
 * init_threepoint: table has (maxsteps+2) rows x ncols, row-major.
 * Replaces the host-side table access of potaccess.h:323-354,465-495,591-621. */
int  imdcu_set_table(imdcu_ctx *ctx, int which, int ncols, int maxsteps,
                     const double *begin, const double *end, const double *step,
                     const double *invstep, const int *len, const double *table);
/* Table interpolation of the whole run.  The reference chooses it at compile time (make-target
 * substrings 4point / spline -> -DFOURPOINT / -DSPLINE, Makefile:917-923; potaccess.h:24-36):
 *   IMDCU_INTERP_3POINT  PAIR_INT2 / VAL_FUNC2 / DERIV_FUNC2     (potaccess.h:323-354,465-495,591-621), default
 *   IMDCU_INTERP_4POINT  PAIR_INT3 / VAL_FUNC3 / DERIV_FUNC3     (:365-413,503-540,630-667)
 *   IMDCU_INTERP_SPLINE  PAIR_INT_SP / VAL_FUNC_SP / DERIV_FUNC_SP (:421-457,547-583,675-711)
 * Call before imdcu_set_table; the tables must then be the ones init_fourpoint / init_spline
 * finished (imd_potential.c:1175-1248), and in spline mode every table needs its second-derivative
 * table (pot_table_t.table2, same shape as table) through imdcu_set_table2. */
#define IMDCU_INTERP_3POINT 0
#define IMDCU_INTERP_4POINT 1
#define IMDCU_INTERP_SPLINE 2
int  imdcu_set_interp(imdcu_ctx *ctx, int mode);
int  imdcu_set_table2(imdcu_ctx *ctx, int which, const double *table2);
/* ntypes, timestep, restrictions[3*nvtypes] (globals.h; imd_integrate.c:208-213) */
int  imdcu_set_params(imdcu_ctx *ctx, int ntypes, double timestep,
                      int nvtypes, const double *restrictions);
/* PDECAY rear absorber of the integrator (imd_integrate.c:233-239, 709-716) */
int  imdcu_set_pdecay(imdcu_ctx *ctx, int enabled, double xipdecay,
                      double ramp_start, double ramp_end);

/* Kernel-selection knobs (no reference counterpart; results are the same for every setting, the
 * parity tests run all of them): "lpa" lanes per atom of the EAM passes (0 = chosen by system size,
 * else 1, 2, 4, 8), "fused_table" (1: interleaved pair+rho rows in pass 1 when both tables share their
 * r^2 grid), "ttm_overlap" (1: imdcu_run_ttm runs the lattice kernel beside EAM pass 2),
 * "ttm_lat_smem", "ttm3d_tiled".  The IMDCU_LPA / IMDCU_FUSED_TABLE / ... environment variables set
 * the defaults when the context is created. */
int  imdcu_set_tuning(imdcu_ctx *ctx, const char *key, int value);

/* ------------------------------------------------------------------ atoms */
/* Upload all owned atoms (AoS pos/mom as ORT/IMPULS, makros.h:143-145).  Invalidates
 * the neighbour list (have_valid_nbl = 0, globals.h:179). */
int  imdcu_upload_atoms(imdcu_ctx *ctx, int64_t n, const int32_t *id, const int32_t *type,
                        const int32_t *vtype, const double *mass,
                        const double *pos, const double *mom);
int64_t imdcu_num_atoms(const imdcu_ctx *ctx);
/* New positions and momenta for the SAME atoms in the device's current order (the order
 * imdcu_download_atoms returns).  For hosts that keep the integrator or apply external
 * modifications; the neighbour list stays valid until check_nblist says otherwise. */
int  imdcu_update_atoms(imdcu_ctx *ctx, const double *pos, const double *mom);
/* Download in the device's current (cell-sorted) order; any pointer may be NULL.
 * This is the "sync to host" the IMD writers need (imd_io_3d.c:721-979). */
int  imdcu_download_atoms(imdcu_ctx *ctx, int32_t *id, int32_t *type, double *mass,
                          double *pos, double *mom, double *force, double *epot,
                          double *rho, double *dF, int32_t *numneighs, int32_t *cell3);

/* VSORTE of the owned atoms, same order as imdcu_download_atoms (write_atoms_config prints it, imd_io_3d.c:778) */
int  imdcu_download_vtypes(imdcu_ctx *ctx, int32_t *vtype);

/* Spatial distributions on the device (write_distrib, imd_distrib.c:33-218): the per-rank part of
 * make_distrib_density (:377-411) and of make_write_distrib_select's atom loop (:678-716) with dist_Ekin_fun
 * (:219), dist_Epot_fun (:230) and dist_vxavg_fun (:364).  dim / ll / ur are dist_dim, dist_ll, dist_ur; the bin
 * of an atom is (int)(dim / (ur - ll) * (x - ll)) per direction with C's truncation, flat index
 * (nx * dim[1] + ny) * dim[2] + nz, atoms outside are skipped.  Output: per-bin SUMS over this rank's atoms
 * (count, p.p / 2m, POTENG, p_x / m), accumulated in double and returned as float like the reference's arrays
 * before its MPI_Reduce; any pointer may be NULL.  Division by the count, the reduction over ranks and the file
 * formats stay the host's (:720-800).  dist_int output then downloads 4 dist_size floats instead of every atom. */
int  imdcu_distrib(imdcu_ctx *ctx, const int dim[3], const double ll[3], const double ur[3],
                   float *num, float *ekin, float *epot, float *vxavg);

/* ------------------------------------------------------------------ hot path */
/* calc_forces (imd_forces_nbl.c:295-1054): rebuilds cells + neighbour list when the
 * list is invalid (fix_cells, make_nblist), then pair/density pass, embedding, EAM
 * force pass.  out8 = {tot_pot_energy, virial, vir_xx, vir_yy, vir_zz, vir_xy, vir_yz,
 * vir_zx} as in imd_forces_nbl.c:997-1016 (only the first two are filled: no P_AXIAL). */
int  imdcu_calc_forces(imdcu_ctx *ctx, int steps, double out8[8]);
/* move_atoms_nve (imd_integrate.c:35-561) fused with check_nblist
 * (imd_forces_nbl.c:1062-1092).  out8[0] = tot_kin_energy; *have_valid_nbl as the
 * reference's global after check_nblist. */
int  imdcu_move_atoms_nve(imdcu_ctx *ctx, double out8[8], int *have_valid_nbl);
/* the main-loop segment calc_forces -> move_atoms -> check_nblist
 * (imd_main_3d.c:406,564,812) for nsteps steps without returning to the host code;
 * energies[2*s] = tot_pot_energy, energies[2*s+1] = tot_kin_energy (may be NULL). */
int  imdcu_run_nve(imdcu_ctx *ctx, int first_step, int nsteps, double *energies);

/* ------------------------------------------------------------------ non-reflecting boundary */
/* The NRB state init_nrb leaves per atom (imd_nrb.c:125-768; cell members isnrbbnd, nrbid[12],
 * refpos: makros.h:192-200,257), for every boundary atom (bnd = NRBBND 1..3) and every atom one of
 * them lists as a lattice-site neighbour (bnd = 0):
 *   id[nent]        atom numbers (NUMMER; the reference counts from 1, 0 marks an unset site)
 *   bnd[nent]       NRBBND
 *   nrbi[nent][12]  NRBI: atom numbers of the twelve fcc sites, 0 = unset (rows of bnd = 0 are ignored)
 *   refpos[nent][3] REF_POS
 * nrb_k is the parameter of that name (imd_param.c:2109; init_nrb takes sqrt(nrb_k / 26.9815)).
 * Call after imdcu_upload_atoms.  From then on imdcu_calc_forces(steps) runs nrb_forces
 * (imd_nrb.c:792-1129; call site imd_forces_nbl.c:1019-1021, steps > 1 only) and both integrators skip
 * the kick of boundary atoms (imd_integrate.c:233-259, 712-751).  startup_images != 0 also reproduces
 * what the reference does to boundary momenta for steps <= 1 (its send_forces adds the momenta of the
 * buffer-cell copies, imd_comm_force_3d.c:1046-1052,1157-1168: momentum times the number of periodic
 * y/z images of the atom's cell, for cpu_dim = (P,1,1)).  nent = 0 switches NRB off.
 * Several GPUs: every rank passes the full table; a boundary atom and its listed neighbours must be
 * owned by the same rank (they are a/2 apart in x; slabs are cells wide) -- otherwise the next
 * list build or imdcu_synchronize returns an error. */
int  imdcu_set_nrb(imdcu_ctx *ctx, double nrb_k, int64_t nent, const int32_t *id, const int32_t *bnd,
                   const int32_t *nrbi, const double *refpos, int startup_images);

/* ------------------------------------------------------------------ two-temperature model (1-D) */
/* Run-time parameters of the 1-D TTM: the keywords of imd_param.c:1979-2092 and the globals
 * init_ttm / tmm_init derive from them (imd_ttm.c:955-1201, imd_tmm.c:98-190).
 * nranks_ref: the reference's result depends on its MPI decomposition (per-rank ghost cells,
 * the virtual-lattice owner rank, imd_ttm.c:1548-1552); the device reproduces a run with
 * num_cpus = nranks_ref (>= 1, must divide ttmdimx).  The laser source is the analytic
 * Beer-Lambert term documented in DESIGN.md (the reference's transfer-matrix solver do_tmm
 * needs data files that are not in its tree): src_mu [1/Angstrom], src_abs = absorbed fraction. */
typedef struct {
  int    ttmdimx, nranks_ref;
  int    fd_n_timesteps, fd_min_atoms, vlatdim, vlatbuffer;
  double vlatdens, atomic_weight, atomic_charge, Temin;
  double I0, lambda, laser_t_0, laser_sigma_t, tmm_laser_threshold, src_mu, src_abs;
} imdcu_ttm_params;

#define IMDCU_EOS_CVE 0   /* EOS_cve_from_r_te.txt    : (rho [kg/m^3], Te [K])       -> J/(K kg) */
#define IMDCU_EOS_EE  1   /* EOS_ee_from_r_tesqrt.txt : (rho [kg/m^3], sqrt(Te [K])) -> J/kg     */
#define IMDCU_EOS_PHASE 2 /* EOS_phase_from_r_ti.txt  : (rho [kg/m^3], Ti [K])       -> phase id; only do_tmm reads it */
/* One EOS interpolation table (nn_read_table, imd_interpol.c:1170-1271).  The point set must be
 * rectilinear: x[nx], y[ny], z[nx*ny] with x the outer index.  diag[(nx-1)*(ny-1)] tells which
 * diagonal the reference's Delaunay triangulation (libnn delaunay_build) put into each grid
 * rectangle: 0 = (i,j)-(i+1,j+1), 1 = (i+1,j)-(i,j+1); NULL = all 0.  Replaces
 * lpi_interpolate_point at imd_ttm.c:2517,2568. */
int  imdcu_ttm_set_eos(imdcu_ctx *ctx, int which, int nx, int ny, const double *x, const double *y,
                       const double *z, const unsigned char *diag);
/* The reference's own laser source instead of the analytic one: do_tmm (imd_tmm.c:193-445), the
 * transfer-matrix solution of the Helmholtz equation over the FD cells with the wide-range permittivity
 * of tmm_get_epsilon (:503-570), evaluated in every diffusion sub-step.  eps_bb[nbb][3] = the rows of
 * alu_eps_bb.dat (lambda [um] ascending, Re, Im), K12[nk][2] = the rows of K12.dat (nk >= 1500), as
 * tmm_init reads them (:117-118); tmm_threshold = tmm_absorption_threshold; laser_t_1 / laser_sigma_t1 =
 * the second pulse (0 0: none).  Needs IMDCU_EOS_PHASE.  Call before imdcu_ttm_init. */
int  imdcu_ttm_set_tmm(imdcu_ctx *ctx, double tmm_threshold, double laser_t_1, double laser_sigma_t1,
                       int nbb, const double *eps_bb, int nk, const double *K12);
/* init_ttm + tmm_init (imd.c:237,245).  Needs box, tables, params and both EOS tables. */
int  imdcu_ttm_init(imdcu_ctx *ctx, const imdcu_ttm_params *p);
/* calc_ttm (imd_ttm.c:75-156) for MD step `steps`, after imdcu_calc_forces of the same step
 * (it consumes numneighs).  out16 (may be NULL: no host synchronisation) = {diff_substeps,
 * max_dt_ttm, last_active_cell_global, cur_vlattice_proc, Eabs, tot_elec_energy, fd_h.x, fd_vol,
 * neighvol, laser_active, tmm_time, error flag, 0...}. */
int  imdcu_calc_ttm(imdcu_ctx *ctx, int steps, double out16[16]);
/* calc_forces(steps) and calc_ttm() of one main-loop iteration (imd_main_3d.c:406, :534) in one call, so that the
 * lattice kernel of calc_ttm runs beside the second EAM pass: same results as the two calls above, ~1.3 ms less per
 * step at 4 M atoms.  For hosts that keep the reference's three call sites: do this in calc_forces and let calc_ttm
 * hand out the cached out16 (integration/imd_cuda.c). */
int  imdcu_calc_forces_ttm(imdcu_ctx *ctx, int steps, double out8[8], double out16[16]);
/* move_atoms_ttm (imd_integrate.c:581-811, TTM1D [+PDECAY]) fused with check_nblist */
int  imdcu_move_atoms_ttm(imdcu_ctx *ctx, double out8[8], int *have_valid_nbl);
/* main-loop segment calc_forces -> calc_ttm -> move_atoms_ttm -> check_nblist
 * (imd_main_3d.c:406,534,564,812); energies[2*s], [2*s+1] = tot_pot_energy, tot_kin_energy */
int  imdcu_run_ttm(imdcu_ctx *ctx, int first_step, int nsteps, double *energies);
/* FD lattice l1 after the last calc_ttm, assembled over the emulated ranks: cells[G*13] =
 * {natoms, natoms_old, temp, md_temp, xi, dens, fd_k, fd_g, Z, ne, Ce, U, source} per cell;
 * xiarr[G]; vcom[3*G]; vlat[vlatdim*6] = {temp, md_temp, fd_k, fd_g, Ce, Z}.  Any may be NULL.
 * This is what ttm_writeout (imd_ttm.c:1683) needs on the host. */
int  imdcu_download_ttm(imdcu_ctx *ctx, double *cells, double *xiarr, double *vcom, double *vlat);

/* ------------------------------------------------------------------ 3-D TTM lattice sweep */
/* The 3-D variant of the FD step (imd_ttm_3D.c; BASELINE.json configs[4]).  The lattice lives on
 * the device as SoA fields of ttm_Element (types.h:441-492).  All host arrays are the EXTENDED
 * lattice (nx+2)*(ny+2)*(nz+2), z fastest, exactly l1[i][j][k] with its ghost layer.
 * imdcu_ttm3d_init  : the lattice set-up of init_ttm (imd_ttm_3D.c:1040-1083): interior size,
 *                     fd_h, pbc_dirs, fd_min_atoms.  Needs only imdcu_create.
 * imdcu_ttm3d_upload: fills both lattices (l1, l2) from host fields; NULL leaves a field as is.
 * imdcu_ttm3d_diff  : nsweeps x { do_DIFF(tau) (imd_ttm_3D.c:1603-1829); ttm_fill_ghost_layers()
 *                     (:3314-3504) } = the FD sub-step loop of calc_ttm (:193-200) with the
 *                     wide-range properties frozen.  diff_substeps enters xi_fac (:1614).  Open
 *                     directions get ghost natoms = 0 (the reference's surface-rank branch).
 *                     *Eabs (may be NULL) = Eabs_local accumulated since init (:1648-1652).
 * Several GPUs: after imdcu_comm_init(nranks > 1) the lattice given here is this rank's x slab
 *                     (cpu_dim = N 1 1): nx local planes, pbc[0] is the global flag; the x ghost
 *                     planes then come from the x-neighbour ranks over NCCL after every sweep
 *                     (the reference's mpi_xplane_block exchange, :3328-3341), Eabs stays per rank.
 * imdcu_ttm3d_fill_ghost_layers: ttm_fill_ghost_layers() on l1 alone, as calc_ttm calls it ahead of
 *                     the sub-step loop (:67); uploaded ghost values are otherwise used as given.
 * imdcu_ttm3d_download: natoms / temp / U / xi of l1 (any may be NULL). */
int  imdcu_ttm3d_init(imdcu_ctx *ctx, int nx, int ny, int nz, const double h[3], const int pbc[3],
                      int fd_min_atoms);
int  imdcu_ttm3d_upload(imdcu_ctx *ctx, const int *natoms, const double *temp, const double *fd_k,
                        const double *fd_g, const double *Ce, const double *md_temp,
                        const double *source, const double *U, const double *xi, const double *dens);
int  imdcu_ttm3d_diff(imdcu_ctx *ctx, double tau, int diff_substeps, int laser_active, int nsweeps,
                      double *Eabs);
int  imdcu_ttm3d_fill_ghost_layers(imdcu_ctx *ctx);
int  imdcu_ttm3d_download(imdcu_ctx *ctx, int *natoms, double *temp, double *U, double *xi);

/* ------------------------------------------------------------------ inspection (parity) */
/* geometry chosen by imdcu_set_box: global_cell_dim and cellsz */
int  imdcu_get_cell_grid(const imdcu_ctx *ctx, int gdim[3], double *cellsz);
/* neighbour list as (id_i, id_j) pairs, every unordered pair once (i.e. the half list
 * of make_nblist up to order); call with pairs==NULL to get the count in *npairs. */
int  imdcu_get_nblist_pairs(imdcu_ctx *ctx, int64_t *npairs, int32_t *pairs);
int  imdcu_nbl_count(const imdcu_ctx *ctx);            /* globals.h:178 nbl_count */
/* > 0: since the last call a table was evaluated below its first argument (is_short, potaccess.h:331-333; the
 * reference warns "short distance" after calc_forces, imd_forces_nbl.c:673,992); < 0: error code */
int  imdcu_short_distance(imdcu_ctx *ctx);
int  imdcu_invalidate_nblist(imdcu_ctx *ctx);           /* have_valid_nbl = 0      */
/* kernel launches issued by this context so far (for bench.py's gpu_launches) */
int64_t imdcu_launch_count(const imdcu_ctx *ctx);
/* per-phase device time in ms accumulated with CUDA events when enabled:
 * 0 bin/sort, 1 nbl build, 2 ghost refresh, 3 pass1, 4 embed, 5 pass2, 6 integrate,
 * 7 ttm; launches[k] = kernels launched in that phase */
#define IMDCU_NPHASE 8
int  imdcu_set_profiling(imdcu_ctx *ctx, int enabled);
int  imdcu_get_phase_times(imdcu_ctx *ctx, double ms[IMDCU_NPHASE], int64_t launches[IMDCU_NPHASE]);
/* the CUDA stream all work of this context is enqueued on (cudaStream_t as void*) */
void *imdcu_stream(const imdcu_ctx *ctx);
int  imdcu_synchronize(imdcu_ctx *ctx);

/* FP64-FMA rate this GPU sustains, in GFLOP/s (a short kernel of independent DFMA chains on every SM): the
 * denominator of the FP64-pipe fraction reported beside the HBM fraction for the EAM passes */
int  imdcu_probe_fp64(imdcu_ctx *ctx, double *gflops);

/* single-function probe of the device table access (parity test of potaccess.h):
 * evaluates PAIR_INT2 on `which` for n arguments on the GPU */
int  imdcu_eval_table(imdcu_ctx *ctx, int which, int col, int64_t n, const double *arg,
                      double *val, double *grad);

/* imdcu_eval_table for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid
 * cudaMalloc.  Same arguments (host pointers), same kernel, same results as imdcu_eval_table, but the call
 * makes no cudaMalloc / cudaFree and calls no allocating library routine: the 3 n doubles of device staging
 * live in a file-scope device array of the library, 3 x IMDCU_EVAL_NOALLOC_MAX doubles = 1.5 MiB that every
 * GPU which loads libimdcu.so holds until the process exits.
 *   - n > IMDCU_EVAL_NOALLOC_MAX is rejected with IMDCU_EINVAL and a message naming the limit.
 *   - The first call on a device looks up the array's device address (once per device); make that call
 *     outside a capture.
 *   - All calls on one GPU share the array, which keeps the last call's contents; every element read is
 *     written earlier in the same call, so nothing needs clearing.  Calls that OVERLAP on the GPU (several
 *     contexts, host threads or streams on the same device) race on it: serialise them.  The call
 *     synchronises its stream before returning. */
#define IMDCU_EVAL_NOALLOC_MAX 65536
int  imdcu_eval_table_noalloc(imdcu_ctx *ctx, int which, int col, int64_t n, const double *arg,
                              double *val, double *grad);

/* single-function probe of the device log/exp/pow/div/sqrt used by the wide-range models
 * (imd_ttm.c:2191-2399 call libm): op 0 log(a), 1 exp(a), 2 pow(a,b), 3 a/b, 4 sqrt(a), 5 1/a */
int  imdcu_eval_ttm_math(imdcu_ctx *ctx, int op, int64_t n, const double *a, const double *b, double *out);

#ifdef __cplusplus
}
#endif
#endif /* IMDCU_H */
