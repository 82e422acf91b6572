/* integration/imd_cuda.c -- the IMD-side half of the drop-in: the host stays IMD's own C program.
 *
 * This file is compiled WITH THE REFERENCE'S HEADERS and linked with the reference's own host objects
 * (parameter parser, potential readers, generators, writers, timers, main loop: compiled in place from the
 * reference tree by integration/Makefile, nothing copied) and with libimdcu.so.  It takes the place of
 *
 *     imd_forces_nbl.c   calc_forces, check_nblist, make_nblist, estimate_nblist_size, deallocate_nblist
 *     imd_integrate.c    move_atoms_nve (:35-561), move_atoms_ttm (:581-811)   [the other ensembles stay]
 *     imd_ttm.c          init_ttm (:955), calc_ttm (:75), ttm_writeout (:1683), ttm_read (:1983)
 *     imd_tmm.c          tmm_init (:98)                       [do_tmm runs inside the device lattice kernel]
 *     imd_nrb.c          nrb_forces (:792-1129) and its exchanges    [init_nrb, nrb_writerestart stay; -DNRB builds]
 *
 * behind the reference's own names and call sites: calc_forces(steps) imd_main_3d.c:406, calc_ttm() :534,
 * move_atoms() :564 (function pointer set in imd_param.c:382,439), check_nblist() :812, init_ttm / tmm_init
 * imd.c:237,245, ttm_writeout imd_main_3d.c:735.  Make-target spelling: the reference selects sources by
 * substrings of the target name (Makefile:141-291,425-460); this is the substring "cuda"
 * (imd_eam_nve_nbl_cuda, imd_mpi_eam_ttm_tmm_nbl_cuda, imd_mpi_eam_ttm_tmm_nrb_nbl_cuda), see integration/Makefile and INTEGRATION.md.
 *
 * Data flow: atoms live on the device from the first calc_forces on.  cell_array (types.h:188-349) is
 * refilled by imdcu_sync_to_host() before every writer (write_config, write_config_select, write_distrib,
 * write_pictures are entered through the wrappers at the end of this file), so .chkpt, .force and picture
 * files are written by the reference's own writers in the reference's own formats; .eng needs only the
 * global sums, .ttm only the FD lattice, which is assembled on rank 0 from the device, and the distribution
 * files (.dens .Ekin .Epot .vxavg) only per-bin sums, which imdcu_distrib makes on the device.
 * There is no CPU fallback: every entry point either runs on the GPU or stops with imderror().
 */
#include "imd.h"
#include "imdcu.h"
#ifdef TTM
#include "nn_interpol/nn.h"
#include "nn_interpol/delaunay.h"
#endif

static imdcu_ctx *cu = NULL;
static int cu_resident = 0;          /* atoms have been handed to the device            */
static int cu_host_stale = 0;        /* the device is ahead of cell_array               */
static long cu_syncs = 0;

/* IMD_CUDA_TIMING=1: wall time spent inside the three hot entry points, printed by rank 0 when the program ends */
#include <sys/time.h>
static double cu_t_forces = 0, cu_t_ttm = 0, cu_t_move = 0, cu_t_first = 0;
static long   cu_n_forces = 0;
static int    cu_timing = -1;
static double cu_now(void) { struct timeval tv; gettimeofday(&tv, NULL); return tv.tv_sec + 1e-6 * tv.tv_usec; }
static void cu_report(void)
{
  if (cu_timing > 0 && myid == 0 && cu_n_forces > 1)
    fprintf(stderr, "IMD_CUDA_TIMING calls %ld: first calc_forces %.3f s; per step calc_forces %.3f ms, calc_ttm %.3f ms, "
            "move_atoms %.3f ms\n", cu_n_forces, cu_t_first, 1e3 * cu_t_forces / (cu_n_forces - 1),
            1e3 * cu_t_ttm / cu_n_forces, 1e3 * cu_t_move / cu_n_forces);
}

static void cu_fail(const char *where, int rc)
{
  char msg[1024];
  snprintf(msg, sizeof msg, "%s: %s (imdcu code %d)", where, imdcu_last_error(cu), rc);
  error(msg);                                        /* imderror, imd_misc.c:78-98 */
}
#define CK(call) do { int rc_ = (call); if (rc_ != IMDCU_OK) cu_fail(#call, rc_); } while (0)

#ifdef NRB
static void cu_nrb_restore(cell *p, int i);          /* the NRB members of an atom refilled into cell_array */
static void cu_nrb_resolve(void);
#endif

static void *cu_alloc(size_t bytes)
{
  void *p = malloc(bytes > 0 ? bytes : 1);
  if (p == NULL) error("imd_cuda: out of host memory");
  return p;
}

/* ------------------------------------------------------------------ TTM set-up state (host side) */
#ifdef TTM
static int     cu_ttm_analytic = 0;          /* analytic Beer-Lambert source instead of do_tmm              */
static double  cu_src_mu = 1.0 / 80.0, cu_src_abs = 0.1;
static double *cu_bb = NULL, *cu_k12 = NULL; /* rows of alu_eps_bb.dat / K12.dat (tmm_init, imd_tmm.c:117-118) */
static int     cu_bb_rows = 0, cu_k12_rows = 0;
static int     cu_ttm_P = 1;                 /* rank layout of the reference the FD lattice reproduces */
static double  cu_ttm_out[16];               /* calc_ttm's scalars, computed together with calc_forces          */
static int     cu_ttm_cached = 0;

/* nn_read_table leaves a scattered point set + the Delaunay triangulation of libnn.  Our tables are
 * rectilinear (is x js header): hand x[], y[], z[][] and, per grid rectangle, which diagonal the
 * triangulation chose (the reference's lpi_interpolate_point interpolates on those triangles). */
static void cu_eos_to_device(int which, nn_interp *nn, const char *fname)
{
  FILE *f;
  int is = 0, js = 0, i, j, t, k;
  double *x, *y, *z;
  unsigned char *diag;
  /* the grid shape is in the file header only (imd_interpol.c:1195-1196); every rank can read it */
  f = fopen(fname, "r");
  if (f == NULL || fscanf(f, "%d %d", &is, &js) != 2) error_str("imd_cuda: cannot read the header of %s", (char *) fname);
  fclose(f);
  if (is < 2 || js < 2 || is * js != nn->npoints) error_str("imd_cuda: %s is not an is x js table", (char *) fname);
  x = cu_alloc(is * sizeof *x); y = cu_alloc(js * sizeof *y); z = cu_alloc((size_t) is * js * sizeof *z);
  diag = calloc((size_t) (is - 1) * (js - 1), 1);
  for (i = 0; i < is; i++) x[i] = nn->points[i * js].x;
  for (j = 0; j < js; j++) y[j] = nn->points[j].y;
  for (i = 0; i < is * js; i++) z[i] = nn->points[i].z;
  for (t = 0; t < nn->d->ntriangles; t++) {
    int *v = nn->d->triangles[t].vids, imin = is, jmin = js, has00 = 0, has11 = 0;
    for (k = 0; k < 3; k++) { if (v[k] / js < imin) imin = v[k] / js; if (v[k] % js < jmin) jmin = v[k] % js; }
    for (k = 0; k < 3; k++) {
      int di = v[k] / js - imin, dj = v[k] % js - jmin;
      if (di > 1 || dj > 1) error_str("imd_cuda: the triangulation of %s is not a split of the grid rectangles", (char *) fname);
      if (di == 0 && dj == 0) has00 = 1;
      if (di == 1 && dj == 1) has11 = 1;
    }
    if (imin < is - 1 && jmin < js - 1) diag[imin * (js - 1) + jmin] = (has00 && has11) ? 0 : 1;
  }
  CK(imdcu_ttm_set_eos(cu, which, is, js, x, y, z, diag));
  free(x); free(y); free(z); free(diag);
}

/* a whitespace-separated table of numbers, `cols` per row (alu_eps_bb.dat: 3, K12.dat: 2) */
static double *cu_read_rows(const char *fname, int cols, int *rows)
{
  FILE *f = fopen(fname, "r");
  double v, *a = NULL;
  long n = 0, cap = 0;
  if (f == NULL) return NULL;
  while (fscanf(f, "%lf", &v) == 1) {
    if (n == cap) { cap = cap ? 2 * cap : 4096; a = realloc(a, cap * sizeof *a); if (a == NULL) error("imd_cuda: out of memory"); }
    a[n++] = v;
  }
  fclose(f);
  *rows = (int) (n / cols);
  return a;
}
#endif /* TTM */

/* ------------------------------------------------------------------ attach: everything the device needs */
static void cu_attach(void)
{
  double box[9];
  int pbc[3], cd[3] = {1, 1, 1}, ndev = 0, dev;
  const char *e;
  if (cu != NULL) return;
  CK(imdcu_device_count(&ndev));
  dev = myid % (ndev > 0 ? ndev : 1);                 /* one rank per GPU, ranks of a node in order */
  if ((e = getenv("IMDCU_DEVICE")) != NULL) dev = atoi(e);
  { int rc = imdcu_create(&cu, dev); if (rc != IMDCU_OK) cu_fail("imdcu_create", rc); }
#ifdef MPI
  if (cpu_dim.y != 1 || cpu_dim.z != 1)
    error("imd_cuda: the device path decomposes along x only: use cpu_dim N 1 1 (one rank per GPU)");
  {
    char id[128];
    memset(id, 0, sizeof id);
    if (num_cpus > 1) {
      if (myid == 0 && imdcu_comm_unique_id(id) != IMDCU_OK) error("imd_cuda: ncclGetUniqueId failed");
      MPI_Bcast(id, 128, MPI_CHAR, 0, cpugrid);
    }
    CK(imdcu_comm_init(cu, id, myid, num_cpus));
  }
#endif
  /* make-target substrings 4point / spline (Makefile:917-923) are a mode of the library */
#if defined(FOURPOINT)
  CK(imdcu_set_interp(cu, IMDCU_INTERP_4POINT));
#elif defined(SPLINE)
  CK(imdcu_set_interp(cu, IMDCU_INTERP_SPLINE));
#endif
  /* pot_table_t (types.h:537-549) as read_pot_table + init_threepoint left it (imd_potential.c:161-466,1260) */
  CK(imdcu_set_table(cu, IMDCU_TAB_PAIR, pair_pot.ncols, pair_pot.maxsteps, pair_pot.begin, pair_pot.end,
                     pair_pot.step, pair_pot.invstep, pair_pot.len, pair_pot.table));
  CK(imdcu_set_table(cu, IMDCU_TAB_RHO, rho_h_tab.ncols, rho_h_tab.maxsteps, rho_h_tab.begin, rho_h_tab.end,
                     rho_h_tab.step, rho_h_tab.invstep, rho_h_tab.len, rho_h_tab.table));
  CK(imdcu_set_table(cu, IMDCU_TAB_EMBED, embed_pot.ncols, embed_pot.maxsteps, embed_pot.begin, embed_pot.end,
                     embed_pot.step, embed_pot.invstep, embed_pot.len, embed_pot.table));
#ifdef SPLINE
  CK(imdcu_set_table2(cu, IMDCU_TAB_PAIR, pair_pot.table2));
  CK(imdcu_set_table2(cu, IMDCU_TAB_RHO, rho_h_tab.table2));
  CK(imdcu_set_table2(cu, IMDCU_TAB_EMBED, embed_pot.table2));
#endif
  box[0] = box_x.x; box[1] = box_x.y; box[2] = box_x.z;
  box[3] = box_y.x; box[4] = box_y.y; box[5] = box_y.z;
  box[6] = box_z.x; box[7] = box_z.y; box[8] = box_z.z;
  pbc[0] = pbc_dirs.x; pbc[1] = pbc_dirs.y; pbc[2] = pbc_dirs.z;
#ifdef TTM
  cd[0] = fd_ext.x; cd[1] = fd_ext.y; cd[2] = fd_ext.z;     /* init_cells: cd = cpu_dim * fd_ext (imd_geom_3d.c:168-176) */
#endif
  CK(imdcu_set_box(cu, box, pbc, nbl_margin, nbl_size, cd));
  {
    int g[3]; double cs;
    CK(imdcu_get_cell_grid(cu, g, &cs));
    if (g[0] != global_cell_dim.x || g[1] != global_cell_dim.y || g[2] != global_cell_dim.z) {
      char msg[256];
      snprintf(msg, sizeof msg, "imd_cuda: the device cell grid %d %d %d differs from init_cells' %d %d %d", g[0], g[1], g[2],
               global_cell_dim.x, global_cell_dim.y, global_cell_dim.z);
      error(msg);
    }
  }
  CK(imdcu_set_params(cu, ntypes, timestep, vtypes, (double *) restrictions));
#ifdef PDECAY
  CK(imdcu_set_pdecay(cu, 1, xipdecay, ramp_start, ramp_end));
#endif
#ifdef TTM
  {
    imdcu_ttm_params tp;
    memset(&tp, 0, sizeof tp);
    tp.ttmdimx = ttmdimx; tp.nranks_ref = cu_ttm_P;
    tp.fd_n_timesteps = fd_n_timesteps; tp.fd_min_atoms = fd_min_atoms; tp.vlatdim = vlatdim; tp.vlatbuffer = vlatbuffer;
    tp.vlatdens = vlatdens; tp.atomic_weight = atomic_weight; tp.atomic_charge = atomic_charge; tp.Temin = Temin;
    tp.I0 = I0; tp.lambda = lambda; tp.laser_t_0 = laser_t_0; tp.laser_sigma_t = laser_sigma_t;
    tp.tmm_laser_threshold = tmm_laser_threshold; tp.src_mu = cu_src_mu; tp.src_abs = cu_src_abs;
    cu_eos_to_device(IMDCU_EOS_CVE, &intp_cve_from_r_te, "../EOS_cve_from_r_te.txt");
    cu_eos_to_device(IMDCU_EOS_EE, &intp_ee_from_r_tesqrt, "../EOS_ee_from_r_tesqrt.txt");
    if (!cu_ttm_analytic) {
      cu_eos_to_device(IMDCU_EOS_PHASE, &intp_phase_from_r_ti, "../EOS_phase_from_r_ti.txt");
      CK(imdcu_ttm_set_tmm(cu, tmm_absorption_threshold, laser_t_1, laser_sigma_t1, cu_bb_rows, cu_bb, cu_k12_rows, cu_k12));
    }
    CK(imdcu_ttm_init(cu, &tp));
  }
#endif
  if (myid == 0) {
    printf("imd_cuda: hot path on CUDA device %d through libimdcu.so (ABI %d)%s\n", dev, imdcu_abi_version(),
           num_cpus > 1 ? ", x slabs over NCCL" : "");
    fflush(stdout);
  }
}

/* ------------------------------------------------------------------ cell_array -> device */
static void cu_upload(void)
{
  long n = 0, m = 0;
  int k, i;
  int32_t *id, *type, *vtype;
  double *mass, *pos, *mom;
  for (k = 0; k < NCELLS; k++) n += CELLPTR(k)->n;
  id = cu_alloc(n * sizeof *id); type = cu_alloc(n * sizeof *type); vtype = cu_alloc(n * sizeof *vtype);
  mass = cu_alloc(n * sizeof *mass); pos = cu_alloc(3 * n * sizeof *pos); mom = cu_alloc(3 * n * sizeof *mom);
  for (k = 0; k < NCELLS; k++) {
    cell *p = CELLPTR(k);
    for (i = 0; i < p->n; i++, m++) {
      id[m] = NUMMER(p, i); type[m] = SORTE(p, i); vtype[m] = VSORTE(p, i); mass[m] = MASSE(p, i);
      pos[3 * m] = ORT(p, i, X); pos[3 * m + 1] = ORT(p, i, Y); pos[3 * m + 2] = ORT(p, i, Z);
      mom[3 * m] = IMPULS(p, i, X); mom[3 * m + 1] = IMPULS(p, i, Y); mom[3 * m + 2] = IMPULS(p, i, Z);
    }
  }
  CK(imdcu_upload_atoms(cu, n, id, type, vtype, mass, pos, mom));
  free(id); free(type); free(vtype); free(mass); free(pos); free(mom);
  cu_resident = 1;
  cu_host_stale = 0;
}

/* ------------------------------------------------------------------ device -> cell_array
 * Called before every writer.  Every cell is emptied and refilled: an atom goes to the cell cell_coord
 * (imd_geom_3d.c:1077) gives for its current position, like fix_cells would put it (imd_fix_cells_3d.c:43).
 * An atom that has left this rank's slab since the device's last cell sort stays with this rank (clamped to
 * the nearest boundary cell) -- the writers only need every atom exactly once. */
void imdcu_sync_to_host(void)
{
  long n, m;
  int k;
  int32_t *id, *type, *vtype, *nn;
  double *mass, *pos, *mom, *frc, *epot, *rho, *dF;
  if (cu == NULL || !cu_resident || !cu_host_stale) return;
  n = (long) imdcu_num_atoms(cu);
  id = cu_alloc(n * sizeof *id); type = cu_alloc(n * sizeof *type); vtype = cu_alloc(n * sizeof *vtype);
  nn = cu_alloc(n * sizeof *nn); mass = cu_alloc(n * sizeof *mass);
  pos = cu_alloc(3 * n * sizeof *pos); mom = cu_alloc(3 * n * sizeof *mom); frc = cu_alloc(3 * n * sizeof *frc);
  epot = cu_alloc(n * sizeof *epot); rho = cu_alloc(n * sizeof *rho); dF = cu_alloc(n * sizeof *dF);
  CK(imdcu_download_atoms(cu, id, type, mass, pos, mom, frc, epot, rho, dF, nn, NULL));
  CK(imdcu_download_vtypes(cu, vtype));
  for (k = 0; k < nallcells; k++) cell_array[k].n = 0;
  for (m = 0; m < n; m++) {
    ivektor c = cell_coord(pos[3 * m], pos[3 * m + 1], pos[3 * m + 2]);
    cell *to;
    int i;
    c = local_cell_coord(c);
    if (c.x < 1) c.x = 1; else if (c.x > cell_dim.x - 2) c.x = cell_dim.x - 2;
    if (c.y < 1) c.y = 1; else if (c.y > cell_dim.y - 2) c.y = cell_dim.y - 2;
    if (c.z < 1) c.z = 1; else if (c.z > cell_dim.z - 2) c.z = cell_dim.z - 2;
    to = PTR_3D_VV(cell_array, c, cell_dim);
    if (to->n >= to->n_max) alloc_cell(to, to->n_max + incrsz);
    i = to->n++;
    NUMMER(to, i) = id[m]; SORTE(to, i) = type[m]; VSORTE(to, i) = vtype[m]; MASSE(to, i) = mass[m];
    ORT(to, i, X) = pos[3 * m]; ORT(to, i, Y) = pos[3 * m + 1]; ORT(to, i, Z) = pos[3 * m + 2];
    IMPULS(to, i, X) = mom[3 * m]; IMPULS(to, i, Y) = mom[3 * m + 1]; IMPULS(to, i, Z) = mom[3 * m + 2];
    KRAFT(to, i, X) = frc[3 * m]; KRAFT(to, i, Y) = frc[3 * m + 1]; KRAFT(to, i, Z) = frc[3 * m + 2];
    POTENG(to, i) = epot[m];
    EAM_RHO(to, i) = rho[m]; EAM_DF(to, i) = dF[m];
#ifdef TTM
    NUMNEIGHS(to, i) = nn[m];
#endif
#ifdef NRB
    cu_nrb_restore(to, i);
#endif
  }
#ifdef NRB
  cu_nrb_resolve();
#endif
  free(id); free(type); free(vtype); free(nn); free(mass); free(pos); free(mom); free(frc); free(epot); free(rho); free(dF);
  cu_host_stale = 0;
  cu_syncs++;
}

/* ------------------------------------------------------------------ the hot-path entry points */

#ifdef NRB
/* ------------------------------------------------------------------ non-reflecting boundary (make-target substring nrb)
 * The one-time setup stays the reference's own: init_nrb (imd_nrb.c:125-768, compiled in place) marks the boundary
 * atoms, finds their lattice-site neighbours by atom number and sets the reference positions -- the reference calls it
 * from make_nblist (imd_forces_nbl.c:281-282), i.e. in the first calc_forces with the buffer cells filled.  What it
 * leaves in cell_array goes to the device once (imdcu_set_nrb); nrb_forces (imd_nrb.c:792-1129) and the integrator
 * hooks (imd_integrate.c:233-259,712-751) then run there.  The table is kept on the host too, sorted by atom number,
 * so that imdcu_sync_to_host can give the atoms their NRB members back for nrb_writerestart (imd_main_3d.c:707-709). */
typedef struct { int id, bnd, neigh, nrbi[12]; double ref[3]; } cu_nrb_ent;
static cu_nrb_ent *cu_nrb_tab = NULL;
static long cu_nrb_n = 0;
static double cu_nrb_k = 0.0;
/* REF_POS of EVERY atom is part of the checkpoint (-DREFPOS, "refpos_x refpos_y refpos_z"): the atoms this rank owned
 * when init_nrb set them, sorted by number; an atom that has come from another slab since is asked for over MPI */
typedef struct { int id; double ref[3]; } cu_ref_ent;
static cu_ref_ent *cu_ref_tab = NULL;
static long cu_ref_n = 0;
typedef struct { cell *p; int i; } cu_ref_miss;
static cu_ref_miss *cu_miss = NULL;
static long cu_miss_n = 0, cu_miss_cap = 0;

static int cu_ref_cmp(const void *a, const void *b)
{
  const int x = ((const cu_ref_ent *) a)->id, y = ((const cu_ref_ent *) b)->id;
  return (x > y) - (x < y);
}

static int cu_nrb_cmp(const void *a, const void *b)
{
  const int x = ((const cu_nrb_ent *) a)->id, y = ((const cu_nrb_ent *) b)->id;
  return (x > y) - (x < y);
}

static void cu_nrb_setup(void)
{
  cu_nrb_ent *loc;
  long nloc = 0, cap = 1024, m;
  int k, i, j;
  cu_nrb_k = nrbk;                                   /* the parameter nrb_k; init_nrb replaces it by sqrt(nrb_k / m), :158 */
#ifdef MPI
  setup_buffers();                                   /* message buffers first (imd_forces_nbl.c:304) */
#endif
  send_cells(copy_cell, pack_cell, unpack_cell);     /* buffer cells, as calc_forces fills them (imd_forces_nbl.c:311) */
  init_nrb();
  loc = cu_alloc(cap * sizeof *loc);
  for (k = 0, cu_ref_n = 0; k < ncells; k++) cu_ref_n += CELLPTR(k)->n;
  cu_ref_tab = cu_alloc((size_t) (cu_ref_n > 0 ? cu_ref_n : 1) * sizeof *cu_ref_tab);
  for (k = 0, m = 0; k < ncells; k++) {
    cell *p = CELLPTR(k);
    for (i = 0; i < p->n; i++, m++) {
      cu_ref_tab[m].id = NUMMER(p, i);
      cu_ref_tab[m].ref[0] = REF_POS(p, i, X); cu_ref_tab[m].ref[1] = REF_POS(p, i, Y); cu_ref_tab[m].ref[2] = REF_POS(p, i, Z);
    }
  }
  qsort(cu_ref_tab, cu_ref_n, sizeof *cu_ref_tab, cu_ref_cmp);
  for (k = 0; k < ncells; k++) {
    cell *p = CELLPTR(k);
    for (i = 0; i < p->n; i++) {
      if (NRBBND(p, i) == 0 && NRBNEIGH(p, i) == 0) continue;
      if (nloc == cap) { cap *= 2; loc = realloc(loc, cap * sizeof *loc); if (loc == NULL) error("imd_cuda: out of host memory"); }
      memset(&loc[nloc], 0, sizeof *loc);
      loc[nloc].id = NUMMER(p, i); loc[nloc].bnd = NRBBND(p, i); loc[nloc].neigh = NRBNEIGH(p, i);
      for (j = 0; j < 12; j++) loc[nloc].nrbi[j] = NRBI(p, i, j);
      loc[nloc].ref[0] = REF_POS(p, i, X); loc[nloc].ref[1] = REF_POS(p, i, Y); loc[nloc].ref[2] = REF_POS(p, i, Z);
      nloc++;
    }
  }
#ifdef MPI
  {                                                  /* every rank hands the full table to the library */
    int mine = (int) nloc, most = 0, r, *counts = cu_alloc(num_cpus * sizeof *counts);
    cu_nrb_ent *pad, *all;
    MPI_Allgather(&mine, 1, MPI_INT, counts, 1, MPI_INT, cpugrid);
    for (r = 0; r < num_cpus; r++) if (counts[r] > most) most = counts[r];
    pad = calloc((size_t) (most > 0 ? most : 1), sizeof *pad);
    all = cu_alloc((size_t) (most > 0 ? most : 1) * num_cpus * sizeof *all);
    memcpy(pad, loc, nloc * sizeof *loc);
    MPI_Allgather(pad, most * (int) sizeof *pad, MPI_CHAR, all, most * (int) sizeof *pad, MPI_CHAR, cpugrid);
    cu_nrb_n = 0;
    for (r = 0; r < num_cpus; r++) cu_nrb_n += counts[r];
    cu_nrb_tab = cu_alloc((size_t) (cu_nrb_n > 0 ? cu_nrb_n : 1) * sizeof *cu_nrb_tab);
    for (r = 0, m = 0; r < num_cpus; r++) { memcpy(cu_nrb_tab + m, all + (size_t) r * most, counts[r] * sizeof *all); m += counts[r]; }
    free(pad); free(all); free(counts); free(loc);
  }
#else
  cu_nrb_tab = loc; cu_nrb_n = nloc;
#endif
  qsort(cu_nrb_tab, cu_nrb_n, sizeof *cu_nrb_tab, cu_nrb_cmp);
}

static void cu_nrb_to_device(void)
{
  int32_t *id = cu_alloc(cu_nrb_n * sizeof *id), *bnd = cu_alloc(cu_nrb_n * sizeof *bnd), *nrbi = cu_alloc(12 * cu_nrb_n * sizeof *nrbi);
  double *ref = cu_alloc(3 * cu_nrb_n * sizeof *ref);
  long m;
  int j;
  for (m = 0; m < cu_nrb_n; m++) {
    id[m] = cu_nrb_tab[m].id; bnd[m] = cu_nrb_tab[m].bnd;
    for (j = 0; j < 12; j++) nrbi[12 * m + j] = cu_nrb_tab[m].nrbi[j];
    for (j = 0; j < 3; j++) ref[3 * m + j] = cu_nrb_tab[m].ref[j];
  }
  /* startup_images = 1: what the reference's send_forces does to the boundary momenta in steps 0 and 1 (imdcu.h) */
  CK(imdcu_set_nrb(cu, cu_nrb_k, cu_nrb_n, id, bnd, nrbi, ref, 1));
  if (myid == 0) printf("imd_cuda: non-reflecting boundary on the device, %ld boundary and neighbour atoms, nrb_k %g\n", cu_nrb_n, cu_nrb_k);
  free(id); free(bnd); free(nrbi); free(ref);
}

/* the NRB members of one atom of cell_array (imdcu_sync_to_host) */
static void cu_nrb_restore(cell *p, int i)
{
  cu_nrb_ent key, *e;
  int j;
  key.id = NUMMER(p, i);
  e = cu_nrb_n > 0 ? bsearch(&key, cu_nrb_tab, cu_nrb_n, sizeof *cu_nrb_tab, cu_nrb_cmp) : NULL;
  NRBBND(p, i) = e ? e->bnd : 0;
  NRBNEIGH(p, i) = e ? e->neigh : 0;
  for (j = 0; j < 12; j++) NRBI(p, i, j) = e ? e->nrbi[j] : 0;
  if (e) { REF_POS(p, i, X) = e->ref[0]; REF_POS(p, i, Y) = e->ref[1]; REF_POS(p, i, Z) = e->ref[2]; }
  else {
    cu_ref_ent rk, *r;
    rk.id = key.id;
    r = cu_ref_n > 0 ? bsearch(&rk, cu_ref_tab, cu_ref_n, sizeof *cu_ref_tab, cu_ref_cmp) : NULL;
    if (r) { REF_POS(p, i, X) = r->ref[0]; REF_POS(p, i, Y) = r->ref[1]; REF_POS(p, i, Z) = r->ref[2]; }
    else {                                           /* arrived from another rank: cu_nrb_resolve asks its first owner */
      if (cu_miss_n == cu_miss_cap) {
        cu_miss_cap = cu_miss_cap ? 2 * cu_miss_cap : 256;
        cu_miss = realloc(cu_miss, cu_miss_cap * sizeof *cu_miss);
        if (cu_miss == NULL) error("imd_cuda: out of host memory");
      }
      cu_miss[cu_miss_n].p = p; cu_miss[cu_miss_n].i = i; cu_miss_n++;
    }
  }
}

/* collective, at the end of imdcu_sync_to_host: reference positions of the atoms that changed their rank */
static void cu_nrb_resolve(void)
{
#ifdef MPI
  int mine = (int) cu_miss_n, most = 0, r, q, *counts, *ask, *all;
  double *ans, *sum;
  if (num_cpus == 1) { if (cu_miss_n > 0) error("imd_cuda: an atom without reference position"); return; }
  counts = cu_alloc(num_cpus * sizeof *counts);
  MPI_Allgather(&mine, 1, MPI_INT, counts, 1, MPI_INT, cpugrid);
  for (r = 0; r < num_cpus; r++) if (counts[r] > most) most = counts[r];
  if (most > 0) {
    static int told = 0;
    if (!told && myid == 0) { printf("imd_cuda: reference positions of atoms that changed their rank are fetched from their first owner\n"); told = 1; }
    ask = calloc(most, sizeof *ask); all = cu_alloc((size_t) most * num_cpus * sizeof *all);
    for (q = 0; q < mine; q++) ask[q] = NUMMER(cu_miss[q].p, cu_miss[q].i);
    MPI_Allgather(ask, most, MPI_INT, all, most, MPI_INT, cpugrid);
    ans = calloc((size_t) 4 * most * num_cpus, sizeof *ans); sum = cu_alloc((size_t) 4 * most * num_cpus * sizeof *sum);
    for (r = 0; r < num_cpus; r++)
      for (q = 0; q < counts[r]; q++) {
        cu_ref_ent rk, *e;
        rk.id = all[r * most + q];
        e = cu_ref_n > 0 ? bsearch(&rk, cu_ref_tab, cu_ref_n, sizeof *cu_ref_tab, cu_ref_cmp) : NULL;
        if (e) { double *a = ans + 4 * ((size_t) r * most + q); a[0] = e->ref[0]; a[1] = e->ref[1]; a[2] = e->ref[2]; a[3] = 1.0; }
      }
    MPI_Allreduce(ans, sum, 4 * most * num_cpus, MPI_DOUBLE, MPI_SUM, cpugrid);
    for (q = 0; q < mine; q++) {
      const double *a = sum + 4 * ((size_t) myid * most + q);
      if (a[3] != 1.0) error("imd_cuda: an atom without reference position");
      REF_POS(cu_miss[q].p, cu_miss[q].i, X) = a[0]; REF_POS(cu_miss[q].p, cu_miss[q].i, Y) = a[1]; REF_POS(cu_miss[q].p, cu_miss[q].i, Z) = a[2];
    }
    free(ask); free(all); free(ans); free(sum);
  }
  free(counts);
#else
  if (cu_miss_n > 0) error("imd_cuda: an atom without reference position");
#endif
  cu_miss_n = 0;
}
#endif /* NRB */

void calc_forces(int steps)                          /* replaces imd_forces_nbl.c:295-1054 */
{
  double out8[8];
  double t0;
  if (cu_timing < 0) { cu_timing = getenv("IMD_CUDA_TIMING") != NULL; if (cu_timing) atexit(cu_report); }
  t0 = cu_timing ? cu_now() : 0.0;
#ifdef NRB
  if (!cu_resident && !have_valid_nrb) cu_nrb_setup();             /* host only */
#endif
  cu_attach();
#ifdef NRB
  if (!cu_resident) { cu_upload(); cu_nrb_to_device(); }
#else
  if (!cu_resident) cu_upload();
#endif
  if (!have_valid_nbl) CK(imdcu_invalidate_nblist(cu));   /* the host asked for a new list (imd_main_3d.c:273) */
#ifdef TTM
  /* calc_ttm() follows at imd_main_3d.c:534 with nothing in between that touches the atoms: both in one call, so that
   * the lattice kernel runs beside the second EAM pass; calc_ttm() below hands out the cached scalars */
  CK(imdcu_calc_forces_ttm(cu, steps, out8, cu_ttm_out));
  cu_ttm_cached = 1;
#else
  CK(imdcu_calc_forces(cu, steps, out8));
#endif
  tot_pot_energy = out8[0]; virial = out8[1];
#ifdef P_AXIAL
  vir_xx = out8[2]; vir_yy = out8[3]; vir_zz = out8[4];
#endif
  if (imdcu_short_distance(cu) > 0 && myid == 0) fprintf(stderr, "Short distance, pair, step %d!\n", steps);
  have_valid_nbl = 1;
  nbl_count = imdcu_nbl_count(cu);
  nfc++;
  cu_host_stale = 1;
  if (cu_timing) { if (cu_n_forces == 0) cu_t_first = cu_now() - t0; else cu_t_forces += cu_now() - t0; cu_n_forces++; }
#ifdef FORCE
  /* write_config_select(.., "force", ..) runs right after calc_forces (imd_main_3d.c:430): the wrapper syncs */
#endif
}

void move_atoms_nve(void)                            /* replaces imd_integrate.c:35-561 */
{
  double out8[8];
  int valid = 0;
  if (cu == NULL || !cu_resident) error("imd_cuda: move_atoms_nve before the first calc_forces");
  CK(imdcu_move_atoms_nve(cu, out8, &valid));
  tot_kin_energy = out8[0];
  have_valid_nbl = valid;                            /* check_nblist is fused into the integrator kernel */
  cu_host_stale = 1;
}

void check_nblist(void)                              /* imd_forces_nbl.c:1062-1092: have_valid_nbl is already set */
{
}

/* the list lives on the device; these reference entry points have nothing left to do on the host */
void make_nblist(void) { have_valid_nbl = 0; }
int  estimate_nblist_size(void) { return 0; }
void deallocate_nblist(void) { have_valid_nbl = 0; }

#ifndef TTM
void move_atoms_ttm(void)                            /* as the reference's own stub for a non-TTM binary (imd_integrate.c:814) */
{
  if (myid == 0) error("the chosen ensemble TTM is not supported by this binary");
}
#endif

#ifdef TTM
/* init_ttm (imd_ttm.c:955-1201): derived globals other host code reads, the arrays ttm_writeout prints, the EOS
 * tables through the reference's own reader.  The lattice itself is set up on the device at attach time. */
void init_ttm(void)
{
  const char *e;
  int i;
  global_fd_dim.x = ttmdimx;
  cu_ttm_P = num_cpus;
  if ((e = getenv("IMDCU_TTM_REF_RANKS")) != NULL) cu_ttm_P = atoi(e);   /* reproduce a reference run on P ranks */
  if (cu_ttm_P < 1 || global_fd_dim.x % cu_ttm_P != 0) error("global_fd_dim.x mod num_cpus != 0.");
  local_fd_dim.x = global_fd_dim.x / cu_ttm_P + 2;
  local_fd_dim.y = 3; local_fd_dim.z = 3;
  fd_h.x = box_x.x / global_fd_dim.x; fd_h.y = box_y.y; fd_h.z = box_z.z;
  fd_vol = fd_h.x * fd_h.y * fd_h.z;
  max_dt_ttm = timestep / ((double) fd_n_timesteps);
  neighvol = pow(sqrt(pair_pot.end[0]), 3.0) * 4.0 / 3.0 * M_PI;
  xiarr_global = cu_alloc(global_fd_dim.x * sizeof(double));
  vcomxglobal = cu_alloc(global_fd_dim.x * sizeof(double));
  vcomyglobal = cu_alloc(global_fd_dim.x * sizeof(double));
  vcomzglobal = cu_alloc(global_fd_dim.x * sizeof(double));
  for (i = 0; i < global_fd_dim.x; i++) xiarr_global[i] = vcomxglobal[i] = vcomyglobal[i] = vcomzglobal[i] = 0.0;
  if (myid == 0) {
    printf("***************************************************\n");
    printf("*     TWO-TEMPERATURE MODEL (device resident)     *\n");
    printf("***************************************************\n");
    printf("Global FD cell array dimension: %d \n", global_fd_dim.x);
    printf("fd_h.x:%f A, fd_h.y:%f A,fd_h.z:%f A\n", fd_h.x, fd_h.y, fd_h.z);
    printf("FD lattice laid out like a reference run on %d rank(s)\n", cu_ttm_P);
  }
  nn_read_table(&intp_cve_from_r_te, "../EOS_cve_from_r_te.txt");
  nn_read_table(&intp_ee_from_r_tesqrt, "../EOS_ee_from_r_tesqrt.txt");
  if (imdrestart > 0)
    error("imd_cuda: restart of the FD lattice from a .ttm file (ttm_read, imd_ttm.c:1983) is not provided: the reference's own TTM restart resets the electronic state (DESIGN.md 9.4)");
}

/* tmm_init (imd_tmm.c:98-190): the two data files of do_tmm.  Without them (they are missing blobs upstream)
 * IMD_SRC_MU / IMD_SRC_ABS select the analytic Beer-Lambert source the parity builds of the reference use. */
int tmm_init(void)
{
  const char *e;
  laser_sigma_t_squared = laser_sigma_t * laser_sigma_t;
  laser_sigma_t1_squared = laser_sigma_t1 * laser_sigma_t1;
  omega_laser = 2.0 * M_PI * c0 / lambda;
  laser_spot_area = (fd_h.y * 1e-10 * fd_h.z * 1e-10);
  laser_active = false;
  if ((e = getenv("IMD_SRC_MU")) != NULL) { cu_ttm_analytic = 1; cu_src_mu = atof(e); }
  if ((e = getenv("IMD_SRC_ABS")) != NULL) { cu_ttm_analytic = 1; cu_src_abs = atof(e); }
  if (!cu_ttm_analytic) {
    cu_bb = cu_read_rows("../alu_eps_bb.dat", 3, &cu_bb_rows);
    cu_k12 = cu_read_rows("../K12.dat", 2, &cu_k12_rows);
    if (cu_bb == NULL || cu_k12 == NULL)
      error("imd_cuda: ../alu_eps_bb.dat / ../K12.dat not found (set IMD_SRC_MU and IMD_SRC_ABS for the analytic source)");
    nn_read_table(&intp_phase_from_r_ti, "../EOS_phase_from_r_ti.txt");
  }
  if (myid == 0)
    printf("imd_cuda: laser source = %s, I0:%.4e W/m^2, t0:%.4e, sigma_t:%.4e\n",
           cu_ttm_analytic ? "analytic Beer-Lambert" : "transfer-matrix method (do_tmm on the device)", I0, laser_t_0, laser_sigma_t);
  return 0;
}

void calc_ttm(void)                                  /* replaces imd_ttm.c:75-156 */
{
  double *o = cu_ttm_out;
  double t0 = cu_timing > 0 ? cu_now() : 0.0;
  if (cu == NULL || !cu_resident) error("imd_cuda: calc_ttm before the first calc_forces");
  if (!cu_ttm_cached) CK(imdcu_calc_ttm(cu, steps, o));     /* a call without calc_forces in front: the separate entry */
  cu_ttm_cached = 0;
  diff_substeps = (int) o[0]; max_dt_ttm = o[1];
  last_active_cell_global = last_active_cell_local = (int) o[2];
  cur_vlattice_proc = (int) o[3];
  Eabs_global = o[4]; Eabs_local = myid == 0 ? o[4] : 0.0;
  /* write_eng_file sums tot_elec_energy_local over the ranks (imd_io.c:2345); the device lattice is replicated */
  tot_elec_energy_local = myid == 0 ? o[5] : 0.0;
  laser_active = (o[9] != 0.0);
  tmm_time = o[10];
  if (cu_timing > 0) cu_t_ttm += cu_now() - t0;
}

void move_atoms_ttm(void)                            /* replaces imd_integrate.c:581-811 */
{
  double out8[8];
  int valid = 0;
  double t0 = cu_timing > 0 ? cu_now() : 0.0;
  if (cu == NULL || !cu_resident) error("imd_cuda: move_atoms_ttm before the first calc_forces");
  CK(imdcu_move_atoms_ttm(cu, out8, &valid));
  tot_kin_energy = out8[0];
  have_valid_nbl = valid;
  cu_host_stale = 1;
  if (cu_timing > 0) cu_t_move += cu_now() - t0;
}

/* ttm_writeout (imd_ttm.c:1683-1969): <outfiles>.<number>.ttm, one line per FD cell, then the virtual lattice
 * (natoms = -2).  The device lattice is complete on every rank, so rank 0 writes it without a gather. */
void ttm_writeout(int number)
{
  const int G = global_fd_dim.x, per = G / cu_ttm_P;
  double *cells, *xi, *vcom, *vlat;
  char fname[512];
  FILE *out;
  int i;
  if (cu == NULL || !cu_resident) return;
  cells = cu_alloc((size_t) G * 13 * sizeof *cells); xi = cu_alloc(G * sizeof *xi);
  vcom = cu_alloc((size_t) 3 * G * sizeof *vcom); vlat = cu_alloc((size_t) (vlatdim > 0 ? vlatdim : 1) * 6 * sizeof *vlat);
  CK(imdcu_download_ttm(cu, cells, xi, vcom, vlat));
  for (i = 0; i < G; i++) {
    xiarr_global[i] = xi[i]; vcomxglobal[i] = vcom[i]; vcomyglobal[i] = vcom[G + i]; vcomzglobal[i] = vcom[2 * G + i];
  }
  if (myid == 0) {
    snprintf(fname, sizeof fname, "%s.%d.ttm", outfilename, number);
    out = fopen(fname, "w");
    if (out == NULL) error("Cannot open ttm file for writing.\n");
    fprintf(out, "#x y z natoms temp md_temp U xi source dens vx vy vz fd_k fd_g Z proc Ce\n");
    for (i = 0; i < G; i++) {
      const double *c = cells + 13 * (size_t) i;     /* natoms natoms_old temp md_temp xi dens fd_k fd_g Z ne Ce U source */
      fprintf(out, "%d %d %d %d %e %e %e %e %e %e %e %e %e %e %e %e %d %f\n", i, 0, 0, (int) c[0], c[2], c[3], c[11], c[4],
              c[12], c[5], vcom[i], vcom[G + i], vcom[2 * G + i], c[6], c[7], c[8], i / per, c[10]);
    }
    for (i = 0; i < vlatdim; i++) {
      const double *v = vlat + 6 * (size_t) i;       /* temp md_temp fd_k fd_g Ce Z */
      fprintf(out, "%d %d %d %d %e %e %e %e %e %e %e %e %e %e %e %e %d %f\n", i + last_active_cell_global + 1, 0, 0, -2,
              v[0], v[1], 0.0, 0.0, 0.0, vlatdens, 0.0, 0.0, 0.0, v[2], v[3], v[5], cur_vlattice_proc, v[4]);
    }
    fclose(out);
  }
  free(cells); free(xi); free(vcom); free(vlat);
}

void ttm_read(int number)
{
  error("imd_cuda: ttm_read (restart of the FD lattice) is not provided: the reference's own TTM restart resets the electronic state (DESIGN.md 9.4)");
}
#endif /* TTM */

/* ------------------------------------------------------------------ writers: cell_array must be current.
 * Linked with -Wl,--wrap=write_config,--wrap=write_config_select,--wrap=write_distrib,--wrap=write_pictures;
 * a maintainer editing the reference would put the one call in front of these calls in imd_main_3d.c:702-735
 * under #ifdef IMDCUDA instead. */
void __real_write_config(int fzhlr, int steps);
void __wrap_write_config(int fzhlr, int steps) { imdcu_sync_to_host(); __real_write_config(fzhlr, steps); }
void __real_write_config_select(int fzhlr, char *suffix, void (*write_atoms_fun)(FILE *), void (*write_header_fun)(FILE *));
void __wrap_write_config_select(int fzhlr, char *suffix, void (*write_atoms_fun)(FILE *), void (*write_header_fun)(FILE *))
{
  imdcu_sync_to_host();
  __real_write_config_select(fzhlr, suffix, write_atoms_fun, write_header_fun);
}
/* ------------------------------------------------------------------ write_distrib (imd_distrib.c:33-218)
 * dist_dens / dist_Ekin / dist_Epot / dist_vxavg need per-bin sums only: imdcu_distrib makes them on the device, and
 * dist_size floats per quantity come back instead of every atom.  The files stay the reference's:
 * headers by its write_distrib_header (:805), the density file by its write_distrib_density (:434, which reads
 * num_1 / dist_size and fills dat_1, the globals of imd_distrib.c), the averaged quantities normalised and laid out
 * as make_write_distrib_select does (:721-800).  Sums over the ranks are the reference's MPI_Reduce of float arrays.
 * Any other dist_*_flag (press*, shock*, mdtemp) selects per-atom data this path does not carry: cell_array is
 * refilled and the reference's own write_distrib runs.  IMD_CUDA_HOST_DISTRIB=1 forces that path for every flag. */
void __real_write_distrib(int steps);
extern float *dat_1, *num_1;
extern int dist_size;

static int cu_dist_on_device(void)
{
  const char *e = getenv("IMD_CUDA_HOST_DISTRIB");
  if (e != NULL && atoi(e) > 0) return 0;
  if (cu == NULL || !cu_resident) return 0;
  if (dist_mdtemp_flag) return 0;
#ifdef STRESS_TENS
  if (dist_press_flag || dist_presstens_flag) return 0;
#endif
#ifdef SHOCK
  if (dist_Ekin_long_flag || dist_Ekin_trans_flag || dist_Ekin_comp_flag || dist_shock_shear_flag ||
      dist_shear_aniso_flag || dist_pressoff_flag) return 0;
#endif
#ifdef DISLOC
  if (Epot_diff == 1) return 0;
#endif
  return 1;
}

/* one averaged quantity: sum over ranks, divide by the count where there are atoms, write file and minmax line */
static void cu_write_dist_select(float *sum, float *red, int mode, int fzhlr, char *suffix)
{
  FILE *out;
  char fname[255];
  float lo = 1e+10, hi = -1e+10;
  int i, r = 0, s = 0, t = 0;
#ifdef MPI
  MPI_Reduce(sum, red, dist_size, MPI_FLOAT, MPI_SUM, 0, cpugrid);
#else
  red = sum;
#endif
  if (myid != 0) return;
  for (i = 0; i < dist_size; i++)
    if (num_1[i] > 0.0) {
      red[i] /= num_1[i];
      lo = MIN(lo, red[i]);
      hi = MAX(hi, red[i]);
    }
  sprintf(fname, "%s.%u.%s", outfilename, fzhlr, suffix);
  out = fopen(fname, "w");
  if (NULL == out) error("Cannot open distribution file.");
  write_distrib_header(out, mode, 1, suffix);
  if (mode == DIST_FORMAT_BINARY) {
    if (fwrite(red, sizeof(float), dist_size, out) != (size_t) dist_size) warning("distribution write incomplete!");
  }
  else if (mode == DIST_FORMAT_ASCII || mode == DIST_FORMAT_ASCII_COORD) {
    for (i = 0; i < dist_size; i++) {
      if (mode == DIST_FORMAT_ASCII_COORD) {
        fprintf(out, "%d %d %d", r, s, t++);
        if (t == dist_dim.z) { t = 0; s++; }
        if (s == dist_dim.y) { s = 0; r++; }
      }
      fprintf(out, " %e\n", red[i]);
    }
  }
  else error("unknown distribution output format");
  fclose(out);
  sprintf(fname, "%s.minmax.%s", outfilename, suffix);
  out = fopen(fname, "a");
  if (NULL == out) error("Cannot open minmax file.");
  fprintf(out, "%d  %e %e\n", fzhlr, lo, hi);
  fclose(out);
}

static void cu_write_distrib(int steps)
{
  static int told = 0;
  int dim[3], fzhlr;
  double ll[3], ur[3];
  float *loc, *red;
  is_big_endian = endian();
  if (0.0 == dist_ur.x) {                            /* backup if dist_ur is not set, :41-47 */
    dist_ur.x = box_x.x; dist_ur.y = box_y.y; dist_ur.z = box_z.z;
  }
  dim[0] = dist_dim.x; dim[1] = dist_dim.y; dim[2] = dist_dim.z;
  ll[0] = dist_ll.x; ll[1] = dist_ll.y; ll[2] = dist_ll.z;
  ur[0] = dist_ur.x; ur[1] = dist_ur.y; ur[2] = dist_ur.z;
  dist_size = dist_dim.x * dist_dim.y * dist_dim.z;
  fzhlr = steps / dist_int;
  loc = cu_alloc((size_t) 4 * dist_size * sizeof *loc);        /* this rank: count, Ekin, Epot, v_x sums */
  red = cu_alloc((size_t) 2 * dist_size * sizeof *red);        /* rank 0: summed count | one summed quantity */
  CK(imdcu_distrib(cu, dim, ll, ur, loc, loc + dist_size, loc + 2 * dist_size, loc + 3 * dist_size));
#ifdef MPI
  MPI_Reduce(loc, red, dist_size, MPI_FLOAT, MPI_SUM, 0, cpugrid);     /* make_distrib_density, :423 */
  num_1 = red;
#else
  num_1 = loc;
#endif
  dat_1 = red + dist_size;
  if (!told && myid == 0) {
    printf("imd_cuda: distributions (%d bins) are summed on the device\n", dist_size);
    told = 1;
  }
  if (dist_Ekin_flag) cu_write_dist_select(loc + dist_size, dat_1, dist_Ekin_flag, fzhlr, "Ekin");
  if (dist_Epot_flag) cu_write_dist_select(loc + 2 * dist_size, dat_1, dist_Epot_flag, fzhlr, "Epot");
#if defined(SHOCK) || defined(TTM)
  if (dist_vxavg_flag) cu_write_dist_select(loc + 3 * dist_size, dat_1, dist_vxavg_flag, fzhlr, "vxavg");
#endif
  if (myid == 0 && dist_dens_flag) write_distrib_density(dist_dens_flag, fzhlr);
  num_1 = dat_1 = NULL;
  free(loc); free(red);
}

void __wrap_write_distrib(int steps)
{
  if (cu_dist_on_device()) { cu_write_distrib(steps); return; }
  imdcu_sync_to_host();
  __real_write_distrib(steps);
}
void __real_write_pictures(int steps);
void __wrap_write_pictures(int steps) { imdcu_sync_to_host(); __real_write_pictures(steps); }
