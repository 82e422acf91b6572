/* oracle/ref_hooks_ttm.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Link-time hooks for the UNMODIFIED reference's TTM build (-DMPI -DTTM -DTTM1D -DTMM over
 * shmpi/).  Linked with  -Wl,--wrap=tmm_init,--wrap=do_tmm,--wrap=calc_ttm .
 *
 *  __wrap_tmm_init / __wrap_do_tmm : the reference's Helmholtz/transfer-matrix laser source
 *      (imd_tmm.c:98-190, 193-445) needs the data files alu_eps_bb.dat and K12.dat, which
 *      are not in the tree (.MISSING_LARGE_BLOBS).  They are replaced by the analytic
 *      Beer-Lambert source documented in DESIGN.md ("TTM source term"):
 *          I_t  = I0 * exp(-(tmm_time - laser_t_0)^2 / laser_sigma_t^2)           (imd_tmm.c:197)
 *          S_i  = I_t * A * mu * exp(-mu * depth_i) * 1e10 * 6.3538562638e-26     (imd_tmm.c:427-428)
 *      for active cells, depth_i = (i_global - i_first_active) * fd_h.x, mu in 1/Angstrom
 *      (env IMD_SRC_MU, default 1/80), A = absorbed fraction (env IMD_SRC_ABS, default 0.1).
 *      The laser_active switch keeps the reference's field threshold (imd_tmm.c:200-208).
 *      Unlike the reference there is no "steps < 2" dead time.
 *  __wrap_calc_ttm : calls the reference's calc_ttm (imd_ttm.c:75) and then dumps the FD
 *      lattice and the arrays the integrator consumes, for the steps in $IMD_DUMP_STEPS.
 *
 * No reference arithmetic is changed; reference state is read through its own headers.
 */
#include "imd.h"
#include <stdlib.h>
#include <string.h>

#if defined(TTM) && defined(TTM1D)

#define RHOMIN_H 2

static double src_mu = 1.0 / 80.0, src_abs = 0.1;

int __wrap_tmm_init(void)
{
  const char *e;
  laser_sigma_t_squared = laser_sigma_t * laser_sigma_t;
  omega_laser = 2.0 * M_PI * c0 / lambda;
  laser_spot_area = (fd_h.y * 1e-10 * fd_h.z * 1e-10);
  if ((e = getenv("IMD_SRC_MU")) != NULL) src_mu = atof(e);
  if ((e = getenv("IMD_SRC_ABS")) != NULL) src_abs = atof(e);
  laser_active = false;
  if (myid == 0)
    printf("ref_hooks_ttm: analytic Beer-Lambert source, mu=%.6e 1/A, A=%.4f, I0=%.4e, t0=%.4e, sigma_t=%.4e\n",
           src_mu, src_abs, I0, laser_t_0, laser_sigma_t);
  return 0;
}

int __wrap_do_tmm(real dt)
{
  int i, first_local = 1 << 30, first;
  double Imp = 3.769911184307751e+02;
  I_t = I0 * exp(-pow(tmm_time - laser_t_0, 2) / laser_sigma_t_squared);
  if (sqrt(2.0 * I_t * Imp) < tmm_laser_threshold * sqrt(2 * I0 * Imp)) {
    laser_active = false;
    for (i = 1; i < local_fd_dim.x - 1; i++) l1[i].source = l2[i].source = 0.0;
    return 0;
  }
  laser_active = true;
  for (i = 1; i < local_fd_dim.x - 1; i++) {
    int ig = (i - 1) + myid * (local_fd_dim.x - 2);
    if (l1[i].natoms >= fd_min_atoms && l1[i].dens >= RHOMIN_H) { if (ig < first_local) first_local = ig; }
  }
  MPI_Allreduce(&first_local, &first, 1, MPI_INT, MPI_MIN, cpugrid);
  for (i = 1; i < local_fd_dim.x - 1; i++) {
    int ig = (i - 1) + myid * (local_fd_dim.x - 2);
    double s = 0.0;
    if (l1[i].natoms >= fd_min_atoms && l1[i].dens >= RHOMIN_H) {
      double depth = (double) (ig - first) * fd_h.x;
      s = I_t * src_abs * src_mu * exp(-src_mu * depth);
      s *= 1e10;
      s *= 6.3538562638e-26;
    }
    l1[i].source = l2[i].source = s;
  }
  return 0;
}

static int want_step(int s)
{
  const char *p = getenv("IMD_DUMP_STEPS");
  if (p == NULL) return 0;
  while (*p) {
    char *e; long v = strtol(p, &e, 10);
    if (e == p) break;
    if (v == s) return 1;
    p = (*e == ',') ? e + 1 : e;
  }
  return 0;
}

void __real_calc_ttm(void);

void __wrap_calc_ttm(void)
{
  const char *dir = getenv("IMD_DUMP_DIR");
  __real_calc_ttm();
  if (dir == NULL || !want_step(steps)) return;
  {
    char fn[1024];
    FILE *f;
    int i, L = local_fd_dim.x - 2, G = global_fd_dim.x;
    double hdr[16];
    snprintf(fn, sizeof fn, "%s/s%d.ttm.r%d.bin", dir, steps, myid);
    f = fopen(fn, "wb");
    if (f == NULL) { perror(fn); exit(3); }
    memset(hdr, 0, sizeof hdr);
    hdr[0] = G; hdr[1] = L; hdr[2] = myid; hdr[3] = num_cpus; hdr[4] = diff_substeps; hdr[5] = max_dt_ttm;
    hdr[6] = last_active_cell_global; hdr[7] = cur_vlattice_proc; hdr[8] = Eabs_local;
    hdr[9] = tot_elec_energy_local; hdr[10] = fd_h.x; hdr[11] = fd_vol; hdr[12] = neighvol;
    hdr[13] = laser_active ? 1.0 : 0.0; hdr[14] = tmm_time; hdr[15] = vlatdim;
    fwrite(hdr, sizeof(double), 16, f);
    /* 13 doubles per local cell */
    for (i = 1; i <= L; i++) {
      double r[13];
      r[0] = l1[i].natoms; r[1] = l1[i].natoms_old; r[2] = l1[i].temp; r[3] = l1[i].md_temp;
      r[4] = l1[i].xi; r[5] = l1[i].dens; r[6] = l1[i].fd_k; r[7] = l1[i].fd_g; r[8] = l1[i].Z;
      r[9] = l1[i].ne; r[10] = l1[i].Ce; r[11] = l1[i].U; r[12] = l1[i].source;
      fwrite(r, sizeof(double), 13, f);
    }
    fwrite(xiarr_global, sizeof(double), G, f);
    fwrite(vcomxglobal, sizeof(double), G, f);
    fwrite(vcomyglobal, sizeof(double), G, f);
    fwrite(vcomzglobal, sizeof(double), G, f);
    for (i = 0; i < G; i++) { double v = natomsglobal[i]; fwrite(&v, sizeof v, 1, f); }
    for (i = 0; i < G; i++) { double v = fluxfromleftglobal[i]; fwrite(&v, sizeof v, 1, f); }
    for (i = 0; i < G; i++) { double v = fluxfromrightglobal[i]; fwrite(&v, sizeof v, 1, f); }
    for (i = 0; i < vlatdim; i++) {
      double r[6];
      r[0] = vlattice1[i].temp; r[1] = vlattice1[i].md_temp; r[2] = vlattice1[i].fd_k;
      r[3] = vlattice1[i].fd_g; r[4] = vlattice1[i].Ce; r[5] = vlattice1[i].Z;
      fwrite(r, sizeof(double), 6, f);
    }
    fclose(f);
  }
}

#endif /* TTM && TTM1D */
