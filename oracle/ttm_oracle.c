/* oracle/ttm_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's 1-D
 * TTM (imd_ttm.c) and move_atoms_ttm; see ttm_oracle.h for scope and parity status. */
#include "ttm_oracle.h"

#include <complex.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* physical constants with the reference's digits (globals.h:324-330, 1190-1192) */
static const double ECONST = 8.841941282883075e-12;
static const double HBAR = 1.054571800139113e-34;
static const double EMASS = 9.109383560000000e-31;
static const double BOLTZMAN = 1.380648520000000e-23;
static const double ECHARGE = 1.602176620000000e-19;
static const double AMU = 1.660539040200000e-27;
static const double J2eV = 6.2415090744607626077762409809304458998869658961709711e18;
static const double eV2J = 1.602176634E-19;
#define RHOMIN 2                       /* imd_ttm.c:69 */

#define DMAX(a, b) ((a) > (b) ? (a) : (b))
#define DMIN(a, b) ((a) < (b) ? (a) : (b))
#define DABS(a) ((a) > 0 ? (a) : -(a))

/* ------------------------------------------------------------------ EOS tables
 * The reference interpolates linearly on a Delaunay triangulation of the table points
 * (nn library: delaunay_build + lpi_build, imd_interpol.c:1263-1267; lpi_interpolate_point,
 * imd_ttm.c:2517,2568).  For a rectilinear point set every triangle is half of a grid
 * rectangle; which half is recorded in diag[] (from oracle/nn_probe).  Linear interpolation
 * on a triangle = the plane through its three vertices. */
static int find_cell(const double *a, int n, double v)
{
  int lo = 0, hi = n - 1;                         /* a[lo] <= v <= a[hi] expected */
  if (!(v >= a[0]) || !(v <= a[n - 1])) return -1;
  while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (a[mid] <= v) lo = mid; else hi = mid; }
  return lo;
}

double orc_eos_interp(const orc_eos *e, double x, double y)
{
  int i = find_cell(e->x, e->nx, x), j = find_cell(e->y, e->ny, y);
  if (i < 0 || j < 0) return NAN;                  /* outside the hull: lpi gives NaN */
  const int ny = e->ny;
  const double x0 = e->x[i], x1 = e->x[i + 1], y0 = e->y[j], y1 = e->y[j + 1];
  const double z00 = e->z[i * ny + j], z01 = e->z[i * ny + j + 1];
  const double z10 = e->z[(i + 1) * ny + j], z11 = e->z[(i + 1) * ny + j + 1];
  const double u = (x - x0) / (x1 - x0), v = (y - y0) / (y1 - y0);
  const int d = e->diag ? e->diag[i * (ny - 1) + j] : 0;
  if (d == 0) {                                    /* diagonal (i,j)-(i+1,j+1) */
    if (v <= u) return z00 + u * (z10 - z00) + v * (z11 - z10);   /* triangle 00,10,11 */
    return z00 + u * (z11 - z01) + v * (z01 - z00);               /* triangle 00,01,11 */
  }
  /* diagonal (i+1,j)-(i,j+1) */
  if (u + v <= 1.0) return z00 + u * (z10 - z00) + v * (z01 - z00);   /* triangle 00,10,01 */
  return z11 + (1.0 - u) * (z01 - z11) + (1.0 - v) * (z10 - z11);    /* triangle 11,01,10 */
}

/* EOS_ee_from_r_te (imd_ttm.c:2483-2529): the range checks clamp local copies only AFTER the
 * query point was stored, so an out-of-range argument still reaches lpi and yields NaN. */
static double eos_ee_from_r_te(const orc_eos *ee, double r, double t)
{
  return orc_eos_interp(ee, r, sqrt(t));
}
/* EOS_cve_from_r_te (imd_ttm.c:2531-2592) */
static double eos_cve_from_r_te(const orc_eos *cve, double r, double t)
{
  double z = orc_eos_interp(cve, r, t);
  z *= r; z *= 11604.5; z *= 1e-30; z *= J2eV;
  return z;
}

/* fminbnd (fminbnd3.c:6-106) with the reference's macro semantics: SIGNUM is an
 * unparenthesised conditional (makros.h:338), so "p *= -SIGNUM(q)" multiplies by +sign(q). */
static double eeminfun(const orc_eos *ee, double x, double r, double e)
{
  double a = eos_ee_from_r_te(ee, r, x) - e;
  return DABS(a);
}
static double fminbnd_ref(const orc_eos *ee, double a, double b, double rho, double eng)
{
  const double tolx = 1e-5;
  const double c = 0.5 * (3 - sqrt(5));
  double v = a + c * (b - a), w = v, x = v, e = 0;
  double fv = eeminfun(ee, x, rho, eng), fw = fv, fval = fv, fu = 0.0;
  const int imax = 500;
  double xm, tol, r, q, p, d = 0.0, u;
  const double sqrteps = 2.220446049250313e-16;
  int dogs;
  for (int i = 0; i < imax + 1; i++) {
    xm = 0.5 * (a + b);
    tol = 2.0 * sqrteps * DABS(x) + tolx / 3.0;
    if (DABS(x - xm) <= (2 * tol - 0.5 * (b - a)) || i == imax) return x;
    if (DABS(e) > tol) {
      dogs = 0;
      r = (x - w) * (fval - fv);
      q = (x - v) * (fval - fw);
      p = (x - v) * q - (x - w) * r;
      q = 2 * (q - r);
      p *= (-(q > 0)) ? 1 : ((q < 0) ? -1 : 0);
      q = DABS(q);
      r = e;
      e = d;
      if (DABS(p) < DABS(0.5 * q * r) && p > q * (a - x) && p < q * (b - x)) {
        d = p / q;
        u = x + d;
        if (DMIN(u - a, b - u) < 2 * tol) d = tol * ((xm - x > 0) ? 1 : (((xm - x < 0) ? -1 : 0) + (xm == x)));
      } else dogs = 1;
    } else dogs = 1;
    if (dogs == 1) {
      e = (x >= xm ? a - x : b - x);
      d = c * e;
    }
    u = x + DMAX(DABS(d), tol) * ((d > 0) ? 1 : (((d < 0) ? -1 : 0) + (d == 0)));
    fu = eeminfun(ee, u, rho, eng);
    if (fu < fval) {
      if (u < x) b = x; else a = x;
      v = w; fv = fw; w = x; fw = fval; x = u; fval = fu;
    } else {
      if (u < x) a = u; else b = u;
      if (fu <= fw || w == x) { v = w; fv = fw; w = u; fw = fu; }
      else if (fu <= fv || v == x || v == w) { v = u; fv = fu; }
    }
  }
  return -1;
}

/* EOS_te_from_r_ee (imd_ttm.c:2699-2724) */
double orc_eos_te_from_r_ee(const orc_eos *ee, double r, double e)
{
  if (r < ee->x[0] || r > ee->x[ee->nx - 1]) { r = DMIN(r, ee->x[ee->nx - 1]); r = DMAX(r, ee->x[0]); }
  double a = pow(ee->y[0], 2.0), b = pow(ee->y[ee->ny - 1], 2.0);
  return fminbnd_ref(ee, a, b, r, e);
}

/* ------------------------------------------------------------------ wide-range models (imd_ttm.c:2191-2399) */
static double fermi_E(double ne) { return HBAR * HBAR * pow(3.0 * M_PI * M_PI * ne, 2.0 / 3.0) / 2.0 / EMASS; }
static double fermi_T(double ne) { return 2.0 * fermi_E(ne) / (3.0 * BOLTZMAN); }
static double omega_pl(double ne) { return sqrt(ne * ECHARGE * ECHARGE / EMASS / ECONST); }
static double r0_(double ni) { return pow(3.0 / (4 * M_PI * ni), 1.0 / 3.0); }
static double bMin(double Z, double Te)
{
  return DMAX(Z * ECHARGE * ECHARGE / (BOLTZMAN * Te), HBAR / sqrt(2 * EMASS * BOLTZMAN * Te));
}
static double bMax(double omega_las, double Z, double ni, double Te, int laser_active)
{
  double teff = sqrt(Te * Te + fermi_T(ni * Z) * fermi_T(ni * Z));
  if (laser_active) return DMAX(r0_(ni), sqrt(BOLTZMAN * teff / (EMASS)) / DMAX(omega_las, omega_pl(ni * Z)));
  return DMAX(r0_(ni), sqrt(BOLTZMAN * teff / (EMASS)) / omega_pl(ni * Z));
}
static double coulomb_log(double omega_las, double Z, double ni, double Te, int la)
{
  return DMAX(2.0, log(bMax(omega_las, Z, ni, Te, la) / bMin(Z, Te)));
}
double orc_ttm_mean_charge(double temp, double rho, double z0, double am)
{
  temp *= 1e-3; rho *= 1e-3;
  const double YA1 = 0.003323, YA2 = 0.971832, YA3 = 0.926148e-5, YA4 = 3.10165, YB0 = -1.7630, YB1 = 1.43175,
               YB2 = 0.315463, YC1 = 0.036666667, YC2 = 0.983333333, YAL = 0.143139e2, YBE = 0.6624e0;
  double t0 = temp / pow(z0, 4. / 3.);            /* the reference uses the global atomic_charge == z0 */
  double r0 = rho / (z0 * am);
  double ytf = t0 / (1.0 + t0);
  double ya = YA1 * pow(t0, YA2) + YA3 * pow(t0, YA4);
  double yb = -exp(YB0 + YB1 * ytf + YB2 * ytf * ytf);
  double yc = YC1 * ytf + YC2;
  double yq1 = ya * pow(r0, yb);
  double yq = pow(pow(r0, yc) + pow(yq1, yc), 1.0 / yc);
  double yx = YAL * pow(yq, YBE);
  return z0 * yx / (1.0 + yx + sqrt(1.0 + 2.0 * yx));
}
static double numet(double A1, double A2, double Te, double Ti, double TF)
{
  return 1.309202957843054e+11 * (A1 * Ti + A2 * Te * Te / TF);
}
static double numax(double A3, double vf, double Te, double ni)
{
  return A3 / r0_(ni) * sqrt(vf * vf + BOLTZMAN * Te / EMASS);
}
static double nupl(double omega_las, double Z, double ni, double ne, double Te, int la)
{
  double coulomb = coulomb_log(omega_las, Z, ni, Te, la);
  return 1.863839656495274e-40 * Z * ne * coulomb / pow(BOLTZMAN * Te, 1.5);
}
double orc_ttm_gamma(double Te, double Ti, double Ne, double Z, double omega_laser, int la)
{
  Te *= 11604.5; Ti *= 11604.5;
  const double A1g = 50, A2g = 20, A3g = 0.25, m_atom = 26.9815;
  double Ni = Ne / Z;
  double EF = fermi_E(Ne);
  double TF = 4.828648689433765e+22 * EF;
  double VF = 1.481734876966785e+15 * sqrt(EF);
  double nu_eff = DMIN(numet(A1g, A2g, Te, Ti, TF), DMIN(numax(A3g, VF, Te, Ni), nupl(omega_laser, Z, Ni, Ne, Te, la)));
  return 2.783313120645740e-74 / (m_atom * AMU) * Ne * nu_eff;
}
double orc_ttm_kappa(double Te, double Ti, double Ne, double Z, double omega_laser, int la)
{
  Te *= 11604.5; Ti *= 11604.5;
  const double A1t = 2.95, A2t = 0.5, A3t = 0.16, A4t = 1.2;
  double Ni = Ne / Z;
  double EF = fermi_E(Ne);
  double TF = 4.828648689433765e+22 * EF;
  double VF = 1.481734876966785e+15 * sqrt(EF);
  double nu_eff = DMIN(numet(A1t, A2t, Te, Ti, TF), numax(A3t, VF, Te, Ni));
  double coulomb = coulomb_log(omega_laser, Z, Ni, Te, la);
  double kappa_met = 6.884236239621913e-16 * Ne * Te / nu_eff;
  double kappa_pl = 4.428788911416808e+43 * pow(BOLTZMAN * Te, 2.5) / Z / coulomb;
  double kappa_wr = kappa_pl + (kappa_met - kappa_pl) * exp(-A4t * Te / TF);
  return kappa_wr * 0.073768115942029;
}

/* ------------------------------------------------------------------ state */
typedef struct {                     /* ttm_Element (types.h:441-492), TTM1D fields used on this path */
  int    natoms, natoms_old;
  double temp, xi, md_temp, source, dens, fd_k, fd_g, Z, Ce, ne, U;
} tcell;

struct orc_ttm {
  orc_ttm_params p;
  orc_eos cve, ee;
  int G, P, L, LD;                   /* LD = L + 2 = local_fd_dim.x */
  tcell *l1, *l2;                    /* [P*LD], swapped by pointer like the reference's l1/l2 */
  tcell *v1, *v2;                    /* virtual lattice (one copy: only its owner rank ever touches it) */
  double *xiarr_local;               /* [P*LD] */
  double *xiarr_global, *vcomx, *vcomy, *vcomz, *mdtemp;   /* [G] */
  int    *natomsglobal, *totneighs, *fluxl, *fluxr;        /* [G] */
  double fd_hx, fd_vol, neighvol, omega_laser, laser_sigma_t_squared;
  double max_dt_ttm, tmm_time, Eabs, tot_elec_energy;
  int    diff_substeps, laser_active, steps;
  int    last_active_cell_global, cur_vlattice_proc;
  int    use_tmm;                    /* orc_ttm_set_tmm called */
  orc_tmm_params tmm;
  double *tmm_Qabs;                  /* [G], persists between calls like the reference's global array */
  double tmm_refl, tmm_trans;
  char   err[512];
};

#define CELL(lat, r, i) ((lat)[(r) * t->LD + (i)])
#define ACTIVE_GT(c) ((c).natoms >= t->p.fd_min_atoms && (c).dens > RHOMIN)    /* "dens > RHOMIN" sites  */
#define ACTIVE_GE(c) ((c).natoms >= t->p.fd_min_atoms && (c).dens >= RHOMIN)   /* "dens >= RHOMIN" sites */

const char *orc_ttm_error(const orc_ttm *t) { return t->err; }

orc_ttm *orc_ttm_create(const orc_ttm_params *p, const orc_eos *cve, const orc_eos *ee)
{
  if (p->G < 1 || p->P < 1 || p->G % p->P != 0) return NULL;      /* imd_ttm.c:964-967 */
  orc_ttm *t = calloc(1, sizeof *t);
  t->p = *p; t->cve = *cve; t->ee = *ee;
  t->G = p->G; t->P = p->P; t->L = p->G / p->P; t->LD = t->L + 2;
  const int nl = t->P * t->LD;
  t->l1 = calloc(nl, sizeof(tcell)); t->l2 = calloc(nl, sizeof(tcell));
  t->v1 = calloc(p->vlatdim > 0 ? p->vlatdim : 1, sizeof(tcell));
  t->v2 = calloc(p->vlatdim > 0 ? p->vlatdim : 1, sizeof(tcell));
  t->xiarr_local = calloc(nl, sizeof(double));
  t->xiarr_global = calloc(t->G, sizeof(double));
  t->vcomx = calloc(t->G, sizeof(double)); t->vcomy = calloc(t->G, sizeof(double)); t->vcomz = calloc(t->G, sizeof(double));
  t->mdtemp = calloc(t->G, sizeof(double));
  t->natomsglobal = calloc(t->G, sizeof(int)); t->totneighs = calloc(t->G, sizeof(int));
  t->fluxl = calloc(t->G, sizeof(int)); t->fluxr = calloc(t->G, sizeof(int));
  /* init_ttm (imd_ttm.c:975-981, 1036-1041) */
  t->fd_hx = p->box_x / t->G;
  t->fd_vol = t->fd_hx * p->box_y * p->box_z;
  t->max_dt_ttm = p->timestep / ((double) p->fd_n_timesteps);
  t->neighvol = pow(sqrt(p->r2_cut_pair), 3.0) * 4.0 / 3.0 * M_PI;
  for (int i = 0; i < p->vlatdim; i++) {
    t->v1[i].dens = t->v2[i].dens = p->vlatdens;
    t->v1[i].temp = t->v2[i].temp = 0.0264;
    t->v1[i].md_temp = t->v2[i].md_temp = 0.0264;
  }
  /* tmm_init as replaced by oracle/ref_hooks_ttm.c */
  t->laser_sigma_t_squared = p->laser_sigma_t * p->laser_sigma_t;
  t->omega_laser = 2.0 * M_PI * 3.0e8 / p->lambda;               /* c0 = 3.0e8, globals.h:499 */
  t->diff_substeps = 100;                                         /* globals.h:1187 */
  t->last_active_cell_global = 0;
  return t;
}

void orc_ttm_set_tmm(orc_ttm *t, const orc_tmm_params *q)
{
  t->use_tmm = 1;
  t->tmm = *q;
  free(t->tmm_Qabs);
  t->tmm_Qabs = calloc(t->G, sizeof(double));
}

void orc_ttm_destroy(orc_ttm *t)
{
  if (!t) return;
  free(t->tmm_Qabs);
  free(t->l1); free(t->l2); free(t->v1); free(t->v2); free(t->xiarr_local); free(t->xiarr_global);
  free(t->vcomx); free(t->vcomy); free(t->vcomz); free(t->mdtemp);
  free(t->natomsglobal); free(t->totneighs); free(t->fluxl); free(t->fluxr);
  free(t);
}

static int iglobal_of(const orc_ttm *t, int r, int i) { return (i - 1) + r * t->L; }
static int slab_of(const orc_ttm *t, double x)
{
  int ig = (int) (x / t->fd_hx);                                 /* imd_ttm.c:277-279 */
  ig = ig < t->G - 1 ? ig : t->G - 1;
  ig = ig > 0 ? ig : 0;
  return ig;
}
static double dens_from_neigh(const orc_ttm *t, int totneigh, int natoms)
{
  return ((double) totneigh) / ((double) natoms) * t->p.atomic_weight / t->neighvol * 1660.53907;   /* :418 */
}

/* update_fd (imd_ttm.c:161-548) */
static void update_fd(orc_ttm *t, long n, const double *pos, const double *mom, const double *mass,
                      const int *type, const int *numneighs, int *fdi)
{
  const int G = t->G, L = t->L;
  for (int g = 0; g < G; g++) {
    t->natomsglobal[g] = t->totneighs[g] = t->fluxl[g] = t->fluxr[g] = 0;
    t->vcomx[g] = t->vcomy[g] = t->vcomz[g] = t->mdtemp[g] = 0.0;
    t->xiarr_global[g] = 0.0;
  }
  int last_active_cell_local = -5000;
  for (int r = 0; r < t->P; r++) {
    CELL(t->l1, r, 0).temp = CELL(t->l1, r, L + 1).temp = 0;                      /* :241-242 */
    for (int i = 1; i <= L; i++) {                                                 /* :248-264 */
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      a->source = b->source = 0.0;
      a->natoms_old = b->natoms_old = a->natoms;
      a->natoms = b->natoms = 0;
      a->xi = b->xi = 0.0;
      t->xiarr_local[r * t->LD + i] = 0.0;
      a->dens = b->dens = 0;
    }
  }
  for (long k = 0; k < n; k++) {                                                   /* :269-318 */
    if (type[k] != 0) continue;
    const int ig = slab_of(t, pos[3 * k]);
    t->natomsglobal[ig]++;
    t->totneighs[ig] += numneighs[k];
    t->vcomx[ig] += mom[3 * k]; t->vcomy[ig] += mom[3 * k + 1]; t->vcomz[ig] += mom[3 * k + 2];
    if (t->steps > 0 && fdi[k] != ig) {
      if (fdi[k] > ig) t->fluxr[ig]++; else if (fdi[k] < ig) t->fluxl[ig]++;
    }
    fdi[k] = ig;
  }
  for (int g = 0; g < G; g++) {                                                    /* :341-363 */
    double totmass = (double) t->natomsglobal[g] * t->p.atomic_weight;
    if (totmass > 0) { t->vcomx[g] /= totmass; t->vcomy[g] /= totmass; t->vcomz[g] /= totmass; }
    else t->vcomx[g] = t->vcomy[g] = t->vcomz[g] = 0.0;
  }
  for (long k = 0; k < n; k++) {                                                   /* :368-394 */
    if (type[k] != 0) continue;
    const int ig = slab_of(t, pos[3 * k]);
    const double m = mass[k];
    double d;
    d = mom[3 * k] / m - t->vcomx[ig];     t->mdtemp[ig] += m * (d * d);
    d = mom[3 * k + 1] / m - t->vcomy[ig]; t->mdtemp[ig] += m * (d * d);
    d = mom[3 * k + 2] / m - t->vcomz[ig]; t->mdtemp[ig] += m * (d * d);
  }
  for (int r = 0; r < t->P; r++)                                                   /* :401-472 */
    for (int i = 1; i <= L; i++) {
      const int g = iglobal_of(t, r, i);
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      a->natoms = t->natomsglobal[g]; b->natoms = a->natoms;
      if (a->natoms > 0) {
        a->dens = dens_from_neigh(t, t->totneighs[g], a->natoms);
        if (a->dens == 0) a->dens = (double) a->natoms * t->p.atomic_weight / t->fd_vol * 1660.53907;
        a->md_temp = t->mdtemp[g];
        a->md_temp /= (3.0 * (double) a->natoms);
        if (ACTIVE_GT(*a)) { int c = g - t->p.vlatbuffer; if (c > last_active_cell_local) last_active_cell_local = c; }
      } else { a->dens = 0.0; a->md_temp = 0.0; a->natoms = b->natoms = 0; }
      b->md_temp = a->md_temp; b->dens = a->dens;
      if (i == 1) {                                                                /* ghost cells, :451-471 */
        tcell *ga = &CELL(t->l1, r, 0), *gb = &CELL(t->l2, r, 0);
        ga->natoms = g - 1 >= 0 ? t->natomsglobal[g - 1] : 0;
        if (ga->natoms > 0) ga->dens = dens_from_neigh(t, t->totneighs[g - 1], ga->natoms);
        gb->natoms = ga->natoms; gb->dens = ga->dens;
      } else if (i == L) {
        tcell *ga = &CELL(t->l1, r, L + 1), *gb = &CELL(t->l2, r, L + 1);
        ga->natoms = g + 1 < G ? t->natomsglobal[g + 1] : 0;
        if (ga->natoms > 0) ga->dens = dens_from_neigh(t, t->totneighs[g + 1], ga->natoms);
        gb->natoms = ga->natoms; gb->dens = ga->dens;
      }
    }
  if (t->steps < 1)                                                                /* :495-505 */
    for (int r = 0; r < t->P; r++)
      for (int i = 1; i <= L; i++) {
        tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
        a->natoms_old = b->natoms_old = a->natoms;
        if (ACTIVE_GT(*a)) a->temp = b->temp = a->md_temp;
      }
  if (t->steps == 0) t->last_active_cell_global = last_active_cell_local;          /* :516-518 (MAX over ranks) */
  t->cur_vlattice_proc = (int) (((double) t->last_active_cell_global * t->P) / ((double) G));   /* :521 */
  for (int r = 0; r < t->P; r++)                                                   /* :536-541 */
    for (int i = 1; i <= L; i++)
      if (iglobal_of(t, r, i) > t->last_active_cell_global) CELL(t->l1, r, i).natoms = CELL(t->l2, r, i).natoms = -1;
}

static void swap_lattices(orc_ttm *t) { tcell *x = t->l1; t->l1 = t->l2; t->l2 = x; }

/* do_COMMFLUX (imd_ttm.c:858-949): ghost U from the x-neighbour ranks */
static void do_commflux(orc_ttm *t)
{
  const int L = t->L, P = t->P;
  for (int r = 0; r < P; r++) {
    if (r > 0) CELL(t->l1, r, 0).U = CELL(t->l1, r - 1, L).U;
    if (r < P - 1) CELL(t->l1, r, L + 1).U = CELL(t->l1, r + 1, 1).U;
  }
}

/* do_ADV (imd_ttm.c:1208-1301) */
static void do_adv(orc_ttm *t, double tau)
{
  if (t->steps < 1) return;
  const int L = t->L, G = t->G;
  do_commflux(t);
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= L; i++) {
      const int g = iglobal_of(t, r, i);
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      const double Nold = (double) a->natoms_old, Nnew = (double) a->natoms;
      if (Nnew > 0) {
        const double fr_g = (double) t->fluxr[g], fl_g = (double) t->fluxl[g];
        const double fl_g1 = g + 1 < G ? (double) t->fluxl[g + 1] : 0.0;
        const double fr_gm = g - 1 >= 0 ? (double) t->fluxr[g - 1] : 0.0;
        b->U = a->U * Nold / Nnew + tau * (+fr_g * CELL(t->l1, r, i + 1).U - fl_g1 * a->U
                                           + fl_g * CELL(t->l1, r, i - 1).U - fr_gm * a->U) / Nnew;
        if (ACTIVE_GT(*a))
          b->temp = orc_eos_te_from_r_ee(&t->ee, a->dens, b->U / (26.9815 * AMU * J2eV)) / 11604.5;
      } else if (Nnew < 1) { b->U = 0.0; b->temp = 0.0; }
    }
  swap_lattices(t);
}

/* do_cell_activation (imd_ttm.c:1311-1425) */
static void do_cell_activation(orc_ttm *t)
{
  const int L = t->L, mn = t->p.fd_min_atoms;
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= L; i++) {
      const int g = iglobal_of(t, r, i);
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      if (a->natoms_old >= mn && a->natoms < mn && a->dens > RHOMIN) a->xi = 0.0;
      else if (a->natoms_old < mn && a->natoms >= mn && a->dens > RHOMIN) {
        if (g > t->last_active_cell_global) continue;
        if (isnan(a->temp) != 0 || a->temp <= t->p.Temin) {
          int n_neighbors = 0;
          double E_el_neighbors = 0.0;
          const tcell *xp = &CELL(t->l1, r, i + 1), *xm = &CELL(t->l1, r, i - 1);
          if (ACTIVE_GE(*xp)) { E_el_neighbors += xp->temp * xp->temp; n_neighbors++; }
          if (ACTIVE_GE(*xm)) { E_el_neighbors += xm->temp * xm->temp; n_neighbors++; }
          if (n_neighbors != 0) {
            a->temp = sqrt(E_el_neighbors / ((double) n_neighbors));
            b->temp = a->temp;
            if (a->temp < t->p.Temin) b->temp = a->temp = a->md_temp;
          } else { a->temp = a->md_temp; b->temp = a->temp; }
          a->U = b->U = eos_ee_from_r_te(&t->ee, a->dens, a->temp * 11604.5) * 26.9815 * AMU * J2eV;
        }
      }
    }
}

static void fill_props(orc_ttm *t, tcell *c)                     /* the vlattice part, imd_ttm.c:828-844 */
{
  c->Z = orc_ttm_mean_charge(c->temp * 11604.5, c->dens, t->p.atomic_charge, t->p.atomic_weight);
  c->ne = c->Z * c->dens / (t->p.atomic_weight * AMU);
  c->fd_k = orc_ttm_kappa(c->temp, c->md_temp, c->ne, c->Z, t->omega_laser, t->laser_active);
  c->Ce = eos_cve_from_r_te(&t->cve, c->dens, c->temp * 11604.5);
  c->fd_g = orc_ttm_gamma(c->temp, c->md_temp, c->ne, c->Z, t->omega_laser, t->laser_active);
}

/* do_FILLMESH (imd_ttm.c:629-852) */
static int do_fillmesh(orc_ttm *t)
{
  const int L = t->L;
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= L; i++) {
      const int g = iglobal_of(t, r, i);
      if (g > t->last_active_cell_global) continue;
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      if (ACTIVE_GT(*a)) {
        a->Z = orc_ttm_mean_charge(a->temp * 11604.5, a->dens, t->p.atomic_charge, t->p.atomic_weight);
        a->ne = a->Z * a->dens / (t->p.atomic_weight * AMU);
        b->ne = a->ne;
        a->Ce = eos_cve_from_r_te(&t->cve, a->dens, a->temp * 11604.5);
        b->Ce = a->Ce;
        if (a->Ce < 0 || isnan(a->Ce)) { snprintf(t->err, sizeof t->err, "ERROR during CfromT in FILLMESH (cell %d, Te %g K, dens %g)", g, a->temp * 11604.5, a->dens); return -1; }
        a->fd_k = orc_ttm_kappa(a->temp, a->md_temp, a->ne, a->Z, t->omega_laser, t->laser_active);
        b->fd_k = a->fd_k;
        if (isnan(a->fd_k) != 0 || a->fd_k < 0) { snprintf(t->err, sizeof t->err, "fd_k is NaN (cell %d)", g); return -1; }
        a->fd_g = orc_ttm_gamma(a->temp, a->md_temp, a->ne, a->Z, t->omega_laser, t->laser_active);
        b->fd_g = a->fd_g;
        if (isnan(a->fd_g) != 0 || a->fd_g < 0) { snprintf(t->err, sizeof t->err, "fd_g is NaN (cell %d)", g); return -1; }
        if (t->steps < 1) {                                                       /* :765-807 */
          a->U = eos_ee_from_r_te(&t->ee, a->dens, a->temp * 11604.5) * 26.9815 * AMU * J2eV;
          b->U = a->U;
          double echeck = eos_ee_from_r_te(&t->ee, a->dens, a->temp * 11604.5) * 26.9815 * AMU * J2eV;
          double tcheck = orc_eos_te_from_r_ee(&t->ee, a->dens, echeck / 26.9815 / AMU * eV2J) / 11604.5;
          double tinit = a->temp;
          if (DABS(tcheck - tinit) > tinit * 0.01) { snprintf(t->err, sizeof t->err, "EOS Plausibility check failed, Tinit %g TfromU %g", tinit, tcheck); return -1; }
        }
      } else { a->fd_k = 0.0; a->fd_g = 0.0; a->Z = 0.0; a->ne = 0.0; a->Ce = 0.0; a->temp = 0.0; }
    }
  if (t->cur_vlattice_proc >= 0 && t->cur_vlattice_proc < t->P)
  for (int i = 0; i < t->p.vlatdim; i++) {                                        /* :825-845 */
    fill_props(t, &t->v1[i]);
    t->v2[i].Z = t->v1[i].Z; t->v2[i].ne = t->v1[i].ne; t->v2[i].fd_k = t->v1[i].fd_k; t->v2[i].fd_g = t->v1[i].fd_g;
  }
  return 0;
}

/* ttm_fill_ghost_layers (imd_ttm.c:3264-3356) */
static void fill_ghost_layers(orc_ttm *t)
{
  const int L = t->L, P = t->P;
  for (int r = 0; r < P; r++) {
    if (r < P - 1) { CELL(t->l1, r, L + 1).temp = CELL(t->l1, r + 1, 1).temp; CELL(t->l1, r, L + 1).fd_k = CELL(t->l1, r + 1, 1).fd_k; }
    if (r > 0) { CELL(t->l1, r, 0).temp = CELL(t->l1, r - 1, L).temp; CELL(t->l1, r, 0).fd_k = CELL(t->l1, r - 1, L).fd_k; }
    if (r == 0) CELL(t->l1, r, 0).natoms = 0;
    if (r == P - 1) CELL(t->l1, r, L + 1).natoms = 0;      /* for P == 1 the reference cannot run; both ends open */
  }
}

/* CFL_maxdt (imd_ttm.c:2444-2480) */
static void cfl_maxdt(orc_ttm *t)
{
  double maxdttmp = 1e9;
  const double dxsq = t->fd_hx * t->fd_hx;
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= t->L; i++) {
      const tcell *a = &CELL(t->l1, r, i);
      if (ACTIVE_GT(*a)) {
        const tcell *xp = &CELL(t->l1, r, i + 1);
        const tcell *m = ACTIVE_GE(*xp) ? xp : a;
        double khalf = (a->fd_k + m->fd_k);
        maxdttmp = DMIN(maxdttmp, a->Ce * dxsq / khalf);
      }
    }
  if (t->steps > 0) t->max_dt_ttm = maxdttmp; else t->max_dt_ttm = t->p.timestep / ((double) t->p.fd_n_timesteps);
  t->max_dt_ttm *= 0.9;
}


/* ------------------------------------------------------------------ transfer-matrix source (imd_tmm.c) */
/* tmm_K1 / tmm_K2 (imd_tmm.c:667-696); note (int) nu * 30 - 1: the cast binds to nu alone */
static double tmm_K(const orc_ttm *t, double nu, int which)
{
  const double *K12 = t->tmm.K12;
  if (nu < 1.0 / 30.0) return 1;
  if (nu < 50) {
    int i = (int) nu * 30 - 1;
    if (i < 0) i = 0;
    if (i > 1498) i = 1498;
    return K12[2 * i + which] + (30 * nu - i - 1) * (K12[2 * (i + 1) + which] - K12[2 * i + which]);
  }
  if (which == 0) return 315.0 / 8.0 / (nu * nu) - 675675.0 / 64.0 / (nu * nu * nu * nu);
  return 6.0 / (nu * nu) - 720.0 / (nu * nu * nu * nu);
}

/* nupl as called from tmm_get_epsilon: the file-level nupl/coulomb_log of imd_ttm.c with the global laser_active */
static double nupl_(double omega_las, double Z, double ni, double ne, double Te, int la)
{
  return nupl(omega_las, Z, ni, ne, Te, la);
}

/* tmm_get_epsilon (imd_tmm.c:503-570) */
static void tmm_get_epsilon(const orc_ttm *t, double lambda, double Te, double Ti, double Z, double Ne,
                            double *re, double *img)
{
  Te *= 11604.52500617; Ti *= 11604.52500617;
  const double complex un = 1, im = I;
  const double LIGHTSPEED = 2.998e8;                       /* globals.h:325 */
  double omega_las = (2.0 * M_PI * LIGHTSPEED / lambda);
  double Ncr = ECONST * EMASS * omega_las * omega_las / ECHARGE / ECHARGE;    /* ne_cr, imd_ttm.c:2219 */
  double Ni = Ne / Z;
  double EF = fermi_E(Ne);
  double TF = 2.0 * EF / (3.0 * BOLTZMAN);
  double VF = sqrt(2 * EF / EMASS);
  const double A1p = 4.41, A2p = 0.8, A3p = 0.7, A4p = 0.2;
  const double *bb = t->tmm.eps_bb;
  int il = 0, ir = 0, id;
  for (id = 0; lambda * 1e6 > bb[3 * id]; id++) { il = id; ir = id + 1; }
  double nu_met = numet(A1p, A2p, Te, Ti, TF);
  double nu_max = numax(A3p, VF, Te, Ni);
  double nu_pl = nupl_(omega_las, Z, Ni, Ne, Te, t->laser_active);
  double ksi = 0.75 * sqrt(M_PI) * nu_pl / omega_las;
  double complex ebl = bb[3 * il + 1] + I * bb[3 * il + 2], ebr = bb[3 * ir + 1] + I * bb[3 * ir + 2];
  double complex epsilon_bb = ebl + (ebr - ebl) / (bb[3 * ir] - bb[3 * il]) * (lambda * 1e6 - bb[3 * il]);
  int phase = ((int) orc_eos_interp(&t->tmm.phase, Ni * 26.9815 * AMU, Ti));
  if (fabs(phase) == 3 || fabs(phase) == 4 || fabs(phase) == 5) epsilon_bb = 0 * I * 0;
  double complex epsilon_met = epsilon_bb + un - Ne / Ncr / (un + im * DMIN(nu_met, nu_max) / omega_las);
  double complex epsilon_pl = un - Ne / Ncr * (tmm_K(t, ksi, 0) - im * nu_pl / omega_las * tmm_K(t, ksi, 1));
  double complex epsilon_wr = epsilon_pl + (epsilon_met - epsilon_pl) * exp(-A4p * Te / TF);
  *re = creal(epsilon_wr);
  *img = cimag(epsilon_wr);
}

/* EE (imd_tmm.c:72-98) */
static double tmm_EE(double z, double delta, double complex kl, double complex Bplus, double complex Bminus)
{
  double complex im = 0 + I * 1;
  double phi1 = creal(delta * z * kl * im);
  double phi2 = cimag(delta * z * kl * im);
  double complex eiphi = exp(phi1) * cos(phi2) + I * exp(phi1) * sin(phi2);
  double complex Ez = (Bplus * eiphi + Bminus / eiphi);
  return creal(Ez) * creal(Ez) + cimag(Ez) * cimag(Ez) + 0 * 0 + 0 * 0;
}

/* Runge5 (imd_tmm.c:37-69): adaptive Simpson-like quadrature of EE over z in [0,1] */
static double tmm_runge5(double delta, double complex kl, double complex Bplus, double complex Bminus)
{
  double dx = 1, cur_xpos = 0, result = 0;
  const double errval = 1e-5;
  double k1, k3, k4, k5;
  while (cur_xpos < 1 && dx > 1e-5) {
    k1 = dx / 3.0 * tmm_EE(cur_xpos, delta, kl, Bplus, Bminus);
    k3 = dx / 3.0 * tmm_EE(cur_xpos + dx / 3.0, delta, kl, Bplus, Bminus);
    k4 = dx / 3.0 * tmm_EE(cur_xpos + 0.5 * dx, delta, kl, Bplus, Bminus);
    k5 = dx / 3.0 * tmm_EE(cur_xpos + dx, delta, kl, Bplus, Bminus);
    double ERR = k1 - 4.5 * k3 + 4 * k4 - 0.5 * k5;
    if (ERR < 5.0 / 32 * errval) { result += 0.5 * (k1 + 4 * k4 + k5); cur_xpos += dx; dx *= 1.1; }
    else if (ERR > 5 * errval) { dx *= 0.5; continue; }
    else { result += 0.5 * (k1 + 4 * k4 + k5); cur_xpos += dx; }
    if (cur_xpos + dx > 1) dx = 1 - cur_xpos;
  }
  return result;
}

static void tmm_matmul(const double complex *a, const double complex *b, double complex *p)
{
  p[0] = a[0] * b[0] + a[2] * b[1];
  p[1] = a[1] * b[0] + a[3] * b[1];
  p[2] = a[0] * b[2] + a[2] * b[3];
  p[3] = a[1] * b[2] + a[3] * b[3];
}


/* do_tmm (imd_tmm.c:193-445).  The P ranks of the reference fill disjoint parts of the permittivity
 * arrays, rank 0 reduces them and solves, MPI_Scatter hands every rank its part of tmm_Qabs: here the
 * arrays are global from the start.  Kept as in the reference: nothing happens for steps < 2; when the
 * field is below the threshold the function returns WITHOUT touching the sources (they keep their last
 * values); only tmm_Qabs[0..ecut) is rewritten, entries beyond keep what an earlier call left. */
static int do_tmm_ref(orc_ttm *t)
{
  const double Imp = 3.769911184307751e+02;
  const int G = t->G, L = t->L;
  double I_t = t->p.I0 * exp(-pow(t->tmm_time - t->p.laser_t_0, 2) / t->laser_sigma_t_squared);
  I_t += t->p.I0 * exp(-pow(t->tmm_time - t->tmm.laser_t_1, 2) / (t->tmm.laser_sigma_t1 * t->tmm.laser_sigma_t1));
  if (t->steps < 2) return 0;
  if (sqrt(2.0 * I_t * Imp) < t->p.tmm_laser_threshold * sqrt(2 * t->p.I0 * Imp)) { t->laser_active = 0; return 0; }
  t->laser_active = 1;

  double *eps_re = calloc(G, sizeof(double)), *eps_im = calloc(G, sizeof(double));
  int *active = calloc(G, sizeof(int));
  double complex *km = calloc(G, sizeof(double complex));
  for (int r = 0; r < t->P; r++) for (int i = 1; i <= L; i++) {
    const tcell *c = &CELL(t->l1, r, i);
    const int ig = iglobal_of(t, r, i);
    if (c->natoms < t->p.fd_min_atoms || c->dens < RHOMIN) continue;
    tmm_get_epsilon(t, t->p.lambda, c->temp, c->md_temp, c->Z, c->ne, &eps_re[ig], &eps_im[ig]);
    active[ig] = 1;
  }
  double k0 = 2.0 * M_PI / t->p.lambda;
  k0 /= 1e10;
  double complex Cinv[4], Pmat[4], tmpmat[4], tmpmat2[4], Emat[4];
  double complex kl, kr, eiphi, BR = 0.0 + I * 0, BT = 0.0 + I * 0, B0 = 1.0 + I * 0;
  double phi1, phi2, dx;
  int ecut = 0;
  const double ecut_thresh = exp(-t->tmm.tmm_threshold);
  km[0] = k0;
  for (int i = 1; i < G - 1; i++)
    km[i] = active[i] == 0 ? k0 : k0 * csqrt(eps_re[i] + I * eps_im[i]);
  Emat[0] = 1.0 + I * 0; Emat[1] = 0; Emat[2] = 0; Emat[3] = 1.0 + I * 0.0;
  int tooshort = 1;
  for (int i = 0; i < G - 1; i++) {
    dx = i == 0 ? 1e10 : t->fd_hx;
    kl = km[i]; kr = km[i + 1];
    phi1 = creal(dx * kl * I); phi2 = cimag(dx * kl * I);
    eiphi = exp(phi1) * cos(phi2) + I * exp(phi1) * sin(phi2);
    Pmat[0] = eiphi; Pmat[2] = 1.0 / eiphi; Pmat[1] = eiphi; Pmat[3] = -Pmat[2];
    Cinv[0] = 0.5; Cinv[2] = 0.5 * kl / kr; Cinv[1] = 0.5; Cinv[3] = -Cinv[2];
    tmm_matmul(Cinv, Pmat, tmpmat);
    tmm_matmul(tmpmat, Emat, tmpmat2);
    for (int j = 0; j < 4; j++) Emat[j] = tmpmat2[j];
    BR = -Emat[1] * B0 / Emat[3];
    BT = Emat[0] * B0 + Emat[2] * BR;
    if (creal(BT) * creal(BT) + cimag(BT) * cimag(BT) < ecut_thresh) { tooshort = 0; ecut = i + 1; break; }
  }
  if (tooshort) {
    snprintf(t->err, sizeof t->err, "Sample is too short for TMM.Consider reducing tmm_threshold.");
    free(eps_re); free(eps_im); free(active); free(km);
    return -1;
  }
  {
    double complex tmpr0 = BR / B0;
    t->tmm_refl = creal(tmpr0) * creal(tmpr0) + cimag(tmpr0) * cimag(tmpr0);
    t->tmm_trans = creal(BT) * creal(BT) + cimag(BT) * cimag(BT);
  }
  double complex *Bplus = calloc(ecut + 1, sizeof(double complex)), *Bminus = calloc(ecut + 1, sizeof(double complex));
  double complex Fvec[2], tmpvec[2];
  Fvec[0] = B0; Fvec[1] = BR; Bplus[0] = B0; Bminus[0] = BR;
  for (int i = 0; i < ecut; i++) {
    dx = i == 0 ? 1e10 : t->fd_hx;
    kl = km[i]; kr = km[i + 1];
    phi1 = creal(dx * kl * I); phi2 = cimag(dx * kl * I);
    eiphi = exp(phi1) * cos(phi2) + I * exp(phi1) * sin(phi2);
    Pmat[0] = eiphi; Pmat[2] = 1.0 / eiphi; Pmat[1] = eiphi; Pmat[3] = -Pmat[2];
    Cinv[0] = 0.5; Cinv[2] = 0.5 * kl / kr; Cinv[1] = 0.5; Cinv[3] = -Cinv[2];
    tmm_matmul(Cinv, Pmat, tmpmat);
    tmpvec[0] = tmpmat[0] * Fvec[0] + tmpmat[2] * Fvec[1];       /* matmul on a 2-vector: first column only */
    tmpvec[1] = tmpmat[1] * Fvec[0] + tmpmat[3] * Fvec[1];
    Fvec[0] = tmpvec[0]; Fvec[1] = tmpvec[1];
    Bplus[i + 1] = Fvec[0]; Bminus[i + 1] = Fvec[1];
  }
  for (int i = 0; i < ecut; i++) {
    dx = i == 0 ? 1e10 : t->fd_hx;
    kl = km[i];
    t->tmm_Qabs[i] = I_t * k0 * eps_im[i] * tmm_runge5(dx, kl, Bplus[i], Bminus[i]);
    t->tmm_Qabs[i] *= 1e10;
    t->tmm_Qabs[i] *= 6.3538562638e-26;
  }
  free(Bplus); free(Bminus);
  for (int r = 0; r < t->P; r++) for (int i = 1; i <= L; i++)
    CELL(t->l1, r, i).source = CELL(t->l2, r, i).source = t->tmm_Qabs[iglobal_of(t, r, i)];
  free(eps_re); free(eps_im); free(active); free(km);
  return 0;
}

/* probes for the host test of imd_b200/csrc/tmm.cuh (tests/test_tmm_cpu.py) */
void orc_tmm_probe_epsilon(orc_ttm *t, double Te, double Ti, double Z, double Ne, double out2[2])
{
  const int la = t->laser_active;
  t->laser_active = 1;                     /* do_tmm sets it before the permittivities are evaluated */
  tmm_get_epsilon(t, t->p.lambda, Te, Ti, Z, Ne, &out2[0], &out2[1]);
  t->laser_active = la;
}
double orc_tmm_probe_runge5(double delta, const double kl[2], const double bp[2], const double bm[2])
{
  return tmm_runge5(delta, kl[0] + I * kl[1], bp[0] + I * bp[1], bm[0] + I * bm[1]);
}
/* one interface: tmpmat = Cinv * Pmat (imd_tmm.c:299-313), out[8] = re/im of the four entries */
void orc_tmm_probe_interface(double dx, const double kl_[2], const double kr_[2], double out[8])
{
  double complex kl = kl_[0] + I * kl_[1], kr = kr_[0] + I * kr_[1], Cinv[4], Pmat[4], M[4];
  double phi1 = creal(dx * kl * I), phi2 = cimag(dx * kl * I);
  double complex eiphi = exp(phi1) * cos(phi2) + I * exp(phi1) * sin(phi2);
  Pmat[0] = eiphi; Pmat[2] = 1.0 / eiphi; Pmat[1] = eiphi; Pmat[3] = -Pmat[2];
  Cinv[0] = 0.5; Cinv[2] = 0.5 * kl / kr; Cinv[1] = 0.5; Cinv[3] = -Cinv[2];
  tmm_matmul(Cinv, Pmat, M);
  for (int j = 0; j < 4; j++) { out[2 * j] = creal(M[j]); out[2 * j + 1] = cimag(M[j]); }
}
void orc_tmm_probe_csqrt(const double z[2], double out[2])
{
  double complex r = csqrt(z[0] + I * z[1]);
  out[0] = creal(r); out[1] = cimag(r);
}

/* the analytic source that replaces do_tmm on both sides (oracle/ref_hooks_ttm.c) */
static void do_source(orc_ttm *t)
{
  const double Imp = 3.769911184307751e+02;
  const double I_t = t->p.I0 * exp(-pow(t->tmm_time - t->p.laser_t_0, 2) / t->laser_sigma_t_squared);
  const int L = t->L;
  if (sqrt(2.0 * I_t * Imp) < t->p.tmm_laser_threshold * sqrt(2 * t->p.I0 * Imp)) {
    t->laser_active = 0;
    for (int r = 0; r < t->P; r++) for (int i = 1; i <= L; i++) CELL(t->l1, r, i).source = CELL(t->l2, r, i).source = 0.0;
    return;
  }
  t->laser_active = 1;
  int first = 1 << 30;
  for (int r = 0; r < t->P; r++) for (int i = 1; i <= L; i++)
    if (ACTIVE_GE(CELL(t->l1, r, i))) { int g = iglobal_of(t, r, i); if (g < first) first = g; }
  for (int r = 0; r < t->P; r++) for (int i = 1; i <= L; i++) {
    double s = 0.0;
    if (ACTIVE_GE(CELL(t->l1, r, i))) {
      double depth = (double) (iglobal_of(t, r, i) - first) * t->fd_hx;
      s = I_t * t->p.src_abs * t->p.src_mu * exp(-t->p.src_mu * depth);
      s *= 1e10; s *= 6.3538562638e-26;
    }
    CELL(t->l1, r, i).source = CELL(t->l2, r, i).source = s;
  }
}

/* do_DIFF (imd_ttm.c:1473-1678) */
static int do_diff(orc_ttm *t, double tau)
{
  const int L = t->L;
  const double xi_fac = 26.9815 * AMU / 3.0 * 1e30 / ((double) t->diff_substeps);
  const double invxsq = 1.0 / (t->fd_hx * t->fd_hx);
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= L; i++) {
      const int g = iglobal_of(t, r, i);
      if (g > t->last_active_cell_global) continue;
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      if (t->laser_active) t->Eabs += a->source * t->fd_vol * tau;
      if (a->natoms < t->p.fd_min_atoms || a->dens < RHOMIN) continue;
      const tcell *cm = &CELL(t->l1, r, i - 1), *cp = &CELL(t->l1, r, i + 1);
      const tcell *xmin = (cm->natoms < t->p.fd_min_atoms || cm->dens < RHOMIN) ? a : cm;
      const tcell *xmax = (cp->natoms < t->p.fd_min_atoms || cp->dens < RHOMIN) ? a : cp;
      double xmaxTe = xmax->temp, xminTe = xmin->temp, xmaxk = xmax->fd_k, xmink = xmin->fd_k;
      if (t->cur_vlattice_proc == r) { xmaxTe = t->v1[0].temp; xmaxk = t->v1[0].fd_k; }   /* :1548-1552 */
      const double Ce = a->Ce;
      b->temp = tau / Ce * ((((a->fd_k + xmaxk) / 2 * (xmaxTe - a->temp) * invxsq)
                             - ((a->fd_k + xmink) / 2 * (a->temp - xminTe) * invxsq))
                            - a->fd_g * (a->temp - a->md_temp) + a->source) + a->temp;
      b->U = a->U + (b->temp - a->temp) * Ce * t->fd_vol / ((double) a->natoms);
      a->xi += (b->temp - a->md_temp) * xi_fac * a->fd_g / a->md_temp / a->dens;
      b->xi = a->xi;
      t->xiarr_local[r * t->LD + i] += (b->temp - a->md_temp) * xi_fac * a->fd_g / a->md_temp / a->dens;
      if (isnan(b->temp) != 0 || isinf(b->temp) != 0) { snprintf(t->err, sizeof t->err, "Te got NaN or Inf in diffloop (cell %d)", g); return -1; }
    }
  if (t->cur_vlattice_proc >= 0 && t->cur_vlattice_proc < t->P) {                 /* virtual lattice, :1602-1657 */
    const int rv = t->cur_vlattice_proc;
    const int ilocal = t->last_active_cell_global + 1 - rv * L;
    double Ci = 8.589449886e+02;
    Ci *= t->p.vlatdens; Ci *= 11604.5; Ci *= 1e-30; Ci *= J2eV;
    for (int i = 0; i < t->p.vlatdim; i++) {
      double xminTe, xmink, xmaxTe, xmaxk;
      if (i == 0) { xminTe = CELL(t->l1, rv, ilocal).temp; xmink = CELL(t->l1, rv, ilocal).fd_k; }
      else { xminTe = t->v1[i - 1].temp; xmink = t->v1[i - 1].fd_k; }
      if (i < t->p.vlatdim - 1) { xmaxk = t->v1[i + 1].fd_k; xmaxTe = t->v1[i + 1].temp; }
      else { xmaxk = t->v1[i].fd_k; xmaxTe = t->v1[i].temp; }
      const double Ce = t->v1[i].Ce;
      t->v2[i].temp = tau / Ce * (((t->v1[i].fd_k + xmaxk) / 2 * (xmaxTe - t->v1[i].temp) * invxsq)
                                  - ((t->v1[i].fd_k + xmink) / 2 * (t->v1[i].temp - xminTe) * invxsq)
                                  - t->v1[i].fd_g * (t->v1[i].temp - t->v1[i].md_temp)) + t->v1[i].temp;
      double dT = tau / Ci * t->v1[i].fd_g * (t->v1[i].temp - t->v1[i].md_temp);
      t->v2[i].md_temp = t->v1[i].md_temp + dT;
    }
  }
  swap_lattices(t);
  { tcell *x = t->v1; t->v1 = t->v2; t->v2 = x; }
  return 0;
}

/* calc_ttm (imd_ttm.c:75-156) */
int orc_ttm_step(orc_ttm *t, int steps, long n, const double *pos, const double *mom, const double *mass,
                 const int *type, const int *numneighs, int *fdi)
{
  t->steps = steps;
  t->err[0] = 0;
  update_fd(t, n, pos, mom, mass, type, numneighs, fdi);
  do_adv(t, 1.0);
  do_cell_activation(t);
  if (do_fillmesh(t)) return -1;
  fill_ghost_layers(t);
  cfl_maxdt(t);
  { int q = (int) (t->p.timestep / t->max_dt_ttm); t->diff_substeps = t->p.fd_n_timesteps > q ? t->p.fd_n_timesteps : q; }
  const double tau = t->p.timestep / ((double) t->diff_substeps);
  for (int s = 0; s < t->diff_substeps; s++) {
    if (t->use_tmm) { if (do_tmm_ref(t)) return -1; } else do_source(t);
    t->tmm_time += tau * 10.18 * 1.0e-15;
    if (do_diff(t, tau)) return -1;
    if (do_fillmesh(t)) return -1;
    fill_ghost_layers(t);
  }
  for (int r = 0; r < t->P; r++)                                                   /* MPI_Allgather, :124 */
    for (int i = 1; i <= t->L; i++) t->xiarr_global[iglobal_of(t, r, i)] = t->xiarr_local[r * t->LD + i];
  t->tot_elec_energy = 0;                                                          /* :131-148 */
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= t->L; i++) {
      tcell *a = &CELL(t->l1, r, i), *b = &CELL(t->l2, r, i);
      if (ACTIVE_GE(*a)) {
        a->U = eos_ee_from_r_te(&t->ee, a->dens, a->temp * 11604.5) * 26.9815 * AMU * J2eV;
        b->U = a->U;
        t->tot_elec_energy += a->U * ((double) a->natoms);
      }
    }
  return 0;
}

/* ------------------------------------------------------------------ results */
void orc_ttm_get_cells(const orc_ttm *t, double *f)
{
  for (int r = 0; r < t->P; r++)
    for (int i = 1; i <= t->L; i++) {
      const tcell *c = &CELL(t->l1, r, i);
      double *o = f + (size_t) iglobal_of(t, r, i) * ORC_TTM_NFIELD;
      o[0] = c->natoms; o[1] = c->natoms_old; o[2] = c->temp; o[3] = c->md_temp; o[4] = c->xi; o[5] = c->dens;
      o[6] = c->fd_k; o[7] = c->fd_g; o[8] = c->Z; o[9] = c->ne; o[10] = c->Ce; o[11] = c->U; o[12] = c->source;
    }
}
void orc_ttm_get_arrays(const orc_ttm *t, double *xiarr, double *vcom, double *natoms, double *fluxl, double *fluxr)
{
  for (int g = 0; g < t->G; g++) {
    if (xiarr) xiarr[g] = t->xiarr_global[g];
    if (vcom) { vcom[g] = t->vcomx[g]; vcom[t->G + g] = t->vcomy[g]; vcom[2 * t->G + g] = t->vcomz[g]; }
    if (natoms) natoms[g] = t->natomsglobal[g];
    if (fluxl) fluxl[g] = t->fluxl[g];
    if (fluxr) fluxr[g] = t->fluxr[g];
  }
}
void orc_ttm_get_scalars(const orc_ttm *t, double out[16])
{
  memset(out, 0, 16 * sizeof(double));
  out[0] = t->diff_substeps; out[1] = t->max_dt_ttm; out[2] = t->last_active_cell_global; out[3] = t->cur_vlattice_proc;
  out[4] = t->Eabs; out[5] = t->tot_elec_energy; out[6] = t->fd_hx; out[7] = t->fd_vol; out[8] = t->neighvol;
  out[9] = t->laser_active; out[10] = t->tmm_time;
}
void orc_ttm_get_vlattice(const orc_ttm *t, double *v)
{
  for (int i = 0; i < t->p.vlatdim; i++) {
    const tcell *c = &t->v1[i];
    double *o = v + 6 * i;
    o[0] = c->temp; o[1] = c->md_temp; o[2] = c->fd_k; o[3] = c->fd_g; o[4] = c->Ce; o[5] = c->Z;
  }
}

/* move_atoms_ttm (imd_integrate.c:581-811), TTM1D [+PDECAY]; VLATTICE is not defined in that
 * translation unit, so its "i_global >= last_active_cell_global" branch is not compiled. */
double orc_move_atoms_ttm(const orc_ttm *t, long n, double *pos, double *mom, const double *force,
                          const double *mass, const int *type, const int *vtype, const double *restrictions,
                          int pdecay, double xipdecay, double ramp_start, double ramp_end)
{
  double tot_kin_energy = 0.0;
  double a = 1.0 / (ramp_end - ramp_start);
  a *= a;
  const double dt = t->p.timestep;
  for (long k = 0; k < n; k++) {
    int ig = 0;
    double fd_xi;
    double *x = pos + 3 * k, *p = mom + 3 * k;
    const double m = mass[k];
    if (type[k] == 0) { ig = slab_of(t, x[0]); fd_xi = t->xiarr_global[ig]; } else fd_xi = 0.0;
    const double kin1 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    const double *rs = restrictions + 3 * vtype[k];
    double fx = force[3 * k] * rs[0], fy = force[3 * k + 1] * rs[1], fz = force[3 * k + 2] * rs[2];
    if (pdecay && x[0] > ramp_start)
      fx -= (p[0] / m) * xipdecay * a * (x[0] - ramp_start) * (x[0] - ramp_start);
    p[0] += dt * (fx + fd_xi * m * (p[0] / m - t->vcomx[ig]));
    p[1] += dt * (fy + fd_xi * m * (p[1] / m - t->vcomy[ig]));
    p[2] += dt * (fz + fd_xi * m * (p[2] / m - t->vcomz[ig]));
    const double kin2 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    tot_kin_energy += (kin1 + kin2) / (4 * m);
    const double tmp = dt / m;
    x[0] += tmp * p[0]; x[1] += tmp * p[1]; x[2] += tmp * p[2];
  }
  return tot_kin_energy;
}
