/* oracle/eam_oracle.c -- TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * Scalar CPU restatement of the reference's cell / neighbour-list / EAM2 / NVE path:
 *   make_box, init_cells        imd_geom_3d.c:52-108, 118-330
 *   do_boundaries               imd_main_3d.c:2018-2105
 *   cell_coord                  imd_geom_3d.c:1077-1097
 *   send_cells(copy_cell)       imd_comm_force_3d.c:222-396, 726-760   (serial branches)
 *   make_cell_lists (AR half)   imd_geom_3d.c:846-975
 *   make_nblist                 imd_forces_nbl.c:136-287
 *   calc_forces                 imd_forces_nbl.c:295-1054
 *   send_forces                 imd_comm_force_3d.c:569-715            (serial branches)
 *   move_atoms_nve              imd_integrate.c:35-561
 *   check_nblist                imd_forces_nbl.c:1062-1092
 * Written from scratch; the data structures (CSR over cells, flat arrays) are ours.
 * PINNED: tests/test_oracle_vs_reference.py compares it with dumps of the unmodified
 * reference binary (the .npz fixtures under tests/golden).
 */
#include "oracle.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- tables */

static inline const double *row_ptr(const orc_table *pt, int k, int col, int inc)
{
  return pt->table + (long) k * inc + col;
}

/* potaccess.h:323-354 */
void orc_pair_int2(const orc_table *pt, int col, int inc, double r2,
                   double *pot, double *grad, int *is_short)
{
  double r2a, istep, chi, p0, p1, p2, dv, d2v;
  const double *ptr;
  int k;
  r2a = (r2 < pt->end[col]) ? r2 : pt->end[col];
  r2a = r2a - pt->begin[col];
  if (r2a < 0) { r2a = 0; *is_short = 1; }
  istep = pt->invstep[col];
  r2a = r2a * istep;
  k = (int) r2a;
  chi = r2a - k;
  ptr = row_ptr(pt, k, col, inc);
  p0 = ptr[0]; p1 = ptr[inc]; p2 = ptr[2 * inc];
  dv = p1 - p0;
  d2v = p2 - 2 * p1 + p0;
  *pot = p0 + chi * dv + 0.5 * chi * (chi - 1) * d2v;
  *grad = 2 * istep * (dv + (chi - 0.5) * d2v);
}

/* potaccess.h:465-495 */
void orc_val_func2(const orc_table *pt, int col, int inc, double r2,
                   double *val, int *is_short)
{
  double g;
  orc_pair_int2(pt, col, inc, r2, val, &g, is_short);
}

/* potaccess.h:591-621 */
void orc_deriv_func2(const orc_table *pt, int col, int inc, double r2,
                     double *grad, int *is_short)
{
  double v;
  orc_pair_int2(pt, col, inc, r2, &v, grad, is_short);
}

/* imd_potential.c:1260-1276: continue the last interpolation parabola by two rows */
void orc_init_threepoint(orc_table *pt)
{
  int col, nc = pt->ncols;
  for (col = 0; col < nc; col++) {
    double *y = pt->table + col;
    int n = pt->len[col];
    y[(long) n * nc]       = 3 * y[(long) (n - 1) * nc] - 3 * y[(long) (n - 2) * nc] + y[(long) (n - 3) * nc];
    y[(long) (n + 1) * nc] = 6 * y[(long) (n - 1) * nc] - 8 * y[(long) (n - 2) * nc] + 3 * y[(long) (n - 3) * nc];
  }
}

/* imd_potential.c:1175-1191 */
void orc_init_fourpoint(orc_table *pt)
{
  int col, nc = pt->ncols;
  for (col = 0; col < nc; col++) {
    double *y = pt->table + col;
    long n = pt->len[col];
    y[n * nc]       =  4 * y[(n - 1) * nc] -  6 * y[(n - 2) * nc] +  4 * y[(n - 3) * nc] -     y[(n - 4) * nc];
    y[(n + 1) * nc] = 10 * y[(n - 1) * nc] - 20 * y[(n - 2) * nc] + 15 * y[(n - 3) * nc] - 4 * y[(n - 4) * nc];
  }
}

/* imd_potential.c:1201-1248 */
void orc_init_spline(orc_table *pt, int radial)
{
  int col, nc = pt->ncols;
  long size = pt->maxsteps + 2, i, k;
  double *u = (double *) calloc((size_t) size, sizeof(double));
  for (col = 0; col < nc; col++) {
    double *y2 = pt->table2 + col, *y = pt->table + col;
    long n = pt->len[col];
    double step = pt->step[col], p, qn, un;
    y2[0] = u[0] = 0;
    for (i = 1; i < n - 1; i++) {
      p = 0.5 * y2[(i - 1) * nc] + 2.0;
      y2[i * nc] = -0.5 / p;
      u[i] = (y[(i + 1) * nc] - 2 * y[i * nc] + y[(i - 1) * nc]) / step;
      u[i] = (6.0 * u[i] / (2 * step) - 0.5 * u[i - 1]) / p;
    }
    if (radial) { qn = 0.5; un = (3.0 / step) * (y[(n - 2) * nc] - y[(n - 1) * nc]) / step; }
    else qn = un = 0.0;
    y2[(n - 1) * nc] = (un - qn * u[n - 2]) / (qn * y2[(n - 2) * nc] + 1.0);
    for (k = n - 2; k >= 0; k--) y2[k * nc] = y2[k * nc] * y2[(k + 1) * nc] + u[k];
    y[n * nc]  = 2 * y[(n - 1) * nc] - y[(n - 2) * nc] + (step * step) * y2[(n - 1) * nc];
    y2[n * nc] = 2 * y2[(n - 1) * nc] - y2[(n - 2) * nc];
  }
  free(u);
}

/* potaccess.h:365-413 (value and twice the derivative, cubic through rows k-1..k+2) */
static void pair_int3(const orc_table *pt, int col, int inc, double r2,
                      double *pot, double *grad, int *is_short)
{
  double r2a, istep, chi, p0, p1, p2, p3;
  double fac0, fac1, fac2, fac3, dfac0, dfac1, dfac2, dfac3;
  const double *ptr;
  int k;
  r2a = (r2 < pt->end[col]) ? r2 : pt->end[col];
  r2a = r2a - pt->begin[col];
  if (r2a < 0) { r2a = 0; *is_short = 1; }
  istep = pt->invstep[col];
  r2a = r2a * istep;
  k = (int) r2a; if (k < 1) k = 1;
  chi = r2a - k;
  fac0 = -(1.0 / 6.0) * chi * (chi - 1.0) * (chi - 2.0);
  fac1 =        0.5 * (chi * chi - 1.0) * (chi - 2.0);
  fac2 =       -0.5 * chi * (chi + 1.0) * (chi - 2.0);
  fac3 =  (1.0 / 6.0) * chi * (chi * chi - 1.0);
  dfac0 = -(1.0 / 6.0) * ((3.0 * chi - 6.0) * chi + 2.0);
  dfac1 =        0.5 * ((3.0 * chi - 4.0) * chi - 1.0);
  dfac2 =       -0.5 * ((3.0 * chi - 2.0) * chi - 2.0);
  dfac3 =    1.0 / 6.0 * (3.0 * chi * chi - 1.0);
  ptr = row_ptr(pt, k - 1, col, inc);
  p0 = ptr[0]; p1 = ptr[inc]; p2 = ptr[2 * inc]; p3 = ptr[3 * inc];
  *pot = fac0 * p0 + fac1 * p1 + fac2 * p2 + fac3 * p3;
  *grad = 2 * istep * (dfac0 * p0 + dfac1 * p1 + dfac2 * p2 + dfac3 * p3);
}

/* potaccess.h:421-457 */
static void pair_int_sp(const orc_table *pt, int col, int inc, double r2,
                        double *pot, double *grad, int *is_short)
{
  double r2a, a, b, a2, b2, istep, step, st6, p1, p2, d21, d22;
  long k;
  r2a = (r2 < pt->end[col]) ? r2 : pt->end[col];
  r2a = r2a - pt->begin[col];
  if (r2a < 0) { r2a = 0; *is_short = 1; }
  istep = pt->invstep[col];
  step = pt->step[col];
  r2a = r2a * istep;
  k = (int) r2a;
  b = r2a - k;
  a = 1.0 - b;
  k = k * inc + col;
  p1 = pt->table[k]; d21 = pt->table2[k];
  k += inc;
  p2 = pt->table[k]; d22 = pt->table2[k];
  a2 = a * a - 1;
  b2 = b * b - 1;
  st6 = step / 6;
  *pot = a * p1 + b * p2 + (a * a2 * d21 + b * b2 * d22) * st6 * step;
  *grad = 2 * ((p2 - p1) * istep + ((3 * b2 + 2) * d22 - (3 * a2 + 2) * d21) * st6);
}

static int g_interp = 0;
void orc_set_interp(int mode) { g_interp = mode; }

void orc_pair_int(const orc_table *pt, int col, int inc, double r2, double *pot, double *grad, int *is_short)
{
  if (g_interp == 1) pair_int3(pt, col, inc, r2, pot, grad, is_short);
  else if (g_interp == 2) pair_int_sp(pt, col, inc, r2, pot, grad, is_short);
  else orc_pair_int2(pt, col, inc, r2, pot, grad, is_short);
}
/* VAL_FUNC* and DERIV_FUNC* evaluate the same expressions as the PAIR_INT* of their mode */
void orc_val_func(const orc_table *pt, int col, int inc, double r2, double *val, int *is_short)
{
  double g;
  orc_pair_int(pt, col, inc, r2, val, &g, is_short);
}
void orc_deriv_func(const orc_table *pt, int col, int inc, double r2, double *grad, int *is_short)
{
  double v;
  orc_pair_int(pt, col, inc, r2, &v, grad, is_short);
}

/* ---------------------------------------------------------------- state */

orc_state *orc_create(void)
{
  return (orc_state *) calloc(1, sizeof(orc_state));
}

static void free_lists(orc_state *s)
{
  free(s->cl_off); free(s->cell_atom); free(s->flat_pos); free(s->atom_cell);
  free(s->tl); free(s->tb); free(s->row_atom);
  s->cl_off = s->cell_atom = s->atom_cell = s->tl = s->tb = s->row_atom = NULL;
  s->flat_pos = NULL;
}

void orc_destroy(orc_state *s)
{
  if (!s) return;
  free_lists(s);
  free(s);
}

static void cross(const double *u, const double *v, double *w)
{
  w[0] = u[1] * v[2] - u[2] * v[1];
  w[1] = u[2] * v[0] - u[0] * v[2];
  w[2] = u[0] * v[1] - u[1] * v[0];
}

static double dot3(const double *a, const double *b)
{
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

int orc_setup(orc_state *s, const double box[9], const int pbc[3],
              double r2_cut, double nbl_margin, const int divisor[3])
{
  double volume, scale;
  int d, k;
  memcpy(s->box, box, sizeof s->box);
  memcpy(s->pbc, pbc, sizeof s->pbc);
  /* make_box: tbox_j with box_i . tbox_j = delta_ij */
  cross(box + 3, box + 6, s->tbox + 0);
  cross(box + 6, box + 0, s->tbox + 3);
  cross(box + 0, box + 3, s->tbox + 6);
  volume = dot3(box, s->tbox);
  if (volume == 0) return -1;
  for (k = 0; k < 9; k++) s->tbox[k] /= volume;
  for (d = 0; d < 3; d++) s->height[d] = 1.0 / dot3(s->tbox + 3 * d, s->tbox + 3 * d);
  /* init_cells: margin added once, then integer number of cells per direction */
  s->cellsz = (sqrt(r2_cut) + nbl_margin) * (sqrt(r2_cut) + nbl_margin);
  for (d = 0; d < 3; d++) {
    int cd = divisor ? divisor[d] : 1;
    scale = sqrt(1.0 * s->cellsz / s->height[d]);
    s->gdim[d] = (int) (1.0 / scale);
    if (0 != s->gdim[d] % cd) s->gdim[d] = ((int) (s->gdim[d] / cd)) * cd;
    if (s->gdim[d] < cd) return -2;
    s->cdim[d] = s->gdim[d] + 2;
  }
  return 0;
}

static inline int cidx(const orc_state *s, int i, int j, int k)
{
  return (i * s->cdim[1] + j) * s->cdim[2] + k;
}

/* the forward ghost copies in the order send_cells performs them (serial branches);
   returns the number of (from,to,shiftvec) triples, fills arrays when non-NULL */
static int ghost_plan(const orc_state *s, int *from, int *to, double *shift)
{
  int cx = s->cdim[0], cy = s->cdim[1], cz = s->cdim[2];
  int i, j, n = 0, d;
  const double *bx = s->box, *by = s->box + 3, *bz = s->box + 6;
#define ADD(f, t, sg, vec)                                                   \
  do { if (from) { from[n] = (f); to[n] = (t);                               \
         for (d = 0; d < 3; d++) shift[3 * n + d] = (sg) * (vec)[d]; } n++; } while (0)
  /* up/down */
  if (s->pbc[2])
    for (i = 1; i < cx - 1; ++i)
      for (j = 1; j < cy - 1; ++j) {
        ADD(cidx(s, i, j, 1), cidx(s, i, j, cz - 1), 1.0, bz);
        ADD(cidx(s, i, j, cz - 2), cidx(s, i, j, 0), -1.0, bz);
      }
  /* north/south */
  if (s->pbc[1])
    for (i = 1; i < cx - 1; ++i)
      for (j = 0; j < cz; ++j) {
        ADD(cidx(s, i, 1, j), cidx(s, i, cy - 1, j), 1.0, by);
        ADD(cidx(s, i, cy - 2, j), cidx(s, i, 0, j), -1.0, by);
      }
  /* east only (AR) */
  if (s->pbc[0])
    for (i = 0; i < cy; ++i)
      for (j = 0; j < cz; ++j)
        ADD(cidx(s, 1, i, j), cidx(s, cx - 1, i, j), 1.0, bx);
#undef ADD
  return n;
}

void orc_refresh_ghosts(orc_state *s, const double *pos)
{
  int nall = s->cdim[0] * s->cdim[1] * s->cdim[2];
  int np = ghost_plan(s, NULL, NULL, NULL);
  int *from = (int *) malloc(sizeof(int) * (np + 1)), *to = (int *) malloc(sizeof(int) * (np + 1));
  double *sh = (double *) malloc(sizeof(double) * 3 * (np + 1));
  int c, q, a, d;
  (void) nall;
  ghost_plan(s, from, to, sh);
  /* owned atoms */
  for (c = 0; c < nall; c++) {
    int i = c / (s->cdim[1] * s->cdim[2]), j = (c / s->cdim[2]) % s->cdim[1], k = c % s->cdim[2];
    if (i == 0 || j == 0 || k == 0 || i == s->cdim[0] - 1 || j == s->cdim[1] - 1 || k == s->cdim[2] - 1)
      continue;
    for (a = s->cl_off[c]; a < s->cl_off[c + 1]; a++)
      for (d = 0; d < 3; d++) s->flat_pos[3 * a + d] = pos[3 * (long) s->cell_atom[a] + d];
  }
  for (q = 0; q < np; q++) {
    int f0 = s->cl_off[from[q]], t0 = s->cl_off[to[q]], cnt = s->cl_off[from[q] + 1] - f0;
    for (a = 0; a < cnt; a++)
      for (d = 0; d < 3; d++)
        s->flat_pos[3 * (t0 + a) + d] = s->flat_pos[3 * (f0 + a) + d] + sh[3 * q + d];
  }
  free(from); free(to); free(sh);
}

int orc_fix_cells(orc_state *s, long n, double *pos)
{
  int nall = s->cdim[0] * s->cdim[1] * s->cdim[2];
  int *cnt, *fill;
  int np, *from, *to, q, c, d;
  long a;
  free_lists(s);
  s->n = n;
  s->atom_cell = (int *) malloc(sizeof(int) * 3 * (n + 1));
  cnt = (int *) calloc(nall + 1, sizeof(int));
  /* do_boundaries: x, then y, then z, each with the already updated position */
  for (a = 0; a < n; a++) {
    double *r = pos + 3 * a;
    for (d = 0; d < 3; d++) {
      if (s->pbc[d] == 1) {
        const double *t = s->tbox + 3 * d, *b = s->box + 3 * d;
        double i = -floor(r[0] * t[0] + r[1] * t[1] + r[2] * t[2]);
        r[0] += i * b[0]; r[1] += i * b[1]; r[2] += i * b[2];
      }
    }
    /* cell_coord */
    for (d = 0; d < 3; d++) {
      const double *t = s->tbox + 3 * d;
      int cc = (int) (s->gdim[d] * (r[0] * t[0] + r[1] * t[1] + r[2] * t[2]));
      if (cc >= s->gdim[d]) cc = s->gdim[d] - 1; else if (cc < 0) cc = 0;
      s->atom_cell[3 * a + d] = cc;
    }
    cnt[cidx(s, s->atom_cell[3 * a] + 1, s->atom_cell[3 * a + 1] + 1, s->atom_cell[3 * a + 2] + 1)]++;
  }
  /* ghost cell occupancies follow their source cells, in send_cells order */
  np = ghost_plan(s, NULL, NULL, NULL);
  from = (int *) malloc(sizeof(int) * (np + 1)); to = (int *) malloc(sizeof(int) * (np + 1));
  { double *sh = (double *) malloc(sizeof(double) * 3 * (np + 1)); ghost_plan(s, from, to, sh); free(sh); }
  for (q = 0; q < np; q++) cnt[to[q]] = cnt[from[q]];
  /* cl_off = exclusive prefix over ALL cells (imd_forces_nbl.c:165-170) */
  s->cl_off = (int *) malloc(sizeof(int) * (nall + 1));
  s->cl_off[0] = 0;
  for (c = 0; c < nall; c++) s->cl_off[c + 1] = s->cl_off[c] + cnt[c];
  s->n_flat = s->cl_off[nall];
  s->cell_atom = (int *) malloc(sizeof(int) * (s->n_flat + 1));
  s->flat_pos = (double *) malloc(sizeof(double) * 3 * (s->n_flat + 1));
  fill = (int *) calloc(nall + 1, sizeof(int));
  for (a = 0; a < n; a++) {
    c = cidx(s, s->atom_cell[3 * a] + 1, s->atom_cell[3 * a + 1] + 1, s->atom_cell[3 * a + 2] + 1);
    s->cell_atom[s->cl_off[c] + fill[c]++] = (int) a;
  }
  for (q = 0; q < np; q++) {
    int f0 = s->cl_off[from[q]], t0 = s->cl_off[to[q]], k, m = cnt[from[q]];
    for (k = 0; k < m; k++) s->cell_atom[t0 + k] = s->cell_atom[f0 + k];
  }
  free(cnt); free(fill); free(from); free(to);
  orc_refresh_ghosts(s, pos);
  return 0;
}

/* ---------------------------------------------------------------- neighbour list */

/* half stencil of make_cell_lists (imd_geom_3d.c:917-975), -1 = excluded */
static int half_stencil(const orc_state *s, int i, int j, int k, int nq[14])
{
  int l, m, n, nn = 1;
  int np = cidx(s, i, j, k);
  nq[0] = np;
  for (l = 0; l <= 1; ++l)
    for (m = -l; m <= 1; ++m)
      for (n = (l == 0 ? -m : -l); n <= 1; ++n) {
        int r = i + l, q = j + m, t = k + n, qq, ip[3], ok = 1, d;
        qq = cidx(s, r, q, t);
        if (qq == np) continue;
        r -= 1; q -= 1; t -= 1;               /* global coords (serial: my_coord = 0) */
        ip[0] = (r < 0) ? -1 : (r > s->gdim[0] - 1);
        ip[1] = (q < 0) ? -1 : (q > s->gdim[1] - 1);
        ip[2] = (t < 0) ? -1 : (t > s->gdim[2] - 1);
        for (d = 0; d < 3; d++) if (!(s->pbc[d] == 1 || ip[d] == 0)) ok = 0;
        nq[nn++] = ok ? qq : -1;
      }
  return nn;
}

int orc_make_nblist(orc_state *s)
{
  long cap = 1024, tn = 0, row = 0;
  int i, j, k;
  free(s->tl); free(s->tb); free(s->row_atom);
  s->tl = (int *) malloc(sizeof(int) * (s->n + 2));
  s->row_atom = (int *) malloc(sizeof(int) * (s->n + 1));
  s->tb = (int *) malloc(sizeof(int) * cap);
  s->tl[0] = 0;
  for (i = 1; i < s->cdim[0] - 1; ++i)
    for (j = 1; j < s->cdim[1] - 1; ++j)
      for (k = 1; k < s->cdim[2] - 1; ++k) {
        int nq[14], nn, c1 = cidx(s, i, j, k), a, m;
        nn = half_stencil(s, i, j, k, nq);
        for (a = s->cl_off[c1]; a < s->cl_off[c1 + 1]; a++) {
          const double *d1 = s->flat_pos + 3 * (long) a;
          for (m = 0; m < nn; m++) {
            int c2 = nq[m], b, bstart;
            if (c2 < 0) continue;
            bstart = (c2 == c1) ? a + 1 : s->cl_off[c2];
            for (b = bstart; b < s->cl_off[c2 + 1]; b++) {
              const double *q = s->flat_pos + 3 * (long) b;
              double dx = q[0] - d1[0], dy = q[1] - d1[1], dz = q[2] - d1[2];
              double r2 = dx * dx + dy * dy + dz * dz;
              if (r2 < s->cellsz) {
                if (tn >= cap) { cap *= 2; s->tb = (int *) realloc(s->tb, sizeof(int) * cap); }
                s->tb[tn++] = b;
              }
            }
          }
          s->row_atom[row] = s->cell_atom[a];
          s->tl[++row] = (int) tn;
        }
      }
  s->tn = tn;
  return 0;
}

long orc_export_pairs(const orc_state *s, int *pairs)
{
  long r, m;
  if (pairs)
    for (r = 0; r < s->n; r++)
      for (m = s->tl[r]; m < s->tl[r + 1]; m++) {
        pairs[2 * m] = s->row_atom[r];
        pairs[2 * m + 1] = s->cell_atom[s->tb[m]];
      }
  return s->tn;
}

/* ---------------------------------------------------------------- forces */

/* reverse ghost accumulation in send_forces order: x, then y, then z */
static void ghost_add_back(const orc_state *s, double *v, int ncomp)
{
  int np = ghost_plan(s, NULL, NULL, NULL);
  int *from = (int *) malloc(sizeof(int) * (np + 1)), *to = (int *) malloc(sizeof(int) * (np + 1));
  double *sh = (double *) malloc(sizeof(double) * 3 * (np + 1));
  int q, a, d;
  ghost_plan(s, from, to, sh);
  for (q = np - 1; q >= 0; q--) {
    int f0 = s->cl_off[from[q]], t0 = s->cl_off[to[q]], cnt = s->cl_off[from[q] + 1] - f0;
    for (a = 0; a < cnt; a++)
      for (d = 0; d < ncomp; d++) v[ncomp * (f0 + a) + d] += v[ncomp * (t0 + a) + d];
  }
  free(from); free(to); free(sh);
}

static void ghost_copy_fwd(const orc_state *s, double *v)
{
  int np = ghost_plan(s, NULL, NULL, NULL);
  int *from = (int *) malloc(sizeof(int) * (np + 1)), *to = (int *) malloc(sizeof(int) * (np + 1));
  double *sh = (double *) malloc(sizeof(double) * 3 * (np + 1));
  int q, a;
  ghost_plan(s, from, to, sh);
  for (q = 0; q < np; q++) {
    int f0 = s->cl_off[from[q]], t0 = s->cl_off[to[q]], cnt = s->cl_off[from[q] + 1] - f0;
    for (a = 0; a < cnt; a++) v[t0 + a] = v[f0 + a];
  }
  free(from); free(to); free(sh);
}

int orc_calc_forces(orc_state *s, const double *pos, const int *type, int ntypes,
                    const orc_table *pair, const orc_table *rho, const orc_table *embed,
                    double *force, double *epot, double *rho_out, double *dF,
                    int *numneighs, double out8[8])
{
  long nf = s->n_flat, r, m;
  double *kraft = (double *) calloc(3 * nf + 3, sizeof(double));
  double *poteng = (double *) calloc(nf + 1, sizeof(double));
  double *erho = (double *) calloc(nf + 1, sizeof(double));
  double *edF = (double *) calloc(nf + 1, sizeof(double));
  double *nnb = (double *) calloc(nf + 1, sizeof(double));
  double tot_pot_energy = 0.0, virial = 0.0;
  int is_short = 0, idummy = 0, inc = ntypes * ntypes;
  int i, j, k;

  orc_refresh_ghosts(s, pos);                      /* send_cells(copy_cell) :311 */

  /* pass 1: pair + host density (:419-672) */
  r = 0;
  for (i = 1; i < s->cdim[0] - 1; ++i)
    for (j = 1; j < s->cdim[1] - 1; ++j)
      for (k = 1; k < s->cdim[2] - 1; ++k) {
        int c1 = cidx(s, i, j, k), a;
        for (a = s->cl_off[c1]; a < s->cl_off[c1 + 1]; a++, r++) {
          const double *d1 = s->flat_pos + 3 * (long) a;
          double ff[3] = {0, 0, 0}, ee = 0.0, eam_r = 0.0;
          int it = type[s->cell_atom[a]], nbttm = 0;
          for (m = s->tl[r]; m < s->tl[r + 1]; m++) {
            int b = s->tb[m];
            const double *q = s->flat_pos + 3 * (long) b;
            double d[3], r2, pot, grad, rho_h = 0.0, f[3];
            int jt = type[s->cell_atom[b]], col = it * ntypes + jt, col2 = jt * ntypes + it;
            d[0] = q[0] - d1[0]; d[1] = q[1] - d1[1]; d[2] = q[2] - d1[2];
            r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            if (r2 <= pair->end[0] && it == 0 && jt == 0) { nbttm++; nnb[b] += 1; }   /* :473-481 */
            if (r2 <= pair->end[col]) {
              orc_pair_int(pair, col, inc, r2, &pot, &grad, &is_short);
              tot_pot_energy += pot;
              f[0] = d[0] * grad; f[1] = d[1] * grad; f[2] = d[2] * grad;
              kraft[3 * b] -= f[0]; kraft[3 * b + 1] -= f[1]; kraft[3 * b + 2] -= f[2];
              ff[0] += f[0]; ff[1] += f[1]; ff[2] += f[2];
              pot *= 0.5;
              ee += pot;
              poteng[b] += pot;
              virial -= r2 * grad;
            }
            if (r2 < rho->end[col]) {
              orc_val_func(rho, col, inc, r2, &rho_h, &is_short);
              eam_r += rho_h;
            }
            if (it == jt) {
              if (r2 < rho->end[col]) erho[b] += rho_h;
            } else if (r2 < rho->end[col2]) {
              orc_val_func(rho, col2, inc, r2, &rho_h, &is_short);
              erho[b] += rho_h;
            }
          }
          kraft[3 * a] += ff[0]; kraft[3 * a + 1] += ff[1]; kraft[3 * a + 2] += ff[2];
          poteng[a] += ee;
          erho[a] += eam_r;
          nnb[a] += nbttm;
        }
      }

  ghost_add_back(s, erho, 1);                      /* send_forces(add_rho) :745 */

  /* embedding energy and derivative (:748-783) */
  for (i = 1; i < s->cdim[0] - 1; ++i)
    for (j = 1; j < s->cdim[1] - 1; ++j)
      for (k = 1; k < s->cdim[2] - 1; ++k) {
        int c1 = cidx(s, i, j, k), a;
        for (a = s->cl_off[c1]; a < s->cl_off[c1 + 1]; a++) {
          double pot;
          orc_pair_int(embed, type[s->cell_atom[a]], ntypes, erho[a], &pot, &edF[a], &idummy);
          poteng[a] += pot;
          tot_pot_energy += pot;
        }
      }

  ghost_copy_fwd(s, edF);                          /* send_cells(copy_dF) :786 */

  /* pass 2: embedding forces (:790-985) */
  r = 0;
  for (i = 1; i < s->cdim[0] - 1; ++i)
    for (j = 1; j < s->cdim[1] - 1; ++j)
      for (k = 1; k < s->cdim[2] - 1; ++k) {
        int c1 = cidx(s, i, j, k), a;
        for (a = s->cl_off[c1]; a < s->cl_off[c1 + 1]; a++, r++) {
          const double *d1 = s->flat_pos + 3 * (long) a;
          double ff[3] = {0, 0, 0};
          int it = type[s->cell_atom[a]];
          for (m = s->tl[r]; m < s->tl[r + 1]; m++) {
            int b = s->tb[m];
            const double *q = s->flat_pos + 3 * (long) b;
            double d[3], r2, f[3];
            int jt = type[s->cell_atom[b]], col1 = jt * ntypes + it, col2 = it * ntypes + jt;
            d[0] = q[0] - d1[0]; d[1] = q[1] - d1[1]; d[2] = q[2] - d1[2];
            r2 = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
            if ((r2 < rho->end[col1]) || (r2 < rho->end[col2])) {
              double grad, rho_i_strich, rho_j_strich;
              orc_deriv_func(rho, col1, inc, r2, &rho_i_strich, &is_short);
              if (col1 == col2) rho_j_strich = rho_i_strich;
              else orc_deriv_func(rho, col2, inc, r2, &rho_j_strich, &is_short);
              grad = 0.5 * (edF[a] * rho_j_strich + edF[b] * rho_i_strich);
              f[0] = d[0] * grad; f[1] = d[1] * grad; f[2] = d[2] * grad;
              kraft[3 * b] -= f[0]; kraft[3 * b + 1] -= f[1]; kraft[3 * b + 2] -= f[2];
              ff[0] += f[0]; ff[1] += f[1]; ff[2] += f[2];
              virial -= d[0] * f[0] + d[1] * f[1] + d[2] * f[2];
            }
          }
          kraft[3 * a] += ff[0]; kraft[3 * a + 1] += ff[1]; kraft[3 * a + 2] += ff[2];
        }
      }

  /* send_forces(add_forces) :1052 -- kraft, pot_eng, numneighs of ghosts onto owners */
  ghost_add_back(s, kraft, 3);
  ghost_add_back(s, poteng, 1);
  ghost_add_back(s, nnb, 1);

  for (i = 1; i < s->cdim[0] - 1; ++i)
    for (j = 1; j < s->cdim[1] - 1; ++j)
      for (k = 1; k < s->cdim[2] - 1; ++k) {
        int c1 = cidx(s, i, j, k), a;
        for (a = s->cl_off[c1]; a < s->cl_off[c1 + 1]; a++) {
          long o = s->cell_atom[a];
          force[3 * o] = kraft[3 * a]; force[3 * o + 1] = kraft[3 * a + 1]; force[3 * o + 2] = kraft[3 * a + 2];
          epot[o] = poteng[a];
          rho_out[o] = erho[a];
          dF[o] = edF[a];
          if (numneighs) numneighs[o] = (int) nnb[a];
        }
      }
  memset(out8, 0, 8 * sizeof(double));
  out8[0] = tot_pot_energy;
  out8[1] = virial;
  out8[2] = (double) is_short;
  free(kraft); free(poteng); free(erho); free(edF); free(nnb);
  return 0;
}

/* ---------------------------------------------------------------- integrator */

double orc_move_atoms_nve(long n, double *pos, double *mom, const double *force_in,
                          const double *mass, const int *vtype,
                          const double *restrictions, double timestep,
                          int pdecay, double xipdecay, double ramp_start, double ramp_end)
{
  double tot_kin_energy = 0.0, a = 0.0;
  long i;
  if (pdecay) { a = 1.0 / (ramp_end - ramp_start); a *= a; }
  for (i = 0; i < n; i++) {
    double *p = mom + 3 * i, *x = pos + 3 * i, f[3], k1, k2, tmp, m = mass[i];
    const double *rs = restrictions + 3 * (vtype ? vtype[i] : 0);
    k1 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    f[0] = force_in[3 * i] * rs[0]; f[1] = force_in[3 * i + 1] * rs[1]; f[2] = force_in[3 * i + 2] * rs[2];
    if (pdecay && x[0] > ramp_start)
      f[0] -= (p[0] / m) * xipdecay * a * (x[0] - ramp_start) * (x[0] - ramp_start);
    p[0] += timestep * f[0]; p[1] += timestep * f[1]; p[2] += timestep * f[2];
    k2 = p[0] * p[0] + p[1] * p[1] + p[2] * p[2];
    tot_kin_energy += (k1 + k2) / (4 * m);
    tmp = timestep / m;
    x[0] += tmp * p[0]; x[1] += tmp * p[1]; x[2] += tmp * p[2];
  }
  return tot_kin_energy;
}

double orc_max_displacement2(long n, const double *pos, const double *ref)
{
  double max1 = 0.0;
  long i;
  for (i = 0; i < n; i++) {
    double dx = pos[3 * i] - ref[3 * i], dy = pos[3 * i + 1] - ref[3 * i + 1], dz = pos[3 * i + 2] - ref[3 * i + 2];
    double r2 = dx * dx + dy * dy + dz * dz;
    if (r2 > max1) max1 = r2;
  }
  return max1;
}
