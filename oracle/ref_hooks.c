/* oracle/ref_hooks.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Link-time observation hook for the UNMODIFIED reference (fmqeisfeld/IMD).
 * The reference has no per-atom dump of cell ids or neighbour lists (SURVEY.md
 * section 4), so the oracle binary is linked with
 *     -Wl,--wrap=calc_forces
 * and this file supplies __wrap_calc_forces(): it records the inputs, calls the
 * reference's own calc_forces (imd_forces_nbl.c:295) and records the outputs,
 * reading the reference's global state (globals.h:99-119,178-188) through the
 * reference's own headers (-I/root/reference at build time).  Nothing from the
 * reference is copied here; no reference arithmetic is changed.
 *
 * Dumps go to $IMD_DUMP_DIR as raw little-endian arrays
 *     s<step>.<name>.bin      (np.fromfile-readable; see tests/golden/make_golden.py)
 * for the steps listed in $IMD_DUMP_STEPS (comma separated).
 */
#include "imd.h"
#include <stdlib.h>
#include <string.h>

void __real_calc_forces(int steps);

static int want_step(int steps)
{
  const char *s = getenv("IMD_DUMP_STEPS");
  if (s == NULL) return 0;
  while (*s) {
    char *e;
    long v = strtol(s, &e, 10);
    if (e == s) break;
    if (v == steps) return 1;
    s = (*e == ',') ? e + 1 : e;
  }
  return 0;
}

static FILE *dump_open(const char *dir, int steps, const char *name)
{
  char fn[1024];
  FILE *f;
#ifdef MPI
  snprintf(fn, sizeof fn, "%s/s%d.%s.r%d.bin", dir, steps, name, myid);   /* one file per rank */
#else
  snprintf(fn, sizeof fn, "%s/s%d.%s.bin", dir, steps, name);
#endif
  f = fopen(fn, "wb");
  if (f == NULL) { perror(fn); exit(3); }
  return f;
}

/* real cell that a (possibly ghost) local cell is an image of */
static int home_cell(int c)
{
  int i = c / (cell_dim.y * cell_dim.z);
  int j = (c / cell_dim.z) % cell_dim.y;
  int k = c % cell_dim.z;
  if (i == 0) i = cell_dim.x - 2; else if (i == cell_dim.x - 1) i = 1;
  if (j == 0) j = cell_dim.y - 2; else if (j == cell_dim.y - 1) j = 1;
  if (k == 0) k = cell_dim.z - 2; else if (k == cell_dim.z - 1) k = 1;
  return i * cell_dim.y * cell_dim.z + j * cell_dim.z + k;
}

static void dump_atoms(const char *dir, int steps, const char *tag)
{
  char nm[64];
  FILE *fid, *fpos, *fmom, *ftyp, *fmas;
  int k, i;
  snprintf(nm, sizeof nm, "%s_id", tag);   fid  = dump_open(dir, steps, nm);
  snprintf(nm, sizeof nm, "%s_pos", tag);  fpos = dump_open(dir, steps, nm);
  snprintf(nm, sizeof nm, "%s_mom", tag);  fmom = dump_open(dir, steps, nm);
  snprintf(nm, sizeof nm, "%s_type", tag); ftyp = dump_open(dir, steps, nm);
  snprintf(nm, sizeof nm, "%s_mass", tag); fmas = dump_open(dir, steps, nm);
  for (k = 0; k < ncells; k++) {
    cell *p = CELLPTR(k);
    for (i = 0; i < p->n; i++) {
      int    id = NUMMER(p, i), ty = SORTE(p, i);
      double x[3], m[3], ma = MASSE(p, i);
      x[0] = ORT(p, i, X);    x[1] = ORT(p, i, Y);    x[2] = ORT(p, i, Z);
      m[0] = IMPULS(p, i, X); m[1] = IMPULS(p, i, Y); m[2] = IMPULS(p, i, Z);
      fwrite(&id, sizeof id, 1, fid);
      fwrite(x, sizeof(double), 3, fpos);
      fwrite(m, sizeof(double), 3, fmom);
      fwrite(&ty, sizeof ty, 1, ftyp);
      fwrite(&ma, sizeof ma, 1, fmas);
    }
  }
  fclose(fid); fclose(fpos); fclose(fmom); fclose(ftyp); fclose(fmas);
}

void __wrap_calc_forces(int steps)
{
  const char *dir = getenv("IMD_DUMP_DIR");
  int dump = (dir != NULL) && want_step(steps);
  int nbl_before = nbl_count;
  int k, i;

  if (dump) dump_atoms(dir, steps, "pre");

  __real_calc_forces(steps);

  if (!dump) return;

  dump_atoms(dir, steps, "post");
  {
    FILE *ff = dump_open(dir, steps, "post_force");
    FILE *fe = dump_open(dir, steps, "post_epot");
    FILE *fr = dump_open(dir, steps, "post_rho");
    FILE *fd = dump_open(dir, steps, "post_dF");
    FILE *fc = dump_open(dir, steps, "post_cell");
#ifdef TTM
    FILE *fn = dump_open(dir, steps, "post_numneighs");
#endif
    for (k = 0; k < ncells; k++) {
      int   c = cnbrs[k].np;
      cell *p = cell_array + c;
      /* global cell coordinates of this (inner) cell, serial or MPI */
      int gc[3];
      gc[0] = c / (cell_dim.y * cell_dim.z) - 1 + my_coord.x * (cell_dim.x - 2);
      gc[1] = (c / cell_dim.z) % cell_dim.y - 1 + my_coord.y * (cell_dim.y - 2);
      gc[2] = c % cell_dim.z               - 1 + my_coord.z * (cell_dim.z - 2);
      for (i = 0; i < p->n; i++) {
        double f[3], e = POTENG(p, i), r = EAM_RHO(p, i), d = EAM_DF(p, i);
        f[0] = KRAFT(p, i, X); f[1] = KRAFT(p, i, Y); f[2] = KRAFT(p, i, Z);
        fwrite(f, sizeof(double), 3, ff);
        fwrite(&e, sizeof e, 1, fe);
        fwrite(&r, sizeof r, 1, fr);
        fwrite(&d, sizeof d, 1, fd);
        fwrite(gc, sizeof(int), 3, fc);
#ifdef TTM
        { int nn = NUMNEIGHS(p, i); fwrite(&nn, sizeof nn, 1, fn); }
#endif
      }
    }
    fclose(ff); fclose(fe); fclose(fr); fclose(fd); fclose(fc);
#ifdef TTM
    fclose(fn);
#endif
  }

#ifdef NRB
  /* per-atom NRB state as init_nrb left it (imd_nrb.c:125-768): boundary type, neighbour flag, the
   * numbers of the 12 lattice sites, reference position; post_mom above is the momentum AFTER
   * nrb_forces (imd_forces_nbl.c:1019-1021) */
  {
    FILE *fb = dump_open(dir, steps, "post_nrbbnd");
    FILE *fg = dump_open(dir, steps, "post_nrbneigh");
    FILE *fi = dump_open(dir, steps, "post_nrbi");
    FILE *fr = dump_open(dir, steps, "post_refpos");
    for (k = 0; k < ncells; k++) {
      cell *p = CELLPTR(k);
      for (i = 0; i < p->n; i++) {
        int b = NRBBND(p, i), g = NRBNEIGH(p, i), id12[12], r;
        double x[3];
        for (r = 0; r < 12; r++) id12[r] = NRBI(p, i, r);
        x[0] = REF_POS(p, i, X); x[1] = REF_POS(p, i, Y); x[2] = REF_POS(p, i, Z);
        fwrite(&b, sizeof b, 1, fb); fwrite(&g, sizeof g, 1, fg);
        fwrite(id12, sizeof(int), 12, fi); fwrite(x, sizeof(double), 3, fr);
      }
    }
    fclose(fb); fclose(fg); fclose(fi); fclose(fr);
  }
#endif

  /* half neighbour list as (id_i, id_j) pairs, only when it was rebuilt now */
  if (nbl_count != nbl_before) {
    FILE *fp = dump_open(dir, steps, "post_nblpairs");
    int n = 0, m;
    for (k = 0; k < ncells; k++) {
      cell *p = cell_array + cnbrs[k].np;
      for (i = 0; i < p->n; i++) {
        for (m = tl[n]; m < tl[n + 1]; m++) {
          int   c  = cl_num[tb[m]];
          int   j  = tb[m] - cl_off[c];
          cell *q  = cell_array + home_cell(c);
          int   pr[2];
          pr[0] = NUMMER(p, i);
          pr[1] = NUMMER(q, j);
          fwrite(pr, sizeof(int), 2, fp);
        }
        n++;
      }
    }
    fclose(fp);
  }

  {
    FILE *fs = dump_open(dir, steps, "post_scalars");
    double s[16];
    memset(s, 0, sizeof s);
    s[0] = tot_pot_energy; s[1] = virial;
    s[2] = (double) (nbl_count != nbl_before);
    s[3] = (double) natoms;
    s[4] = global_cell_dim.x; s[5] = global_cell_dim.y; s[6] = global_cell_dim.z;
    s[7] = cellsz;
    s[8] = box_x.x; s[9] = box_y.y; s[10] = box_z.z;
    s[11] = tot_kin_energy;
    s[12] = (double) last_nbl_len;
    fwrite(s, sizeof(double), 16, fs);
    fclose(fs);
  }
}
