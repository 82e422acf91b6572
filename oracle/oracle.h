/* oracle/oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C, scalar FP64, IEEE, no FMA contraction) of the algorithm of
 * the reference's hot path, written from scratch; each function cites the reference
 * file:line it follows.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  The product library
 * (imd_b200/libimdcu.so) never links or calls it.
 *
 * Parity status: EAM/nbl/NVE part is PINNED against the unmodified reference binary
 * built from /root/reference (oracle/_ref, see tests/golden/make_golden.py and
 * tests/test_oracle_vs_reference.py).  TTM part: see ttm_oracle.c header.
 */
#ifndef IMD_ORACLE_H
#define IMD_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* tabulated function set, same meaning as pot_table_t (types.h:537-549) */
typedef struct {
  int     ncols;
  int     maxsteps;
  double *begin, *end, *step, *invstep;
  int    *len;
  double *table;            /* (maxsteps+2) rows x ncols, row-major, padded rows filled */
  double *table2;           /* spline mode: second derivatives (init_spline), else NULL */
} orc_table;

typedef struct {
  /* box (imd_geom_3d.c:52-108) */
  double box[9];            /* rows: box_x, box_y, box_z */
  double tbox[9];
  double height[3];
  int    pbc[3];
  /* cells (imd_geom_3d.c:118-330) */
  double cellsz;            /* (r_cut + nbl_margin)^2 */
  int    gdim[3];           /* global_cell_dim */
  int    cdim[3];           /* cell_dim = gdim + 2 (serial) */
  /* atoms, in caller order */
  long   n;
  /* per-cell lists built by orc_fix_cells: CSR over ALL cells incl. ghosts */
  int   *cl_off;            /* [nallcells+1] */
  int   *cell_atom;         /* [n_flat] -> owner atom index (for ghosts: image source) */
  double *flat_pos;         /* [3*n_flat] (ghosts: shifted copies) */
  int    n_flat;
  int   *atom_cell;         /* [3*n] global cell coords of each owned atom */
  /* half neighbour list (imd_forces_nbl.c:136-287) */
  int   *tl;                /* [n+1], row r = r-th owned atom in cell traversal order */
  int   *tb;                /* flat ids */
  int   *row_atom;          /* [n] owner atom index of row r */
  long   tn;
} orc_state;

/* potaccess.h:323-354 / 465-495 / 591-621 (3-point, quadratic in r^2) */
void orc_pair_int2(const orc_table *pt, int col, int inc, double r2,
                   double *pot, double *grad, int *is_short);
void orc_val_func2(const orc_table *pt, int col, int inc, double r2,
                   double *val, int *is_short);
void orc_deriv_func2(const orc_table *pt, int col, int inc, double r2,
                     double *grad, int *is_short);
/* imd_potential.c:1260-1276 */
void orc_init_threepoint(orc_table *pt);

/* The reference picks its table interpolation at compile time (potaccess.h:24-36, make-target
   substrings 4point / spline); here it is a mode of the library: 0 three-point (default),
   1 four-point (PAIR_INT3 / VAL_FUNC3 / DERIV_FUNC3, potaccess.h:365-413,503-540,630-667),
   2 spline (PAIR_INT_SP / VAL_FUNC_SP / DERIV_FUNC_SP, :421-457,547-583,675-711). */
void orc_set_interp(int mode);
void orc_pair_int(const orc_table *pt, int col, int inc, double r2,
                  double *pot, double *grad, int *is_short);
void orc_val_func(const orc_table *pt, int col, int inc, double r2, double *val, int *is_short);
void orc_deriv_func(const orc_table *pt, int col, int inc, double r2, double *grad, int *is_short);
/* imd_potential.c:1175-1191 and :1201-1248 (table2 must hold (maxsteps+2)*ncols doubles) */
void orc_init_fourpoint(orc_table *pt);
void orc_init_spline(orc_table *pt, int radial);

orc_state *orc_create(void);
void orc_destroy(orc_state *s);

/* make_box + init_cells; cellsz_in = max r2_end of the radial tables, divisor = cd */
int  orc_setup(orc_state *s, const double box[9], const int pbc[3],
               double r2_cut, double nbl_margin, const int divisor[3]);
/* do_boundaries + fix_cells + send_cells(copy_cell): wraps pos IN PLACE, bins atoms */
int  orc_fix_cells(orc_state *s, long n, double *pos);
/* send_cells only: refresh ghost copies from current pos (no rebinning) */
void orc_refresh_ghosts(orc_state *s, const double *pos);
/* make_nblist */
int  orc_make_nblist(orc_state *s);
/* calc_forces passes over the existing list; out arrays in caller atom order
   out8 = {tot_pot_energy, virial, 0...} ; numneighs may be NULL */
int  orc_calc_forces(orc_state *s, const double *pos, const int *type, int ntypes,
                     const orc_table *pair, const orc_table *rho, const orc_table *embed,
                     double *force, double *epot, double *rho_out, double *dF,
                     int *numneighs, double out8[8]);
/* move_atoms_nve (imd_integrate.c:35-561, plain NVE branch [+PDECAY]) */
double orc_move_atoms_nve(long n, double *pos, double *mom, const double *force,
                          const double *mass, const int *vtype,
                          const double *restrictions, double timestep,
                          int pdecay, double xipdecay, double ramp_start, double ramp_end);
/* check_nblist (imd_forces_nbl.c:1062-1092): returns max |r - r_ref|^2 */
double orc_max_displacement2(long n, const double *pos, const double *ref);

/* export the half list as owner-atom index pairs (i, j) */
long orc_export_pairs(const orc_state *s, int *pairs /* [2*tn] or NULL */);

#ifdef __cplusplus
}
#endif
#endif
