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

#define IMDCU_ABI_VERSION 1

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
/* device: CUDA ordinal.  Replaces nothing in the reference (there is no device). */
int  imdcu_create(imdcu_ctx **ctx, int device);
int  imdcu_destroy(imdcu_ctx *ctx);
const char *imdcu_last_error(const imdcu_ctx *ctx);   /* ctx may be NULL: create errors */

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
/* ntypes, timestep, restrictions[3*nvtypes] (globals.h; imd_integrate.c:208-213) */
int  imdcu_set_params(imdcu_ctx *ctx, int ntypes, double timestep,
                      int nvtypes, const double *restrictions);
/* PDECAY rear absorber of the integrator (imd_integrate.c:233-239, 709-716) */
int  imdcu_set_pdecay(imdcu_ctx *ctx, int enabled, double xipdecay,
                      double ramp_start, double ramp_end);

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

/* ------------------------------------------------------------------ inspection (parity) */
/* geometry chosen by imdcu_set_box: global_cell_dim and cellsz */
int  imdcu_get_cell_grid(const imdcu_ctx *ctx, int gdim[3], double *cellsz);
/* neighbour list as (id_i, id_j) pairs, every unordered pair once (i.e. the half list
 * of make_nblist up to order); call with pairs==NULL to get the count in *npairs. */
int  imdcu_get_nblist_pairs(imdcu_ctx *ctx, int64_t *npairs, int32_t *pairs);
int  imdcu_nbl_count(const imdcu_ctx *ctx);            /* globals.h:178 nbl_count */
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

/* single-function probe of the device table access (parity test of potaccess.h):
 * evaluates PAIR_INT2 on `which` for n arguments on the GPU */
int  imdcu_eval_table(imdcu_ctx *ctx, int which, int col, int64_t n, const double *arg,
                      double *val, double *grad);

#ifdef __cplusplus
}
#endif
#endif /* IMDCU_H */
