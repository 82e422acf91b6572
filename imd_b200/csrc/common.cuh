// imd_b200/csrc/common.cuh -- shared declarations of libimdcu.so (sm_100a only).
//
// Device-resident state of one rank: atoms in cell-sorted SoA/packed arrays, an extended
// cell grid (one ghost layer, like the reference's cell_dim = local + 2,
// imd_geom_3d.c:262-274) whose ghost cells hold shifted image atoms, a FULL neighbour
// list in a tile-padded ("sliced ELL", 32 atoms per slice) layout, and the potential
// tables re-packed as per-interval coefficient rows.
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <chrono>

#include "../../include/imdcu.h"

#define IMDCU_TILE 32            // atoms per neighbour-list slice (= warp size)

// ----------------------------------------------------------------------------- errors
#define CU_CHECK(ctx, expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      (ctx)->set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " at " +     \
                       __FILE__ + ":" + std::to_string(__LINE__));                       \
      return IMDCU_ECUDA;                                                                \
    }                                                                                    \
  } while (0)

#define RET_IF(expr)                                                                     \
  do { int _r = (expr); if (_r != IMDCU_OK) return _r; } while (0)

// ----------------------------------------------------------------------------- buffers
template <typename T>
struct DevBuf {
  T     *p   = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t n, double slack = 1.0) {
    if (n <= cap) return cudaSuccess;
    size_t want = (size_t) (n * slack) + 64;
    T *q = nullptr;
    cudaError_t e = cudaMalloc(&q, want * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p) cudaFree(p);
    p = q; cap = want;
    return cudaSuccess;
  }
  cudaError_t ensure_exact(size_t n) {          // capacity >= n, no slack added
    if (n <= cap) return cudaSuccess;
    T *q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p) cudaFree(p);
    p = q; cap = n;
    return cudaSuccess;
  }
  // grow but keep the first `keep` elements
  cudaError_t ensure_keep(size_t n, size_t keep, cudaStream_t st, double slack = 1.0) {
    if (n <= cap) return cudaSuccess;
    size_t want = (size_t) (n * slack) + 64;
    T *q = nullptr;
    cudaError_t e = cudaMalloc(&q, want * sizeof(T));
    if (e != cudaSuccess) return e;
    if (p && keep) {
      e = cudaMemcpyAsync(q, p, keep * sizeof(T), cudaMemcpyDeviceToDevice, st);
      if (e != cudaSuccess) return e;
      cudaStreamSynchronize(st);
    }
    if (p) cudaFree(p);
    p = q; cap = want;
    return cudaSuccess;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// ----------------------------------------------------------------------------- tables
// One tabulated function set on the device.  Row k of column c holds the three numbers
// the reference's 3-point macros derive from table rows k..k+2 (potaccess.h:344-349):
//   {p0, dv = p1-p0, d2v = p2-2*p1+p0}
// computed once on the host with the same operations, so a lookup is one 24-byte read.
struct DevTable {
  int     ncols = 0, nrows = 0;       // nrows = maxsteps
  double *coef = nullptr;             // [ncols][nrows][3]
  double *begin = nullptr, *end = nullptr, *invstep = nullptr, *step = nullptr;   // [ncols] (device)
  double *raw = nullptr, *raw2 = nullptr;   // 4-point / spline modes: table and table2 as handed in, [(nrows+2)][ncols]
  std::vector<double> h_begin, h_end, h_step, h_invstep, h_table;
  std::vector<int>    h_len;
  int     maxsteps = 0;
  bool    set = false;
  size_t  coef_bytes() const { return (size_t) ncols * nrows * 3 * sizeof(double); }
};

struct TabView {                      // passed by value to kernels
  const double *coef;                 // global coefficient rows
  const double *begin, *end, *invstep, *step;
  const double *raw, *raw2;           // 4-point / spline modes (row-major, row stride ncols)
  int ncols, nrows;
};

// ----------------------------------------------------------------------------- geometry
struct Geom {                         // passed by value to kernels
  double box[9];                      // rows box_x, box_y, box_z
  double tbox[9];
  int    pbc[3];
  int    gdim[3];                     // global_cell_dim
  int    ldim[3];                     // cells owned by this rank (x-slab decomposition)
  int    edim[3];                     // ldim + 2
  int    xoff;                        // first owned global x cell layer
  double cellsz;                      // (r_cut + margin)^2
};

// ghost-cell plan entry (host built at set_box; stage order z, y, x like send_cells,
// imd_comm_force_3d.c:270-396)
struct GhostCell {
  int cell;        // extended index of the ghost cell
  int src;         // extended index of the source cell (real or earlier-stage ghost)
  int code;        // shift of this stage: trits (sx+1) + 3*(sy+1) + 9*(sz+1)
};

struct Reduce8 { double v[8]; };

// ----------------------------------------------------------------------------- TTM (1-D)
// One FD cell, the fields of ttm_Element (types.h:441-492) the TTM1D path touches.
struct TCell {
  int    natoms, natoms_old;
  double temp, xi, md_temp, source, dens, fd_k, fd_g, Z, Ce, ne, U;
};

// rectilinear EOS table + diagonal flags of the reference's Delaunay triangulation
struct EosView {
  int nx, ny;
  const double *x, *y, *z;
  const unsigned char *diag;
};

struct TtmState {
  bool   inited = false;
  size_t grid_smem_set = 0;          // dynamic smem opted in for k_ttm_grid
  cudaStream_t stream2 = nullptr;    // k_ttm_grid beside k_eam_pass2 (imdcu_ttm_fork / imdcu_ttm_join)
  cudaEvent_t  ev_fork = nullptr, ev_join = nullptr;
  bool   forked = false;
  imdcu_ttm_params p;
  int    G = 0, P = 1, L = 0, LD = 0;
  double fd_hx = 0, fd_vol = 0, neighvol = 0, omega_laser = 0, sigma_t_sq = 0;
  // EOS tables (device)
  DevBuf<double> eos_buf[3];
  DevBuf<unsigned char> eos_diag[3];
  int    eos_nx[3] = {0, 0, 0}, eos_ny[3] = {0, 0, 0};
  bool   eos_set[3] = {false, false, false};
  // do_tmm (imdcu_ttm_set_tmm): data tables, scratch and the persistent tmm_Qabs
  bool   use_tmm = false;
  double tmm_threshold = 0, laser_t_1 = 0, laser_sigma_t1 = 0;
  int    tmm_nbb = 0, tmm_nk = 0;
  DevBuf<double> tmm_tab, tmm_buf;
  // lattices: lat[0], lat[1] = [P*LD] TCell each; vlat[0], vlat[1] = [vlatdim]
  DevBuf<TCell> lat, vlat;
  // slab sums: ints {natoms, totneighs, fluxl, fluxr}[G], doubles {px,py,pz,mdtemp}[G]
  DevBuf<int>    isum;
  DevBuf<double> dsum;
  DevBuf<double> xiarr;             // [G] xiarr_global
  DevBuf<double> xiloc;             // [P*LD] xiarr_local
  DevBuf<double> vcom;              // [3*G]
  DevBuf<double> src_e;             // [P*LD] exp(-mu*depth) per cell
  DevBuf<double> winv;              // [P*L + vlatdim][8] density-only terms of a cell, per MD step
  DevBuf<int>    fdi, fdi_alt;      // per atom: slab of the previous step (p->fdi)
  // device scalars: see ttm.cu (TtmScal)
  void  *d_scal = nullptr;
  void  *h_scal = nullptr;          // pinned mirror
};

// ----------------------------------------------------------------------------- TTM (3-D sweep, ttm3d.cu)
struct Ttm3dState {
  bool   inited = false;
  int    nx = 0, ny = 0, nz = 0, pbc[3] = {0, 0, 0}, fd_min_atoms = 0;
  double h[3] = {0, 0, 0};
  long long pz = 0, py = 0, total = 0;          // pitches and size of one field (elements)
  int    cur = 0;                                // which temp buffer is l1
  DevBuf<int>    natoms;
  DevBuf<double> temp[2], fd_k, fd_g, Ce, md_temp, source, U, xi, dens, partials;
  double *d_eabs = nullptr;
};

// ----------------------------------------------------------------------------- NRB (nrb.cu)
#define IMDCU_VTYPE_NRB (1 << 30)     // vtype bit: NRBBND > 0, the integrators leave the atom without a kick
struct NrbState {
  bool   active = false, startup_images = false;
  int    nent = 0, maxid = 0;
  double k = 0.0;                     // sqrt(nrb_k / 26.9815), imd_nrb.c:149
  DevBuf<int>    ent_bnd, ent_nrbi, ent_index, slot_of_id;   // [nent], [12 nent], [nent], [maxid + 1]
  DevBuf<double> ent_ref, ent_newp;                           // [3 nent]
};

// ----------------------------------------------------------------------------- context
struct imdcu_ctx {
  int          device = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t up_stream = nullptr;          // imdcu_update_atoms: the momenta travel beside the force passes
  cudaEvent_t  ev_up = nullptr, ev_up0 = nullptr;
  bool         up_pending = false;
  bool         force_sums_pending = false;   // multi-GPU: Epot/virial sums ride on the integrator's all-reduce
  std::string  err;
  bool         err_agreed = false;           // multi-GPU: every rank knows about the error being returned (multi.cu)
  int          short_warn = 0;               // is_short of the EAM passes seen since the last imdcu_short_distance()
  int          num_sms = 148;
  size_t       smem_optin = 0;

  // parameters
  int    ntypes = 1, nvtypes = 1;
  int    interp = IMDCU_INTERP_3POINT;       // table interpolation of the run (imdcu_set_interp)
  double timestep = 0.0, nbl_margin = 0.0, nbl_size = 1.3;
  std::vector<double> restrictions;          // host copy
  DevBuf<double> d_restrictions;
  int    pdecay = 0; double xipdecay = 0, ramp_start = 0, ramp_end = 1;
  double *fused_coef = nullptr;              // pair + rho rows interleaved (context.cu: imdcu_fused_table)
  int    fused_state = 0;                    // 0 not looked at, 1 built, -1 tables do not share a grid
  // IMDCU_DEBUG_REBUILD=1: wall-clock marks along the rebuild path (imdcu_dbg_mark synchronises at each), printed per rank
  int    dbg_rebuild = 0;
  std::chrono::steady_clock::time_point dbg_t;
  std::string dbg_line;
  std::vector<int> xcut;                     // [nranks+1] global x cell layers where the slabs begin (context.cu)
  double *rep4_tab = nullptr;                // raw {pair, rho} rows, four bank-staggered copies (imdcu_rep4_table)
  int    rep4_state = 0;
  size_t rep4_bytes = 0;
  bool   ttm_isolate = true;                 // forked k_ttm_grid takes a whole SM (IMDCU_TTM_ISOLATE)
  size_t ttm_isolate_smem = 0;               // dynamic smem request that keeps k_eam_pass2 CTAs off that SM
  int    fuse_ttm_steps = -1;                // >= 0: imdcu_eam_forces forks calc_ttm(steps) between its passes
  bool   box_set = false, params_set = false;
  // kernel-selection knobs (imdcu_set_tuning; the IMDCU_* environment variables give the defaults at create)
  struct Tune {
    int lpa = 0;                             // lanes per atom of the EAM passes: 0 = by system size, 1/2/4/8
    int nbl_filter = 1;                      // neighbour-list build: 1 float classifier with the exact test in the uncertain
                                             // shell (k_nbl_build2, default), 0 exact tests only (k_nbl_build): same list
    int rep4_block = 896;                    // threads per CTA of k_eam_pass1_rep4 (896 or 768)
    int fused_table = 2;                     // pass 1 when the pair and rho tables share their grid: 1 k_eam_pass1_fused, 2 k_eam_pass1_rep4, 0 general kernel
    int ttm_overlap = 1;                     // imdcu_run_ttm: lattice kernel beside EAM pass 2
    int ttm_lat_smem = 1;                    // k_ttm_grid keeps the lattices in shared memory
    int ttm3d_variant = 9;                   // direct 3-D sweep kernel: brick shape / CTAs per SM / loads up front (ttm3d.cu: T3D_VARIANTS)
    int ttm3d_tiled = 0;                     // 3-D sweep: 2.5-D tiled kernel instead of the direct one
  } tune;

  Geom   geom;
  DevTable tab[3];

  // atoms: [0,n) real, [n, n+ng) ghosts
  int64_t n = 0, ng = 0;
  DevBuf<double4> pos4, pos4_alt;            // x,y,z,dF
  DevBuf<double4> mom4, mom4_alt;            // px,py,pz,mass
  DevBuf<double4> frc4;                      // fx,fy,fz,epot
  DevBuf<double>  rho, dF, refx, refy, refz;
  DevBuf<int>     type, type_alt;            // [n+ng]
  DevBuf<int>     vtype, vtype_alt, id, id_alt, numneighs;
  DevBuf<int>     key, key_alt, perm, perm_alt;
  DevBuf<unsigned char> cub_tmp;
  DevBuf<double>  stage;                     // host<->device staging (atoms.cu)
  DevBuf<double>  dist_sums;                 // spatial distributions (distrib.cu): [4][dist_size] sums
  DevBuf<float>   dist_out;                  //   the same as float, as they cross the ABI

  // cells (extended grid)
  int  ncell_ext = 0;
  DevBuf<int> cell_begin, cell_count;        // [ncell_ext]
  std::vector<GhostCell> gplan;              // host
  int  gstage_cells[4] = {0, 0, 0, 0};       // plan index ranges per stage z,y,x
  DevBuf<GhostCell> d_gplan;
  DevBuf<int> gcell_off;                     // [nplan+1] exclusive scan of ghost cell counts
  int64_t gstage_atoms[4] = {0, 0, 0, 0};    // ghost atom ranges per stage (relative to n)
  DevBuf<int> ghost_src;                     // [ng] immediate source index
  DevBuf<int> ghost_code;                    // [ng] stage shift code
  DevBuf<int> ghost_root;                    // [ng] owning real atom (local)
  DevBuf<int> ghost_tcode;                   // [ng] total shift code
  DevBuf<double> ghost_raw;                  // [3*ng] root's unshifted position at build

  // neighbour list
  bool have_valid_nbl = false;
  int  nbl_count = 0;
  int  nbl_cap = 0;                          // slots per atom
  DevBuf<int> nbl_idx;                       // [ntiles*cap*32]
  DevBuf<int> nbl_num;                       // [n]
  DevBuf<float4> posf;                       // [n+ng] single-precision positions for the first phase of the list build
  unsigned int *d_xmax = nullptr;            // bits of the largest |coordinate| (float), measured at every rebuild
  int  last_max_neigh = 0;

  // reductions / flags (device + pinned host mirror)
  DevBuf<double> partials;                   // per-block partial sums
  double *d_results = nullptr;               // [32] doubles
  double *h_results = nullptr;               // pinned
  unsigned int *d_ticket = nullptr;          // last-block tickets
  int *d_flags = nullptr;                    // [8] ints: overflow, maxneigh, is_short ...
  int *h_flags = nullptr;                    // pinned

  TtmState ttm;
  Ttm3dState ttm3d;
  NrbState nrb;

  // multi-GPU (x-slab decomposition, multi.cu); nranks == 1: everything below is unused
  struct Comm {
    void *nccl = nullptr;                    // ncclComm_t
    int   rank = 0, nranks = 1;
    int   west = -1, east = -1;              // x-neighbour ranks, -1: open end
    int   wrap_w = 0, wrap_e = 0;            // message to west/east crosses the periodic x boundary
    int   nplane = 0;                        // cells per yz plane of the extended grid (ey*ez)
    int   q_east = 0, q_west = 0;            // first plan entry of the ghost planes x = ex-1 / x = 0
    int   plane_cnt[2] = {0, 0};             // atoms in my west / east plane (packed layout: west first)
    int   nsend[2] = {0, 0}, nrecv[2] = {0, 0};   // [0] west plane / from east, [1] east plane / from west
    int64_t recv_off[2] = {0, 0};            // ghost atom offset (relative to n) of the two received planes
    DevBuf<int>     send_idx;                // [nsend[0] + nsend[1]] local indices, plane-cell order
    DevBuf<double4> send_buf;                // packed records
    DevBuf<int>     cnt_send, cnt_recv;      // [2*nplane] per-cell occupancies
    DevBuf<double>  build_send, build_recv;  // build-time metadata records
    DevBuf<double>  mig_send, mig_recv;      // migration records
    int  *d_migcnt = nullptr, *h_migcnt = nullptr;   // [4]
  } comm;
  DevBuf<int> ghost_gid;                     // [ng] atom id of every image atom (pair export)

  // bookkeeping
  int64_t launches = 0;
  bool    profiling = false;
  double  phase_ms[IMDCU_NPHASE] = {0};
  int64_t phase_launches[IMDCU_NPHASE] = {0};
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int     cur_phase = -1;

  void set_error(const std::string &m) { err = m; }
  int  fail(int code, const std::string &m) { err = m; return code; }
  void count_launch(int k = 1) { launches += k; if (cur_phase >= 0) phase_launches[cur_phase] += k; }
};

struct PhaseTimer {                 // CUDA-event bracket on the context's stream
  imdcu_ctx *c; int phase;
  PhaseTimer(imdcu_ctx *ctx, int ph) : c(ctx), phase(ph) {
    c->cur_phase = ph;
    if (c->profiling) cudaEventRecord(c->ev0, c->stream);
  }
  ~PhaseTimer() {
    if (c->profiling) {
      cudaEventRecord(c->ev1, c->stream);
      cudaEventSynchronize(c->ev1);
      float ms = 0; cudaEventElapsedTime(&ms, c->ev0, c->ev1);
      c->phase_ms[phase] += ms;
    }
    c->cur_phase = -1;
  }
};

// momenta uploaded by imdcu_update_atoms on the second stream must have landed before anything reads or
// writes mom4 (cell sort, nrb_forces, update_fd, integrators, downloads): those call this first
static inline int imdcu_join_upload(imdcu_ctx *c)
{
  if (c->up_pending) {
    c->up_pending = false;
    CU_CHECK(c, cudaStreamWaitEvent(c->stream, c->ev_up, 0));
  }
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- internal API
int imdcu_rebuild_cells(imdcu_ctx *ctx);                 // cells.cu: wrap, bin, sort, ghosts
int imdcu_refresh_ghosts(imdcu_ctx *ctx, bool at_build); // cells.cu
int imdcu_build_nblist(imdcu_ctx *ctx);                  // nblist.cu
int imdcu_eam_forces(imdcu_ctx *ctx, double out8[8]);    // eam.cu
int imdcu_integrate_nve(imdcu_ctx *ctx, double out8[8], int *valid);   // integrate.cu
int imdcu_integrate_ttm(imdcu_ctx *ctx, double out8[8], int *valid);   // integrate.cu
int imdcu_ttm_step(imdcu_ctx *ctx, int steps, double out16[16]);       // ttm.cu
// multi.cu
int  imdcu_multi_migrate(imdcu_ctx *ctx, int *n_total);      // after k_wrap_key: movers out, arrivals appended
int  imdcu_multi_ghost_counts(imdcu_ctx *ctx);               // x-stage ghost cell occupancies from the neighbours
int  imdcu_multi_ghost_build(imdcu_ctx *ctx);                // x-stage ghost atoms + metadata at a rebuild
int  imdcu_multi_ghost_refresh(imdcu_ctx *ctx);              // x-stage positions / F' every step
int  imdcu_multi_allreduce(imdcu_ctx *ctx, void *buf, int count, int kind);   // kind 0 sum f64, 1 max f64, 2 sum i32
int  imdcu_multi_allreduce2(imdcu_ctx *ctx, void *buf_a, int count_a, int kind_a, void *buf_b, int count_b, int kind_b);
void imdcu_multi_release(imdcu_ctx *ctx);
int  imdcu_multi_agree(imdcu_ctx *ctx, int local_rc);        // rank-local error -> the same return on every rank
void imdcu_multi_abort(imdcu_ctx *ctx);                      // un-agreed failure: abort the communicator
__device__ __host__ inline int imdcu_combine_code(int pc, int code)
{
  return (pc % 3 - 1 + code % 3 - 1 + 1) + 3 * ((pc / 3) % 3 - 1 + (code / 3) % 3 - 1 + 1) + 9 * (pc / 9 - 1 + code / 9 - 1 + 1);
}
void imdcu_dbg_mark(imdcu_ctx *ctx, const char *label, bool flush = false);   // context.cu
const double *imdcu_fused_table(imdcu_ctx *ctx);                       // context.cu
const double *imdcu_rep4_table(imdcu_ctx *ctx, size_t *bytes);         // context.cu
void imdcu_ttm_release(imdcu_ctx *ctx);                                // ttm.cu
int  imdcu_ttm_fork(imdcu_ctx *ctx, int steps);                        // ttm.cu
int  imdcu_ttm_join(imdcu_ctx *ctx, double out16[16]);
void imdcu_ttm3d_release(imdcu_ctx *ctx);                              // ttm3d.cu
void imdcu_nrb_release(imdcu_ctx *ctx);                                // nrb.cu
int  imdcu_nrb_reindex(imdcu_ctx *ctx, bool mark);                     // after a cell sort
int  imdcu_nrb_forces(imdcu_ctx *ctx, int steps);                      // nrb_forces (imd_nrb.c:792)
int  imdcu_check_flags(imdcu_ctx *ctx);                                // nblist.cu: device-side error flags
int  imdcu_multi_sendrecv_x(imdcu_ctx *ctx, int west, int east, const void *s_w, const void *s_e, void *r_e,
                            void *r_w, size_t count, int is_int);         // multi.cu
TabView imdcu_tab_view(const DevTable &t);

// host -> device copy ordered with the context's (non-blocking) stream: a plain cudaMemcpy from pageable
// memory may return before its last DMA chunk has landed, and kernels on another stream would race it
static inline cudaError_t imdcu_h2d(imdcu_ctx *c, void *dst, const void *src, size_t bytes)
{
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(c->stream);
}

static inline int div_up(int64_t a, int64_t b) { return (int) ((a + b - 1) / b); }
