// imd_b200/csrc/api.cu -- the hot-path entry points of the C ABI (include/imdcu.h)
#include "device.cuh"
#include <mutex>

static int check_ready(imdcu_ctx *c)
{
  if (!c) return IMDCU_EINVAL;
  if (!c->box_set) return c->fail(IMDCU_EINVAL, "imdcu_set_box has not been called");
  if (!c->params_set) return c->fail(IMDCU_EINVAL, "imdcu_set_params has not been called");
  for (int t = 0; t < 3; t++)
    if (!c->tab[t].set) return c->fail(IMDCU_EINVAL, "potential tables not set");
  if (c->tab[IMDCU_TAB_PAIR].ncols != c->ntypes * c->ntypes || c->tab[IMDCU_TAB_RHO].ncols != c->ntypes * c->ntypes ||
      c->tab[IMDCU_TAB_EMBED].ncols != c->ntypes)
    return c->fail(IMDCU_EINVAL, "table column counts do not match ntypes");
  cudaSetDevice(c->device);
  return IMDCU_OK;
}

// Leaving a hot-path entry point with an error: work forked to the second stream is joined, and on several
// GPUs a failure the other ranks do not know about aborts the communicator (their pending collectives then
// fail instead of hanging; multi.cu).
static int leave(imdcu_ctx *c, int rc)
{
  if (rc == IMDCU_OK) return rc;
  const std::string msg = c->err;
  if (c->ttm.forked) { cudaStreamSynchronize(c->ttm.stream2); c->ttm.forked = false; }
  c->fuse_ttm_steps = -1;
  imdcu_multi_abort(c);
  c->err = msg;
  return rc;
}

static int calc_forces_impl(imdcu_ctx *c, int steps, double out8[8]);

// calc_forces (imd_forces_nbl.c:295-1054)
extern "C" int imdcu_calc_forces(imdcu_ctx *c, int steps, double out8[8])
{
  RET_IF(check_ready(c));
  return leave(c, calc_forces_impl(c, steps, out8));
}

static int calc_forces_impl(imdcu_ctx *c, int steps, double out8[8])
{
  if (!c->have_valid_nbl) {
    RET_IF(imdcu_rebuild_cells(c));              // fix_cells (:304)
    RET_IF(imdcu_refresh_ghosts(c, true));       // send_cells(copy_cell) (:311)
    RET_IF(imdcu_build_nblist(c));               // make_nblist (:315)
    RET_IF(imdcu_nrb_reindex(c, false));
  } else {
    RET_IF(imdcu_refresh_ghosts(c, false));
  }
  // nrb_forces (:1019-1021) reads positions, momenta and the list, not the forces: it runs first, so
  // that calc_ttm (forked between the EAM passes by imdcu_run_ttm) sees the new boundary momenta
  RET_IF(imdcu_nrb_forces(c, steps));
  return imdcu_eam_forces(c, out8);
}

// move_atoms_nve + check_nblist (imd_main_3d.c:564, :812)
extern "C" int imdcu_move_atoms_nve(imdcu_ctx *c, double out8[8], int *have_valid_nbl)
{
  RET_IF(check_ready(c));
  return leave(c, imdcu_integrate_nve(c, out8, have_valid_nbl));
}

// main-loop segment (imd_main_3d.c:151-916 restricted to the hot calls)
static int run_nve_impl(imdcu_ctx *c, int first_step, int nsteps, double *energies)
{
  double f8[8], m8[8];
  int valid = 0;
  for (int s = 0; s < nsteps; s++) {
    RET_IF(calc_forces_impl(c, first_step + s, energies ? f8 : nullptr));
    RET_IF(imdcu_integrate_nve(c, m8, &valid));
    if (energies) { energies[2 * s] = f8[0]; energies[2 * s + 1] = m8[0]; }
  }
  return IMDCU_OK;
}

extern "C" int imdcu_run_nve(imdcu_ctx *c, int first_step, int nsteps, double *energies)
{
  RET_IF(check_ready(c));
  return leave(c, run_nve_impl(c, first_step, nsteps, energies));
}

// calc_ttm (imd_main_3d.c:534 -> imd_ttm.c:75)
extern "C" int imdcu_calc_ttm(imdcu_ctx *c, int steps, double out16[16])
{
  RET_IF(check_ready(c));
  return leave(c, imdcu_ttm_step(c, steps, out16));
}

// calc_forces + calc_ttm of one main-loop iteration in ONE call (imd_main_3d.c:406 and :534; nothing that touches the
// atoms lies between them): calc_ttm's atom kernels run between the two EAM passes and its lattice kernel beside the
// second one, as in imdcu_run_ttm.  Same inputs and results as imdcu_calc_forces followed by imdcu_calc_ttm.
extern "C" int imdcu_calc_forces_ttm(imdcu_ctx *c, int steps, double out8[8], double out16[16])
{
  RET_IF(check_ready(c));
  if (!c->ttm.inited) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init has not been called");
  if (c->profiling || !c->tune.ttm_overlap) {
    const int rc = calc_forces_impl(c, steps, out8);
    if (rc != IMDCU_OK) return leave(c, rc);
    return leave(c, imdcu_ttm_step(c, steps, out16));
  }
  c->fuse_ttm_steps = steps;
  const int rc = calc_forces_impl(c, steps, out8);
  c->fuse_ttm_steps = -1;
  if (rc != IMDCU_OK) return leave(c, rc);
  return leave(c, imdcu_ttm_join(c, out16));
}

// move_atoms_ttm + check_nblist (imd_main_3d.c:564, :812)
extern "C" int imdcu_move_atoms_ttm(imdcu_ctx *c, double out8[8], int *have_valid_nbl)
{
  RET_IF(check_ready(c));
  return leave(c, imdcu_integrate_ttm(c, out8, have_valid_nbl));
}

// main-loop segment with the TTM (imd_main_3d.c:406,534,564,812)
static int run_ttm_impl(imdcu_ctx *c, int first_step, int nsteps, double *energies);

extern "C" int imdcu_run_ttm(imdcu_ctx *c, int first_step, int nsteps, double *energies)
{
  RET_IF(check_ready(c));
  return leave(c, run_ttm_impl(c, first_step, nsteps, energies));
}

static int run_ttm_impl(imdcu_ctx *c, int first_step, int nsteps, double *energies)
{
  double f8[8], m8[8], t16[16];
  int valid = 0;
  if (!c->ttm.inited) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init has not been called");
  // calc_ttm runs between the two EAM passes and its lattice kernel beside pass 2 (same inputs and
  // results as the sequential calc_forces; calc_ttm; only the schedule differs).  With per-phase
  // profiling on, the sequential order is kept so that the phase times stay separable.
  const bool fuse = !c->profiling && c->tune.ttm_overlap;
  for (int s = 0; s < nsteps; s++) {
    double *t_out = (s == nsteps - 1) ? t16 : nullptr;                    // errors surface at the end
    if (fuse) {
      c->fuse_ttm_steps = first_step + s;
      const int rc = calc_forces_impl(c, first_step + s, energies ? f8 : nullptr);
      c->fuse_ttm_steps = -1;
      RET_IF(rc);
      RET_IF(imdcu_ttm_join(c, t_out));
    } else {
      RET_IF(calc_forces_impl(c, first_step + s, energies ? f8 : nullptr));
      RET_IF(imdcu_ttm_step(c, first_step + s, t_out));
    }
    RET_IF(imdcu_integrate_ttm(c, m8, &valid));
    if (energies) { energies[2 * s] = f8[0]; energies[2 * s + 1] = m8[0]; }
  }
  return IMDCU_OK;
}

// is_short of the EAM passes since the last call (the reference's "short distance" warning,
// imd_forces_nbl.c:673,992); reads the device flags and clears the count
extern "C" int imdcu_short_distance(imdcu_ctx *c)
{
  if (!c) return IMDCU_EINVAL;
  cudaSetDevice(c->device);
  const int rc = imdcu_check_flags(c);
  if (rc != IMDCU_OK) return rc;
  const int w = c->short_warn;
  c->short_warn = 0;
  return w;
}

extern "C" int imdcu_nbl_count(const imdcu_ctx *c) { return c ? c->nbl_count : -1; }

extern "C" int imdcu_invalidate_nblist(imdcu_ctx *c)
{
  if (!c) return IMDCU_EINVAL;
  c->have_valid_nbl = false;
  return IMDCU_OK;
}

extern "C" int64_t imdcu_launch_count(const imdcu_ctx *c) { return c ? c->launches : -1; }

extern "C" int imdcu_set_profiling(imdcu_ctx *c, int enabled)
{
  if (!c) return IMDCU_EINVAL;
  c->profiling = enabled != 0;
  for (int k = 0; k < IMDCU_NPHASE; k++) { c->phase_ms[k] = 0; c->phase_launches[k] = 0; }
  return IMDCU_OK;
}

extern "C" int imdcu_get_phase_times(imdcu_ctx *c, double ms[IMDCU_NPHASE], int64_t launches[IMDCU_NPHASE])
{
  if (!c) return IMDCU_EINVAL;
  for (int k = 0; k < IMDCU_NPHASE; k++) {
    if (ms) ms[k] = c->phase_ms[k];
    if (launches) launches[k] = c->phase_launches[k];
  }
  return IMDCU_OK;
}

extern "C" void *imdcu_stream(const imdcu_ctx *c) { return c ? (void *) c->stream : nullptr; }

extern "C" int imdcu_synchronize(imdcu_ctx *c)
{
  if (!c) return IMDCU_EINVAL;
  cudaSetDevice(c->device);
  RET_IF(imdcu_join_upload(c));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  return imdcu_check_flags(c);
}

// ----------------------------------------------------------------------------- table probe
__global__ void k_eval_table(TabView v, int interp, int col, int64_t n, const double *__restrict__ arg,
                             double *__restrict__ val, double *__restrict__ grad)
{
  int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double p, g; int is_short = 0;
  if (interp == IMDCU_INTERP_4POINT) raw_pair_int<1>(raw_col(v, col), arg[i], p, g, is_short);
  else if (interp == IMDCU_INTERP_SPLINE) raw_pair_int<2>(raw_col(v, col), arg[i], p, g, is_short);
  else pair_int2(tab_col(v, v.coef, col), arg[i], p, g, is_short);
  val[i] = p; grad[i] = g;
}

extern "C" int imdcu_eval_table(imdcu_ctx *c, int which, int col, int64_t n, const double *arg,
                                double *val, double *grad)
{
  if (!c || which < 0 || which > 2 || !c->tab[which].set || col < 0 || col >= c->tab[which].ncols || n < 0)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_eval_table: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  if (n == 0) return IMDCU_OK;
  if (c->interp == IMDCU_INTERP_SPLINE && !c->tab[which].raw2)       // before the allocation: nothing to release
    return c->fail(IMDCU_EINVAL, "imdcu_eval_table: spline mode needs imdcu_set_table2");
  double *d = nullptr;
  CU_CHECK(c, cudaMalloc(&d, 3 * n * sizeof(double)));
  CU_CHECK(c, cudaMemcpyAsync(d, arg, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_eval_table<<<div_up(n, 256), 256, 0, c->stream>>>(imdcu_tab_view(c->tab[which]), c->interp, col, n, d, d + n, d + 2 * n);
  c->count_launch();
  CU_CHECK(c, cudaMemcpyAsync(val, d + n, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaMemcpyAsync(grad, d + 2 * n, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  cudaFree(d);
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- table probe, allocation-free
// imdcu_eval_table without its per-call cudaMalloc / cudaFree, for callers that may not allocate (code under
// CUDA graph capture, tight serving loops, harnesses that forbid it).  The staging that imdcu_eval_table takes
// from cudaMalloc (argument, value, gradient: 3 n doubles) is a file-scope device array of fixed size instead:
// 3 x IMDCU_EVAL_NOALLOC_MAX doubles = 1.5 MiB, held by every GPU that loads this library until the process exits.
//   * In HOST code `g_eval_scratch` is only a handle of the symbol, not a device address: the address is asked
//     for once per device, with that device current (eval_scratch_address), and that pointer goes to the kernel.
//   * Thirds are fixed (arg at 0, val at MAX, grad at 2 MAX) and every element the kernel or the copies read is
//     written earlier in the same call, so nothing has to start at zero although the array keeps the last call's
//     contents.
//   * Every call on a GPU shares the one array: calls that overlap on the GPU (two contexts, two host threads,
//     two streams) race.  This entry point synchronises its stream before it returns.
static __device__ __align__(256) double g_eval_scratch[3 * (size_t) IMDCU_EVAL_NOALLOC_MAX];

static const int kMaxScratchDevices = 64;
static double    *g_eval_scratch_ptr[kMaxScratchDevices];      // per device, resolved once
static std::mutex g_eval_scratch_mutex;

// the device must be current
static cudaError_t eval_scratch_address(int device, double **out)
{
  std::lock_guard<std::mutex> lock(g_eval_scratch_mutex);
  if (g_eval_scratch_ptr[device] == nullptr) {
    double *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **) &p, g_eval_scratch);
    if (e != cudaSuccess) return e;
    g_eval_scratch_ptr[device] = p;
  }
  *out = g_eval_scratch_ptr[device];
  return cudaSuccess;
}

extern "C" int imdcu_eval_table_noalloc(imdcu_ctx *c, int which, int col, int64_t n, const double *arg,
                                        double *val, double *grad)
{
  if (!c || which < 0 || which > 2 || !c->tab[which].set || col < 0 || col >= c->tab[which].ncols || n < 0 ||
      (n > 0 && (!arg || !val || !grad)))
    return c ? c->fail(IMDCU_EINVAL, "imdcu_eval_table_noalloc: bad argument") : IMDCU_EINVAL;
  if (n > IMDCU_EVAL_NOALLOC_MAX)
    return c->fail(IMDCU_EINVAL, "imdcu_eval_table_noalloc: n = " + std::to_string((long long) n) +
                   " exceeds the static scratch of IMDCU_EVAL_NOALLOC_MAX = " + std::to_string(IMDCU_EVAL_NOALLOC_MAX) +
                   " arguments per call (split the call, or use imdcu_eval_table)");
  if (c->interp == IMDCU_INTERP_SPLINE && !c->tab[which].raw2)
    return c->fail(IMDCU_EINVAL, "imdcu_eval_table_noalloc: spline mode needs imdcu_set_table2");
  if (c->device < 0 || c->device >= kMaxScratchDevices)
    return c->fail(IMDCU_EINVAL, "imdcu_eval_table_noalloc: device ordinal beyond the per-device address table");
  CU_CHECK(c, cudaSetDevice(c->device));
  if (n == 0) return IMDCU_OK;
  double *d = nullptr;
  CU_CHECK(c, eval_scratch_address(c->device, &d));
  double *d_arg = d, *d_val = d + IMDCU_EVAL_NOALLOC_MAX, *d_grad = d + 2 * (size_t) IMDCU_EVAL_NOALLOC_MAX;
  CU_CHECK(c, cudaMemcpyAsync(d_arg, arg, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  k_eval_table<<<div_up(n, 256), 256, 0, c->stream>>>(imdcu_tab_view(c->tab[which]), c->interp, col, n, d_arg, d_val, d_grad);
  CU_CHECK(c, cudaGetLastError());
  c->count_launch();
  CU_CHECK(c, cudaMemcpyAsync(val, d_val, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaMemcpyAsync(grad, d_grad, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- FP64 peak probe
// Dependent-free DFMA chains, eight per thread, every SM full: the FP64-FMA rate this GPU sustains, the
// denominator of the FP64 fraction bench.py reports for the EAM passes (SURVEY 8d asks for HBM AND FP64).
__global__ void __launch_bounds__(1024)
k_fp64_probe(double *out, int iters, double a, double b)
{
  double x0 = threadIdx.x * 1e-3, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 12345.678) out[0] = s;                       // keeps the chains alive, never true for these inputs
}

extern "C" int imdcu_probe_fp64(imdcu_ctx *c, double *gflops)
{
  if (!c || !gflops) return IMDCU_EINVAL;
  cudaSetDevice(c->device);
  const int iters = 1 << 15, grid = c->num_sms * 2, block = 1024;
  double best = 0.0;
  for (int rep = 0; rep < 4; rep++) {
    CU_CHECK(c, cudaEventRecord(c->ev0, c->stream));
    k_fp64_probe<<<grid, block, 0, c->stream>>>(c->d_results + 24, iters, 0.999999, 1e-7);
    c->count_launch();
    CU_CHECK(c, cudaEventRecord(c->ev1, c->stream));
    CU_CHECK(c, cudaEventSynchronize(c->ev1));
    float ms = 0; CU_CHECK(c, cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    const double gf = 2.0 * 8.0 * (double) iters * grid * block / (ms * 1e-3) / 1e9;
    if (rep > 0 && gf > best) best = gf;
  }
  *gflops = best;
  return IMDCU_OK;
}
