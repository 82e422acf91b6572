// imd_b200/csrc/context.cu -- life cycle, set-up and host<->device staging of libimdcu.so
#include "common.cuh"

#include <cmath>

static thread_local std::string g_create_error;

extern "C" int imdcu_abi_version(void) { return IMDCU_ABI_VERSION; }

extern "C" const char *imdcu_last_error(const imdcu_ctx *ctx)
{
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int imdcu_device_count(int *n)
{
  if (!n) return IMDCU_EINVAL;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (libimdcu has no CPU fallback)";
    *n = 0;
    return IMDCU_ENODEVICE;
  }
  *n = ndev;
  return IMDCU_OK;
}

extern "C" int imdcu_create(imdcu_ctx **out, int device)
{
  if (!out) return IMDCU_EINVAL;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                     " (libimdcu has no CPU fallback)";
    return IMDCU_ENODEVICE;
  }
  if (device < 0 || device >= ndev) {
    g_create_error = "device ordinal out of range";
    return IMDCU_EINVAL;
  }
  imdcu_ctx *c = new imdcu_ctx();
  c->device = device;
#define CREATE_CHECK(expr)                                                              \
  do { cudaError_t _e = (expr); if (_e != cudaSuccess) {                                \
         g_create_error = std::string(#expr) + ": " + cudaGetErrorString(_e);           \
         delete c; return IMDCU_ECUDA; } } while (0)
  CREATE_CHECK(cudaSetDevice(device));
  cudaDeviceProp prop;
  CREATE_CHECK(cudaGetDeviceProperties(&prop, device));
  c->num_sms = prop.multiProcessorCount;
  c->smem_optin = prop.sharedMemPerBlockOptin;
  {
    // an SM holds sharedMemPerMultiprocessor bytes, every CTA reserves 1 KB on top of its request; a CTA of
    // k_eam_pass2 needs >= 24 KB (one table): asking for all but 16 KB keeps every such CTA away
    const char *e = getenv("IMDCU_TTM_ISOLATE");
    c->ttm_isolate = !(e && atoi(e) == 0);
    size_t want = prop.sharedMemPerMultiprocessor > 18 * 1024 ? prop.sharedMemPerMultiprocessor - 18 * 1024 : 0;
    c->ttm_isolate_smem = want < c->smem_optin ? want : c->smem_optin;
  }
  {
    auto env_int = [](const char *name, int dflt) { const char *e = getenv(name); return e && *e ? atoi(e) : dflt; };
    c->tune.lpa = env_int("IMDCU_LPA", 0);
    c->tune.fused_table = env_int("IMDCU_FUSED_TABLE", 2);
    c->tune.rep4_block = env_int("IMDCU_REP4_BLOCK", 896);
    c->tune.nbl_filter = env_int("IMDCU_NBL_FILTER", 1);
    c->dbg_rebuild = getenv("IMDCU_DEBUG_REBUILD") != nullptr && strcmp(getenv("IMDCU_DEBUG_REBUILD"), "0") != 0;
    c->tune.ttm3d_variant = env_int("IMDCU_TTM3D_VARIANT", 9);
    c->tune.ttm_overlap = env_int("IMDCU_TTM_OVERLAP", 1);
    c->tune.ttm_lat_smem = env_int("IMDCU_TTM_LAT_SMEM", 1);
    const char *e3 = getenv("IMDCU_TTM3D");
    c->tune.ttm3d_tiled = e3 && !strcmp(e3, "tiled");
  }
  CREATE_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CREATE_CHECK(cudaStreamCreateWithFlags(&c->up_stream, cudaStreamNonBlocking));
  CREATE_CHECK(cudaEventCreateWithFlags(&c->ev_up, cudaEventDisableTiming));
  CREATE_CHECK(cudaEventCreateWithFlags(&c->ev_up0, cudaEventDisableTiming));
  CREATE_CHECK(cudaEventCreate(&c->ev0));
  CREATE_CHECK(cudaEventCreate(&c->ev1));
  CREATE_CHECK(cudaMalloc(&c->d_results, 32 * sizeof(double)));
  CREATE_CHECK(cudaMemset(c->d_results, 0, 32 * sizeof(double)));
  CREATE_CHECK(cudaMallocHost(&c->h_results, 32 * sizeof(double)));
  CREATE_CHECK(cudaMalloc(&c->d_ticket, 8 * sizeof(unsigned int)));
  CREATE_CHECK(cudaMemset(c->d_ticket, 0, 8 * sizeof(unsigned int)));
  CREATE_CHECK(cudaMalloc(&c->d_xmax, sizeof(unsigned int)));
  CREATE_CHECK(cudaMalloc(&c->d_flags, 8 * sizeof(int)));
  CREATE_CHECK(cudaMemset(c->d_flags, 0, 8 * sizeof(int)));
  CREATE_CHECK(cudaMallocHost(&c->h_flags, 8 * sizeof(int)));
#undef CREATE_CHECK
  c->restrictions.assign(3, 1.0);
  *out = c;
  return IMDCU_OK;
}

extern "C" int imdcu_destroy(imdcu_ctx *c)
{
  if (!c) return IMDCU_EINVAL;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  cudaStreamSynchronize(c->up_stream);
  cudaEventDestroy(c->ev_up); cudaEventDestroy(c->ev_up0); cudaStreamDestroy(c->up_stream);
  c->pos4.release(); c->pos4_alt.release(); c->mom4.release(); c->mom4_alt.release();
  c->frc4.release(); c->rho.release(); c->dF.release(); c->refx.release(); c->refy.release(); c->refz.release();
  c->type.release(); c->type_alt.release(); c->vtype.release(); c->vtype_alt.release();
  c->id.release(); c->id_alt.release(); c->numneighs.release();
  c->key.release(); c->key_alt.release(); c->perm.release(); c->perm_alt.release();
  c->dist_sums.release(); c->dist_out.release();
  c->cub_tmp.release(); c->stage.release(); c->cell_begin.release(); c->cell_count.release();
  c->d_gplan.release(); c->gcell_off.release(); c->ghost_src.release(); c->ghost_code.release();
  c->ghost_root.release(); c->ghost_tcode.release(); c->ghost_raw.release();
  c->nbl_idx.release(); c->nbl_num.release(); c->partials.release(); c->d_restrictions.release();
  if (c->fused_coef) cudaFree(c->fused_coef);
  if (c->rep4_tab) cudaFree(c->rep4_tab);
  imdcu_ttm_release(c);
  imdcu_ttm3d_release(c); imdcu_nrb_release(c);
  imdcu_multi_release(c);
  c->ghost_gid.release();
  for (int t = 0; t < 3; t++) {
    if (c->tab[t].coef) cudaFree(c->tab[t].coef);
    if (c->tab[t].begin) cudaFree(c->tab[t].begin);
    if (c->tab[t].raw) cudaFree(c->tab[t].raw);
    if (c->tab[t].raw2) cudaFree(c->tab[t].raw2);
  }
  cudaFree(c->d_results); cudaFreeHost(c->h_results);
  cudaFree(c->d_xmax); c->posf.release();
  cudaFree(c->d_ticket); cudaFree(c->d_flags); cudaFreeHost(c->h_flags);
  cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1);
  cudaStreamDestroy(c->stream);
  delete c;
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- tables

TabView imdcu_tab_view(const DevTable &t)
{
  TabView v;
  v.coef = t.coef; v.begin = t.begin; v.end = t.end; v.invstep = t.invstep; v.step = t.step;
  v.raw = t.raw; v.raw2 = t.raw2;
  v.ncols = t.ncols; v.nrows = t.nrows;
  return v;
}

extern "C" int imdcu_set_table(imdcu_ctx *c, int which, int ncols, int maxsteps,
                               const double *begin, const double *end, const double *step,
                               const double *invstep, const int *len, const double *table)
{
  if (!c || which < 0 || which > 2 || ncols < 1 || maxsteps < 3 || !begin || !end || !step ||
      !invstep || !len || !table)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_set_table: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  DevTable &t = c->tab[which];
  if (c->fused_coef) { cudaFree(c->fused_coef); c->fused_coef = nullptr; }
  c->fused_state = 0;
  if (c->rep4_tab) { cudaFree(c->rep4_tab); c->rep4_tab = nullptr; }
  c->rep4_state = 0;
  if (t.coef) { cudaFree(t.coef); t.coef = nullptr; }
  if (t.begin) { cudaFree(t.begin); t.begin = nullptr; }
  if (t.raw) { cudaFree(t.raw); t.raw = nullptr; }
  if (t.raw2) { cudaFree(t.raw2); t.raw2 = nullptr; }
  t.ncols = ncols; t.nrows = maxsteps; t.maxsteps = maxsteps;
  t.h_begin.assign(begin, begin + ncols); t.h_end.assign(end, end + ncols);
  t.h_step.assign(step, step + ncols); t.h_invstep.assign(invstep, invstep + ncols);
  t.h_len.assign(len, len + ncols);
  t.h_table.assign(table, table + (size_t) (maxsteps + 2) * ncols);
  // coefficient rows: same operations as the macro body (potaccess.h:344-349)
  std::vector<double> coef((size_t) ncols * maxsteps * 3);
  for (int col = 0; col < ncols; col++)
    for (int k = 0; k < maxsteps; k++) {
      volatile double p0 = table[(size_t) k * ncols + col];
      volatile double p1 = table[(size_t) (k + 1) * ncols + col];
      volatile double p2 = table[(size_t) (k + 2) * ncols + col];
      volatile double dv = p1 - p0;
      volatile double two_p1 = 2 * p1;
      volatile double t1 = p2 - two_p1;
      volatile double d2v = t1 + p0;
      double *o = &coef[((size_t) col * maxsteps + k) * 3];
      o[0] = p0; o[1] = dv; o[2] = d2v;
    }
  CU_CHECK(c, cudaMalloc(&t.coef, coef.size() * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.coef, coef.data(), coef.size() * sizeof(double)));
  std::vector<double> info(4 * (size_t) ncols);
  for (int i = 0; i < ncols; i++) {
    info[i] = begin[i]; info[ncols + i] = end[i]; info[2 * ncols + i] = invstep[i]; info[3 * ncols + i] = step[i];
  }
  CU_CHECK(c, cudaMalloc(&t.begin, info.size() * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.begin, info.data(), info.size() * sizeof(double)));
  t.end = t.begin + ncols; t.invstep = t.begin + 2 * ncols; t.step = t.begin + 3 * ncols;
  if (c->interp != IMDCU_INTERP_3POINT) {      // these modes read the table as the reference does
    CU_CHECK(c, cudaMalloc(&t.raw, t.h_table.size() * sizeof(double)));
    CU_CHECK(c, imdcu_h2d(c, t.raw, t.h_table.data(), t.h_table.size() * sizeof(double)));
  }
  t.set = true;
  return IMDCU_OK;
}

extern "C" int imdcu_set_interp(imdcu_ctx *c, int mode)
{
  if (!c || mode < IMDCU_INTERP_3POINT || mode > IMDCU_INTERP_SPLINE)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_set_interp: bad mode") : IMDCU_EINVAL;
  for (int t = 0; t < 3; t++)
    if (c->tab[t].set && mode != c->interp)
      return c->fail(IMDCU_EINVAL, "imdcu_set_interp: call before imdcu_set_table");
  c->interp = mode;
  return IMDCU_OK;
}

extern "C" int imdcu_set_table2(imdcu_ctx *c, int which, const double *table2)
{
  if (!c || which < 0 || which > 2 || !table2)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_set_table2: bad argument") : IMDCU_EINVAL;
  DevTable &t = c->tab[which];
  if (!t.set || c->interp != IMDCU_INTERP_SPLINE)
    return c->fail(IMDCU_EINVAL, "imdcu_set_table2: spline mode and imdcu_set_table first");
  cudaSetDevice(c->device);
  if (t.raw2) { cudaFree(t.raw2); t.raw2 = nullptr; }
  const size_t n = (size_t) (t.maxsteps + 2) * t.ncols;
  CU_CHECK(c, cudaMalloc(&t.raw2, n * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.raw2, table2, n * sizeof(double)));
  return IMDCU_OK;
}

// pair and rho rows interleaved, {p0,dv,d2v | r0,drv,d2rv} per r^2 interval, when both tables are
// single-column and tabulated on the same grid (k_eam_pass1_fused); nullptr otherwise
const double *imdcu_fused_table(imdcu_ctx *c)
{
  if (!c->tune.fused_table) return nullptr;
  if (c->fused_state != 0) return c->fused_state > 0 ? c->fused_coef : nullptr;
  c->fused_state = -1;
  const bool off = !c->tune.fused_table;
  const DevTable &p = c->tab[IMDCU_TAB_PAIR], &r = c->tab[IMDCU_TAB_RHO];
  if (off || !p.set || !r.set || p.ncols != 1 || r.ncols != 1 || p.nrows != r.nrows) return nullptr;
  if (p.h_begin[0] != r.h_begin[0] || p.h_end[0] != r.h_end[0] || p.h_invstep[0] != r.h_invstep[0] ||
      p.h_len[0] != r.h_len[0])
    return nullptr;
  const size_t n = (size_t) p.nrows;
  std::vector<double> pc(n * 3), rc(n * 3), f(n * 6);
  if (cudaMemcpy(pc.data(), p.coef, n * 3 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return nullptr;
  if (cudaMemcpy(rc.data(), r.coef, n * 3 * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) return nullptr;
  for (size_t k = 0; k < n; k++)
    for (int q = 0; q < 3; q++) { f[6 * k + q] = pc[3 * k + q]; f[6 * k + 3 + q] = rc[3 * k + q]; }
  if (cudaMalloc(&c->fused_coef, f.size() * sizeof(double)) != cudaSuccess) { c->fused_coef = nullptr; return nullptr; }
  if (imdcu_h2d(c, c->fused_coef, f.data(), f.size() * sizeof(double)) != cudaSuccess) return nullptr;
  c->fused_state = 1;
  return c->fused_coef;
}

// Raw pair and rho values {pair[k], rho[k]} as 16-byte rows, FOUR copies interleaved in every 128-byte line:
// row k of copy r sits at 16-byte slot 8 (k >> 1) + 2 r + (k & 1).  A lane reads copy (lane & 3), so the shared-memory
// bank group of its three loads (rows k, k+1, k+2 of PAIR_INT2, potaccess.h:344-346) is fixed by the lane and the
// parity of k instead of being random in k: two lanes of a quarter warp share a copy and collide only when their
// rows have the same parity (k_eam_pass1_rep4).  Same conditions as imdcu_fused_table; *bytes = size to stage.
const double *imdcu_rep4_table(imdcu_ctx *c, size_t *bytes)
{
  *bytes = c->rep4_bytes;
  if (c->rep4_state != 0) return c->rep4_state > 0 ? c->rep4_tab : nullptr;
  c->rep4_state = -1;
  const DevTable &p = c->tab[IMDCU_TAB_PAIR], &r = c->tab[IMDCU_TAB_RHO];
  if (!p.set || !r.set || p.ncols != 1 || r.ncols != 1 || p.nrows != r.nrows) return nullptr;
  if (p.h_begin[0] != r.h_begin[0] || p.h_end[0] != r.h_end[0] || p.h_invstep[0] != r.h_invstep[0] ||
      p.h_len[0] != r.h_len[0])
    return nullptr;
  const size_t rows = (size_t) p.maxsteps + 2, lines = (rows + 1) / 2;
  std::vector<double> f(lines * 16, 0.0);
  for (size_t k = 0; k < rows; k++)
    for (int rep = 0; rep < 4; rep++) {
      const size_t slot = 8 * (k >> 1) + 2 * rep + (k & 1);
      f[2 * slot] = p.h_table[k]; f[2 * slot + 1] = r.h_table[k];
    }
  if (cudaMalloc(&c->rep4_tab, f.size() * sizeof(double)) != cudaSuccess) { c->rep4_tab = nullptr; return nullptr; }
  if (imdcu_h2d(c, c->rep4_tab, f.data(), f.size() * sizeof(double)) != cudaSuccess) return nullptr;
  c->rep4_bytes = f.size() * sizeof(double);
  *bytes = c->rep4_bytes;
  c->rep4_state = 1;
  return c->rep4_tab;
}

// Diagnostic (IMDCU_DEBUG_REBUILD=1): time since the previous mark, after synchronising the stream, so that a rebuild
// can be attributed to its parts on every rank; label == nullptr starts a new line.
void imdcu_dbg_mark(imdcu_ctx *c, const char *label, bool flush)
{
  if (!c->dbg_rebuild) return;
  cudaStreamSynchronize(c->stream);
  const auto now = std::chrono::steady_clock::now();
  if (label) {
    char buf[96];
    snprintf(buf, sizeof buf, " %s %.3f", label, std::chrono::duration<double, std::milli>(now - c->dbg_t).count());
    c->dbg_line += buf;
  } else c->dbg_line.clear();
  c->dbg_t = std::chrono::steady_clock::now();
  if (flush) {
    // IMDCU_DEBUG_REBUILD=<prefix>: one file per rank (<prefix>.rank<r>), otherwise stderr (the ranks of a node share it)
    const char *pre = getenv("IMDCU_DEBUG_REBUILD");
    FILE *f = stderr;
    if (pre && pre[0] && strcmp(pre, "1") != 0) {
      const std::string path = std::string(pre) + ".rank" + std::to_string(c->comm.rank);
      f = fopen(path.c_str(), "a");
      if (!f) f = stderr;
    }
    fprintf(f, "REBUILD rank %d n %lld ng %lld cap %d:%s\n", c->comm.rank, (long long) c->n, (long long) c->ng, c->nbl_cap,
            c->dbg_line.c_str());
    if (f != stderr) fclose(f);
    c->dbg_line.clear();
  }
}

// ----------------------------------------------------------------------------- tuning
extern "C" int imdcu_set_tuning(imdcu_ctx *c, const char *key, int value)
{
  if (!c || !key) return IMDCU_EINVAL;
  const std::string k(key);
  if (k == "lpa") {
    if (value != 0 && value != 1 && value != 2 && value != 4 && value != 8) return c->fail(IMDCU_EINVAL, "lpa: 0, 1, 2, 4 or 8");
    c->tune.lpa = value;
  } else if (k == "fused_table") {
    if (value < 0 || value > 2) return c->fail(IMDCU_EINVAL, "fused_table: 0 (general kernel), 1 (48-byte rows), 2 (raw rows, four bank-staggered replicas)");
    c->tune.fused_table = value;
  }
  else if (k == "nbl_filter") c->tune.nbl_filter = value != 0;
  else if (k == "rep4_block") {
    if (value != 896 && value != 768) return c->fail(IMDCU_EINVAL, "rep4_block: 896 or 768");
    c->tune.rep4_block = value;
  }
  else if (k == "ttm_overlap") c->tune.ttm_overlap = value != 0;
  else if (k == "ttm_lat_smem") c->tune.ttm_lat_smem = value != 0;
  else if (k == "ttm3d_tiled") c->tune.ttm3d_tiled = value != 0;
  else if (k == "ttm3d_variant") c->tune.ttm3d_variant = value;
  else return c->fail(IMDCU_EINVAL, "imdcu_set_tuning: unknown key " + k);
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- params

extern "C" int imdcu_set_params(imdcu_ctx *c, int ntypes, double timestep, int nvtypes,
                                const double *restrictions)
{
  if (!c || ntypes < 1 || nvtypes < 1)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_set_params: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  c->ntypes = ntypes; c->timestep = timestep; c->nvtypes = nvtypes;
  c->restrictions.assign(3 * (size_t) nvtypes, 1.0);
  if (restrictions) c->restrictions.assign(restrictions, restrictions + 3 * (size_t) nvtypes);
  CU_CHECK(c, c->d_restrictions.ensure(c->restrictions.size()));
  CU_CHECK(c, imdcu_h2d(c, c->d_restrictions.p, c->restrictions.data(),
                         c->restrictions.size() * sizeof(double)));
  c->params_set = true;
  return IMDCU_OK;
}

extern "C" int imdcu_set_pdecay(imdcu_ctx *c, int enabled, double xipdecay, double ramp_start,
                                double ramp_end)
{
  if (!c) return IMDCU_EINVAL;
  c->pdecay = enabled; c->xipdecay = xipdecay; c->ramp_start = ramp_start; c->ramp_end = ramp_end;
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- box / cells

static void cross3(const double *u, const double *v, double *w)
{
  w[0] = u[1] * v[2] - u[2] * v[1];
  w[1] = u[2] * v[0] - u[0] * v[2];
  w[2] = u[0] * v[1] - u[1] * v[0];
}
static double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

static inline int eidx(const Geom &g, int i, int j, int k) { return (i * g.edim[1] + j) * g.edim[2] + k; }

// stage order z, y, x as send_cells (imd_comm_force_3d.c:270-396); unlike the reference's
// AR scheme both x sides are filled because the device list is a full list.
static void build_ghost_plan(imdcu_ctx *c)
{
  Geom &g = c->geom;
  const int ex = g.edim[0], ey = g.edim[1], ez = g.edim[2];
  auto code = [](int sx, int sy, int sz) { return (sx + 1) + 3 * (sy + 1) + 9 * (sz + 1); };
  c->gplan.clear();
  c->gstage_cells[0] = 0;
  if (g.pbc[2])
    for (int i = 1; i < ex - 1; i++)
      for (int j = 1; j < ey - 1; j++) {
        c->gplan.push_back({eidx(g, i, j, ez - 1), eidx(g, i, j, 1), code(0, 0, +1)});
        c->gplan.push_back({eidx(g, i, j, 0), eidx(g, i, j, ez - 2), code(0, 0, -1)});
      }
  c->gstage_cells[1] = (int) c->gplan.size();
  if (g.pbc[1])
    for (int i = 1; i < ex - 1; i++)
      for (int k = 0; k < ez; k++) {
        c->gplan.push_back({eidx(g, i, ey - 1, k), eidx(g, i, 1, k), code(0, +1, 0)});
        c->gplan.push_back({eidx(g, i, 0, k), eidx(g, i, ey - 2, k), code(0, -1, 0)});
      }
  c->gstage_cells[2] = (int) c->gplan.size();
  if (c->comm.nranks > 1) {
    // decomposed along x: the two ghost planes are filled from the neighbour ranks (multi.cu);
    // src = -1 marks a remote plane; each plane is one contiguous block of plan entries
    imdcu_ctx::Comm &m = c->comm;
    m.nplane = ey * ez;
    const int N = m.nranks, r = m.rank;
    m.east = r + 1 < N ? r + 1 : (g.pbc[0] ? 0 : -1);
    m.west = r - 1 >= 0 ? r - 1 : (g.pbc[0] ? N - 1 : -1);
    m.wrap_e = (r == N - 1 && g.pbc[0]) ? 1 : 0;
    m.wrap_w = (r == 0 && g.pbc[0]) ? 1 : 0;
    m.q_east = (int) c->gplan.size();
    if (m.east >= 0)
      for (int j = 0; j < ey; j++)
        for (int k = 0; k < ez; k++) c->gplan.push_back({eidx(g, ex - 1, j, k), -1, code(m.wrap_e ? +1 : 0, 0, 0)});
    m.q_west = (int) c->gplan.size();
    if (m.west >= 0)
      for (int j = 0; j < ey; j++)
        for (int k = 0; k < ez; k++) c->gplan.push_back({eidx(g, 0, j, k), -1, code(m.wrap_w ? -1 : 0, 0, 0)});
  } else if (g.pbc[0])
    for (int j = 0; j < ey; j++)
      for (int k = 0; k < ez; k++) {
        c->gplan.push_back({eidx(g, ex - 1, j, k), eidx(g, 1, j, k), code(+1, 0, 0)});
        c->gplan.push_back({eidx(g, 0, j, k), eidx(g, ex - 2, j, k), code(-1, 0, 0)});
      }
  c->gstage_cells[3] = (int) c->gplan.size();
}

static int apply_slabs(imdcu_ctx *c);

extern "C" int imdcu_set_box(imdcu_ctx *c, const double box[9], const int pbc[3],
                             double nbl_margin, double nbl_size, const int divisor[3])
{
  if (!c || !box || !pbc) return c ? c->fail(IMDCU_EINVAL, "imdcu_set_box: bad argument") : IMDCU_EINVAL;
  if (!c->tab[IMDCU_TAB_PAIR].set || !c->tab[IMDCU_TAB_RHO].set)
    return c->fail(IMDCU_EINVAL, "imdcu_set_box: set the pair and rho tables first (cellsz comes from them)");
  cudaSetDevice(c->device);
  Geom &g = c->geom;
  memcpy(g.box, box, sizeof g.box);
  for (int d = 0; d < 3; d++) g.pbc[d] = pbc[d];
  // make_box (imd_geom_3d.c:61-84)
  cross3(box + 3, box + 6, g.tbox + 0);
  cross3(box + 6, box + 0, g.tbox + 3);
  cross3(box + 0, box + 3, g.tbox + 6);
  volatile double volume = dot3(box, g.tbox);
  if (volume == 0) return c->fail(IMDCU_EINVAL, "Box Edges are parallel.");
  for (int k = 0; k < 9; k++) g.tbox[k] /= volume;
  double height[3];
  for (int d = 0; d < 3; d++) height[d] = 1.0 / dot3(g.tbox + 3 * d, g.tbox + 3 * d);
  // cellsz: max r2 end of the radial tables (imd_potential.c:406) + margin (imd_geom_3d.c:129-131)
  double cellsz = 0.0;
  for (int t = 0; t < 2; t++)
    for (int i = 0; i < c->tab[t].ncols; i++) cellsz = std::max(cellsz, c->tab[t].h_end[i]);
  volatile double rc = sqrt((double) cellsz) + nbl_margin;
  g.cellsz = rc * rc;
  c->nbl_margin = nbl_margin;
  c->nbl_size = nbl_size > 1.0 ? nbl_size : 1.3;
  for (int d = 0; d < 3; d++) {
    int cd = divisor ? divisor[d] : 1;
    if (cd < 1) cd = 1;
    if (d == 0) cd *= c->comm.nranks;            // cpu_dim.x of the reference (imd_geom_3d.c:168-176)
    volatile double scale = sqrt(1.0 * g.cellsz / height[d]);
    g.gdim[d] = (int) (1.0 / scale);
    if (0 != g.gdim[d] % cd) g.gdim[d] = ((int) (g.gdim[d] / cd)) * cd;
    if (g.gdim[d] < cd || g.gdim[d] < 1)
      return c->fail(IMDCU_EINVAL, "global_cell_dim too small for the cutoff (imd_geom_3d.c:186-199)");
    g.ldim[d] = g.gdim[d];
    g.edim[d] = g.ldim[d] + 2;
  }
  // x slabs of equal cell-layer count, rank r owns [r*l, (r+1)*l) (imd_geom_mpi_3d.c:119-128); imdcu_comm_set_cuts
  // replaces them by slabs of equal atom count
  c->xcut.assign((size_t) c->comm.nranks + 1, 0);
  for (int r = 0; r <= c->comm.nranks; r++) c->xcut[r] = (int) ((long long) r * g.gdim[0] / c->comm.nranks);
  return apply_slabs(c);
}

// this rank's slab [xcut[rank], xcut[rank+1]) of global x cell layers -> local cell grid, ghost plan, cell arrays
static int apply_slabs(imdcu_ctx *c)
{
  Geom &g = c->geom;
  g.xoff = 0;
  g.ldim[0] = g.gdim[0];
  if (c->comm.nranks > 1) {
    g.xoff = c->xcut[c->comm.rank];
    g.ldim[0] = c->xcut[c->comm.rank + 1] - g.xoff;
    if (g.ldim[0] < 1) return c->fail(IMDCU_EINVAL, "x slab too thin");
  }
  g.edim[0] = g.ldim[0] + 2;
  c->ncell_ext = g.edim[0] * g.edim[1] * g.edim[2];
  build_ghost_plan(c);
  CU_CHECK(c, c->cell_begin.ensure(c->ncell_ext + 1));
  CU_CHECK(c, c->cell_count.ensure(c->ncell_ext + 1));
  CU_CHECK(c, c->d_gplan.ensure(c->gplan.size() + 1));
  CU_CHECK(c, c->gcell_off.ensure(c->gplan.size() + 2));
  if (!c->gplan.empty())
    CU_CHECK(c, imdcu_h2d(c, c->d_gplan.p, c->gplan.data(), c->gplan.size() * sizeof(GhostCell)));
  c->box_set = true;
  c->have_valid_nbl = false;
  return IMDCU_OK;
}

// Slab boundaries by atom count (the static part of what the reference's load balancer does for its
// x direction, imd_main_3d.c:864-914): cut r is put on the cell-layer boundary where the running atom count comes
// closest to r/nranks of the total, every slab keeping at least min_layers layers.  Pure host arithmetic, every
// rank computes the same cuts from the same (all-reduced) histogram.
extern "C" int imdcu_balance_cuts(int nranks, int gdimx, const int64_t *atoms_per_layer, int min_layers, int *xcut)
{
  if (nranks < 1 || gdimx < 1 || !atoms_per_layer || !xcut) return IMDCU_EINVAL;
  if (min_layers < 1) min_layers = 1;
  if ((long long) nranks * min_layers > gdimx) return IMDCU_EINVAL;
  std::vector<long long> pre((size_t) gdimx + 1, 0);
  for (int l = 0; l < gdimx; l++) {
    if (atoms_per_layer[l] < 0) return IMDCU_EINVAL;
    pre[l + 1] = pre[l] + atoms_per_layer[l];
  }
  const long long total = pre[gdimx];
  xcut[0] = 0; xcut[nranks] = gdimx;
  for (int r = 1; r < nranks; r++) {
    const int lo = xcut[r - 1] + min_layers, hi = gdimx - (nranks - r) * min_layers;
    int best = lo;
    // |pre[cut] * nranks - r * total| is minimal; the first such layer boundary on ties
    long long bestd = -1;
    for (int cut = lo; cut <= hi; cut++) {
      long long d = pre[cut] * nranks - (long long) r * total;
      if (d < 0) d = -d;
      if (bestd < 0 || d < bestd) { bestd = d; best = cut; }
    }
    xcut[r] = best;
  }
  return IMDCU_OK;
}

// Use these slab boundaries (global x cell layers, xcut[0] = 0 < ... < xcut[nranks] = global_cell_dim.x) instead of
// equal ones.  After imdcu_set_box, before the atoms are uploaded; the same array on every rank.
extern "C" int imdcu_comm_set_cuts(imdcu_ctx *c, const int *xcut)
{
  if (!c || !xcut) return c ? c->fail(IMDCU_EINVAL, "imdcu_comm_set_cuts: bad argument") : IMDCU_EINVAL;
  if (!c->box_set) return c->fail(IMDCU_EINVAL, "imdcu_comm_set_cuts: call imdcu_set_box first");
  if (c->n > 0) return c->fail(IMDCU_EINVAL, "imdcu_comm_set_cuts: atoms are already uploaded");
  const int nr = c->comm.nranks;
  if (xcut[0] != 0 || xcut[nr] != c->geom.gdim[0]) return c->fail(IMDCU_EINVAL, "imdcu_comm_set_cuts: xcut[0] = 0 and xcut[nranks] = global_cell_dim.x required");
  for (int r = 0; r < nr; r++)
    if (xcut[r + 1] <= xcut[r]) return c->fail(IMDCU_EINVAL, "imdcu_comm_set_cuts: every slab needs at least one cell layer");
  cudaSetDevice(c->device);
  c->xcut.assign(xcut, xcut + nr + 1);
  return apply_slabs(c);
}

extern "C" int imdcu_get_cell_grid(const imdcu_ctx *c, int gdim[3], double *cellsz)
{
  if (!c || !c->box_set) return IMDCU_EINVAL;
  for (int d = 0; d < 3; d++) gdim[d] = c->geom.gdim[d];
  if (cellsz) *cellsz = c->geom.cellsz;
  return IMDCU_OK;
}
