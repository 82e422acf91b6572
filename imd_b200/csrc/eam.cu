// imd_b200/csrc/eam.cu -- the three EAM2 passes of calc_forces (imd_forces_nbl.c:419-985)
// as two persistent FP64 kernels over the full neighbour list.
//
//   k_eam_pass1 : pair energy/force (PAIR_INT on pair_pot, :485-507), host density
//                 (VAL_FUNC on rho_h_tab, :557-572), TTM neighbour count (:473-481) and --
//                 because a full list completes rho_i inside the thread -- the embedding
//                 step F(rho), F'(rho) (:748-783) fused into the tail.
//   k_eam_pass2 : embedding forces 1/2 (F'_i rho'_j + F'_j rho'_i) d   (:790-985)
//
// What the reference does with scatter (KRAFT(q,j) -= ...) and three reverse/forward halo
// exchanges (send_forces(add_rho), send_cells(copy_dF), send_forces(add_forces)) is done
// here by gathers only: image atoms are read-only, so only positions and F' travel.
//
// Layout: pos4 = {x,y,z,F'} one 32-byte sector per gathered neighbour; neighbour slots are
// read coalesced (nblist.cu); coefficient tables {p0,dv,d2v} are staged into shared memory
// by 1-D bulk TMA copies (cp.async.bulk + mbarrier) once per persistent CTA.
// Bound: FP64 pipe (about 45 DP instructions per pair and pass), not HBM; no tensor cores
// (nothing here is a dense contraction).
#include "device.cuh"

#ifndef EAM_BLOCK
#define EAM_BLOCK 512
#endif
#ifndef EAM_MINB
#define EAM_MINB 2
#endif
#ifndef EAM_UNROLL
#define EAM_UNROLL 2
#endif
#define EAM_PRAGMA_(x) _Pragma(#x)
#define EAM_PRAGMA_UNROLL(n) EAM_PRAGMA_(unroll n)

struct EamArgs {
  int n, ntypes, cap;
  double4 *pos4;                 // [n+ng] x,y,z,dF
  const int *type;               // [n+ng]
  const int *nbl_idx, *nbl_num;
  double4 *frc4;                 // [n] fx,fy,fz,epot
  double *rho;                   // [n]
  double *dF;                    // [n] F'(rho) out of pass 1
  int *numneighs;                // [n]
  TabView pair, rhot, emb;
  const double *fused;           // [nrows][6] {p0,dv,d2v | r0,drv,d2rv}: pair + rho rows on a common r^2 grid, or null
  double *partials; unsigned int *ticket; double *results; int *flags;
  int count_ttm;                 // fill numneighs
};

template <int LPA>
__device__ __forceinline__ double sub_sum(double v)
{
#pragma unroll
  for (int o = LPA / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// LPA lanes cooperate on one atom (1 for big systems, more to fill the GPU on small ones);
// SMEM: tables staged in shared memory, else read through L1/L2; MULTI: ntypes > 1.
template <int LPA, bool SMEM, bool MULTI>
__global__ void __launch_bounds__(EAM_BLOCK, EAM_MINB)
k_eam_pass1(EamArgs a)
{
  extern __shared__ __align__(16) double s_tab[];
  __shared__ __align__(8) unsigned long long s_bar;
  const double *pair_c = a.pair.coef, *rho_c = a.rhot.coef;
  if (SMEM) {
    const size_t pb = (size_t) a.pair.ncols * a.pair.nrows * 3 * sizeof(double);
    const size_t rb = (size_t) a.rhot.ncols * a.rhot.nrows * 3 * sizeof(double);
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, (unsigned) (pb + rb));
      for (size_t off = 0; off < pb; off += 65536)
        bulk_g2s((char *) s_tab + off, (const char *) a.pair.coef + off,
                 (unsigned) (pb - off < 65536 ? pb - off : 65536), &s_bar);
      for (size_t off = 0; off < rb; off += 65536)
        bulk_g2s((char *) s_tab + pb + off, (const char *) a.rhot.coef + off,
                 (unsigned) (rb - off < 65536 ? rb - off : 65536), &s_bar);
    }
    mbar_wait(&s_bar, 0);
    pair_c = s_tab;
    rho_c = s_tab + (size_t) a.pair.ncols * a.pair.nrows * 3;
  }
  TabCol tp0 = tab_col(a.pair, pair_c, 0), tr0 = tab_col(a.rhot, rho_c, 0);
  const double pair_end0 = a.pair.end[0];

  const int apb = EAM_BLOCK / LPA;                       // atoms per CTA iteration
  const int sub = threadIdx.x % LPA;
  double acc[2] = {0.0, 0.0};                            // epot, virial
  int is_short = 0;

  for (int base = blockIdx.x * apb; base < a.n; base += gridDim.x * apb) {
    const int i = base + threadIdx.x / LPA;
    const bool live = i < a.n;
    double fx = 0, fy = 0, fz = 0, ee = 0, vir = 0, rh = 0;
    int nn = 0, it = 0;
    if (live) {
      const double4 pi = a.pos4[i];
      const int num = a.nbl_num[i];
      const int *slot = a.nbl_idx + ((size_t) (i >> 5) * a.cap) * 32 + (i & 31);
      if (MULTI) it = a.type[i];
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = sub; k < num; k += LPA) {
        const int j = slot[(size_t) k * 32];
        const double4 pj = ld256(&a.pos4[j]);
        const double dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
        const double r2 = r2_exact(dx, dy, dz);
        TabCol tp = tp0, tr = tr0;
        int jt = 0;
        if (MULTI) {
          jt = a.type[j];
          const int col = it * a.ntypes + jt;
          tp = tab_col(a.pair, pair_c, col);
          tr = tab_col(a.rhot, rho_c, col);
        }
        if (a.count_ttm && r2 <= pair_end0 && it == 0 && jt == 0) nn++;
        if (r2 <= tp.end) {
          double pot, grad;
          pair_int2(tp, r2, pot, grad, is_short);
          fx += dx * grad; fy += dy * grad; fz += dz * grad;
          ee += 0.5 * pot;
          vir -= 0.5 * (r2 * grad);
        }
        if (r2 < tr.end) rh += val_func2(tr, r2, is_short);
      }
    }
    if (LPA > 1) {
      fx = sub_sum<LPA>(fx); fy = sub_sum<LPA>(fy); fz = sub_sum<LPA>(fz);
      ee = sub_sum<LPA>(ee); vir = sub_sum<LPA>(vir); rh = sub_sum<LPA>(rh);
      nn = (int) sub_sum<LPA>((double) nn);
    }
    if (live && sub == 0) {
      // embedding energy and its derivative (imd_forces_nbl.c:758-760); table via L1/L2
      double pot, dF; int idummy = 0;
      TabCol te = tab_col(a.emb, a.emb.coef, it);
      pair_int2(te, rh, pot, dF, idummy);
      ee += pot;
      a.frc4[i] = make_double4(fx, fy, fz, ee);
      a.rho[i] = rh;
      a.dF[i] = dF;                // merged into pos4.w by k_set_dF before the image refresh
      if (a.count_ttm) a.numneighs[i] = nn;
      acc[0] += ee; acc[1] += vir;
    }
  }
  if (is_short) a.flags[2] = 1;
  grid_reduce<2, 0>(acc, a.partials, a.ticket, a.results);
}

// Pass 1 for one species whose pair and rho tables share their r^2 grid (begin, end, step, length --
// the usual case: both files are tabulated together).  One index computation and ONE 48-byte row
// {p0,dv,d2v | r0,drv,d2rv} per pair, read as three 16-byte shared-memory loads, instead of two
// 24-byte rows read as six 8-byte loads: the kernel is bound by the L1 data pipe, and this is a
// third fewer shared-memory wavefronts.  Same operations per function as PAIR_INT2 / VAL_FUNC2.
template <int LPA>
__global__ void __launch_bounds__(EAM_BLOCK, EAM_MINB)
k_eam_pass1_fused(EamArgs a)
{
  extern __shared__ __align__(16) double s_tab[];
  __shared__ __align__(8) unsigned long long s_bar;
  {
    const size_t fb = (size_t) a.pair.nrows * 6 * sizeof(double);
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, (unsigned) fb);
      for (size_t off = 0; off < fb; off += 65536)
        bulk_g2s((char *) s_tab + off, (const char *) a.fused + off, (unsigned) (fb - off < 65536 ? fb - off : 65536), &s_bar);
    }
    mbar_wait(&s_bar, 0);
  }
  const TabCol tp = tab_col(a.pair, s_tab, 0);           // begin, end, invstep (identical for the rho column)
  const double rho_end = a.rhot.end[0];
  const double two_invstep = 2 * tp.invstep;
  const int apb = EAM_BLOCK / LPA;
  const int sub = threadIdx.x % LPA;
  double acc[2] = {0.0, 0.0};
  int is_short = 0;
  for (int base = blockIdx.x * apb; base < a.n; base += gridDim.x * apb) {
    const int i = base + threadIdx.x / LPA;
    const bool live = i < a.n;
    double fx = 0, fy = 0, fz = 0, ee = 0, vir = 0, rh = 0;
    int nn = 0;
    if (live) {
      const double4 pi = a.pos4[i];
      const int num = a.nbl_num[i];
      const int *slot = a.nbl_idx + ((size_t) (i >> 5) * a.cap) * 32 + (i & 31);
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = sub; k < num; k += LPA) {
        const int j = slot[(size_t) k * 32];
        const double4 pj = ld256(&a.pos4[j]);
        const double dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
        const double r2 = r2_exact(dx, dy, dz);
        if (r2 <= tp.end) {                                // pair: r2 <= end, rho: r2 < end (same end)
          if (a.count_ttm) nn++;
          int kk; double chi;
          tab_index(tp, r2, kk, chi, is_short);
          const double2 *row = reinterpret_cast<const double2 *>(s_tab + 6 * (size_t) kk);
          const double2 q0 = row[0], q1 = row[1], q2 = row[2];   // {p0,dv} {d2v,r0} {drv,d2rv}
          const double h = 0.5 * chi * (chi - 1);
          const double pot = q0.x + chi * q0.y + h * q1.x;
          const double grad = two_invstep * (q0.y + (chi - 0.5) * q1.x);
          fx += dx * grad; fy += dy * grad; fz += dz * grad;
          ee += 0.5 * pot;
          vir -= 0.5 * (r2 * grad);
          if (r2 < rho_end) rh += q1.y + chi * q2.x + h * q2.y;
        }
      }
    }
    if (LPA > 1) {
      fx = sub_sum<LPA>(fx); fy = sub_sum<LPA>(fy); fz = sub_sum<LPA>(fz);
      ee = sub_sum<LPA>(ee); vir = sub_sum<LPA>(vir); rh = sub_sum<LPA>(rh);
      nn = (int) sub_sum<LPA>((double) nn);
    }
    if (live && sub == 0) {
      double pot, dF; int idummy = 0;
      TabCol te = tab_col(a.emb, a.emb.coef, 0);
      pair_int2(te, rh, pot, dF, idummy);
      ee += pot;
      a.frc4[i] = make_double4(fx, fy, fz, ee);
      a.rho[i] = rh;
      a.dF[i] = dF;
      if (a.count_ttm) a.numneighs[i] = nn;
      acc[0] += ee; acc[1] += vir;
    }
  }
  if (is_short) a.flags[2] = 1;
  grid_reduce<2, 0>(acc, a.partials, a.ticket, a.results);
}

template <int LPA, bool SMEM, bool MULTI>
__global__ void __launch_bounds__(EAM_BLOCK, EAM_MINB)
k_eam_pass2(EamArgs a)
{
  extern __shared__ __align__(16) double s_tab[];
  __shared__ __align__(8) unsigned long long s_bar;
  const double *rho_c = a.rhot.coef;
  if (SMEM) {
    const size_t rb = (size_t) a.rhot.ncols * a.rhot.nrows * 3 * sizeof(double);
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, (unsigned) rb);
      for (size_t off = 0; off < rb; off += 65536)
        bulk_g2s((char *) s_tab + off, (const char *) a.rhot.coef + off,
                 (unsigned) (rb - off < 65536 ? rb - off : 65536), &s_bar);
    }
    mbar_wait(&s_bar, 0);
    rho_c = s_tab;
  }
  const TabCol tr0 = tab_col(a.rhot, rho_c, 0);
  const int apb = EAM_BLOCK / LPA;
  const int sub = threadIdx.x % LPA;
  double acc[1] = {0.0};
  int is_short = 0;

  for (int base = blockIdx.x * apb; base < a.n; base += gridDim.x * apb) {
    const int i = base + threadIdx.x / LPA;
    const bool live = i < a.n;
    double fx = 0, fy = 0, fz = 0, vir = 0;
    if (live) {
      const double4 pi = a.pos4[i];                      // .w = F'(rho_i)
      const int num = a.nbl_num[i];
      const int *slot = a.nbl_idx + ((size_t) (i >> 5) * a.cap) * 32 + (i & 31);
      int it = 0;
      if (MULTI) it = a.type[i];
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = sub; k < num; k += LPA) {
        const int j = slot[(size_t) k * 32];
        const double4 pj = ld256(&a.pos4[j]);
        const double dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
        const double r2 = r2_exact(dx, dy, dz);
        if (!MULTI) {
          if (r2 < tr0.end) {
            const double rs = deriv_func2(tr0, r2, is_short);
            const double grad = 0.5 * (pi.w * rs + pj.w * rs);
            const double gx = dx * grad, gy = dy * grad, gz = dz * grad;
            fx += gx; fy += gy; fz += gz;
            vir -= 0.5 * (dx * gx + dy * gy + dz * gz);
          }
        } else {
          const int jt = a.type[j];
          const int col1 = jt * a.ntypes + it, col2 = it * a.ntypes + jt;
          const TabCol t1 = tab_col(a.rhot, rho_c, col1), t2 = tab_col(a.rhot, rho_c, col2);
          if (r2 < t1.end || r2 < t2.end) {
            const double ri = deriv_func2(t1, r2, is_short);
            const double rj = (col1 == col2) ? ri : deriv_func2(t2, r2, is_short);
            const double grad = 0.5 * (pi.w * rj + pj.w * ri);
            const double gx = dx * grad, gy = dy * grad, gz = dz * grad;
            fx += gx; fy += gy; fz += gz;
            vir -= 0.5 * (dx * gx + dy * gy + dz * gz);
          }
        }
      }
    }
    if (LPA > 1) {
      fx = sub_sum<LPA>(fx); fy = sub_sum<LPA>(fy); fz = sub_sum<LPA>(fz); vir = sub_sum<LPA>(vir);
    }
    if (live && sub == 0) {
      double4 f = a.frc4[i];
      f.x += fx; f.y += fy; f.z += fz;
      a.frc4[i] = f;
      acc[0] += vir;
    }
  }
  if (is_short) a.flags[3] = 1;
  grid_reduce<1, 0>(acc, a.partials, a.ticket + 1, a.results + 2);
}

__global__ void __launch_bounds__(256)
k_set_dF(int n, const double *__restrict__ dF, double4 *__restrict__ pos4)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pos4[i].w = dF[i];
}

// ----------------------------------------------------------------------------- dispatch

template <typename K>
static int launch_cfg(imdcu_ctx *c, K kern, size_t smem, int *grid)
{
  if (smem > 48 * 1024)
    CU_CHECK(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
  int per_sm = 0;
  CU_CHECK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, EAM_BLOCK, smem));
  if (per_sm < 1) return c->fail(IMDCU_ECUDA, "EAM kernel does not fit on an SM");
  *grid = c->num_sms * per_sm;
  return IMDCU_OK;
}

template <int LPA, bool SMEM, bool MULTI>
static int run_passes(imdcu_ctx *c, EamArgs &a, size_t smem1, size_t smem2)
{
  int g1 = 0, g2 = 0;
  const int apb = EAM_BLOCK / LPA;
  const bool fused = SMEM && !MULTI && a.fused;
  if (fused) RET_IF(launch_cfg(c, k_eam_pass1_fused<LPA>, smem1, &g1));
  else RET_IF(launch_cfg(c, k_eam_pass1<LPA, SMEM, MULTI>, smem1, &g1));
  RET_IF(launch_cfg(c, k_eam_pass2<LPA, SMEM, MULTI>, smem2, &g2));
  g1 = std::min(g1, div_up(a.n, apb));
  g2 = std::min(g2, div_up(a.n, apb));
  CU_CHECK(c, c->partials.ensure((size_t) std::max(g1, g2) * 4 + 64));
  a.partials = c->partials.p;
  {
    PhaseTimer pt(c, 3);
    if (fused) k_eam_pass1_fused<LPA><<<g1, EAM_BLOCK, smem1, c->stream>>>(a);
    else k_eam_pass1<LPA, SMEM, MULTI><<<g1, EAM_BLOCK, smem1, c->stream>>>(a);
    c->count_launch();
  }
  {
    PhaseTimer pt(c, 4);
    k_set_dF<<<div_up(a.n, 256), 256, 0, c->stream>>>(a.n, a.dF, a.pos4);
    c->count_launch();
  }
  // calc_ttm overlapped with pass 2 (imdcu_run_ttm): numneighs of this step exist now; the lattice
  // kernel goes to a second stream and pass 2 leaves one CTA slot of the machine to it
  const bool fuse = c->fuse_ttm_steps >= 0;
  if (fuse) {
    RET_IF(imdcu_ttm_fork(c, c->fuse_ttm_steps));
    // the lattice CTA gets an SM of its own (ttm.cu: ttm_grid_part), otherwise one CTA slot
    if (g2 == c->num_sms * (g2 / c->num_sms) && g2 > c->num_sms) g2 -= c->ttm_isolate ? g2 / c->num_sms : 1;
  }
  RET_IF(imdcu_refresh_ghosts(c, false));          // F' (and positions again) to the images
  {
    PhaseTimer pt(c, 5);
    k_eam_pass2<LPA, SMEM, MULTI><<<g2, EAM_BLOCK, smem2, c->stream>>>(a);
    c->count_launch();
  }
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}

template <int LPA>
static int run_lpa(imdcu_ctx *c, EamArgs &a)
{
  const size_t pb = c->tab[IMDCU_TAB_PAIR].coef_bytes(), rb = c->tab[IMDCU_TAB_RHO].coef_bytes();
  const bool multi = c->ntypes > 1;
  // two CTAs of 512 threads per SM when the tables allow it (<= ~110 KB per CTA)
  const bool smem_ok = (pb + rb + 1024 <= c->smem_optin) && (pb % 16 == 0) && (rb % 16 == 0);
  if (smem_ok) {
    if (multi) return run_passes<LPA, true, true>(c, a, pb + rb, rb);
    return run_passes<LPA, true, false>(c, a, pb + rb, rb);
  }
  if (multi) return run_passes<LPA, false, true>(c, a, 0, 0);
  return run_passes<LPA, false, false>(c, a, 0, 0);
}

int imdcu_eam_forces(imdcu_ctx *c, double out8[8])
{
  EamArgs a;
  a.n = (int) c->n; a.ntypes = c->ntypes; a.cap = c->nbl_cap;
  a.pos4 = c->pos4.p; a.type = c->type.p;
  a.nbl_idx = c->nbl_idx.p; a.nbl_num = c->nbl_num.p;
  a.frc4 = c->frc4.p; a.rho = c->rho.p; a.dF = c->dF.p; a.numneighs = c->numneighs.p;
  a.pair = imdcu_tab_view(c->tab[IMDCU_TAB_PAIR]);
  a.rhot = imdcu_tab_view(c->tab[IMDCU_TAB_RHO]);
  a.emb = imdcu_tab_view(c->tab[IMDCU_TAB_EMBED]);
  a.fused = imdcu_fused_table(c);
  a.partials = nullptr; a.ticket = c->d_ticket; a.results = c->d_results; a.flags = c->d_flags;
  a.count_ttm = 1;
  if (a.n > 0) {
    // lanes per atom: keep at least ~2 resident waves of threads on the machine
    const int64_t want = (int64_t) c->num_sms * 2048;
    int rc, force_lpa = 0;
    if (const char *e = getenv("IMDCU_LPA")) force_lpa = atoi(e);      // tuning knob
    if (force_lpa == 1) rc = run_lpa<1>(c, a);
    else if (force_lpa == 2) rc = run_lpa<2>(c, a);
    else if (force_lpa == 4) rc = run_lpa<4>(c, a);
    else if (force_lpa == 8) rc = run_lpa<8>(c, a);
    else if (a.n >= want) rc = run_lpa<1>(c, a);
    else if (a.n * 2 >= want) rc = run_lpa<2>(c, a);
    else if (a.n * 4 >= want) rc = run_lpa<4>(c, a);
    else rc = run_lpa<8>(c, a);
    RET_IF(rc);
  } else {
    CU_CHECK(c, cudaMemsetAsync(c->d_results, 0, 4 * sizeof(double), c->stream));
    if (c->fuse_ttm_steps >= 0) RET_IF(imdcu_ttm_fork(c, c->fuse_ttm_steps));
    if (c->comm.nranks > 1) RET_IF(imdcu_refresh_ghosts(c, false));   // the neighbours still expect the F' exchange
  }
  RET_IF(imdcu_multi_allreduce(c, c->d_results, 3, 0));               // MPI_Allreduce of imd_forces_nbl.c:997-1016
  if (out8) {
    CU_CHECK(c, cudaMemcpyAsync(c->h_results, c->d_results, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < 8; k++) out8[k] = 0.0;
    out8[0] = c->h_results[0];                              // tot_pot_energy
    out8[1] = c->h_results[1] + c->h_results[2];            // virial (pair + embedding part)
  }
  return IMDCU_OK;
}
