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

// CTA shapes (measured on a B200, 2.05 M Al atoms, profiles/r01g_eam_variants.txt): ONE CTA per SM with
// one copy of the tables and the rest of the 228 KB as L1 beats two CTAs of 512 threads; pass 1 likes
// 72 registers (896 threads), pass 2 needs no more than 64 (1024 threads).
#ifndef EAM_BLOCK1             /* pass 1 */
#define EAM_BLOCK1 896
#endif
#ifndef EAM_MINB1
#define EAM_MINB1 1
#endif
#ifndef EAM_BLOCK2             /* pass 2 */
#define EAM_BLOCK2 1024
#endif
#ifndef EAM_MINB2
#define EAM_MINB2 1
#endif
#ifndef EAM_UNROLL
#define EAM_UNROLL 2
#endif
#ifndef EAM_IDX_CS
#define EAM_IDX_CS 0           /* 1: list slots with ld.global.cs (streaming, evict first) */
#endif
#ifndef EAM_GATHER_NC
#define EAM_GATHER_NC 0        /* 1: pass 2 gathers through the non-coherent path (ld.global.nc) */
#endif
#if EAM_IDX_CS
#define LD_SLOT(p) __ldcs(p)
#else
#define LD_SLOT(p) (*(p))
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
  const double *rep4;            // raw {pair[k], rho[k]} rows, four bank-staggered copies (context.cu: imdcu_rep4_table), or null
  unsigned rep4_bytes;
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

// Software pipeline over one lane's share of a neighbour row.  A pair costs a chain of two dependent
// loads (list slot from HBM, then the gathered 32-byte record) ahead of ~45 FP64 instructions; written
// as a plain loop the compiler keeps that chain serial per iteration and only the eight warps of a
// scheduler overlap it.  Here the slot indices run three iterations ahead and the record of pair k+1
// is requested before pair k is evaluated.
template <int LPA, bool MULTI, bool NC = false>
struct NbrPipe {
  const int *sp;            // my first slot
  int cnt;                  // my number of slots
  int j2, j3, tn;           // indices of iterations k+1, k+2; type of the record in flight
  double4 pn;               // record in flight
  static constexpr size_t ST = (size_t) LPA * 32;
  __device__ __forceinline__ NbrPipe(const double4 *pos4, const int *type, const int *slot, int num, int sub)
  {
    sp = slot + (size_t) sub * 32;
    cnt = num > sub ? (num - sub + LPA - 1) / LPA : 0;
    const int j1 = cnt > 0 ? LD_SLOT(sp) : 0;
    j2 = cnt > 1 ? LD_SLOT(sp + ST) : 0;
    j3 = cnt > 2 ? LD_SLOT(sp + 2 * ST) : 0;
    tn = 0;
    if (cnt > 0) { pn = rec(pos4, j1); if (MULTI) tn = type[j1]; }
    else pn = make_double4(0, 0, 0, 0);
  }
  __device__ __forceinline__ double4 rec(const double4 *pos4, int j)
  {
    return NC ? ld256nc(&pos4[j]) : ld256(&pos4[j]);
  }
  __device__ __forceinline__ double4 next(const double4 *pos4, const int *type, int k, int &jt)
  {
    const double4 pj = pn;
    if (MULTI) jt = tn;
    if (k + 1 < cnt) { pn = rec(pos4, j2); if (MULTI) tn = type[j2]; }
    j2 = j3;
    if (k + 3 < cnt) j3 = LD_SLOT(sp + (size_t) (k + 3) * ST);
    return pj;
  }
};

// LPA lanes cooperate on one atom (1 for big systems, more to fill the GPU on small ones);
// SMEM: tables staged in shared memory, else read through L1/L2; MULTI: ntypes > 1.
// INTERP: 0 three-point rows (the hot configuration), 1 four-point, 2 spline (raw tables through L1)
template <int LPA, bool SMEM, bool MULTI, int INTERP>
__global__ void __launch_bounds__(EAM_BLOCK1, EAM_MINB1)
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

  const int apb = EAM_BLOCK1 / LPA;                       // atoms per CTA iteration
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
      NbrPipe<LPA, MULTI> np(a.pos4, a.type, slot, num, sub);
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = 0; k < np.cnt; k++) {
        int jt = 0;
        const double4 pj = np.next(a.pos4, a.type, k, jt);
        const double dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
        const double r2 = r2_exact(dx, dy, dz);
        if (a.count_ttm && r2 <= pair_end0 && it == 0 && jt == 0) nn++;
        if (INTERP == 0) {
          TabCol tp = tp0, tr = tr0;
          if (MULTI) {
            const int col = it * a.ntypes + jt;
            tp = tab_col(a.pair, pair_c, col);
            tr = tab_col(a.rhot, rho_c, col);
          }
          if (r2 <= tp.end) {
            double pot, grad;
            pair_int2(tp, r2, pot, grad, is_short);
            fx += dx * grad; fy += dy * grad; fz += dz * grad;
            ee += 0.5 * pot;
            vir -= 0.5 * (r2 * grad);
          }
          if (r2 < tr.end) rh += val_func2(tr, r2, is_short);
        } else {
          const int col = it * a.ntypes + jt;
          const RawCol cp = raw_col(a.pair, col), cr = raw_col(a.rhot, col);
          double pot, grad;
          if (r2 <= cp.end) {
            raw_pair_int<INTERP == 0 ? 1 : INTERP>(cp, r2, pot, grad, is_short);
            fx += dx * grad; fy += dy * grad; fz += dz * grad;
            ee += 0.5 * pot;
            vir -= 0.5 * (r2 * grad);
          }
          if (r2 < cr.end) { raw_pair_int<INTERP == 0 ? 1 : INTERP>(cr, r2, pot, grad, is_short); rh += pot; }
        }
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
      if (INTERP == 0) pair_int2(tab_col(a.emb, a.emb.coef, it), rh, pot, dF, idummy);
      else raw_pair_int<INTERP == 0 ? 1 : INTERP>(raw_col(a.emb, it), rh, pot, dF, idummy);
      ee += pot;
      a.frc4[i] = make_double4(fx, fy, fz, ee);
      a.rho[i] = rh;
      a.dF[i] = dF;
      // Pass 2 gathers F' with the position.  This store overlaps, in time, with other CTAs' 256-bit loads of the same
      // record.  In the PTX memory model a vector access is a set of scalar accesses: those to x, y, z never meet a
      // store during this kernel (nothing writes them), and the one to w meets this store but pass 1 never uses the
      // w it loaded -- old or new, the value is dead.  Pass 2 reads w only after this kernel has finished.
      a.pos4[i].w = dF;
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
__global__ void __launch_bounds__(EAM_BLOCK1, EAM_MINB1)
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
  const int apb = EAM_BLOCK1 / LPA;
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
      NbrPipe<LPA, false> np(a.pos4, a.type, slot, num, sub);
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = 0; k < np.cnt; k++) {
        int jt_unused = 0;
        const double4 pj = np.next(a.pos4, a.type, k, jt_unused);
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
      a.pos4[i].w = dF;            // benign overlap with other CTAs' record loads: see k_eam_pass1
      if (a.count_ttm) a.numneighs[i] = nn;
      acc[0] += ee; acc[1] += vir;
    }
  }
  if (is_short) a.flags[2] = 1;
  grid_reduce<2, 0>(acc, a.partials, a.ticket, a.results);
}

// Pass 1 as k_eam_pass1_fused, but with fewer shared-memory bank conflicts: the table holds the RAW values
// {pair[k], rho[k]} as 16-byte rows in four interleaved copies (imdcu_rep4_table), a lane reads rows k, k+1, k+2 of
// copy (lane & 3) and forms dv = p1 - p0, d2v = p2 - 2 p1 + p0 itself -- the macro's own operations
// (potaccess.h:344-349), so the results equal the precomputed rows bit for bit.  Slot of row k in copy r:
// 8 (k >> 1) + 2 r + (k & 1), i.e. the bank group of a load is 2 r + parity: two lanes of a quarter warp share a copy
// and conflict only when their rows have equal parity, where 48-byte rows at random k collide ~2.1-fold.
template <int LPA, int BLOCK>
__global__ void __launch_bounds__(BLOCK, 1)
k_eam_pass1_rep4(EamArgs a)
{
  extern __shared__ __align__(16) double s_tab[];
  __shared__ __align__(8) unsigned long long s_bar;
  {
    const size_t fb = a.rep4_bytes;
    if (threadIdx.x == 0) mbar_init(&s_bar, 1);
    __syncthreads();
    if (threadIdx.x == 0) {
      mbar_expect_tx(&s_bar, (unsigned) fb);
      for (size_t off = 0; off < fb; off += 65536)
        bulk_g2s((char *) s_tab + off, (const char *) a.rep4 + off, (unsigned) (fb - off < 65536 ? fb - off : 65536), &s_bar);
    }
    mbar_wait(&s_bar, 0);
  }
  const TabCol tp = tab_col(a.pair, s_tab, 0);           // begin, end, invstep (identical for the rho column)
  const double rho_end = a.rhot.end[0];
  const double two_invstep = 2 * tp.invstep;
  const double2 *tab = reinterpret_cast<const double2 *>(s_tab) + 2 * (threadIdx.x & 3);   // my copy
  const int apb = BLOCK / LPA;
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
      NbrPipe<LPA, false> np(a.pos4, a.type, slot, num, sub);
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = 0; k < np.cnt; k++) {
        int jt_unused = 0;
        const double4 pj = np.next(a.pos4, a.type, k, jt_unused);
        const double dx = pj.x - pi.x, dy = pj.y - pi.y, dz = pj.z - pi.z;
        const double r2 = r2_exact(dx, dy, dz);
        if (r2 <= tp.end) {                                // pair: r2 <= end, rho: r2 < end (same end)
          if (a.count_ttm) nn++;
          int kk; double chi;
          tab_index(tp, r2, kk, chi, is_short);
          const int odd = kk & 1;
          const double2 *row = tab + 8 * (kk >> 1) + odd;
          const double2 t0 = row[0], t1 = row[odd ? 7 : 1], t2 = row[8];   // rows kk, kk+1, kk+2: {pair, rho}
          const double dv = t1.x - t0.x, d2v = (t2.x - 2 * t1.x) + t0.x;
          const double h = 0.5 * chi * (chi - 1);
          const double pot = t0.x + chi * dv + h * d2v;
          const double grad = two_invstep * (dv + (chi - 0.5) * d2v);
          fx += dx * grad; fy += dy * grad; fz += dz * grad;
          ee += 0.5 * pot;
          vir -= 0.5 * (r2 * grad);
          if (r2 < rho_end) {
            const double drv = t1.y - t0.y, d2rv = (t2.y - 2 * t1.y) + t0.y;
            rh += t0.y + chi * drv + h * d2rv;
          }
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
      a.pos4[i].w = dF;            // benign overlap with other CTAs' record loads: see k_eam_pass1
      if (a.count_ttm) a.numneighs[i] = nn;
      acc[0] += ee; acc[1] += vir;
    }
  }
  if (is_short) a.flags[2] = 1;
  grid_reduce<2, 0>(acc, a.partials, a.ticket, a.results);
}

template <int LPA, bool SMEM, bool MULTI, int INTERP>
__global__ void __launch_bounds__(EAM_BLOCK2, EAM_MINB2)
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
  const int apb = EAM_BLOCK2 / LPA;
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
      NbrPipe<LPA, MULTI, EAM_GATHER_NC != 0> np(a.pos4, a.type, slot, num, sub);
EAM_PRAGMA_UNROLL(EAM_UNROLL)
      for (int k = 0; k < np.cnt; k++) {
        int jt = 0;
        const double4 pj = np.next(a.pos4, a.type, k, jt);
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
        } else if (INTERP != 0) {
          const int col1 = jt * a.ntypes + it, col2 = it * a.ntypes + jt;
          const RawCol t1 = raw_col(a.rhot, col1), t2 = raw_col(a.rhot, col2);
          if (r2 < t1.end || r2 < t2.end) {
            double v, ri, rj;
            raw_pair_int<INTERP == 0 ? 1 : INTERP>(t1, r2, v, ri, is_short);
            if (col1 == col2) rj = ri; else raw_pair_int<INTERP == 0 ? 1 : INTERP>(t2, r2, v, rj, is_short);
            const double grad = 0.5 * (pi.w * rj + pj.w * ri);
            const double gx = dx * grad, gy = dy * grad, gz = dz * grad;
            fx += gx; fy += gy; fz += gz;
            vir -= 0.5 * (dx * gx + dy * gy + dz * gz);
          }
        } else {
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

// ----------------------------------------------------------------------------- dispatch

template <typename K>
static int launch_cfg(imdcu_ctx *c, K kern, int block, size_t smem, int *grid)
{
  if (smem > 48 * 1024)
    CU_CHECK(c, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
  int per_sm = 0;
  CU_CHECK(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, block, smem));
  if (per_sm < 1) return c->fail(IMDCU_ECUDA, "EAM kernel does not fit on an SM");
  *grid = c->num_sms * per_sm;
  return IMDCU_OK;
}

template <int LPA, bool SMEM, bool MULTI, int INTERP = 0>
static int run_passes(imdcu_ctx *c, EamArgs &a, size_t smem1, size_t smem2)
{
  int g1 = 0, g2 = 0;
  const bool fused = SMEM && !MULTI && INTERP == 0 && a.fused;
  const bool rep4 = fused && a.rep4 && a.rep4_bytes + 1024 <= c->smem_optin;
  // block size of the rep4 kernel: 896 threads leave 72 registers (40 bytes spilled), 768 threads 80 registers (no spill); ptxas budgets registers
  // for the thread count rounded up to a multiple of 128, so nothing in between helps
  const int b4 = (LPA == 1 && c->tune.rep4_block == 768) ? 768 : EAM_BLOCK1;
  if (rep4) {
    smem1 = a.rep4_bytes;
    if (LPA == 1 && b4 == 768) RET_IF(launch_cfg(c, k_eam_pass1_rep4<1, 768>, 768, smem1, &g1));
    else RET_IF(launch_cfg(c, k_eam_pass1_rep4<LPA, EAM_BLOCK1>, EAM_BLOCK1, smem1, &g1));
  }
  else if (fused) RET_IF(launch_cfg(c, k_eam_pass1_fused<LPA>, EAM_BLOCK1, smem1, &g1));
  else RET_IF(launch_cfg(c, k_eam_pass1<LPA, SMEM, MULTI, INTERP>, EAM_BLOCK1, smem1, &g1));
  RET_IF(launch_cfg(c, k_eam_pass2<LPA, SMEM, MULTI, INTERP>, EAM_BLOCK2, smem2, &g2));
  g1 = std::min(g1, div_up(a.n, (rep4 ? b4 : EAM_BLOCK1) / LPA));
  g2 = std::min(g2, div_up(a.n, EAM_BLOCK2 / LPA));
  CU_CHECK(c, c->partials.ensure((size_t) std::max(g1, g2) * 4 + 64));
  a.partials = c->partials.p;
  {
    PhaseTimer pt(c, 3);
    if (rep4 && LPA == 1 && b4 == 768) k_eam_pass1_rep4<1, 768><<<g1, 768, smem1, c->stream>>>(a);
    else if (rep4) k_eam_pass1_rep4<LPA, EAM_BLOCK1><<<g1, EAM_BLOCK1, smem1, c->stream>>>(a);
    else if (fused) k_eam_pass1_fused<LPA><<<g1, EAM_BLOCK1, smem1, c->stream>>>(a);
    else k_eam_pass1<LPA, SMEM, MULTI, INTERP><<<g1, EAM_BLOCK1, smem1, c->stream>>>(a);
    c->count_launch();
  }
  // calc_ttm overlapped with pass 2 (imdcu_run_ttm): numneighs of this step exist now; the lattice
  // kernel goes to a second stream and pass 2 leaves one CTA slot of the machine to it
  const bool fuse = c->fuse_ttm_steps >= 0;
  if (fuse) {
    RET_IF(imdcu_ttm_fork(c, c->fuse_ttm_steps));
    // the lattice CTA gets an SM of its own (ttm.cu: ttm_grid_part), otherwise one CTA slot
    if (g2 == c->num_sms * (g2 / c->num_sms) && g2 >= c->num_sms) g2 -= c->ttm_isolate ? g2 / c->num_sms : 1;
  }
  RET_IF(imdcu_refresh_ghosts(c, false));          // F' (and positions again) to the images
  {
    PhaseTimer pt(c, 5);
    k_eam_pass2<LPA, SMEM, MULTI, INTERP><<<g2, EAM_BLOCK2, smem2, c->stream>>>(a);
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
  if (c->interp == IMDCU_INTERP_4POINT) return run_passes<LPA, false, true, 1>(c, a, 0, 0);
  if (c->interp == IMDCU_INTERP_SPLINE) {
    for (int t = 0; t < 3; t++)
      if (!c->tab[t].raw2) return c->fail(IMDCU_EINVAL, "spline mode: imdcu_set_table2 missing for a table");
    return run_passes<LPA, false, true, 2>(c, a, 0, 0);
  }
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
  a.fused = c->interp == IMDCU_INTERP_3POINT ? imdcu_fused_table(c) : nullptr;
  a.rep4 = nullptr; a.rep4_bytes = 0;
  if (a.fused && c->tune.fused_table == 2) {
    size_t rb = 0;
    a.rep4 = imdcu_rep4_table(c, &rb);
    a.rep4_bytes = (unsigned) rb;
  }
  a.partials = nullptr; a.ticket = c->d_ticket; a.results = c->d_results; a.flags = c->d_flags;
  a.count_ttm = 1;
  if (a.n > 0) {
    // lanes per atom: keep at least ~2 resident waves of threads on the machine
    const int64_t want = (int64_t) c->num_sms * 2048;
    int rc;
    const int force_lpa = c->tune.lpa;                                 // imdcu_set_tuning("lpa")
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
  // MPI_Allreduce of imd_forces_nbl.c:997-1016.  When the caller does not read the sums now (the main-loop
  // entry points without an energy buffer) they are reduced together with the integrator's sums, one
  // all-reduce per step instead of three.
  if (out8 || c->comm.nranks <= 1) { RET_IF(imdcu_multi_allreduce(c, c->d_results, 3, 0)); c->force_sums_pending = false; }
  else c->force_sums_pending = true;
  if (out8) {
    CU_CHECK(c, cudaMemcpyAsync(c->h_results, c->d_results, 4 * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(c, cudaStreamSynchronize(c->stream));
    for (int k = 0; k < 8; k++) out8[k] = 0.0;
    out8[0] = c->h_results[0];                              // tot_pot_energy
    out8[1] = c->h_results[1] + c->h_results[2];            // virial (pair + embedding part)
  }
  return IMDCU_OK;
}
