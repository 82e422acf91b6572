// imd_b200/csrc/nblist.cu -- Verlet neighbour list build (replaces make_nblist and
// estimate_nblist_size, imd_forces_nbl.c:75-287).
//
// Differences in FORM, not in content: the reference builds a CSR half list over the
// 14-cell half stencil (imd_geom_3d.c:917-922); here every atom gets its FULL list over
// the 27-cell stencil (no scatter, no atomics in the force passes), stored in slices of
// 32 atoms:  idx[(tile*cap + k)*32 + lane]  so that a warp reads slot k of its 32 atoms
// as one 128-byte line.  The unordered pair SET equals the reference's bit for bit:
// the predicate r2 < cellsz is evaluated with the reference's rounding AND in the
// reference's orientation -- for a pair that the reference sees as (owner i, image of j)
// the distance is ((x_j + shift) - x_i) also when the pair is met from j's side.
#include "device.cuh"

template <bool COUNT_ONLY>
__global__ void __launch_bounds__(128)
k_nbl_build(Geom g, int n, const double4 *__restrict__ pos4, const int *__restrict__ key,
            const int *__restrict__ cell_begin, const int *__restrict__ cell_count,
            const int *__restrict__ gtcode, const double *__restrict__ graw,
            int cap, int *__restrict__ idx, int *__restrict__ num, int *__restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int cnt = 0;
  if (i < n) {
    const double4 pi = pos4[i];
    const int c = key[i];
    const int ey = g.edim[1], ez = g.edim[2];
    int *slot = idx + ((size_t) (i >> 5) * cap) * 32 + lane;
    for (int l = -1; l <= 1; l++)
      for (int m = -1; m <= 1; m++)
        for (int q = -1; q <= 1; q++) {
          const int c2 = c + (l * ey + m) * ez + q;
          const int b = cell_begin[c2], e = b + cell_count[c2];
          const bool lexneg = (l < 0) || (l == 0 && m < 0) || (l == 0 && m == 0 && q < 0);
          if (!(lexneg && e > n)) {
            // plain candidates, four records requested ahead of their distance tests
            for (int j = b; j < e; j += 4) {
              const int j1 = min(j + 1, e - 1), j2 = min(j + 2, e - 1), j3 = min(j + 3, e - 1);
              const double4 q0 = ld256(&pos4[j]), q1 = ld256(&pos4[j1]), q2 = ld256(&pos4[j2]), q3 = ld256(&pos4[j3]);
              const double r0 = r2_exact(q0.x - pi.x, q0.y - pi.y, q0.z - pi.z);
              const double r1 = r2_exact(q1.x - pi.x, q1.y - pi.y, q1.z - pi.z);
              const double r2 = r2_exact(q2.x - pi.x, q2.y - pi.y, q2.z - pi.z);
              const double r3 = r2_exact(q3.x - pi.x, q3.y - pi.y, q3.z - pi.z);
              if (j != i && r0 < g.cellsz) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j; cnt++; }
              if (j + 1 < e && j + 1 != i && r1 < g.cellsz) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 1; cnt++; }
              if (j + 2 < e && j + 2 != i && r2 < g.cellsz) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 2; cnt++; }
              if (j + 3 < e && j + 3 != i && r3 < g.cellsz) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 3; cnt++; }
            }
            continue;
          }
          for (int j = b; j < e; j++) {
            if (j == i) continue;
            double dx, dy, dz;
            if (j >= n && lexneg) {
              // the reference meets this pair from the other atom: image of i minus raw j
              double x = pi.x, y = pi.y, z = pi.z;
              apply_shift(g, gtcode[j - n], -1.0, x, y, z);
              const double *r = graw + 3 * (size_t) (j - n);
              dx = x - r[0]; dy = y - r[1]; dz = z - r[2];
            } else {
              const double4 pj = ld256(&pos4[j]);
              dx = pj.x - pi.x; dy = pj.y - pi.y; dz = pj.z - pi.z;
            }
            const double r2 = r2_exact(dx, dy, dz);
            if (r2 < g.cellsz) {
              if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j;
              cnt++;
            }
          }
        }
    num[i] = cnt;
  }
  int mx = cnt;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0 && mx > 0) atomicMax(&flags[1], mx);
}

// ----------------------------------------------------------------------------- build with a float classifier
// k_nbl_build is bound by FP64: ~27 x 18 candidates per Al atom, eight double-precision operations each (1.5 ms of
// FP64 pipe time per 4 M atoms), to keep one in eleven.  k_nbl_build2 classifies every candidate in SINGLE precision:
//     r2_float <  cellsz - margin   the pair is in range            (no FP64)
//     r2_float >= cellsz + margin   it is not                       (no FP64)
//     otherwise                     the reference's exact test decides (r2 < cellsz with SPROD3D rounding, makros.h:361)
// The margin bounds the error of the float test,
//     |r2_float - r2| <= 2 sqrt(3) r 2^-23 X + 2^-21.3 r^2     for |coordinates| <= X and r <= sqrt(cellsz) + a little,
// and X is MEASURED at every rebuild (k_posf), so atoms that fly far out of an open box only widen the shell.  Inside
// the bound the float decision equals the exact one, so the list -- order included -- is k_nbl_build's.  The shell holds
// ~5e-5 of the candidates: the exact branch is rare enough that its divergence does not show.  Pairs the reference meets
// from the image side (lexneg ghost cells, boundary cells only) keep the exact path.
__global__ void __launch_bounds__(256)
k_posf(int ntot, const double4 *__restrict__ pos4, float4 *__restrict__ posf, unsigned int *__restrict__ xmax_bits)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  float m = 0.0f;
  if (i < ntot) {
    const double4 p = pos4[i];
    posf[i] = make_float4((float) p.x, (float) p.y, (float) p.z, 0.0f);
    m = (float) fmax(fmax(fabs(p.x), fabs(p.y)), fabs(p.z));
    m = m * 1.0000002f;                                   // the conversion may have rounded down
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.0f) atomicMax(xmax_bits, __float_as_uint(m));   // non-negative floats order as ints
}

template <bool COUNT_ONLY>
__global__ void __launch_bounds__(128)
k_nbl_build2(Geom g, int n, const double4 *__restrict__ pos4, const float4 *__restrict__ posf,
             const unsigned int *__restrict__ xmax_bits, const int *__restrict__ key,
             const int *__restrict__ cell_begin, const int *__restrict__ cell_count,
             const int *__restrict__ gtcode, const double *__restrict__ graw,
             int cap, int *__restrict__ idx, int *__restrict__ num, int *__restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  int cnt = 0;
  if (i < n) {
    const double xmax = (double) __uint_as_float(*xmax_bits);
    const double rc = sqrt(g.cellsz);
    // 4 > 2 sqrt(3) and 2^-20 > 2^-21.3: the bound with room to spare; lo is rounded down, hi up
    const double marg = 4.0 * rc * 1.1920928955078125e-7 * xmax + 9.5367431640625e-7 * g.cellsz;
    const float lo = (float) ((g.cellsz - marg) * 0.999999), hi = (float) ((g.cellsz + marg) * 1.000001);
    const double4 pi = pos4[i];
    const float4 pf = posf[i];
    const int c = key[i];
    const int ey = g.edim[1], ez = g.edim[2];
    int *slot = idx + ((size_t) (i >> 5) * cap) * 32 + lane;
    // in the shell between lo and hi the exact test decides
    auto in_range = [&](float r2f, int j) -> bool {
      if (r2f < lo) return true;
      if (!(r2f < hi)) return false;
      const double4 pj = ld256(&pos4[j]);
      return r2_exact(pj.x - pi.x, pj.y - pi.y, pj.z - pi.z) < g.cellsz;
    };
    for (int l = -1; l <= 1; l++)
      for (int m = -1; m <= 1; m++)
        for (int q = -1; q <= 1; q++) {
          const int c2 = c + (l * ey + m) * ez + q;
          const int b = cell_begin[c2], e = b + cell_count[c2];
          const bool lexneg = (l < 0) || (l == 0 && m < 0) || (l == 0 && m == 0 && q < 0);
          if (!(lexneg && e > n)) {
            for (int j = b; j < e; j += 4) {               // four candidates per round
              const int j1 = min(j + 1, e - 1), j2 = min(j + 2, e - 1), j3 = min(j + 3, e - 1);
              const float4 q0 = posf[j], q1 = posf[j1], q2 = posf[j2], q3 = posf[j3];
              float dx = q0.x - pf.x, dy = q0.y - pf.y, dz = q0.z - pf.z;
              const float r0 = dx * dx + dy * dy + dz * dz;
              dx = q1.x - pf.x; dy = q1.y - pf.y; dz = q1.z - pf.z;
              const float r1 = dx * dx + dy * dy + dz * dz;
              dx = q2.x - pf.x; dy = q2.y - pf.y; dz = q2.z - pf.z;
              const float r2 = dx * dx + dy * dy + dz * dz;
              dx = q3.x - pf.x; dy = q3.y - pf.y; dz = q3.z - pf.z;
              const float r3 = dx * dx + dy * dy + dz * dz;
              if (j != i && r0 < hi && in_range(r0, j)) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j; cnt++; }
              if (j + 1 < e && j + 1 != i && r1 < hi && in_range(r1, j + 1)) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 1; cnt++; }
              if (j + 2 < e && j + 2 != i && r2 < hi && in_range(r2, j + 2)) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 2; cnt++; }
              if (j + 3 < e && j + 3 != i && r3 < hi && in_range(r3, j + 3)) { if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j + 3; cnt++; }
            }
            continue;
          }
          for (int j = b; j < e; j++) {                    // boundary cells with images: the exact path of k_nbl_build
            if (j == i) continue;
            double dx, dy, dz;
            if (j >= n && lexneg) {
              double x = pi.x, y = pi.y, z = pi.z;
              apply_shift(g, gtcode[j - n], -1.0, x, y, z);
              const double *r = graw + 3 * (size_t) (j - n);
              dx = x - r[0]; dy = y - r[1]; dz = z - r[2];
            } else {
              const double4 pj = ld256(&pos4[j]);
              dx = pj.x - pi.x; dy = pj.y - pi.y; dz = pj.z - pi.z;
            }
            if (r2_exact(dx, dy, dz) < g.cellsz) {
              if (!COUNT_ONLY && cnt < cap) slot[(size_t) cnt * 32] = j;
              cnt++;
            }
          }
        }
    num[i] = cnt;
  }
  int mx = cnt;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if (lane == 0 && mx > 0) atomicMax(&flags[1], mx);
}

// error flags that kernels raise between two list builds (the build resets them)
int imdcu_check_flags(imdcu_ctx *c)
{
  CU_CHECK(c, cudaMemcpyAsync(c->h_flags, c->d_flags, 8 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  // is_short of the EAM passes (potaccess.h:331-333: argument below the table begin, clamped): the reference
  // prints "short distance" after calc_forces (imd_forces_nbl.c:673,992); here it is kept as a warning count
  if (c->h_flags[2] || c->h_flags[3]) {
    c->short_warn += 1;
    CU_CHECK(c, cudaMemsetAsync(c->d_flags + 2, 0, 2 * sizeof(int), c->stream));
  }
  if (c->h_flags[4]) return c->fail(IMDCU_EINVAL, "an atom crossed more than one x slab between two cell sorts");
  if (c->h_flags[5]) return c->fail(IMDCU_EINVAL, "NRB: a listed neighbour of a boundary atom lives on another GPU");
  if (c->h_flags[6]) return c->fail(IMDCU_EINVAL, "NRB: a listed neighbour of a boundary atom has no entry");
  return IMDCU_OK;
}

static int build_nblist_local(imdcu_ctx *c)
{
  const int n = (int) c->n;
  cudaStream_t st = c->stream;
  RET_IF(imdcu_check_flags(c));
  imdcu_dbg_mark(c, "n.flags");
  if (n == 0) return IMDCU_OK;
  const int nb = div_up(n, 128);
  const size_t ntiles = (size_t) div_up(n, 32);
  const bool two_phase = c->tune.nbl_filter != 0;
  if (two_phase) {                           // single-precision copies of owned atoms and images, max |coordinate|
    const int ntot = n + (int) c->ng;
    CU_CHECK(c, c->posf.ensure((size_t) ntot + 64, 1.1));     // headroom: the ghost count changes at every rebuild
    CU_CHECK(c, cudaMemsetAsync(c->d_xmax, 0, sizeof(unsigned int), st));
    k_posf<<<div_up(ntot, 256), 256, 0, st>>>(ntot, c->pos4.p, c->posf.p, c->d_xmax);
    c->count_launch();
  }
  auto run = [&](bool count_only) -> int {
    CU_CHECK(c, cudaMemsetAsync(c->d_flags, 0, 8 * sizeof(int), st));
    if (two_phase && count_only)
      k_nbl_build2<true><<<nb, 128, 0, st>>>(c->geom, n, c->pos4.p, c->posf.p, c->d_xmax, c->key.p, c->cell_begin.p,
                                              c->cell_count.p, c->ghost_tcode.p, c->ghost_raw.p, 0, nullptr, c->nbl_num.p, c->d_flags);
    else if (two_phase)
      k_nbl_build2<false><<<nb, 128, 0, st>>>(c->geom, n, c->pos4.p, c->posf.p, c->d_xmax, c->key.p, c->cell_begin.p,
                                               c->cell_count.p, c->ghost_tcode.p, c->ghost_raw.p, c->nbl_cap, c->nbl_idx.p,
                                               c->nbl_num.p, c->d_flags);
    else if (count_only)
      k_nbl_build<true><<<nb, 128, 0, st>>>(c->geom, n, c->pos4.p, c->key.p, c->cell_begin.p, c->cell_count.p,
                                             c->ghost_tcode.p, c->ghost_raw.p, 0, nullptr, c->nbl_num.p, c->d_flags);
    else
      k_nbl_build<false><<<nb, 128, 0, st>>>(c->geom, n, c->pos4.p, c->key.p, c->cell_begin.p, c->cell_count.p,
                                              c->ghost_tcode.p, c->ghost_raw.p, c->nbl_cap, c->nbl_idx.p,
                                              c->nbl_num.p, c->d_flags);
    c->count_launch();
    CU_CHECK(c, cudaMemcpyAsync(c->h_flags, c->d_flags, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CU_CHECK(c, cudaStreamSynchronize(st));
    return IMDCU_OK;
  };
  auto grow = [&](int maxn) -> int {
    // same policy as the reference: nbl_size * measured need (imd_forces_nbl.c:181-190)
    int cap = (int) (c->nbl_size * maxn) + 1;
    cap = (cap + 3) & ~3;
    c->nbl_cap = cap;
    CU_CHECK(c, c->nbl_idx.ensure(ntiles * (size_t) cap * 32, 1.03));
    return IMDCU_OK;
  };
  if (c->nbl_cap == 0) {                     // first build: estimate_nblist_size
    RET_IF(run(true));
    RET_IF(grow(c->h_flags[1]));
  }
  // 3 % headroom: with migration the number of owned atoms creeps up, and every new maximum would otherwise
  // reallocate the whole list (cudaMalloc + cudaFree of ~1.6 GB at 4 M atoms) in the middle of a rebuild
  CU_CHECK(c, c->nbl_idx.ensure(ntiles * (size_t) c->nbl_cap * 32, 1.03));
  imdcu_dbg_mark(c, "n.alloc");
  RET_IF(run(false));
  imdcu_dbg_mark(c, "n.build");
  if (c->h_flags[1] > c->nbl_cap) {          // "neighbor table full": grow and redo once
    RET_IF(grow(c->h_flags[1]));
    RET_IF(run(false));
    if (c->h_flags[1] > c->nbl_cap) return c->fail(IMDCU_EOVERFLOW, "neighbor table full - increase nbl_size");
  }
  c->last_max_neigh = c->h_flags[1];
  return IMDCU_OK;
}

int imdcu_build_nblist(imdcu_ctx *c)
{
  PhaseTimer pt(c, 1);
  // the build itself runs no collective, so a rank may fail or redo it locally; the outcome is agreed on once,
  // here, so that a rank with a raised flag or a full table does not leave its peers in the next halo exchange
  const int rc_local = build_nblist_local(c);
  imdcu_dbg_mark(c, "nbl_build");
  RET_IF(imdcu_multi_agree(c, rc_local));
  imdcu_dbg_mark(c, "agree", true);
  c->have_valid_nbl = true;
  c->nbl_count++;
  return IMDCU_OK;
}

// every unordered pair once, as atom ids (parity tests; not a hot path)
extern "C" int imdcu_get_nblist_pairs(imdcu_ctx *c, int64_t *npairs, int32_t *pairs)
{
  if (!c || !npairs) return IMDCU_EINVAL;
  if (!c->have_valid_nbl) return c->fail(IMDCU_EINVAL, "no valid neighbour list");
  cudaSetDevice(c->device);
  const int64_t n = c->n, ng = c->ng;
  const int cap = c->nbl_cap;
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  std::vector<int> num((size_t) n), id((size_t) n), gid((size_t) ng + 1);
  std::vector<int> idx((size_t) div_up(n, 32) * cap * 32);
  CU_CHECK(c, cudaMemcpy(num.data(), c->nbl_num.p, n * sizeof(int), cudaMemcpyDeviceToHost));
  CU_CHECK(c, cudaMemcpy(id.data(), c->id.p, n * sizeof(int), cudaMemcpyDeviceToHost));
  if (ng) CU_CHECK(c, cudaMemcpy(gid.data(), c->ghost_gid.p, ng * sizeof(int), cudaMemcpyDeviceToHost));
  CU_CHECK(c, cudaMemcpy(idx.data(), c->nbl_idx.p, idx.size() * sizeof(int), cudaMemcpyDeviceToHost));
  int64_t np = 0;
  for (int64_t i = 0; i < n; i++) {
    const size_t base = ((size_t) (i >> 5) * cap) * 32 + (i & 31);
    for (int k = 0; k < num[i]; k++) {
      int j = idx[base + (size_t) k * 32];
      const int idj = j < n ? id[j] : gid[j - n];
      if (id[i] < idj || (id[i] == idj && j >= n && i < j)) {
        if (pairs) { pairs[2 * np] = id[i]; pairs[2 * np + 1] = idj; }
        np++;
      }
    }
  }
  *npairs = np;
  return IMDCU_OK;
}
