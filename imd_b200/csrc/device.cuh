// imd_b200/csrc/device.cuh -- device-side helpers (sm_100a).
#pragma once
#include "common.cuh"

// r2 with the reference's rounding: SPROD3D = ((dx*dx) + (dy*dy)) + (dz*dz), no FMA
// contraction (makros.h:361).  Integer results (neighbour sets, numneighs) depend on it.
__device__ __forceinline__ double r2_exact(double dx, double dy, double dz)
{
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// position + image shift in the order the reference's staged copies add the box vectors
// (z stage, then y, then x: imd_comm_force_3d.c:270-396, copy_cell :754-756).  A zero
// trit adds +0.0, which is exact.
__device__ __forceinline__ void apply_shift(const Geom &g, int code, double sign,
                                            double &x, double &y, double &z)
{
  double fx = sign * (double) (code % 3 - 1);
  double fy = sign * (double) ((code / 3) % 3 - 1);
  double fz = sign * (double) (code / 9 - 1);
  x = __dadd_rn(x, __dmul_rn(fz, g.box[6])); y = __dadd_rn(y, __dmul_rn(fz, g.box[7])); z = __dadd_rn(z, __dmul_rn(fz, g.box[8]));
  x = __dadd_rn(x, __dmul_rn(fy, g.box[3])); y = __dadd_rn(y, __dmul_rn(fy, g.box[4])); z = __dadd_rn(z, __dmul_rn(fy, g.box[5]));
  x = __dadd_rn(x, __dmul_rn(fx, g.box[0])); y = __dadd_rn(y, __dmul_rn(fx, g.box[1])); z = __dadd_rn(z, __dmul_rn(fx, g.box[2]));
}

// one 256-bit load of a packed atom record {x,y,z,w} (sm_100: LDG.E.256).  A gathered
// neighbour is exactly one 32-byte sector, so this is one request and one sector.
__device__ __forceinline__ double4 ld256(const double4 *p)
{
  double4 v;
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}

// the same through the non-coherent path; only for arrays no thread writes while the kernel runs
__device__ __forceinline__ double4 ld256nc(const double4 *p)
{
  double4 v;
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w) : "l"(p));
  return v;
}

// ----------------------------------------------------------------------------- x slabs
// Global x cell layer -> layer index relative to this rank's slab [xoff, xoff + ldim): [0, ldim) owned, < 0 belongs
// to the west side, >= ldim to the east side.  Across a periodic x boundary the offset is taken modulo gdim.x and an
// outside layer is given to the NEARER side (atoms move less than a layer between two sorts); slabs may have any
// widths (imdcu_comm_set_cuts), also more than half of the box.
__device__ __forceinline__ int slab_local_x(const Geom &g, int cx)
{
  int li = cx - g.xoff;
  if (g.pbc[0]) {
    li %= g.gdim[0];
    if (li < 0) li += g.gdim[0];
    if (li >= g.ldim[0] && g.gdim[0] - li < li - g.ldim[0] + 1) li -= g.gdim[0];
  }
  return li;
}

// ----------------------------------------------------------------------------- tables
// 3-point (quadratic in the argument) table access, potaccess.h:323-354.  `c` points to
// the coefficient rows {p0, dv, d2v} of one column (shared or global memory).
struct TabCol {
  const double *c;
  double begin, end, invstep;
};

__device__ __forceinline__ void tab_index(const TabCol &t, double r2, int &k, double &chi, int &is_short)
{
  double r2a = fmin(r2, t.end);
  r2a = r2a - t.begin;
  if (r2a < 0) { r2a = 0; is_short = 1; }
  r2a = r2a * t.invstep;
  k = (int) r2a;
  chi = r2a - (double) k;
}

// PAIR_INT2: value and (1/r) d/dr = 2 d/dr2
__device__ __forceinline__ void pair_int2(const TabCol &t, double r2, double &pot, double &grad, int &is_short)
{
  int k; double chi;
  tab_index(t, r2, k, chi, is_short);
  const double *p = t.c + 3 * (size_t) k;
  double p0 = p[0], dv = p[1], d2v = p[2];
  pot = p0 + chi * dv + 0.5 * chi * (chi - 1) * d2v;
  grad = 2 * t.invstep * (dv + (chi - 0.5) * d2v);
}

// VAL_FUNC2 (potaccess.h:465-495)
__device__ __forceinline__ double val_func2(const TabCol &t, double r2, int &is_short)
{
  int k; double chi;
  tab_index(t, r2, k, chi, is_short);
  const double *p = t.c + 3 * (size_t) k;
  return p[0] + chi * p[1] + 0.5 * chi * (chi - 1) * p[2];
}

// DERIV_FUNC2 (potaccess.h:591-621)
__device__ __forceinline__ double deriv_func2(const TabCol &t, double r2, int &is_short)
{
  int k; double chi;
  tab_index(t, r2, k, chi, is_short);
  const double *p = t.c + 3 * (size_t) k;
  return 2 * t.invstep * (p[1] + (chi - 0.5) * p[2]);
}

__device__ __forceinline__ TabCol tab_col(const TabView &v, const double *coef_base, int col)
{
  TabCol t;
  t.c = coef_base + (size_t) col * v.nrows * 3;
  t.begin = v.begin[col]; t.end = v.end[col]; t.invstep = v.invstep[col];
  return t;
}

// The other two interpolation modes of potaccess.h read the table as the reference does (rows
// k-1..k+2, or rows k, k+1 of table and table2), straight from global memory through L1: they are
// the reference's compile-time options FOURPOINT / SPLINE, not the hot configuration.
struct RawCol {
  const double *t, *t2;      // column `col` of row 0
  double begin, end, invstep, step;
  int inc;
};
__device__ __forceinline__ RawCol raw_col(const TabView &v, int col)
{
  RawCol c;
  c.t = v.raw + col; c.t2 = v.raw2 ? v.raw2 + col : nullptr;
  c.begin = v.begin[col]; c.end = v.end[col]; c.invstep = v.invstep[col]; c.step = v.step[col];
  c.inc = v.ncols;
  return c;
}
// INTERP 1: PAIR_INT3 (potaccess.h:365-413), INTERP 2: PAIR_INT_SP (:421-457); VAL_FUNC* / DERIV_FUNC*
// of a mode evaluate the same expressions, callers drop what they do not need
template <int INTERP>
__device__ __forceinline__ void raw_pair_int(const RawCol &c, double r2, double &pot, double &grad, int &is_short)
{
  double r2a = fmin(r2, c.end);
  r2a = r2a - c.begin;
  if (r2a < 0) { r2a = 0; is_short = 1; }
  const double istep = c.invstep;
  r2a = r2a * istep;
  if (INTERP == 1) {
    int k = (int) r2a; if (k < 1) k = 1;
    const double chi = r2a - (double) k;
    const double fac0 = -(1.0 / 6.0) * chi * (chi - 1.0) * (chi - 2.0);
    const double fac1 = 0.5 * (chi * chi - 1.0) * (chi - 2.0);
    const double fac2 = -0.5 * chi * (chi + 1.0) * (chi - 2.0);
    const double fac3 = (1.0 / 6.0) * chi * (chi * chi - 1.0);
    const double dfac0 = -(1.0 / 6.0) * ((3.0 * chi - 6.0) * chi + 2.0);
    const double dfac1 = 0.5 * ((3.0 * chi - 4.0) * chi - 1.0);
    const double dfac2 = -0.5 * ((3.0 * chi - 2.0) * chi - 2.0);
    const double dfac3 = 1.0 / 6.0 * (3.0 * chi * chi - 1.0);
    const double *p = c.t + (size_t) (k - 1) * c.inc;
    const double p0 = __ldg(p), p1 = __ldg(p + c.inc), p2 = __ldg(p + 2 * c.inc), p3 = __ldg(p + 3 * c.inc);
    pot = fac0 * p0 + fac1 * p1 + fac2 * p2 + fac3 * p3;
    grad = 2 * istep * (dfac0 * p0 + dfac1 * p1 + dfac2 * p2 + dfac3 * p3);
  } else {
    const int k = (int) r2a;
    const double b = r2a - (double) k, a = 1.0 - b;
    const size_t o = (size_t) k * c.inc;
    const double p1 = __ldg(c.t + o), d21 = __ldg(c.t2 + o), p2 = __ldg(c.t + o + c.inc), d22 = __ldg(c.t2 + o + c.inc);
    const double a2 = a * a - 1, b2 = b * b - 1, st6 = c.step / 6;
    pot = a * p1 + b * p2 + (a * a2 * d21 + b * b2 * d22) * st6 * c.step;
    grad = 2 * ((p2 - p1) * istep + ((3 * b2 + 2) * d22 - (3 * a2 + 2) * d21) * st6);
  }
}

// ----------------------------------------------------------------------------- reductions
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Deterministic grid reduction of NV per-thread values: warp shuffle -> shared -> one
// partial row per block -> the last block to finish (ticket) adds the rows in block
// order and writes results[0..NV).  OP 0 = sum, 1 = max.  All threads must call.
template <int NV, int OP>
__device__ __forceinline__ void grid_reduce(double (&v)[NV], double *partials, unsigned int *ticket,
                                            double *results)
{
  __shared__ double s_part[NV][32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int q = 0; q < NV; q++) {
    double w = OP == 0 ? warp_sum(v[q]) : warp_max(v[q]);
    if (lane == 0) s_part[q][warp] = w;
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int q = 0; q < NV; q++) {
      double w = lane < nwarp ? s_part[q][lane] : (OP == 0 ? 0.0 : -1.0e300);
      w = OP == 0 ? warp_sum(w) : warp_max(w);
      if (lane == 0) partials[(size_t) blockIdx.x * NV + q] = w;
    }
  }
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last) {
    __threadfence();
    if (warp == 0) {
#pragma unroll
      for (int q = 0; q < NV; q++) {
        double acc = OP == 0 ? 0.0 : -1.0e300;
        for (unsigned int b = lane; b < gridDim.x; b += 32) {
          double w = partials[(size_t) b * NV + q];
          acc = OP == 0 ? acc + w : fmax(acc, w);
        }
        acc = OP == 0 ? warp_sum(acc) : warp_max(acc);
        if (lane == 0) results[q] = acc;
      }
      if (lane == 0) *ticket = 0;     // re-arm for the next launch
    }
  }
}

// ----------------------------------------------------------------------------- bulk copy (TMA, 1-D)
// cp.async.bulk global -> shared with mbarrier completion (SASS: UBLKCP).  Used to stage
// the coefficient tables; bytes must be a multiple of 16 and both addresses 16-aligned.
__device__ __forceinline__ unsigned smem_u32(const void *p)
{
  return (unsigned) __cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned phase)
{
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}

// stage `bytes` from global to shared: one thread issues bulk copies in <= 64 KiB pieces,
// everybody waits on the barrier.  Call from all threads of the CTA.
__device__ __forceinline__ void stage_table(void *dst, const void *src, size_t bytes,
                                            unsigned long long *bar, unsigned phase)
{
  if (threadIdx.x == 0) {
    size_t off = 0;
    mbar_expect_tx(bar, (unsigned) bytes);          // one arrival, whole transaction count
    while (off < bytes) {
      unsigned chunk = (unsigned) ((bytes - off) < 65536 ? (bytes - off) : 65536);
      bulk_g2s((char *) dst + off, (const char *) src + off, chunk, bar);
      off += chunk;
    }
  }
  __syncthreads();
  mbar_wait(bar, phase);
}
