// imd_b200/csrc/ttm3d.cu -- the 3-D TTM electron-temperature diffusion sweep on the device:
// do_DIFF (imd_ttm_3D.c:1603-1829) + ttm_fill_ghost_layers (:3314-3504) on a resident lattice.
//
// The reference keeps an AoS lattice of ttm_Element (168 B, types.h:441-492) twice (l1/l2) and
// sweeps it with a 7-point conservative variable-kappa stencil; only temp, U, xi change.  Here
// the lattice is SoA on the extended grid (nx+2, ny+2, nz+2), z fastest, each z row pitched so
// that the first interior cell starts a 128-byte line.  temp is double-buffered (neighbours are
// read); U and xi are updated in place (only the own cell is read), which is what the
// reference's node2.U = node.U + ..., node.xi += ..., node2.xi = node.xi amount to when both
// lattices start equal and inactive cells are never written.
//
// k_ttm3d_diff: 2.5-D blocking.  A CTA owns a 32(z) x 16(y) column tile and marches along x.
// temp / fd_k / natoms of planes i-1, i, i+1 sit in registers for the own column; the plane i
// values are shared through a double-buffered shared-memory tile with a one-cell halo, so each
// lattice value is read from global memory once per sweep (plus the tile halo): the kernel is
// HBM bound at 108 B per cell update: 4 + 10 x 8 read, 3 x 8 written.
// Compiled -fmad=false, operation order of the reference: results are bit-identical to it.
#include "device.cuh"

namespace {

constexpr int T3_BX = 32, T3_BY = 16;            // tile: z x y threads
constexpr int T3_OFF = 15;                       // row offset: interior k = 1 sits at a multiple of 16 doubles

struct T3Args {
  int nx, ny, nz;                                // interior cells
  long long pz, py;                              // pitches in elements: row (z), plane (y * z)
  int chunk;                                     // x planes per CTA
  int fd_min_atoms, laser_active;
  double tau, fd_vol, xi_fac, invxsq, invysq, invzsq;
  const int *natoms;
  const double *temp_in; double *temp_out;
  const double *fd_k, *fd_g, *Ce, *md_temp, *source, *dens;
  double *U, *xi;
  double *partials;                              // Eabs partial per CTA
};

__device__ __forceinline__ long long t3_idx(const T3Args &a, int i, int j, int k)
{
  return (long long) i * a.py + (long long) j * a.pz + T3_OFF + k;
}

__global__ void __launch_bounds__(T3_BX * T3_BY, 2)
k_ttm3d_diff(T3Args a)
{
  __shared__ double s_t[2][T3_BY + 2][T3_BX + 2];
  __shared__ double s_k[2][T3_BY + 2][T3_BX + 2];
  __shared__ int    s_n[2][T3_BY + 2][T3_BX + 2];
  __shared__ double s_red[T3_BX * T3_BY / 32];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int k = blockIdx.x * T3_BX + tx + 1, j = blockIdx.y * T3_BY + ty + 1;
  const int i0 = blockIdx.z * a.chunk + 1;
  const int i1 = min(i0 + a.chunk, a.nx + 1);
  const bool inside = k <= a.nz && j <= a.ny;
  const int mn = a.fd_min_atoms;
  // which halo cell of the tile this thread also fetches (edges of the tile; ragged tiles end at the lattice ghost)
  const bool h_jm = ty == 0, h_jp = (ty == T3_BY - 1 || j == a.ny), h_km = tx == 0, h_kp = (tx == T3_BX - 1 || k == a.nz);

  double tm = 0, tc = 0, tp = 0, km = 0, kc = 0, kp = 0;
  int nm = 0, nc = 0, np = 0;
  if (inside) {
    const long long pm = t3_idx(a, i0 - 1, j, k), pc = t3_idx(a, i0, j, k);
    tm = a.temp_in[pm]; km = a.fd_k[pm]; nm = a.natoms[pm];
    tc = a.temp_in[pc]; kc = a.fd_k[pc]; nc = a.natoms[pc];
  }
  double eabs = 0.0;
  int buf = 0;
  for (int i = i0; i < i1; i++, buf ^= 1) {
    const long long c = inside ? t3_idx(a, i, j, k) : 0;
    // issue the loads of this plane's own-cell fields and of plane i+1 before the tile exchange
    double ce = 0, g = 0, ti = 0, src = 0, u = 0, xi = 0, dens = 0;
    if (inside) {
      const long long pn = c + a.py;
      tp = a.temp_in[pn]; kp = a.fd_k[pn]; np = a.natoms[pn];
      src = a.source[c];
      if (nc >= mn) { ce = a.Ce[c]; g = a.fd_g[c]; ti = a.md_temp[c]; u = a.U[c]; xi = a.xi[c]; dens = a.dens[c]; }
      s_t[buf][ty + 1][tx + 1] = tc; s_k[buf][ty + 1][tx + 1] = kc; s_n[buf][ty + 1][tx + 1] = nc;
      if (h_jm) { const long long q = c - a.pz; s_t[buf][ty][tx + 1] = a.temp_in[q]; s_k[buf][ty][tx + 1] = a.fd_k[q]; s_n[buf][ty][tx + 1] = a.natoms[q]; }
      if (h_jp) { const long long q = c + a.pz; s_t[buf][ty + 2][tx + 1] = a.temp_in[q]; s_k[buf][ty + 2][tx + 1] = a.fd_k[q]; s_n[buf][ty + 2][tx + 1] = a.natoms[q]; }
      if (h_km) { const long long q = c - 1; s_t[buf][ty + 1][tx] = a.temp_in[q]; s_k[buf][ty + 1][tx] = a.fd_k[q]; s_n[buf][ty + 1][tx] = a.natoms[q]; }
      if (h_kp) { const long long q = c + 1; s_t[buf][ty + 1][tx + 2] = a.temp_in[q]; s_k[buf][ty + 1][tx + 2] = a.fd_k[q]; s_n[buf][ty + 1][tx + 2] = a.natoms[q]; }
    }
    __syncthreads();                               // one barrier per plane: the tile is double-buffered
    if (inside) {
      if (a.laser_active) eabs += src * a.fd_vol * a.tau;                      // :1648-1652
      if (nc >= mn) {                                                           // :1660
        // a neighbour below fd_min_atoms is replaced by the cell itself (:1661-1682)
        const bool oxm = nm >= mn, oxp = np >= mn;
        const bool oym = s_n[buf][ty][tx + 1] >= mn, oyp = s_n[buf][ty + 2][tx + 1] >= mn;
        const bool ozm = s_n[buf][ty + 1][tx] >= mn, ozp = s_n[buf][ty + 1][tx + 2] >= mn;
        const double xmaxTe = oxp ? tp : tc, xminTe = oxm ? tm : tc;
        const double ymaxTe = oyp ? s_t[buf][ty + 2][tx + 1] : tc, yminTe = oym ? s_t[buf][ty][tx + 1] : tc;
        const double zmaxTe = ozp ? s_t[buf][ty + 1][tx + 2] : tc, zminTe = ozm ? s_t[buf][ty + 1][tx] : tc;
        const double xmaxk = oxp ? kp : kc, xmink = oxm ? km : kc;
        const double ymaxk = oyp ? s_k[buf][ty + 2][tx + 1] : kc, ymink = oym ? s_k[buf][ty][tx + 1] : kc;
        const double zmaxk = ozp ? s_k[buf][ty + 1][tx + 2] : kc, zmink = ozm ? s_k[buf][ty + 1][tx] : kc;
        // :1715-1737, same association
        const double t2 = a.tau / ce *
            ((((kc + xmaxk) / 2 * (xmaxTe - tc) * a.invxsq)
              - ((kc + xmink) / 2 * (tc - xminTe) * a.invxsq)
              + ((kc + ymaxk) / 2 * (ymaxTe - tc) * a.invysq)
              - ((kc + ymink) / 2 * (tc - yminTe) * a.invysq)
              + ((kc + zmaxk) / 2 * (zmaxTe - tc) * a.invzsq)
              - ((kc + zmink) / 2 * (tc - zminTe) * a.invzsq))
             - g * (tc - ti)
             + src) + tc;
        a.temp_out[c] = t2;
        a.U[c] = u + (t2 - tc) * ce * a.fd_vol / ((double) nc);                 // :1741
        a.xi[c] = xi + (t2 - ti) * a.xi_fac * g / ti / dens;                    // :1744-1745
      } else {
        a.temp_out[c] = tc;                        // inactive: both lattices keep their (equal) value
      }
    }
    tm = tc; km = kc; nm = nc;
    tc = tp; kc = kp; nc = np;
  }
  // Eabs: fixed-order reduction inside the CTA, one partial per CTA (summed in CTA order afterwards)
  if (a.partials) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) eabs += __shfl_down_sync(0xffffffffu, eabs, o);
    const int tid = ty * T3_BX + tx;
    if ((tid & 31) == 0) s_red[tid >> 5] = eabs;
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int w = 0; w < T3_BX * T3_BY / 32; w++) s += s_red[w];
      a.partials[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
    }
  }
}

// Direct variant: one thread per cell, no marching, no shared memory.  All loads of a cell (own fields
// and the six neighbours' temp / fd_k / natoms) are independent and issued together; neighbour reuse
// is left to L1 (z, y within the CTA's 32 x 4 x 4 brick) and L2 (the 126 MB L2 holds many x planes),
// latency is hidden by 2048 resident threads per SM instead of by a software pipeline.  Same
// arithmetic, same results; which variant runs is decided by measurement (profiles/).
constexpr int T3D_BX = 32;

// loads the compiler may not sink into the `active cell` branch (SPEC variants: every load of a cell is in flight before
// natoms[c] is known, one memory round trip per thread instead of two; an inactive cell wastes its neighbour loads)
__device__ __forceinline__ double t3_ld(const double *p)
{
  double v; asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p)); return v;
}
__device__ __forceinline__ int t3_ldi(const int *p)
{
  int v; asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}

// BY x BZ rows of 32 cells per CTA, MINB CTAs per SM (sets the register budget), SPEC as above.  Same arithmetic in
// every instantiation; which one runs is decided by measurement (imdcu_set_tuning("ttm3d_variant")).
template <int BY, int BZ, int MINB, bool SPEC>
__global__ void __launch_bounds__(T3D_BX * BY * BZ, MINB)
k_ttm3d_diff_direct(T3Args a)
{
  __shared__ double s_red[T3D_BX * BY * BZ / 32];
  const int k = blockIdx.x * T3D_BX + threadIdx.x + 1;
  const int j = blockIdx.y * BY + threadIdx.y + 1;
  const int i = blockIdx.z * BZ + threadIdx.z + 1;
  const bool inside = k <= a.nz && j <= a.ny && i <= a.nx;
  const int mn = a.fd_min_atoms;
  double eabs = 0.0;
  if (inside) {
    const long long c = t3_idx(a, i, j, k);
    const long long qxm = c - a.py, qxp = c + a.py, qym = c - a.pz, qyp = c + a.pz, qzm = c - 1, qzp = c + 1;
    int nc, nxm, nxp, nym, nyp, nzm, nzp;
    double tc, src, txm, txp, tym, typ, tzm, tzp, kc, kxm, kxp, kym, kyp, kzm, kzp, ce, g, ti, u, xi, dens;
    if (SPEC) {
      nc = t3_ldi(a.natoms + c); tc = t3_ld(a.temp_in + c); src = t3_ld(a.source + c);
      nxm = t3_ldi(a.natoms + qxm); nxp = t3_ldi(a.natoms + qxp); nym = t3_ldi(a.natoms + qym);
      nyp = t3_ldi(a.natoms + qyp); nzm = t3_ldi(a.natoms + qzm); nzp = t3_ldi(a.natoms + qzp);
      txm = t3_ld(a.temp_in + qxm); txp = t3_ld(a.temp_in + qxp); tym = t3_ld(a.temp_in + qym);
      typ = t3_ld(a.temp_in + qyp); tzm = t3_ld(a.temp_in + qzm); tzp = t3_ld(a.temp_in + qzp);
      kc = t3_ld(a.fd_k + c);
      kxm = t3_ld(a.fd_k + qxm); kxp = t3_ld(a.fd_k + qxp); kym = t3_ld(a.fd_k + qym);
      kyp = t3_ld(a.fd_k + qyp); kzm = t3_ld(a.fd_k + qzm); kzp = t3_ld(a.fd_k + qzp);
      ce = t3_ld(a.Ce + c); g = t3_ld(a.fd_g + c); ti = t3_ld(a.md_temp + c); u = t3_ld(a.U + c);
      xi = t3_ld(a.xi + c); dens = t3_ld(a.dens + c);
    } else {
      nc = a.natoms[c]; tc = a.temp_in[c]; src = a.source[c];
    }
    if (a.laser_active) eabs = src * a.fd_vol * a.tau;
    if (nc >= mn) {
      if (!SPEC) {
        nxm = a.natoms[qxm]; nxp = a.natoms[qxp]; nym = a.natoms[qym]; nyp = a.natoms[qyp]; nzm = a.natoms[qzm]; nzp = a.natoms[qzp];
        txm = a.temp_in[qxm]; txp = a.temp_in[qxp]; tym = a.temp_in[qym]; typ = a.temp_in[qyp]; tzm = a.temp_in[qzm]; tzp = a.temp_in[qzp];
        kc = a.fd_k[c];
        kxm = a.fd_k[qxm]; kxp = a.fd_k[qxp]; kym = a.fd_k[qym]; kyp = a.fd_k[qyp]; kzm = a.fd_k[qzm]; kzp = a.fd_k[qzp];
        ce = a.Ce[c]; g = a.fd_g[c]; ti = a.md_temp[c]; u = a.U[c]; xi = a.xi[c]; dens = a.dens[c];
      }
      const bool oxm = nxm >= mn, oxp = nxp >= mn, oym = nym >= mn, oyp = nyp >= mn, ozm = nzm >= mn, ozp = nzp >= mn;
      const double xmaxTe = oxp ? txp : tc, xminTe = oxm ? txm : tc;
      const double ymaxTe = oyp ? typ : tc, yminTe = oym ? tym : tc;
      const double zmaxTe = ozp ? tzp : tc, zminTe = ozm ? tzm : tc;
      const double xmaxk = oxp ? kxp : kc, xmink = oxm ? kxm : kc;
      const double ymaxk = oyp ? kyp : kc, ymink = oym ? kym : kc;
      const double zmaxk = ozp ? kzp : kc, zmink = ozm ? kzm : kc;
      const double t2 = a.tau / ce *
          ((((kc + xmaxk) / 2 * (xmaxTe - tc) * a.invxsq)
            - ((kc + xmink) / 2 * (tc - xminTe) * a.invxsq)
            + ((kc + ymaxk) / 2 * (ymaxTe - tc) * a.invysq)
            - ((kc + ymink) / 2 * (tc - yminTe) * a.invysq)
            + ((kc + zmaxk) / 2 * (zmaxTe - tc) * a.invzsq)
            - ((kc + zmink) / 2 * (tc - zminTe) * a.invzsq))
           - g * (tc - ti)
           + src) + tc;
      a.temp_out[c] = t2;
      a.U[c] = u + (t2 - tc) * ce * a.fd_vol / ((double) nc);
      a.xi[c] = xi + (t2 - ti) * a.xi_fac * g / ti / dens;
    } else {
      a.temp_out[c] = tc;
    }
  }
  if (a.partials) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) eabs += __shfl_down_sync(0xffffffffu, eabs, o);
    const int tid = (threadIdx.z * BY + threadIdx.y) * T3D_BX + threadIdx.x;
    if ((tid & 31) == 0) s_red[tid >> 5] = eabs;
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int w = 0; w < T3D_BX * BY * BZ / 32; w++) s += s_red[w];
      a.partials[(blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
    }
  }
}

// the instantiations that exist (ttm3d_variant): {BY, BZ, MINB, SPEC}
static const int T3D_VARIANTS[][4] = {
  {4, 4, 2, 0},      // 0: 512 threads, 64 registers (a few spilled): the round-1 kernel
  {4, 4, 2, 1},      // 1: the same with every load issued before natoms[c] is known
  {4, 2, 3, 0},      // 2: 256 threads, 3 CTAs per SM: 80 registers, 768 resident threads
  {4, 2, 3, 1},      // 3
  {4, 2, 4, 1},      // 4: 256 threads, 4 CTAs per SM: 64 registers
  {2, 2, 6, 1},      // 5: 128 threads, 6 CTAs per SM: 80 registers
  {4, 2, 2, 1},      // 6: 256 threads, 2 CTAs per SM: up to 128 registers, 512 resident threads
  {4, 1, 8, 1},      // 7: 128 threads (one x plane of 4 rows), 8 CTAs per SM: 64 registers
  {2, 4, 4, 1},      // 8: 256 threads as 2 rows x 4 planes
  {8, 1, 4, 1},      // 9: 256 threads as 8 rows of one plane                              <- default (measured best)
  {4, 2, 5, 1},      // 10: 256 threads, 5 CTAs per SM: 48 registers, 1280 resident threads
  {2, 2, 8, 1},      // 11: 128 threads as 2 rows x 2 planes, 8 CTAs per SM
};
static const int T3D_NVARIANTS = (int) (sizeof(T3D_VARIANTS) / sizeof(T3D_VARIANTS[0]));

static void t3d_launch(int variant, dim3 grid, dim3 block, cudaStream_t st, const T3Args &a)
{
  switch (variant) {
    case 1: k_ttm3d_diff_direct<4, 4, 2, true><<<grid, block, 0, st>>>(a); break;
    case 2: k_ttm3d_diff_direct<4, 2, 3, false><<<grid, block, 0, st>>>(a); break;
    case 3: k_ttm3d_diff_direct<4, 2, 3, true><<<grid, block, 0, st>>>(a); break;
    case 4: k_ttm3d_diff_direct<4, 2, 4, true><<<grid, block, 0, st>>>(a); break;
    case 5: k_ttm3d_diff_direct<2, 2, 6, true><<<grid, block, 0, st>>>(a); break;
    case 6: k_ttm3d_diff_direct<4, 2, 2, true><<<grid, block, 0, st>>>(a); break;
    case 7: k_ttm3d_diff_direct<4, 1, 8, true><<<grid, block, 0, st>>>(a); break;
    case 8: k_ttm3d_diff_direct<2, 4, 4, true><<<grid, block, 0, st>>>(a); break;
    case 9: k_ttm3d_diff_direct<8, 1, 4, true><<<grid, block, 0, st>>>(a); break;
    case 10: k_ttm3d_diff_direct<4, 2, 5, true><<<grid, block, 0, st>>>(a); break;
    case 11: k_ttm3d_diff_direct<2, 2, 8, true><<<grid, block, 0, st>>>(a); break;
    default: k_ttm3d_diff_direct<4, 4, 2, false><<<grid, block, 0, st>>>(a); break;
  }
}

__global__ void k_ttm3d_sum(const double *__restrict__ partials, int n, double *__restrict__ eabs)
{
  __shared__ double s[256];
  double v = 0.0;
  for (int q = threadIdx.x; q < n; q += 256) v += partials[q];
  s[threadIdx.x] = v;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if ((int) threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o]; __syncthreads(); }
  if (threadIdx.x == 0) *eabs += s[0];
}

// ttm_fill_ghost_layers (:3314-3504): faces only (every plane type spans the interior of the other two
// directions).  Periodic: natoms, temp, fd_k of the opposite interior plane; open: natoms = 0.
// One thread per face cell; the six faces are disjoint, so one launch does all of them.
struct T3Ghost {
  int nx, ny, nz, pbc[3];
  int xmode[2];                                  // low / high x ghost plane: 0 open, 1 periodic (local copy), 2 filled by the x-neighbour rank
  long long pz, py;
  int *natoms; double *temp, *fd_k;
};

__global__ void __launch_bounds__(256)
k_ttm3d_ghosts(T3Ghost a)
{
  const long long fx = (long long) a.ny * a.nz, fy = (long long) a.nx * a.nz, fz = (long long) a.nx * a.ny;
  const long long total = 2 * (fx + fy + fz);
  for (long long q = (long long) blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long long) gridDim.x * blockDim.x) {
    int ax, side, u, v;
    long long r = q;
    if (r < 2 * fx) { ax = 0; side = (int) (r / fx); r %= fx; u = (int) (r / a.nz); v = (int) (r % a.nz); }
    else if ((r -= 2 * fx) < 2 * fy) { ax = 1; side = (int) (r / fy); r %= fy; u = (int) (r / a.nz); v = (int) (r % a.nz); }
    else { r -= 2 * fy; ax = 2; side = (int) (r / fz); r %= fz; u = (int) (r / a.ny); v = (int) (r % a.ny); }
    int gi, gj, gk, si, sj, sk;                    // ghost cell, source cell
    if (ax == 0) { gj = sj = u + 1; gk = sk = v + 1; gi = side ? a.nx + 1 : 0; si = side ? 1 : a.nx; }
    else if (ax == 1) { gi = si = u + 1; gk = sk = v + 1; gj = side ? a.ny + 1 : 0; sj = side ? 1 : a.ny; }
    else { gi = si = u + 1; gj = sj = v + 1; gk = side ? a.nz + 1 : 0; sk = side ? 1 : a.nz; }
    const long long g = (long long) gi * a.py + (long long) gj * a.pz + T3_OFF + gk;
    const long long s = (long long) si * a.py + (long long) sj * a.pz + T3_OFF + sk;
    const int mode = ax == 0 ? a.xmode[side] : a.pbc[ax];
    if (mode == 1) { a.natoms[g] = a.natoms[s]; a.temp[g] = a.temp[s]; a.fd_k[g] = a.fd_k[s]; }
    else if (mode == 0) a.natoms[g] = 0;
  }
}

}  // namespace

// ----------------------------------------------------------------------------- host side

void imdcu_ttm3d_release(imdcu_ctx *c)
{
  Ttm3dState &t = c->ttm3d;
  t.natoms.release(); t.temp[0].release(); t.temp[1].release(); t.fd_k.release(); t.fd_g.release(); t.Ce.release();
  t.md_temp.release(); t.source.release(); t.U.release(); t.xi.release(); t.dens.release(); t.partials.release();
  if (t.d_eabs) cudaFree(t.d_eabs);
  t.d_eabs = nullptr; t.inited = false;
}

// the lattice set-up part of init_ttm (imd_ttm_3D.c:1040-1083): sizes, lattice constants, two lattices
extern "C" int imdcu_ttm3d_init(imdcu_ctx *c, int nx, int ny, int nz, const double h[3], const int pbc[3],
                                int fd_min_atoms)
{
  if (!c || !h || !pbc || nx < 1 || ny < 1 || nz < 1 || !(h[0] > 0) || !(h[1] > 0) || !(h[2] > 0))
    return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm3d_init: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  Ttm3dState &t = c->ttm3d;
  t.nx = nx; t.ny = ny; t.nz = nz; t.fd_min_atoms = fd_min_atoms;
  for (int d = 0; d < 3; d++) { t.h[d] = h[d]; t.pbc[d] = pbc[d] != 0; }
  t.pz = ((long long) nz + 2 + T3_OFF + 15) / 16 * 16;
  t.py = t.pz * (ny + 2);
  t.total = t.py * (nx + 2);
  const size_t n = (size_t) t.total;
  CU_CHECK(c, t.natoms.ensure_exact(n));
  DevBuf<double> *all[] = { &t.temp[0], &t.temp[1], &t.fd_k, &t.fd_g, &t.Ce, &t.md_temp, &t.source, &t.U, &t.xi, &t.dens };
  for (DevBuf<double> *b : all) {
    CU_CHECK(c, b->ensure_exact(n));
    CU_CHECK(c, cudaMemsetAsync(b->p, 0, n * sizeof(double), c->stream));
  }
  CU_CHECK(c, cudaMemsetAsync(t.natoms.p, 0, n * sizeof(int), c->stream));
  if (!t.d_eabs) CU_CHECK(c, cudaMalloc(&t.d_eabs, sizeof(double)));
  CU_CHECK(c, cudaMemsetAsync(t.d_eabs, 0, sizeof(double), c->stream));
  t.cur = 0;
  t.inited = true;
  return IMDCU_OK;
}

// host arrays are the extended lattice (nx+2)(ny+2)(nz+2), z fastest, like l1[i][j][k]
template <typename T>
static int t3_copy(imdcu_ctx *c, T *dev, T *host, bool up)
{
  const Ttm3dState &t = c->ttm3d;
  const size_t w = (size_t) (t.nz + 2) * sizeof(T), rows = (size_t) (t.nx + 2) * (t.ny + 2);
  if (up) CU_CHECK(c, cudaMemcpy2DAsync(dev + T3_OFF, (size_t) t.pz * sizeof(T), host, w, w, rows, cudaMemcpyHostToDevice, c->stream));
  else CU_CHECK(c, cudaMemcpy2DAsync(host, w, dev + T3_OFF, (size_t) t.pz * sizeof(T), w, rows, cudaMemcpyDeviceToHost, c->stream));
  return IMDCU_OK;
}

extern "C" int imdcu_ttm3d_upload(imdcu_ctx *c, const int *natoms, const double *temp, const double *fd_k,
                                  const double *fd_g, const double *Ce, const double *md_temp,
                                  const double *source, const double *U, const double *xi, const double *dens)
{
  if (!c || !c->ttm3d.inited) return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm3d_init has not been called") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  Ttm3dState &t = c->ttm3d;
  if (natoms) RET_IF(t3_copy(c, t.natoms.p, const_cast<int *>(natoms), true));
  if (temp) {                                    // both lattices start from the same values (init_ttm)
    RET_IF(t3_copy(c, t.temp[0].p, const_cast<double *>(temp), true));
    RET_IF(t3_copy(c, t.temp[1].p, const_cast<double *>(temp), true));
  }
  struct { double *d; const double *h; } f[] = { { t.fd_k.p, fd_k }, { t.fd_g.p, fd_g }, { t.Ce.p, Ce }, { t.md_temp.p, md_temp },
                                                 { t.source.p, source }, { t.U.p, U }, { t.xi.p, xi }, { t.dens.p, dens } };
  for (auto &e : f) if (e.h) RET_IF(t3_copy(c, e.d, const_cast<double *>(e.h), true));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  return IMDCU_OK;
}

extern "C" int imdcu_ttm3d_download(imdcu_ctx *c, int *natoms, double *temp, double *U, double *xi)
{
  if (!c || !c->ttm3d.inited) return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm3d_init has not been called") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  Ttm3dState &t = c->ttm3d;
  if (natoms) RET_IF(t3_copy(c, t.natoms.p, natoms, false));
  if (temp) RET_IF(t3_copy(c, t.temp[t.cur].p, temp, false));
  if (U) RET_IF(t3_copy(c, t.U.p, U, false));
  if (xi) RET_IF(t3_copy(c, t.xi.p, xi, false));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  return IMDCU_OK;
}

// x-neighbour ranks of the slab decomposition (cpu_dim = N 1 1): -1 at an open end
static void t3_neighbours(const imdcu_ctx *c, int *west, int *east)
{
  const int r = c->comm.rank, N = c->comm.nranks, px = c->ttm3d.pbc[0];
  *east = r + 1 < N ? r + 1 : (px ? 0 : -1);
  *west = r - 1 >= 0 ? r - 1 : (px ? N - 1 : -1);
}

static int t3_ghosts(imdcu_ctx *c)
{
  Ttm3dState &t = c->ttm3d;
  T3Ghost g;
  g.nx = t.nx; g.ny = t.ny; g.nz = t.nz; g.pz = t.pz; g.py = t.py;
  for (int d = 0; d < 3; d++) g.pbc[d] = t.pbc[d];
  int west = -1, east = -1;
  const bool multi = c->comm.nranks > 1;
  if (multi) t3_neighbours(c, &west, &east);
  g.xmode[0] = multi ? (west >= 0 ? 2 : 0) : t.pbc[0];
  g.xmode[1] = multi ? (east >= 0 ? 2 : 0) : t.pbc[0];
  g.natoms = t.natoms.p; g.fd_k = t.fd_k.p; g.temp = t.temp[t.cur].p;
  const long long faces = 2 * ((long long) t.ny * t.nz + (long long) t.nx * t.nz + (long long) t.nx * t.ny);
  const int gblocks = (int) std::min<long long>((faces + 255) / 256, (long long) c->num_sms * 8);
  k_ttm3d_ghosts<<<gblocks, 256, 0, c->stream>>>(g);
  c->count_launch();
  if (multi) {
    // whole planes travel (their y/z ghost rows ride along): plane 1 -> west's plane nx+1, plane nx -> east's
    // plane 0; the members the reference's mpi_element carries: natoms, temp, fd_k (imd_ttm_3D.c:2990-3006)
    const size_t n = (size_t) t.py, lo = (size_t) t.py, hi = (size_t) t.py * t.nx, gh = (size_t) t.py * (t.nx + 1);
    double *T = t.temp[t.cur].p, *K = t.fd_k.p;
    int *N = t.natoms.p;
    RET_IF(imdcu_multi_sendrecv_x(c, west, east, T + lo, T + hi, T + gh, T, n, 0));
    RET_IF(imdcu_multi_sendrecv_x(c, west, east, K + lo, K + hi, K + gh, K, n, 0));
    RET_IF(imdcu_multi_sendrecv_x(c, west, east, N + lo, N + hi, N + gh, N, n, 1));
  }
  return IMDCU_OK;
}

// ttm_fill_ghost_layers (imd_ttm_3D.c:3314-3504) on l1, as calc_ttm calls it ahead of the sub-step loop (:67)
extern "C" int imdcu_ttm3d_fill_ghost_layers(imdcu_ctx *c)
{
  if (!c || !c->ttm3d.inited) return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm3d_init has not been called") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  RET_IF(t3_ghosts(c));
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}

// nsweeps x { do_DIFF(tau); ttm_fill_ghost_layers(); } -- the FD sub-step loop of calc_ttm
// (imd_ttm_3D.c:193-200) with the wide-range properties frozen.  Eabs (if given) receives
// Eabs_local accumulated since imdcu_ttm3d_init.
extern "C" int imdcu_ttm3d_diff(imdcu_ctx *c, double tau, int diff_substeps, int laser_active, int nsweeps,
                                double *Eabs)
{
  if (!c || !c->ttm3d.inited) return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm3d_init has not been called") : IMDCU_EINVAL;
  if (diff_substeps < 1 || nsweeps < 0) return c->fail(IMDCU_EINVAL, "imdcu_ttm3d_diff: bad argument");
  cudaSetDevice(c->device);
  Ttm3dState &t = c->ttm3d;
  T3Args a;
  a.nx = t.nx; a.ny = t.ny; a.nz = t.nz; a.pz = t.pz; a.py = t.py;
  a.fd_min_atoms = t.fd_min_atoms; a.laser_active = laser_active != 0;
  a.tau = tau;
  a.fd_vol = t.h[0] * t.h[1] * t.h[2];
  a.xi_fac = 26.9815 * 1.660539040200000e-27 / 3.0 * 1e30 / ((double) diff_substeps);      // :1614
  a.invxsq = 1.0 / (t.h[0] * t.h[0]); a.invysq = 1.0 / (t.h[1] * t.h[1]); a.invzsq = 1.0 / (t.h[2] * t.h[2]);
  a.natoms = t.natoms.p; a.fd_k = t.fd_k.p; a.fd_g = t.fd_g.p; a.Ce = t.Ce.p; a.md_temp = t.md_temp.p;
  a.source = t.source.p; a.dens = t.dens.p; a.U = t.U.p; a.xi = t.xi.p;
  // variant: IMDCU_TTM3D=tiled selects the 2.5-D shared-memory kernel, default is the direct one
  const bool tiled = c->tune.ttm3d_tiled != 0;
  const int variant = (c->tune.ttm3d_variant >= 0 && c->tune.ttm3d_variant < T3D_NVARIANTS) ? c->tune.ttm3d_variant : 0;
  dim3 grid, block;
  if (tiled) {
    // x chunk: enough CTAs for >= 4 waves of 148 SMs x 4 resident CTAs, at least 8 planes per CTA
    const int gx = div_up(t.nz, T3_BX), gy = div_up(t.ny, T3_BY);
    int chunk = t.nx;
    const long long want = (long long) c->num_sms * 16;
    if ((long long) gx * gy < want) {
      const int nch = (int) std::min<long long>(div_up((int) want, gx * gy), std::max(1, t.nx / 8));
      chunk = div_up(t.nx, std::max(1, nch));
    }
    a.chunk = chunk;
    grid = dim3(gx, gy, div_up(t.nx, chunk)); block = dim3(T3_BX, T3_BY);
  } else {
    a.chunk = 0;
    const int *tv = T3D_VARIANTS[variant];
    grid = dim3(div_up(t.nz, T3D_BX), div_up(t.ny, tv[0]), div_up(t.nx, tv[1])); block = dim3(T3D_BX, tv[0], tv[1]);
    if (grid.y > 65535 || grid.z > 65535) return c->fail(IMDCU_EINVAL, "imdcu_ttm3d_diff: lattice too large for the launch grid");
  }
  const int nblocks = (int) (grid.x * grid.y * grid.z);
  a.partials = nullptr;
  if (a.laser_active) { CU_CHECK(c, t.partials.ensure((size_t) nblocks)); a.partials = t.partials.p; }
  for (int s = 0; s < nsweeps; s++) {
    a.temp_in = t.temp[t.cur].p; a.temp_out = t.temp[1 - t.cur].p;
    {
      PhaseTimer pt(c, 7);
      if (tiled) k_ttm3d_diff<<<grid, block, 0, c->stream>>>(a);
      else t3d_launch(variant, grid, block, c->stream, a);
      c->count_launch();
    }
    if (a.partials) { k_ttm3d_sum<<<1, 256, 0, c->stream>>>(t.partials.p, nblocks, t.d_eabs); c->count_launch(); }
    t.cur = 1 - t.cur;                             // l3 = l1; l1 = l2; l2 = l3 (:1822-1824)
    RET_IF(t3_ghosts(c));
  }
  CU_CHECK(c, cudaGetLastError());
  if (Eabs) {
    CU_CHECK(c, cudaMemcpyAsync(Eabs, t.d_eabs, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(c, cudaStreamSynchronize(c->stream));
  }
  return IMDCU_OK;
}
