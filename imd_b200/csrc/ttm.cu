// imd_b200/csrc/ttm.cu -- the 1-D two-temperature model on the device: calc_ttm
// (imd_ttm.c:75-156) and everything it calls.
//
//   k_ttm_atoms1 : update_fd's first atom loop (imd_ttm.c:269-318): atoms -> FD slabs
//                  (count, sum numneighs, sum p, left/right fluxes vs the stored slab index)
//   k_ttm_atoms2 : update_fd's second atom loop (:368-394): sum m (p/m - v_com)^2
//   k_ttm_grid   : the G-sized rest of calc_ttm in ONE launch: reconstruct l1/l2 (:401-541),
//                  do_ADV incl. the Brent inverse of the EOS (:1208-1301, fminbnd3.c),
//                  do_cell_activation (:1311-1425), do_FILLMESH (:629-852), CFL_maxdt
//                  (:2444-2480) and the whole sub-step loop {source, do_DIFF (:1473-1678),
//                  do_FILLMESH, ttm_fill_ghost_layers (:3264-3356)} -- diff_substeps is decided
//                  on the device, so the >= fd_n_timesteps sub-steps need no host round trip.
//
// The FD lattice is tiny (G = ttmdimx cells, L2/L1 resident); its cost is latency (~15 pow per
// cell and sub-step), so one CTA works on it while the atom kernels are HBM-streaming kernels.
// The reference's result depends on its MPI decomposition; the lattice keeps the reference's
// layout for P = nranks_ref emulated ranks: P x (L+2) cells with ghost cells, two lattices
// swapped by index (l1/l2).  Compiled with -fmad=false: the Brent iteration must see the
// same IEEE operations as the host code.
#include "device.cuh"
#include "ttm_math.cuh"
#include "ttm_models.cuh"

#include <cmath>

#define RHOMIN 2
namespace {
__constant__ const double C_AMU = 1.660539040200000e-27;
__constant__ const double C_J2eV = 6.2415090744607626077762409809304458998869658961709711e18;
__constant__ const double C_eV2J = 1.602176634E-19;
}

#define DMAX(a, b) ((a) > (b) ? (a) : (b))
#define DMIN(a, b) ((a) < (b) ? (a) : (b))
#define DABS(a) ((a) > 0 ? (a) : -(a))

// ----------------------------------------------------------------------------- EOS tables
__device__ __forceinline__ int eos_find(const double *a, int n, double v)
{
  if (!(v >= a[0]) || !(v <= a[n - 1])) return -1;
  int lo = 0, hi = n - 1;
  while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (a[mid] <= v) lo = mid; else hi = mid; }
  return lo;
}

// linear interpolation on the reference's triangulation (lpi_interpolate_point)
__device__ double eos_interp(const EosView &e, double x, double y)
{
  const int i = eos_find(e.x, e.nx, x), j = eos_find(e.y, e.ny, y);
  if (i < 0 || j < 0) return nan("");
  const int ny = e.ny;
  const double x0 = e.x[i], x1 = e.x[i + 1], y0 = e.y[j], y1 = e.y[j + 1];
  const double z00 = e.z[i * ny + j], z01 = e.z[i * ny + j + 1];
  const double z10 = e.z[(i + 1) * ny + j], z11 = e.z[(i + 1) * ny + j + 1];
  const double u = (x - x0) / (x1 - x0), v = (y - y0) / (y1 - y0);
  const int d = e.diag[i * (ny - 1) + j];
  if (d == 0) {
    if (v <= u) return z00 + u * (z10 - z00) + v * (z11 - z10);
    return z00 + u * (z11 - z01) + v * (z01 - z00);
  }
  if (u + v <= 1.0) return z00 + u * (z10 - z00) + v * (z01 - z00);
  return z11 + (1.0 - u) * (z01 - z11) + (1.0 - v) * (z10 - z11);
}

// The same interpolation with the x bracket (i, u) given and the y bracket searched from a hint: during
// the sub-step loop a cell's density is constant (x side evaluated once per MD step, eos_locate_x) and
// its Te moves by less than a table row per sub-step (the walk below ends after 0-1 steps).  Index
// semantics are eos_find's, so the result is bit-identical to eos_interp(e, x, y).
__device__ __forceinline__ void eos_locate_x(const EosView &e, double x, int &i, double &u)
{
  i = eos_find(e.x, e.nx, x);
  u = 0.0;
  if (i >= 0) { const double x0 = e.x[i], x1 = e.x[i + 1]; u = (x - x0) / (x1 - x0); }
}
__device__ __forceinline__ double eos_interp_at(const EosView &e, int i, double u, double y, int &jh)
{
  const int ny = e.ny;
  if (i < 0 || !(y >= e.y[0]) || !(y <= e.y[ny - 1])) return nan("");
  int j = jh < 0 ? 0 : (jh > ny - 2 ? ny - 2 : jh);
  while (j > 0 && y < e.y[j]) j--;
  while (j < ny - 2 && y >= e.y[j + 1]) j++;
  jh = j;
  const double y0 = e.y[j], y1 = e.y[j + 1];
  const double z00 = e.z[i * ny + j], z01 = e.z[i * ny + j + 1];
  const double z10 = e.z[(i + 1) * ny + j], z11 = e.z[(i + 1) * ny + j + 1];
  const double v = (y - y0) / (y1 - y0);
  const int d = e.diag[i * (ny - 1) + j];
  if (d == 0) {
    if (v <= u) return z00 + u * (z10 - z00) + v * (z11 - z10);
    return z00 + u * (z11 - z01) + v * (z01 - z00);
  }
  if (u + v <= 1.0) return z00 + u * (z10 - z00) + v * (z01 - z00);
  return z11 + (1.0 - u) * (z01 - z11) + (1.0 - v) * (z10 - z11);
}

__device__ __forceinline__ double eos_ee_from_r_te(const EosView &ee, double r, double t)
{
  return eos_interp(ee, r, sqrt(t));                                   // imd_ttm.c:2483-2529
}
__device__ __forceinline__ double eos_cve_from_r_te(const EosView &cve, double r, double t)
{
  double z = eos_interp(cve, r, t);                                    // imd_ttm.c:2531-2592
  z *= r; z *= 11604.5; z *= 1e-30; z *= C_J2eV;
  return z;
}
__device__ __forceinline__ double eeminfun(const EosView &ee, double x, double r, double e)
{
  double a = eos_ee_from_r_te(ee, r, x) - e;
  return DABS(a);
}

// fminbnd (fminbnd3.c:6-106) incl. the effect of the unparenthesised SIGNUM macro (makros.h:338)
__device__ double fminbnd_ref(const EosView &ee, double a, double b, double rho, double eng)
{
  const double tolx = 1e-5;
  const double c = 0.5 * (3 - sqrt(5.0));
  double v = a + c * (b - a), w = v, x = v, e = 0;
  double fv = eeminfun(ee, x, rho, eng), fw = fv, fval = fv, fu = 0.0;
  const int imax = 500;
  double xm, tol, r, q, p, d = 0.0, u;
  const double sqrteps = 2.220446049250313e-16;
  int dogs;
  for (int i = 0; i < imax + 1; i++) {
    xm = 0.5 * (a + b);
    tol = 2.0 * sqrteps * DABS(x) + tolx / 3.0;
    if (DABS(x - xm) <= (2 * tol - 0.5 * (b - a)) || i == imax) return x;
    if (DABS(e) > tol) {
      dogs = 0;
      r = (x - w) * (fval - fv);
      q = (x - v) * (fval - fw);
      p = (x - v) * q - (x - w) * r;
      q = 2 * (q - r);
      p *= (q > 0) ? 1 : ((q < 0) ? -1 : 0);
      q = DABS(q);
      r = e;
      e = d;
      if (DABS(p) < DABS(0.5 * q * r) && p > q * (a - x) && p < q * (b - x)) {
        d = p / q;
        u = x + d;
        if (DMIN(u - a, b - u) < 2 * tol) d = tol * ((xm - x > 0) ? 1 : (((xm - x < 0) ? -1 : 0) + (xm == x)));
      } else dogs = 1;
    } else dogs = 1;
    if (dogs == 1) {
      e = (x >= xm ? a - x : b - x);
      d = c * e;
    }
    u = x + DMAX(DABS(d), tol) * ((d > 0) ? 1 : (((d < 0) ? -1 : 0) + (d == 0)));
    fu = eeminfun(ee, u, rho, eng);
    if (fu < fval) {
      if (u < x) b = x; else a = x;
      v = w; fv = fw; w = x; fw = fval; x = u; fval = fu;
    } else {
      if (u < x) a = u; else b = u;
      if (fu <= fw || w == x) { v = w; fv = fw; w = u; fw = fu; }
      else if (fu <= fv || v == x || v == w) { v = u; fv = fu; }
    }
  }
  return -1;
}

// EOS_te_from_r_ee (imd_ttm.c:2699-2724)
__device__ double eos_te_from_r_ee(const EosView &ee, double r, double e)
{
  if (r < ee.x[0] || r > ee.x[ee.nx - 1]) { r = DMIN(r, ee.x[ee.nx - 1]); r = DMAX(r, ee.x[0]); }
  const double a = ee.y[0] * ee.y[0], b = ee.y[ee.ny - 1] * ee.y[ee.ny - 1];     // pow(y, 2.0) is exact y*y
  return fminbnd_ref(ee, a, b, r, e);
}

#ifndef TMM_SKIP
#define TMM_SKIP 0            /* timing experiments on the do_tmm sections, 0 = product */
#endif
#include "tmm.cuh"

// ----------------------------------------------------------------------------- device state
struct TtmScal {                     // persists between MD steps (device), mirrored to the host on demand
  int    cur, vcur;                  // which lattice is l1 / which virtual lattice is vlattice1
  int    last_active_cell_global, cur_vlattice_proc;
  int    diff_substeps, laser_active, err, pad;
  double max_dt_ttm, tmm_time, Eabs, tot_elec_energy;
};

struct TtmArgs {
  int G, P, L, LD, steps, ncell_tot;
  imdcu_ttm_params p;
  double timestep, fd_hx, fd_vol, neighvol, omega_laser, sigma_t_sq;
  EosView cve, ee;
  int eos_smem, lat_smem;            // EOS tables / the lattices are staged into dynamic shared memory
  TCell *lat, *vlat;                 // [2][P*LD], [2][vlatdim]
  int *isum; double *dsum;           // slab sums, see k_ttm_atoms1
  double *xiarr, *xiloc, *vcom, *src_e, *winv;   // winv: [ncell_tot][WINV_N], see fill_inv
  TtmScal *scal;
  // do_tmm (tmm.cuh)
  int use_tmm, pad2_;
  double tmm_threshold, laser_t_1, sigma_t1_sq;
  TmmTables tmm;
  EosView phase;                     // global memory
  double *tmm_buf;                   // [TMM_BUF_PER_CELL][G+2]: eps_re, eps_im, active, km(2), M(8), Bplus(2), Bminus(2), Qabs, phase flag
};
#define TMM_BUF_PER_CELL 19

// ----------------------------------------------------------------------------- atom kernels
// slab index of an atom (imd_ttm.c:277-279)
__device__ __forceinline__ int slab_of(double x, double fd_hx, int G)
{
  int ig = (int) (x / fd_hx);
  ig = ig < G - 1 ? ig : G - 1;
  ig = ig > 0 ? ig : 0;
  return ig;
}

#define TTM_ANT 256                  // threads per CTA in the atom -> slab kernels
#define TTM_AU 4                     // atoms per lane in flight

__device__ __forceinline__ int warp_sum_i(int v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// update_fd, first atom loop.  Atoms are cell-sorted (x slowest), so a contiguous range of atoms lies in
// one slab, or -- where a slab boundary cuts through a plane of cells -- alternates between TWO.  Every
// warp owns a contiguous range (lane-strided, coalesced) and every lane sums its atoms in registers, in
// two slots (current slab, previous slab); a third slab evicts the older slot by atomics.  At the end
// the warp reduces the slots of its two lowest slabs by shuffles, the CTA merges its warps in shared
// memory and ONE thread issues the atomics.  With a few CTAs per SM a slab address sees some thousand
// atomics per step.  (Measured: seven shared-memory atomics per atom, and later one set of global
// atomics per slab CHANGE of a lane, both cost 0.3 ms per 2 M atoms -- same-address atomics serialise.)
// isum = {natoms, totneighs, fluxl, fluxr}[G], dsum = {px, py, pz, mdtemp}[G]
struct SlabAcc { int ig, n, nn, fl, fr; double px, py, pz; };

__device__ __forceinline__ SlabAcc slab_fresh(int ig) { SlabAcc s = { ig, 0, 0, 0, 0, 0.0, 0.0, 0.0 }; return s; }

__device__ __forceinline__ void slab_flush(const SlabAcc &s, int G, int *isum, double *dsum)
{
  if (s.ig < 0 || s.n == 0) return;
  atomicAdd(&isum[s.ig], s.n); atomicAdd(&isum[G + s.ig], s.nn);
  if (s.fl) atomicAdd(&isum[2 * G + s.ig], s.fl);
  if (s.fr) atomicAdd(&isum[3 * G + s.ig], s.fr);
  atomicAdd(&dsum[s.ig], s.px); atomicAdd(&dsum[G + s.ig], s.py); atomicAdd(&dsum[2 * G + s.ig], s.pz);
}

__device__ __forceinline__ void slab_add(SlabAcc &r, const SlabAcc &q)
{
  r.n += q.n; r.nn += q.nn; r.fl += q.fl; r.fr += q.fr; r.px += q.px; r.py += q.py; r.pz += q.pz;
}

// make slot a the one of slab ig (b keeps the other slab; a third slab evicts b)
__device__ __forceinline__ void slab_select(SlabAcc &a, SlabAcc &b, int ig, int G, int *isum, double *dsum)
{
  if (ig == a.ig) return;
  if (ig != b.ig) { slab_flush(b, G, isum, dsum); b = slab_fresh(ig); }
  const SlabAcc t = a; a = b; b = t;
}

// the part of a lane's two slots that belongs to slab `target`, summed over the warp
__device__ __forceinline__ SlabAcc slab_warp_sum(const SlabAcc &a, const SlabAcc &b, int target)
{
  const bool ua = a.ig == target, ub = b.ig == target;
  SlabAcc t;
  t.ig = target;
  t.n = warp_sum_i(ua ? a.n : ub ? b.n : 0); t.nn = warp_sum_i(ua ? a.nn : ub ? b.nn : 0);
  t.fl = warp_sum_i(ua ? a.fl : ub ? b.fl : 0); t.fr = warp_sum_i(ua ? a.fr : ub ? b.fr : 0);
  t.px = warp_sum(ua ? a.px : ub ? b.px : 0.0); t.py = warp_sum(ua ? a.py : ub ? b.py : 0.0);
  t.pz = warp_sum(ua ? a.pz : ub ? b.pz : 0.0);
  return t;
}

__device__ __forceinline__ int warp_min_slab(int a, int b)
{
  int m = a >= 0 ? a : 0x7fffffff;
  if (b >= 0 && b < m) m = b;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const int v = __shfl_xor_sync(0xffffffffu, m, o); m = v < m ? v : m; }
  return m;                                              // 0x7fffffff: no lane holds an atom
}

// the contiguous atom range of this warp: [b, e)
__device__ __forceinline__ void warp_range(int n, long long &b, long long &e)
{
  const long long nw = (long long) gridDim.x * (TTM_ANT / 32);
  const long long per = ((n + nw - 1) / nw + 31) / 32 * 32;
  b = ((long long) blockIdx.x * (TTM_ANT / 32) + (threadIdx.x >> 5)) * per;
  e = b + per < n ? b + per : n;
}

__global__ void __launch_bounds__(TTM_ANT)
k_ttm_atoms1(int n, int G, double fd_hx, int steps, const double4 *__restrict__ pos4,
             const double4 *__restrict__ mom4, const int *__restrict__ type, const int *__restrict__ numneighs,
             int *__restrict__ fdi, int *__restrict__ isum, double *__restrict__ dsum)
{
  __shared__ SlabAcc s_w[2 * (TTM_ANT / 32)];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long b, e;
  warp_range(n, b, e);
  SlabAcc s = slab_fresh(-1), o = slab_fresh(-1);
  for (long long i0 = b + lane; i0 < e; i0 += 32 * TTM_AU) {
    int ty[TTM_AU], nn[TTM_AU], old[TTM_AU];
    double x[TTM_AU];
    double4 p[TTM_AU];
#pragma unroll
    for (int u = 0; u < TTM_AU; u++) {                   // all loads of the batch first
      const long long i = i0 + u * 32;
      const bool ok = i < e;
      ty[u] = ok ? type[i] : -1;
      x[u] = ok ? pos4[i].x : 0.0;
      p[u] = ok ? mom4[i] : make_double4(0.0, 0.0, 0.0, 1.0);
      nn[u] = ok ? numneighs[i] : 0;
      old[u] = ok && steps > 0 ? fdi[i] : 0;
    }
#pragma unroll
    for (int u = 0; u < TTM_AU; u++) {
      if (ty[u] != 0) continue;
      const int ig = slab_of(x[u], fd_hx, G);
      slab_select(s, o, ig, G, isum, dsum);
      if (steps > 0) { if (old[u] > ig) s.fr++; else if (old[u] < ig) s.fl++; }
      fdi[i0 + u * 32] = ig;
      s.n++; s.nn += nn[u];
      s.px += p[u].x; s.py += p[u].y; s.pz += p[u].z;
    }
  }
  // warp: the two lowest slabs by shuffle reduction, whatever else a lane holds by its own atomics
  const int lo = warp_min_slab(s.ig, o.ig), hi = lo == 0x7fffffff ? lo : lo + 1;
  if (s.ig >= 0 && s.ig != lo && s.ig != hi) slab_flush(s, G, isum, dsum);
  if (o.ig >= 0 && o.ig != lo && o.ig != hi) slab_flush(o, G, isum, dsum);
  SlabAcc t0 = slab_warp_sum(s, o, lo), t1 = slab_warp_sum(s, o, hi);
  if (lo == 0x7fffffff) t0.ig = t1.ig = -1;
  if (lane == 0) { s_w[2 * warp] = t0; s_w[2 * warp + 1] = t1; }
  __syncthreads();
  // CTA: warp ranges are consecutive in atom order, so the same two-slot merge works over the warps
  if (threadIdx.x == 0) {
    SlabAcc r = slab_fresh(-1), q = slab_fresh(-1);
    for (int w = 0; w < 2 * (TTM_ANT / 32); w++) {
      const SlabAcc v = s_w[w];
      if (v.ig < 0 || v.n == 0) continue;
      slab_select(r, q, v.ig, G, isum, dsum);
      slab_add(r, v);
    }
    slab_flush(r, G, isum, dsum); slab_flush(q, G, isum, dsum);
  }
}

// update_fd, second atom loop: md_temp sums with v_com = sum p / (natoms * atomic_weight); same scheme
struct TempAcc { int ig; double t, vx, vy, vz; };

__device__ __forceinline__ void temp_select(TempAcc &a, TempAcc &b, int ig, int G, double atomic_weight,
                                            const int *isum, double *dsum)
{
  if (ig == a.ig) return;
  if (ig != b.ig) {
    if (b.ig >= 0) atomicAdd(&dsum[3 * G + b.ig], b.t);
    const double totmass = (double) isum[ig] * atomic_weight;        // > 0: an atom is in the slab
    b.ig = ig; b.t = 0.0;
    b.vx = dsum[ig] / totmass; b.vy = dsum[G + ig] / totmass; b.vz = dsum[2 * G + ig] / totmass;
  }
  const TempAcc t = a; a = b; b = t;
}

__global__ void __launch_bounds__(TTM_ANT)
k_ttm_atoms2(int n, int G, double fd_hx, double atomic_weight, const double4 *__restrict__ mom4,
             const int *__restrict__ type, const int *__restrict__ fdi, const int *__restrict__ isum,
             double *__restrict__ dsum)
{
  __shared__ int s_ig[2 * (TTM_ANT / 32)];
  __shared__ double s_t[2 * (TTM_ANT / 32)];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  long long b, e;
  warp_range(n, b, e);
  TempAcc s = { -1, 0.0, 0.0, 0.0, 0.0 }, o = s;
  for (long long i0 = b + lane; i0 < e; i0 += 32 * TTM_AU) {
    int ty[TTM_AU], g[TTM_AU];
    double4 p[TTM_AU];
#pragma unroll
    for (int u = 0; u < TTM_AU; u++) {
      const long long i = i0 + u * 32;
      const bool ok = i < e;
      ty[u] = ok ? type[i] : -1;
      g[u] = ok ? fdi[i] : 0;                            // set by k_ttm_atoms1 from the same positions
      p[u] = ok ? mom4[i] : make_double4(0.0, 0.0, 0.0, 1.0);
    }
#pragma unroll
    for (int u = 0; u < TTM_AU; u++) {
      if (ty[u] != 0) continue;
      temp_select(s, o, g[u], G, atomic_weight, isum, dsum);
      const double m = p[u].w;
      double d, t = 0.0;
      d = p[u].x / m - s.vx; t += m * (d * d);
      d = p[u].y / m - s.vy; t += m * (d * d);
      d = p[u].z / m - s.vz; t += m * (d * d);
      s.t += t;
    }
  }
  const int lo = warp_min_slab(s.ig, o.ig), hi = lo == 0x7fffffff ? lo : lo + 1;
  if (s.ig >= 0 && s.ig != lo && s.ig != hi) atomicAdd(&dsum[3 * G + s.ig], s.t);
  if (o.ig >= 0 && o.ig != lo && o.ig != hi) atomicAdd(&dsum[3 * G + o.ig], o.t);
  const double t0 = warp_sum(s.ig == lo ? s.t : o.ig == lo ? o.t : 0.0);
  const double t1 = warp_sum(s.ig == hi ? s.t : o.ig == hi ? o.t : 0.0);
  const unsigned h1 = __ballot_sync(0xffffffffu, s.ig == hi || o.ig == hi);
  if (lane == 0) {
    s_ig[2 * warp] = lo == 0x7fffffff ? -1 : lo; s_t[2 * warp] = t0;
    s_ig[2 * warp + 1] = lo == 0x7fffffff || !h1 ? -1 : hi; s_t[2 * warp + 1] = t1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int ia = -1, ib = -1;
    double ta = 0.0, tb = 0.0;
    for (int w = 0; w < 2 * (TTM_ANT / 32); w++) {
      const int ig = s_ig[w];
      if (ig < 0) continue;
      if (ig != ia) {
        if (ig != ib) { if (ib >= 0) atomicAdd(&dsum[3 * G + ib], tb); ib = ig; tb = 0.0; }
        const int ii = ia; ia = ib; ib = ii;
        const double tt = ta; ta = tb; tb = tt;
      }
      ta += s_t[w];
    }
    if (ia >= 0) atomicAdd(&dsum[3 * G + ia], ta);
    if (ib >= 0) atomicAdd(&dsum[3 * G + ib], tb);
  }
}

// ----------------------------------------------------------------------------- grid kernel
#define TTM_NT 256

__device__ __forceinline__ bool active_gt(const TCell &c, int mn) { return c.natoms >= mn && c.dens > RHOMIN; }
__device__ __forceinline__ bool active_ge(const TCell &c, int mn) { return c.natoms >= mn && c.dens >= RHOMIN; }

// block-wide reductions (deterministic tree); every thread of the CTA must call
__device__ double block_reduce(double v, int op /*0 sum, 1 min*/, double *s_red)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double t = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 0 ? v + t : fmin(v, t);
  }
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  double r = op == 0 ? 0.0 : 1e300;
  for (int k = 0; k < nw; k++) r = op == 0 ? r + s_red[k] : fmin(r, s_red[k]);
  return r;
}
__device__ int block_reduce_int(int v, int op /*1 min, 2 max*/, int *s_red)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int t = __shfl_xor_sync(0xffffffffu, v, o);
    v = op == 1 ? min(v, t) : max(v, t);
  }
  __syncthreads();
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  int r = s_red[0];
  for (int k = 1; k < nw; k++) r = op == 1 ? min(r, s_red[k]) : max(r, s_red[k]);
  return r;
}

// per cell and MD step: what depends on the cell's density only (constant during the sub-step loop):
// the density terms of the wide-range models and the x bracket of the Ce table
#define WINV_N 8
__device__ __forceinline__ void fill_inv(const TtmArgs &a, const EosView &cve, const TCell &c, double *wi)
{
  const WideRangeInv v = wide_range_inv(c.dens, a.p.atomic_charge, a.p.atomic_weight);
  int i; double u;
  eos_locate_x(cve, c.dens, i, u);
  wi[0] = v.lr0; wi[1] = v.ni; wi[2] = v.r0n; wi[3] = v.g_r0n; wi[4] = v.t_r0n;
  wi[5] = (double) i; wi[6] = u; wi[7] = 0.0;
}

// wide-range properties of one cell (the body of do_FILLMESH's active branch / the vlattice loop)
__device__ __forceinline__ void fill_props(const TtmArgs &a, const EosView &cve, TCell &c, int la, double z0p43,
                                           double *wi)
{
  WideRangeInv v;
  v.lr0 = wi[0]; v.ni = wi[1]; v.r0n = wi[2]; v.g_r0n = wi[3]; v.t_r0n = wi[4];
  int jh = (int) wi[7];
  double ce = eos_interp_at(cve, (int) wi[5], wi[6], c.temp * 11604.5, jh);       // EOS_cve_from_r_te, imd_ttm.c:2531-2592
  wi[7] = (double) jh;
  ce *= c.dens; ce *= 11604.5; ce *= 1e-30; ce *= C_J2eV;
  const WideRange w = wide_range(c.temp, c.md_temp, v, a.p.atomic_charge, z0p43, a.omega_laser, la);
  c.Z = w.Z; c.ne = w.ne; c.Ce = ce; c.fd_k = w.kappa; c.fd_g = w.gamma;
}

// do_FILLMESH (imd_ttm.c:629-852) for the calling thread's cells + the virtual lattice.  Lattice
// cells and virtual cells share ONE call site of fill_props, so a warp holding both kinds does
// not run the expensive part twice.
__device__ __forceinline__ void fillmesh(const TtmArgs &a, const EosView &cve, const EosView &ee, double z0p43,
                                         TCell *l1, TCell *l2, TCell *v1, TCell *v2, int last_active,
                                         int rv, int la, int *err, bool first)
{
  const int ncell = a.P * a.L, mn = a.p.fd_min_atoms;
  for (int c = threadIdx.x; c < ncell + a.p.vlatdim; c += blockDim.x) {
    TCell *t = nullptr, *t2 = nullptr;
    const bool virt = c >= ncell;
    if (!virt) {
      if (c <= last_active) {
        const int idx = (c / a.L) * a.LD + (c % a.L) + 1;
        if (active_gt(l1[idx], mn)) { t = &l1[idx]; t2 = &l2[idx]; }
        else { TCell &x = l1[idx]; x.fd_k = 0.0; x.fd_g = 0.0; x.Z = 0.0; x.ne = 0.0; x.Ce = 0.0; x.temp = 0.0; }
      }
    } else if (rv >= 0 && rv < a.P) { t = &v1[c - ncell]; t2 = &v2[c - ncell]; }   // :825-845
    if (!t) continue;
    double *wi = a.winv + (size_t) c * WINV_N;
    if (first) fill_inv(a, cve, *t, wi);                                   // first FILLMESH of this MD step
    fill_props(a, cve, *t, la, z0p43, wi);
    TCell &x = *t, &y = *t2;
    if (virt) { y.Z = x.Z; y.ne = x.ne; y.fd_k = x.fd_k; y.fd_g = x.fd_g; continue; }
    y.ne = x.ne; y.Ce = x.Ce; y.fd_k = x.fd_k; y.fd_g = x.fd_g;
    if (x.Ce < 0 || isnan(x.Ce)) atomicMax(err, 1);
    if (isnan(x.fd_k) || x.fd_k < 0) atomicMax(err, 2);
    if (isnan(x.fd_g) || x.fd_g < 0) atomicMax(err, 3);
    if (a.steps < 1) {                                                     // :765-807
      x.U = eos_ee_from_r_te(ee, x.dens, x.temp * 11604.5) * 26.9815 * C_AMU * C_J2eV;
      y.U = x.U;
      const double echeck = eos_ee_from_r_te(ee, x.dens, x.temp * 11604.5) * 26.9815 * C_AMU * C_J2eV;
      const double tcheck = eos_te_from_r_ee(ee, x.dens, echeck / 26.9815 / C_AMU * C_eV2J) / 11604.5;
      const double tinit = x.temp;
      if (DABS(tcheck - tinit) > tinit * 0.01) atomicMax(err, 4);
    }
  }
}

// ttm_fill_ghost_layers (imd_ttm.c:3264-3356): one thread per emulated rank
__device__ void fill_ghosts(const TtmArgs &a, TCell *l1)
{
  const int L = a.L, LD = a.LD, P = a.P;
  for (int r = threadIdx.x; r < P; r += blockDim.x) {
    if (r < P - 1) { l1[r * LD + L + 1].temp = l1[(r + 1) * LD + 1].temp; l1[r * LD + L + 1].fd_k = l1[(r + 1) * LD + 1].fd_k; }
    if (r > 0) { l1[r * LD].temp = l1[(r - 1) * LD + L].temp; l1[r * LD].fd_k = l1[(r - 1) * LD + L].fd_k; }
    if (r == 0) l1[r * LD].natoms = 0;
    if (r == P - 1) l1[r * LD + L + 1].natoms = 0;
  }
}

__global__ void __launch_bounds__(TTM_NT)
k_ttm_grid(TtmArgs a)
{
  extern __shared__ double s_dyn[];
  __shared__ double s_red[32];
  __shared__ int s_ired[32];
  __shared__ int s_last_active, s_rv, s_substeps, s_ecut;
  __shared__ double s_maxdt;
  const int G = a.G, P = a.P, L = a.L, LD = a.LD, ncell = P * L, mn = a.p.fd_min_atoms;
  const int nl = P * LD;
  TtmScal *sc = a.scal;
  int cur = sc->cur, vcur = sc->vcur;
  TCell *l1 = a.lat + cur * nl, *l2 = a.lat + (1 - cur) * nl;
  TCell *v1 = a.vlat + vcur * a.p.vlatdim, *v2 = a.vlat + (1 - vcur) * a.p.vlatdim;
  const int *natomsg = a.isum, *totneigh = a.isum + G, *fluxl = a.isum + 2 * G, *fluxr = a.isum + 3 * G;
  int *err = &sc->err;
  // EOS tables: staged into shared memory when they fit (the binary searches and the four corner
  // loads of every interpolation are dependent loads on the critical path of each sub-step)
  EosView cve = a.cve, ee = a.ee;
  size_t dyn_used = 0;
  if (a.eos_smem) {
    double *d = s_dyn;
    EosView *vs[2] = { &cve, &ee };
    for (int w = 0; w < 2; w++) {
      EosView &v = *vs[w];
      const int nd = v.nx + v.ny + v.nx * v.ny;
      for (int k = threadIdx.x; k < nd; k += blockDim.x) d[k] = v.x[k];       // x | y | z are contiguous
      v.x = d; v.y = d + v.nx; v.z = d + v.nx + v.ny;
      d += nd;
    }
    unsigned char *b = reinterpret_cast<unsigned char *>(d);
    for (int w = 0; w < 2; w++) {
      EosView &v = *vs[w];
      const int ng = (v.nx - 1) * (v.ny - 1);
      for (int k = threadIdx.x; k < ng; k += blockDim.x) b[k] = v.diag[k];
      v.diag = b;
      b += (ng + 7) & ~7;
    }
    __syncthreads();
    dyn_used = (size_t) (b - reinterpret_cast<unsigned char *>(s_dyn));
  }
  // Both lattices and both virtual lattices live in shared memory for the whole launch: every sub-step reads the
  // cells the previous one wrote, and written lines are not kept in L1, so from global memory each of those
  // reads is an L2 round trip on the critical path (long-scoreboard stalls 5.2 per issue at G = 207).
  TCell *g_lat = a.lat, *g_vlat = a.vlat;
  double *xiloc = a.xiloc, *tmm_base = a.tmm_buf;
  const int nvl = a.p.vlatdim > 0 ? a.p.vlatdim : 1;
  if (a.lat_smem) {
    TCell *sl = reinterpret_cast<TCell *>(reinterpret_cast<unsigned char *>(s_dyn) + ((dyn_used + 15) & ~(size_t) 15));
    TCell *sv = sl + 2 * nl;
    for (int k = threadIdx.x; k < 2 * nl; k += blockDim.x) sl[k] = g_lat[k];
    for (int k = threadIdx.x; k < 2 * nvl; k += blockDim.x) sv[k] = g_vlat[k];
    l1 = sl + cur * nl; l2 = sl + (1 - cur) * nl;
    v1 = sv + vcur * a.p.vlatdim; v2 = sv + (1 - vcur) * a.p.vlatdim;
    g_lat = sl; g_vlat = sv;                     // base pointers of whichever copy is live
    // per-cell xi sums of this step (thread private in effect, but read-modify-write in every sub-step) and the
    // scratch arrays of do_tmm; of the latter only tmm_Qabs outlives the launch
    xiloc = reinterpret_cast<double *>(sv + 2 * nvl);
    if (a.use_tmm) {
      tmm_base = xiloc + nl;
      double *qs = tmm_base + 17 * (size_t) (a.G + 2);
      const double *qg = a.tmm_buf + 17 * (size_t) (a.G + 2);
      for (int k = threadIdx.x; k < a.G + 2; k += blockDim.x) qs[k] = qg[k];
    }
    __syncthreads();
  }
  const double z0p43 = pow(a.p.atomic_charge, 4. / 3.);                       // MeanCharge's pow(z0, 4/3)

  // ---------------------------------------------------------------- update_fd, grid part (:241-264, 341-363, 401-541)
  for (int r = threadIdx.x; r < P; r += blockDim.x) l1[r * LD].temp = l1[r * LD + L + 1].temp = 0;
  int last_local = -5000;
  for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
    const int r = c / L, i = c % L + 1, g = c, idx = r * LD + i;
    TCell &x = l1[idx], &y = l2[idx];
    const double totmass = (double) natomsg[g] * a.p.atomic_weight;
    if (totmass > 0) { a.vcom[g] = a.dsum[g] / totmass; a.vcom[G + g] = a.dsum[G + g] / totmass; a.vcom[2 * G + g] = a.dsum[2 * G + g] / totmass; }
    else a.vcom[g] = a.vcom[G + g] = a.vcom[2 * G + g] = 0.0;
    a.xiarr[g] = 0.0;
    x.source = y.source = 0.0;
    x.natoms_old = y.natoms_old = x.natoms;
    x.xi = y.xi = 0.0;
    xiloc[idx] = 0.0;
    x.natoms = natomsg[g]; y.natoms = x.natoms;
    if (x.natoms > 0) {
      x.dens = ((double) totneigh[g]) / ((double) x.natoms) * a.p.atomic_weight / a.neighvol * 1660.53907;
      if (x.dens == 0) x.dens = (double) x.natoms * a.p.atomic_weight / a.fd_vol * 1660.53907;
      x.md_temp = a.dsum[3 * G + g];
      x.md_temp /= (3.0 * (double) x.natoms);
      if (active_gt(x, mn)) last_local = max(last_local, g - a.p.vlatbuffer);
    } else { x.dens = 0.0; x.md_temp = 0.0; x.natoms = y.natoms = 0; }
    y.md_temp = x.md_temp; y.dens = x.dens;
    if (i == 1) {
      TCell &ga = l1[r * LD], &gb = l2[r * LD];
      ga.natoms = g - 1 >= 0 ? natomsg[g - 1] : 0;
      if (ga.natoms > 0) ga.dens = ((double) totneigh[g - 1]) / ((double) ga.natoms) * a.p.atomic_weight / a.neighvol * 1660.53907;
      gb.natoms = ga.natoms; gb.dens = ga.dens;
    } else if (i == L) {
      TCell &ga = l1[r * LD + L + 1], &gb = l2[r * LD + L + 1];
      ga.natoms = g + 1 < G ? natomsg[g + 1] : 0;
      if (ga.natoms > 0) ga.dens = ((double) totneigh[g + 1]) / ((double) ga.natoms) * a.p.atomic_weight / a.neighvol * 1660.53907;
      gb.natoms = ga.natoms; gb.dens = ga.dens;
    }
    if (a.steps < 1) {
      x.natoms_old = y.natoms_old = x.natoms;
      if (active_gt(x, mn)) x.temp = y.temp = x.md_temp;
    }
  }
  last_local = block_reduce_int(last_local, 2, s_ired);
  if (threadIdx.x == 0) {
    if (a.steps == 0) sc->last_active_cell_global = last_local;
    s_last_active = sc->last_active_cell_global;
    s_rv = (int) (((double) s_last_active * P) / ((double) G));
    sc->cur_vlattice_proc = s_rv;
  }
  __syncthreads();
  const int last_active = s_last_active, rv = s_rv;
  for (int c = threadIdx.x; c < ncell; c += blockDim.x)
    if (c > last_active) { const int idx = (c / L) * LD + c % L + 1; l1[idx].natoms = l2[idx].natoms = -1; }
  __syncthreads();

  // ---------------------------------------------------------------- do_ADV (:1208-1301)
  if (a.steps >= 1) {
    for (int r = threadIdx.x; r < P; r += blockDim.x) {                   // do_COMMFLUX
      if (r > 0) l1[r * LD].U = l1[(r - 1) * LD + L].U;
      if (r < P - 1) l1[r * LD + L + 1].U = l1[(r + 1) * LD + 1].U;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
      const int g = c, idx = (c / L) * LD + c % L + 1;
      const TCell &x = l1[idx]; TCell &y = l2[idx];
      const double Nold = (double) x.natoms_old, Nnew = (double) x.natoms;
      if (Nnew > 0) {
        const double fr_g = (double) fluxr[g], fl_g = (double) fluxl[g];
        const double fl_g1 = g + 1 < G ? (double) fluxl[g + 1] : 0.0;
        const double fr_gm = g - 1 >= 0 ? (double) fluxr[g - 1] : 0.0;
        y.U = x.U * Nold / Nnew + 1.0 * (+fr_g * l1[idx + 1].U - fl_g1 * x.U + fl_g * l1[idx - 1].U - fr_gm * x.U) / Nnew;
        if (active_gt(x, mn)) y.temp = eos_te_from_r_ee(ee, x.dens, y.U / (26.9815 * C_AMU * C_J2eV)) / 11604.5;
      } else if (Nnew < 1) { y.U = 0.0; y.temp = 0.0; }
    }
    __syncthreads();
    cur = 1 - cur;
    { TCell *t = l1; l1 = l2; l2 = t; }
  }

  // ---------------------------------------------------------------- do_cell_activation (:1311-1425), in cell order
  if (threadIdx.x == 0) {
    for (int c = 0; c < ncell; c++) {
      const int idx = (c / L) * LD + c % L + 1;
      TCell &x = l1[idx], &y = l2[idx];
      if (x.natoms_old >= mn && x.natoms < mn && x.dens > RHOMIN) x.xi = 0.0;
      else if (x.natoms_old < mn && x.natoms >= mn && x.dens > RHOMIN) {
        if (c > last_active) continue;
        if (isnan(x.temp) || x.temp <= a.p.Temin) {
          int nn = 0; double e_nb = 0.0;
          const TCell &xp = l1[idx + 1], &xm = l1[idx - 1];
          if (active_ge(xp, mn)) { e_nb += xp.temp * xp.temp; nn++; }
          if (active_ge(xm, mn)) { e_nb += xm.temp * xm.temp; nn++; }
          if (nn != 0) {
            x.temp = sqrt(e_nb / ((double) nn)); y.temp = x.temp;
            if (x.temp < a.p.Temin) y.temp = x.temp = x.md_temp;
          } else { x.temp = x.md_temp; y.temp = x.temp; }
          x.U = y.U = eos_ee_from_r_te(ee, x.dens, x.temp * 11604.5) * 26.9815 * C_AMU * C_J2eV;
        }
      }
    }
  }
  __syncthreads();

  // ---------------------------------------------------------------- do_FILLMESH, ghosts, CFL_maxdt (:103-108)
  int la = sc->laser_active;
  fillmesh(a, cve, ee, z0p43, l1, l2, v1, v2, last_active, rv, la, err, true);
  __syncthreads();
  fill_ghosts(a, l1);
  __syncthreads();
  {
    double maxdttmp = 1e9;
    const double dxsq = a.fd_hx * a.fd_hx;
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
      const int idx = (c / L) * LD + c % L + 1;
      const TCell &x = l1[idx];
      if (active_gt(x, mn)) {
        const TCell &xp = l1[idx + 1];
        const double khalf = x.fd_k + (active_ge(xp, mn) ? xp.fd_k : x.fd_k);
        maxdttmp = fmin(maxdttmp, x.Ce * dxsq / khalf);
      }
    }
    maxdttmp = block_reduce(maxdttmp, 1, s_red);
    if (threadIdx.x == 0) {
      double m = a.steps > 0 ? maxdttmp : a.timestep / ((double) a.p.fd_n_timesteps);
      m *= 0.9;
      s_maxdt = m;
      const int q = (int) (a.timestep / m);
      s_substeps = a.p.fd_n_timesteps > q ? a.p.fd_n_timesteps : q;
    }
    __syncthreads();
  }
  const int diff_substeps = s_substeps;
  const double tau = a.timestep / ((double) diff_substeps);

  // first active cell: depth origin of the source; natoms/dens do not change during the sub-steps
  int first = 1 << 30;
  for (int c = threadIdx.x; c < ncell; c += blockDim.x)
    if (active_ge(l1[(c / L) * LD + c % L + 1], mn)) first = min(first, c);
  first = block_reduce_int(first, 1, s_ired);
  for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
    const int idx = (c / L) * LD + c % L + 1;
    a.src_e[idx] = active_ge(l1[idx], mn) ? exp(-a.p.src_mu * ((double) (c - first) * a.fd_hx)) : -1.0;
  }

  // ---------------------------------------------------------------- sub-step loop (:111-121)
  double tmm_time = sc->tmm_time, eabs = 0.0;
  const double xi_fac = 26.9815 * C_AMU / 3.0 * 1e30 / ((double) diff_substeps);
  const double invxsq = 1.0 / (a.fd_hx * a.fd_hx);
  const double Imp = 3.769911184307751e+02;
  const double la_threshold = a.p.tmm_laser_threshold * sqrt(2 * a.p.I0 * Imp);      // imd_tmm.c:200-208, loop invariant
  double Ci = 8.589449886e+02;
  Ci *= a.p.vlatdens; Ci *= 11604.5; Ci *= 1e-30; Ci *= C_J2eV;
  const double tau_ci = tau / Ci;
  // do_tmm scratch (global memory, L1/L2 resident): see TtmArgs::tmm_buf
  const int GG = a.G + 2;
  double *q_re = tmm_base, *q_im = q_re + GG, *q_act = q_im + GG, *q_km = q_act + GG, *q_M = q_km + 2 * GG,
         *q_Bp = q_M + 8 * GG, *q_Bm = q_Bp + 2 * GG, *q_Q = q_Bm + 2 * GG;
  // the vacuum in front of the sample is "cell 0" with dx = 1e10 A (imd_tmm.c:297): its phase factor
  // exp(i k0 1e10) -- a sincos of 7.85e6 rad -- is the same in every sub-step
  cplx eiphi0 = cmk(1.0, 0.0), eps_bb = cmk(0.0, 0.0);
  double *q_ph = q_Q + GG;           // 1 where the phase table removes the interband term
  if (a.use_tmm) {
    double k0 = 2.0 * TMC_PI / a.p.lambda; k0 /= 1e10;
    eiphi0 = tmm_eiphi(1e10, cmk(k0, 0.0));
    eps_bb = tmm_eps_bb(a.tmm, a.p.lambda);
    // phase = (int) EOS_phase_from_r_ti(Ni 26.9815 AMU, Ti [K]) (:548): Ni = ne/Z = dens/(A AMU) and Ti do not change
    // during the sub-steps, so the lookup is made once per MD step
    for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
      const TCell &x = l1[(c / L) * LD + c % L + 1];
      double flag = 0.0;
      if (!(x.natoms < mn || x.dens < RHOMIN)) {
        const double Ni = x.ne / x.Z;
        double pr = Ni * 26.9815 * C_AMU, pt = x.md_temp * 11604.52500617;      // EOS_phase_from_r_ti clamps (DEBUG_LEVEL > 0)
        pr = DMAX(DMIN(pr, a.phase.x[a.phase.nx - 1]), a.phase.x[0]);
        pt = DMAX(DMIN(pt, a.phase.y[a.phase.ny - 1]), a.phase.y[0]);
        int ph = (int) eos_interp(a.phase, pr, pt);
        ph = ph < 0 ? -ph : ph;
        flag = (ph == 3 || ph == 4 || ph == 5) ? 1.0 : 0.0;
      }
      q_ph[c] = flag;
    }
    __syncthreads();
  }
  for (int s = 0; s < diff_substeps; s++) {
    double I_t;
    if (!a.use_tmm) {
      // analytic source (stands in for do_tmm where its data files are missing): I_t and the laser_active
      // switch of imd_tmm.c:197-208
      const double dtl = tmm_time - a.p.laser_t_0;
      I_t = a.p.I0 * pexp(pdiv(-(dtl * dtl), a.sigma_t_sq));
      la = !(sqrt(2.0 * I_t * Imp) < la_threshold);
    } else {
      // do_tmm (imd_tmm.c:193-445).  Nothing happens for steps < 2; below the field threshold the sources keep
      // their last values; only tmm_Qabs[0, ecut) is rewritten.
      const double d0 = tmm_time - a.p.laser_t_0, d1 = tmm_time - a.laser_t_1;
      I_t = a.p.I0 * exp(-(d0 * d0) / a.sigma_t_sq);
      I_t += a.p.I0 * exp(-(d1 * d1) / a.sigma_t1_sq);
      if (a.steps >= 2) {
        if (sqrt(2.0 * I_t * Imp) < la_threshold) la = 0;
        else {
          la = 1;
          const int G = a.G;
          double k0 = 2.0 * TMC_PI / a.p.lambda;
          k0 /= 1e10;
          // permittivity of the active cells (every rank its own cells, then MPI_Reduce, :209-252)
          for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
            const TCell &x = l1[(c / L) * LD + c % L + 1];
            if (x.natoms < mn || x.dens < RHOMIN) { q_re[c] = 0.0; q_im[c] = 0.0; q_act[c] = 0.0; continue; }
#if TMM_SKIP == 1
            const cplx e = cmk(-40.0 - x.temp, 12.0 + x.temp);
#else
            const cplx e = tmm_get_epsilon(a.tmm, a.p.lambda, x.temp, x.md_temp, x.Z, x.ne,
                                           q_ph[c] != 0.0 ? cmk(0.0, 0.0) : eps_bb);
#endif
            q_re[c] = e.re; q_im[c] = e.im; q_act[c] = 1.0;
          }
          __syncthreads();
          for (int c = threadIdx.x; c < G; c += blockDim.x) {                      // km (:278-287); km[G-1] stays 0
            cplx k = cmk(k0, 0.0);
            if (c == G - 1) k = cmk(0.0, 0.0);
            else if (c > 0 && q_act[c] != 0.0) k = cscale(k0, csqrtc(cmk(q_re[c], q_im[c])));
            q_km[2 * c] = k.re; q_km[2 * c + 1] = k.im;
          }
          __syncthreads();
          // Cinv * Pmat of every interface, in parallel: the transcendental part of loops 2 and 3
          for (int i = threadIdx.x; i < G - 1; i += blockDim.x) {
            cplx M[4];
            const cplx kl = cmk(q_km[2 * i], q_km[2 * i + 1]);
#if TMM_SKIP == 2
            M[0] = kl; M[1] = cmk(0.5, 0.1); M[2] = cmk(0.1, 0.2); M[3] = cmk(0.9, 0.0);
#else
            tmm_interface(i == 0 ? eiphi0 : tmm_eiphi(a.fd_hx, kl), kl, cmk(q_km[2 * i + 2], q_km[2 * i + 3]), M);
#endif
#pragma unroll
            for (int j = 0; j < 4; j++) { q_M[8 * i + 2 * j] = M[j].re; q_M[8 * i + 2 * j + 1] = M[j].im; }
          }
          __syncthreads();
          if (threadIdx.x < 32) {
            // loops 2 and 3 (:292-334, :353-385) are prefix products E_i = M_i ... M_0 of the interface matrices: warp 0
            // scans them 32 cells at a time (five shuffle steps instead of 32 dependent 2x2 complex products), tests
            // |BT_i|^2 = |det E_i|^2 / |E_i[3]|^2 < exp(-tmm_threshold) on every lane, and gets the amplitudes of loop 3
            // as E_i (B0, BR)^T.  Same products as the reference's recurrences, associated as a tree.
            const int lane = threadIdx.x;
            const double ecut_thresh = exp(-a.tmm_threshold);
            cplx C[4] = { cmk(1.0, 0.0), cmk(0.0, 0.0), cmk(0.0, 0.0), cmk(1.0, 0.0) };      // carry: product of earlier chunks
            cplx BR = cmk(0.0, 0.0);
            int ecut = -1;
            for (int base = 0; base < G - 1 && ecut < 0; base += 32) {
              const int i = base + lane;
              const bool valid = i < G - 1;
              cplx Pm[4] = { cmk(1.0, 0.0), cmk(0.0, 0.0), cmk(0.0, 0.0), cmk(1.0, 0.0) };
              if (valid) {
                const double *m = q_M + 8 * i;
                Pm[0] = cmk(m[0], m[1]); Pm[1] = cmk(m[2], m[3]); Pm[2] = cmk(m[4], m[5]); Pm[3] = cmk(m[6], m[7]);
              }
#pragma unroll
              for (int o = 1; o < 32; o <<= 1) {
                cplx O[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                  O[j].re = __shfl_up_sync(0xffffffffu, Pm[j].re, o);
                  O[j].im = __shfl_up_sync(0xffffffffu, Pm[j].im, o);
                }
                if (lane >= o) {                                                       // P_i <- P_i * P_(i-o)
                  const cplx n0 = cadd(cmul(Pm[0], O[0]), cmul(Pm[2], O[1])), n1 = cadd(cmul(Pm[1], O[0]), cmul(Pm[3], O[1]));
                  const cplx n2 = cadd(cmul(Pm[0], O[2]), cmul(Pm[2], O[3])), n3 = cadd(cmul(Pm[1], O[2]), cmul(Pm[3], O[3]));
                  Pm[0] = n0; Pm[1] = n1; Pm[2] = n2; Pm[3] = n3;
                }
              }
              // E_i = P_i * carry
              const cplx E0 = cadd(cmul(Pm[0], C[0]), cmul(Pm[2], C[1])), E1 = cadd(cmul(Pm[1], C[0]), cmul(Pm[3], C[1]));
              const cplx E2 = cadd(cmul(Pm[0], C[2]), cmul(Pm[2], C[3])), E3 = cadd(cmul(Pm[1], C[2]), cmul(Pm[3], C[3]));
              const cplx det = csub(cmul(E0, E3), cmul(E2, E1));
              const bool hit = valid && cabs2(det) < ecut_thresh * cabs2(E3);
              const unsigned mask = __ballot_sync(0xffffffffu, hit);
              const int src = mask ? __ffs(mask) - 1 : 31;
              if (valid) {                                                              // E_i replaces M_i (read back below)
                double *m = q_M + 8 * i;
                m[0] = E0.re; m[1] = E0.im; m[2] = E1.re; m[3] = E1.im; m[4] = E2.re; m[5] = E2.im; m[6] = E3.re; m[7] = E3.im;
              }
              const cplx br = cdivc(cneg(E1), E3);                                      // -Emat[1] B0 / Emat[3], B0 = 1
              BR.re = __shfl_sync(0xffffffffu, br.re, src); BR.im = __shfl_sync(0xffffffffu, br.im, src);
              if (mask) ecut = base + src + 1;
              else {
                C[0].re = __shfl_sync(0xffffffffu, E0.re, 31); C[0].im = __shfl_sync(0xffffffffu, E0.im, 31);
                C[1].re = __shfl_sync(0xffffffffu, E1.re, 31); C[1].im = __shfl_sync(0xffffffffu, E1.im, 31);
                C[2].re = __shfl_sync(0xffffffffu, E2.re, 31); C[2].im = __shfl_sync(0xffffffffu, E2.im, 31);
                C[3].re = __shfl_sync(0xffffffffu, E3.re, 31); C[3].im = __shfl_sync(0xffffffffu, E3.im, 31);
              }
            }
            if (ecut < 0) {                                                             // "Sample is too short for TMM"
              if (lane == 0) atomicMax(err, 6);
              ecut = 0;
            }
            __syncwarp();
            if (lane == 0) { q_Bp[0] = 1.0; q_Bp[1] = 0.0; q_Bm[0] = BR.re; q_Bm[1] = BR.im; s_ecut = ecut; }
            for (int i = lane; i < ecut; i += 32) {                                     // Bplus/Bminus[i+1] = E_i (B0, BR)^T
              const double *m = q_M + 8 * i;
              const cplx F0 = cadd(cmk(m[0], m[1]), cmul(cmk(m[4], m[5]), BR)), F1 = cadd(cmk(m[2], m[3]), cmul(cmk(m[6], m[7]), BR));
              q_Bp[2 * i + 2] = F0.re; q_Bp[2 * i + 3] = F0.im; q_Bm[2 * i + 2] = F1.re; q_Bm[2 * i + 3] = F1.im;
            }
          }
          __syncthreads();
          // loop 4 (:389-408): absorbed power per cell; a cell without absorption gives I_t k0 0 (finite) = 0
          for (int i = threadIdx.x; i < s_ecut; i += blockDim.x) {
            double R = 0.0;
            if (q_im[i] != 0.0 && TMM_SKIP != 4)
              R = tmm_runge5(i == 0 ? 1e10 : a.fd_hx, cmk(q_km[2 * i], q_km[2 * i + 1]), cmk(q_Bp[2 * i], q_Bp[2 * i + 1]),
                             cmk(q_Bm[2 * i], q_Bm[2 * i + 1]));
            double Q = I_t * k0 * q_im[i] * R;
            Q *= 1e10;
            Q *= 6.3538562638e-26;
            q_Q[i] = Q;
          }
          __syncthreads();
          for (int c = threadIdx.x; c < ncell; c += blockDim.x) {                  // MPI_Scatter + node.source (:414-420)
            const int idx = (c / L) * LD + c % L + 1;
            l1[idx].source = l2[idx].source = q_Q[c];
          }
        }
      }
    }
    tmm_time += tau * 10.18 * 1.0e-15;
    for (int c = threadIdx.x; c < ncell + a.p.vlatdim; c += blockDim.x) {
      if (c < ncell) {
        const int r = c / L, idx = r * LD + c % L + 1;
        TCell &x = l1[idx], &y = l2[idx];
        double src = 0.0;
        if (a.use_tmm) src = x.source;
        else {
          if (la && a.src_e[idx] >= 0.0) { src = I_t * a.p.src_abs * a.p.src_mu * a.src_e[idx]; src *= 1e10; src *= 6.3538562638e-26; }
          x.source = y.source = src;
        }
        if (c > last_active) continue;                                   // do_DIFF (:1506-1598)
        if (la) eabs += src * a.fd_vol * tau;
        if (x.natoms < mn || x.dens < RHOMIN) continue;
        const TCell &cm = l1[idx - 1], &cp = l1[idx + 1];
        const TCell &xmin = (cm.natoms < mn || cm.dens < RHOMIN) ? x : cm;
        const TCell &xmax = (cp.natoms < mn || cp.dens < RHOMIN) ? x : cp;
        double xmaxTe = xmax.temp, xminTe = xmin.temp, xmaxk = xmax.fd_k, xmink = xmin.fd_k;
        if (rv == r) { xmaxTe = v1[0].temp; xmaxk = v1[0].fd_k; }           // :1548-1552
        const double Ce = x.Ce;
        y.temp = pdiv(tau, Ce) * ((((x.fd_k + xmaxk) / 2 * (xmaxTe - x.temp) * invxsq)
                              - ((x.fd_k + xmink) / 2 * (x.temp - xminTe) * invxsq))
                             - x.fd_g * (x.temp - x.md_temp) + src) + x.temp;
        y.U = x.U + pdiv((y.temp - x.temp) * Ce * a.fd_vol, (double) x.natoms);
        const double dxi = pdiv(pdiv((y.temp - x.md_temp) * xi_fac * x.fd_g, x.md_temp), x.dens);
        x.xi += dxi;
        y.xi = x.xi;
        xiloc[idx] += dxi;
        if (isnan(y.temp) || isinf(y.temp)) atomicMax(err, 5);
      } else if (rv >= 0 && rv < P) {                                      // virtual lattice (:1602-1657)
        const int i = c - ncell, nv = a.p.vlatdim;
        const int ilocal = last_active + 1 - rv * L;
        double xminTe, xmink, xmaxTe, xmaxk;
        if (i == 0) { xminTe = l1[rv * LD + ilocal].temp; xmink = l1[rv * LD + ilocal].fd_k; }
        else { xminTe = v1[i - 1].temp; xmink = v1[i - 1].fd_k; }
        if (i < nv - 1) { xmaxk = v1[i + 1].fd_k; xmaxTe = v1[i + 1].temp; }
        else { xmaxk = v1[i].fd_k; xmaxTe = v1[i].temp; }
        const double Ce = v1[i].Ce;
        v2[i].temp = pdiv(tau, Ce) * (((v1[i].fd_k + xmaxk) / 2 * (xmaxTe - v1[i].temp) * invxsq)
                                 - ((v1[i].fd_k + xmink) / 2 * (v1[i].temp - xminTe) * invxsq)
                                 - v1[i].fd_g * (v1[i].temp - v1[i].md_temp)) + v1[i].temp;
        const double dT = tau_ci * v1[i].fd_g * (v1[i].temp - v1[i].md_temp);
        v2[i].md_temp = v1[i].md_temp + dT;
      }
    }
    __syncthreads();
    cur = 1 - cur; vcur = 1 - vcur;
    { TCell *t = l1; l1 = l2; l2 = t; t = v1; v1 = v2; v2 = t; }
    fillmesh(a, cve, ee, z0p43, l1, l2, v1, v2, last_active, rv, la, err, false);
    __syncthreads();
    fill_ghosts(a, l1);
    __syncthreads();
  }

  // ---------------------------------------------------------------- xi gather, U from the EOS (:124-148)
  double eel = 0.0;
  for (int c = threadIdx.x; c < ncell; c += blockDim.x) {
    const int idx = (c / L) * LD + c % L + 1;
    TCell &x = l1[idx], &y = l2[idx];
    a.xiarr[c] = xiloc[idx];
    if (active_ge(x, mn)) {
      x.U = eos_ee_from_r_te(ee, x.dens, x.temp * 11604.5) * 26.9815 * C_AMU * C_J2eV;
      y.U = x.U;
      eel += x.U * ((double) x.natoms);
    }
  }
  eel = block_reduce(eel, 0, s_red);
  eabs = block_reduce(eabs, 0, s_red);
  if (a.lat_smem) {
    __syncthreads();
    for (int k = threadIdx.x; k < 2 * nl; k += blockDim.x) a.lat[k] = g_lat[k];
    for (int k = threadIdx.x; k < 2 * nvl; k += blockDim.x) a.vlat[k] = g_vlat[k];
    if (a.use_tmm) {
      const double *qs = tmm_base + 17 * (size_t) (a.G + 2);
      double *qg = a.tmm_buf + 17 * (size_t) (a.G + 2);
      for (int k = threadIdx.x; k < a.G + 2; k += blockDim.x) qg[k] = qs[k];
    }
  }
  if (threadIdx.x == 0) {
    sc->cur = cur; sc->vcur = vcur;
    sc->diff_substeps = diff_substeps; sc->laser_active = la;
    sc->max_dt_ttm = s_maxdt; sc->tmm_time = tmm_time;
    sc->Eabs += eabs; sc->tot_elec_energy = eel;
  }
}

// ----------------------------------------------------------------------------- host side

static EosView eos_view(const TtmState &t, int w)
{
  EosView v;
  v.nx = t.eos_nx[w]; v.ny = t.eos_ny[w];
  v.x = t.eos_buf[w].p; v.y = v.x + v.nx; v.z = v.y + v.ny;
  v.diag = t.eos_diag[w].p;
  return v;
}

extern "C" int imdcu_ttm_set_eos(imdcu_ctx *c, int which, int nx, int ny, const double *x, const double *y,
                                 const double *z, const unsigned char *diag)
{
  if (!c || which < 0 || which > 2 || nx < 2 || ny < 2 || !x || !y || !z)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm_set_eos: bad argument") : IMDCU_EINVAL;
  for (int i = 1; i < nx; i++) if (!(x[i] > x[i - 1])) return c->fail(IMDCU_EINVAL, "imdcu_ttm_set_eos: x axis is not increasing (the table must be rectilinear)");
  for (int j = 1; j < ny; j++) if (!(y[j] > y[j - 1])) return c->fail(IMDCU_EINVAL, "imdcu_ttm_set_eos: y axis is not increasing (the table must be rectilinear)");
  cudaSetDevice(c->device);
  TtmState &t = c->ttm;
  const size_t nd = (size_t) nx + ny + (size_t) nx * ny, ng = (size_t) (nx - 1) * (ny - 1);
  CU_CHECK(c, t.eos_buf[which].ensure(nd));
  CU_CHECK(c, t.eos_diag[which].ensure(ng));
  CU_CHECK(c, imdcu_h2d(c, t.eos_buf[which].p, x, nx * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.eos_buf[which].p + nx, y, ny * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.eos_buf[which].p + nx + ny, z, (size_t) nx * ny * sizeof(double)));
  if (diag) {
    for (size_t k = 0; k < ng; k++) if (diag[k] > 1) return c->fail(IMDCU_EINVAL, "imdcu_ttm_set_eos: diag entries must be 0 or 1");
    CU_CHECK(c, imdcu_h2d(c, t.eos_diag[which].p, diag, ng));
  } else CU_CHECK(c, cudaMemset(t.eos_diag[which].p, 0, ng));
  t.eos_nx[which] = nx; t.eos_ny[which] = ny; t.eos_set[which] = true;
  return IMDCU_OK;
}

extern "C" int imdcu_ttm_set_tmm(imdcu_ctx *c, double tmm_threshold, double laser_t_1, double laser_sigma_t1,
                                 int nbb, const double *eps_bb, int nk, const double *K12)
{
  if (!c || nbb < 2 || nk < 1500 || !eps_bb || !K12)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm_set_tmm: bad argument (K12 needs 1500 rows, imd_tmm.c:674)") : IMDCU_EINVAL;
  for (int i = 1; i < nbb; i++)
    if (!(eps_bb[3 * i] > eps_bb[3 * (i - 1)])) return c->fail(IMDCU_EINVAL, "imdcu_ttm_set_tmm: eps_bb wavelengths must ascend");
  cudaSetDevice(c->device);
  TtmState &t = c->ttm;
  CU_CHECK(c, t.tmm_tab.ensure(3 * (size_t) nbb + 2 * (size_t) nk));
  CU_CHECK(c, imdcu_h2d(c, t.tmm_tab.p, eps_bb, 3 * (size_t) nbb * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.tmm_tab.p + 3 * (size_t) nbb, K12, 2 * (size_t) nk * sizeof(double)));
  t.tmm_nbb = nbb; t.tmm_nk = nk;
  t.tmm_threshold = tmm_threshold; t.laser_t_1 = laser_t_1; t.laser_sigma_t1 = laser_sigma_t1;
  t.use_tmm = true;
  return IMDCU_OK;
}

// init_ttm (imd_ttm.c:955-1201) + tmm_init
extern "C" int imdcu_ttm_init(imdcu_ctx *c, const imdcu_ttm_params *p)
{
  if (!c || !p) return c ? c->fail(IMDCU_EINVAL, "imdcu_ttm_init: bad argument") : IMDCU_EINVAL;
  if (!c->box_set || !c->params_set || !c->tab[IMDCU_TAB_PAIR].set)
    return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: set box, params and tables first");
  TtmState &t = c->ttm;
  if (!t.eos_set[0] || !t.eos_set[1]) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: EOS tables not set");
  if (t.use_tmm && !t.eos_set[2]) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: do_tmm needs the phase table (IMDCU_EOS_PHASE)");
  if (t.use_tmm) {
    const double lam_um = p->lambda * 1e6;          // tmm_get_epsilon's table walk must end inside the table
    std::vector<double> l0(1), l1(1);
    CU_CHECK(c, cudaMemcpy(l0.data(), t.tmm_tab.p, sizeof(double), cudaMemcpyDeviceToHost));
    CU_CHECK(c, cudaMemcpy(l1.data(), t.tmm_tab.p + 3 * (size_t) (t.tmm_nbb - 1), sizeof(double), cudaMemcpyDeviceToHost));
    if (!(lam_um > l0[0]) || !(lam_um <= l1[0])) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: lambda outside the eps_bb table");
  }
  if (p->ttmdimx < 1 || p->nranks_ref < 1 || p->ttmdimx % p->nranks_ref != 0)
    return c->fail(IMDCU_EINVAL, "global_fd_dim.x mod num_cpus != 0.");            // imd_ttm.c:964-967
  if (p->fd_n_timesteps < 1 || p->vlatdim < 0 || p->vlatdim > 4096)
    return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: bad fd_n_timesteps / vlatdim");
  const Geom &g = c->geom;
  if (g.box[1] != 0 || g.box[2] != 0 || g.box[3] != 0 || g.box[5] != 0 || g.box[6] != 0 || g.box[7] != 0)
    return c->fail(IMDCU_EINVAL, "imdcu_ttm_init: the 1-D TTM needs an orthorhombic box");
  cudaSetDevice(c->device);
  t.p = *p;
  t.G = p->ttmdimx; t.P = p->nranks_ref; t.L = t.G / t.P; t.LD = t.L + 2;
  t.fd_hx = g.box[0] / t.G;
  t.fd_vol = t.fd_hx * g.box[4] * g.box[8];
  t.neighvol = pow(sqrt(c->tab[IMDCU_TAB_PAIR].h_end[0]), 3.0) * 4.0 / 3.0 * M_PI;
  t.sigma_t_sq = p->laser_sigma_t * p->laser_sigma_t;
  t.omega_laser = 2.0 * M_PI * 3.0e8 / p->lambda;
  const int nl = t.P * t.LD, nv = p->vlatdim > 0 ? p->vlatdim : 1;
  CU_CHECK(c, t.lat.ensure(2 * (size_t) nl));
  CU_CHECK(c, t.vlat.ensure(2 * (size_t) nv));
  CU_CHECK(c, t.isum.ensure(4 * (size_t) t.G)); CU_CHECK(c, t.dsum.ensure(4 * (size_t) t.G));
  CU_CHECK(c, t.xiarr.ensure(t.G)); CU_CHECK(c, t.xiloc.ensure(nl)); CU_CHECK(c, t.vcom.ensure(3 * (size_t) t.G));
  CU_CHECK(c, t.src_e.ensure(nl));
  if (t.use_tmm) {
    CU_CHECK(c, t.tmm_buf.ensure(TMM_BUF_PER_CELL * (size_t) (t.G + 2)));
    CU_CHECK(c, cudaMemset(t.tmm_buf.p, 0, TMM_BUF_PER_CELL * (size_t) (t.G + 2) * sizeof(double)));
  }
  CU_CHECK(c, t.winv.ensure(((size_t) t.P * t.L + nv) * 8));
  CU_CHECK(c, cudaMemset(t.lat.p, 0, 2 * (size_t) nl * sizeof(TCell)));
  CU_CHECK(c, cudaMemset(t.xiarr.p, 0, t.G * sizeof(double)));
  CU_CHECK(c, cudaMemset(t.xiloc.p, 0, nl * sizeof(double)));
  CU_CHECK(c, cudaMemset(t.vcom.p, 0, 3 * (size_t) t.G * sizeof(double)));
  std::vector<TCell> hv(2 * (size_t) nv);
  memset(hv.data(), 0, hv.size() * sizeof(TCell));
  for (int k = 0; k < 2 * nv; k++) { hv[k].dens = p->vlatdens; hv[k].temp = 0.0264; hv[k].md_temp = 0.0264; }   // :1036-1041
  CU_CHECK(c, imdcu_h2d(c, t.vlat.p, hv.data(), hv.size() * sizeof(TCell)));
  if (!t.d_scal) {
    CU_CHECK(c, cudaMalloc(&t.d_scal, sizeof(TtmScal)));
    CU_CHECK(c, cudaMallocHost(&t.h_scal, sizeof(TtmScal)));
  }
  TtmScal s0;
  memset(&s0, 0, sizeof s0);
  s0.diff_substeps = 100;                                                         // globals.h:1187
  s0.max_dt_ttm = c->timestep / ((double) p->fd_n_timesteps);
  CU_CHECK(c, imdcu_h2d(c, t.d_scal, &s0, sizeof s0));
  if (c->n > 0) {
    CU_CHECK(c, t.fdi.ensure(c->key.cap)); CU_CHECK(c, t.fdi_alt.ensure(c->key.cap));
    CU_CHECK(c, cudaMemset(t.fdi.p, 0, c->key.cap * sizeof(int)));
  }
  t.inited = true;
  return IMDCU_OK;
}

void imdcu_ttm_release(imdcu_ctx *c)
{
  if (c->ttm.stream2) { cudaStreamDestroy(c->ttm.stream2); cudaEventDestroy(c->ttm.ev_fork); cudaEventDestroy(c->ttm.ev_join); c->ttm.stream2 = nullptr; }
  TtmState &t = c->ttm;
  for (int w = 0; w < 2; w++) { t.eos_buf[w].release(); t.eos_diag[w].release(); }
  t.lat.release(); t.vlat.release(); t.isum.release(); t.dsum.release(); t.xiarr.release(); t.xiloc.release();
  t.tmm_tab.release(); t.tmm_buf.release(); t.eos_buf[2].release(); t.eos_diag[2].release();
  t.vcom.release(); t.src_e.release(); t.winv.release(); t.fdi.release(); t.fdi_alt.release();
  if (t.d_scal) cudaFree(t.d_scal);
  if (t.h_scal) cudaFreeHost(t.h_scal);
  t.d_scal = t.h_scal = nullptr;
}

static int ttm_read_scalars(imdcu_ctx *c, double out16[16])
{
  TtmState &t = c->ttm;
  CU_CHECK(c, cudaMemcpyAsync(t.h_scal, t.d_scal, sizeof(TtmScal), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  const TtmScal *s = (const TtmScal *) t.h_scal;
  for (int k = 0; k < 16; k++) out16[k] = 0.0;
  out16[0] = s->diff_substeps; out16[1] = s->max_dt_ttm; out16[2] = s->last_active_cell_global;
  out16[3] = s->cur_vlattice_proc; out16[4] = s->Eabs; out16[5] = s->tot_elec_energy;
  out16[6] = t.fd_hx; out16[7] = t.fd_vol; out16[8] = t.neighvol; out16[9] = s->laser_active;
  out16[10] = s->tmm_time; out16[11] = s->err;
  if (s->err) {
    static const char *msg[] = {"", "ERROR during CfromT in FILLMESH", "fd_k is NaN", "fd_g is NaN",
                                "EOS Plausibility check failed, TfromU != Tinit", "Te got NaN or Inf in diffloop",
                                "Sample is too short for TMM.Consider reducing tmm_threshold."};
    return c->fail(IMDCU_EINVAL, std::string("calc_ttm: ") + msg[s->err < 7 ? s->err : 0]);
  }
  return IMDCU_OK;
}

// update_fd's atom loops and their reductions (imd_ttm.c:269-395), on the context's stream
static int ttm_atoms_part(imdcu_ctx *c, int steps)
{
  TtmState &t = c->ttm;
  cudaStream_t st = c->stream;
  const int n = (int) c->n, G = t.G;
  RET_IF(imdcu_join_upload(c));
  if (t.fdi.cap < (size_t) n) {
    CU_CHECK(c, t.fdi.ensure(c->key.cap)); CU_CHECK(c, t.fdi_alt.ensure(c->key.cap));
    CU_CHECK(c, cudaMemsetAsync(t.fdi.p, 0, c->key.cap * sizeof(int), st));
  }
  CU_CHECK(c, cudaMemsetAsync(t.isum.p, 0, 4 * (size_t) G * sizeof(int), st));
  CU_CHECK(c, cudaMemsetAsync(t.dsum.p, 0, 4 * (size_t) G * sizeof(double), st));
  // a few CTAs per SM, every warp a contiguous atom range of at least 32 * TTM_AU atoms
  const int nb = std::max(1, std::min(c->num_sms * 4, div_up(n, TTM_ANT * TTM_AU)));
  if (n > 0) {
    k_ttm_atoms1<<<nb, TTM_ANT, 0, st>>>(n, G, t.fd_hx, steps, c->pos4.p, c->mom4.p, c->type.p, c->numneighs.p,
                                     t.fdi.p, t.isum.p, t.dsum.p);
    c->count_launch();
  }
  // the 7 MPI_Allreduce of imd_ttm.c:325-334 as two calls on the packed buffers; afterwards every
  // GPU holds the global slab sums and advances the identical (replicated) FD lattice
  RET_IF(imdcu_multi_allreduce2(c, t.isum.p, 4 * G, 2, t.dsum.p, 3 * G, 0));      // one NCCL launch for both
  if (n > 0) {
    k_ttm_atoms2<<<nb, TTM_ANT, 0, st>>>(n, G, t.fd_hx, t.p.atomic_weight, c->mom4.p, c->type.p, t.fdi.p, t.isum.p,
                                     t.dsum.p);
    c->count_launch();
  }
  RET_IF(imdcu_multi_allreduce(c, t.dsum.p + 3 * (size_t) G, G, 0));   // imd_ttm.c:395
  return IMDCU_OK;
}

// the G-sized rest of calc_ttm: one launch of k_ttm_grid on stream `st`
static int ttm_grid_part(imdcu_ctx *c, int steps, cudaStream_t st)
{
  TtmState &t = c->ttm;
  const int G = t.G;
  TtmArgs a;
  a.G = G; a.P = t.P; a.L = t.L; a.LD = t.LD; a.steps = steps; a.ncell_tot = t.P * t.L + t.p.vlatdim;
  a.p = t.p;
  a.timestep = c->timestep; a.fd_hx = t.fd_hx; a.fd_vol = t.fd_vol; a.neighvol = t.neighvol;
  a.omega_laser = t.omega_laser; a.sigma_t_sq = t.sigma_t_sq;
  a.cve = eos_view(t, 0); a.ee = eos_view(t, 1);
  a.lat = t.lat.p; a.vlat = t.vlat.p; a.isum = t.isum.p; a.dsum = t.dsum.p;
  a.xiarr = t.xiarr.p; a.xiloc = t.xiloc.p; a.vcom = t.vcom.p; a.src_e = t.src_e.p; a.winv = t.winv.p;
  a.scal = (TtmScal *) t.d_scal;
  a.use_tmm = t.use_tmm ? 1 : 0;
  a.tmm_threshold = t.tmm_threshold; a.laser_t_1 = t.laser_t_1; a.sigma_t1_sq = t.laser_sigma_t1 * t.laser_sigma_t1;
  a.tmm.nbb = t.tmm_nbb; a.tmm.nk = t.tmm_nk; a.tmm.eps_bb = t.tmm_tab.p; a.tmm.K12 = t.tmm_tab.p + 3 * (size_t) t.tmm_nbb;
  a.tmm_buf = t.tmm_buf.p;
  if (t.use_tmm) a.phase = eos_view(t, 2); else a.phase = a.cve;
  int nt = ((a.ncell_tot + 31) / 32) * 32;
  if (nt > TTM_NT) nt = TTM_NT;
  if (nt < 32) nt = 32;
  // dynamic shared memory for the two EOS tables (axes, values, diagonal flags)
  size_t smem = 0;
  for (int w = 0; w < 2; w++) {
    const size_t nx = t.eos_nx[w], ny = t.eos_ny[w];
    smem += (nx + ny + nx * ny) * sizeof(double) + (((nx - 1) * (ny - 1) + 7) & ~(size_t) 7);
  }
  a.eos_smem = smem <= 200 * 1024;
  if (!a.eos_smem) smem = 0;
  {
    const bool off = !c->tune.ttm_lat_smem;
    const size_t nvl = t.p.vlatdim > 0 ? t.p.vlatdim : 1;
    const size_t lat_bytes = (2 * (size_t) t.P * t.LD + 2 * nvl) * sizeof(TCell) + 16 + (size_t) t.P * t.LD * sizeof(double) +
                             (t.use_tmm ? TMM_BUF_PER_CELL * (size_t) (t.G + 2) * sizeof(double) : 0);
    a.lat_smem = !off && smem + lat_bytes <= 210 * 1024;
    if (a.lat_smem) smem += lat_bytes;
  }
  // Beside k_eam_pass2 (imdcu_ttm_fork) the lattice CTA asks for so much shared memory that no CTA of
  // the force kernel fits next to it: sharing an SM with 16 warps that keep the L1 data pipe 80 % busy
  // stretches every dependent load and barrier of this latency chain.  Pass 2 leaves that SM alone
  // (eam.cu: run_passes).  IMDCU_TTM_ISOLATE=0 switches it off.
  if (st != c->stream && c->ttm_isolate && smem < c->ttm_isolate_smem) smem = c->ttm_isolate_smem;
  if (smem > t.grid_smem_set) {
    CU_CHECK(c, cudaFuncSetAttribute(k_ttm_grid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    t.grid_smem_set = smem;
  }
  k_ttm_grid<<<1, nt, smem, st>>>(a);
  c->count_launch();
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}

// calc_ttm (imd_ttm.c:75-156)
int imdcu_ttm_step(imdcu_ctx *c, int steps, double out16[16])
{
  TtmState &t = c->ttm;
  if (!t.inited) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init has not been called");
  {
    PhaseTimer pt(c, 7);
    RET_IF(ttm_atoms_part(c, steps));
    RET_IF(ttm_grid_part(c, steps, c->stream));
  }
  if (out16) return ttm_read_scalars(c, out16);
  return IMDCU_OK;
}

// calc_ttm overlapped with the second EAM pass (used by imdcu_run_ttm): called between the passes,
// after numneighs of this step exist.  The atom loops run on the context's stream; the lattice
// kernel (one CTA, ~2 ms of dependent FP64 latency, imd_ttm.c:401-541 + :103-148) is forked to a
// second stream and runs beside k_eam_pass2, which leaves one CTA slot free for it.
int imdcu_ttm_fork(imdcu_ctx *c, int steps)
{
  TtmState &t = c->ttm;
  if (!t.inited) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init has not been called");
  if (!t.stream2) {
    CU_CHECK(c, cudaStreamCreateWithFlags(&t.stream2, cudaStreamNonBlocking));
    CU_CHECK(c, cudaEventCreateWithFlags(&t.ev_fork, cudaEventDisableTiming));
    CU_CHECK(c, cudaEventCreateWithFlags(&t.ev_join, cudaEventDisableTiming));
  }
  RET_IF(ttm_atoms_part(c, steps));
  CU_CHECK(c, cudaEventRecord(t.ev_fork, c->stream));
  CU_CHECK(c, cudaStreamWaitEvent(t.stream2, t.ev_fork, 0));
  RET_IF(ttm_grid_part(c, steps, t.stream2));
  CU_CHECK(c, cudaEventRecord(t.ev_join, t.stream2));
  t.forked = true;
  return IMDCU_OK;
}

int imdcu_ttm_join(imdcu_ctx *c, double out16[16])
{
  TtmState &t = c->ttm;
  if (t.forked) { CU_CHECK(c, cudaStreamWaitEvent(c->stream, t.ev_join, 0)); t.forked = false; }
  if (out16) return ttm_read_scalars(c, out16);
  return IMDCU_OK;
}

// single-function probe of ttm_math.cuh on the device (parity tests): op 0 plog(a), 1 pexp(a),
// 2 ppow(a,b), 3 pdiv(a,b), 4 psqrt(a), 5 prcp(a)
__global__ void k_ttm_math_probe(int op, int64_t n, const double *__restrict__ a, const double *__restrict__ b,
                                 double *__restrict__ out)
{
  const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = a[i], y = b ? b[i] : 0.0;
  double r;
  switch (op) {
    case 0: r = plog(x); break;
    case 1: r = pexp(x); break;
    case 2: r = ppow(x, y); break;
    case 3: r = pdiv(x, y); break;
    case 4: r = psqrt(x); break;
    default: r = prcp(x); break;
  }
  out[i] = r;
}

extern "C" int imdcu_eval_ttm_math(imdcu_ctx *c, int op, int64_t n, const double *a, const double *b, double *out)
{
  if (!c || op < 0 || op > 5 || n < 0 || !a || !out || ((op == 2 || op == 3) && !b))
    return c ? c->fail(IMDCU_EINVAL, "imdcu_eval_ttm_math: bad argument") : IMDCU_EINVAL;
  if (n == 0) return IMDCU_OK;
  cudaSetDevice(c->device);
  double *d = nullptr;
  CU_CHECK(c, cudaMalloc(&d, 3 * (size_t) n * sizeof(double)));
  imdcu_h2d(c, d, a, n * sizeof(double));
  if (b) imdcu_h2d(c, d + n, b, n * sizeof(double));
  k_ttm_math_probe<<<div_up(n, 256), 256, 0, c->stream>>>(op, n, d, b ? d + n : nullptr, d + 2 * n);
  cudaStreamSynchronize(c->stream);
  const cudaError_t e = cudaMemcpy(out, d + 2 * n, n * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(d);
  CU_CHECK(c, e);
  return IMDCU_OK;
}

extern "C" int imdcu_download_ttm(imdcu_ctx *c, double *cells, double *xiarr, double *vcom, double *vlat)
{
  if (!c || !c->ttm.inited) return c ? c->fail(IMDCU_EINVAL, "TTM not initialised") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  TtmState &t = c->ttm;
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  TtmScal s;
  CU_CHECK(c, cudaMemcpy(&s, t.d_scal, sizeof s, cudaMemcpyDeviceToHost));
  const int nl = t.P * t.LD, nv = t.p.vlatdim;
  if (cells) {
    std::vector<TCell> h((size_t) nl);
    CU_CHECK(c, cudaMemcpy(h.data(), t.lat.p + (size_t) s.cur * nl, nl * sizeof(TCell), cudaMemcpyDeviceToHost));
    for (int r = 0; r < t.P; r++)
      for (int i = 1; i <= t.L; i++) {
        const TCell &x = h[(size_t) r * t.LD + i];
        double *o = cells + (size_t) ((i - 1) + r * t.L) * 13;
        o[0] = x.natoms; o[1] = x.natoms_old; o[2] = x.temp; o[3] = x.md_temp; o[4] = x.xi; o[5] = x.dens;
        o[6] = x.fd_k; o[7] = x.fd_g; o[8] = x.Z; o[9] = x.ne; o[10] = x.Ce; o[11] = x.U; o[12] = x.source;
      }
  }
  if (xiarr) CU_CHECK(c, cudaMemcpy(xiarr, t.xiarr.p, t.G * sizeof(double), cudaMemcpyDeviceToHost));
  if (vcom) CU_CHECK(c, cudaMemcpy(vcom, t.vcom.p, 3 * (size_t) t.G * sizeof(double), cudaMemcpyDeviceToHost));
  if (vlat && nv > 0) {
    std::vector<TCell> h((size_t) nv);
    CU_CHECK(c, cudaMemcpy(h.data(), t.vlat.p + (size_t) s.vcur * nv, nv * sizeof(TCell), cudaMemcpyDeviceToHost));
    for (int i = 0; i < nv; i++) {
      double *o = vlat + 6 * (size_t) i;
      o[0] = h[i].temp; o[1] = h[i].md_temp; o[2] = h[i].fd_k; o[3] = h[i].fd_g; o[4] = h[i].Ce; o[5] = h[i].Z;
    }
  }
  return IMDCU_OK;
}
