// imd_b200/csrc/integrate.cu -- leap-frog NVE step fused with the neighbour-list validity
// check: move_atoms_nve (imd_integrate.c:35-561, the plain-NVE branch incl. the PDECAY rear
// absorber :233-239) + check_nblist (imd_forces_nbl.c:1062-1092) in one pass over the atoms.
// HBM-bound streaming kernel: 32 (mom) + 32 (force) + 32 (pos) + 24 (ref) + 4 (vtype) bytes
// read, 64 bytes written per atom.
#include "device.cuh"

struct MoveArgs {
  int n;
  double4 *pos4; double4 *mom4; const double4 *frc4;
  const double *refx, *refy, *refz;
  const int *vtype; const double *restrictions;
  double timestep;
  int pdecay; double xipdecay, ramp_start, pd_a;
  double *partials; unsigned int *ticket; double *results;
};

__global__ void __launch_bounds__(256)
k_move_nve(MoveArgs a)
{
  double ekin[1] = {0.0}, dmax[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
    double4 p = a.mom4[i];
    const double4 f4 = a.frc4[i];
    double4 x = a.pos4[i];
    const int vt = a.vtype[i];
    const double *rs = a.restrictions + 3 * (vt & ~IMDCU_VTYPE_NRB);
    const double m = p.w;
    const double k1 = p.x * p.x + p.y * p.y + p.z * p.z;
    if (!(vt & IMDCU_VTYPE_NRB)) {                       // NRB boundary atoms: no kick (imd_integrate.c:233-259)
      double fx = f4.x * rs[0], fy = f4.y * rs[1], fz = f4.z * rs[2];
      if (a.pdecay && x.x > a.ramp_start)
        fx -= (p.x / m) * a.xipdecay * a.pd_a * (x.x - a.ramp_start) * (x.x - a.ramp_start);
      p.x += a.timestep * fx; p.y += a.timestep * fy; p.z += a.timestep * fz;
    }
    const double k2 = p.x * p.x + p.y * p.y + p.z * p.z;
    ekin[0] += (k1 + k2) / (4 * m);
    const double tmp = a.timestep / m;
    x.x += tmp * p.x; x.y += tmp * p.y; x.z += tmp * p.z;
    a.mom4[i] = p;
    a.pos4[i] = x;
    const double dx = x.x - a.refx[i], dy = x.y - a.refy[i], dz = x.z - a.refz[i];
    dmax[0] = fmax(dmax[0], r2_exact(dx, dy, dz));
  }
  grid_reduce<1, 0>(ekin, a.partials, a.ticket, a.results);
  grid_reduce<1, 1>(dmax, a.partials + gridDim.x, a.ticket + 1, a.results + 1);
}

// move_atoms_ttm (imd_integrate.c:581-811, TTM1D [+PDECAY]): as above plus the electron-phonon
// coupling force xi * m * (p/m - v_com) with xi, v_com of the atom's FD slab (:649-667, 738-747).
// xiarr/vcom are G-sized and cache resident; +12 B/atom (type, slab arithmetic on x).
__global__ void __launch_bounds__(256)
k_move_ttm(MoveArgs a, const int *__restrict__ type, int G, double fd_hx,
           const double *__restrict__ xiarr, const double *__restrict__ vcom)
{
  double ekin[1] = {0.0}, dmax[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
    double4 p = a.mom4[i];
    const double4 f4 = a.frc4[i];
    double4 x = a.pos4[i];
    const int vt = a.vtype[i];
    const double *rs = a.restrictions + 3 * (vt & ~IMDCU_VTYPE_NRB);
    const double m = p.w;
    int ig = 0;
    double xi = 0.0;
    if (type[i] == 0) {
      ig = (int) (x.x / fd_hx);
      ig = ig < G - 1 ? ig : G - 1;
      ig = ig > 0 ? ig : 0;
      xi = xiarr[ig];
    }
    const double k1 = p.x * p.x + p.y * p.y + p.z * p.z;
    if (!(vt & IMDCU_VTYPE_NRB)) {                       // NRB boundary atoms: no kick, no heating (:712-751)
      double fx = f4.x * rs[0], fy = f4.y * rs[1], fz = f4.z * rs[2];
      if (a.pdecay && x.x > a.ramp_start)
        fx -= (p.x / m) * a.xipdecay * a.pd_a * (x.x - a.ramp_start) * (x.x - a.ramp_start);
      const double vx = vcom[ig], vy = vcom[G + ig], vz = vcom[2 * G + ig];
      p.x += a.timestep * (fx + xi * m * (p.x / m - vx));
      p.y += a.timestep * (fy + xi * m * (p.y / m - vy));
      p.z += a.timestep * (fz + xi * m * (p.z / m - vz));
    }
    const double k2 = p.x * p.x + p.y * p.y + p.z * p.z;
    ekin[0] += (k1 + k2) / (4 * m);
    const double tmp = a.timestep / m;
    x.x += tmp * p.x; x.y += tmp * p.y; x.z += tmp * p.z;
    a.mom4[i] = p;
    a.pos4[i] = x;
    const double dx = x.x - a.refx[i], dy = x.y - a.refy[i], dz = x.z - a.refz[i];
    dmax[0] = fmax(dmax[0], r2_exact(dx, dy, dz));
  }
  grid_reduce<1, 0>(ekin, a.partials, a.ticket, a.results);
  grid_reduce<1, 1>(dmax, a.partials + gridDim.x, a.ticket + 1, a.results + 1);
}

static int integrate(imdcu_ctx *c, double out8[8], int *valid, bool ttm);

// results[5] = this rank's max displacement^2 -> results[6] = 1 if it asks for a new list
__global__ void k_nbl_flag(double *results, double lim)
{
  results[3] = 0.0;
  results[6] = results[5] > lim ? 1.0 : 0.0;
}

int imdcu_integrate_nve(imdcu_ctx *c, double out8[8], int *valid) { return integrate(c, out8, valid, false); }
int imdcu_integrate_ttm(imdcu_ctx *c, double out8[8], int *valid)
{
  if (!c->ttm.inited) return c->fail(IMDCU_EINVAL, "imdcu_ttm_init has not been called");
  return integrate(c, out8, valid, true);
}

static int integrate(imdcu_ctx *c, double out8[8], int *valid, bool ttm)
{
  RET_IF(imdcu_join_upload(c));
  PhaseTimer pt(c, 6);
  MoveArgs a;
  a.n = (int) c->n;
  a.pos4 = c->pos4.p; a.mom4 = c->mom4.p; a.frc4 = c->frc4.p;
  a.refx = c->refx.p; a.refy = c->refy.p; a.refz = c->refz.p;
  a.vtype = c->vtype.p; a.restrictions = c->d_restrictions.p;
  a.timestep = c->timestep;
  a.pdecay = c->pdecay; a.xipdecay = c->xipdecay; a.ramp_start = c->ramp_start;
  { double t = 1.0 / (c->ramp_end - c->ramp_start); a.pd_a = t * t; }
  const int grid = std::max(1, std::min(div_up(a.n, 256), c->num_sms * 8));
  CU_CHECK(c, c->partials.ensure((size_t) grid * 4 + 64));
  a.partials = c->partials.p; a.ticket = c->d_ticket + 2; a.results = c->d_results + 4;
  if (ttm) k_move_ttm<<<grid, 256, 0, c->stream>>>(a, c->type.p, c->ttm.G, c->ttm.fd_hx, c->ttm.xiarr.p, c->ttm.vcom.p);
  else k_move_nve<<<grid, 256, 0, c->stream>>>(a);
  c->count_launch();
  CU_CHECK(c, cudaGetLastError());
  // check_nblist: invalidate when max displacement^2 > (0.5*nbl_margin)^2 on any rank
  const double lim = (0.5 * c->nbl_margin) * (0.5 * c->nbl_margin);
  const bool multi = c->comm.nranks > 1;
  if (multi) {
    // tot_kin_energy (imd_integrate.c:790-808) and check_nblist's MPI_MAX (imd_forces_nbl.c:1085) in ONE sum:
    // the maximum is only compared with the limit, so every rank contributes a 0/1 flag instead; pending
    // force sums (slots 0..2) ride along.  Slots 3 and 5 are reduced too and not used.
    k_nbl_flag<<<1, 1, 0, c->stream>>>(c->d_results, lim);
    c->count_launch();
    if (c->force_sums_pending) RET_IF(imdcu_multi_allreduce(c, c->d_results, 7, 0));
    else RET_IF(imdcu_multi_allreduce(c, c->d_results + 4, 3, 0));
    c->force_sums_pending = false;
  }
  CU_CHECK(c, cudaMemcpyAsync(c->h_results + 4, c->d_results + 4, 3 * sizeof(double),
                              cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  if (out8) { for (int k = 0; k < 8; k++) out8[k] = 0.0; out8[0] = c->h_results[4]; }
  if (multi ? c->h_results[6] > 0.5 : c->h_results[5] > lim) c->have_valid_nbl = false;
  if (valid) *valid = c->have_valid_nbl ? 1 : 0;
  return IMDCU_OK;
}
