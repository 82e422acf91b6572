// imd_b200/csrc/atoms.cu -- host <-> device staging of per-atom state.
// Host arrays cross the ABI in the reference's AoS form (ORT/IMPULS/KRAFT: 3 doubles per
// atom, makros.h:143-145); they are copied raw (cudaMemcpyAsync, full speed when the caller's
// memory is pinned) into a device staging buffer and (un)packed by a kernel.
#include "common.cuh"

static int ensure_atom_capacity(imdcu_ctx *c, int64_t n)
{
  const double slack = 1.1;
  CU_CHECK(c, c->mom4.ensure(n, slack));      CU_CHECK(c, c->mom4_alt.ensure(n, slack));
  CU_CHECK(c, c->frc4.ensure(n, slack));      CU_CHECK(c, c->rho.ensure(n, slack));
  CU_CHECK(c, c->dF.ensure(n, slack));
  CU_CHECK(c, c->refx.ensure(n, slack));      CU_CHECK(c, c->refy.ensure(n, slack));
  CU_CHECK(c, c->refz.ensure(n, slack));
  CU_CHECK(c, c->vtype.ensure(n, slack));     CU_CHECK(c, c->vtype_alt.ensure(n, slack));
  CU_CHECK(c, c->id.ensure(n, slack));        CU_CHECK(c, c->id_alt.ensure(n, slack));
  CU_CHECK(c, c->numneighs.ensure(n, slack)); CU_CHECK(c, c->nbl_num.ensure(n, slack));
  CU_CHECK(c, c->key.ensure(n, slack));       CU_CHECK(c, c->key_alt.ensure(n, slack));
  CU_CHECK(c, c->perm.ensure(n, slack));      CU_CHECK(c, c->perm_alt.ensure(n, slack));
  // arrays that also hold ghosts start with room for 50 % ghosts and grow on demand
  CU_CHECK(c, c->pos4.ensure(n + n / 2 + 1024));  CU_CHECK(c, c->pos4_alt.ensure_exact(c->pos4.cap));
  CU_CHECK(c, c->type.ensure(n + n / 2 + 1024));  CU_CHECK(c, c->type_alt.ensure_exact(c->type.cap));
  // headroom like the arrays above: on several GPUs the atom count of a rank creeps up by migration, and growing this
  // 320 MB buffer (cudaMalloc + cudaFree) in the middle of a list rebuild stalls every rank for tens of milliseconds
  CU_CHECK(c, c->stage.ensure(10 * (size_t) n + 64, 1.25));
  return IMDCU_OK;
}

// staging layout (doubles): [0,3n) pos, [3n,6n) mom, [6n,7n) mass   (in)
//                           [0,3n) pos, [3n,6n) mom, [6n,9n) force, [9n,10n) epot (out)
__global__ void __launch_bounds__(256)
k_unpack_in(int n, const double *__restrict__ st, int have_mass, double4 *__restrict__ pos4,
            double4 *__restrict__ mom4)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double *p = st + 3 * (size_t) i, *m = st + 3 * (size_t) n + 3 * (size_t) i;
  double4 x = pos4[i], q = mom4[i];
  x.x = p[0]; x.y = p[1]; x.z = p[2];
  q.x = m[0]; q.y = m[1]; q.z = m[2];
  if (have_mass) { q.w = st[6 * (size_t) n + i]; x.w = 0.0; }
  pos4[i] = x; mom4[i] = q;
}

__global__ void __launch_bounds__(256)
k_pack_out(int n, const double4 *__restrict__ pos4, const double4 *__restrict__ mom4,
           const double4 *__restrict__ frc4, double *__restrict__ st)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 x = pos4[i], q = mom4[i], f = frc4[i];
  double *p = st + 3 * (size_t) i;
  p[0] = x.x; p[1] = x.y; p[2] = x.z;
  p += 3 * (size_t) n; p[0] = q.x; p[1] = q.y; p[2] = q.z;
  p += 3 * (size_t) n; p[0] = f.x; p[1] = f.y; p[2] = f.z;
  st[9 * (size_t) n + i] = f.w;
}

__global__ void __launch_bounds__(256)
k_pack_w(int n, const double4 *__restrict__ a, double *__restrict__ out)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i].w;
}

extern "C" int imdcu_upload_atoms(imdcu_ctx *c, int64_t n, const int32_t *id, const int32_t *type,
                                  const int32_t *vtype, const double *mass, const double *pos,
                                  const double *mom)
{
  if (!c || n < 0 || !pos || !mom || !mass)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_upload_atoms: bad argument") : IMDCU_EINVAL;
  if (n >= (int64_t) 1 << 30) return c->fail(IMDCU_EINVAL, "imdcu_upload_atoms: more than 2^30 atoms per rank");
  cudaSetDevice(c->device);
  RET_IF(imdcu_join_upload(c));
  CU_CHECK(c, cudaStreamSynchronize(c->up_stream));          // buffers may be reallocated below
  RET_IF(ensure_atom_capacity(c, n));
  std::vector<int> ht((size_t) n), hv((size_t) n), hi((size_t) n);
  for (int64_t i = 0; i < n; i++) {
    ht[i] = type ? type[i] : 0;
    hv[i] = vtype ? vtype[i] : (c->nvtypes > ht[i] ? ht[i] : 0);
    hi[i] = id ? id[i] : (int) i;
    if (ht[i] < 0 || ht[i] >= c->ntypes) return c->fail(IMDCU_EINVAL, "atom type out of range");
    if (hv[i] < 0 || hv[i] >= c->nvtypes) return c->fail(IMDCU_EINVAL, "atom vtype out of range");
  }
  cudaStream_t s = c->stream;
  double *st = c->stage.p;
  CU_CHECK(c, cudaMemcpyAsync(st, pos, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(st + 3 * n, mom, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(st + 6 * n, mass, n * sizeof(double), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->type.p, ht.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->vtype.p, hv.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->id.p, hi.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  if (n) { k_unpack_in<<<div_up(n, 256), 256, 0, s>>>((int) n, st, 1, c->pos4.p, c->mom4.p); c->count_launch(); }
  CU_CHECK(c, cudaMemsetAsync(c->frc4.p, 0, n * sizeof(double4), s));
  CU_CHECK(c, cudaMemsetAsync(c->rho.p, 0, n * sizeof(double), s));
  CU_CHECK(c, cudaMemsetAsync(c->numneighs.p, 0, n * sizeof(int), s));
  CU_CHECK(c, cudaMemsetAsync(c->key.p, 0, n * sizeof(int), s));
  if (c->ttm.inited) {                          // fresh atoms: no stored FD slab yet
    CU_CHECK(c, c->ttm.fdi.ensure(c->key.cap)); CU_CHECK(c, c->ttm.fdi_alt.ensure(c->key.cap));
    CU_CHECK(c, cudaMemsetAsync(c->ttm.fdi.p, 0, c->key.cap * sizeof(int), s));
  }
  CU_CHECK(c, cudaStreamSynchronize(s));
  c->n = n; c->ng = 0;
  c->have_valid_nbl = false;
  if (c->nrb.active) { RET_IF(imdcu_nrb_reindex(c, true)); CU_CHECK(c, cudaStreamSynchronize(s)); }
  return IMDCU_OK;
}

// the two halves of k_unpack_in for imdcu_update_atoms (st: 3 doubles per atom)
__global__ void __launch_bounds__(256)
k_unpack_xyz(int n, const double *__restrict__ st, double4 *__restrict__ dst)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double *p = st + 3 * (size_t) i;
  double4 x = dst[i];
  x.x = p[0]; x.y = p[1]; x.z = p[2];
  dst[i] = x;
}

// positions and momenta of the SAME atoms in the device's current order (the order of the
// last imdcu_download_atoms): the neighbour list stays valid, check_nblist decides later.
extern "C" int imdcu_update_atoms(imdcu_ctx *c, const double *pos, const double *mom)
{
  if (!c || !pos || !mom) return c ? c->fail(IMDCU_EINVAL, "imdcu_update_atoms: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  const int64_t n = c->n;
  if (n == 0) return IMDCU_OK;
  cudaStream_t s = c->stream, u = c->up_stream;
  double *st = c->stage.p;
  RET_IF(imdcu_join_upload(c));
  // calc_forces needs the positions only: the momenta go over a second stream and are awaited by the
  // first thing that touches mom4 (imdcu_join_upload), so their copy runs beside the force passes
  CU_CHECK(c, cudaEventRecord(c->ev_up0, s));                 // staging and mom4 are free once earlier work is done
  CU_CHECK(c, cudaStreamWaitEvent(u, c->ev_up0, 0));
  CU_CHECK(c, cudaMemcpyAsync(st, pos, 3 * n * sizeof(double), cudaMemcpyHostToDevice, s));
  k_unpack_xyz<<<div_up(n, 256), 256, 0, s>>>((int) n, st, c->pos4.p);
  CU_CHECK(c, cudaMemcpyAsync(st + 3 * n, mom, 3 * n * sizeof(double), cudaMemcpyHostToDevice, u));
  k_unpack_xyz<<<div_up(n, 256), 256, 0, u>>>((int) n, st + 3 * n, c->mom4.p);
  CU_CHECK(c, cudaEventRecord(c->ev_up, u));
  c->up_pending = true;
  c->count_launch(2);
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}

extern "C" int64_t imdcu_num_atoms(const imdcu_ctx *c) { return c ? c->n : -1; }

extern "C" int imdcu_download_atoms(imdcu_ctx *c, int32_t *id, int32_t *type, double *mass,
                                    double *pos, double *mom, double *force, double *epot,
                                    double *rho, double *dF, int32_t *numneighs, int32_t *cell3)
{
  if (!c) return IMDCU_EINVAL;
  cudaSetDevice(c->device);
  const int64_t n = c->n;
  cudaStream_t s = c->stream;
  if (n == 0) return IMDCU_OK;
  RET_IF(imdcu_join_upload(c));
  double *st = c->stage.p;
  const int nb = div_up(n, 256);
  if (pos || mom || force || epot) {
    k_pack_out<<<nb, 256, 0, s>>>((int) n, c->pos4.p, c->mom4.p, c->frc4.p, st);
    c->count_launch();
    if (pos) CU_CHECK(c, cudaMemcpyAsync(pos, st, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (mom) CU_CHECK(c, cudaMemcpyAsync(mom, st + 3 * n, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (force) CU_CHECK(c, cudaMemcpyAsync(force, st + 6 * n, 3 * n * sizeof(double), cudaMemcpyDeviceToHost, s));
    if (epot) CU_CHECK(c, cudaMemcpyAsync(epot, st + 9 * n, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CU_CHECK(c, cudaStreamSynchronize(s));
  }
  if (dF) {
    k_pack_w<<<nb, 256, 0, s>>>((int) n, c->pos4.p, st); c->count_launch();
    CU_CHECK(c, cudaMemcpyAsync(dF, st, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CU_CHECK(c, cudaStreamSynchronize(s));
  }
  if (mass) {
    k_pack_w<<<nb, 256, 0, s>>>((int) n, c->mom4.p, st); c->count_launch();
    CU_CHECK(c, cudaMemcpyAsync(mass, st, n * sizeof(double), cudaMemcpyDeviceToHost, s));
    CU_CHECK(c, cudaStreamSynchronize(s));
  }
  if (rho) CU_CHECK(c, cudaMemcpyAsync(rho, c->rho.p, n * sizeof(double), cudaMemcpyDeviceToHost, s));
  if (id) CU_CHECK(c, cudaMemcpyAsync(id, c->id.p, n * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (type) CU_CHECK(c, cudaMemcpyAsync(type, c->type.p, n * sizeof(int), cudaMemcpyDeviceToHost, s));
  if (numneighs) CU_CHECK(c, cudaMemcpyAsync(numneighs, c->numneighs.p, n * sizeof(int), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaStreamSynchronize(s));
  if (cell3) {
    // key = extended cell index at the last rebuild -> global cell coordinates
    std::vector<int> k((size_t) n);
    CU_CHECK(c, cudaMemcpy(k.data(), c->key.p, n * sizeof(int), cudaMemcpyDeviceToHost));
    const Geom &g = c->geom;
    for (int64_t i = 0; i < n; i++) {
      int e = k[i];
      cell3[3 * i] = e / (g.edim[1] * g.edim[2]) - 1 + g.xoff;
      cell3[3 * i + 1] = (e / g.edim[2]) % g.edim[1] - 1;
      cell3[3 * i + 2] = e % g.edim[2] - 1;
    }
  }
  return IMDCU_OK;
}

// VSORTE of every owned atom in the device's current order (the writers print it, imd_io_3d.c:778); the
// internal NRB bit is masked off
extern "C" int imdcu_download_vtypes(imdcu_ctx *c, int32_t *vtype)
{
  if (!c || !vtype) return c ? c->fail(IMDCU_EINVAL, "imdcu_download_vtypes: bad argument") : IMDCU_EINVAL;
  cudaSetDevice(c->device);
  const int64_t n = c->n;
  if (n == 0) return IMDCU_OK;
  CU_CHECK(c, cudaMemcpyAsync(vtype, c->vtype.p, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  for (int64_t i = 0; i < n; i++) vtype[i] &= ~IMDCU_VTYPE_NRB;
  return IMDCU_OK;
}
