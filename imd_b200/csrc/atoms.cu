// imd_b200/csrc/atoms.cu -- host <-> device staging of per-atom state
#include "common.cuh"

static int ensure_atom_capacity(imdcu_ctx *c, int64_t n)
{
  const double slack = 1.1;
  CU_CHECK(c, c->mom4.ensure(n, slack));      CU_CHECK(c, c->mom4_alt.ensure(n, slack));
  CU_CHECK(c, c->frc4.ensure(n, slack));      CU_CHECK(c, c->rho.ensure(n, slack));
  CU_CHECK(c, c->refx.ensure(n, slack));      CU_CHECK(c, c->refy.ensure(n, slack));
  CU_CHECK(c, c->refz.ensure(n, slack));
  CU_CHECK(c, c->vtype.ensure(n, slack));     CU_CHECK(c, c->vtype_alt.ensure(n, slack));
  CU_CHECK(c, c->id.ensure(n, slack));        CU_CHECK(c, c->id_alt.ensure(n, slack));
  CU_CHECK(c, c->numneighs.ensure(n, slack)); CU_CHECK(c, c->nbl_num.ensure(n, slack));
  CU_CHECK(c, c->key.ensure(n, slack));       CU_CHECK(c, c->key_alt.ensure(n, slack));
  CU_CHECK(c, c->perm.ensure(n, slack));      CU_CHECK(c, c->perm_alt.ensure(n, slack));
  // arrays that also hold ghosts start with room for 50 % ghosts and grow on demand
  CU_CHECK(c, c->pos4.ensure(n + n / 2 + 1024));  CU_CHECK(c, c->pos4_alt.ensure(n + 1024, slack));
  CU_CHECK(c, c->type.ensure(n + n / 2 + 1024));  CU_CHECK(c, c->type_alt.ensure(n + 1024, slack));
  return IMDCU_OK;
}

extern "C" int imdcu_upload_atoms(imdcu_ctx *c, int64_t n, const int32_t *id, const int32_t *type,
                                  const int32_t *vtype, const double *mass, const double *pos,
                                  const double *mom)
{
  if (!c || n < 0 || !pos || !mom || !mass)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_upload_atoms: bad argument") : IMDCU_EINVAL;
  if (n >= (int64_t) 1 << 30) return c->fail(IMDCU_EINVAL, "imdcu_upload_atoms: more than 2^30 atoms per rank");
  cudaSetDevice(c->device);
  RET_IF(ensure_atom_capacity(c, n));
  std::vector<double4> hp((size_t) n), hm((size_t) n);
  std::vector<int> ht((size_t) n), hv((size_t) n), hi((size_t) n);
  for (int64_t i = 0; i < n; i++) {
    hp[i] = make_double4(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2], 0.0);
    hm[i] = make_double4(mom[3 * i], mom[3 * i + 1], mom[3 * i + 2], mass[i]);
    ht[i] = type ? type[i] : 0;
    hv[i] = vtype ? vtype[i] : ht[i];
    hi[i] = id ? id[i] : (int) i;
    if (ht[i] < 0 || ht[i] >= c->ntypes) return c->fail(IMDCU_EINVAL, "atom type out of range");
    if (hv[i] < 0 || hv[i] >= c->nvtypes) return c->fail(IMDCU_EINVAL, "atom vtype out of range");
  }
  cudaStream_t s = c->stream;
  CU_CHECK(c, cudaMemcpyAsync(c->pos4.p, hp.data(), n * sizeof(double4), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->mom4.p, hm.data(), n * sizeof(double4), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->type.p, ht.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->vtype.p, hv.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemcpyAsync(c->id.p, hi.data(), n * sizeof(int), cudaMemcpyHostToDevice, s));
  CU_CHECK(c, cudaMemsetAsync(c->frc4.p, 0, n * sizeof(double4), s));
  CU_CHECK(c, cudaMemsetAsync(c->rho.p, 0, n * sizeof(double), s));
  CU_CHECK(c, cudaMemsetAsync(c->numneighs.p, 0, n * sizeof(int), s));
  CU_CHECK(c, cudaMemsetAsync(c->key.p, 0, n * sizeof(int), s));
  CU_CHECK(c, cudaStreamSynchronize(s));
  c->n = n; c->ng = 0;
  c->have_valid_nbl = false;
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
  CU_CHECK(c, cudaStreamSynchronize(s));
  std::vector<double4> h((size_t) n);
  if (pos || dF) {
    CU_CHECK(c, cudaMemcpy(h.data(), c->pos4.p, n * sizeof(double4), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; i++) {
      if (pos) { pos[3 * i] = h[i].x; pos[3 * i + 1] = h[i].y; pos[3 * i + 2] = h[i].z; }
      if (dF) dF[i] = h[i].w;
    }
  }
  if (mom || mass) {
    CU_CHECK(c, cudaMemcpy(h.data(), c->mom4.p, n * sizeof(double4), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; i++) {
      if (mom) { mom[3 * i] = h[i].x; mom[3 * i + 1] = h[i].y; mom[3 * i + 2] = h[i].z; }
      if (mass) mass[i] = h[i].w;
    }
  }
  if (force || epot) {
    CU_CHECK(c, cudaMemcpy(h.data(), c->frc4.p, n * sizeof(double4), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; i++) {
      if (force) { force[3 * i] = h[i].x; force[3 * i + 1] = h[i].y; force[3 * i + 2] = h[i].z; }
      if (epot) epot[i] = h[i].w;
    }
  }
  if (rho) CU_CHECK(c, cudaMemcpy(rho, c->rho.p, n * sizeof(double), cudaMemcpyDeviceToHost));
  if (id) CU_CHECK(c, cudaMemcpy(id, c->id.p, n * sizeof(int), cudaMemcpyDeviceToHost));
  if (type) CU_CHECK(c, cudaMemcpy(type, c->type.p, n * sizeof(int), cudaMemcpyDeviceToHost));
  if (numneighs) CU_CHECK(c, cudaMemcpy(numneighs, c->numneighs.p, n * sizeof(int), cudaMemcpyDeviceToHost));
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
