// imd_b200/csrc/nrb.cu -- non-reflecting boundary (imd_nrb.c, after Comput. Mech. 50:645-655 (2012); the
// reference supports fcc (100) surfaces: +x "type 1", -y "type 2", +y "type 3").
//
// Hot path: nrb_forces (imd_nrb.c:792-1129), called at the end of calc_forces for steps > 1
// (imd_forces_nbl.c:1019-1021): the momentum of every boundary atom is REPLACED by
//     p = m * (-k c_self u_self)  +  sum_r 26.9815 * (k c_r u_r - v_r / 4)
// over those of its (up to 12) listed lattice-site neighbours r that are in its neighbour list, with
// u = position - reference position (minimum image in z, and in y if periodic), v_r = the neighbour's
// momentum of BEFORE the call / 26.9815, k = sqrt(nrb_k / 26.9815) (:149) and the direction factors of
// :860-871, :942-953.  The integrators then leave boundary atoms without a kick
// (imd_integrate.c:233-259, 712-751; integrate.cu reads the flag from a bit of vtype).
//
// Setup stays with IMD: init_nrb (:125-768) runs once on the host and its per-atom result (NRBBND,
// NRBI[12], REF_POS) comes in through imdcu_set_nrb for the boundary atoms and their listed neighbours.
// The device keeps that table by ENTRY, a dense atom-number -> entry map, and per entry the current
// local index of the atom (rebuilt after every cell sort).  One thread per boundary entry walks the
// atom's own neighbour list (the device list is a full list: the pair (bnd, neigh) is in it exactly when
// the reference's half list holds it in either orientation), so a step costs two launches over some
// ten thousand threads.  All contributions of an atom are summed by one thread in list order.
//
// What is NOT reproduced: nrb_test_forces (imd_forces_nbl.c:1297-1345, a left-over test that adds
// drand48() forces near x = 1100 A for steps < 300).  What IS reproduced on request (startup_images): for
// steps <= 1, where nrb_forces is not called, the reference's send_forces still adds the momenta of the
// buffer-cell copies of a boundary atom to the atom (imd_comm_force_3d.c:1046-1052, 1157-1168), i.e.
// multiplies its momentum by the number of periodic images of its cell in the rank's cell array.
#include "device.cuh"

#define NRB_MASS 26.9815              // hard-coded in nrb_forces (:805) for the neighbour terms
#define NRB_SQRT2HALF 0.70710678118655   // globals.h:426
#define NRB_DBLSQRT2  2.82842712474619   // globals.h:427

struct NrbArgs {
  int nent, n, maxid, cap;
  double k;
  double boxy, boxz, invy, invz;
  int pby, pbz;
  const int *ent_bnd, *ent_nrbi, *ent_index, *slot_of_id;
  const double *ent_ref;
  double *ent_newp;
  const double4 *pos4;
  double4 *mom4;
  const int *id, *ghost_gid, *ghost_root, *nbl_idx, *nbl_num;
  int *flags;                          // [5]: neighbour on another GPU, [6]: neighbour without reference position
};

__device__ __forceinline__ double nrb_minimg(double d, double box, double inv, int pbc)
{
  if (!pbc) return d;
  const double t = d * inv;            // makros.h:103-106: NINT rounds half away from zero
  const int n = (int) (t < 0 ? (t - 0.5) : (t + 0.5));
  return d - box * n;
}

__global__ void __launch_bounds__(128)
k_nrb_forces(NrbArgs a)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.nent) return;
  const int bnd = a.ent_bnd[s];
  const int i = a.ent_index[s];
  if (bnd <= 0 || i < 0) return;
  int site[12];
#pragma unroll
  for (int r = 0; r < 12; r++) site[r] = a.ent_nrbi[12 * s + r];
  const double4 xi = a.pos4[i];
  const double mi = a.mom4[i].w;
  const double k = a.k;
  // self term (:846-872)
  double ux = xi.x - a.ent_ref[3 * s], uy = xi.y - a.ent_ref[3 * s + 1], uz = xi.z - a.ent_ref[3 * s + 2];
  uz = nrb_minimg(uz, a.boxz, a.invz, a.pbz);
  uy = nrb_minimg(uy, a.boxy, a.invy, a.pby);
  double dx, dy, dz;
  if (bnd == 1) { dx = -k * 4.0 * ux; dy = -k * NRB_DBLSQRT2 * uy; dz = -k * NRB_DBLSQRT2 * uz; }
  else          { dx = -k * NRB_DBLSQRT2 * ux; dy = -k * uy * 4; dz = -k * NRB_DBLSQRT2 * uz; }
  double px = 0.0, py = 0.0, pz = 0.0;
  px += dx * mi; py += dy * mi; pz += dz * mi;
  // neighbour terms (:905-975): every list entry whose atom number is one of the twelve sites
  const int num = a.nbl_num[i];
  const int *slot = a.nbl_idx + ((size_t) (i >> 5) * a.cap) * 32 + (i & 31);
  for (int m = 0; m < num; m++) {
    const int j = slot[(size_t) m * 32];
    const int idj = j < a.n ? a.id[j] : a.ghost_gid[j - a.n];
#pragma unroll
    for (int r = 0; r < 12; r++) {
      if (site[r] != idj) continue;
      const int jr = j < a.n ? j : a.ghost_root[j - a.n];
      const int e = idj >= 0 && idj <= a.maxid ? a.slot_of_id[idj] : -1;
      if (jr < 0) { a.flags[5] = 1; continue; }
      if (e < 0) { a.flags[6] = 1; continue; }
      const double4 xj = a.pos4[j];                      // the image the list holds, like the buffer-cell copy
      const double4 pj = a.mom4[jr];                     // IMPULS_ALT: nothing has been overwritten yet
      double Ux = xj.x - a.ent_ref[3 * e], Uy = xj.y - a.ent_ref[3 * e + 1], Uz = xj.z - a.ent_ref[3 * e + 2];
      Uz = nrb_minimg(Uz, a.boxz, a.invz, a.pbz);
      Uy = nrb_minimg(Uy, a.boxy, a.invy, a.pby);
      const double Vx = pj.x / NRB_MASS, Vy = pj.y / NRB_MASS, Vz = pj.z / NRB_MASS;
      double ex, ey, ez;
      if (bnd == 1) { ex = k * Ux; ey = k * NRB_SQRT2HALF * Uy; ez = k * NRB_SQRT2HALF * Uz; }
      else          { ex = k * NRB_SQRT2HALF * Ux; ey = k * Uy; ez = k * NRB_SQRT2HALF * Uz; }
      ex -= 0.25 * Vx * 1.0; ey -= 0.25 * Vy * 1.0; ez -= 0.25 * Vz * 1.0;
      px += ex * NRB_MASS; py += ey * NRB_MASS; pz += ez * NRB_MASS;
    }
  }
  a.ent_newp[3 * s] = px; a.ent_newp[3 * s + 1] = py; a.ent_newp[3 * s + 2] = pz;
}

__global__ void __launch_bounds__(128)
k_nrb_apply(NrbArgs a)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.nent) return;
  const int i = a.ent_index[s];
  if (a.ent_bnd[s] <= 0 || i < 0) return;
  double4 p = a.mom4[i];
  p.x = a.ent_newp[3 * s]; p.y = a.ent_newp[3 * s + 1]; p.z = a.ent_newp[3 * s + 2];
  a.mom4[i] = p;
}

// steps <= 1 of the reference: momentum times the number of images of the atom's cell (see the header)
__global__ void __launch_bounds__(128)
k_nrb_startup(NrbArgs a, Geom g, const int *__restrict__ key)
{
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= a.nent) return;
  const int i = a.ent_index[s];
  if (a.ent_bnd[s] <= 0 || i < 0) return;
  const int c = key[i];                                  // extended cell index of the last cell sort
  const int cz = c % g.edim[2] - 1, cy = (c / g.edim[2]) % g.edim[1] - 1;
  int mult = 1;
  if (g.pbc[1]) mult *= 1 + (cy == 0) + (cy == g.gdim[1] - 1);
  if (g.pbc[2]) mult *= 1 + (cz == 0) + (cz == g.gdim[2] - 1);
  if (mult == 1) return;
  double4 p = a.mom4[i];
  // the copies are added one after the other (add_forces); for 2 and 4 images the sums are exact doublings
  double sx = p.x, sy = p.y, sz = p.z;
  for (int t = 1; t < mult; t++) { sx += p.x; sy += p.y; sz += p.z; }
  p.x = sx; p.y = sy; p.z = sz;
  a.mom4[i] = p;
}

// vtype bit of the boundary atoms + current local index of every entry
__global__ void __launch_bounds__(256)
k_nrb_index(int n, int maxid, const int *__restrict__ id, const int *__restrict__ slot_of_id,
            const int *__restrict__ ent_bnd, int *__restrict__ ent_index, int *__restrict__ vtype, int mark)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int a = id[i];
  const int s = a >= 0 && a <= maxid ? slot_of_id[a] : -1;
  if (s >= 0) ent_index[s] = i;
  if (mark) {
    const int v = vtype[i] & ~IMDCU_VTYPE_NRB;
    vtype[i] = s >= 0 && ent_bnd[s] > 0 ? v | IMDCU_VTYPE_NRB : v;
  }
}

__global__ void k_nrb_clear(int n, int *__restrict__ vtype)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) vtype[i] &= ~IMDCU_VTYPE_NRB;
}

void imdcu_nrb_release(imdcu_ctx *c)
{
  NrbState &t = c->nrb;
  t.ent_bnd.release(); t.ent_nrbi.release(); t.ent_index.release(); t.slot_of_id.release();
  t.ent_ref.release(); t.ent_newp.release();
  t.active = false; t.nent = 0;
}

// after every cell sort (and at set time): where the entries' atoms are now
int imdcu_nrb_reindex(imdcu_ctx *c, bool mark)
{
  NrbState &t = c->nrb;
  if (!t.active) return IMDCU_OK;
  CU_CHECK(c, cudaMemsetAsync(t.ent_index.p, 0xff, (size_t) t.nent * sizeof(int), c->stream));
  const int n = (int) c->n;
  if (n > 0) {
    k_nrb_index<<<div_up(n, 256), 256, 0, c->stream>>>(n, t.maxid, c->id.p, t.slot_of_id.p, t.ent_bnd.p,
                                                      t.ent_index.p, c->vtype.p, mark ? 1 : 0);
    c->count_launch();
    CU_CHECK(c, cudaGetLastError());
  }
  return IMDCU_OK;
}

extern "C" int imdcu_set_nrb(imdcu_ctx *c, double nrb_k, int64_t nent, const int32_t *id, const int32_t *bnd,
                             const int32_t *nrbi, const double *refpos, int startup_images)
{
  if (!c) return IMDCU_EINVAL;
  if (nent == 0) {                                       // switch off: every atom is kicked again
    if (c->nrb.active && c->n > 0) {
      k_nrb_clear<<<div_up(c->n, 256), 256, 0, c->stream>>>((int) c->n, c->vtype.p);
      c->count_launch();
      CU_CHECK(c, cudaStreamSynchronize(c->stream));
    }
    imdcu_nrb_release(c);
    return IMDCU_OK;
  }
  if (nent < 0 || !id || !bnd || !nrbi || !refpos || !(nrb_k > 0))
    return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: bad arguments");
  if (c->n <= 0) return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: upload the atoms first");
  NrbState &t = c->nrb;
  int maxid = 0;
  for (int64_t s = 0; s < nent; s++) {
    if (id[s] < 0) return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: negative atom number");
    if (bnd[s] < 0 || bnd[s] > 3) return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: boundary type must be 0..3");
    maxid = std::max(maxid, (int) id[s]);
  }
  std::vector<int> map((size_t) maxid + 1, -1);
  for (int64_t s = 0; s < nent; s++) {
    if (map[id[s]] >= 0) return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: atom number listed twice");
    map[id[s]] = (int) s;
  }
  // every listed site of a boundary atom needs its own entry (reference position)
  for (int64_t s = 0; s < nent; s++) {
    if (bnd[s] <= 0) continue;
    for (int r = 0; r < 12; r++) {
      const int q = nrbi[12 * s + r];
      if (q == 0) continue;                              // unset site (atom numbers start at 1, imd_generate.c)
      if (q < 0 || q > maxid || map[q] < 0)
        return c->fail(IMDCU_EINVAL, "imdcu_set_nrb: a listed neighbour has no entry (reference position)");
    }
  }
  t.nent = (int) nent; t.maxid = maxid; t.startup_images = startup_images != 0;
  t.k = sqrt(nrb_k / NRB_MASS);                          // imd_nrb.c:149
  CU_CHECK(c, t.ent_bnd.ensure(nent)); CU_CHECK(c, t.ent_nrbi.ensure(12 * nent)); CU_CHECK(c, t.ent_index.ensure(nent));
  CU_CHECK(c, t.ent_ref.ensure(3 * nent)); CU_CHECK(c, t.ent_newp.ensure(3 * nent));
  CU_CHECK(c, t.slot_of_id.ensure((size_t) maxid + 1));
  CU_CHECK(c, imdcu_h2d(c, t.ent_bnd.p, bnd, nent * sizeof(int)));
  CU_CHECK(c, imdcu_h2d(c, t.ent_nrbi.p, nrbi, 12 * nent * sizeof(int)));
  CU_CHECK(c, imdcu_h2d(c, t.ent_ref.p, refpos, 3 * nent * sizeof(double)));
  CU_CHECK(c, imdcu_h2d(c, t.slot_of_id.p, map.data(), map.size() * sizeof(int)));
  t.active = true;
  RET_IF(imdcu_nrb_reindex(c, true));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  return IMDCU_OK;
}

// nrb_forces(steps) on the current positions, momenta and neighbour list
int imdcu_nrb_forces(imdcu_ctx *c, int steps)
{
  NrbState &t = c->nrb;
  if (!t.active) return IMDCU_OK;
  if (steps <= 1 && !t.startup_images) return IMDCU_OK;
  RET_IF(imdcu_join_upload(c));
  NrbArgs a;
  a.nent = t.nent; a.n = (int) c->n; a.maxid = t.maxid; a.cap = c->nbl_cap;
  a.k = t.k;
  a.boxy = c->geom.box[4]; a.boxz = c->geom.box[8];
  a.invy = 1.0 / a.boxy; a.invz = 1.0 / a.boxz;          // imd_nrb.c:139-140
  a.pby = c->geom.pbc[1]; a.pbz = c->geom.pbc[2];
  a.ent_bnd = t.ent_bnd.p; a.ent_nrbi = t.ent_nrbi.p; a.ent_index = t.ent_index.p; a.slot_of_id = t.slot_of_id.p;
  a.ent_ref = t.ent_ref.p; a.ent_newp = t.ent_newp.p;
  a.pos4 = c->pos4.p; a.mom4 = c->mom4.p;
  a.id = c->id.p; a.ghost_gid = c->ghost_gid.p; a.ghost_root = c->ghost_root.p;
  a.nbl_idx = c->nbl_idx.p; a.nbl_num = c->nbl_num.p;
  a.flags = c->d_flags;
  const int nb = div_up(t.nent, 128);
  if (steps <= 1) {
    k_nrb_startup<<<nb, 128, 0, c->stream>>>(a, c->geom, c->key.p);
    c->count_launch();
  } else {
    k_nrb_forces<<<nb, 128, 0, c->stream>>>(a);
    k_nrb_apply<<<nb, 128, 0, c->stream>>>(a);
    c->count_launch(2);
  }
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}
