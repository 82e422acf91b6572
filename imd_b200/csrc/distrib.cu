// imd_b200/csrc/distrib.cu -- spatial distributions on the device (imd_distrib.c:33-218, 377-411,
// 678-716): per-bin count, kinetic energy, potential energy and v_x of the atoms, so that `dist_int` output needs a
// download of dist_size numbers instead of every atom.
//
// The reference loops over the atoms of its cells, finds the bin by (int)(scale * (x - ll)) per direction and adds
// into float arrays (make_distrib_density, make_write_distrib_select with dist_Ekin_fun / dist_Epot_fun /
// dist_vxavg_fun); MPI_Reduce sums the ranks' arrays and rank 0 divides by the count.  Here one kernel does the
// per-rank part: sums in double, returned as float.  Normalisation, the reduction over ranks and the file formats
// stay the host's (imd_distrib.c:720-800).
#include "device.cuh"

struct DistArgs {
  int    dim[3];
  double ll[3], scale[3];
  int    size;
};

// flat bin of a position or -1 (imd_distrib.c:399-408; the conversion truncates toward zero as C's does, so an atom
// less than one bin width below ll lands in bin 0)
__device__ __forceinline__ int dist_bin(const DistArgs &a, double x, double y, double z)
{
  const int nx = (int) (a.scale[0] * (x - a.ll[0]));
  if (nx < 0 || nx >= a.dim[0]) return -1;
  const int ny = (int) (a.scale[1] * (y - a.ll[1]));
  if (ny < 0 || ny >= a.dim[1]) return -1;
  const int nz = (int) (a.scale[2] * (z - a.ll[2]));
  if (nz < 0 || nz >= a.dim[2]) return -1;
  return (nx * a.dim[1] + ny) * a.dim[2] + nz;
}

// Atoms are sorted by cell, so consecutive atoms mostly share a bin: every thread walks DIST_RUN consecutive atoms
// and keeps the sums of the current bin in registers, adding them to the histogram only when the bin changes.
// SMEM: the histogram of the CTA lives in shared memory (4 * size doubles) and is added to global memory once per
// CTA; otherwise (many bins, little contention) the runs go straight to global memory.
#define DIST_RUN 16

template <bool SMEM>
__global__ void __launch_bounds__(256)
k_distrib(int n, const double4 *__restrict__ pos4, const double4 *__restrict__ mom4, const double4 *__restrict__ frc4,
          DistArgs a, double *__restrict__ sums)
{
  extern __shared__ double s_hist[];
  double *hist = SMEM ? s_hist : sums;
  if (SMEM) {
    for (int k = threadIdx.x; k < 4 * a.size; k += blockDim.x) s_hist[k] = 0.0;
    __syncthreads();
  }
  const long long first = ((long long) blockIdx.x * blockDim.x + threadIdx.x) * DIST_RUN;
  int cur = -1;
  double cnt = 0.0, ek = 0.0, ep = 0.0, vx = 0.0;
  for (int r = 0; r < DIST_RUN; r++) {
    const long long i = first + r;
    if (i >= n) break;
    const double4 x = pos4[i];
    const int b = dist_bin(a, x.x, x.y, x.z);
    if (b != cur) {
      if (cur >= 0) {
        atomicAdd(hist + cur, cnt); atomicAdd(hist + a.size + cur, ek);
        atomicAdd(hist + 2 * a.size + cur, ep); atomicAdd(hist + 3 * a.size + cur, vx);
      }
      cur = b; cnt = ek = ep = vx = 0.0;
    }
    if (b < 0) continue;
    const double4 p = mom4[i];
    cnt += 1.0;
    ek += (p.x * p.x + p.y * p.y + p.z * p.z) / (2 * p.w);     // SPRODN(IMPULS,IMPULS) / (2 MASSE), :221
    ep += frc4[i].w;                                           // POTENG, :237
    vx += p.x / p.w;                                           // IMPULS(X) / MASSE, :366
  }
  if (cur >= 0) {
    atomicAdd(hist + cur, cnt); atomicAdd(hist + a.size + cur, ek);
    atomicAdd(hist + 2 * a.size + cur, ep); atomicAdd(hist + 3 * a.size + cur, vx);
  }
  if (SMEM) {
    __syncthreads();
    for (int k = threadIdx.x; k < 4 * a.size; k += blockDim.x)
      if (s_hist[k] != 0.0) atomicAdd(sums + k, s_hist[k]);
  }
}

__global__ void k_dist_to_float(int n, const double *__restrict__ in, float *__restrict__ out)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (float) in[i];
}

// Rank-local per-bin SUMS over the owned atoms (what the reference holds in num_2 / dat_1 before its MPI_Reduce):
// num, ekin, epot, vxavg each [dim0 * dim1 * dim2] floats or NULL.
extern "C" int imdcu_distrib(imdcu_ctx *c, const int dim[3], const double ll[3], const double ur[3],
                             float *num, float *ekin, float *epot, float *vxavg)
{
  if (!c || !dim || !ll || !ur) return c ? c->fail(IMDCU_EINVAL, "imdcu_distrib: bad argument") : IMDCU_EINVAL;
  DistArgs a;
  long long size = 1;
  for (int d = 0; d < 3; d++) {
    if (dim[d] < 1 || !(ur[d] > ll[d])) return c->fail(IMDCU_EINVAL, "imdcu_distrib: dist_dim >= 1 and dist_ur > dist_ll required");
    a.dim[d] = dim[d];
    a.ll[d] = ll[d];
    a.scale[d] = dim[d] / (ur[d] - ll[d]);                     // scalex = dist_dim.x / (dist_ur.x - dist_ll.x), :385
    size *= dim[d];
  }
  if (size > (1ll << 27)) return c->fail(IMDCU_EINVAL, "imdcu_distrib: more than 2^27 bins");
  a.size = (int) size;
  cudaSetDevice(c->device);
  RET_IF(imdcu_join_upload(c));
  cudaStream_t s = c->stream;
  CU_CHECK(c, c->dist_sums.ensure(4 * (size_t) size));
  CU_CHECK(c, c->dist_out.ensure(4 * (size_t) size));
  CU_CHECK(c, cudaMemsetAsync(c->dist_sums.p, 0, 4 * size * sizeof(double), s));
  const int64_t n = c->n;
  if (n > 0) {
    const int nb = div_up(div_up(n, (int64_t) DIST_RUN), 256);
    const size_t smem = 4 * (size_t) size * sizeof(double);
    if (smem <= 32 * 1024)
      k_distrib<true><<<nb, 256, smem, s>>>((int) n, c->pos4.p, c->mom4.p, c->frc4.p, a, c->dist_sums.p);
    else
      k_distrib<false><<<nb, 256, 0, s>>>((int) n, c->pos4.p, c->mom4.p, c->frc4.p, a, c->dist_sums.p);
    CU_CHECK(c, cudaGetLastError());
    c->count_launch();
  }
  k_dist_to_float<<<div_up(4 * size, 256ll), 256, 0, s>>>((int) (4 * size), c->dist_sums.p, c->dist_out.p);
  c->count_launch();
  float *dst[4] = {num, ekin, epot, vxavg};
  for (int q = 0; q < 4; q++)
    if (dst[q]) CU_CHECK(c, cudaMemcpyAsync(dst[q], c->dist_out.p + q * size, size * sizeof(float), cudaMemcpyDeviceToHost, s));
  CU_CHECK(c, cudaStreamSynchronize(s));
  return IMDCU_OK;
}
