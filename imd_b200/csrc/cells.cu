// imd_b200/csrc/cells.cu -- cell binning by device radix sort + ghost (image) atoms.
//
// Replaces do_boundaries (imd_main_3d.c:2018-2105), fix_cells (imd_fix_cells_3d.c:43-306),
// cell_coord (imd_geom_3d.c:1077-1097) and the serial branches of send_cells/copy_cell
// (imd_comm_force_3d.c:222-396, 726-760).  Instead of moving atoms between per-cell
// malloc'ed arrays, atoms are sorted by their cell key; a cell is a [begin,count) range.
#include "device.cuh"

#include <cub/cub.cuh>

// ----------------------------------------------------------------------------- kernels

// PBC wrap (x, then y, then z, each on the updated position) + cell key.
__global__ void __launch_bounds__(256)
k_wrap_key(Geom g, int n, int nranks, int ncell_ext, double4 *__restrict__ pos4, int *__restrict__ key,
           int *__restrict__ perm)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 p = pos4[i];
  double x = p.x, y = p.y, z = p.z;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    if (g.pbc[d] == 1) {
      const double *t = g.tbox + 3 * d, *b = g.box + 3 * d;
      // SPRODX: (x*t.x + y*t.y) + z*t.z, no contraction
      double s = __dadd_rn(__dadd_rn(__dmul_rn(x, t[0]), __dmul_rn(y, t[1])), __dmul_rn(z, t[2]));
      double f = -floor(s);
      x = __dadd_rn(x, __dmul_rn(f, b[0]));
      y = __dadd_rn(y, __dmul_rn(f, b[1]));
      z = __dadd_rn(z, __dmul_rn(f, b[2]));
    }
  }
  int cc[3];
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const double *t = g.tbox + 3 * d;
    double s = __dadd_rn(__dadd_rn(__dmul_rn(x, t[0]), __dmul_rn(y, t[1])), __dmul_rn(z, t[2]));
    int c = (int) __dmul_rn((double) g.gdim[d], s);
    if (c >= g.gdim[d]) c = g.gdim[d] - 1; else if (c < 0) c = 0;
    cc[d] = c;
  }
  p.x = x; p.y = y; p.z = z;
  pos4[i] = p;
  // local extended index.  Decomposed runs: an atom whose cell layer belongs to the west/east
  // neighbour gets the sentinel key ncell_ext / ncell_ext+1 (sorted to the end, sent away by
  // multi.cu); across a periodic x boundary the layer distance is taken modulo gdim.x.
  int li = cc[0] - g.xoff;
  perm[i] = i;
  if (nranks > 1) {
    li = slab_local_x(g, cc[0]);
    if (li < 0) { key[i] = ncell_ext; return; }
    if (li >= g.ldim[0]) { key[i] = ncell_ext + 1; return; }
  } else { if (li < 0) li = 0; else if (li >= g.ldim[0]) li = g.ldim[0] - 1; }
  key[i] = ((li + 1) * g.edim[1] + (cc[1] + 1)) * g.edim[2] + (cc[2] + 1);
}

__global__ void __launch_bounds__(256)
k_gather(int n, const int *__restrict__ perm,
         const double4 *__restrict__ pos_in, const double4 *__restrict__ mom_in,
         const int *__restrict__ type_in, const int *__restrict__ vtype_in, const int *__restrict__ id_in,
         double4 *__restrict__ pos_out, double4 *__restrict__ mom_out,
         int *__restrict__ type_out, int *__restrict__ vtype_out, int *__restrict__ id_out,
         double *__restrict__ refx, double *__restrict__ refy, double *__restrict__ refz,
         const int *__restrict__ fdi_in, int *__restrict__ fdi_out)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int s = perm[i];
  if (fdi_in) fdi_out[i] = fdi_in[s];              // p->fdi travels with the atom (imd_ttm.c:313)
  double4 p = pos_in[s];
  pos_out[i] = p;
  mom_out[i] = mom_in[s];
  type_out[i] = type_in[s];
  vtype_out[i] = vtype_in[s];
  id_out[i] = id_in[s];
  refx[i] = p.x; refy[i] = p.y; refz[i] = p.z;       // NBL_POS = ORT (imd_forces_nbl.c:141-153)
}

// [begin,count) of every real cell from the sorted keys
__global__ void __launch_bounds__(256)
k_cell_bounds(int n, const int *__restrict__ key, int *__restrict__ cell_begin, int *__restrict__ cell_count)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int k = key[i];
  if (i == 0 || key[i - 1] != k) cell_begin[k] = i;
  if (i == n - 1 || key[i + 1] != k) cell_count[k] = i + 1;      // end for now
}
__global__ void __launch_bounds__(256)
k_cell_fix_count(int ncell, const int *__restrict__ cell_begin, int *__restrict__ cell_count)
{
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncell) return;
  int e = cell_count[c];
  cell_count[c] = e > 0 ? e - cell_begin[c] : 0;
}

// ghost cell occupancy = occupancy of its source cell (one stage at a time)
__global__ void __launch_bounds__(256)
k_ghost_counts(int q0, int q1, const GhostCell *__restrict__ plan, int *__restrict__ cell_count,
               int *__restrict__ gcount)
{
  int q = q0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= q1) return;
  if (plan[q].src < 0) { gcount[q] = cell_count[plan[q].cell]; return; }   // remote plane: set by multi.cu
  int c = cell_count[plan[q].src];
  cell_count[plan[q].cell] = c;
  gcount[q] = c;
}

__global__ void __launch_bounds__(256)
k_ghost_begin(int nplan, int n, const GhostCell *__restrict__ plan, const int *__restrict__ goff,
              int *__restrict__ cell_begin)
{
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nplan) return;
  cell_begin[plan[q].cell] = n + goff[q];
}

// per ghost atom: immediate source, stage code, root atom and accumulated code.
// One warp per ghost cell; run stage by stage so that root/tcode of earlier stages exist.
__global__ void __launch_bounds__(256)
k_ghost_fill(int q0, int q1, int n, const GhostCell *__restrict__ plan, const int *__restrict__ goff,
             const int *__restrict__ cell_begin, int *__restrict__ gsrc, int *__restrict__ gcode,
             int *__restrict__ groot, int *__restrict__ gtcode, int *__restrict__ type,
             const int *__restrict__ id, int *__restrict__ ggid)
{
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  int q = q0 + warp;
  if (q >= q1) return;
  GhostCell gc = plan[q];
  if (gc.src < 0) return;                           // remote plane: filled by multi.cu
  int cnt = goff[q + 1] - goff[q];
  int sb = cell_begin[gc.src], tb = goff[q];
  for (int o = lane; o < cnt; o += 32) {
    int s = sb + o, t = tb + o;
    gsrc[t] = s;
    gcode[t] = gc.code;
    int sx = gc.code % 3 - 1, sy = (gc.code / 3) % 3 - 1, sz = gc.code / 9 - 1;
    if (s < n) { groot[t] = s; gtcode[t] = gc.code; ggid[t] = id[s]; }
    else {
      int pc = gtcode[s - n];
      groot[t] = groot[s - n];
      gtcode[t] = (pc % 3 - 1 + sx + 1) + 3 * ((pc / 3) % 3 - 1 + sy + 1) + 9 * (pc / 9 - 1 + sz + 1);
      ggid[t] = ggid[s - n];
    }
    type[n + t] = type[s];
  }
}

// image refresh of one stage: pos4[ghost] = {pos[src] + stage shift, dF[src]}
// (copy_cell, imd_comm_force_3d.c:754-756, and copy_dF :1185-1213 in one go)
template <bool AT_BUILD>
__global__ void __launch_bounds__(256)
k_ghost_refresh(Geom g, int t0, int t1, int n, const int *__restrict__ gsrc, const int *__restrict__ gcode,
                double4 *__restrict__ pos4, double *__restrict__ graw)
{
  int t = t0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= t1) return;
  int s = gsrc[t], code = gcode[t];
  double4 p = pos4[s];
  if (AT_BUILD) {
    // unshifted position of the owning atom, needed once for the orientation rule of
    // the neighbour predicate (nblist.cu)
    if (s < n) { graw[3 * (size_t) t] = p.x; graw[3 * (size_t) t + 1] = p.y; graw[3 * (size_t) t + 2] = p.z; }
    else { const double *r = graw + 3 * (size_t) (s - n);
           graw[3 * (size_t) t] = r[0]; graw[3 * (size_t) t + 1] = r[1]; graw[3 * (size_t) t + 2] = r[2]; }
  }
  apply_shift(g, code, 1.0, p.x, p.y, p.z);
  pos4[n + t] = p;
}

// ----------------------------------------------------------------------------- host side

template <typename T> static inline void swap_buf(DevBuf<T> &a, DevBuf<T> &b)
{
  DevBuf<T> t = a; a = b; b = t;
}

int imdcu_refresh_ghosts(imdcu_ctx *c, bool at_build)
{
  PhaseTimer pt(c, 2);
  const int n = (int) c->n;
  const bool multi = c->comm.nranks > 1;
  for (int s = 0; s < 3; s++) {
    if (s == 2 && multi) {                         // x stage: the planes come from the neighbour ranks
      if (!at_build) RET_IF(imdcu_multi_ghost_refresh(c));
      continue;                                    // (at a rebuild imdcu_multi_ghost_build filled them)
    }
    int t0 = (int) c->gstage_atoms[s], t1 = (int) c->gstage_atoms[s + 1];
    if (t1 <= t0) continue;
    if (at_build)
      k_ghost_refresh<true><<<div_up(t1 - t0, 256), 256, 0, c->stream>>>(
          c->geom, t0, t1, n, c->ghost_src.p, c->ghost_code.p, c->pos4.p, c->ghost_raw.p);
    else
      k_ghost_refresh<false><<<div_up(t1 - t0, 256), 256, 0, c->stream>>>(
          c->geom, t0, t1, n, c->ghost_src.p, c->ghost_code.p, c->pos4.p, nullptr);
    c->count_launch();
  }
  CU_CHECK(c, cudaGetLastError());
  if (at_build) imdcu_dbg_mark(c, "g.refresh_yz");
  if (at_build && multi) RET_IF(imdcu_multi_ghost_build(c));
  if (at_build) imdcu_dbg_mark(c, "g.build_x");
  return IMDCU_OK;
}

int imdcu_rebuild_cells(imdcu_ctx *c)
{
  RET_IF(imdcu_join_upload(c));
  PhaseTimer pt(c, 0);
  int n = (int) c->n;
  cudaStream_t st = c->stream;
  const bool multi = c->comm.nranks > 1;
  if (n == 0 && !multi) { c->ng = 0; return IMDCU_OK; }
  imdcu_dbg_mark(c, nullptr);

  if (n > 0) {
    k_wrap_key<<<div_up(n, 256), 256, 0, st>>>(c->geom, n, c->comm.nranks, c->ncell_ext, c->pos4.p, c->key.p, c->perm.p);
    c->count_launch();
  }
  int n_sort = n;                                   // elements to sort (stayers + movers + arrivals)
  imdcu_dbg_mark(c, "wrap_key");
  if (multi) RET_IF(imdcu_multi_migrate(c, &n_sort));
  imdcu_dbg_mark(c, "migrate");
  const int n_new = (int) c->n;                     // owned atoms after migration (set by multi_migrate)

  // radix sort (key, perm) over just the bits a cell index (+ 2 sentinels) needs
  int bits = 1; while ((1 << bits) < c->ncell_ext + 2) bits++;
  size_t tmp_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, c->key.p, c->key_alt.p, c->perm.p, c->perm_alt.p,
                                  n_sort, 0, bits, st);
  size_t scan_bytes = 0;
  const int nplan = (int) c->gplan.size();
  cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, c->gcell_off.p, c->gcell_off.p, nplan + 1, st);
  // headroom: the sort's temporary storage grows with the element count, and on several GPUs that count reaches a new
  // maximum at many rebuilds (arrivals); without slack each of them is a cudaMalloc + cudaFree that the other ranks wait for
  CU_CHECK(c, c->cub_tmp.ensure(std::max(tmp_bytes, scan_bytes) + 256, 1.25));
  if (n_sort > 0) {
    CU_CHECK(c, cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, tmp_bytes, c->key.p, c->key_alt.p, c->perm.p,
                                                c->perm_alt.p, n_sort, 0, bits, st));
    c->count_launch(4);   // cub: histogram + onesweep passes (approximate count)
  }
  n = n_new;
  const int nb = div_up(std::max(n, 1), 256);

  CU_CHECK(c, c->pos4_alt.ensure_exact(c->pos4.cap));   // same capacity on both sides of the swap
  CU_CHECK(c, c->type_alt.ensure_exact(c->type.cap));
  const bool have_fdi = c->ttm.inited && c->ttm.fdi.cap >= (size_t) n_sort;
  if (n > 0) {
    k_gather<<<nb, 256, 0, st>>>(n, c->perm_alt.p, c->pos4.p, c->mom4.p, c->type.p, c->vtype.p, c->id.p,
                                 c->pos4_alt.p, c->mom4_alt.p, c->type_alt.p, c->vtype_alt.p, c->id_alt.p,
                                 c->refx.p, c->refy.p, c->refz.p, have_fdi ? c->ttm.fdi.p : nullptr, c->ttm.fdi_alt.p);
    c->count_launch();
  }
  if (have_fdi) swap_buf(c->ttm.fdi, c->ttm.fdi_alt);
  swap_buf(c->pos4, c->pos4_alt); swap_buf(c->mom4, c->mom4_alt); swap_buf(c->type, c->type_alt);
  swap_buf(c->vtype, c->vtype_alt); swap_buf(c->id, c->id_alt); swap_buf(c->key, c->key_alt);

  CU_CHECK(c, cudaMemsetAsync(c->cell_begin.p, 0, (c->ncell_ext + 1) * sizeof(int), st));
  CU_CHECK(c, cudaMemsetAsync(c->cell_count.p, 0, (c->ncell_ext + 1) * sizeof(int), st));
  if (n > 0) {
    k_cell_bounds<<<nb, 256, 0, st>>>(n, c->key.p, c->cell_begin.p, c->cell_count.p);
    c->count_launch();
  }
  k_cell_fix_count<<<div_up(c->ncell_ext, 256), 256, 0, st>>>(c->ncell_ext, c->cell_begin.p, c->cell_count.p);
  c->count_launch();
  imdcu_dbg_mark(c, "sort+gather+bounds");

  // ghost cells: counts stage by stage, offsets by one exclusive scan over the plan
  c->ng = 0;
  for (int s = 0; s < 4; s++) c->gstage_atoms[s] = 0;
  if (nplan > 0) {
    CU_CHECK(c, cudaMemsetAsync(c->gcell_off.p, 0, (nplan + 2) * sizeof(int), st));
    for (int s = 0; s < 3; s++) {
      int q0 = c->gstage_cells[s], q1 = c->gstage_cells[s + 1];
      if (q1 > q0) {
        if (s == 2 && multi) RET_IF(imdcu_multi_ghost_counts(c));      // needs the y/z ghost counts of the planes
        k_ghost_counts<<<div_up(q1 - q0, 256), 256, 0, st>>>(q0, q1, c->d_gplan.p, c->cell_count.p, c->gcell_off.p);
        c->count_launch();
      }
    }
    imdcu_dbg_mark(c, "g.counts");
    CU_CHECK(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, scan_bytes, c->gcell_off.p, c->gcell_off.p,
                                              nplan + 1, st));
    c->count_launch();
    int h_off[4];
    for (int s = 0; s < 4; s++)
      CU_CHECK(c, cudaMemcpyAsync(&h_off[s], c->gcell_off.p + c->gstage_cells[s], sizeof(int),
                                  cudaMemcpyDeviceToHost, st));
    CU_CHECK(c, cudaStreamSynchronize(st));
    imdcu_dbg_mark(c, "g.scan");
    for (int s = 0; s < 4; s++) c->gstage_atoms[s] = h_off[s];
    c->ng = h_off[3];
    const int64_t nt = c->n + c->ng;
    CU_CHECK(c, c->pos4.ensure_keep(nt, n, st, 1.1));
    CU_CHECK(c, c->type.ensure_keep(nt, n, st, 1.1));
    CU_CHECK(c, c->ghost_src.ensure(c->ng, 1.1));  CU_CHECK(c, c->ghost_code.ensure(c->ng, 1.1));
    CU_CHECK(c, c->ghost_root.ensure(c->ng, 1.1)); CU_CHECK(c, c->ghost_tcode.ensure(c->ng, 1.1));
    CU_CHECK(c, c->ghost_gid.ensure(c->ng, 1.1));
    CU_CHECK(c, c->ghost_raw.ensure(3 * c->ng, 1.1));
    imdcu_dbg_mark(c, "g.alloc");
    k_ghost_begin<<<div_up(nplan, 256), 256, 0, st>>>(nplan, n, c->d_gplan.p, c->gcell_off.p, c->cell_begin.p);
    c->count_launch();
    for (int s = 0; s < 3; s++) {
      int q0 = c->gstage_cells[s], q1 = c->gstage_cells[s + 1];
      if (q1 > q0 && !(s == 2 && multi)) {
        k_ghost_fill<<<div_up((int64_t) (q1 - q0) * 32, 256), 256, 0, st>>>(
            q0, q1, n, c->d_gplan.p, c->gcell_off.p, c->cell_begin.p, c->ghost_src.p, c->ghost_code.p,
            c->ghost_root.p, c->ghost_tcode.p, c->type.p, c->id.p, c->ghost_gid.p);
        c->count_launch();
      }
    }
  }
  CU_CHECK(c, cudaGetLastError());
  c->cur_phase = -1;
  return IMDCU_OK;
}
