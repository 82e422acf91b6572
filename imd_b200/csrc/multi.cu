// imd_b200/csrc/multi.cu -- x-slab domain decomposition over the GPUs of one box.
//
// Replaces the MPI layer of the hot path: setup_mpi_topology (cpu_dim = N 1 1), send_cells with
// copy_cell/pack_cell/unpack_cell (imd_comm_force_3d.c:222-396, 726-1000) and copy_dF/pack_dF
// (:1185-1364), send_atoms of fix_cells (imd_fix_cells_3d.c:43-306) and the MPI_Allreduce calls
// of calc_forces / move_atoms / check_nblist / update_fd.  Transport: ncclSend/ncclRecv pairs
// between x-neighbours and ncclAllReduce, all on the context's stream (NVLink 5 / NVSwitch).
//
// Because the device list is a FULL list, image atoms are read-only: only the two forward
// exchanges exist (positions before pass 1, F' before pass 2, both as the packed 32-byte
// {x,y,z,F'} record); the reference's reverse adds (send_forces: add_rho, add_forces) vanish.
// Both x sides carry a ghost plane (the reference's AR scheme has one).  The planes that
// travel include the y/z image cells of the layer, exactly like the reference's staged
// z,y,x exchange.
#include "device.cuh"

#include <cub/cub.cuh>
#include <dlfcn.h>
#include <nccl.h>

// NCCL is bound at run time, on the first imdcu_comm_* call: single-GPU users never load it, and
// inside a process that already carries an NCCL (e.g. the one bundled with PyTorch, same SONAME)
// that copy is reused instead of pulling a second, different libnccl.so.2 into the process.
namespace {
struct NcclApi {
  bool ok = false;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*CommAbort)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
} g_nccl;

bool nccl_load(std::string *why)
{
  if (g_nccl.ok) return true;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);        // already in the process?
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) { if (why) *why = std::string("cannot load libnccl.so.2: ") + dlerror(); return false; }
#define NCCL_SYM(field, name)                                                            \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, name));               \
  if (!g_nccl.field) { if (why) *why = std::string("libnccl lacks ") + name; return false; }
  NCCL_SYM(GetUniqueId, "ncclGetUniqueId") NCCL_SYM(CommInitRank, "ncclCommInitRank")
  NCCL_SYM(CommDestroy, "ncclCommDestroy") NCCL_SYM(CommAbort, "ncclCommAbort") NCCL_SYM(Send, "ncclSend") NCCL_SYM(Recv, "ncclRecv")
  NCCL_SYM(AllReduce, "ncclAllReduce") NCCL_SYM(GroupStart, "ncclGroupStart") NCCL_SYM(GroupEnd, "ncclGroupEnd")
  NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef NCCL_SYM
  g_nccl.ok = true;
  return true;
}
}
#define ncclGetUniqueId g_nccl.GetUniqueId
#define ncclCommInitRank g_nccl.CommInitRank
#define ncclCommDestroy g_nccl.CommDestroy
#define ncclCommAbort g_nccl.CommAbort
#define ncclSend g_nccl.Send
#define ncclRecv g_nccl.Recv
#define ncclAllReduce g_nccl.AllReduce
#define ncclGroupStart g_nccl.GroupStart
#define ncclGroupEnd g_nccl.GroupEnd
#define ncclGetErrorString g_nccl.GetErrorString

#define NCCL_CHECK(ctx, expr)                                                            \
  do {                                                                                   \
    ncclResult_t _r = (expr);                                                            \
    if (_r != ncclSuccess) {                                                             \
      (ctx)->set_error(std::string(#expr) + ": " + ncclGetErrorString(_r));              \
      return IMDCU_ENCCL;                                                                \
    }                                                                                    \
  } while (0)

static inline ncclComm_t comm_of(imdcu_ctx *c) { return (ncclComm_t) c->comm.nccl; }

extern "C" int imdcu_comm_unique_id(char id[128])
{
  ncclUniqueId u;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  if (!id || !nccl_load(nullptr) || ncclGetUniqueId(&u) != ncclSuccess) return IMDCU_ENCCL;
  memcpy(id, &u, 128);
  return IMDCU_OK;
}

extern "C" int imdcu_comm_init(imdcu_ctx *c, const char id[128], int rank, int nranks)
{
  if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks)
    return c ? c->fail(IMDCU_EINVAL, "imdcu_comm_init: bad argument") : IMDCU_EINVAL;
  if (c->box_set) return c->fail(IMDCU_EINVAL, "imdcu_comm_init must precede imdcu_set_box");
  cudaSetDevice(c->device);
  c->comm.rank = rank; c->comm.nranks = nranks;
  if (nranks == 1) return IMDCU_OK;
  { std::string why; if (!nccl_load(&why)) return c->fail(IMDCU_ENCCL, why); }
  ncclUniqueId u;
  memcpy(&u, id, 128);
  ncclComm_t comm;
  NCCL_CHECK(c, ncclCommInitRank(&comm, nranks, u, rank));
  c->comm.nccl = (void *) comm;
  CU_CHECK(c, cudaMalloc(&c->comm.d_migcnt, 8 * sizeof(int)));
  CU_CHECK(c, cudaMallocHost(&c->comm.h_migcnt, 8 * sizeof(int)));
  return IMDCU_OK;
}

extern "C" int imdcu_comm_rank(const imdcu_ctx *c, int *rank, int *nranks, int *xoff, int *ldimx)
{
  if (!c) return IMDCU_EINVAL;
  if (rank) *rank = c->comm.rank;
  if (nranks) *nranks = c->comm.nranks;
  if (xoff) *xoff = c->geom.xoff;
  if (ldimx) *ldimx = c->geom.ldim[0];
  return IMDCU_OK;
}

void imdcu_multi_release(imdcu_ctx *c)
{
  imdcu_ctx::Comm &m = c->comm;
  m.send_idx.release(); m.send_buf.release(); m.cnt_send.release(); m.cnt_recv.release();
  m.build_send.release(); m.build_recv.release(); m.mig_send.release(); m.mig_recv.release();
  if (m.d_migcnt) cudaFree(m.d_migcnt);
  if (m.h_migcnt) cudaFreeHost(m.h_migcnt);
  if (m.nccl && g_nccl.ok) ncclCommDestroy((ncclComm_t) m.nccl);
  m.nccl = nullptr; m.d_migcnt = m.h_migcnt = nullptr;
}

int imdcu_multi_allreduce(imdcu_ctx *c, void *buf, int count, int kind)
{
  if (c->comm.nranks <= 1 || count <= 0) return IMDCU_OK;
  if (!c->comm.nccl) return c->fail(IMDCU_ENCCL, "the NCCL communicator was aborted after an error on this rank");
  const ncclDataType_t dt = (kind == 2 || kind == 3) ? ncclInt32 : ncclFloat64;
  const ncclRedOp_t op = (kind == 1 || kind == 3) ? ncclMax : ncclSum;
  NCCL_CHECK(c, ncclAllReduce(buf, buf, (size_t) count, dt, op, comm_of(c), c->stream));
  return IMDCU_OK;
}

// Two independent all-reduces as ONE NCCL launch (grouped collectives are fused): the integer and the double slab sums
// of update_fd (imd_ttm.c:325-334), which the reference reduces with eight MPI_Allreduce calls.
int imdcu_multi_allreduce2(imdcu_ctx *c, void *buf_a, int count_a, int kind_a, void *buf_b, int count_b, int kind_b)
{
  if (c->comm.nranks <= 1) return IMDCU_OK;
  if (!c->comm.nccl) return c->fail(IMDCU_ENCCL, "the NCCL communicator was aborted after an error on this rank");
  const ncclDataType_t da = (kind_a == 2 || kind_a == 3) ? ncclInt32 : ncclFloat64, db = (kind_b == 2 || kind_b == 3) ? ncclInt32 : ncclFloat64;
  const ncclRedOp_t oa = (kind_a == 1 || kind_a == 3) ? ncclMax : ncclSum, ob = (kind_b == 1 || kind_b == 3) ? ncclMax : ncclSum;
  NCCL_CHECK(c, ncclGroupStart());
  if (count_a > 0) NCCL_CHECK(c, ncclAllReduce(buf_a, buf_a, (size_t) count_a, da, oa, comm_of(c), c->stream));
  if (count_b > 0) NCCL_CHECK(c, ncclAllReduce(buf_b, buf_b, (size_t) count_b, db, ob, comm_of(c), c->stream));
  NCCL_CHECK(c, ncclGroupEnd());
  return IMDCU_OK;
}

// Error agreement: a rank-local failure (an atom left through an open end, a full neighbour table, a raised
// device flag) must not let this rank leave the step while its peers queue the next collective -- that would
// hang the job instead of failing it.  Every rank contributes its local code; all ranks return together.
// Used on the rebuild path only (one tiny all-reduce + the host sync that path already has).
int imdcu_multi_agree(imdcu_ctx *c, int local_rc)
{
  imdcu_ctx::Comm &m = c->comm;
  if (m.nranks <= 1) return local_rc;
  if (!m.nccl) return local_rc != IMDCU_OK ? local_rc : c->fail(IMDCU_ENCCL, "the NCCL communicator was aborted");
  const std::string local_msg = c->err;
  m.h_migcnt[7] = local_rc != IMDCU_OK ? 1 : 0;
  CU_CHECK(c, cudaMemcpyAsync(m.d_migcnt + 7, m.h_migcnt + 7, sizeof(int), cudaMemcpyHostToDevice, c->stream));
  RET_IF(imdcu_multi_allreduce(c, m.d_migcnt + 7, 1, 3));
  CU_CHECK(c, cudaMemcpyAsync(m.h_migcnt + 7, m.d_migcnt + 7, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CU_CHECK(c, cudaStreamSynchronize(c->stream));
  c->err_agreed = true;
  if (local_rc != IMDCU_OK) { c->err = local_msg; return local_rc; }
  if (m.h_migcnt[7]) return c->fail(IMDCU_EINVAL, "another rank reported an error in this step (see its message)");
  c->err_agreed = false;
  return IMDCU_OK;
}

// A failure that was NOT agreed on (a CUDA or NCCL error in the middle of a step): tear the communicator down so
// that the peers' pending collectives fail instead of waiting forever.
void imdcu_multi_abort(imdcu_ctx *c)
{
  imdcu_ctx::Comm &m = c->comm;
  if (m.nranks > 1 && m.nccl && g_nccl.ok && !c->err_agreed) { ncclCommAbort((ncclComm_t) m.nccl); m.nccl = nullptr; }
  c->err_agreed = false;
}

// one exchange with the x-neighbours: what goes west is received by the west neighbour as ITS
// east plane and vice versa.  *_w: my west plane -> west; *_e: my east plane -> east;
// r_e: from east (its west plane), r_w: from west (its east plane).  Counts in elements of dt.
static int exchange(imdcu_ctx *c, const void *s_w, size_t n_sw, const void *s_e, size_t n_se,
                    void *r_e, size_t n_re, void *r_w, size_t n_rw, ncclDataType_t dt)
{
  const imdcu_ctx::Comm &m = c->comm;
  if (!m.nccl) return c->fail(IMDCU_ENCCL, "the NCCL communicator was aborted after an error on this rank");
  NCCL_CHECK(c, ncclGroupStart());
  if (m.west >= 0 && n_sw) NCCL_CHECK(c, ncclSend(s_w, n_sw, dt, m.west, comm_of(c), c->stream));
  if (m.east >= 0 && n_se) NCCL_CHECK(c, ncclSend(s_e, n_se, dt, m.east, comm_of(c), c->stream));
  if (m.east >= 0 && n_re) NCCL_CHECK(c, ncclRecv(r_e, n_re, dt, m.east, comm_of(c), c->stream));
  if (m.west >= 0 && n_rw) NCCL_CHECK(c, ncclRecv(r_w, n_rw, dt, m.west, comm_of(c), c->stream));
  NCCL_CHECK(c, ncclGroupEnd());
  return IMDCU_OK;
}

// plain x-neighbour exchange of equally sized buffers (3-D FD lattice planes, ttm3d.cu):
// s_w -> west, s_e -> east, r_e <- east, r_w <- west; a neighbour < 0 is an open end.
int imdcu_multi_sendrecv_x(imdcu_ctx *c, int west, int east, const void *s_w, const void *s_e, void *r_e, void *r_w,
                           size_t count, int is_int)
{
  if (c->comm.nranks <= 1) return IMDCU_OK;
  const ncclDataType_t dt = is_int ? ncclInt32 : ncclFloat64;
  NCCL_CHECK(c, ncclGroupStart());
  if (west >= 0) NCCL_CHECK(c, ncclSend(s_w, count, dt, west, comm_of(c), c->stream));
  if (east >= 0) NCCL_CHECK(c, ncclSend(s_e, count, dt, east, comm_of(c), c->stream));
  if (east >= 0) NCCL_CHECK(c, ncclRecv(r_e, count, dt, east, comm_of(c), c->stream));
  if (west >= 0) NCCL_CHECK(c, ncclRecv(r_w, count, dt, west, comm_of(c), c->stream));
  NCCL_CHECK(c, ncclGroupEnd());
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- migration
#define MIG_REC 10        // doubles per migrating atom: pos4, mom4, {type,vtype}, {id,fdi}

__global__ void __launch_bounds__(256)
k_mig_count(int n, int ncell_ext, const int *__restrict__ key, int *__restrict__ cnt)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int w = 0, e = 0;
  if (i < n) { w = key[i] == ncell_ext; e = key[i] == ncell_ext + 1; }
  w = __syncthreads_count(w);
  e = __syncthreads_count(e);
  if (threadIdx.x == 0) { if (w) atomicAdd(&cnt[0], w); if (e) atomicAdd(&cnt[1], e); }
}

// leaver flags for the slot scan: a record's slot is the number of leavers to the same side in front of it in
// the (deterministic) cell-sorted order -- an atomic counter would make the arrival order, and with it the
// intra-cell order and every FP64 summation order downstream, vary from run to run
__global__ void __launch_bounds__(256)
k_mig_flags(int n, int ncell_ext, const int *__restrict__ key, int *__restrict__ fw, int *__restrict__ fe)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = key[i];
  fw[i] = k == ncell_ext; fe[i] = k == ncell_ext + 1;
}

__global__ void __launch_bounds__(256)
k_mig_pack(int n, int ncell_ext, const int *__restrict__ key, const double4 *__restrict__ pos4,
           const double4 *__restrict__ mom4, const int *__restrict__ type, const int *__restrict__ vtype,
           const int *__restrict__ id, const int *__restrict__ fdi, double *__restrict__ rec_w,
           double *__restrict__ rec_e, const int *__restrict__ slot_w, const int *__restrict__ slot_e)
{
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int k = key[i];
  if (k < ncell_ext) return;
  const int side = k - ncell_ext;
  const int slot = side == 0 ? slot_w[i] : slot_e[i];
  double *o = (side == 0 ? rec_w : rec_e) + (size_t) slot * MIG_REC;
  const double4 p = pos4[i], q = mom4[i];
  o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = p.w;
  o[4] = q.x; o[5] = q.y; o[6] = q.z; o[7] = q.w;
  int2 a = make_int2(type[i], vtype[i]), b = make_int2(id[i], fdi ? fdi[i] : 0);
  o[8] = *reinterpret_cast<double *>(&a);
  o[9] = *reinterpret_cast<double *>(&b);
}

// arrivals are appended behind the current atoms; their positions were wrapped by the sender
__global__ void __launch_bounds__(256)
k_mig_unpack(Geom g, int n0, int narr, int ncell_ext, const double *__restrict__ rec, double4 *__restrict__ pos4,
             double4 *__restrict__ mom4, int *__restrict__ type, int *__restrict__ vtype, int *__restrict__ id,
             int *__restrict__ fdi, int *__restrict__ key, int *__restrict__ perm, int *__restrict__ flags)
{
  int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= narr) return;
  const double *o = rec + (size_t) a * MIG_REC;
  const int i = n0 + a;
  pos4[i] = make_double4(o[0], o[1], o[2], o[3]);
  mom4[i] = make_double4(o[4], o[5], o[6], o[7]);
  double d8 = o[8], d9 = o[9];
  const int2 t = *reinterpret_cast<int2 *>(&d8), u = *reinterpret_cast<int2 *>(&d9);
  type[i] = t.x; vtype[i] = t.y; id[i] = u.x;
  if (fdi) fdi[i] = u.y;
  int cc[3];
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const double *tb = g.tbox + 3 * d;
    double s = __dadd_rn(__dadd_rn(__dmul_rn(o[0], tb[0]), __dmul_rn(o[1], tb[1])), __dmul_rn(o[2], tb[2]));
    int cidx = (int) __dmul_rn((double) g.gdim[d], s);
    if (cidx >= g.gdim[d]) cidx = g.gdim[d] - 1; else if (cidx < 0) cidx = 0;
    cc[d] = cidx;
  }
  int li = slab_local_x(g, cc[0]);
  if (li < 0 || li >= g.ldim[0]) { flags[4] = 1; li = li < 0 ? 0 : g.ldim[0] - 1; }   // moved more than one slab
  key[i] = ((li + 1) * g.edim[1] + (cc[1] + 1)) * g.edim[2] + (cc[2] + 1);
  perm[i] = i;
}

template <typename T>
static cudaError_t grow_keep(DevBuf<T> &b, size_t need, size_t keep, cudaStream_t st)
{
  return b.ensure_keep(need, keep, st, 1.2);
}

int imdcu_multi_migrate(imdcu_ctx *c, int *n_total)
{
  imdcu_ctx::Comm &m = c->comm;
  cudaStream_t st = c->stream;
  const int n = (int) c->n;
  int *dc = m.d_migcnt, *hc = m.h_migcnt;
  CU_CHECK(c, cudaMemsetAsync(dc, 0, 8 * sizeof(int), st));
  if (n > 0) { k_mig_count<<<div_up(n, 256), 256, 0, st>>>(n, c->ncell_ext, c->key.p, dc); c->count_launch(); }
  // how many come in: my west-goers are the west neighbour's arrivals from the east, and so on
  RET_IF(exchange(c, dc + 0, 1, dc + 1, 1, dc + 4, 1, dc + 5, 1, ncclInt32));
  CU_CHECK(c, cudaMemcpyAsync(hc, dc, 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_CHECK(c, cudaStreamSynchronize(st));
  imdcu_dbg_mark(c, "mig.counts");
  const int out_w = hc[0], out_e = hc[1];
  const int in_e = m.east >= 0 ? hc[4] : 0, in_w = m.west >= 0 ? hc[5] : 0;
  const int narr = in_e + in_w;
  {
    int rc = IMDCU_OK;
    if ((out_w && m.west < 0) || (out_e && m.east < 0)) rc = c->fail(IMDCU_EINVAL, "atom left the box through an open x end");
    RET_IF(imdcu_multi_agree(c, rc));                 // all ranks leave together, nobody waits in the payload exchange
    imdcu_dbg_mark(c, "mig.agree");
  }
  if (out_w + out_e + narr > 0) {
    CU_CHECK(c, m.mig_send.ensure((size_t) (out_w + out_e) * MIG_REC + 16, 1.5));
    CU_CHECK(c, m.mig_recv.ensure((size_t) narr * MIG_REC + 16, 1.5));
    double *rw = m.mig_send.p, *re = m.mig_send.p + (size_t) out_w * MIG_REC;
    const bool have_fdi = c->ttm.inited && c->ttm.fdi.cap >= (size_t) n;
    if (out_w + out_e > 0) {
      // slots by exclusive scan of the leaver flags (scratch: the sort's output buffers, free until the sort)
      int *fw = c->key_alt.p, *fe = c->perm_alt.p;
      size_t scan_bytes = 0;
      cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, fw, fw, n, st);
      CU_CHECK(c, c->cub_tmp.ensure(scan_bytes + 256, 1.25));
      k_mig_flags<<<div_up(n, 256), 256, 0, st>>>(n, c->ncell_ext, c->key.p, fw, fe);
      CU_CHECK(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, scan_bytes, fw, fw, n, st));
      CU_CHECK(c, cub::DeviceScan::ExclusiveSum(c->cub_tmp.p, scan_bytes, fe, fe, n, st));
      k_mig_pack<<<div_up(n, 256), 256, 0, st>>>(n, c->ncell_ext, c->key.p, c->pos4.p, c->mom4.p, c->type.p, c->vtype.p,
                                                 c->id.p, have_fdi ? c->ttm.fdi.p : nullptr, rw, re, fw, fe);
      c->count_launch(4);
    }
    RET_IF(exchange(c, rw, (size_t) out_w * MIG_REC, re, (size_t) out_e * MIG_REC, m.mig_recv.p, (size_t) in_e * MIG_REC,
                    m.mig_recv.p + (size_t) in_e * MIG_REC, (size_t) in_w * MIG_REC, ncclFloat64));
    if (narr > 0) {
      const size_t need = (size_t) n + narr;
      CU_CHECK(c, grow_keep(c->pos4, need, n, st)); CU_CHECK(c, grow_keep(c->mom4, need, n, st));
      CU_CHECK(c, grow_keep(c->type, need, n, st)); CU_CHECK(c, grow_keep(c->vtype, need, n, st));
      CU_CHECK(c, grow_keep(c->id, need, n, st));   CU_CHECK(c, grow_keep(c->key, need, n, st));
      CU_CHECK(c, grow_keep(c->perm, need, n, st));
      if (c->ttm.inited) { CU_CHECK(c, grow_keep(c->ttm.fdi, need, n, st)); CU_CHECK(c, c->ttm.fdi_alt.ensure(need, 1.2)); }
      CU_CHECK(c, c->mom4_alt.ensure(need, 1.2)); CU_CHECK(c, c->vtype_alt.ensure(need, 1.2));
      CU_CHECK(c, c->id_alt.ensure(need, 1.2));   CU_CHECK(c, c->key_alt.ensure(need, 1.2));
      CU_CHECK(c, c->perm_alt.ensure(need, 1.2)); CU_CHECK(c, c->frc4.ensure(need, 1.2));
      CU_CHECK(c, c->rho.ensure(need, 1.2));      CU_CHECK(c, c->dF.ensure(need, 1.2));
      CU_CHECK(c, c->refx.ensure(need, 1.2));     CU_CHECK(c, c->refy.ensure(need, 1.2));
      CU_CHECK(c, c->refz.ensure(need, 1.2));     CU_CHECK(c, c->numneighs.ensure(need, 1.2));
      CU_CHECK(c, c->nbl_num.ensure(need, 1.2));  CU_CHECK(c, c->stage.ensure(10 * need + 64, 1.2));
      k_mig_unpack<<<div_up(narr, 256), 256, 0, st>>>(c->geom, n, narr, c->ncell_ext, m.mig_recv.p, c->pos4.p, c->mom4.p,
                                                      c->type.p, c->vtype.p, c->id.p,
                                                      c->ttm.inited ? c->ttm.fdi.p : nullptr, c->key.p, c->perm.p,
                                                      c->d_flags);
      c->count_launch();
    }
  }
  *n_total = n + narr;
  c->n = (int64_t) n - out_w - out_e + narr;
  return IMDCU_OK;
}

// ----------------------------------------------------------------------------- ghost planes
__device__ __forceinline__ int plane_cell(const Geom &g, int side, int p)
{
  // side 0: my west plane (x layer 1), side 1: my east plane (x layer ldim); p = j*ez + k
  const int xl = side == 0 ? 1 : g.ldim[0];
  return xl * g.edim[1] * g.edim[2] + p;
}

__global__ void __launch_bounds__(256)
k_plane_counts(Geom g, int nplane, const int *__restrict__ cell_count, int *__restrict__ cnt_send)
{
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= 2 * nplane) return;
  cnt_send[q] = cell_count[plane_cell(g, q / nplane, q % nplane)];
}

// occupancies of my ghost planes: x = ex-1 from the east neighbour, x = 0 from the west neighbour
__global__ void __launch_bounds__(256)
k_plane_set_counts(Geom g, int nplane, int have_e, int have_w, const int *__restrict__ cnt_recv,
                   int *__restrict__ cell_count)
{
  int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= 2 * nplane) return;
  const int side = q / nplane, p = q % nplane;
  if (side == 0 && have_e) cell_count[(g.edim[0] - 1) * g.edim[1] * g.edim[2] + p] = cnt_recv[q];
  if (side == 1 && have_w) cell_count[p] = cnt_recv[q];
}

int imdcu_multi_ghost_counts(imdcu_ctx *c)
{
  imdcu_ctx::Comm &m = c->comm;
  cudaStream_t st = c->stream;
  const int np = m.nplane;
  CU_CHECK(c, m.cnt_send.ensure(4 * (size_t) np + 8)); CU_CHECK(c, m.cnt_recv.ensure(2 * (size_t) np + 8));
  CU_CHECK(c, cudaMemsetAsync(m.cnt_recv.p, 0, 2 * (size_t) np * sizeof(int), st));
  k_plane_counts<<<div_up(2 * np, 256), 256, 0, st>>>(c->geom, np, c->cell_count.p, m.cnt_send.p);
  c->count_launch();
  RET_IF(exchange(c, m.cnt_send.p, np, m.cnt_send.p + np, np, m.cnt_recv.p, np, m.cnt_recv.p + np, np, ncclInt32));
  k_plane_set_counts<<<div_up(2 * np, 256), 256, 0, st>>>(c->geom, np, m.east >= 0, m.west >= 0, m.cnt_recv.p,
                                                          c->cell_count.p);
  c->count_launch();
  // totals and send offsets on the host (rebuilds only)
  std::vector<int> hs(2 * (size_t) np), hr(2 * (size_t) np);
  CU_CHECK(c, cudaMemcpyAsync(hs.data(), m.cnt_send.p, hs.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_CHECK(c, cudaMemcpyAsync(hr.data(), m.cnt_recv.p, hr.size() * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU_CHECK(c, cudaStreamSynchronize(st));
  std::vector<int> off(2 * (size_t) np + 1);
  int acc = 0;
  for (int q = 0; q < 2 * np; q++) { if (q == np) m.nsend[0] = acc; off[q] = acc; acc += hs[q]; }
  off[2 * np] = acc;
  m.nsend[1] = acc - m.nsend[0];
  m.plane_cnt[0] = m.nsend[0]; m.plane_cnt[1] = m.nsend[1];
  if (m.west < 0) m.nsend[0] = 0;          // nothing travels through an open end (offsets stay valid)
  if (m.east < 0) m.nsend[1] = 0;
  m.nrecv[0] = m.nrecv[1] = 0;
  for (int p = 0; p < np; p++) { m.nrecv[0] += hr[p]; m.nrecv[1] += hr[np + p]; }
  // send offsets live behind the counts: cnt_send[2np .. 4np]
  CU_CHECK(c, cudaMemcpyAsync(m.cnt_send.p + 2 * np, off.data(), (2 * (size_t) np + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
  CU_CHECK(c, cudaStreamSynchronize(st));
  return IMDCU_OK;
}

// local indices of the atoms of both send planes, in plane-cell order (one warp per cell)
__global__ void __launch_bounds__(256)
k_plane_index(Geom g, int nplane, const int *__restrict__ cell_begin, const int *__restrict__ cell_count,
              const int *__restrict__ send_off, int *__restrict__ send_idx)
{
  const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (q >= 2 * nplane) return;
  const int cell = plane_cell(g, q / nplane, q % nplane);
  const int b = cell_begin[cell], cnt = cell_count[cell], o = send_off[q];
  for (int k = lane; k < cnt; k += 32) send_idx[o + k] = b + k;
}

#define BUILD_REC 9     // pos4 (shifted), raw[3], {tcode,type}, {id,0}

__global__ void __launch_bounds__(256)
k_pack_build(Geom g, int n, int ns_w, int ns_tot, int code_w, int code_e, const int *__restrict__ send_idx,
             const double4 *__restrict__ pos4, const int *__restrict__ type, const int *__restrict__ id,
             const int *__restrict__ gtcode, const double *__restrict__ graw, const int *__restrict__ ggid,
             double *__restrict__ rec)
{
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ns_tot) return;
  const int i = send_idx[k];
  double4 p = pos4[i];
  double *o = rec + (size_t) k * BUILD_REC;
  int tcode = 13, gid;
  if (i < n) { o[4] = p.x; o[5] = p.y; o[6] = p.z; gid = id[i]; }
  else { const double *r = graw + 3 * (size_t) (i - n); o[4] = r[0]; o[5] = r[1]; o[6] = r[2]; tcode = gtcode[i - n]; gid = ggid[i - n]; }
  apply_shift(g, k < ns_w ? code_w : code_e, 1.0, p.x, p.y, p.z);
  o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = p.w;
  int2 a = make_int2(tcode, type[i]), b = make_int2(gid, 0);
  o[7] = *reinterpret_cast<double *>(&a);
  o[8] = *reinterpret_cast<double *>(&b);
}

__global__ void __launch_bounds__(256)
k_unpack_build(int n, int t0, int cnt, int xcode, const double *__restrict__ rec, double4 *__restrict__ pos4,
               int *__restrict__ type, int *__restrict__ gsrc, int *__restrict__ gcode, int *__restrict__ groot,
               int *__restrict__ gtcode, double *__restrict__ graw, int *__restrict__ ggid)
{
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= cnt) return;
  const double *o = rec + (size_t) k * BUILD_REC;
  const int t = t0 + k;
  pos4[n + t] = make_double4(o[0], o[1], o[2], o[3]);
  graw[3 * (size_t) t] = o[4]; graw[3 * (size_t) t + 1] = o[5]; graw[3 * (size_t) t + 2] = o[6];
  double d7 = o[7], d8 = o[8];
  const int2 a = *reinterpret_cast<int2 *>(&d7), b = *reinterpret_cast<int2 *>(&d8);
  gtcode[t] = imdcu_combine_code(a.x, xcode);
  type[n + t] = a.y;
  ggid[t] = b.x;
  gsrc[t] = -1; gcode[t] = xcode; groot[t] = -1;
}

static inline int xcode(int sx) { return (sx + 1) + 3 + 9; }

int imdcu_multi_ghost_build(imdcu_ctx *c)
{
  imdcu_ctx::Comm &m = c->comm;
  cudaStream_t st = c->stream;
  const int n = (int) c->n, np = m.nplane;
  const int west_cnt = m.plane_cnt[0], tot = m.plane_cnt[0] + m.plane_cnt[1];
  CU_CHECK(c, m.send_idx.ensure((size_t) tot + 8, 1.2));
  CU_CHECK(c, m.send_buf.ensure((size_t) tot + 8, 1.2));
  k_plane_index<<<div_up((int64_t) 2 * np * 32, 256), 256, 0, st>>>(c->geom, np, c->cell_begin.p, c->cell_count.p,
                                                                   m.cnt_send.p + 2 * np, m.send_idx.p);
  c->count_launch();
  // west plane records occupy [0, west_cnt), east plane records [west_cnt, tot)
  CU_CHECK(c, m.build_send.ensure((size_t) tot * BUILD_REC + 16, 1.2));
  CU_CHECK(c, m.build_recv.ensure((size_t) (m.nrecv[0] + m.nrecv[1]) * BUILD_REC + 16, 1.2));
  if (tot > 0) {
    k_pack_build<<<div_up(tot, 256), 256, 0, st>>>(c->geom, n, west_cnt, tot, xcode(m.wrap_w ? +1 : 0), xcode(m.wrap_e ? -1 : 0),
                                                   m.send_idx.p, c->pos4.p, c->type.p, c->id.p, c->ghost_tcode.p,
                                                   c->ghost_raw.p, c->ghost_gid.p, m.build_send.p);
    c->count_launch();
  }
  m.recv_off[0] = c->gstage_atoms[2];
  m.recv_off[1] = c->gstage_atoms[2] + m.nrecv[0];
  RET_IF(exchange(c, m.build_send.p, (size_t) m.nsend[0] * BUILD_REC, m.build_send.p + (size_t) west_cnt * BUILD_REC,
                  (size_t) m.nsend[1] * BUILD_REC, m.build_recv.p, (size_t) m.nrecv[0] * BUILD_REC,
                  m.build_recv.p + (size_t) m.nrecv[0] * BUILD_REC, (size_t) m.nrecv[1] * BUILD_REC, ncclFloat64));
  if (m.nrecv[0] > 0) {
    k_unpack_build<<<div_up(m.nrecv[0], 256), 256, 0, st>>>(n, (int) m.recv_off[0], m.nrecv[0], xcode(m.wrap_e ? +1 : 0),
        m.build_recv.p, c->pos4.p, c->type.p, c->ghost_src.p, c->ghost_code.p, c->ghost_root.p, c->ghost_tcode.p,
        c->ghost_raw.p, c->ghost_gid.p);
    c->count_launch();
  }
  if (m.nrecv[1] > 0) {
    k_unpack_build<<<div_up(m.nrecv[1], 256), 256, 0, st>>>(n, (int) m.recv_off[1], m.nrecv[1], xcode(m.wrap_w ? -1 : 0),
        m.build_recv.p + (size_t) m.nrecv[0] * BUILD_REC, c->pos4.p, c->type.p, c->ghost_src.p, c->ghost_code.p,
        c->ghost_root.p, c->ghost_tcode.p, c->ghost_raw.p, c->ghost_gid.p);
    c->count_launch();
  }
  CU_CHECK(c, cudaGetLastError());
  return IMDCU_OK;
}

__global__ void __launch_bounds__(256)
k_pack_plane(Geom g, int ns_w, int ns_tot, int code_w, int code_e, const int *__restrict__ send_idx,
             const double4 *__restrict__ pos4, double4 *__restrict__ buf)
{
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ns_tot) return;
  double4 p = pos4[send_idx[k]];
  apply_shift(g, k < ns_w ? code_w : code_e, 1.0, p.x, p.y, p.z);
  buf[k] = p;
}

// every step (positions) and between the passes (F' in .w): send_cells(copy_cell) / send_cells(copy_dF)
int imdcu_multi_ghost_refresh(imdcu_ctx *c)
{
  imdcu_ctx::Comm &m = c->comm;
  cudaStream_t st = c->stream;
  const int n = (int) c->n, west_cnt = m.plane_cnt[0], tot_all = m.plane_cnt[0] + m.plane_cnt[1];
  if (tot_all > 0) {
    k_pack_plane<<<div_up(tot_all, 256), 256, 0, st>>>(c->geom, west_cnt, tot_all, xcode(m.wrap_w ? +1 : 0),
                                                       xcode(m.wrap_e ? -1 : 0), m.send_idx.p, c->pos4.p, m.send_buf.p);
    c->count_launch();
  }
  // double4 records travel as 4 doubles each
  RET_IF(exchange(c, m.send_buf.p, (size_t) m.nsend[0] * 4, m.send_buf.p + west_cnt, (size_t) m.nsend[1] * 4,
                  c->pos4.p + n + m.recv_off[0], (size_t) m.nrecv[0] * 4, c->pos4.p + n + m.recv_off[1],
                  (size_t) m.nrecv[1] * 4, ncclFloat64));
  return IMDCU_OK;
}
