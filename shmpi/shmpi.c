/* shmpi/shmpi.c -- a stand-in for an MPI library in an image that has none: fork + shared-memory MPI subset.
 * See mpi.h in this directory for scope and semantics.  Written from scratch; it contains
 * nothing of the reference. */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <sched.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

/* ------------------------------------------------------------------ shared state */
typedef struct {
  volatile uint64_t head;            /* bytes ever written (producer)  */
  char pad0[56];
  volatile uint64_t tail;            /* bytes ever consumed (consumer) */
  char pad1[56];
} ring_ctl;

typedef struct {
  volatile int abort_flag;
  volatile int bar_count;
  volatile int bar_sense;
  int np;
  size_t ring_bytes;
} shared_hdr;

static shared_hdr *g_sh = NULL;
static ring_ctl   *g_ctl = NULL;     /* [np*np], index src*np+dst */
static char       *g_ring = NULL;    /* [np*np][ring_bytes]       */
static int g_rank = 0, g_np = 1, g_inited = 0;
static int g_bar_sense = 0;
static pid_t *g_children = NULL;

typedef struct { int tag; int nbytes; } msg_hdr;

/* messages popped from a ring while looking for another tag */
typedef struct umsg { struct umsg *next; int tag, nbytes; char *data; } umsg;
static umsg **g_unexp_head = NULL, **g_unexp_tail = NULL;      /* per source */

static void die(const char *m)
{
  fprintf(stderr, "shmpi[%d]: %s\n", g_rank, m);
  if (g_sh) g_sh->abort_flag = 1;
  _exit(3);
}

static inline void spin_pause(unsigned *n)
{
  if (g_sh->abort_flag) _exit(3);
  if (++*n > 64) { sched_yield(); *n = 0; }
  else __builtin_ia32_pause();
}

/* ------------------------------------------------------------------ datatypes */
#define SHMPI_FIRST_DERIVED 32
#define SHMPI_MAX_TYPES 256
typedef struct { int nblocks; long *off; int *len; long extent; int size; } dtype;
static dtype g_types[SHMPI_MAX_TYPES];
static int g_ntypes = SHMPI_FIRST_DERIVED;

static int basic_size(MPI_Datatype t)
{
  switch (t) {
    case MPI_CHAR: case MPI_BYTE: case MPI_PACKED: case MPI_UNSIGNED_CHAR: return 1;
    case MPI_SHORT: return 2;
    case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: return 4;
    case MPI_LONG: case MPI_DOUBLE: return 8;
    case MPI_DOUBLE_COMPLEX: return 16;
    case MPI_LB: case MPI_UB: return 0;
    default: return -1;
  }
}
static int type_size(MPI_Datatype t)
{
  if (t < SHMPI_FIRST_DERIVED) { int s = basic_size(t); if (s < 0) die("unknown datatype"); return s; }
  if (t >= g_ntypes) die("bad derived datatype");
  return g_types[t].size;
}
static long type_extent(MPI_Datatype t)
{
  if (t < SHMPI_FIRST_DERIVED) return basic_size(t);
  return g_types[t].extent;
}
/* append the blocks of `t` placed at byte offset `base` to a growing list */
static void type_blocks(MPI_Datatype t, long base, long **off, int **len, int *n, int *cap)
{
  if (t < SHMPI_FIRST_DERIVED) {
    int s = basic_size(t);
    if (s <= 0) return;
    if (*n > 0 && (*off)[*n - 1] + (*len)[*n - 1] == base) { (*len)[*n - 1] += s; return; }   /* merge */
    if (*n == *cap) { *cap = *cap ? 2 * *cap : 16; *off = realloc(*off, *cap * sizeof(long)); *len = realloc(*len, *cap * sizeof(int)); }
    (*off)[*n] = base; (*len)[*n] = s; (*n)++;
    return;
  }
  const dtype *d = &g_types[t];
  for (int b = 0; b < d->nblocks; b++) {
    long o = base + d->off[b];
    if (*n > 0 && (*off)[*n - 1] + (*len)[*n - 1] == o) { (*len)[*n - 1] += d->len[b]; continue; }
    if (*n == *cap) { *cap = *cap ? 2 * *cap : 16; *off = realloc(*off, *cap * sizeof(long)); *len = realloc(*len, *cap * sizeof(int)); }
    (*off)[*n] = o; (*len)[*n] = d->len[b]; (*n)++;
  }
}
static MPI_Datatype new_type(long *off, int *len, int n, long extent)
{
  if (g_ntypes >= SHMPI_MAX_TYPES) die("too many datatypes");
  dtype *d = &g_types[g_ntypes];
  d->nblocks = n; d->off = off; d->len = len; d->extent = extent; d->size = 0;
  for (int b = 0; b < n; b++) d->size += len[b];
  return g_ntypes++;
}

int MPI_Address(void *location, MPI_Aint *address) { *address = (MPI_Aint) location; return 0; }

int MPI_Type_struct(int count, int *blocklens, MPI_Aint *displs, MPI_Datatype *types, MPI_Datatype *newtype)
{
  long *off = NULL; int *len = NULL; int n = 0, cap = 0;
  long lb = 0, ub = 0; int have_ub = 0, first = 1;
  for (int i = 0; i < count; i++) {
    if (types[i] == MPI_UB) { ub = displs[i]; have_ub = 1; continue; }
    if (types[i] == MPI_LB) { lb = displs[i]; continue; }
    long ext = type_extent(types[i]);
    for (int k = 0; k < blocklens[i]; k++) type_blocks(types[i], displs[i] + k * ext, &off, &len, &n, &cap);
    long end = displs[i] + blocklens[i] * ext;
    if (!have_ub && (first || end > ub)) ub = end;
    first = 0;
  }
  if (!have_ub) ub = (ub + 7) & ~7L;            /* natural alignment epsilon */
  *newtype = new_type(off, len, n, ub - lb);
  return 0;
}
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
  long *off = NULL; int *len = NULL; int n = 0, cap = 0;
  long ext = type_extent(oldtype);
  for (int k = 0; k < count; k++) type_blocks(oldtype, k * ext, &off, &len, &n, &cap);
  *newtype = new_type(off, len, n, count * ext);
  return 0;
}
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype)
{
  long *off = NULL; int *len = NULL; int n = 0, cap = 0;
  long ext = type_extent(oldtype);
  for (int c = 0; c < count; c++)
    for (int k = 0; k < blocklen; k++) type_blocks(oldtype, ((long) c * stride + k) * ext, &off, &len, &n, &cap);
  *newtype = new_type(off, len, n, (((long) count - 1) * stride + blocklen) * ext);
  return 0;
}
int MPI_Type_commit(MPI_Datatype *type) { (void) type; return 0; }
int MPI_Type_free(MPI_Datatype *type) { (void) type; return 0; }
int MPI_Type_size(MPI_Datatype type, int *size) { *size = type_size(type); return 0; }
int MPI_Type_extent(MPI_Datatype type, MPI_Aint *extent) { *extent = type_extent(type); return 0; }

static void pack_bytes(char *dst, const void *buf, int count, MPI_Datatype t)
{
  if (t < SHMPI_FIRST_DERIVED) { memcpy(dst, buf, (size_t) count * basic_size(t)); return; }
  const dtype *d = &g_types[t];
  for (int e = 0; e < count; e++)
    for (int b = 0; b < d->nblocks; b++) {
      memcpy(dst, (const char *) buf + e * d->extent + d->off[b], d->len[b]);
      dst += d->len[b];
    }
}
static void unpack_bytes(void *buf, const char *src, int nbytes, MPI_Datatype t)
{
  if (t < SHMPI_FIRST_DERIVED) { memcpy(buf, src, nbytes); return; }
  const dtype *d = &g_types[t];
  int left = nbytes;
  for (int e = 0; left > 0; e++)
    for (int b = 0; b < d->nblocks && left > 0; b++) {
      int l = d->len[b] < left ? d->len[b] : left;
      memcpy((char *) buf + e * d->extent + d->off[b], src, l);
      src += l; left -= l;
    }
}
int MPI_Pack_size(int incount, MPI_Datatype type, MPI_Comm comm, int *size)
{ (void) comm; *size = incount * type_size(type); return 0; }
int MPI_Pack(void *inbuf, int incount, MPI_Datatype type, void *outbuf, int outsize, int *position, MPI_Comm comm)
{
  (void) comm;
  int nb = incount * type_size(type);
  if (*position + nb > outsize) die("MPI_Pack: buffer too small");
  pack_bytes((char *) outbuf + *position, inbuf, incount, type);
  *position += nb;
  return 0;
}
int MPI_Unpack(void *inbuf, int insize, int *position, void *outbuf, int outcount, MPI_Datatype type, MPI_Comm comm)
{
  (void) comm;
  int nb = outcount * type_size(type);
  if (*position + nb > insize) die("MPI_Unpack: buffer too small");
  unpack_bytes(outbuf, (char *) inbuf + *position, nb, type);
  *position += nb;
  return 0;
}

/* ------------------------------------------------------------------ start-up */
static void on_sigchld(int sig)
{
  (void) sig;
  int st; pid_t p;
  while ((p = waitpid(-1, &st, WNOHANG)) > 0) {
    if (!(WIFEXITED(st) && WEXITSTATUS(st) == 0)) { if (g_sh) g_sh->abort_flag = 1; }
    if (g_children) for (int i = 1; i < g_np; i++) if (g_children[i] == p) g_children[i] = 0;
  }
}

int MPI_Init_thread(int *argc, char ***argv, int required, int *provided)
{
  (void) argc; (void) argv; (void) required;
  if (provided) *provided = MPI_THREAD_FUNNELED;
  if (g_inited) return 0;
  const char *e = getenv("SHMPI_NP");
  g_np = e ? atoi(e) : 1;
  if (g_np < 1 || g_np > 256) { fprintf(stderr, "shmpi: bad SHMPI_NP\n"); exit(2); }
  e = getenv("SHMPI_RING_MB");
  size_t ring = (size_t) (e ? atoi(e) : 16) << 20;
  size_t npair = (size_t) g_np * g_np;
  size_t total = 4096 + npair * sizeof(ring_ctl) + npair * ring;
  char *base = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
  if (base == MAP_FAILED) { perror("shmpi: mmap"); exit(2); }
  g_sh = (shared_hdr *) base;
  g_sh->abort_flag = 0; g_sh->bar_count = 0; g_sh->bar_sense = 0; g_sh->np = g_np; g_sh->ring_bytes = ring;
  g_ctl = (ring_ctl *) (base + 4096);
  g_ring = base + 4096 + npair * sizeof(ring_ctl);
  g_children = calloc(g_np, sizeof(pid_t));
  fflush(NULL);
  g_rank = 0;
  if (g_np > 1) {
    struct sigaction sa; memset(&sa, 0, sizeof sa); sa.sa_handler = on_sigchld; sa.sa_flags = SA_RESTART;
    sigaction(SIGCHLD, &sa, NULL);
  }
  for (int r = 1; r < g_np; r++) {
    pid_t p = fork();
    if (p < 0) { perror("shmpi: fork"); exit(2); }
    if (p == 0) { g_rank = r; signal(SIGCHLD, SIG_DFL); free(g_children); g_children = NULL; break; }
    g_children[r] = p;
  }
  g_unexp_head = calloc(g_np, sizeof(umsg *));
  g_unexp_tail = calloc(g_np, sizeof(umsg *));
  g_inited = 1;
  return 0;
}
int MPI_Init(int *argc, char ***argv) { return MPI_Init_thread(argc, argv, 0, NULL); }

int MPI_Finalize(void)
{
  if (!g_inited) return 0;
  MPI_Barrier(MPI_COMM_WORLD);
  fflush(NULL);
  if (g_rank != 0) _exit(0);
  /* rank 0 waits for the others */
  signal(SIGCHLD, SIG_DFL);
  for (int r = 1; r < g_np; r++)
    if (g_children[r] > 0) { int st; waitpid(g_children[r], &st, 0); }
  g_inited = 0;
  return 0;
}

int MPI_Abort(MPI_Comm comm, int code)
{
  (void) comm;
  fflush(NULL);
  if (g_sh) g_sh->abort_flag = 1;
  if (g_rank == 0 && g_children) for (int r = 1; r < g_np; r++) if (g_children[r] > 0) kill(g_children[r], SIGKILL);
  _exit(code ? code : 1);
}

int MPI_Comm_rank(MPI_Comm comm, int *rank) { (void) comm; *rank = g_rank; return 0; }
int MPI_Comm_size(MPI_Comm comm, int *size) { (void) comm; *size = g_np; return 0; }
double MPI_Wtime(void)
{
  struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int MPI_Barrier(MPI_Comm comm)
{
  (void) comm;
  if (g_np == 1) return 0;
  g_bar_sense = !g_bar_sense;
  if (__atomic_add_fetch(&g_sh->bar_count, 1, __ATOMIC_ACQ_REL) == g_np) {
    g_sh->bar_count = 0;
    __atomic_store_n(&g_sh->bar_sense, g_bar_sense, __ATOMIC_RELEASE);
  } else {
    unsigned n = 0;
    while (__atomic_load_n(&g_sh->bar_sense, __ATOMIC_ACQUIRE) != g_bar_sense) spin_pause(&n);
  }
  return 0;
}

/* ------------------------------------------------------------------ cartesian topology */
#define SHMPI_MAX_COMMS 16
static struct { int ndims, dims[3], periods[3]; } g_cart[SHMPI_MAX_COMMS];
static int g_ncomm = 1;

int MPI_Cart_create(MPI_Comm comm, int ndims, int *dims, int *periods, int reorder, MPI_Comm *newcomm)
{
  (void) comm; (void) reorder;
  if (g_ncomm >= SHMPI_MAX_COMMS || ndims > 3) die("MPI_Cart_create: limits");
  int prod = 1;
  g_cart[g_ncomm].ndims = ndims;
  for (int d = 0; d < ndims; d++) { g_cart[g_ncomm].dims[d] = dims[d]; g_cart[g_ncomm].periods[d] = periods[d]; prod *= dims[d]; }
  if (prod != g_np) die("MPI_Cart_create: grid size != number of ranks");
  *newcomm = g_ncomm++;
  return 0;
}
int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int *coords)
{
  if (comm <= 0 || comm >= g_ncomm) die("MPI_Cart_coords: not a cartesian communicator");
  int nd = g_cart[comm].ndims;
  for (int d = nd - 1; d >= 0; d--) {
    if (d < maxdims) coords[d] = rank % g_cart[comm].dims[d];
    rank /= g_cart[comm].dims[d];
  }
  return 0;
}
int MPI_Cart_rank(MPI_Comm comm, int *coords, int *rank)
{
  if (comm <= 0 || comm >= g_ncomm) die("MPI_Cart_rank: not a cartesian communicator");
  int nd = g_cart[comm].ndims, r = 0;
  for (int d = 0; d < nd; d++) {
    int c = coords[d], n = g_cart[comm].dims[d];
    if (c < 0 || c >= n) {
      if (!g_cart[comm].periods[d]) die("MPI_Cart_rank: coordinate out of range in a non-periodic dimension");
      c = ((c % n) + n) % n;
    }
    r = r * n + c;
  }
  *rank = r;
  return 0;
}

/* ------------------------------------------------------------------ point to point */
static void ring_write(int dst, int tag, const char *data, int nbytes)
{
  ring_ctl *c = &g_ctl[(size_t) g_rank * g_np + dst];
  char *ring = g_ring + ((size_t) g_rank * g_np + dst) * g_sh->ring_bytes;
  const size_t cap = g_sh->ring_bytes;
  const size_t need = sizeof(msg_hdr) + (((size_t) nbytes + 7) & ~(size_t) 7);
  if (need > cap) {
    fprintf(stderr, "shmpi[%d]: message of %d bytes (tag %d, to rank %d) does not fit the ring of %zu bytes\n", g_rank, nbytes, tag, dst, cap);
    die("message larger than the ring: raise SHMPI_RING_MB");
  }
  unsigned n = 0;
  uint64_t head = c->head;
  while (cap - (head - __atomic_load_n(&c->tail, __ATOMIC_ACQUIRE)) < need) spin_pause(&n);
  msg_hdr h = { tag, nbytes };
  size_t pos = head % cap;                      /* cap and all sizes are multiples of 8: header never wraps */
  memcpy(ring + pos, &h, sizeof h);
  pos = (pos + sizeof h) % cap;
  size_t first = cap - pos < (size_t) nbytes ? cap - pos : (size_t) nbytes;
  memcpy(ring + pos, data, first);
  if ((size_t) nbytes > first) memcpy(ring, data + first, nbytes - first);
  __atomic_store_n(&c->head, head + need, __ATOMIC_RELEASE);
}

/* pop the next message from src's ring into a fresh umsg; returns NULL if the ring is empty */
static umsg *ring_pop(int src)
{
  ring_ctl *c = &g_ctl[(size_t) src * g_np + g_rank];
  char *ring = g_ring + ((size_t) src * g_np + g_rank) * g_sh->ring_bytes;
  const size_t cap = g_sh->ring_bytes;
  uint64_t tail = c->tail;
  if (__atomic_load_n(&c->head, __ATOMIC_ACQUIRE) == tail) return NULL;
  msg_hdr h;
  size_t pos = tail % cap;
  memcpy(&h, ring + pos, sizeof h);
  pos = (pos + sizeof h) % cap;
  umsg *m = malloc(sizeof(umsg));
  m->next = NULL; m->tag = h.tag; m->nbytes = h.nbytes; m->data = malloc(h.nbytes > 0 ? h.nbytes : 1);
  size_t first = cap - pos < (size_t) h.nbytes ? cap - pos : (size_t) h.nbytes;
  memcpy(m->data, ring + pos, first);
  if ((size_t) h.nbytes > first) memcpy(m->data + first, ring, h.nbytes - first);
  __atomic_store_n(&c->tail, tail + sizeof h + (((size_t) h.nbytes + 7) & ~(size_t) 7), __ATOMIC_RELEASE);
  return m;
}

/* find (and unlink) a matching message from src: unexpected list first, then the ring */
static umsg *match_from(int src, int tag)
{
  umsg *prev = NULL;
  for (umsg *m = g_unexp_head[src]; m; prev = m, m = m->next)
    if (tag == MPI_ANY_TAG ? m->tag >= 0 : m->tag == tag) {
      if (prev) prev->next = m->next; else g_unexp_head[src] = m->next;
      if (g_unexp_tail[src] == m) g_unexp_tail[src] = prev;
      return m;
    }
  for (;;) {
    umsg *m = ring_pop(src);
    if (!m) return NULL;
    if (tag == MPI_ANY_TAG ? m->tag >= 0 : m->tag == tag) return m;
    if (g_unexp_tail[src]) g_unexp_tail[src]->next = m; else g_unexp_head[src] = m;
    g_unexp_tail[src] = m;
  }
}

static void do_send(const void *buf, int count, MPI_Datatype type, int dest, int tag)
{
  if (dest == MPI_PROC_NULL) return;
  if (dest < 0 || dest >= g_np) die("send: bad destination rank");
  int nb = count * type_size(type);
  if (type < SHMPI_FIRST_DERIVED) { ring_write(dest, tag, buf, nb); return; }
  char *tmp = malloc(nb > 0 ? nb : 1);
  pack_bytes(tmp, buf, count, type);
  ring_write(dest, tag, tmp, nb);
  free(tmp);
}

static void do_recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Status *st)
{
  if (src == MPI_PROC_NULL) { if (st) { st->MPI_SOURCE = src; st->MPI_TAG = tag; st->shmpi_nbytes = 0; } return; }
  umsg *m = NULL; unsigned n = 0; int from = src;
  if (src == MPI_ANY_SOURCE) {
    for (;;) {
      for (int s = 0; s < g_np && !m; s++) { m = match_from(s, tag); if (m) from = s; }
      if (m) break;
      spin_pause(&n);
    }
  } else {
    if (src < 0 || src >= g_np) die("recv: bad source rank");
    while (!(m = match_from(src, tag))) spin_pause(&n);
  }
  int cap = count * type_size(type);
  if (m->nbytes > cap) die("recv: message truncated");
  unpack_bytes(buf, m->data, m->nbytes, type);
  if (st) { st->MPI_SOURCE = from; st->MPI_TAG = m->tag; st->MPI_ERROR = 0; st->shmpi_nbytes = m->nbytes; }
  free(m->data); free(m);
}

int MPI_Send(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm)
{ (void) comm; do_send(buf, count, type, dest, tag); return 0; }
int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Status *st)
{ (void) comm; do_recv(buf, count, type, src, tag, st); return 0; }
int MPI_Sendrecv(void *sbuf, int scount, MPI_Datatype stype, int dest, int stag,
                 void *rbuf, int rcount, MPI_Datatype rtype, int src, int rtag, MPI_Comm comm, MPI_Status *st)
{ (void) comm; do_send(sbuf, scount, stype, dest, stag); do_recv(rbuf, rcount, rtype, src, rtag, st); return 0; }
int MPI_Get_count(MPI_Status *st, MPI_Datatype type, int *count)
{ int s = type_size(type); *count = s ? st->shmpi_nbytes / s : 0; return 0; }

/* nonblocking: sends are eager (done at once); receives complete in Wait */
#define SHMPI_MAX_REQ 1024
static struct { int used; void *buf; int count, src, tag; MPI_Datatype type; } g_req[SHMPI_MAX_REQ];

int MPI_Isend(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *rq)
{ (void) comm; do_send(buf, count, type, dest, tag); *rq = MPI_REQUEST_NULL; return 0; }
int MPI_Irecv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Request *rq)
{
  (void) comm;
  for (int i = 0; i < SHMPI_MAX_REQ; i++)
    if (!g_req[i].used) {
      g_req[i].used = 1; g_req[i].buf = buf; g_req[i].count = count; g_req[i].src = src; g_req[i].tag = tag; g_req[i].type = type;
      *rq = i; return 0;
    }
  die("too many pending requests");
  return 1;
}
int MPI_Wait(MPI_Request *rq, MPI_Status *st)
{
  if (*rq == MPI_REQUEST_NULL) return 0;
  int i = *rq;
  do_recv(g_req[i].buf, g_req[i].count, g_req[i].type, g_req[i].src, g_req[i].tag, st);
  g_req[i].used = 0; *rq = MPI_REQUEST_NULL;
  return 0;
}
int MPI_Waitall(int n, MPI_Request *rq, MPI_Status *st)
{ for (int i = 0; i < n; i++) MPI_Wait(&rq[i], st ? &st[i] : NULL); return 0; }
int MPI_Waitany(int n, MPI_Request *rq, int *index, MPI_Status *st)
{
  for (int i = 0; i < n; i++) if (rq[i] != MPI_REQUEST_NULL) { *index = i; return MPI_Wait(&rq[i], st); }
  *index = -1;
  return 0;
}

/* ------------------------------------------------------------------ collectives */
#define TAG_BCAST   (-101)
#define TAG_REDUCE  (-102)
#define TAG_GATHER  (-103)
#define TAG_SCATTER (-104)
#define TAG_SCAN    (-105)

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm)
{
  (void) comm;
  if (g_np == 1) return 0;
  if (g_rank == root) { for (int r = 0; r < g_np; r++) if (r != root) do_send(buf, count, type, r, TAG_BCAST); }
  else do_recv(buf, count, type, root, TAG_BCAST, NULL);
  return 0;
}

#define COMBINE(T)                                                                      \
  do { T *a = (T *) acc; const T *b = (const T *) in;                                   \
    for (int i = 0; i < count; i++) {                                                   \
      switch (op) {                                                                     \
        case MPI_SUM: a[i] = a[i] + b[i]; break;                                        \
        case MPI_PROD: a[i] = a[i] * b[i]; break;                                       \
        case MPI_MAX: if (b[i] > a[i]) a[i] = b[i]; break;                              \
        case MPI_MIN: if (b[i] < a[i]) a[i] = b[i]; break;                              \
        default: die("unsupported reduction op"); } } } while (0)

static void combine(void *acc, const void *in, int count, MPI_Datatype type, MPI_Op op)
{
  switch (type) {
    case MPI_INT: COMBINE(int); break;
    case MPI_LONG: COMBINE(long); break;
    case MPI_DOUBLE: COMBINE(double); break;
    case MPI_FLOAT: COMBINE(float); break;
    case MPI_SHORT: COMBINE(short); break;
    case MPI_CHAR: COMBINE(char); break;
    case MPI_UNSIGNED: COMBINE(unsigned); break;
    default: die("unsupported reduction datatype");
  }
}

int MPI_Reduce(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm)
{
  (void) comm;
  size_t nb = (size_t) count * type_size(type);
  if (g_rank != root) { do_send(sbuf, count, type, root, TAG_REDUCE); return 0; }
  char *tmp = malloc(nb ? nb : 1), *acc = malloc(nb ? nb : 1);
  /* combine in rank order 0,1,2,... whatever the root is */
  for (int r = 0; r < g_np; r++) {
    const void *in;
    if (r == root) in = sbuf; else { do_recv(tmp, count, type, r, TAG_REDUCE, NULL); in = tmp; }
    if (r == 0) memcpy(acc, in, nb); else combine(acc, in, count, type, op);
  }
  memcpy(rbuf, acc, nb);
  free(tmp); free(acc);
  return 0;
}
int MPI_Allreduce(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
  MPI_Reduce(sbuf, rbuf, count, type, op, 0, comm);
  return MPI_Bcast(rbuf, count, type, 0, comm);
}
int MPI_Exscan(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{
  (void) comm;
  size_t nb = (size_t) count * type_size(type);
  char *acc = malloc(nb ? nb : 1);
  if (g_rank > 0) { do_recv(acc, count, type, g_rank - 1, TAG_SCAN, NULL); memcpy(rbuf, acc, nb); combine(acc, sbuf, count, type, op); }
  else memcpy(acc, sbuf, nb);
  if (g_rank < g_np - 1) do_send(acc, count, type, g_rank + 1, TAG_SCAN);
  free(acc);
  return 0;
}
int MPI_Gather(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
               int root, MPI_Comm comm)
{
  (void) comm;
  if (g_rank != root) { do_send(sbuf, scount, stype, root, TAG_GATHER); return 0; }
  long ext = type_extent(rtype);
  for (int r = 0; r < g_np; r++) {
    char *dst = (char *) rbuf + (long) r * rcount * ext;
    if (r == root) {
      int nb = scount * type_size(stype);
      char *tmp = malloc(nb ? nb : 1);
      pack_bytes(tmp, sbuf, scount, stype); unpack_bytes(dst, tmp, nb, rtype); free(tmp);
    } else do_recv(dst, rcount, rtype, r, TAG_GATHER, NULL);
  }
  return 0;
}
int MPI_Allgather(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
                  MPI_Comm comm)
{
  MPI_Gather(sbuf, scount, stype, rbuf, rcount, rtype, 0, comm);
  return MPI_Bcast(rbuf, rcount * g_np, rtype, 0, comm);
}
int MPI_Scatter(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
                int root, MPI_Comm comm)
{
  (void) comm;
  if (g_rank == root) {
    long ext = type_extent(stype);
    for (int r = 0; r < g_np; r++) {
      char *src = (char *) sbuf + (long) r * scount * ext;
      if (r == root) {
        int nb = scount * type_size(stype);
        char *tmp = malloc(nb ? nb : 1);
        pack_bytes(tmp, src, scount, stype); unpack_bytes(rbuf, tmp, nb, rtype); free(tmp);
      } else do_send(src, scount, stype, r, TAG_SCATTER);
    }
  } else do_recv(rbuf, rcount, rtype, root, TAG_SCATTER, NULL);
  return 0;
}

/* ------------------------------------------------------------------ memory, MPI-IO stubs */
int MPI_Alloc_mem(MPI_Aint size, MPI_Info info, void *baseptr)
{ (void) info; void *p = malloc(size > 0 ? size : 1); *(void **) baseptr = p; return p ? 0 : 1; }
int MPI_Free_mem(void *base) { free(base); return 0; }

static int no_io(void) { die("MPI-IO is not provided by shmpi"); return 1; }
int MPI_File_open(MPI_Comm c, char *n, int a, MPI_Info i, MPI_File *f) { (void) c; (void) n; (void) a; (void) i; (void) f; return no_io(); }
int MPI_File_close(MPI_File *f) { (void) f; return no_io(); }
int MPI_File_set_view(MPI_File f, MPI_Offset d, MPI_Datatype e, MPI_Datatype t, char *r, MPI_Info i)
{ (void) f; (void) d; (void) e; (void) t; (void) r; (void) i; return no_io(); }
int MPI_File_write_at(MPI_File f, MPI_Offset o, void *b, int c, MPI_Datatype t, MPI_Status *s)
{ (void) f; (void) o; (void) b; (void) c; (void) t; (void) s; return no_io(); }
int MPI_File_read_at(MPI_File f, MPI_Offset o, void *b, int c, MPI_Datatype t, MPI_Status *s)
{ (void) f; (void) o; (void) b; (void) c; (void) t; (void) s; return no_io(); }
int MPI_File_write_all(MPI_File f, void *b, int c, MPI_Datatype t, MPI_Status *s)
{ (void) f; (void) b; (void) c; (void) t; (void) s; return no_io(); }
int MPI_File_read_all(MPI_File f, void *b, int c, MPI_Datatype t, MPI_Status *s)
{ (void) f; (void) b; (void) c; (void) t; (void) s; return no_io(); }
