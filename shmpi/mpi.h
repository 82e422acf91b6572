/* shmpi/mpi.h -- a stand-in for an MPI library in an image that has none (never part of libimdcu.so).
 *
 * "shmpi": a small single-node MPI subset over fork() + shared memory, written from
 * scratch so that the UNMODIFIED reference (fmqeisfeld/IMD) can be built with -DMPI in an
 * image that has no MPI installation.  It is what lets the reference's MPI-only code paths
 * run here: imd_ttm.c (needs >= 2 ranks), send_cells/send_forces
 * (imd_comm_force_3d.c) and the multi-core CPU baseline of bench.py.
 *
 * Only what the reference calls is provided (see SURVEY.md 8c for the list).  Semantics:
 *   - ranks are processes forked inside MPI_Init/MPI_Init_thread; their number comes from
 *     the environment variable SHMPI_NP (default 1);
 *   - every send is eager/buffered into a per-(src,dst) ring in shared memory, so
 *     MPI_Send never waits for the matching receive (ring size: SHMPI_RING_MB, default 16);
 *   - receives match (source, tag) in order, MPI_ANY_SOURCE / MPI_ANY_TAG supported;
 *   - reductions combine in rank order (deterministic);
 *   - communicators other than MPI_COMM_WORLD only carry a cartesian topology.
 */
#ifndef SHMPI_MPI_H
#define SHMPI_MPI_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPI_VERSION 2
#define MPI_SUBVERSION 0
#define SHMPI 1

typedef int  MPI_Comm;
typedef int  MPI_Datatype;
typedef int  MPI_Op;
typedef int  MPI_Request;
typedef int  MPI_Info;
typedef int  MPI_File;
typedef long MPI_Aint;
typedef long MPI_Offset;

typedef struct {
  int MPI_SOURCE;
  int MPI_TAG;
  int MPI_ERROR;
  int shmpi_nbytes;
} MPI_Status;

#define MPI_SUCCESS      0
#define MPI_COMM_WORLD   0
#define MPI_COMM_NULL   (-1)
#define MPI_INFO_NULL    0
#define MPI_ANY_SOURCE  (-1)
#define MPI_ANY_TAG     (-1)
#define MPI_PROC_NULL   (-2)
#define MPI_REQUEST_NULL (-1)
#define MPI_THREAD_SINGLE     0
#define MPI_THREAD_FUNNELED   1
#define MPI_STATUS_IGNORE   ((MPI_Status *) 0)
#define MPI_STATUSES_IGNORE ((MPI_Status *) 0)

/* basic datatypes (handles < SHMPI_FIRST_DERIVED) */
#define MPI_CHAR     1
#define MPI_SHORT    2
#define MPI_INT      3
#define MPI_LONG     4
#define MPI_FLOAT    5
#define MPI_DOUBLE   6
#define MPI_BYTE     7
#define MPI_PACKED   8
#define MPI_UNSIGNED 9
#define MPI_LB       10
#define MPI_UB       11
#define MPI_UNSIGNED_CHAR 12
#define MPI_DOUBLE_COMPLEX 13
#define MPI_COMPLEX_DOUBLE 13
#define MPI_REAL     5

#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_PROD 4

#define MPI_MODE_RDONLY 1
#define MPI_MODE_WRONLY 2
#define MPI_MODE_CREATE 4

int MPI_Init(int *argc, char ***argv);
int MPI_Init_thread(int *argc, char ***argv, int required, int *provided);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int code);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
double MPI_Wtime(void);
int MPI_Barrier(MPI_Comm comm);

int MPI_Cart_create(MPI_Comm comm, int ndims, int *dims, int *periods, int reorder, MPI_Comm *newcomm);
int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int *coords);
int MPI_Cart_rank(MPI_Comm comm, int *coords, int *rank);

int MPI_Send(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Status *st);
int MPI_Isend(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request *rq);
int MPI_Irecv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Request *rq);
int MPI_Wait(MPI_Request *rq, MPI_Status *st);
int MPI_Waitall(int n, MPI_Request *rq, MPI_Status *st);
int MPI_Waitany(int n, MPI_Request *rq, int *index, MPI_Status *st);
int MPI_Sendrecv(void *sbuf, int scount, MPI_Datatype stype, int dest, int stag,
                 void *rbuf, int rcount, MPI_Datatype rtype, int src, int rtag,
                 MPI_Comm comm, MPI_Status *st);
int MPI_Get_count(MPI_Status *st, MPI_Datatype type, int *count);

int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Reduce(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm);
int MPI_Allreduce(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Exscan(void *sbuf, void *rbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Gather(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
               int root, MPI_Comm comm);
int MPI_Allgather(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
                  MPI_Comm comm);
int MPI_Scatter(void *sbuf, int scount, MPI_Datatype stype, void *rbuf, int rcount, MPI_Datatype rtype,
                int root, MPI_Comm comm);

int MPI_Address(void *location, MPI_Aint *address);
int MPI_Type_struct(int count, int *blocklens, MPI_Aint *displs, MPI_Datatype *types, MPI_Datatype *newtype);
int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype oldtype, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);
int MPI_Type_size(MPI_Datatype type, int *size);
int MPI_Type_extent(MPI_Datatype type, MPI_Aint *extent);
int MPI_Pack_size(int incount, MPI_Datatype type, MPI_Comm comm, int *size);
int MPI_Pack(void *inbuf, int incount, MPI_Datatype type, void *outbuf, int outsize, int *position, MPI_Comm comm);
int MPI_Unpack(void *inbuf, int insize, int *position, void *outbuf, int outcount, MPI_Datatype type, MPI_Comm comm);

int MPI_Alloc_mem(MPI_Aint size, MPI_Info info, void *baseptr);
int MPI_Free_mem(void *base);

/* MPI-IO: declared so that imd_mpiio.c compiles if it is ever enabled; calls abort */
int MPI_File_open(MPI_Comm comm, char *name, int amode, MPI_Info info, MPI_File *fh);
int MPI_File_close(MPI_File *fh);
int MPI_File_set_view(MPI_File fh, MPI_Offset disp, MPI_Datatype etype, MPI_Datatype ftype, char *rep, MPI_Info info);
int MPI_File_write_at(MPI_File fh, MPI_Offset off, void *buf, int count, MPI_Datatype type, MPI_Status *st);
int MPI_File_read_at(MPI_File fh, MPI_Offset off, void *buf, int count, MPI_Datatype type, MPI_Status *st);
int MPI_File_write_all(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Status *st);
int MPI_File_read_all(MPI_File fh, void *buf, int count, MPI_Datatype type, MPI_Status *st);

#ifdef __cplusplus
}
#endif
#endif
