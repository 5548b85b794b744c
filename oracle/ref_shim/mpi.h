/* TEST INFRASTRUCTURE ONLY - single-rank stand-in for <mpi.h>.
 *
 * The reference (MP-Gadget/shenqi) needs MPI to compile; MPI is not installed in
 * this image.  This header lets the reference's own, unmodified sources for the
 * short-range gravity path compile into oracle/_ref/ as a one-rank program.
 * Every collective degenerates to a local copy; point-to-point calls work only
 * from rank 0 to itself and abort for any other partner.
 *
 * Nothing in the product (shenqi_b200/) includes this file.
 */
#ifndef SHENQI_ORACLE_MPI_SHIM_H
#define SHENQI_ORACLE_MPI_SHIM_H

#include <stddef.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef long MPI_Datatype;   /* encodes the byte size of one element */
typedef int MPI_Op;
typedef int MPI_Request;
typedef long MPI_Aint;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 1
#define MPI_COMM_NULL 0
#define MPI_SUCCESS 0
#define MPI_UNDEFINED (-32766)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_REQUEST_NULL (-1)
#define MPI_STATUS_IGNORE ((MPI_Status *) 0)
#define MPI_STATUSES_IGNORE ((MPI_Status *) 0)
#define MPI_IN_PLACE ((void *) 1)
#define MPI_THREAD_FUNNELED 1
#define MPI_MAX_PROCESSOR_NAME 64

#define MPI_DATATYPE_NULL ((MPI_Datatype) 0)
#define MPI_BYTE ((MPI_Datatype) 1)
#define MPI_CHAR ((MPI_Datatype) 1)
#define MPI_INT ((MPI_Datatype) 4)
#define MPI_UNSIGNED ((MPI_Datatype) 4)
#define MPI_FLOAT ((MPI_Datatype) 4)
#define MPI_LONG ((MPI_Datatype) 8)
#define MPI_INT64 ((MPI_Datatype) 8)
#define MPI_INT64_T ((MPI_Datatype) 8)
#define MPI_UINT64 ((MPI_Datatype) 8)
#define MPI_UINT64_T ((MPI_Datatype) 8)
#define MPI_LONG_LONG ((MPI_Datatype) 8)
#define MPI_UNSIGNED_LONG ((MPI_Datatype) 8)
#define MPI_DOUBLE ((MPI_Datatype) 8)
#define MPI_AINT ((MPI_Datatype) 8)

#define MPI_SUM 1
#define MPI_MAX 2
#define MPI_MIN 3
#define MPI_LOR 4
#define MPI_LAND 5
#define MPI_BOR 6

static inline void shim_mpi_unreachable(const char * what)
{
    fprintf(stderr, "oracle mpi shim: %s reached with a single rank\n", what);
    abort();
}

static inline int shim_mpi_copy(const void * sendbuf, void * recvbuf, long count, MPI_Datatype type)
{
    if(sendbuf != MPI_IN_PLACE && sendbuf != recvbuf && count > 0)
        memmove(recvbuf, sendbuf, (size_t) count * (size_t) type);
    return MPI_SUCCESS;
}

static inline int MPI_Init(int * argc, char *** argv) { (void) argc; (void) argv; return MPI_SUCCESS; }
static inline int MPI_Init_thread(int * argc, char *** argv, int required, int * provided)
{ (void) argc; (void) argv; if(provided) *provided = required; return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Initialized(int * flag) { *flag = 1; return MPI_SUCCESS; }
static inline int MPI_Abort(MPI_Comm comm, int code) { (void) comm; fprintf(stderr, "MPI_Abort(%d)\n", code); abort(); return 0; }
static inline int MPI_Comm_rank(MPI_Comm comm, int * rank) { (void) comm; *rank = 0; return MPI_SUCCESS; }
/* One rank - except that a translation unit may set shim_mpi_fake_size to drive PURELY LOCAL functions of the reference that
 * only read the task count (domain_assign_topleaves_balanced, domain.cpp:620-778) for more than one task; no communication
 * works at such a size (every collective is still a local copy). */
static int shim_mpi_fake_size = 0;
static inline int MPI_Comm_size(MPI_Comm comm, int * size) { (void) comm; *size = shim_mpi_fake_size > 0 ? shim_mpi_fake_size : 1; return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm comm) { (void) comm; return MPI_SUCCESS; }
/* the 1 x 1 cartesian task grid of the PM (petapm.cpp:233-238, 338, 1220) */
static inline int MPI_Cart_create(MPI_Comm comm, int ndims, const int * dims, const int * periods, int reorder, MPI_Comm * cart)
{ (void) periods; (void) reorder; for(int i = 0; i < ndims; i++) if(dims[i] != 1) shim_mpi_unreachable("MPI_Cart_create with more than one task"); *cart = comm; return MPI_SUCCESS; }
static inline int MPI_Cart_get(MPI_Comm comm, int maxdims, int * dims, int * periods, int * coords)
{ (void) comm; for(int i = 0; i < maxdims; i++) { dims[i] = 1; periods[i] = 1; coords[i] = 0; } return MPI_SUCCESS; }
static inline int MPI_Cart_rank(MPI_Comm comm, const int * coords, int * rank) { (void) comm; (void) coords; *rank = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_free(MPI_Comm * comm) { *comm = MPI_COMM_NULL; return MPI_SUCCESS; }
static inline double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}
static inline int MPI_Get_processor_name(char * name, int * len) { strcpy(name, "oracle"); *len = 6; return MPI_SUCCESS; }

static inline int MPI_Type_contiguous(int count, MPI_Datatype oldtype, MPI_Datatype * newtype)
{ *newtype = (MPI_Datatype) count * oldtype; return MPI_SUCCESS; }
static inline int MPI_Type_commit(MPI_Datatype * type) { (void) type; return MPI_SUCCESS; }
static inline int MPI_Type_free(MPI_Datatype * type) { *type = MPI_DATATYPE_NULL; return MPI_SUCCESS; }
static inline int MPI_Type_get_extent(MPI_Datatype type, MPI_Aint * lb, MPI_Aint * extent)
{ *lb = 0; *extent = (MPI_Aint) type; return MPI_SUCCESS; }
static inline int MPI_Type_size(MPI_Datatype type, int * size) { *size = (int) type; return MPI_SUCCESS; }

static inline int MPI_Bcast(void * buf, int count, MPI_Datatype type, int root, MPI_Comm comm)
{ (void) buf; (void) count; (void) type; (void) root; (void) comm; return MPI_SUCCESS; }
static inline int MPI_Reduce(const void * sendbuf, void * recvbuf, int count, MPI_Datatype type, MPI_Op op, int root, MPI_Comm comm)
{ (void) op; (void) root; (void) comm; return shim_mpi_copy(sendbuf, recvbuf, count, type); }
static inline int MPI_Allreduce(const void * sendbuf, void * recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm)
{ (void) op; (void) comm; return shim_mpi_copy(sendbuf, recvbuf, count, type); }
static inline int MPI_Alltoall(const void * sendbuf, int sendcount, MPI_Datatype sendtype, void * recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm)
{ (void) recvcount; (void) recvtype; (void) comm; return shim_mpi_copy(sendbuf, recvbuf, sendcount, sendtype); }
static inline int MPI_Alltoallv(const void * sendbuf, const int * sendcounts, const int * sdispls, MPI_Datatype sendtype,
        void * recvbuf, const int * recvcounts, const int * rdispls, MPI_Datatype recvtype, MPI_Comm comm)
{
    (void) recvcounts; (void) recvtype; (void) comm;
    if(sendbuf == MPI_IN_PLACE) return MPI_SUCCESS;
    memmove((char *) recvbuf + (size_t) rdispls[0] * recvtype, (const char *) sendbuf + (size_t) sdispls[0] * sendtype, (size_t) sendcounts[0] * sendtype);
    return MPI_SUCCESS;
}
static inline int MPI_Allgather(const void * sendbuf, int sendcount, MPI_Datatype sendtype, void * recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm)
{ (void) recvcount; (void) recvtype; (void) comm; return shim_mpi_copy(sendbuf, recvbuf, sendcount, sendtype); }
static inline int MPI_Allgatherv(const void * sendbuf, int sendcount, MPI_Datatype sendtype, void * recvbuf, const int * recvcounts, const int * displs, MPI_Datatype recvtype, MPI_Comm comm)
{
    (void) recvcounts; (void) comm;
    if(sendbuf == MPI_IN_PLACE) return MPI_SUCCESS;
    memmove((char *) recvbuf + (size_t) displs[0] * recvtype, sendbuf, (size_t) sendcount * sendtype);
    return MPI_SUCCESS;
}
static inline int MPI_Gather(const void * sendbuf, int sendcount, MPI_Datatype sendtype, void * recvbuf, int recvcount, MPI_Datatype recvtype, int root, MPI_Comm comm)
{ (void) recvcount; (void) recvtype; (void) root; (void) comm; return shim_mpi_copy(sendbuf, recvbuf, sendcount, sendtype); }

static inline int MPI_Gatherv(const void * sendbuf, int sendcount, MPI_Datatype sendtype, void * recvbuf, const int * recvcounts, const int * displs, MPI_Datatype recvtype, int root, MPI_Comm comm)
{
    (void) recvcounts; (void) root; (void) comm;
    if(sendbuf == MPI_IN_PLACE) return MPI_SUCCESS;
    memmove((char *) recvbuf + (size_t) displs[0] * recvtype, sendbuf, (size_t) sendcount * sendtype);
    return MPI_SUCCESS;
}
static inline int MPI_Scatterv(const void * sendbuf, const int * sendcounts, const int * displs, MPI_Datatype sendtype, void * recvbuf, int recvcount, MPI_Datatype recvtype, int root, MPI_Comm comm)
{
    (void) recvcount; (void) recvtype; (void) root; (void) comm;
    if(recvbuf == MPI_IN_PLACE) return MPI_SUCCESS;
    memmove(recvbuf, (const char *) sendbuf + (size_t) displs[0] * sendtype, (size_t) sendcounts[0] * sendtype);
    return MPI_SUCCESS;
}
/* the only partner of rank 0 is rank 0 */
static inline int MPI_Sendrecv(const void * sendbuf, int sendcount, MPI_Datatype sendtype, int dest, int sendtag, void * recvbuf, int recvcount, MPI_Datatype recvtype,
        int source, int recvtag, MPI_Comm comm, MPI_Status * st)
{
    (void) sendtag; (void) recvcount; (void) recvtype; (void) recvtag; (void) comm; (void) st;
    if(dest != 0 || source != 0) shim_mpi_unreachable("MPI_Sendrecv with another rank");
    return shim_mpi_copy(sendbuf, recvbuf, sendcount, sendtype);
}
static inline int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm * newcomm) { (void) color; (void) key; *newcomm = comm; return MPI_SUCCESS; }

/* Point-to-point with oneself (rank 0 -> rank 0), as MPI_Alltoallv_sparse (utils/system.cpp:264-325) does even on one
 * rank: a receive is posted and the matching send (same tag, same translation unit) copies into it, or the other way
 * round through a buffered copy.  Any other partner aborts. */
typedef struct { void * buf; size_t bytes; int tag; int live; } shim_mpi_post;
static shim_mpi_post shim_mpi_recvs[64], shim_mpi_sends[64];
static inline int MPI_Irecv(void * buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Request * req)
{
    (void) comm;
    if(src != 0 && src != MPI_ANY_SOURCE) shim_mpi_unreachable("MPI_Irecv from another rank");
    const size_t bytes = (size_t) count * (size_t) type;
    if(req) *req = 0;
    for(int i = 0; i < 64; i++)
        if(shim_mpi_sends[i].live && (tag == MPI_ANY_TAG || shim_mpi_sends[i].tag == tag)) {
            memcpy(buf, shim_mpi_sends[i].buf, bytes < shim_mpi_sends[i].bytes ? bytes : shim_mpi_sends[i].bytes);
            free(shim_mpi_sends[i].buf);
            shim_mpi_sends[i].live = 0;
            return MPI_SUCCESS;
        }
    for(int i = 0; i < 64; i++)
        if(!shim_mpi_recvs[i].live) {
            shim_mpi_recvs[i].buf = buf; shim_mpi_recvs[i].bytes = bytes; shim_mpi_recvs[i].tag = tag; shim_mpi_recvs[i].live = 1;
            return MPI_SUCCESS;
        }
    shim_mpi_unreachable("MPI_Irecv: more than 64 receives posted");
    return 0;
}
static inline int MPI_Isend(const void * buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request * req)
{
    (void) comm;
    if(dest != 0) shim_mpi_unreachable("MPI_Isend to another rank");
    const size_t bytes = (size_t) count * (size_t) type;
    if(req) *req = 0;
    for(int i = 0; i < 64; i++)
        if(shim_mpi_recvs[i].live && (shim_mpi_recvs[i].tag == MPI_ANY_TAG || shim_mpi_recvs[i].tag == tag)) {
            memcpy(shim_mpi_recvs[i].buf, buf, bytes < shim_mpi_recvs[i].bytes ? bytes : shim_mpi_recvs[i].bytes);
            shim_mpi_recvs[i].live = 0;
            return MPI_SUCCESS;
        }
    for(int i = 0; i < 64; i++)
        if(!shim_mpi_sends[i].live) {
            shim_mpi_sends[i].buf = malloc(bytes ? bytes : 1);
            memcpy(shim_mpi_sends[i].buf, buf, bytes);
            shim_mpi_sends[i].bytes = bytes; shim_mpi_sends[i].tag = tag; shim_mpi_sends[i].live = 1;
            return MPI_SUCCESS;
        }
    shim_mpi_unreachable("MPI_Isend: more than 64 sends pending");
    return 0;
}
static inline int MPI_Send(const void * buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm)
{ (void) buf; (void) count; (void) type; (void) dest; (void) tag; (void) comm; shim_mpi_unreachable("MPI_Send"); return 0; }
static inline int MPI_Recv(void * buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Status * st)
{ (void) buf; (void) count; (void) type; (void) src; (void) tag; (void) comm; (void) st; shim_mpi_unreachable("MPI_Recv"); return 0; }
static inline int MPI_Waitall(int count, MPI_Request * reqs, MPI_Status * sts)
{ (void) count; (void) reqs; (void) sts; return MPI_SUCCESS; }
static inline int MPI_Waitsome(int incount, MPI_Request * reqs, int * outcount, int * indices, MPI_Status * sts)
{ (void) incount; (void) reqs; (void) indices; (void) sts; *outcount = MPI_UNDEFINED; return MPI_SUCCESS; }
static inline int MPI_Wait(MPI_Request * req, MPI_Status * st) { (void) req; (void) st; return MPI_SUCCESS; }

#ifdef __cplusplus
}
#endif
#endif
