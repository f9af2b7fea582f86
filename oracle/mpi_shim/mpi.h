/*
 * oracle/mpi_shim/mpi.h -- TEST/BENCH INFRASTRUCTURE ONLY.
 *
 * A single-node, shared-memory stand-in for the MPI subset the reference's distributed
 * programs use (matFact-mpi.c, matFact-hybrid.c, mpiutil.c: 27 entry points), so that they can
 * be compiled UNMODIFIED from /root/reference and timed on the host cores of a box that has no
 * MPI installation.  Ranks are processes started by oracle/_build/mpirun_shim; they share one
 * anonymous memory file (rings for point-to-point messages, per-communicator staging slots
 * for collectives).  It is not a general MPI: messages between two ranks match in the order
 * they were sent (a tag or communicator mismatch aborts), requests complete inside the call
 * that starts them, derived datatypes are contiguous C structs.  Nothing in the product path
 * uses it.
 */
#ifndef MPI_SHIM_MPI_H
#define MPI_SHIM_MPI_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef ptrdiff_t MPI_Aint;

typedef struct MPI_Status {
	int MPI_SOURCE;
	int MPI_TAG;
	int MPI_ERROR;
	int shim_bytes;
} MPI_Status;

typedef void MPI_User_function(void *invec, void *inoutvec, int *len, MPI_Datatype *datatype);

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15

#define MPI_COMM_NULL ((MPI_Comm)-1)
#define MPI_COMM_WORLD ((MPI_Comm)0)

#define MPI_DATATYPE_NULL ((MPI_Datatype)-1)
#define MPI_CHAR ((MPI_Datatype)0)
#define MPI_BYTE ((MPI_Datatype)1)
#define MPI_INT ((MPI_Datatype)2)
#define MPI_LONG ((MPI_Datatype)3)
#define MPI_FLOAT ((MPI_Datatype)4)
#define MPI_DOUBLE ((MPI_Datatype)5)
#define MPI_UNSIGNED ((MPI_Datatype)6)
#define MPI_LONG_LONG ((MPI_Datatype)7)

#define MPI_OP_NULL ((MPI_Op)-1)
#define MPI_SUM ((MPI_Op)0)
#define MPI_MAX ((MPI_Op)1)
#define MPI_MIN ((MPI_Op)2)
#define MPI_PROD ((MPI_Op)3)

#define MPI_REQUEST_NULL ((MPI_Request)-1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)
#define MPI_IN_PLACE ((void *)-1)
#define MPI_ANY_TAG (-1)
#define MPI_UNDEFINED (-32766)

#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_FUNNELED 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3

/* environment (matFact-mpi.c:538-540,647,686; matFact-hybrid.c:648) */
int MPI_Init(int *argc, char ***argv);
int MPI_Init_thread(int *argc, char ***argv, int required, int *provided);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int errorcode);
double MPI_Wtime(void);

/* communicators and the process grid (mpiutil.c:58-116, matFact-mpi.c:575) */
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm);
int MPI_Comm_free(MPI_Comm *comm);
int MPI_Dims_create(int nnodes, int ndims, int dims[]);
int MPI_Cart_create(MPI_Comm comm, int ndims, const int dims[], const int periods[], int reorder,
                    MPI_Comm *newcomm);
int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int coords[]);
int MPI_Cart_rank(MPI_Comm comm, const int coords[], int *rank);
int MPI_Cart_get(MPI_Comm comm, int maxdims, int dims[], int periods[], int coords[]);

/* datatypes and operations (mpiutil.c:8-43, matFact-mpi.c:660) */
int MPI_Type_create_struct(int count, const int blocklengths[], const MPI_Aint displacements[],
                           const MPI_Datatype types[], MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);
int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op);
int MPI_Op_free(MPI_Op *op);

/* point to point (matFact-mpi.c:357-367,423-447,476,503,513,624) */
int MPI_Send(const void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm,
             MPI_Status *status);

/* collectives (matFact-mpi.c:98,135,207-209,558,598,627,638; matFact-hybrid.c:303-305) */
int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Reduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, int root,
               MPI_Comm comm);
int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op,
                  MPI_Comm comm);
int MPI_Iallreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op,
                   MPI_Comm comm, MPI_Request *request);
int MPI_Wait(MPI_Request *request, MPI_Status *status);
int MPI_Waitall(int count, MPI_Request requests[], MPI_Status statuses[]);
int MPI_Gatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf,
                const int recvcounts[], const int displs[], MPI_Datatype recvtype, int root,
                MPI_Comm comm);

#ifdef __cplusplus
}
#endif
#endif
