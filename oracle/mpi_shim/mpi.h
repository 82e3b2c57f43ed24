/*
 * oracle/mpi_shim/mpi.h -- minimal single-host stand-in for <mpi.h>.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/grow_cpu.c header).  No MPI is installed
 * in this image; this shim implements exactly the 15 calls the reference's
 * 2-mpi-region-growing/region.c makes (region.c:86-102, 106-143, 250-353,
 * 391-440, 537-548, 599) so that the UNMODIFIED file can be compiled and run
 * as the 2-D oracle:  SHIM_NP=4 ./region pic1.bmp   ==   mpirun -np 4 region pic1.bmp
 * Ranks are fork()ed inside MPI_Init and talk through shared-memory byte rings.
 */
#ifndef RG_MPI_SHIM_H
#define RG_MPI_SHIM_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD     0
#define MPI_UNSIGNED_CHAR  1
#define MPI_INT            2
#define MPI_MIN            1
#define MPI_PROC_NULL      (-2)
#define MPI_SUCCESS        0

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Dims_create(int nnodes, int ndims, int *dims);
int MPI_Cart_create(MPI_Comm old, int ndims, int *dims, int *periods, int reorder, MPI_Comm *cart);
int MPI_Cart_coords(MPI_Comm comm, int rank, int maxdims, int *coords);
int MPI_Cart_shift(MPI_Comm comm, int direction, int disp, int *src, int *dst);
int MPI_Cart_rank(MPI_Comm comm, int *coords, int *rank);
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Send(void *buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype type, int src, int tag, MPI_Comm comm, MPI_Status *st);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(void *sendbuf, void *recvbuf, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);

#endif
