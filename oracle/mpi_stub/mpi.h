/*
 * Minimal MPI stand-in used ONLY to build the unmodified reference lda-c
 * (oni-ml/oni-lda-c/lda-estimate.c:21 includes "mpi.h") into oracle/_ref/lda.
 * TEST INFRASTRUCTURE - never linked into the product library.
 *
 * No MPI exists in this image.  This header + mpi_stub.c implement exactly the
 * calls lda-estimate.c makes (Init/Finalize, Comm_rank/size, Barrier,
 * Allreduce SUM on MPI_INT/MPI_DOUBLE, blocking Send/Recv of single values)
 * on top of fork() + an anonymous shared mapping, so the reference can still
 * be run on several host cores:  ONI_MPI_STUB_NP=<ranks> ./lda est ...
 * With ONI_MPI_STUB_NP unset the program is a plain single rank.
 */
#ifndef ONI_ORACLE_MPI_STUB_H
#define ONI_ORACLE_MPI_STUB_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;

typedef struct {
    int MPI_SOURCE;
    int MPI_TAG;
    int MPI_ERROR;
} MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_INT 1
#define MPI_DOUBLE 2
#define MPI_SUM 1
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count,
                  MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Send(const void* buf, int count, MPI_Datatype type, int dest, int tag,
             MPI_Comm comm);
int MPI_Recv(void* buf, int count, MPI_Datatype type, int source, int tag,
             MPI_Comm comm, MPI_Status* status);

#endif
