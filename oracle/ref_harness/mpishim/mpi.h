/* TEST / BASELINE INFRASTRUCTURE ONLY -- a single-node stand-in for the eleven MPI calls the reference's
 * MultiGPU drivers make (MultiGPU/Burgers3d_Baseline/main.c, Tools.c:229-245), so that the UNMODIFIED
 * reference programs can be rebuilt for sm_100 and run at np = 1, 2, 4, 8 on one box without an MPI
 * installation.  `mpirun -np P prog args` becomes `ADV_MPI_NP=P prog args`: MPI_Init
 * forks P-1 children before the program touches CUDA; messages travel through shared-memory mailboxes
 * with one copy in and one copy out, as a shared-memory MPI transport does.  Never part of the product. */
#ifndef ADV_MPI_SHIM_H
#define ADV_MPI_SHIM_H
#ifdef __cplusplus
extern "C" {
#endif

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Errhandler;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
typedef struct adv_mpi_req* MPI_Request;

#define MPI_SUCCESS 0
#define MPI_COMM_WORLD 0
#define MPI_FLOAT 4
#define MPI_DOUBLE 8
#define MPI_ERRORS_RETURN 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm c, int* rank);
int MPI_Comm_size(MPI_Comm c, int* size);
int MPI_Errhandler_set(MPI_Comm c, MPI_Errhandler e);
int MPI_Barrier(MPI_Comm c);
double MPI_Wtime(void);
int MPI_Isend(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request* req);
int MPI_Recv(void* buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status* st);
int MPI_Wait(MPI_Request* req, MPI_Status* st);

#ifdef __cplusplus
}
#endif
#endif
