/* TEST INFRASTRUCTURE ONLY. Minimal stand-in for <mpi.h> so the reference's
 * MultiGPU headers and Tools.c parse without an MPI installation (none exists
 * in this image). No communication happens: the replay harness emulates P
 * ranks inside one process and moves halo buffers with memcpy. */
#ifndef ORACLE_STUB_MPI_H
#define ORACLE_STUB_MPI_H
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Request;
typedef int MPI_Errhandler;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_SUCCESS 0
#define MPI_COMM_WORLD 0
#define MPI_FLOAT 1
#define MPI_DOUBLE 2
#define MPI_REAL 1
#define MPI_ERRORS_RETURN 0
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
static inline int MPI_Init(int*, char***) { return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return MPI_SUCCESS; }
static inline int MPI_Errhandler_set(MPI_Comm, MPI_Errhandler) { return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
static inline double MPI_Wtime(void) { return 0.0; }
#endif
