/* mpi.h — minimal stand-in for the ten MPI calls the reference uses, so that the
 * UNMODIFIED reference sources build and run on an image without any MPI installation.
 * Transport: device buffers by CUDA IPC rendezvous (one device-to-device copy, as a CUDA-aware MPI
 * would do); small host buffers as files in $SHIM_DIR (/dev/shm). Baseline plumbing only — nothing
 * in b200knn uses it. */
#ifndef B200KNN_MPI_SHIM_H
#define B200KNN_MPI_SHIM_H
#ifdef __cplusplus
extern "C" {
#endif
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int unused; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_BYTE 1
#define MPI_INT 4
#define MPI_FLOAT 5
#define MPI_MAX 1
#define MPI_SUCCESS 0
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Isend(const void* buf, int count, MPI_Datatype type, int dest, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Irecv(void* buf, int count, MPI_Datatype type, int source, int tag, MPI_Comm comm, MPI_Request* req);
int MPI_Waitall(int count, MPI_Request* reqs, MPI_Status* statuses);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allreduce(const void* send, void* recv, int count, MPI_Datatype type, MPI_Op op, MPI_Comm comm);
int MPI_Allgather(const void* send, int scount, MPI_Datatype stype, void* recv, int rcount, MPI_Datatype rtype, MPI_Comm comm);
#ifdef __cplusplus
}
#endif
#endif
