/* Minimal single-process MPI stand-in for building the UNMODIFIED reference programs
 * (mpi/mpi_convolution.c, open-mp/omp_convolution.c) on a box without MPI.
 * Exactly the API surface those two files use: 22 functions.
 * Ranks are threads of one process; the reference's main() becomes ref_main()
 * (-Dmain=ref_main) and has no global state, so that is safe.
 * This is baseline tooling only — the product never links it. */
#ifndef B2C_MPI_SHIM_H
#define B2C_MPI_SHIM_H
#include <stddef.h>

typedef int MPI_Comm;
typedef int MPI_Info;
typedef struct { int source, tag, error; } MPI_Status;
typedef struct shim_dtype *MPI_Datatype;
typedef struct shim_req *MPI_Request;
typedef struct shim_file *MPI_File;
typedef long long MPI_Offset;

#define MPI_COMM_WORLD 0
#define MPI_INFO_NULL 0
#define MPI_SUCCESS 0
#define MPI_STATUS_IGNORE ((MPI_Status *)0)
#define MPI_SEEK_SET 0
#define MPI_MODE_RDONLY 1
#define MPI_MODE_CREATE 2
#define MPI_MODE_WRONLY 4

extern struct shim_dtype shim_byte, shim_int, shim_double;
#define MPI_BYTE (&shim_byte)
#define MPI_INT (&shim_int)
#define MPI_DOUBLE (&shim_double)

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int code);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Bcast(void *buf, int count, MPI_Datatype t, int root, MPI_Comm comm);
int MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *t);
int MPI_Type_free(MPI_Datatype *t);
int MPI_Isend(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irecv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Wait(MPI_Request *req, MPI_Status *status);
int MPI_Send(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_File_open(MPI_Comm comm, const char *name, int amode, MPI_Info info, MPI_File *fh);
int MPI_File_seek(MPI_File fh, MPI_Offset off, int whence);
int MPI_File_read(MPI_File fh, void *buf, int count, MPI_Datatype t, MPI_Status *status);
int MPI_File_write(MPI_File fh, const void *buf, int count, MPI_Datatype t, MPI_Status *status);
int MPI_File_close(MPI_File *fh);
#endif
