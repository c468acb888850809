#pragma once
typedef int MPI_Comm;
typedef int MPI_Datatype;
#define MPI_BYTE 1
#define MPI_INT 2
#define MPI_INT64_T 3
int MPI_Comm_rank(MPI_Comm, int*);
int MPI_Comm_size(MPI_Comm, int*);
int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm);
int MPI_Alltoall(const void*, int, MPI_Datatype, void*, int, MPI_Datatype, MPI_Comm);
int MPI_Alltoallv(const void*, const int*, const int*, MPI_Datatype, void*, const int*, const int*, MPI_Datatype,
                  MPI_Comm);
