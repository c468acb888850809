// Threads-as-ranks stand-in for the handful of MPI calls dolfinx_adapter.hpp makes (test infrastructure).
// A communicator is a pointer to a shared context; the calling thread's rank is thread-local.
#pragma once
#include <barrier>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

struct bd_mock_mpi_ctx
{
  int size;
  std::barrier<> bar;
  std::vector<const void*> ptr_a, ptr_b, ptr_c; // per-rank published pointers
  explicit bd_mock_mpi_ctx(int n) : size(n), bar(n), ptr_a(n), ptr_b(n), ptr_c(n) {}
};
typedef bd_mock_mpi_ctx* MPI_Comm;
typedef int MPI_Datatype;
#define MPI_BYTE 1
#define MPI_INT 4
#define MPI_INT64_T 8
inline thread_local int bd_mock_mpi_rank = 0;

inline int MPI_Comm_rank(MPI_Comm, int* r)
{
  *r = bd_mock_mpi_rank;
  return 0;
}
inline int MPI_Comm_size(MPI_Comm c, int* s)
{
  *s = c->size;
  return 0;
}
inline int MPI_Bcast(void* buf, int count, MPI_Datatype dt, int root, MPI_Comm c)
{
  const int me = bd_mock_mpi_rank;
  c->ptr_a[me] = buf;
  c->bar.arrive_and_wait();
  if (me != root)
    std::memcpy(buf, c->ptr_a[root], static_cast<size_t>(count) * dt);
  c->bar.arrive_and_wait();
  return 0;
}
inline int MPI_Alltoallv(const void* sbuf, const int* scount, const int* sdispl, MPI_Datatype dt, void* rbuf,
                         const int* rcount, const int* rdispl, MPI_Datatype, MPI_Comm c)
{
  const int me = bd_mock_mpi_rank;
  c->ptr_a[me] = sbuf;
  c->ptr_b[me] = scount;
  c->ptr_c[me] = sdispl;
  c->bar.arrive_and_wait();
  for (int src = 0; src < c->size; ++src)
  {
    const int n = static_cast<const int*>(c->ptr_b[src])[me];
    if (n != rcount[src])
      return 1; // counts disagree
    const char* from = static_cast<const char*>(c->ptr_a[src]) + static_cast<size_t>(static_cast<const int*>(c->ptr_c[src])[me]) * dt;
    std::memcpy(static_cast<char*>(rbuf) + static_cast<size_t>(rdispl[src]) * dt, from, static_cast<size_t>(n) * dt);
  }
  c->bar.arrive_and_wait();
  return 0;
}
inline int MPI_Alltoall(const void* sbuf, int scount, MPI_Datatype dt, void* rbuf, int, MPI_Datatype, MPI_Comm c)
{
  const int me = bd_mock_mpi_rank;
  c->ptr_a[me] = sbuf;
  c->bar.arrive_and_wait();
  for (int src = 0; src < c->size; ++src)
    std::memcpy(static_cast<char*>(rbuf) + static_cast<size_t>(src) * scount * dt,
                static_cast<const char*>(c->ptr_a[src]) + static_cast<size_t>(me) * scount * dt,
                static_cast<size_t>(scount) * dt);
  c->bar.arrive_and_wait();
  return 0;
}
