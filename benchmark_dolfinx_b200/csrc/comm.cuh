// Inter-GPU plumbing: NCCL (loaded at run time with dlopen) + forward-scatter plan.
#pragma once

#include <vector>

#include "common.cuh"

struct bd_comm
{
  void* nccl = nullptr; // ncclComm_t
  int rank = 0, nranks = 1;
  int device = 0;
};

struct bd_halo
{
  bd_comm* comm = nullptr;
  int scalar_bits = 64;
  std::int64_t n_owned = 0, n_ghost = 0;
  std::vector<int> send_ranks, recv_ranks;
  std::vector<std::int64_t> send_counts, recv_counts;
  std::int64_t n_send_total = 0;
  bd::DeviceArray<std::int32_t> send_idx;
  bd::DeviceArray<char> send_buf;
  // optional unpack (ghosts not ordered by source rank): receive into recv_buf, then
  // x[n_owned + unpack_idx[i]] = recv_buf[i]  (unpack_gpu, reference src/vector.hpp:52-62)
  bd::DeviceArray<std::int32_t> unpack_idx;
  bd::DeviceArray<char> recv_buf;
  bool use_unpack = false;
  cudaStream_t stream = nullptr;      // exchange runs here, overlapped with the lcells kernel
  cudaEvent_t ev_ready = nullptr;     // producer of x done
  cudaEvent_t ev_done = nullptr;      // ghosts of x up to date
  // optional timing of the exchange inside bd_laplacian_apply_halo (bd_halo_profile): per application four
  // events -- main stream before the pack, side stream before / after the NCCL group, main stream after the
  // interior-cell kernel -- from which pack, exchange and exposed (not hidden behind the interior cells) times follow
  bool profile = false;
  std::vector<cudaEvent_t> prof_events;
  size_t prof_used = 0;
};

namespace bd
{
// sum-allreduce of `count` doubles in device memory, in place, stream-ordered
int allreduce_sum_f64(bd_comm* comm, double* buf_dev, size_t count, cudaStream_t stream);
} // namespace bd
