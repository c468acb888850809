// NCCL over NVLink replaces the reference's CUDA-aware MPI:
//   * forward scatter of the input vector (DOLFINx Scatterer driven by scatter_fwd_begin/end with
//     pack/unpack functors: src/vector.hpp:95-149, src/laplacian.hpp:286,320, src/cg.hpp:133-136)
//     -> one pack kernel + grouped ncclSend/ncclRecv; ghosts are ordered by source rank, so each
//     message is received straight into the ghost segment of the vector (no unpack kernel);
//   * MPI_Allreduce of one scalar per dot product (src/vector.hpp:173, src/cg.hpp:76)
//     -> ncclAllReduce on a double that never leaves device memory.
// NCCL is resolved with dlopen at the first use so that single-GPU users need no NCCL at all and
// the library binds to whichever libnccl the host process already loaded (e.g. torch's).
#include "comm.cuh"

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>

#include "operator.cuh"

namespace
{
struct NcclApi
{
  void* handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string error;
};

NcclApi& nccl_api()
{
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names)
    {
      api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.handle)
        break;
    }
    if (!api.handle)
    {
      api.error = std::string("cannot load libnccl: ") + dlerror();
      return;
    }
#define LOAD(field, sym)                                                                           \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, sym));                       \
  if (!api.field)                                                                                  \
  {                                                                                                \
    api.error = std::string("libnccl lacks ") + sym;                                               \
    return;                                                                                        \
  }
    LOAD(GetUniqueId, "ncclGetUniqueId")
    LOAD(CommInitRank, "ncclCommInitRank")
    LOAD(CommDestroy, "ncclCommDestroy")
    LOAD(AllReduce, "ncclAllReduce")
    LOAD(Send, "ncclSend")
    LOAD(Recv, "ncclRecv")
    LOAD(GroupStart, "ncclGroupStart")
    LOAD(GroupEnd, "ncclGroupEnd")
    LOAD(GetErrorString, "ncclGetErrorString")
#undef LOAD
  });
  return api;
}

int nccl_ready()
{
  NcclApi& api = nccl_api();
  if (!api.error.empty())
    return bd::fail(BD_ERR_NCCL, "%s", api.error.c_str());
  return BD_SUCCESS;
}

#define BD_NCCL_CHECK(expr)                                                                        \
  do                                                                                               \
  {                                                                                                \
    ncclResult_t bd_nr_ = (expr);                                                                  \
    if (bd_nr_ != ncclSuccess)                                                                     \
      return bd::fail(BD_ERR_NCCL, "%s:%d: NCCL reports %s for `%s`", __FILE__, __LINE__,          \
                      nccl_api().GetErrorString(bd_nr_), #expr);                                   \
  } while (0)
} // namespace

namespace bd
{
int allreduce_sum_f64(bd_comm* comm, double* buf, size_t count, cudaStream_t stream)
{
  if (!comm || comm->nranks == 1)
    return BD_SUCCESS;
  BD_NCCL_CHECK(nccl_api().AllReduce(buf, buf, count, ncclFloat64, ncclSum,
                                     static_cast<ncclComm_t>(comm->nccl), stream));
  return BD_SUCCESS;
}
} // namespace bd

using namespace bd;

extern "C" int bd_comm_unique_id(void* id_out)
{
  BD_REQUIRE(id_out, "bd_comm_unique_id: null argument");
  static_assert(sizeof(ncclUniqueId) == BD_UNIQUE_ID_BYTES, "ncclUniqueId size");
  BD_CHECK(nccl_ready());
  BD_NCCL_CHECK(nccl_api().GetUniqueId(static_cast<ncclUniqueId*>(id_out)));
  return BD_SUCCESS;
}

extern "C" int bd_comm_create(const void* id, int rank, int nranks, bd_comm** out)
{
  BD_REQUIRE(out, "bd_comm_create: null argument");
  *out = nullptr;
  BD_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
  auto c = new bd_comm;
  c->rank = rank;
  c->nranks = nranks;
  if (cudaGetDevice(&c->device) != cudaSuccess)
  {
    delete c;
    return fail(BD_ERR_CUDA, "cudaGetDevice failed");
  }
  if (nranks > 1)
  {
    if (!id)
    {
      delete c;
      return fail(BD_ERR_INVALID_ARGUMENT, "bd_comm_create: null unique id");
    }
    int rc = nccl_ready();
    if (rc != BD_SUCCESS)
    {
      delete c;
      return rc;
    }
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    ncclComm_t nc = nullptr;
    ncclResult_t r = nccl_api().CommInitRank(&nc, nranks, uid, rank);
    if (r != ncclSuccess)
    {
      delete c;
      return fail(BD_ERR_NCCL, "ncclCommInitRank: %s", nccl_api().GetErrorString(r));
    }
    c->nccl = nc;
  }
  *out = c;
  return BD_SUCCESS;
}

extern "C" int bd_comm_destroy(bd_comm* c)
{
  if (!c)
    return BD_SUCCESS;
  if (c->nccl)
    nccl_api().CommDestroy(static_cast<ncclComm_t>(c->nccl));
  delete c;
  return BD_SUCCESS;
}

extern "C" int bd_halo_create(bd_comm* comm, int scalar_bits, int64_t n_owned, int64_t n_ghost,
                              int n_send, const int32_t* send_ranks, const int64_t* send_counts,
                              const int32_t* send_indices_host, int n_recv,
                              const int32_t* recv_ranks, const int64_t* recv_counts, bd_halo** out)
{
  BD_REQUIRE(out && comm, "bd_halo_create: null argument");
  *out = nullptr;
  BD_REQUIRE(scalar_bits == 32 || scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  BD_REQUIRE(n_owned >= 0 && n_ghost >= 0 && n_send >= 0 && n_recv >= 0, "negative size");
  BD_REQUIRE(n_send == 0 || (send_ranks && send_counts && send_indices_host), "null send plan");
  BD_REQUIRE(n_recv == 0 || (recv_ranks && recv_counts), "null receive plan");
  auto h = new bd_halo;
  auto bail = [&](int rc) {
    bd_halo_destroy(h);
    return rc;
  };
  h->comm = comm;
  h->scalar_bits = scalar_bits;
  h->n_owned = n_owned;
  h->n_ghost = n_ghost;
  std::int64_t tot = 0, rtot = 0;
  for (int i = 0; i < n_send; ++i)
  {
    if (send_ranks[i] < 0 || send_ranks[i] >= comm->nranks || send_counts[i] < 0)
      return bail(fail(BD_ERR_INVALID_ARGUMENT, "bad send plan entry %d", i));
    h->send_ranks.push_back(send_ranks[i]);
    h->send_counts.push_back(send_counts[i]);
    tot += send_counts[i];
  }
  for (int i = 0; i < n_recv; ++i)
  {
    if (recv_ranks[i] < 0 || recv_ranks[i] >= comm->nranks || recv_counts[i] < 0)
      return bail(fail(BD_ERR_INVALID_ARGUMENT, "bad receive plan entry %d", i));
    h->recv_ranks.push_back(recv_ranks[i]);
    h->recv_counts.push_back(recv_counts[i]);
    rtot += recv_counts[i];
  }
  if (rtot != n_ghost)
    return bail(fail(BD_ERR_INVALID_ARGUMENT, "receive counts (%lld) do not add up to n_ghost (%lld)",
                     (long long)rtot, (long long)n_ghost));
  for (std::int64_t i = 0; i < tot; ++i)
    if (send_indices_host[i] < 0 || send_indices_host[i] >= n_owned)
      return bail(fail(BD_ERR_INVALID_ARGUMENT, "send index %lld out of the owned range", (long long)i));
  h->n_send_total = tot;
  int rc;
  if ((rc = h->send_idx.assign(send_indices_host, static_cast<size_t>(tot), false)) != BD_SUCCESS)
    return bail(rc);
  if ((rc = h->send_buf.alloc(static_cast<size_t>(tot > 0 ? tot : 1) * (scalar_bits / 8))) != BD_SUCCESS)
    return bail(rc);
  // Highest priority: the NCCL kernel must get its SM resources before the persistent interior-cell kernel, which
  // is released by the same event and otherwise fills every SM until it ends (the exchange then only starts in that
  // kernel's tail: 1.6 ms exposed per application on 2 GPUs in round 1).
  int prio_lo = 0, prio_hi = 0;
  cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
  if (cudaStreamCreateWithPriority(&h->stream, cudaStreamNonBlocking, prio_hi) != cudaSuccess
      || cudaEventCreateWithFlags(&h->ev_ready, cudaEventDisableTiming) != cudaSuccess
      || cudaEventCreateWithFlags(&h->ev_done, cudaEventDisableTiming) != cudaSuccess)
    return bail(fail(BD_ERR_CUDA, "halo stream/event creation failed"));
  *out = h;
  return BD_SUCCESS;
}

extern "C" int bd_halo_set_unpack(bd_halo* h, const int32_t* ghost_positions_host)
{
  BD_REQUIRE(h, "bd_halo_set_unpack: null plan");
  if (!ghost_positions_host)
  {
    h->use_unpack = false;
    return BD_SUCCESS;
  }
  for (std::int64_t i = 0; i < h->n_ghost; ++i)
    if (ghost_positions_host[i] < 0 || ghost_positions_host[i] >= h->n_ghost)
      return fail(BD_ERR_INVALID_ARGUMENT, "ghost position %lld out of range", (long long)i);
  BD_CHECK(h->unpack_idx.assign(ghost_positions_host, static_cast<size_t>(h->n_ghost), false));
  BD_CHECK(h->recv_buf.alloc(static_cast<size_t>(h->n_ghost > 0 ? h->n_ghost : 1) * (h->scalar_bits / 8)));
  h->use_unpack = true;
  return BD_SUCCESS;
}

extern "C" int bd_halo_destroy(bd_halo* h)
{
  if (!h)
    return BD_SUCCESS;
  for (cudaEvent_t e : h->prof_events)
    cudaEventDestroy(e);
  if (h->ev_ready)
    cudaEventDestroy(h->ev_ready);
  if (h->ev_done)
    cudaEventDestroy(h->ev_done);
  if (h->stream)
    cudaStreamDestroy(h->stream);
  delete h;
  return BD_SUCCESS;
}

namespace
{
// `mark` (optional): events recorded on the side stream right before / after the NCCL group.
int halo_begin_impl(bd_halo* h, void* x_dev, cudaStream_t stream, cudaEvent_t mark0, cudaEvent_t mark1)
{
  const bool multi = h->comm->nranks > 1;
  const size_t ts = static_cast<size_t>(h->scalar_bits / 8);
  // The pack runs on the caller's stream, ahead of the event: the exchange proper (one NCCL kernel) is then
  // ready at the same instant as whatever the caller queues next, and wins the SMs by stream priority.
  if (multi)
    BD_CHECK(bd_pack(h->scalar_bits, h->n_send_total, h->send_idx.ptr, x_dev, h->send_buf.ptr, stream));
  // the exchange starts once everything already queued on `stream` (the producer of x, the pack) is done
  BD_CUDA_CHECK(cudaEventRecord(h->ev_ready, stream));
  BD_CUDA_CHECK(cudaStreamWaitEvent(h->stream, h->ev_ready, 0));
  if (mark0)
    BD_CUDA_CHECK(cudaEventRecord(mark0, h->stream));
  if (multi)
  {
    const ncclDataType_t dt = h->scalar_bits == 64 ? ncclFloat64 : ncclFloat32;
    ncclComm_t nc = static_cast<ncclComm_t>(h->comm->nccl);
    NcclApi& api = nccl_api();
    BD_NCCL_CHECK(api.GroupStart());
    std::int64_t off = 0;
    for (size_t i = 0; i < h->send_ranks.size(); ++i)
    {
      BD_NCCL_CHECK(api.Send(h->send_buf.ptr + off * ts, static_cast<size_t>(h->send_counts[i]), dt,
                             h->send_ranks[i], nc, h->stream));
      off += h->send_counts[i];
    }
    char* ghost_seg = static_cast<char*>(x_dev) + static_cast<size_t>(h->n_owned) * ts;
    char* ghost = h->use_unpack ? h->recv_buf.ptr : ghost_seg;
    off = 0;
    for (size_t i = 0; i < h->recv_ranks.size(); ++i)
    {
      BD_NCCL_CHECK(api.Recv(ghost + off * ts, static_cast<size_t>(h->recv_counts[i]), dt,
                             h->recv_ranks[i], nc, h->stream));
      off += h->recv_counts[i];
    }
    BD_NCCL_CHECK(api.GroupEnd());
    if (h->use_unpack)
      BD_CHECK(bd_unpack(h->scalar_bits, h->n_ghost, h->unpack_idx.ptr, h->recv_buf.ptr, ghost_seg, h->stream));
  }
  if (mark1)
    BD_CUDA_CHECK(cudaEventRecord(mark1, h->stream));
  BD_CUDA_CHECK(cudaEventRecord(h->ev_done, h->stream));
  return BD_SUCCESS;
}
} // namespace

extern "C" int bd_halo_begin(bd_halo* h, void* x_dev, void* stream_)
{
  BD_REQUIRE(h && x_dev, "bd_halo_begin: null argument");
  return halo_begin_impl(h, x_dev, as_stream(stream_), nullptr, nullptr);
}

extern "C" int bd_halo_profile(bd_halo* h, int enable)
{
  BD_REQUIRE(h, "bd_halo_profile: null plan");
  h->profile = enable != 0;
  h->prof_used = 0;
  return BD_SUCCESS;
}

extern "C" int bd_halo_profile_read(bd_halo* h, double* pack_ms, double* exchange_ms, double* exposed_ms,
                                    int64_t* applications)
{
  BD_REQUIRE(h && pack_ms && exchange_ms && exposed_ms && applications, "bd_halo_profile_read: null argument");
  double pk = 0, ex = 0, xp = 0;
  for (size_t i = 0; i + 3 < h->prof_used; i += 4)
  {
    cudaEvent_t* e = h->prof_events.data() + i;
    BD_CUDA_CHECK(cudaEventSynchronize(e[2]));
    BD_CUDA_CHECK(cudaEventSynchronize(e[3]));
    float a = 0, b = 0, c = 0;
    BD_CUDA_CHECK(cudaEventElapsedTime(&a, e[0], e[1])); // main: before pack -> side: exchange may start
    BD_CUDA_CHECK(cudaEventElapsedTime(&b, e[1], e[2])); // side: the NCCL group (+ unpack)
    BD_CUDA_CHECK(cudaEventElapsedTime(&c, e[3], e[2])); // interior kernel done -> ghosts ready (> 0: exposed)
    pk += a;
    ex += b;
    xp += c > 0 ? c : 0;
  }
  *pack_ms = pk;
  *exchange_ms = ex;
  *exposed_ms = xp;
  *applications = static_cast<int64_t>(h->prof_used / 4);
  h->prof_used = 0;
  return BD_SUCCESS;
}

extern "C" int bd_halo_end(bd_halo* h, void* stream)
{
  BD_REQUIRE(h, "bd_halo_end: null argument");
  BD_CUDA_CHECK(cudaStreamWaitEvent(as_stream(stream), h->ev_done, 0));
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_apply_halo(bd_laplacian* op, bd_halo* halo, void* in_dev, void* out_dev,
                                       void* stream)
{
  if (!halo)
    return bd_laplacian_apply(op, in_dev, out_dev, stream);
  BD_REQUIRE(op, "bd_laplacian_apply_halo: null operator");
  BD_REQUIRE(halo->n_owned + halo->n_ghost == op->ndofs && halo->scalar_bits == op->scalar_bits,
             "halo plan does not match the operator");
  // impl_operator (src/laplacian.hpp:281-349): begin scatter, lcells, end scatter, bcells
  if (!halo->profile)
  {
    BD_CHECK(bd_halo_begin(halo, in_dev, stream));
    BD_CHECK(bd_laplacian_apply_part(op, 0, in_dev, out_dev, stream));
    BD_CHECK(bd_halo_end(halo, stream));
    return bd_laplacian_apply_part(op, 1, in_dev, out_dev, stream);
  }
  if (halo->prof_used + 4 > halo->prof_events.size())
    for (int i = 0; i < 4; ++i)
    {
      cudaEvent_t e;
      BD_CUDA_CHECK(cudaEventCreate(&e));
      halo->prof_events.push_back(e);
    }
  cudaEvent_t* e = halo->prof_events.data() + halo->prof_used;
  BD_CUDA_CHECK(cudaEventRecord(e[0], as_stream(stream)));
  BD_CHECK(halo_begin_impl(halo, in_dev, as_stream(stream), e[1], e[2]));
  BD_CHECK(bd_laplacian_apply_part(op, 0, in_dev, out_dev, stream));
  BD_CUDA_CHECK(cudaEventRecord(e[3], as_stream(stream)));
  halo->prof_used += 4;
  BD_CHECK(bd_halo_end(halo, stream));
  return bd_laplacian_apply_part(op, 1, in_dev, out_dev, stream);
}
