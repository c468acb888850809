// Conjugate gradients, replacing cg_solve of the reference (src/cg.hpp:89-169).
//
// Same algorithm and update order (unpreconditioned, `max_iter` iterations when rtol == 0).
// What changes is where things live: every scalar (r.r, p.Ap, alpha, beta) stays in device
// memory, the vector passes are fused (10 scalars of traffic per dof and iteration instead of
// the reference's 13 + a fill), and nothing synchronises with the host, so the whole solve is
// one asynchronous stream of kernels (+ one NCCL allreduce of a double per dot product and one
// forward scatter per operator application on more than one rank; the reference scatters twice,
// src/cg.hpp:133-136 and src/laplacian.hpp:286-320).
#include <cmath>

#include "comm.cuh"
#include "operator.cuh"
#include "vector_ops.cuh"

namespace bd
{
// r = b - y ; p = r ; partial(r.r)       (src/cg.hpp:109-112)
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    cg_init_kernel(std::int64_t n, const T* __restrict__ b, const T* __restrict__ y, T* __restrict__ r,
                   T* __restrict__ p, double* __restrict__ partials)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  double s = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
  {
    const T ri = y[i] * T(-1) + b[i]; // axpy(r, -1, y, b)
    r[i] = ri;
    p[i] = ri;                        // axpy(p, 0, r, r)
    s += static_cast<double>(ri) * static_cast<double>(ri);
  }
  s = block_sum(s);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = s;
}

// alpha = rnorm / pAp ; r -= alpha y ; partial(r.r)                    (src/cg.hpp:145,150-154)
// The other half of the reference's update, x += alpha p (:147), only needs the old p: it rides on the pass that
// replaces p (cg_direction_kernel), which saves one read of p per iteration.  Same operations on the same operands.
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    cg_update_kernel(std::int64_t n, const double* __restrict__ rnorm, const double* __restrict__ pAp,
                     const T* __restrict__ y, T* __restrict__ r, double* __restrict__ partials)
{
  const T malpha = -static_cast<T>(*rnorm / *pAp);
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  double s = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
  {
    const T ri = y[i] * malpha + r[i];
    r[i] = ri;
    s += static_cast<double>(ri) * static_cast<double>(ri);
  }
  s = block_sum(s);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = s;
}

// x += alpha p (src/cg.hpp:147) ; beta = rnorm_new / rnorm ; p = beta p + r   (src/cg.hpp:156,166)
// UPDATE_P = false: only the x update (the iteration that leaves the loop on the tolerance, :162-163)
template <typename T, bool UPDATE_P>
__global__ void __launch_bounds__(kVecBlock)
    cg_direction_kernel(std::int64_t n, const double* __restrict__ rnorm_new, const double* __restrict__ rnorm,
                        const double* __restrict__ pAp, const T* __restrict__ r, T* __restrict__ p,
                        T* __restrict__ x)
{
  const T alpha = static_cast<T>(*rnorm / *pAp);
  const T beta = static_cast<T>(*rnorm_new / *rnorm);
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
  {
    const T pi = p[i];
    x[i] = pi * alpha + x[i];
    if constexpr (UPDATE_P)
      p[i] = pi * beta + r[i];
  }
}

template <typename T>
int cg_impl(bd_laplacian* op, bd_halo* halo, bd_comm* comm, T* x, const T* b, int max_iter, double rtol,
            int* iterations, cudaStream_t stream)
{
  const std::int64_t ntot = op->ndofs;
  const std::int64_t n = halo ? halo->n_owned : ntot;
  // each work vector starts on a 256-byte boundary (the doubles behind them need 8)
  const size_t vbytes = (static_cast<size_t>(ntot) * sizeof(T) + 255) / 256 * 256;
  // workspace: r, y, p (the reference allocates them inside cg_solve, src/cg.hpp:103-105),
  // per-CTA partials and the device scalars
  const size_t need = 3 * vbytes + bd_dot_scratch_bytes() + 8 * sizeof(double);
  if (op->cg_work.count < need)
    BD_CHECK(op->cg_work.alloc(need));
  char* base = op->cg_work.ptr;
  T* r = reinterpret_cast<T*>(base);
  T* y = reinterpret_cast<T*>(base + vbytes);
  T* p = reinterpret_cast<T*>(base + 2 * vbytes);
  double* partials = reinterpret_cast<double*>(base + 3 * vbytes);
  double* scal = partials + kMaxPartials; // [0],[1]: r.r ping-pong, [2]: p.Ap, [3]: rnorm0
  BD_CUDA_CHECK(cudaMemsetAsync(base, 0, 3 * vbytes, stream)); // ghost entries start at 0
  const int grid = vec_grid(n);
  void* s = stream;

  // y = A x  (x needs current ghosts: the reference relies on the caller for this first apply)
  BD_CHECK(bd_laplacian_apply_halo(op, halo, x, y, s));
  cg_init_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, b, y, r, p, partials);
  BD_CUDA_CHECK(cudaGetLastError());
  BD_CHECK(launch_reduce_partials(grid, partials, scal + 0, stream));
  BD_CHECK(allreduce_sum_f64(comm, scal + 0, 1, stream));
  BD_CUDA_CHECK(cudaMemcpyAsync(scal + 3, scal + 0, sizeof(double), cudaMemcpyDeviceToDevice, stream));

  const double rtol2 = rtol * rtol;
  int cur = 0, k = 0;
  while (k < max_iter)
  {
    ++k;
    // y = A p, with the forward scatter of p overlapped with the interior cells
    BD_CHECK(bd_laplacian_apply_halo(op, halo, p, y, s));
    // p.y
    BD_CHECK(bd_dot(op->scalar_bits, n, p, y, scal + 2, partials, s));
    BD_CHECK(allreduce_sum_f64(comm, scal + 2, 1, stream));
    cg_update_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, scal + cur, scal + 2, y, r, partials);
    BD_CUDA_CHECK(cudaGetLastError());
    BD_CHECK(launch_reduce_partials(grid, partials, scal + (cur ^ 1), stream));
    BD_CHECK(allreduce_sum_f64(comm, scal + (cur ^ 1), 1, stream));
    if (rtol2 > 0)
    {
      // the only host round trip, and only when a tolerance is requested (src/cg.hpp:162-163)
      double h[4];
      BD_CUDA_CHECK(cudaMemcpyAsync(h, scal, sizeof(h), cudaMemcpyDeviceToHost, stream));
      BD_CUDA_CHECK(cudaStreamSynchronize(stream));
      if (h[cur ^ 1] / h[3] < rtol2)
      {
        cg_direction_kernel<T, false><<<grid, kVecBlock, 0, stream>>>(n, scal + (cur ^ 1), scal + cur, scal + 2, r,
                                                                     p, x);
        BD_CUDA_CHECK(cudaGetLastError());
        break;
      }
    }
    cg_direction_kernel<T, true><<<grid, kVecBlock, 0, stream>>>(n, scal + (cur ^ 1), scal + cur, scal + 2, r, p, x);
    BD_CUDA_CHECK(cudaGetLastError());
    cur ^= 1;
  }
  if (iterations)
    *iterations = k;
  return BD_SUCCESS;
}
} // namespace bd

using namespace bd;

extern "C" int bd_cg_solve(bd_laplacian* op, bd_halo* halo, bd_comm* comm, void* x_dev,
                           const void* b_dev, int max_iter, double rtol, int* iterations, void* stream)
{
  BD_REQUIRE(op && x_dev && b_dev, "bd_cg_solve: null argument");
  BD_REQUIRE(max_iter >= 0 && rtol >= 0, "bd_cg_solve: negative max_iter or rtol");
  BD_REQUIRE(!halo || (halo->n_owned + halo->n_ghost == op->ndofs && halo->scalar_bits == op->scalar_bits),
             "halo plan does not match the operator");
  BD_REQUIRE(!(comm && comm->nranks > 1) || halo, "a multi-rank solve needs a halo plan");
  // A multi-rank halo plan without a communicator would make every rank use its local dot products
  // (ranks then diverge, and with rtol > 0 leave the loop at different iterations): the plan's own
  // communicator is the one the sums must run over.
  if (!comm && halo && halo->comm && halo->comm->nranks > 1)
    comm = halo->comm;
  BD_REQUIRE(!(comm && halo && halo->comm) || comm->nranks == halo->comm->nranks,
             "bd_cg_solve: communicator and halo plan have different rank counts");
  return op->scalar_bits == 64
             ? cg_impl<double>(op, halo, comm, static_cast<double*>(x_dev),
                               static_cast<const double*>(b_dev), max_iter, rtol, iterations,
                               as_stream(stream))
             : cg_impl<float>(op, halo, comm, static_cast<float*>(x_dev),
                              static_cast<const float*>(b_dev), max_iter, rtol, iterations,
                              as_stream(stream));
}

extern "C" int bd_inner_product(bd_comm* comm, int scalar_bits, int64_t n_owned, const void* a,
                                const void* b, double* result_host)
{
  BD_REQUIRE(result_host, "bd_inner_product: null result pointer");
  double* buf = nullptr;
  BD_CUDA_CHECK(cudaMalloc(&buf, bd_dot_scratch_bytes() + sizeof(double)));
  double* res = buf + kMaxPartials;
  int rc = bd_dot(scalar_bits, n_owned, a, b, res, buf, nullptr);
  if (rc == BD_SUCCESS)
    rc = allreduce_sum_f64(comm, res, 1, nullptr);
  if (rc == BD_SUCCESS)
  {
    cudaError_t e = cudaMemcpy(result_host, res, sizeof(double), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess)
      rc = fail(BD_ERR_CUDA, "bd_inner_product: %s", cudaGetErrorString(e));
  }
  cudaFree(buf);
  return rc;
}

extern "C" int bd_norm(bd_comm* comm, int scalar_bits, int64_t n_owned, const void* a,
                       double* result_host)
{
  BD_CHECK(bd_inner_product(comm, scalar_bits, n_owned, a, a, result_host));
  *result_host = std::sqrt(*result_host);
  return BD_SUCCESS;
}
