// Assembled-CSR check path on the GPU (the reference's --mat_comp).
//
// Replaces, for the comparison run only, DOLFINx's CPU assembly of the bilinear form into a
// MatrixCSR (src/laplacian_solver.cpp:160-191: create_sparsity_pattern, assemble_matrix with the BCs,
// set_diagonal(1.0)) and the device SpMV of src/csr.hpp:28-45,174-217.  The element matrices
// A_e = B^T (kappa G) B are formed densely from the same tables and geometry tensor as the
// matrix-free operator but WITHOUT sum factorisation, scattered as (row, col, value) triplets,
// sorted and merged into CSR with thrust (library code, setup only), and applied with a
// warp-per-row SpMV.  Rows and columns of constrained dofs are dropped and their diagonal set to 1,
// exactly as the reference; only owned rows are assembled (ghost cells complete them).
// This is a correctness check, not a performance path ("slow with large ndofs", src/main.cpp:172-173).
#include <thrust/device_ptr.h>
#include <thrust/execution_policy.h>
#include <thrust/reduce.h>
#include <thrust/sort.h>

#include <vector>

#include "operator.cuh"
#include "stiffness_launch.cuh"

struct bd_csr
{
  int scalar_bits = 64;
  std::int64_t nrows = 0, ncols = 0, nnz = 0;
  bd::DeviceArray<std::int64_t> row_ptr;
  bd::DeviceArray<std::int32_t> cols;
  bd::DeviceArray<char> vals;
};

namespace bd
{
// One CTA per cell: dense element matrix, entries with a constrained row/column or a ghost row get
// key = -1 (dropped later).  Tables: phi [nq][nd], dq = dphi1*phi0 [nq][nd].
template <typename T>
__global__ void __launch_bounds__(256)
    element_matrix_kernel(int nd, int nq, std::int64_t ncells, const T* __restrict__ Gref,
                          const T* __restrict__ constants, const std::int32_t* __restrict__ dofmap,
                          const std::int32_t* __restrict__ cells, const std::int8_t* __restrict__ bc,
                          const double* __restrict__ phi, const double* __restrict__ dq,
                          std::int64_t n_owned, std::int64_t ncols, std::int64_t* __restrict__ keys,
                          T* __restrict__ vals)
{
  const int nd3 = nd * nd * nd, nq3 = nq * nq * nq;
  const std::int64_t per_cell = static_cast<std::int64_t>(nd3) * nd3;
  for (std::int64_t c = blockIdx.x; c < ncells; c += gridDim.x)
  {
    const T* G = Gref + c * 6 * nq3;
    const double kappa = static_cast<double>(constants[c]);
    const std::int32_t* dofs = dofmap + static_cast<std::int64_t>(cells[c]) * nd3;
    for (std::int64_t e = threadIdx.x; e < per_cell; e += blockDim.x)
    {
      const int i = static_cast<int>(e / nd3), j = static_cast<int>(e % nd3);
      const int i0 = i / (nd * nd), i1 = (i / nd) % nd, i2 = i % nd;
      const int j0 = j / (nd * nd), j1 = (j / nd) % nd, j2 = j % nd;
      double s = 0;
      for (int q0 = 0; q0 < nq; ++q0)
        for (int q1 = 0; q1 < nq; ++q1)
          for (int q2 = 0; q2 < nq; ++q2)
          {
            const int q = (q0 * nq + q1) * nq + q2;
            const double pi0 = phi[q0 * nd + i0], pi1 = phi[q1 * nd + i1], pi2 = phi[q2 * nd + i2];
            const double pj0 = phi[q0 * nd + j0], pj1 = phi[q1 * nd + j1], pj2 = phi[q2 * nd + j2];
            const double bi[3] = {dq[q0 * nd + i0] * pi1 * pi2, pi0 * dq[q1 * nd + i1] * pi2,
                                  pi0 * pi1 * dq[q2 * nd + i2]};
            const double bj[3] = {dq[q0 * nd + j0] * pj1 * pj2, pj0 * dq[q1 * nd + j1] * pj2,
                                  pj0 * pj1 * dq[q2 * nd + j2]};
            const double G0 = G[q], G1 = G[nq3 + q], G2 = G[2 * nq3 + q], G3 = G[3 * nq3 + q],
                         G4 = G[4 * nq3 + q], G5 = G[5 * nq3 + q];
            s += bi[0] * (G0 * bj[0] + G1 * bj[1] + G2 * bj[2]) + bi[1] * (G1 * bj[0] + G3 * bj[1] + G4 * bj[2])
                 + bi[2] * (G2 * bj[0] + G4 * bj[1] + G5 * bj[2]);
          }
      const std::int32_t row = dofs[i], col = dofs[j];
      const bool keep = row < n_owned && !bc[row] && !bc[col];
      keys[c * per_cell + e] = keep ? static_cast<std::int64_t>(row) * ncols + col : -1;
      vals[c * per_cell + e] = keep ? static_cast<T>(kappa * s) : T(0);
    }
  }
}

// unit diagonal on constrained owned rows (set_diagonal, src/laplacian_solver.cpp:184)
template <typename T>
__global__ void bc_diagonal_kernel(std::int64_t n_owned, std::int64_t ncols, const std::int8_t* __restrict__ bc,
                                   std::int64_t* __restrict__ keys, T* __restrict__ vals)
{
  const std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n_owned)
  {
    keys[i] = bc[i] ? i * ncols + i : -1;
    vals[i] = bc[i] ? T(1) : T(0);
  }
}

__global__ void split_keys_kernel(std::int64_t nnz, std::int64_t ncols, const std::int64_t* __restrict__ keys,
                                  std::int32_t* __restrict__ cols, std::int64_t* __restrict__ row_count)
{
  const std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < nnz)
  {
    const std::int64_t row = keys[i] / ncols;
    cols[i] = static_cast<std::int32_t>(keys[i] - row * ncols);
    atomicAdd(reinterpret_cast<unsigned long long*>(row_count + row + 1), 1ull);
  }
}

// y[row] = sum_j A[row, j] x[j] for owned rows, one warp per row (src/csr.hpp:28-45)
template <typename T>
__global__ void __launch_bounds__(256)
    spmv_kernel(std::int64_t nrows, const std::int64_t* __restrict__ row_ptr,
                const std::int32_t* __restrict__ cols, const T* __restrict__ vals, const T* __restrict__ x,
                T* __restrict__ y)
{
  const std::int64_t warp = (static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const std::int64_t nwarps = (static_cast<std::int64_t>(gridDim.x) * blockDim.x) >> 5;
  for (std::int64_t row = warp; row < nrows; row += nwarps)
  {
    double s = 0;
    for (std::int64_t k = row_ptr[row] + lane; k < row_ptr[row + 1]; k += 32)
      s += static_cast<double>(vals[k]) * static_cast<double>(x[cols[k]]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0)
      y[row] = static_cast<T>(s);
  }
}

template <typename T>
int csr_assemble_impl(bd_laplacian* op, std::int64_t n_owned, bd_csr* A)
{
  const int nd = op->degree + 1, nq = op->nq, nd3 = nd * nd * nd;
  const std::int64_t nlist = op->n_lcells + op->n_bcells;
  const std::int64_t per_cell = static_cast<std::int64_t>(nd3) * nd3;
  const std::int64_t ntrip = nlist * per_cell + n_owned;
  if (ntrip > (1ll << 31))
    return fail(BD_ERR_INVALID_ARGUMENT, "Too many matrix entries for the check path (%lld triplets)",
                (long long)ntrip);
  // tables in double: phi, dq = dphi1 * phi0
  std::vector<double> phi(nq * nd), dq(nq * nd);
  const T* p0 = reinterpret_cast<const T*>(op->phi0.data());
  const T* d1 = reinterpret_cast<const T*>(op->dphi1.data());
  for (int q = 0; q < nq; ++q)
    for (int i = 0; i < nd; ++i)
    {
      phi[q * nd + i] = p0[q * nd + i];
      double s = 0;
      for (int r = 0; r < nq; ++r)
        s += static_cast<double>(d1[q * nq + r]) * static_cast<double>(p0[r * nd + i]);
      dq[q * nd + i] = s;
    }
  DeviceArray<double> phi_d, dq_d;
  BD_CHECK(phi_d.assign(phi.data(), phi.size(), false));
  BD_CHECK(dq_d.assign(dq.data(), dq.size(), false));
  const void* Gref = nullptr;
  std::int64_t gn = 0;
  BD_CHECK(bd_laplacian_geometry(op, &Gref, &gn));
  DeviceArray<std::int64_t> keys;
  DeviceArray<T> vals;
  BD_CHECK(keys.alloc(static_cast<size_t>(ntrip)));
  BD_CHECK(vals.alloc(static_cast<size_t>(ntrip)));
  if (nlist > 0)
  {
    const int grid = static_cast<int>(nlist < 65535 ? nlist : 65535);
    element_matrix_kernel<T><<<grid, 256>>>(nd, nq, nlist, static_cast<const T*>(Gref),
                                            reinterpret_cast<const T*>(op->cell_constants.ptr),
                                            op->cell_dofmap.ptr, op->cell_list.ptr, op->bc_marker.ptr,
                                            phi_d.ptr, dq_d.ptr, n_owned, A->ncols, keys.ptr, vals.ptr);
    BD_CUDA_CHECK(cudaGetLastError());
  }
  if (n_owned > 0)
  {
    bc_diagonal_kernel<T><<<static_cast<int>((n_owned + 255) / 256), 256>>>(
        n_owned, A->ncols, op->bc_marker.ptr, keys.ptr + nlist * per_cell, vals.ptr + nlist * per_cell);
    BD_CUDA_CHECK(cudaGetLastError());
  }
  // sort by (row, col), merge duplicates
  thrust::device_ptr<std::int64_t> k(keys.ptr);
  thrust::device_ptr<T> v(vals.ptr);
  thrust::stable_sort_by_key(thrust::device, k, k + ntrip, v);
  DeviceArray<std::int64_t> ukeys;
  DeviceArray<T> uvals;
  BD_CHECK(ukeys.alloc(static_cast<size_t>(ntrip)));
  BD_CHECK(uvals.alloc(static_cast<size_t>(ntrip)));
  auto ends = thrust::reduce_by_key(thrust::device, k, k + ntrip, v, thrust::device_ptr<std::int64_t>(ukeys.ptr),
                                    thrust::device_ptr<T>(uvals.ptr));
  std::int64_t nuniq = ends.first - thrust::device_ptr<std::int64_t>(ukeys.ptr);
  // dropped entries share key -1 and sort first
  std::int64_t first_key = 0;
  BD_CUDA_CHECK(cudaMemcpy(&first_key, ukeys.ptr, sizeof(first_key), cudaMemcpyDeviceToHost));
  const std::int64_t skip = (nuniq > 0 && first_key < 0) ? 1 : 0;
  A->nnz = nuniq - skip;
  BD_CHECK(A->row_ptr.alloc(static_cast<size_t>(A->nrows + 1)));
  BD_CUDA_CHECK(cudaMemset(A->row_ptr.ptr, 0, sizeof(std::int64_t) * (A->nrows + 1)));
  BD_CHECK(A->cols.alloc(static_cast<size_t>(A->nnz > 0 ? A->nnz : 1)));
  BD_CHECK(A->vals.alloc(static_cast<size_t>(A->nnz > 0 ? A->nnz : 1) * sizeof(T)));
  if (A->nnz > 0)
  {
    split_keys_kernel<<<static_cast<int>((A->nnz + 255) / 256), 256>>>(A->nnz, A->ncols, ukeys.ptr + skip,
                                                                      A->cols.ptr, A->row_ptr.ptr);
    BD_CUDA_CHECK(cudaGetLastError());
    BD_CUDA_CHECK(cudaMemcpy(A->vals.ptr, uvals.ptr + skip, sizeof(T) * A->nnz, cudaMemcpyDeviceToDevice));
    thrust::device_ptr<std::int64_t> rp(A->row_ptr.ptr);
    thrust::inclusive_scan(thrust::device, rp, rp + A->nrows + 1, rp);
  }
  BD_CUDA_CHECK(cudaDeviceSynchronize());
  return BD_SUCCESS;
}
} // namespace bd

using namespace bd;

extern "C" int bd_csr_assemble(bd_laplacian* op, int64_t n_owned, bd_csr** out)
{
  BD_REQUIRE(op && out, "bd_csr_assemble: null argument");
  *out = nullptr;
  BD_REQUIRE(n_owned >= 0 && n_owned <= op->ndofs, "bd_csr_assemble: bad n_owned");
  auto A = new bd_csr;
  A->scalar_bits = op->scalar_bits;
  A->nrows = n_owned;
  A->ncols = op->ndofs;
  int rc;
  try
  {
    rc = op->scalar_bits == 64 ? csr_assemble_impl<double>(op, n_owned, A)
                               : csr_assemble_impl<float>(op, n_owned, A);
  }
  catch (const std::exception& e)
  {
    rc = fail(BD_ERR_CUDA, "bd_csr_assemble: %s", e.what());
  }
  if (rc != BD_SUCCESS)
  {
    delete A;
    return rc;
  }
  *out = A;
  return BD_SUCCESS;
}

extern "C" int bd_csr_destroy(bd_csr* A)
{
  delete A;
  return BD_SUCCESS;
}

extern "C" int bd_csr_info(bd_csr* A, int64_t* nrows, int64_t* ncols, int64_t* nnz)
{
  BD_REQUIRE(A, "bd_csr_info: null argument");
  if (nrows)
    *nrows = A->nrows;
  if (ncols)
    *ncols = A->ncols;
  if (nnz)
    *nnz = A->nnz;
  return BD_SUCCESS;
}

extern "C" int bd_csr_apply(bd_csr* A, const void* in_dev, void* out_dev, void* stream_)
{
  BD_REQUIRE(A && in_dev && out_dev, "bd_csr_apply: null argument");
  cudaStream_t stream = as_stream(stream_);
  const size_t ts = static_cast<size_t>(A->scalar_bits / 8);
  // set_value(y, 0) then y = A x over owned rows (src/csr.hpp:174-217)
  BD_CUDA_CHECK(cudaMemsetAsync(out_dev, 0, static_cast<size_t>(A->ncols) * ts, stream));
  if (A->nrows == 0)
    return BD_SUCCESS;
  const std::int64_t want = (A->nrows * 32 + 255) / 256;
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 16;
  const int grid = static_cast<int>(want < cap ? want : cap);
  if (A->scalar_bits == 64)
    spmv_kernel<double><<<grid, 256, 0, stream>>>(A->nrows, A->row_ptr.ptr, A->cols.ptr,
                                                  reinterpret_cast<const double*>(A->vals.ptr),
                                                  static_cast<const double*>(in_dev),
                                                  static_cast<double*>(out_dev));
  else
    spmv_kernel<float><<<grid, 256, 0, stream>>>(A->nrows, A->row_ptr.ptr, A->cols.ptr,
                                                 reinterpret_cast<const float*>(A->vals.ptr),
                                                 static_cast<const float*>(in_dev), static_cast<float*>(out_dev));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}
