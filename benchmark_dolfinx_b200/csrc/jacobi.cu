// Jacobi preconditioning for the matrix-free operator.
//
// The reference only has the scaffolding: the inverse diagonal of the *assembled* matrix (compute_diag_inv,
// src/csr.hpp:79-107, one thrust::find_if per row; MatrixOperator::get_diag_inverse, :154-162), pointwise_mult
// (src/vector.hpp:266-285) and the remark "p <- beta*p + M^-1(r)" in cg_solve (src/cg.hpp:165).  Here:
//   * the diagonal is computed matrix-free, cell by cell, from the same tables and geometry as the operator:
//       A_ii = sum_cells sum_q kappa [G00 a0^2 + 2 G10 a0 a1 + 2 G20 a0 a2 + G11 a1^2 + 2 G21 a1 a2 + G22 a2^2],
//       a0 = DPhi[q0][i0] Phi[q1][i1] Phi[q2][i2],  a1 = Phi DPhi Phi,  a2 = Phi Phi DPhi,
//     with the unit diagonal of the constrained rows (src/laplacian_solver.cpp:182-184);
//   * the preconditioned CG keeps the structure of bd_cg_solve: device-resident scalars, two fused vector passes
//     per iteration (z = D^-1 r is never stored), one allreduce per dot product.
#include <cmath>

#include "comm.cuh"
#include "operator.cuh"
#include "vector_ops.cuh"

namespace bd
{
struct DiagTables
{
  double phi[9 * 8];  // Phi[q][i]
  double dphi[9 * 8]; // DPhi[q][i] = sum_r dphi1[q][r] phi0[r][i]
};

// One CTA per cell (grid-stride).  Phase 1: G at the cell's nq^3 points into shared memory (same arithmetic as
// geometry_kernel / the reference, src/geometry_gpu.hpp:89-130).  Phase 2: one thread per dof sums over the points.
template <typename T>
__global__ void __launch_bounds__(128)
    diagonal_kernel(const T* __restrict__ xgeom, const std::int32_t* __restrict__ gdofmap,
                    const T* __restrict__ dphi_geom, const T* __restrict__ weights,
                    const T* __restrict__ cell_constants, const std::int32_t* __restrict__ dofmap,
                    const std::int8_t* __restrict__ bc, const std::int32_t* __restrict__ cells, std::int64_t ncells,
                    int nd, int nq, const __grid_constant__ DiagTables tab, double* __restrict__ diag)
{
  extern __shared__ double gsm[]; // [6][nq^3]
  const int nq3 = nq * nq * nq, nd3 = nd * nd * nd;
  for (std::int64_t pos = blockIdx.x; pos < ncells; pos += gridDim.x)
  {
    const std::int32_t cell = cells[pos];
    const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cell) * 8;
    const double kappa = static_cast<double>(cell_constants[pos]);
    __syncthreads(); // previous cell's phase 2 is done with gsm
    for (int q = threadIdx.x; q < nq3; q += blockDim.x)
    {
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int k = 0; k < 8; ++k)
      {
        const T* xk = xgeom + 3 * static_cast<std::int64_t>(gd[k]);
        const double x0 = xk[0], x1 = xk[1], x2 = xk[2];
        for (int j = 0; j < 3; ++j)
        {
          const double d = dphi_geom[(static_cast<std::int64_t>(j) * nq3 + q) * 8 + k];
          J[0][j] += x0 * d;
          J[1][j] += x1 * d;
          J[2][j] += x2 * d;
        }
      }
      const double K00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
      const double K01 = -J[0][1] * J[2][2] + J[0][2] * J[2][1];
      const double K02 = J[0][1] * J[1][2] - J[0][2] * J[1][1];
      const double K10 = -J[1][0] * J[2][2] + J[1][2] * J[2][0];
      const double K11 = J[0][0] * J[2][2] - J[0][2] * J[2][0];
      const double K12 = -J[0][0] * J[1][2] + J[0][2] * J[1][0];
      const double K20 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
      const double K21 = -J[0][0] * J[2][1] + J[0][1] * J[2][0];
      const double K22 = J[0][0] * J[1][1] - J[0][1] * J[1][0];
      const double detJ = J[0][0] * K00 - J[0][1] * K10 + J[0][2] * K20;
      const double s = kappa * static_cast<double>(weights[q]) / detJ;
      gsm[0 * nq3 + q] = (K00 * K00 + K01 * K01 + K02 * K02) * s;
      gsm[1 * nq3 + q] = (K10 * K00 + K11 * K01 + K12 * K02) * s;
      gsm[2 * nq3 + q] = (K20 * K00 + K21 * K01 + K22 * K02) * s;
      gsm[3 * nq3 + q] = (K10 * K10 + K11 * K11 + K12 * K12) * s;
      gsm[4 * nq3 + q] = (K20 * K10 + K21 * K11 + K22 * K12) * s;
      gsm[5 * nq3 + q] = (K20 * K20 + K21 * K21 + K22 * K22) * s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nd3; i += blockDim.x)
    {
      const std::int32_t dof = dofmap[static_cast<std::int64_t>(cell) * nd3 + i];
      if (bc[dof])
        continue; // unit diagonal, set by the finishing kernel
      const int i0 = i / (nd * nd), i1 = (i / nd) % nd, i2 = i % nd;
      double acc = 0;
      for (int q0 = 0; q0 < nq; ++q0)
      {
        const double p0 = tab.phi[q0 * nd + i0], d0 = tab.dphi[q0 * nd + i0];
        for (int q1 = 0; q1 < nq; ++q1)
        {
          const double p1 = tab.phi[q1 * nd + i1], d1 = tab.dphi[q1 * nd + i1];
          const double a0_01 = d0 * p1, a1_01 = p0 * d1, a2_01 = p0 * p1;
          for (int q2 = 0; q2 < nq; ++q2)
          {
            const double p2 = tab.phi[q2 * nd + i2], d2 = tab.dphi[q2 * nd + i2];
            const double a0 = a0_01 * p2, a1 = a1_01 * p2, a2 = a2_01 * d2;
            const int q = (q0 * nq + q1) * nq + q2;
            acc += gsm[q] * a0 * a0 + gsm[3 * nq3 + q] * a1 * a1 + gsm[5 * nq3 + q] * a2 * a2
                   + 2.0 * (gsm[nq3 + q] * a0 * a1 + gsm[2 * nq3 + q] * a0 * a2 + gsm[4 * nq3 + q] * a1 * a2);
          }
        }
      }
      atomicAdd(diag + dof, acc);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    diag_invert_kernel(std::int64_t n, const double* __restrict__ diag, const std::int8_t* __restrict__ bc,
                       T* __restrict__ dinv)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    dinv[i] = bc[i] ? T(1) : static_cast<T>(1.0 / diag[i]);
}

// w = x * y  (pointwise_mult, src/vector.hpp:266-285)
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    pointwise_mult_kernel(std::int64_t n, const T* __restrict__ x, const T* __restrict__ y, T* __restrict__ w)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    w[i] = x[i] * y[i];
}

// r = b - y ; p = D^-1 r ; partial(r . D^-1 r)
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    pcg_init_kernel(std::int64_t n, const T* __restrict__ b, const T* __restrict__ y, const T* __restrict__ dinv,
                    T* __restrict__ r, T* __restrict__ p, double* __restrict__ partials)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  double s = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
  {
    const T ri = y[i] * T(-1) + b[i];
    const T zi = dinv[i] * ri;
    r[i] = ri;
    p[i] = zi;
    s += static_cast<double>(ri) * static_cast<double>(zi);
  }
  s = block_sum(s);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = s;
}

// alpha = rz / pAp ; x += alpha p ; r -= alpha y ; partial(r . D^-1 r)
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    pcg_update_kernel(std::int64_t n, const double* __restrict__ rz, const double* __restrict__ pAp,
                      const T* __restrict__ p, const T* __restrict__ y, const T* __restrict__ dinv,
                      T* __restrict__ x, T* __restrict__ r, double* __restrict__ partials)
{
  const T alpha = static_cast<T>(*rz / *pAp);
  const T malpha = -alpha;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  double s = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
  {
    x[i] = p[i] * alpha + x[i];
    const T ri = y[i] * malpha + r[i];
    r[i] = ri;
    s += static_cast<double>(ri) * static_cast<double>(dinv[i] * ri);
  }
  s = block_sum(s);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = s;
}

// beta = rz_new / rz ; p = beta p + D^-1 r      ("p <- beta*p + M^-1(r)", src/cg.hpp:165)
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    pcg_direction_kernel(std::int64_t n, const double* __restrict__ rz_new, const double* __restrict__ rz,
                         const T* __restrict__ r, const T* __restrict__ dinv, T* __restrict__ p)
{
  const T beta = static_cast<T>(*rz_new / *rz);
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
    p[i] = p[i] * beta + dinv[i] * r[i];
}

template <typename T>
int diagonal_inverse_impl(bd_laplacian* op, T* dinv, cudaStream_t stream)
{
  const int nd = op->degree + 1, nq = op->nq;
  const std::int64_t nlist = op->n_lcells + op->n_bcells;
  double* diag = nullptr;
  BD_CUDA_CHECK(cudaMalloc(&diag, static_cast<size_t>(op->ndofs > 0 ? op->ndofs : 1) * sizeof(double)));
  auto cleanup = [&](int rc) {
    cudaFree(diag);
    return rc;
  };
  if (cudaMemsetAsync(diag, 0, static_cast<size_t>(op->ndofs) * sizeof(double), stream) != cudaSuccess)
    return cleanup(fail(BD_ERR_CUDA, "bd_laplacian_diagonal_inverse: memset failed"));
  DiagTables tab{};
  const T* phi = reinterpret_cast<const T*>(op->phi0.data());   // [nq][nd]
  const T* dphi = reinterpret_cast<const T*>(op->dphi1.data()); // [nq][nq]
  for (int q = 0; q < nq; ++q)
    for (int i = 0; i < nd; ++i)
    {
      tab.phi[q * nd + i] = static_cast<double>(phi[q * nd + i]);
      double s = 0;
      if (op->is_identity)
        s = static_cast<double>(dphi[q * nq + i]);
      else
        for (int r = 0; r < nq; ++r)
          s += static_cast<double>(dphi[q * nq + r]) * static_cast<double>(phi[r * nd + i]);
      tab.dphi[q * nd + i] = s;
    }
  if (nlist > 0)
  {
    const size_t smem = static_cast<size_t>(6) * nq * nq * nq * sizeof(double);
    auto kernel = diagonal_kernel<T>;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))
        != cudaSuccess)
      return cleanup(fail(BD_ERR_CUDA, "bd_laplacian_diagonal_inverse: shared memory attribute"));
    const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
    const int grid = static_cast<int>(nlist < cap ? nlist : cap);
    kernel<<<grid, 128, smem, stream>>>(reinterpret_cast<const T*>(op->xgeom.ptr), op->geometry_dofmap.ptr,
                                        reinterpret_cast<const T*>(op->dphi_geometry.ptr),
                                        reinterpret_cast<const T*>(op->weights.ptr),
                                        reinterpret_cast<const T*>(op->cell_constants.ptr), op->cell_dofmap.ptr,
                                        op->bc_marker.ptr, op->cell_list.ptr, nlist, nd, nq, tab, diag);
    if (cudaGetLastError() != cudaSuccess)
      return cleanup(fail(BD_ERR_CUDA, "bd_laplacian_diagonal_inverse: launch failed"));
  }
  diag_invert_kernel<T><<<vec_grid(op->ndofs), kVecBlock, 0, stream>>>(op->ndofs, diag, op->bc_marker.ptr, dinv);
  if (cudaGetLastError() != cudaSuccess || cudaStreamSynchronize(stream) != cudaSuccess)
    return cleanup(fail(BD_ERR_CUDA, "bd_laplacian_diagonal_inverse: %s", cudaGetErrorString(cudaGetLastError())));
  return cleanup(BD_SUCCESS);
}

template <typename T>
int pcg_impl(bd_laplacian* op, bd_halo* halo, bd_comm* comm, T* x, const T* b, const T* dinv, int max_iter,
             double rtol, int* iterations, cudaStream_t stream)
{
  const std::int64_t ntot = op->ndofs;
  const std::int64_t n = halo ? halo->n_owned : ntot;
  const size_t vbytes = (static_cast<size_t>(ntot) * sizeof(T) + 255) / 256 * 256;
  const size_t need = 3 * vbytes + bd_dot_scratch_bytes() + 8 * sizeof(double);
  if (op->cg_work.count < need)
    BD_CHECK(op->cg_work.alloc(need));
  char* base = op->cg_work.ptr;
  T* r = reinterpret_cast<T*>(base);
  T* y = reinterpret_cast<T*>(base + vbytes);
  T* p = reinterpret_cast<T*>(base + 2 * vbytes);
  double* partials = reinterpret_cast<double*>(base + 3 * vbytes);
  double* scal = partials + kMaxPartials; // [0],[1]: r.z ping-pong, [2]: p.Ap, [3]: (r.z)_0
  BD_CUDA_CHECK(cudaMemsetAsync(base, 0, 3 * vbytes, stream));
  const int grid = vec_grid(n);
  void* s = stream;

  BD_CHECK(bd_laplacian_apply_halo(op, halo, x, y, s));
  pcg_init_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, b, y, dinv, r, p, partials);
  BD_CUDA_CHECK(cudaGetLastError());
  BD_CHECK(launch_reduce_partials(grid, partials, scal + 0, stream));
  BD_CHECK(allreduce_sum_f64(comm, scal + 0, 1, stream));
  BD_CUDA_CHECK(cudaMemcpyAsync(scal + 3, scal + 0, sizeof(double), cudaMemcpyDeviceToDevice, stream));

  const double rtol2 = rtol * rtol;
  int cur = 0, k = 0;
  while (k < max_iter)
  {
    ++k;
    BD_CHECK(bd_laplacian_apply_halo(op, halo, p, y, s));
    BD_CHECK(bd_dot(op->scalar_bits, n, p, y, scal + 2, partials, s));
    BD_CHECK(allreduce_sum_f64(comm, scal + 2, 1, stream));
    pcg_update_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, scal + cur, scal + 2, p, y, dinv, x, r, partials);
    BD_CUDA_CHECK(cudaGetLastError());
    BD_CHECK(launch_reduce_partials(grid, partials, scal + (cur ^ 1), stream));
    BD_CHECK(allreduce_sum_f64(comm, scal + (cur ^ 1), 1, stream));
    if (rtol2 > 0)
    {
      double h[4];
      BD_CUDA_CHECK(cudaMemcpyAsync(h, scal, sizeof(h), cudaMemcpyDeviceToHost, stream));
      BD_CUDA_CHECK(cudaStreamSynchronize(stream));
      if (h[cur ^ 1] / h[3] < rtol2)
        break;
    }
    pcg_direction_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, scal + (cur ^ 1), scal + cur, r, dinv, p);
    BD_CUDA_CHECK(cudaGetLastError());
    cur ^= 1;
  }
  if (iterations)
    *iterations = k;
  return BD_SUCCESS;
}
} // namespace bd

using namespace bd;

extern "C" int bd_laplacian_diagonal_inverse(bd_laplacian* op, void* diag_inv_dev, void* stream)
{
  BD_REQUIRE(op && diag_inv_dev, "bd_laplacian_diagonal_inverse: null argument");
  return op->scalar_bits == 64 ? diagonal_inverse_impl<double>(op, static_cast<double*>(diag_inv_dev), as_stream(stream))
                               : diagonal_inverse_impl<float>(op, static_cast<float*>(diag_inv_dev), as_stream(stream));
}

extern "C" int bd_pointwise_mult(int scalar_bits, int64_t n, const void* x, const void* y, void* w, void* stream)
{
  BD_REQUIRE(scalar_bits == 32 || scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  BD_REQUIRE(n >= 0, "negative length");
  if (n == 0)
    return BD_SUCCESS;
  BD_REQUIRE(x && y && w, "bd_pointwise_mult: null pointer argument");
  if (scalar_bits == 64)
    pointwise_mult_kernel<double><<<vec_grid(n), kVecBlock, 0, as_stream(stream)>>>(
        n, static_cast<const double*>(x), static_cast<const double*>(y), static_cast<double*>(w));
  else
    pointwise_mult_kernel<float><<<vec_grid(n), kVecBlock, 0, as_stream(stream)>>>(
        n, static_cast<const float*>(x), static_cast<const float*>(y), static_cast<float*>(w));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

extern "C" int bd_cg_solve_jacobi(bd_laplacian* op, bd_halo* halo, bd_comm* comm, void* x_dev, const void* b_dev,
                                  const void* diag_inv_dev, int max_iter, double rtol, int* iterations, void* stream)
{
  BD_REQUIRE(op && x_dev && b_dev && diag_inv_dev, "bd_cg_solve_jacobi: null argument");
  BD_REQUIRE(max_iter >= 0 && rtol >= 0, "bd_cg_solve_jacobi: negative max_iter or rtol");
  BD_REQUIRE(!halo || (halo->n_owned + halo->n_ghost == op->ndofs && halo->scalar_bits == op->scalar_bits),
             "halo plan does not match the operator");
  BD_REQUIRE(!(comm && comm->nranks > 1) || halo, "a multi-rank solve needs a halo plan");
  if (!comm && halo && halo->comm && halo->comm->nranks > 1)
    comm = halo->comm;
  return op->scalar_bits == 64
             ? pcg_impl<double>(op, halo, comm, static_cast<double*>(x_dev), static_cast<const double*>(b_dev),
                                static_cast<const double*>(diag_inv_dev), max_iter, rtol, iterations, as_stream(stream))
             : pcg_impl<float>(op, halo, comm, static_cast<float*>(x_dev), static_cast<const float*>(b_dev),
                               static_cast<const float*>(diag_inv_dev), max_iter, rtol, iterations, as_stream(stream));
}
