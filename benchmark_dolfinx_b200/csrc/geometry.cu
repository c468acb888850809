// Geometry-factor precompute for sm_100a.
//
// Replaces geometry_computation_gpu<T,Q> (reference src/geometry_gpu.hpp:26-132): for every
// cell of a list and every quadrature point, J = sum_k x_k (x) grad phi_k, K = adj(J),
// G = K K^T w / det J, six unique entries stored as planes of nq^3 per cell
// (order 00,10,20,11,21,22), cell index = position in the list.
//
// Launch shape differs from the reference (one CTA per cell, nq^3 threads, which is 8..27
// threads at low degree): the (cell, point) index space is flattened over 256-thread CTAs so
// every warp is full and the six output planes are written as contiguous runs.  The kernel is
// write-bandwidth bound (48 B or 24 B per point out, the 8 vertices of a cell are shared by
// its nq^3 threads through L1).
#include "common.cuh"
#include "stiffness_mma8.cuh"
#include "stiffness_slab.cuh"

namespace bd
{
// LAYOUT 0: the reference's layout G[c][6][nq^3];  1: the batch-interleaved layout the slab
// stiffness kernel streams (stiffness_slab.cuh);  2: accumulator order for the DMMA kernel
// (stiffness_mma8.cuh, nq = 8).
template <typename T, int LAYOUT>
__global__ void __launch_bounds__(256)
    geometry_kernel(const T* __restrict__ xgeom, T* __restrict__ G,
                    const std::int32_t* __restrict__ geometry_dofmap, const T* __restrict__ dphi,
                    const T* __restrict__ weights, const std::int32_t* __restrict__ cells,
                    std::int64_t ncells, int nq)
{
  const int nq3 = nq * nq * nq;
  const std::int64_t total = ncells * nq3;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
       idx < total; idx += stride)
  {
    const std::int64_t c = idx / nq3;
    const int q = static_cast<int>(idx - c * nq3);
    const std::int32_t* gd = geometry_dofmap + static_cast<std::int64_t>(cells[c]) * 8;
    T J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
    for (int k = 0; k < 8; ++k)
    {
      const T* xk = xgeom + 3 * static_cast<std::int64_t>(__ldg(gd + k));
      const T x0 = __ldg(xk), x1 = __ldg(xk + 1), x2 = __ldg(xk + 2);
#pragma unroll
      for (int j = 0; j < 3; ++j)
      {
        const T d = __ldg(dphi + (static_cast<std::int64_t>(j) * nq3 + q) * 8 + k);
        J[0][j] += x0 * d;
        J[1][j] += x1 * d;
        J[2][j] += x2 * d;
      }
    }
    // K = adj(J) = det(J) J^-1   (reference :100-108)
    const T K00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
    const T K01 = -J[0][1] * J[2][2] + J[0][2] * J[2][1];
    const T K02 = J[0][1] * J[1][2] - J[0][2] * J[1][1];
    const T K10 = -J[1][0] * J[2][2] + J[1][2] * J[2][0];
    const T K11 = J[0][0] * J[2][2] - J[0][2] * J[2][0];
    const T K12 = -J[0][0] * J[1][2] + J[0][2] * J[1][0];
    const T K20 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    const T K21 = -J[0][0] * J[2][1] + J[0][1] * J[2][0];
    const T K22 = J[0][0] * J[1][1] - J[0][1] * J[1][0];
    const T detJ = J[0][0] * K00 - J[0][1] * K10 + J[0][2] * K20;
    // one division per point instead of the reference's six (`* w / detJ` per entry, src/geometry_gpu.hpp:113-130):
    // fp64 divisions were 2/3 of this kernel's instructions; the entries differ from the reference's by <= 1.5 ulp
    const T w = __ldg(weights + q);
    const T s = w / detJ;
    const T g0 = (K00 * K00 + K01 * K01 + K02 * K02) * s;
    const T g1 = (K10 * K00 + K11 * K01 + K12 * K02) * s;
    const T g2 = (K20 * K00 + K21 * K01 + K22 * K02) * s;
    const T g3 = (K10 * K10 + K11 * K11 + K12 * K12) * s;
    const T g4 = (K20 * K10 + K21 * K11 + K22 * K12) * s;
    const T g5 = (K20 * K20 + K21 * K21 + K22 * K22) * s;
    if constexpr (LAYOUT == 2)
    {
      const int q0 = q / (nq * nq), q1 = (q / nq) % nq, q2 = q % nq;
      G[mma8_g_index(nq, c, 0, q0, q1, q2)] = g0;
      G[mma8_g_index(nq, c, 1, q0, q1, q2)] = g1;
      G[mma8_g_index(nq, c, 2, q0, q1, q2)] = g2;
      G[mma8_g_index(nq, c, 3, q0, q1, q2)] = g3;
      G[mma8_g_index(nq, c, 4, q0, q1, q2)] = g4;
      G[mma8_g_index(nq, c, 5, q0, q1, q2)] = g5;
    }
    else if constexpr (LAYOUT == 1)
    {
      const int q0 = q / (nq * nq), q1 = (q / nq) % nq, q2 = q % nq;
      const int row = (32 / nq) * nq;
      T* g = G + slab_g_index(nq, c, 0, q0, q1, q2);
      g[0] = g0;
      g[row] = g1;
      g[2 * row] = g2;
      g[3 * row] = g3;
      g[4 * row] = g4;
      g[5 * row] = g5;
    }
    else
    {
      T* g = G + c * (6 * static_cast<std::int64_t>(nq3)) + q;
      g[0] = g0;
      g[nq3] = g1;
      g[2 * nq3] = g2;
      g[3 * nq3] = g3;
      g[4 * nq3] = g4;
      g[5 * nq3] = g5;
    }
  }
}

// Layouts 1 and 2 interleave cells / components / points, so a thread per point writes six runs of nq scalars
// scattered over the block (measured in round 1: 0.77 TB/s, 12 % of the HBM write rate).  Here a CTA owns one
// contiguous block of the layout -- a warp batch of 32/nq cells (layout 1) or one cell (layout 2) -- builds it in
// shared memory and writes it out front to back with 16-byte stores.
template <typename T, int LAYOUT>
__global__ void __launch_bounds__(256)
    geometry_block_kernel(const T* __restrict__ xgeom, T* __restrict__ G,
                          const std::int32_t* __restrict__ geometry_dofmap, const T* __restrict__ dphi,
                          const T* __restrict__ weights, const std::int32_t* __restrict__ cells,
                          std::int64_t ncells, int nq, int cells_per_block, int block_scalars)
{
  static_assert(LAYOUT == 1 || LAYOUT == 2, "block kernel: slab or DMMA layout");
  extern __shared__ __align__(16) unsigned char geom_smem_raw[];
  T* buf = reinterpret_cast<T*>(geom_smem_raw);
  const int nq2 = nq * nq, nq3 = nq2 * nq;
  const std::int64_t nblocks = (ncells + cells_per_block - 1) / cells_per_block;
  for (std::int64_t blk = blockIdx.x; blk < nblocks; blk += gridDim.x)
  {
    const std::int64_t pos0 = blk * cells_per_block;
    const bool partial = pos0 + cells_per_block > ncells;
    if (partial || (LAYOUT == 2 && nq != 8))
      for (int i = threadIdx.x; i < block_scalars; i += blockDim.x) // entries no listed cell / point owns
        buf[i] = T(0);
    __syncthreads();
    for (int idx = threadIdx.x; idx < cells_per_block * nq3; idx += blockDim.x)
    {
      const int c = idx / nq3, q = idx - c * nq3;
      if (pos0 + c >= ncells)
        continue;
      const std::int32_t* gd = geometry_dofmap + static_cast<std::int64_t>(cells[pos0 + c]) * 8;
      T J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
#pragma unroll
      for (int k = 0; k < 8; ++k)
      {
        const T* xk = xgeom + 3 * static_cast<std::int64_t>(__ldg(gd + k));
        const T x0 = __ldg(xk), x1 = __ldg(xk + 1), x2 = __ldg(xk + 2);
#pragma unroll
        for (int j = 0; j < 3; ++j)
        {
          const T d = __ldg(dphi + (static_cast<std::int64_t>(j) * nq3 + q) * 8 + k);
          J[0][j] += x0 * d;
          J[1][j] += x1 * d;
          J[2][j] += x2 * d;
        }
      }
      // same expressions as geometry_kernel (reference src/geometry_gpu.hpp:100-130)
      const T K00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
      const T K01 = -J[0][1] * J[2][2] + J[0][2] * J[2][1];
      const T K02 = J[0][1] * J[1][2] - J[0][2] * J[1][1];
      const T K10 = -J[1][0] * J[2][2] + J[1][2] * J[2][0];
      const T K11 = J[0][0] * J[2][2] - J[0][2] * J[2][0];
      const T K12 = -J[0][0] * J[1][2] + J[0][2] * J[1][0];
      const T K20 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
      const T K21 = -J[0][0] * J[2][1] + J[0][1] * J[2][0];
      const T K22 = J[0][0] * J[1][1] - J[0][1] * J[1][0];
      const T detJ = J[0][0] * K00 - J[0][1] * K10 + J[0][2] * K20;
      const T sc = __ldg(weights + q) / detJ; // one division per point, see geometry_kernel
      T g[6];
      g[0] = (K00 * K00 + K01 * K01 + K02 * K02) * sc;
      g[1] = (K10 * K00 + K11 * K01 + K12 * K02) * sc;
      g[2] = (K20 * K00 + K21 * K01 + K22 * K02) * sc;
      g[3] = (K10 * K10 + K11 * K11 + K12 * K12) * sc;
      g[4] = (K20 * K10 + K21 * K11 + K22 * K12) * sc;
      g[5] = (K20 * K20 + K21 * K21 + K22 * K22) * sc;
      const int q0 = q / nq2, q1 = (q / nq) % nq, q2 = q % nq;
#pragma unroll
      for (int comp = 0; comp < 6; ++comp)
      {
        const std::int64_t li = LAYOUT == 1 ? slab_g_index(nq, c, comp, q0, q1, q2)
                                            : mma8_g_index(nq, c, comp, q0, q1, q2);
        buf[li] = g[comp];
      }
    }
    __syncthreads();
    T* dst = G + blk * static_cast<std::int64_t>(block_scalars);
    constexpr int V = 16 / int(sizeof(T));
    if (block_scalars % V == 0 && (reinterpret_cast<std::uintptr_t>(dst) & 15) == 0)
    {
      const int4* src4 = reinterpret_cast<const int4*>(buf);
      int4* dst4 = reinterpret_cast<int4*>(dst);
      for (int i = threadIdx.x; i < block_scalars / V; i += blockDim.x)
        __stcs(dst4 + i, src4[i]); // streaming store: G is not re-read before the next kernel
    }
    else
      for (int i = threadIdx.x; i < block_scalars; i += blockDim.x)
        dst[i] = buf[i];
    __syncthreads();
  }
}

template <typename T>
int launch_geometry(int nq, const void* xgeom, const std::int32_t* gdofmap, const void* dphi,
                    const void* weights, const std::int32_t* cells, std::int64_t ncells, void* G,
                    cudaStream_t stream, int layout)
{
  if (layout == 1 || layout == 2)
  {
    const int cpb = layout == 1 ? 32 / nq : 1;
    const std::int64_t bs = layout == 1 ? static_cast<std::int64_t>(nq) * nq * 6 * cpb * nq : mma8_cell_scalars(nq);
    const size_t smem = static_cast<size_t>(bs) * sizeof(T);
    const std::int64_t nblocks = (ncells + cpb - 1) / cpb;
    const std::int64_t capb = static_cast<std::int64_t>(sm_count()) * 4;
    const int gridb = static_cast<int>(nblocks < capb ? nblocks : capb);
    if (layout == 1)
    {
      auto kernel = geometry_block_kernel<T, 1>;
      BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
      kernel<<<gridb, 256, smem, stream>>>(static_cast<const T*>(xgeom), static_cast<T*>(G), gdofmap,
                                           static_cast<const T*>(dphi), static_cast<const T*>(weights), cells, ncells,
                                           nq, cpb, static_cast<int>(bs));
    }
    else
    {
      auto kernel = geometry_block_kernel<T, 2>;
      BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
      kernel<<<gridb, 256, smem, stream>>>(static_cast<const T*>(xgeom), static_cast<T*>(G), gdofmap,
                                           static_cast<const T*>(dphi), static_cast<const T*>(weights), cells, ncells,
                                           nq, cpb, static_cast<int>(bs));
    }
    BD_CUDA_CHECK(cudaGetLastError());
    return BD_SUCCESS;
  }
  const int nq3 = nq * nq * nq;
  const std::int64_t total = ncells * nq3;
  const std::int64_t want = (total + 255) / 256;
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
  const int grid = static_cast<int>(want < cap ? want : cap);
#define BD_GEOM_LAUNCH(L)                                                                          \
  geometry_kernel<T, L><<<grid, 256, 0, stream>>>(static_cast<const T*>(xgeom), static_cast<T*>(G),     \
                                                  gdofmap, static_cast<const T*>(dphi),                \
                                                  static_cast<const T*>(weights), cells, ncells, nq)
  if (layout == 2)
    BD_GEOM_LAUNCH(2);
  else if (layout == 1)
    BD_GEOM_LAUNCH(1);
  else
    BD_GEOM_LAUNCH(0);
#undef BD_GEOM_LAUNCH
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

int geometry_compute_layout(int scalar_bits, int nq, const void* xgeom, const std::int32_t* gdofmap,
                            const void* dphi, const void* weights, const std::int32_t* cells,
                            std::int64_t ncells, void* G, cudaStream_t stream, int layout)
{
  if (ncells == 0)
    return BD_SUCCESS;
  return scalar_bits == 64
             ? launch_geometry<double>(nq, xgeom, gdofmap, dphi, weights, cells, ncells, G, stream, layout)
             : launch_geometry<float>(nq, xgeom, gdofmap, dphi, weights, cells, ncells, G, stream, layout);
}

int layout_g_to_reference(int scalar_bits, int nq, const void* Gs, void* G, std::int64_t ncells,
                          cudaStream_t stream, int layout)
{
  if (ncells == 0)
    return BD_SUCCESS;
  const int grid = sm_count() * 8;
  if (layout == 0)
  {
    const size_t bytes = static_cast<size_t>(ncells) * 6 * nq * nq * nq * (scalar_bits / 8);
    BD_CUDA_CHECK(cudaMemcpyAsync(G, Gs, bytes, cudaMemcpyDeviceToDevice, stream));
    return BD_SUCCESS;
  }
  if (layout == 2)
  {
    if (scalar_bits == 64)
      mma8_g_to_reference_kernel<double><<<grid, 256, 0, stream>>>(nq, static_cast<const double*>(Gs),
                                                                  static_cast<double*>(G), ncells);
    else
      mma8_g_to_reference_kernel<float><<<grid, 256, 0, stream>>>(nq, static_cast<const float*>(Gs),
                                                                 static_cast<float*>(G), ncells);
    BD_CUDA_CHECK(cudaGetLastError());
    return BD_SUCCESS;
  }
  if (scalar_bits == 64)
    slab_g_to_reference_kernel<double><<<grid, 256, 0, stream>>>(nq, static_cast<const double*>(Gs),
                                                                static_cast<double*>(G), ncells);
  else
    slab_g_to_reference_kernel<float><<<grid, 256, 0, stream>>>(nq, static_cast<const float*>(Gs),
                                                               static_cast<float*>(G), ncells);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}
} // namespace bd

extern "C" int bd_geometry_compute(int scalar_bits, int nq, const void* xgeom,
                                   const int32_t* geometry_dofmap, const void* dphi_geometry,
                                   const void* weights, const int32_t* cells, int64_t ncells,
                                   void* G_out, void* stream)
{
  BD_REQUIRE(scalar_bits == 32 || scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  if (nq < 2 || nq > 9)
    return bd::fail(BD_ERR_UNSUPPORTED, "Unsupported degree [geometry]");
  BD_REQUIRE(ncells >= 0, "negative cell count");
  if (ncells == 0)
    return BD_SUCCESS;
  BD_REQUIRE(xgeom && geometry_dofmap && dphi_geometry && weights && cells && G_out,
             "bd_geometry_compute: null pointer argument");
  return scalar_bits == 64
             ? bd::launch_geometry<double>(nq, xgeom, geometry_dofmap, dphi_geometry, weights, cells,
                                           ncells, G_out, bd::as_stream(stream), 0)
             : bd::launch_geometry<float>(nq, xgeom, geometry_dofmap, dphi_geometry, weights, cells,
                                          ncells, G_out, bd::as_stream(stream), 0);
}
