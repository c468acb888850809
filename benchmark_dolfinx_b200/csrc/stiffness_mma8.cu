// Launcher of the fp64 tensor-core kernel (stiffness_mma8.cuh).
#include "stiffness_launch.cuh"
#include "stiffness_mma8.cuh"
#include "stiffness_mma8_otf.cuh"


namespace bd
{
template <typename TS, int P, int NQ, bool IDENTITY, int kMma8Warps, bool PREFETCH>
static int launch_mma8(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap)
    return fail(BD_ERR_INVALID_ARGUMENT, "the DMMA kernel needs the masked dofmap");
  auto kernel = stiffness_mma8_kernel<TS, P, NQ, IDENTITY, kMma8Warps, PREFETCH>;
  constexpr int smem = kMma8Warps * Mma8Ring<TS, NQ>::WARP_SMEM;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kMma8Warps * 32, smem));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  Mma8Tables tab{};
  const TS* phi = static_cast<const TS*>(a.phi0_host);   // [NQ][ND]
  const TS* dphi = static_cast<const TS*>(a.dphi1_host); // [NQ][NQ]
  for (int q = 0; q < 8; ++q)
    for (int i = 0; i < 8; ++i)
    {
      tab.phi[q * 8 + i] = (q < NQ && i < ND) ? phi[q * ND + i] : 0.0;
      tab.dphi[q * 8 + i] = (q < NQ && i < NQ) ? dphi[q * NQ + i] : 0.0;
    }
  const std::int64_t want = (a.ncells + kMma8Warps - 1) / kMma8Warps;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(want, max_grid));
  kernel<<<grid, kMma8Warps * 32, smem, a.stream>>>(
      static_cast<const TS*>(a.u), static_cast<TS*>(a.y), static_cast<const TS*>(a.G),
      static_cast<const TS*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

template <typename TS, int P, int NQ, bool IDENTITY>
static int launch_mma8_tuned(const StiffnessArgs& a)
{
  // 4 warps per CTA, next cell's dof indices prefetched (measured, round 1).  Prefetching the input values as
  // well was tried in round 2 and lost by 2.6x (register spills; profiles/r2_experiments_ab.txt): removed.
  return launch_mma8<TS, P, NQ, IDENTITY, 4, true>(a);
}

int launch_stiffness_mma8(int scalar_bits, int degree, bool identity, const StiffnessArgs& a)
{
#define BD_MMA8_CASE(P, NQ, ID)                                                                    \
  if (degree == P && identity == ID)                                                               \
    return scalar_bits == 64 ? launch_mma8_tuned<double, P, NQ, ID>(a) : launch_mma8_tuned<float, P, NQ, ID>(a);
  BD_MMA8_CASE(6, 8, false)
  BD_MMA8_CASE(7, 8, true)
  BD_MMA8_CASE(5, 7, false)
  BD_MMA8_CASE(6, 7, true)
#undef BD_MMA8_CASE
  return fail(BD_ERR_UNSUPPORTED, "the DMMA kernel covers nq = 7 and nq = 8 only");
}

// MINB: 3 CTAs of 4 warps per SM for the affine-cell kernel (168 registers, 16 bytes of spills); the general-cell
// kernel runs uncapped (238 registers, 2 CTAs per SM): capped at 168 it spills and is 2.5x slower (measured).
// FE: a scheduling fence every 4 y-tiles for the affine-cell kernel (measured at Q6 / 500M dofs: a fence per tile
// 34.5, every 2 tiles 37.1, every 4 tiles 37.8 GDoF/s; every 4 or 8 tiles at 2 CTAs per SM 33.7 / 35.5).
template <typename TS, int P, int NQ, bool IDENTITY, bool AFFINE_ONLY, int MINB = 1, bool PREFETCH = true, int FE = 1>
static int launch_mma8_otf(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  constexpr int kWarps = 4;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap || !a.xgeom || !a.gdofmap || !a.w1d || !a.pts1d)
    return fail(BD_ERR_INVALID_ARGUMENT, "on-the-fly geometry needs the masked dofmap, vertices and 1-D rule");
  auto kernel = stiffness_mma8_otf_kernel<TS, P, NQ, IDENTITY, kWarps, AFFINE_ONLY, MINB, PREFETCH, FE>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, kWarps * 32, 0));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  Mma8OtfTables tab{};
  const TS* phi = static_cast<const TS*>(a.phi0_host);   // [NQ][ND]
  const TS* dphi = static_cast<const TS*>(a.dphi1_host); // [NQ][NQ]
  for (int q = 0; q < 8; ++q)
  {
    tab.w[q] = q < NQ ? a.w1d[q] : 0.0;
    tab.pts[q] = q < NQ ? a.pts1d[q] : 0.5;
    for (int i = 0; i < 8; ++i)
    {
      tab.base.phi[q * 8 + i] = (q < NQ && i < ND) ? phi[q * ND + i] : 0.0;
      tab.base.dphi[q * 8 + i] = (q < NQ && i < NQ) ? dphi[q * NQ + i] : 0.0;
    }
  }
  const std::int64_t want = (a.ncells + kWarps - 1) / kWarps;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(want, max_grid));
  kernel<<<grid, kWarps * 32, 0, a.stream>>>(
      static_cast<const TS*>(a.u), static_cast<TS*>(a.y), static_cast<const TS*>(a.xgeom), a.gdofmap,
      static_cast<const TS*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

int launch_stiffness_mma8_otf(int scalar_bits, int degree, bool identity, const StiffnessArgs& a)
{
#define BD_MMA8_OTF_CASE(P, NQ, ID)                                                                \
  if (degree == P && identity == ID)                                                               \
  {                                                                                                \
    if (scalar_bits == 64)                                                                         \
      return a.geom_mode == 1 ? launch_mma8_otf<double, P, NQ, ID, true, 3, true, 4>(a)            \
                              : launch_mma8_otf<double, P, NQ, ID, false, 1>(a);                   \
    return a.geom_mode == 1 ? launch_mma8_otf<float, P, NQ, ID, true, 3, true, 4>(a)               \
                            : launch_mma8_otf<float, P, NQ, ID, false, 1>(a);                      \
  }
  BD_MMA8_OTF_CASE(6, 8, false)
  BD_MMA8_OTF_CASE(7, 8, true)
  BD_MMA8_OTF_CASE(5, 7, false)
  BD_MMA8_OTF_CASE(6, 7, true)
#undef BD_MMA8_OTF_CASE
  return fail(BD_ERR_UNSUPPORTED, "on-the-fly geometry: the DMMA kernel covers nq = 7 and nq = 8 only");
}
} // namespace bd
