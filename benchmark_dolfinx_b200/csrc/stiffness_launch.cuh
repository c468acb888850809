// Host-side launcher shared by the per-(scalar, qmode) instantiation units.
#pragma once

#include "common.cuh"
#include "stiffness.cuh"
#include "stiffness_slab.cuh"
#include "stiffness_pencil.cuh"
#include "stiffness_pencil_bulk.cuh"
#include "stiffness_slab_otf.cuh"
#include "stiffness_affine.cuh"
#include "stiffness_affine_pipe.cuh"

namespace bd
{
// 0 = default choice, 1 = plain slab kernel, 2 = asynchronous-copy slab kernel, 3 = pencil kernel wherever
// it is instantiated (env BD_STIFFNESS_VARIANT = slab | tma | pencil)
int stiffness_variant();

struct StiffnessArgs
{
  const void* u;
  void* y;
  const void* G;
  const void* cell_constants;
  const std::int32_t* dofmap;
  const std::int32_t* cells;
  std::int64_t ncells;
  const std::int8_t* bc;
  const void* phi0_host;  // T[nq*nd]
  const void* dphi1_host; // T[nq*nq]
  cudaStream_t stream;
  const std::int32_t* mdofmap = nullptr; // dofmap with the BC marker in bit 31 (asynchronous-copy kernel)
  bool slab = false; // G is in the slab layout (stiffness_slab.cuh) and the slab kernel is used
  int g_layout = 0;  // 0 reference, 1 slab, 2 DMMA accumulator order
  // geometry on the fly (stiffness_slab_otf.cuh, stiffness_mma8_otf.cuh): no G; 1 = every cell of the list is
  // affine (verified by the caller), 2 = general cells (affine batches still take the shortcut)
  int geom_mode = 0;
  const std::int8_t* merge = nullptr; // per list position: face shared with the next listed cell (build_merge_flags)
  bool g_diagonal = false; // geom_mode 1 and every cell has orthogonal edges: G_cell is diagonal (dof-space kernels)
  const void* xgeom = nullptr;           // T [nnodes][3]
  const std::int32_t* gdofmap = nullptr; // [ncells][8], tensor-product vertex order
  const double* w1d = nullptr;           // host, [nq] 1-D quadrature weights
  const double* pts1d = nullptr;         // host, [nq] 1-D quadrature points
};

template <typename T, int P, int NQ, bool IDENTITY>
int launch_stiffness(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  constexpr int CPB = cells_per_block(NQ);
  constexpr int THREADS = block_threads(NQ);
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  auto kernel = stiffness_kernel<T, P, NQ, IDENTITY>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, THREADS, 0));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  Tables1D<T, ND, NQ> tab;
  const T* phi = static_cast<const T*>(a.phi0_host);
  const T* dphi = static_cast<const T*>(a.dphi1_host);
  for (int i = 0; i < NQ * ND; ++i)
    tab.phi[i] = phi[i];
  for (int i = 0; i < NQ * NQ; ++i)
    tab.dphi[i] = dphi[i];
  const std::int64_t nbatches = (a.ncells + CPB - 1) / CPB;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(nbatches, max_grid));
  kernel<<<grid, THREADS, 0, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.G),
      static_cast<const T*>(a.cell_constants), a.dofmap, a.cells, static_cast<int>(a.ncells), a.bc,
      tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

template <typename T, int ND, int NQ, bool IDENTITY>
void fill_slab_tables(SlabTables<T, ND, NQ>& tab, const void* phi0_host, const void* dphi1_host)
{
  // Phi and DPhi = dphi1 * phi0 (products accumulated in double)
  const T* phi = static_cast<const T*>(phi0_host);
  const T* dphi = static_cast<const T*>(dphi1_host);
  for (int q = 0; q < NQ; ++q)
    for (int i = 0; i < ND; ++i)
    {
      tab.phi[q * ND + i] = phi[q * ND + i];
      if (IDENTITY)
        tab.dphi[q * ND + i] = dphi[q * NQ + i];
      else
      {
        double s = 0;
        for (int r = 0; r < NQ; ++r)
          s += static_cast<double>(dphi[q * NQ + r]) * static_cast<double>(phi[r * ND + i]);
        tab.dphi[q * ND + i] = static_cast<T>(s);
      }
    }
}

// Tuning of the asynchronous-copy kernel per 1-D quadrature size: ring depth, q0 rows per slot,
// warps per CTA and whether the next batch's gather is prefetched into registers.
template <int NQ>
struct SlabTmaTuning
{
  // measured (scripts/gpu_tune*.sh): nq <= 5: 2 slots of a whole q1 slice; nq = 6: 3 half-slice slots
  static constexpr int NSTAGE = NQ <= 5 ? 2 : (NQ == 6 ? 3 : 4);
  static constexpr int Q0C = NQ <= 5 ? NQ : (NQ == 6 ? 3 : (NQ == 8 ? 2 : (NQ == 9 ? 3 : NQ)));
  static constexpr int WARPS = 1;
  static constexpr bool PREFETCH = NQ <= 5; // also nq = 6 with nd = 5 (see launch_stiffness_slab_tma)
};

// Slab kernel with the G stream on asynchronous bulk copies (stiffness_slab.cuh), explicit tuning.
template <typename T, int P, int NQ, bool IDENTITY, int NSTAGE, int Q0C_, int WARPS_, bool PREFETCH_>
int launch_stiffness_slab_tma_cfg(const StiffnessArgs& a)
{
  using C = SlabConfig<T, P, NQ>;
  using TC = SlabTmaConfig<T, P, NQ, NSTAGE, Q0C_, WARPS_>;
  constexpr int ND = P + 1;
  if constexpr (!TC::SUPPORTED)
    return fail(BD_ERR_UNSUPPORTED, "asynchronous-copy slab kernel unsupported for this configuration");
  else
  {
    if (!a.mdofmap)
      return fail(BD_ERR_INVALID_ARGUMENT, "the asynchronous-copy kernel needs the masked dofmap");
    auto kernel = stiffness_slab_tma_kernel<T, P, NQ, IDENTITY, NSTAGE, Q0C_, WARPS_, PREFETCH_>;
    static thread_local int blocks_per_sm[64] = {0};
    int dev = 0;
    BD_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64)
      return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
    if (blocks_per_sm[dev] == 0)
    {
      BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TC::CTA_SMEM));
      int nb = 0;
      BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, TC::THREADS, TC::CTA_SMEM));
      blocks_per_sm[dev] = nb > 0 ? nb : 1;
    }
    SlabTables<T, ND, NQ> tab;
    fill_slab_tables<T, ND, NQ, IDENTITY>(tab, a.phi0_host, a.dphi1_host);
    const std::int64_t nbatches = (a.ncells + C::CW - 1) / C::CW;
    const std::int64_t want = (nbatches + TC::WARPS - 1) / TC::WARPS;
    const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
    const int grid = static_cast<int>(clamp_grid(want, max_grid));
    kernel<<<grid, TC::THREADS, TC::CTA_SMEM, a.stream>>>(
        static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.G),
        static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
    BD_CUDA_CHECK(cudaGetLastError());
    return BD_SUCCESS;
  }
}

template <typename T, int P, int NQ, bool IDENTITY>
int launch_stiffness_slab_tma(const StiffnessArgs& a)
{
  using TU = SlabTmaTuning<NQ>;
  constexpr int NSTAGE = (NQ == 7) ? 2 : TU::NSTAGE;
  // nq <= 5: ring of 2 slots, 1 warp per CTA (7 warps/SM at Q3 fp64) measured ahead of 3 slots / 2 warps per
  // CTA (15.2 vs 16.1 ms at Q3/300M, round 1)
  // Round 2 A/B at Q3 / 300M dofs fp64 (profiles/r2_ring_ab.txt): finer slots with more of them in flight (one q0 row
  // per slot, rings of 4 / 6 / 8 / 10 slots, 8-9 CTAs per SM) are slower (17.2-18.1 GDoF/s) than 2 whole-slice slots
  // (19.8), and so is a third whole-slice slot (18.7): the per-copy cost and the lost occupancy outweigh the deeper queue.
  // the gather prefetch needs 3 * nd^2 extra registers: affordable up to nd = 5
  constexpr bool PF = TU::PREFETCH || (NQ == 6 && P + 1 <= 5);
  return launch_stiffness_slab_tma_cfg<T, P, NQ, IDENTITY, NSTAGE, TU::Q0C, TU::WARPS, PF>(a);
}

template <typename T, int P, int NQ>
constexpr bool slab_tma_supported()
{
  using TU = SlabTmaTuning<NQ>;
  return SlabTmaConfig<T, P, NQ, (NQ == 7) ? 2 : TU::NSTAGE, TU::Q0C, TU::WARPS>::SUPPORTED;
}

// The register-blocked kernel (stiffness_slab.cuh); `a.G` must be in the slab layout.
template <typename T, int P, int NQ, bool IDENTITY>
int launch_stiffness_slab(const StiffnessArgs& a)
{
  using C = SlabConfig<T, P, NQ>;
  constexpr int ND = P + 1;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  // kernel selection (BD_STIFFNESS_VARIANT=slab|tma overrides; measured choices in DESIGN.md)
  if constexpr (slab_tma_supported<T, P, NQ>())
  {
    if (stiffness_variant() != 1 && a.mdofmap)
      return launch_stiffness_slab_tma<T, P, NQ, IDENTITY>(a);
  }
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  auto kernel = stiffness_slab_kernel<T, P, NQ, IDENTITY>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, C::THREADS, C::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  SlabTables<T, ND, NQ> tab;
  fill_slab_tables<T, ND, NQ, IDENTITY>(tab, a.phi0_host, a.dphi1_host);
  const std::int64_t nbatches = (a.ncells + C::CW - 1) / C::CW;
  const std::int64_t want = (nbatches + C::WARPS - 1) / C::WARPS;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(want, max_grid));
  kernel<<<grid, C::THREADS, C::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.G),
      static_cast<const T*>(a.cell_constants), a.dofmap, a.cells, static_cast<int>(a.ncells), a.bc,
      tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// Slab kernel with on-the-fly geometry (stiffness_slab_otf.cuh), nq <= 6.
template <typename T, int P, int NQ, bool IDENTITY, bool AFFINE_ONLY, int MINB = 1>
int launch_stiffness_slab_otf_cfg(const StiffnessArgs& a)
{
  using C = SlabConfig<T, P, NQ>;
  constexpr int ND = P + 1;
  constexpr int WARPS = 1;
  // the gather prefetch needs 3 * nd^2 extra registers: affordable up to nd = 5 (as in the stored-G kernel)
  constexpr bool PF = NQ <= 5 || (NQ == 6 && ND <= 5);
  using OC = SlabOtfConfig<T, P, NQ, WARPS>;
  static_assert(OC::SUPPORTED, "exchange cube exceeds shared memory");
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap || !a.xgeom || !a.gdofmap || !a.w1d || !a.pts1d)
    return fail(BD_ERR_INVALID_ARGUMENT, "on-the-fly geometry needs the masked dofmap, vertices and 1-D rule");
  auto kernel = stiffness_slab_otf_kernel<T, P, NQ, IDENTITY, AFFINE_ONLY, WARPS, PF, MINB>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OC::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, OC::THREADS, OC::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  SlabOtfTables<T, ND, NQ> tab;
  fill_slab_tables<T, ND, NQ, IDENTITY>(tab.fwd, a.phi0_host, a.dphi1_host);
  for (int q = 0; q < NQ; ++q)
  {
    tab.w[q] = static_cast<T>(a.w1d[q]);
    tab.pts[q] = static_cast<T>(a.pts1d[q]);
    for (int i = 0; i < ND; ++i)
    {
      // weights folded into the transposed contractions; with Phi = I they are applied per point instead
      const double wq = IDENTITY ? 1.0 : a.w1d[q];
      tab.bwd.phi[q * ND + i] = static_cast<T>(wq * static_cast<double>(tab.fwd.phi[q * ND + i]));
      tab.bwd.dphi[q * ND + i] = static_cast<T>(wq * static_cast<double>(tab.fwd.dphi[q * ND + i]));
    }
  }
  const std::int64_t nbatches = (a.ncells + C::CW - 1) / C::CW;
  const std::int64_t want = (nbatches + WARPS - 1) / WARPS;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(want, max_grid));
  kernel<<<grid, OC::THREADS, OC::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.xgeom), a.gdofmap,
      static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// Where the dof-space kernel (stiffness_affine.cuh) replaces the quadrature-space on-the-fly kernel on meshes of
// parallelepipeds (measured on one B200, profiles/r2_affine_ab.txt): qmode 1 with degree <= 3 (Q3 fp64 44.7 vs 32.0
// GDoF/s, Q2 34.2 vs 24.3, Q1 14.7 vs 8.0, Q3 fp32 74.1 vs 46.9) and degree 4 in fp32 (70.0 vs 56.6).  Not at
// qmode 0, where Phi = I makes the quadrature-space kernel the cheaper one (Q3 59.9 vs 41.4, Q2 49.2 vs 34.2,
// Q1 19.3 vs 14.7), and not at degree 4 in fp64 (24.3 vs 32.1: its exchange buffer leaves 3 warps per SM).
constexpr bool affine_kernel_used(int scalar_bits, int degree, bool identity)
{
  return !identity && (degree <= 3 || (degree == 4 && scalar_bits == 32));
}

// Host-side 1-D operators M, S, C of the dof-space kernels, from the operator's own tables (in double).
template <typename T, int ND, int NQ, bool IDENTITY>
void fill_affine_tables(AffineTables<T, ND>& tab, const StiffnessArgs& a)
{
  SlabTables<T, ND, NQ> st;
  fill_slab_tables<T, ND, NQ, IDENTITY>(st, a.phi0_host, a.dphi1_host);
  constexpr int NDP = AffineTables<T, ND>::NDP;
  for (int i = 0; i < ND * NDP; ++i)
    tab.M[i] = tab.S[i] = tab.C[i] = tab.CT[i] = T(0);
  for (int r = 0; r < ND; ++r)
    for (int c = 0; c < ND; ++c)
    {
      // M and S from the (r <= c) triangle and mirrored: exactly symmetric, the kernels read their columns as rows
      const int a0 = r < c ? r : c, b0 = r < c ? c : r;
      double m = 0, sv = 0, cv = 0;
      for (int q = 0; q < NQ; ++q)
      {
        const double w = a.w1d[q];
        m += w * st.phi[q * ND + a0] * st.phi[q * ND + b0];
        sv += w * st.dphi[q * ND + a0] * st.dphi[q * ND + b0];
        cv += w * st.dphi[q * ND + r] * st.phi[q * ND + c];
      }
      tab.M[r * NDP + c] = static_cast<T>(m);
      tab.S[r * NDP + c] = static_cast<T>(sv);
      tab.C[r * NDP + c] = static_cast<T>(cv);
      tab.CT[c * NDP + r] = static_cast<T>(cv);
    }
}

// Dof-space kernel for meshes of parallelepipeds (stiffness_affine.cuh), degree <= 3.
template <typename T, int P, int NQ, bool IDENTITY, int MINB, int KCH, bool DIAG = false, bool SE = false>
int launch_stiffness_affine(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  using AC = AffineConfig<T, P, 1, KCH, false, DIAG ? 2 : 4>;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap || !a.xgeom || !a.gdofmap || !a.w1d)
    return fail(BD_ERR_INVALID_ARGUMENT, "the affine-cell kernel needs the masked dofmap, vertices and 1-D rule");
  auto kernel = stiffness_affine_kernel<T, P, 1, MINB, KCH, DIAG, SE>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AC::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, AC::THREADS, AC::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  AffineTables<T, ND> tab;
  fill_affine_tables<T, ND, NQ, IDENTITY>(tab, a);
  const std::int64_t nbatches = (a.ncells + AC::CW - 1) / AC::CW;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(nbatches, max_grid));
  kernel<<<grid, AC::THREADS, AC::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.xgeom), a.gdofmap,
      static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// Dof-space kernel with the asynchronous gather pipeline (stiffness_affine_pipe.cuh).
template <typename T, int P, int NQ, bool IDENTITY, int MINB, bool DIAG, int KCH>
int launch_stiffness_affine_pipe(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  using PC = AffinePipeConfig<T, P, 1, DIAG, KCH>;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap || !a.xgeom || !a.gdofmap || !a.w1d)
    return fail(BD_ERR_INVALID_ARGUMENT, "the affine-cell kernel needs the masked dofmap, vertices and 1-D rule");
  auto kernel = stiffness_affine_pipe_kernel<T, P, 1, MINB, DIAG, KCH>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PC::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, PC::THREADS, PC::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  AffineTables<T, ND> tab;
  fill_affine_tables<T, ND, NQ, IDENTITY>(tab, a);
  const std::int64_t nbatches = (a.ncells + PC::CW - 1) / PC::CW;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(nbatches, max_grid));
  kernel<<<grid, PC::THREADS, PC::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.xgeom), a.gdofmap,
      static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, a.merge, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// Dof-space kernel with the plane in shared memory and column rounds (stiffness_affine_col.cuh), degree 4..7.
template <typename T, int P, int NQ, bool IDENTITY, int MINB, bool COMPACT, bool DIAG = false>
int launch_stiffness_affine_col(const StiffnessArgs& a)
{
  constexpr int ND = P + 1;
  using CC = AffineColConfig<T, P, 1, COMPACT, DIAG>;
  if (a.ncells == 0)
    return BD_SUCCESS;
  if (a.ncells > 0x7fffffffLL)
    return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
  if (!a.mdofmap || !a.xgeom || !a.gdofmap || !a.w1d)
    return fail(BD_ERR_INVALID_ARGUMENT, "the affine-cell kernel needs the masked dofmap, vertices and 1-D rule");
  auto kernel = stiffness_affine_col_kernel<T, P, 1, MINB, COMPACT, DIAG>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CC::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, CC::THREADS, CC::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  AffineTables<T, ND> tab;
  fill_affine_tables<T, ND, NQ, IDENTITY>(tab, a);
  const std::int64_t nbatches = (a.ncells + CC::CW - 1) / CC::CW;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(nbatches, max_grid));
  kernel<<<grid, CC::THREADS, CC::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.xgeom), a.gdofmap,
      static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

template <typename T, int P, int NQ, bool IDENTITY>
int launch_stiffness_slab_otf(const StiffnessArgs& a)
{
  if constexpr (affine_kernel_used(int(sizeof(T)) * 8, P, IDENTITY))
  {
    // every cell a parallelepiped: the dof-space kernel (one exchange round, uncapped registers: measured best,
    // profiles/r2_affine_ab.txt)
    if (a.geom_mode == 1)
    {
      if (a.g_diagonal)
      {
        // bricks.  nd = 4: the asynchronous-gather kernel with merged shared faces (fp64: one exchange column per
        // round and <= 128 registers = 16 warps/SM, 78.9 GDoF/s against 69.9 with the register pipeline; fp32: whole
        // planes, 118.7 against 92.4).  Other nd: the register pipeline (the 4-byte index copies of the asynchronous
        // one lose there: Q2 fp64 50.4 vs 65.6), at degree 4 / fp32 capped to 128 registers (118.5 vs 107.9).
        // profiles/r2_brick_ab.txt
        if constexpr (P == 3 && sizeof(T) == 8)
          return launch_stiffness_affine_pipe<T, P, NQ, IDENTITY, 16, true, 1>(a);
        else if constexpr (P == 3)
          return launch_stiffness_affine_pipe<T, P, NQ, IDENTITY, 1, true, P + 1>(a);
        else
          return launch_stiffness_affine<T, P, NQ, IDENTITY, (P == 4 ? 16 : 1), P + 1, true, true>(a);
      }
      return launch_stiffness_affine<T, P, NQ, IDENTITY, 1, P + 1, false>(a);
    }
  }
  if constexpr (NQ > 6)
    return fail(BD_ERR_UNSUPPORTED, "on-the-fly geometry: the slab kernel covers nq <= 6");
  else
  {
    // Register cap (measured, profiles/r2_onthefly_*_ab.txt): 168 registers = 12 one-warp CTAs per SM wins for the
    // fp64 qmode-1 kernels at nq = 4, 5 (Q3: 30.3 -> 31.2 GDoF/s, Q2: 22.8 -> 24.3); it loses where the kernel then
    // spills (nq = 6: 31.9 -> 22.4) or where Phi = I leaves the registers free anyway (Q4 qmode 0: 69.0 -> 61.5).
    // The general-cell kernel runs uncapped (capped it spills: Q3 perturbed 18.6 -> 15.2 GDoF/s).
    constexpr int MINB = (sizeof(T) == 8 && !IDENTITY && (NQ == 4 || NQ == 5)) ? 12 : 1;
    return a.geom_mode == 1 ? launch_stiffness_slab_otf_cfg<T, P, NQ, IDENTITY, true, MINB>(a)
                            : launch_stiffness_slab_otf_cfg<T, P, NQ, IDENTITY, false, 1>(a);
  }
}

// The pencil kernel (stiffness_pencil.cuh) is instantiated where it wins or is the only kernel without
// spills (A/B in profiles/): qmode 1 with nq = 9, and fp32 with nq = 7, 8.
template <typename T, int NQ, bool IDENTITY>
constexpr bool pencil_instantiated()
{
  return !IDENTITY && (NQ == 9 || (sizeof(T) == 4 && (NQ == 7 || NQ == 8)));
}

template <typename T, int P, int NQ, int CPB, int RD, bool UPF, bool STREAM, bool PAIRED>
int launch_stiffness_pencil_cfg(const StiffnessArgs& a)
{
  using C = PencilConfig<T, P, NQ, CPB, RD>;
  static_assert(C::SUPPORTED, "pencil kernel: shared memory or CTA size out of range");
  static_assert((C::ND - 1) * C::S1J + NQ <= C::S1I, "pencil kernel: S1 rows overlap");
  constexpr int ND = P + 1;
  auto kernel = stiffness_pencil_kernel<T, P, NQ, CPB, RD, UPF, STREAM, PAIRED>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, C::THREADS, C::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  SlabTables<T, ND, NQ> tab;
  fill_slab_tables<T, ND, NQ, false>(tab, a.phi0_host, a.dphi1_host);
  const std::int64_t nbatches = (a.ncells + CPB - 1) / CPB;
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(nbatches, max_grid));
  PencilPairTables<PAIRED ? ND : 1, PAIRED ? NQ : 1> ptab{};
  if constexpr (PAIRED)
  {
    constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
    auto at = [](const T* t, int q, int k) { return (q < NQ && k < ND) ? static_cast<float>(t[q * ND + k]) : 0.f; };
    for (int k = 0; k < ND; ++k)
      for (int qp = 0; qp < NQP; ++qp)
      {
        ptab.fphi[k * NQP + qp] = make_float2(at(tab.phi, 2 * qp, k), at(tab.phi, 2 * qp + 1, k));
        ptab.fdphi[k * NQP + qp] = make_float2(at(tab.dphi, 2 * qp, k), at(tab.dphi, 2 * qp + 1, k));
      }
    for (int q = 0; q < NQ; ++q)
      for (int kp = 0; kp < NDP; ++kp)
      {
        ptab.bphi[q * NDP + kp] = make_float2(at(tab.phi, q, 2 * kp), at(tab.phi, q, 2 * kp + 1));
        ptab.bdphi[q * NDP + kp] = make_float2(at(tab.dphi, q, 2 * kp), at(tab.dphi, q, 2 * kp + 1));
      }
  }
  kernel<<<grid, C::THREADS, C::CTA_SMEM, a.stream>>>(
      static_cast<const T*>(a.u), static_cast<T*>(a.y), static_cast<const T*>(a.G),
      static_cast<const T*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab, ptab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// The experimental fp32 pencil kernel with one bulk copy of G per cell (stiffness_pencil_bulk.cuh).
template <int P, int NQ, bool PAIRED>
int launch_stiffness_pencil_bulk(const StiffnessArgs& a)
{
  using C = PencilBulkConfig<P, NQ>;
  constexpr int ND = P + 1;
  auto kernel = stiffness_pencil_bulk_kernel<P, NQ, PAIRED>;
  static thread_local int blocks_per_sm[64] = {0};
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  if (blocks_per_sm[dev] == 0)
  {
    BD_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, C::CTA_SMEM));
    int nb = 0;
    BD_CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, C::THREADS, C::CTA_SMEM));
    blocks_per_sm[dev] = nb > 0 ? nb : 1;
  }
  SlabTables<float, ND, NQ> tab;
  fill_slab_tables<float, ND, NQ, false>(tab, a.phi0_host, a.dphi1_host);
  PencilPairTables<PAIRED ? ND : 1, PAIRED ? NQ : 1> ptab{};
  if constexpr (PAIRED)
  {
    constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
    auto at = [](const float* t, int q, int k) { return (q < NQ && k < ND) ? t[q * ND + k] : 0.f; };
    for (int k = 0; k < ND; ++k)
      for (int qp = 0; qp < NQP; ++qp)
      {
        ptab.fphi[k * NQP + qp] = make_float2(at(tab.phi, 2 * qp, k), at(tab.phi, 2 * qp + 1, k));
        ptab.fdphi[k * NQP + qp] = make_float2(at(tab.dphi, 2 * qp, k), at(tab.dphi, 2 * qp + 1, k));
      }
    for (int q = 0; q < NQ; ++q)
      for (int kp = 0; kp < NDP; ++kp)
      {
        ptab.bphi[q * NDP + kp] = make_float2(at(tab.phi, q, 2 * kp), at(tab.phi, q, 2 * kp + 1));
        ptab.bdphi[q * NDP + kp] = make_float2(at(tab.dphi, q, 2 * kp), at(tab.dphi, q, 2 * kp + 1));
      }
  }
  const std::int64_t max_grid = static_cast<std::int64_t>(sm_count()) * blocks_per_sm[dev];
  const int grid = static_cast<int>(clamp_grid(a.ncells, max_grid));
  kernel<<<grid, C::THREADS, C::CTA_SMEM, a.stream>>>(
      static_cast<const float*>(a.u), static_cast<float*>(a.y), static_cast<const float*>(a.G),
      static_cast<const float*>(a.cell_constants), a.mdofmap, a.cells, static_cast<int>(a.ncells), tab, ptab);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// `a.G` in the reference layout (g_layout 0), `a.mdofmap` required.
template <typename T, int P, int NQ, bool IDENTITY>
int launch_stiffness_pencil(const StiffnessArgs& a)
{
  if constexpr (!pencil_instantiated<T, NQ, IDENTITY>())
    return fail(BD_ERR_UNSUPPORTED, "the pencil kernel is not built for degree %d, nq %d", P, NQ);
  else
  {
    if (a.ncells == 0)
      return BD_SUCCESS;
    if (a.ncells > 0x7fffffffLL)
      return fail(BD_ERR_INVALID_ARGUMENT, "cell list of %lld entries exceeds int32", (long long)a.ncells);
    if (!a.mdofmap)
      return fail(BD_ERR_INVALID_ARGUMENT, "the pencil kernel needs the masked dofmap");
    // One cell per CTA throughout (measured best for nq = 7, 8, 9).  Choices measured on one B200
    // (profiles/r1_pencil_*, profiles/r2_experiments_ab.txt):
    //   fp32, nq = 8: one bulk copy of the cell's G block per cell (stiffness_pencil_bulk.cuh), 43.3 -> 52.0 GDoF/s;
    //   fp32, nq = 7, 9: G rows requested a whole cell ahead, next cell's input prefetched; FFMA2 contractions at
    //                    nq = 9 (+16 %), scalar at nq = 7;
    //   fp64 (nq = 9): ring of 5 rows restarted per cell, no input prefetch (151 registers, 48 KB, 4 CTAs/SM),
    //                  19.5 -> 21.2 GDoF/s against the 3-row ring.
    if constexpr (sizeof(T) == 4 && PencilBulkConfig<P, NQ>::SUPPORTED)
      return launch_stiffness_pencil_bulk<P, NQ, false>(a);
    else if constexpr (sizeof(T) == 4)
      return launch_stiffness_pencil_cfg<T, P, NQ, 1, NQ + 1, true, true, NQ == 9>(a);
    else
      return launch_stiffness_pencil_cfg<T, P, NQ, 1, 5, false, false, false>(a);
  }
}

// One function per (scalar, qmode), each defined in its own translation unit so
// the 28 kernel instantiations compile in parallel.
int launch_stiffness_f64_q0(int degree, const StiffnessArgs& a);
int launch_stiffness_f64_q1(int degree, const StiffnessArgs& a);
int launch_stiffness_f32_q0(int degree, const StiffnessArgs& a);
int launch_stiffness_f32_q1(int degree, const StiffnessArgs& a);
// on-the-fly geometry, nq <= 6 (same four units, compiled with -DBD_INST_OTF)
int launch_stiffness_otf_f64_q0(int degree, const StiffnessArgs& a);
int launch_stiffness_otf_f64_q1(int degree, const StiffnessArgs& a);
int launch_stiffness_otf_f32_q0(int degree, const StiffnessArgs& a);
int launch_stiffness_otf_f32_q1(int degree, const StiffnessArgs& a);
// in stiffness_affine_col.cu: dof-space kernel for parallelepiped cells, degree 4..7, qmode 1
int launch_stiffness_affine_col_any(int scalar_bits, int degree, const StiffnessArgs& a);
// in stiffness_mma8.cu: on-the-fly geometry for nq = 7, 8
int launch_stiffness_mma8_otf(int scalar_bits, int degree, bool identity, const StiffnessArgs& a);
// The column-round dof-space kernel (stiffness_affine_col.cuh) exists for qmode 1 at degree 4..7 ...
constexpr bool affine_dofspace_exists(int degree, bool identity) { return !identity && degree >= 1 && degree <= 7; }
// ... and is used where it measured faster than the alternatives on one B200 (profiles/r2_affine_col_ab.txt, GDoF/s):
//   fp64: degree 4 46.7 (quadrature-space on the fly 32.1), degree 5 49.3 (stored G, DMMA 24.6), degree 6 40.3 (DMMA on
//         the fly 37.9, stored G 34.7), degree 7 40.7 (stored G, pencils 21.1);
//   fp32: degree 5 74.3 (stored G 32.5), degree 6 68.8 (50.3), degree 7 62.3 (41.3); NOT degree 4, where the
//         register-plane kernel of stiffness_affine.cuh does 76.2.
constexpr bool affine_col_used(int scalar_bits, int degree, bool identity)
{
  if (identity)
    return false;
  return scalar_bits == 64 ? (degree >= 4 && degree <= 7) : (degree == 5 || degree == 6 || degree == 7);
}
// unpadded exchange buffer (2-way bank conflicts on its reads for some nd, more warps per SM): where it measured faster
// or equal
constexpr bool affine_col_compact(int scalar_bits, int degree)
{
  return scalar_bits == 64 ? (degree == 4 || degree == 6) : (degree == 5 || degree == 6);
}
// on-the-fly geometry exists for this (degree, nq)?
inline bool otf_supported(int degree, int nq) { return nq >= 2 && nq <= 8 && (nq == degree + 1 || nq == degree + 2); }
// ... and is the measured-faster choice on a mesh of parallelepipeds (profiles/r2_onthefly_*_ab.txt, r2_affine*_ab.txt,
// one B200) -- which, with the dof-space kernels, is now every (scalar type, degree, qmode):
//   qmode 1: the dof-space kernels (affine_kernel_used / affine_col_used);
//   qmode 0, nq <= 6 (slab kernel): fp64 1.45-1.7x, fp32 1.35-1.5x over streaming G;
//   qmode 0, nq = 7 (DMMA kernel): fp64 59.5 vs 52.9, fp32 60.1 vs 22.3 GDoF/s;
//   qmode 0, nq = 8 (DMMA kernel): fp64 78.2 vs 51.4, fp32 85.7 vs 85.6 (a tie: on the fly for the memory).
inline bool otf_preferred(int scalar_bits, int degree, int nq, bool identity)
{
  if (affine_kernel_used(scalar_bits, degree, identity) || affine_col_used(scalar_bits, degree, identity))
    return true;
  return otf_supported(degree, nq) && (nq <= 6 || nq == 8 || (nq == 7 && identity));
}

// in geometry.cu
// G layouts: 0 = reference [cell][6][nq^3] (pencil kernel), 1 = slab (stiffness_slab.cuh), 2 = DMMA
// accumulator order (stiffness_mma8.cuh; nq = 7, 8)
int geometry_compute_layout(int scalar_bits, int nq, const void* xgeom, const std::int32_t* gdofmap,
                            const void* dphi, const void* weights, const std::int32_t* cells,
                            std::int64_t ncells, void* G, cudaStream_t stream, int layout);
int layout_g_to_reference(int scalar_bits, int nq, const void* Gs, void* G, std::int64_t ncells,
                          cudaStream_t stream, int layout);
// Layout the operator should use for (scalar_bits, nq) and the scalars it needs for ncells cells.
int preferred_g_layout(int scalar_bits, int degree, int nq);
std::int64_t g_layout_size(int layout, int nq, std::int64_t ncells);
// in stiffness_mma8.cu
int launch_stiffness_mma8(int scalar_bits, int degree, bool identity, const StiffnessArgs& a);
// in stiffness.cu: which consecutive cells of a list (in batches of cw) share their k = nd - 1 / k = 0 faces
int build_merge_flags(std::int64_t ncells, int nd, int cw, const std::int32_t* mdofmap, const std::int32_t* cells,
                      std::int8_t* flags, cudaStream_t stream);
// in stiffness.cu: out[i] = dofmap[i] | (bc[dofmap[i]] ? 0x80000000 : 0)
int build_masked_dofmap(std::int64_t n, const std::int32_t* dofmap, const std::int8_t* bc,
                        std::int32_t* out, cudaStream_t stream);

// Dispatch on (scalar_bits, degree, nq, identity); mirrors the reference's
// dispatch and its error strings (src/laplacian.hpp:361-400).
int stiffness_dispatch(int scalar_bits, int degree, int nq, bool identity, const StiffnessArgs& a);
} // namespace bd
