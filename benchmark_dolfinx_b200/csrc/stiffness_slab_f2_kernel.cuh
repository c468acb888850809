// The asynchronous-copy slab kernel (stiffness_slab_tma_kernel, stiffness_slab.cuh) with its three compute
// phases on FFMA2 (stiffness_slab_f2.cuh): fp32, qmode 1, nq <= 5.  Everything outside the three phase
// calls -- the UBLKCP ring for G, the masked dofmap, the prefetches, the scatter -- is the same code as in
// stiffness_slab_tma_kernel (kept as a separate kernel so that the default kernels' code does not change);
// selected with BD_SLAB_F2=1.
#pragma once

#include "stiffness_slab.cuh"

namespace bd
{
template <int P, int NQ, int NSTAGE, int WARPS_, bool PREFETCH>
__global__ void __launch_bounds__(WARPS_ * 32, 16 / WARPS_) // <= 128 registers, as the scalar kernel uses
    stiffness_slab_tma_f2_kernel(const float* __restrict__ u, float* __restrict__ y,
                                 const float* __restrict__ Gs, const float* __restrict__ cell_constants,
                                 const std::int32_t* __restrict__ mdofmap,
                                 const std::int32_t* __restrict__ cells, int ncells,
                                 const __grid_constant__ SlabPairTables<P + 1, NQ> ptab)
{
  using T = float;
  constexpr int Q0C = NQ; // one ring slot = one q1 slice
  static_assert(NQ < kSlabRolledFromNQ, "FFMA2 phases exist for the unrolled outer phases only");
  using C = SlabConfig<T, P, NQ>;
  using TC = SlabTmaConfig<T, P, NQ, NSTAGE, Q0C, WARPS_>;
  constexpr int ND = C::ND, CW = C::CW, RL = C::RL, CUBE = C::CUBE, ROW = C::ROW;
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  constexpr int NCH = TC::NCH;
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / NQ;
  const int m = lane - c * NQ;
  const bool lane_ok = c < CW;
  unsigned char* wbase = smem_raw + warp * TC::WARP_SMEM;
  T* cube = reinterpret_cast<T*>(wbase);
  T* ring = reinterpret_cast<T*>(wbase + TC::CUBE_BYTES);
  std::uint64_t* bars = reinterpret_cast<std::uint64_t*>(wbase + TC::CUBE_BYTES + NSTAGE * TC::SLOT_BYTES);
  const int rbase = (lane_ok ? c : 0) * NQ;

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * TC::WARPS;
  const int first = blockIdx.x * TC::WARPS + warp;
  const int my_batches = first < nbatches ? (nbatches - first + warps_total - 1) / warps_total : 0;
  const int total_slots = my_batches * NQ * NCH;

  if (lane == 0)
  {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s)
      mbar_init(bars + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  const std::uint64_t g_policy = l2_evict_first_policy();
  // slot s of this warp's stream = (batch first + k*warps_total, q1, chunk), s = (k*NQ + q1)*NCH + chunk;
  // within a batch the slots are consecutive in memory
  auto issue = [&](int s) {
    const int k = s / (NQ * NCH), r = s - k * (NQ * NCH);
    const std::int64_t batch = first + static_cast<std::int64_t>(k) * warps_total;
    const T* src = Gs + batch * C::G_BATCH + static_cast<std::int64_t>(r) * TC::SLOT;
    const int st = s % NSTAGE;
    mbar_expect_tx(bars + st, TC::SLOT_BYTES);
    bulk_copy_g2s_hint(ring + st * TC::SLOT, src, TC::SLOT_BYTES, bars + st, g_policy);
  };
  if (lane == 0)
  {
    for (int s = 0; s < NSTAGE - 1 && s < total_slots; ++s)
      issue(s);
  }

  // Cell index and kappa of a batch are loaded a whole batch ahead: ncu showed the warp stalling in
  // order on cells[pos] (address of the dof row) and on cell_constants[pos] right at their first use.
  auto cell_of = [&](int batch) -> std::int32_t {
    const int pos = batch * CW + c;
    const bool x = lane_ok && batch < nbatches && pos < ncells && m < ND;
    return x ? cells[pos] : -1;
  };
  auto kappa_of = [&](int batch) -> T {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cell_constants[pos] : T(0);
  };
  auto row_from = [&](std::int32_t cell) -> const std::int32_t* {
    return cell >= 0 ? mdofmap + static_cast<std::int64_t>(cell) * ND3 + m * ND2 : nullptr;
  };
  auto load_dofs = [&](const std::int32_t* row, std::int32_t (&d)[ND2]) {
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      d[t] = row ? row[t] : kBcBit; // inactive lanes look like constrained dofs: no gather, no scatter
  };
  auto load_values = [&](const std::int32_t* row, const std::int32_t (&d)[ND2], T (&v)[ND2]) {
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      v[t] = row ? u[d[t] & 0x7fffffff] : T(0);
  };

  std::int32_t dof[ND2];
  T ul[ND2];
  std::int32_t cell_cur = my_batches > 0 ? cell_of(first) : -1;
  std::int32_t cell_next = my_batches > 1 ? cell_of(first + warps_total) : -1;
  T kappa = my_batches > 0 ? kappa_of(first) : T(0);
  if constexpr (PREFETCH)
  {
    const std::int32_t* row = row_from(cell_cur);
    load_dofs(row, dof);
    load_values(row, dof, ul);
  }

  int slot = 0;
  for (int k = 0; k < my_batches; ++k)
  {
    const int batch = first + k * warps_total;
    const int pos = batch * CW + c;
    const bool active = lane_ok && pos < ncells;
    const bool xlane = active && m < ND;
    const std::int32_t* row = row_from(cell_cur);
    // two batches ahead / one batch ahead: consumed at the end of this iteration
    const std::int32_t cell_next2 = k + 2 < my_batches ? cell_of(batch + 2 * warps_total) : -1;
    const T kappa_next = k + 1 < my_batches ? kappa_of(batch + warps_total) : T(0);

    if constexpr (!PREFETCH)
    {
      load_dofs(row, dof);
      load_values(row, dof, ul);
    }
    // BC masking (reference :163-169): b[dof] = u[dof], the dof contributes 0 and receives nothing
    if (xlane)
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        if (dof[t] < 0)
        {
          y[dof[t] & 0x7fffffff] = ul[t];
          ul[t] = 0;
        }
        ul[t] *= kappa; // kappa folded into the input: y = A(kappa) u = A(1) (kappa u)
      }
      slab_forward_x_f2<ND, NQ, RL, CUBE>(ptab, ul, cube + m * RL + rbase);
    }
    __syncwarp();

    // next batch's dof indices: in flight during the whole middle phase
    std::int32_t ndof[PREFETCH ? ND2 : 1];
    const std::int32_t* nrow = nullptr;
    if constexpr (PREFETCH)
    {
      nrow = row_from(cell_next);
      load_dofs(nrow, ndof);
    }

    // Middle phase.  When the padded row is at least 32 wide every lane runs it without a branch:
    // the idle lanes of the warp (and teams past the end of the list) work on row entries nobody
    // else uses -- lanes >= ROW sit in the padding columns -- and their results are never scattered.
    constexpr bool ALL_LANES = RL >= 32;
    const int zr = ALL_LANES ? lane : rbase + m;
    const bool zactive = ALL_LANES || active;
#pragma unroll 1
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      T ax[ND], ay[ND], az[ND], bx[ND], by[ND], bz[ND];
      T* src = cube + (q1 * ND) * RL + zr;
      if (zactive)
      {
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          ax[i] = src[i * RL];
          ay[i] = src[i * RL + CUBE];
          az[i] = src[i * RL + 2 * CUBE];
        }
      }
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        bx[i] = 0;
        by[i] = 0;
        bz[i] = 0;
      }
#pragma unroll
      for (int ch = 0; ch < NCH; ++ch, ++slot)
      {
        // refill the slot consumed one step ago with the slot NSTAGE-1 ahead in the stream
        if (lane == 0 && slot + NSTAGE - 1 < total_slots)
          issue(slot + NSTAGE - 1);
        const int st = slot % NSTAGE;
        mbar_wait(bars + st, (slot / NSTAGE) & 1);
        if (zactive)
        {
          const T* gs = ring + st * TC::SLOT + zr;
          slab_z_points_f2<ND, NQ, ROW>(ptab, ax, ay, az, bx, by, bz, gs);
        }
        __syncwarp(); // every lane is done with slot `st` before it is refilled
      }
      if (zactive)
      {
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          src[i * RL] = bx[i];
          src[i * RL + CUBE] = by[i];
          src[i * RL + 2 * CUBE] = bz[i];
        }
      }
    }
    __syncwarp();

    // next batch's input values: in flight during the backward phase
    T nul[PREFETCH ? ND2 : 1];
    if constexpr (PREFETCH)
      load_values(nrow, ndof, nul);

    if (xlane)
    {
      T out[ND2];
      slab_backward_x_f2<ND, NQ, RL, CUBE>(ptab, cube + m * RL + rbase, out);
      if constexpr (PREFETCH)
      {
#pragma unroll
        for (int t = 0; t < ND2; ++t)
          if (dof[t] >= 0)
            atomicAdd(y + dof[t], out[t]); // result unused -> RED.E.ADD (reference :424-425)
      }
      else
      {
#pragma unroll
        for (int t = 0; t < ND2; ++t)
        {
          const std::int32_t d = row[t]; // re-read (cache hit) instead of holding ND2 registers
          if (d >= 0)
            atomicAdd(y + d, out[t]);
        }
      }
    }
    __syncwarp(); // the cube is rewritten by the next batch's forward phase
    if constexpr (PREFETCH)
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        dof[t] = ndof[t];
        ul[t] = nul[t];
      }
    }
    cell_cur = cell_next;
    cell_next = cell_next2;
    kappa = kappa_next;
  }
}
} // namespace bd
