// Register-blocked ("slab") sum-factorised stiffness action for sm_100a -- the operator's hot kernel.
//
// Same mathematics as stiffness_operator_gpu<T,P,Q> of the reference (src/laplacian_gpu.hpp:91-426)
// -- gather with BC masking, interpolation to the quadrature points, reference gradient, G and
// kappa, transposed gradient, transposed interpolation, scatter-add -- but re-associated so that
// almost every multiply-add takes its matrix entry from the constant bank and its vector entry
// from a register:
//
//   u  --(y,z contractions, registers)-->  a_x = Phi_y Phi_z u,  a_y = DPhi_y Phi_z u,  a_z = Phi_y DPhi_z u
//      --(transpose through shared memory)-->
//      --(x contractions, registers)-->     g_x = DPhi_x a_x,  g_y = Phi_x a_y,  g_z = Phi_x a_z
//   f = kappa G g                           (G streamed from HBM, fully coalesced)
//   b_x = DPhi_x^T f_x, b_y = Phi_x^T f_y, b_z = Phi_x^T f_z   --(transpose back)-->
//   y += Phi_y^T Phi_z^T b_x + DPhi_y^T Phi_z^T b_y + Phi_y^T DPhi_z^T b_z
//
// with Phi = phi0 [nq x nd] and DPhi = dphi1 * phi0 (Phi = I, DPhi = dphi1 for qmode 0).
//
// Execution model: a "team" of NQ lanes owns one cell, a warp owns CW = 32 / NQ cells and walks
// over batches of CW consecutive cells of the list.  In the outer ("X") phases lane i of a team
// holds the (y,z) slab of dof-plane i; in the middle ("Z") phase lane q2 holds the quadrature
// points (., ., q2) and streams over q1.  The two layout changes per direction go through a
// per-warp shared-memory buffer and need only __syncwarp(): there is no CTA-wide barrier in the
// kernel, warps are fully independent.  Shared-memory traffic is ~4x lower than with one thread
// per quadrature point or per column (one LDS per ~NQ FMAs instead of one per FMA).
//
// G is kept in a kernel-friendly order (see GLayout): for a batch b, G[b][q1][q0][comp][c*NQ+q2],
// so that every load of the middle phase is one contiguous run of CW*NQ scalars per warp and the
// whole batch is read front to back exactly once.
#pragma once

#include <cstdint>

#include "pair_fma.cuh"

namespace bd
{
template <typename T, int ND, int NQ>
struct SlabTables
{
  T phi[NQ * ND];  // Phi[q][i]
  T dphi[NQ * ND]; // DPhi[q][i] = sum_r dphi1[q][r] phi0[r][i]
};

// Row padding that makes the X-phase accesses (lanes (c,i) -> i*RL + c*NQ) bank-conflict free.
constexpr int slab_row_pad(int nd, int nq, bool is_double)
{
  const int cw = 32 / nq;
  const int row = cw * nq;
  const int nbanks = is_double ? 16 : 32;
  const int group = is_double ? 16 : 32;
  for (int pad = 0; pad < 33; ++pad)
  {
    const int rl = row + pad;
    bool ok = true;
    for (int g0 = 0; g0 < 32 && ok; g0 += group)
    {
      bool used[32] = {};
      for (int l = g0; l < g0 + group && ok; ++l)
      {
        const int c = l / nq, i = l % nq;
        if (c >= cw || i >= nd)
          continue;
        const int b = (i * rl + c * nq) % nbanks;
        if (used[b])
          ok = false;
        used[b] = true;
      }
    }
    if (ok)
      return pad;
  }
  return 0;
}

template <typename T, int P, int NQ>
struct SlabConfig
{
  static constexpr int ND = P + 1;
  static constexpr int CW = 32 / NQ;   // cells per warp
  static constexpr int ROW = CW * NQ;  // active lanes per warp
  static constexpr int RL = ROW + slab_row_pad(ND, NQ, sizeof(T) == 8);
  static constexpr int CUBE = NQ * ND * RL;             // one of the three exchange cubes, per warp
  static constexpr int WARP_SMEM = 3 * CUBE * int(sizeof(T));
  static constexpr int MAX_SMEM = 220 * 1024;
  static constexpr int WARPS = (MAX_SMEM / WARP_SMEM) >= 4 ? 4 : ((MAX_SMEM / WARP_SMEM) < 1 ? 1 : (MAX_SMEM / WARP_SMEM));
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_SMEM;
  // G layout: per batch NQ (q1) x NQ (q0) x 6 x ROW scalars
  static constexpr std::int64_t G_BATCH = std::int64_t(NQ) * NQ * 6 * ROW;
};

// Position of (list position pos, comp, q0, q1, q2) in the slab G layout (depends on nq only).
__host__ __device__ inline std::int64_t slab_g_index(int nq, std::int64_t pos, int comp, int q0, int q1,
                                                     int q2)
{
  const int cw = 32 / nq, row = cw * nq;
  const std::int64_t batch = pos / cw;
  const int c = static_cast<int>(pos - batch * cw);
  return batch * (static_cast<std::int64_t>(nq) * nq * 6 * row)
         + ((static_cast<std::int64_t>(q1) * nq + q0) * 6 + comp) * row + c * nq + q2;
}
// Scalars to allocate for `ncells` list positions (whole batches).
inline std::int64_t slab_g_size(int nq, std::int64_t ncells)
{
  const int cw = 32 / nq;
  return (ncells + cw - 1) / cw * (static_cast<std::int64_t>(nq) * nq * 6 * cw * nq);
}

// ------------------------------------------------------------------------------------------------
// The three compute phases, shared by the two kernels below.
// Cube element (k3, q1, i, row r): cube[k3*CUBE + (q1*ND + i)*RL + r], r = c*NQ + q2.
// ------------------------------------------------------------------------------------------------

// Forward X phase of lane (c, i): y/z contractions of the dof plane `ul` -> a_x, a_y, a_z in the cube.
template <typename T, int ND, int NQ, int RL, int CUBE, bool IDENTITY>
BD_PHASE_FN void slab_forward_x(const SlabTables<T, ND, NQ>& tab, const T (&ul)[ND * ND],
                                               T* __restrict__ dst0)
{
  // dst0 = cube + i*RL + rbase ; element (q1, q2) at dst0[q1*ND*RL + q2]
#pragma unroll
  for (int q2 = 0; q2 < NQ; ++q2)
  {
    T pz[ND], dz[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      if constexpr (IDENTITY)
      {
        pz[j] = ul[j * ND + q2];
        T s = 0;
#pragma unroll
        for (int k = 0; k < ND; ++k)
          s += tab.dphi[q2 * ND + k] * ul[j * ND + k];
        dz[j] = s;
      }
      else
      {
        T s = 0, d = 0;
#pragma unroll
        for (int k = 0; k < ND; ++k)
        {
          s += tab.phi[q2 * ND + k] * ul[j * ND + k];
          d += tab.dphi[q2 * ND + k] * ul[j * ND + k];
        }
        pz[j] = s;
        dz[j] = d;
      }
    }
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      T ax, ay = 0, az;
      if constexpr (IDENTITY)
      {
        ax = pz[q1];
        az = dz[q1];
#pragma unroll
        for (int j = 0; j < ND; ++j)
          ay += tab.dphi[q1 * ND + j] * pz[j];
      }
      else
      {
        ax = 0;
        az = 0;
#pragma unroll
        for (int j = 0; j < ND; ++j)
        {
          ax += tab.phi[q1 * ND + j] * pz[j];
          ay += tab.dphi[q1 * ND + j] * pz[j];
          az += tab.phi[q1 * ND + j] * dz[j];
        }
      }
      T* dst = dst0 + q1 * (ND * RL) + q2;
      dst[0] = ax;
      dst[CUBE] = ay;
      dst[2 * CUBE] = az;
    }
  }
}

// Variants of the two X phases for large ND (nq = 9): the q2 loop is a real loop and the two table
// rows it needs are fetched from the constant bank with a warp-uniform dynamic index.  Fully unrolled,
// these phases are ~10k instructions each at ND = 8 and ncu shows the warps starving on instruction
// fetch ("no instruction" = 3.4 of 7 stall cycles per issue) on top of register spills.
#ifndef BD_SLAB_ROLLED_FROM_NQ
#define BD_SLAB_ROLLED_FROM_NQ 6
#endif
constexpr int kSlabRolledFromNQ = BD_SLAB_ROLLED_FROM_NQ; // rolled X phases from this nq on

template <typename T, int ND, int NQ, int RL, int CUBE>
__device__ __forceinline__ void slab_forward_x_rolled(const SlabTables<T, ND, NQ>& tab, const T (&ul)[ND * ND],
                                                      T* __restrict__ dst0)
{
#pragma unroll 1
  for (int q2 = 0; q2 < NQ; ++q2)
  {
    T pr[ND], dr[ND];
#pragma unroll
    for (int k = 0; k < ND; ++k)
    {
      pr[k] = tab.phi[q2 * ND + k];
      dr[k] = tab.dphi[q2 * ND + k];
    }
    T pz[ND], dz[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      T s = 0, d = 0;
#pragma unroll
      for (int k = 0; k < ND; ++k)
      {
        s += pr[k] * ul[j * ND + k];
        d += dr[k] * ul[j * ND + k];
      }
      pz[j] = s;
      dz[j] = d;
    }
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      T ax = 0, ay = 0, az = 0;
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        ax += tab.phi[q1 * ND + j] * pz[j];
        ay += tab.dphi[q1 * ND + j] * pz[j];
        az += tab.phi[q1 * ND + j] * dz[j];
      }
      T* dst = dst0 + q1 * (ND * RL) + q2;
      dst[0] = ax;
      dst[CUBE] = ay;
      dst[2 * CUBE] = az;
    }
  }
}

template <typename T, int ND, int NQ, int RL, int CUBE>
__device__ __forceinline__ void slab_backward_x_rolled(const SlabTables<T, ND, NQ>& tab,
                                                       const T* __restrict__ src0, T (&out)[ND * ND])
{
#pragma unroll
  for (int t = 0; t < ND * ND; ++t)
    out[t] = 0;
#pragma unroll 1
  for (int q2 = 0; q2 < NQ; ++q2)
  {
    T e1[ND], e2[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      e1[j] = 0;
      e2[j] = 0;
    }
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      const T* src = src0 + q1 * (ND * RL) + q2;
      const T bx = src[0], by = src[CUBE], bz = src[2 * CUBE];
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        e1[j] += tab.phi[q1 * ND + j] * bx;
        e1[j] += tab.dphi[q1 * ND + j] * by;
        e2[j] += tab.phi[q1 * ND + j] * bz;
      }
    }
    T pr[ND], dr[ND];
#pragma unroll
    for (int k = 0; k < ND; ++k)
    {
      pr[k] = tab.phi[q2 * ND + k];
      dr[k] = tab.dphi[q2 * ND + k];
    }
#pragma unroll
    for (int j = 0; j < ND; ++j)
#pragma unroll
      for (int k = 0; k < ND; ++k)
      {
        out[j * ND + k] += pr[k] * e1[j];
        out[j * ND + k] += dr[k] * e2[j];
      }
  }
}

// Z phase of lane (c, q2) for one q1 and the q0 range [Q0B, Q0B + Q0N): x contractions, G, kappa,
// transposed x contractions accumulated into bx, by, bz.  G values come from `g` (global or shared
// memory), element (q0 - Q0B, comp) at g[((q0 - Q0B)*6 + comp)*ROW].
// Q0B is an ordinary argument, but every call site sits in a fully unrolled loop, so after inlining
// all table indices are compile-time constants.
template <typename T, int ND, int NQ, int ROW, bool IDENTITY, int Q0N, bool STREAM_LDG, bool SCALE = true>
BD_PHASE_FN void slab_z_points(const SlabTables<T, ND, NQ>& tab, const T (&ax)[ND],
                                              const T (&ay)[ND], const T (&az)[ND], T (&bx)[ND],
                                              T (&by)[ND], T (&bz)[ND], const T* __restrict__ g, T kappa,
                                              const int Q0B)
{
#pragma unroll
  for (int qq = 0; qq < Q0N; ++qq)
  {
    constexpr int dummy = 0;
    (void)dummy;
    const int q0 = Q0B + qq;
    const T* gp = g + qq * (6 * ROW);
    T G0, G1, G2, G3, G4, G5;
    if constexpr (STREAM_LDG)
    {
      G0 = __ldg(gp + 0 * ROW);
      G1 = __ldg(gp + 1 * ROW);
      G2 = __ldg(gp + 2 * ROW);
      G3 = __ldg(gp + 3 * ROW);
      G4 = __ldg(gp + 4 * ROW);
      G5 = __ldg(gp + 5 * ROW);
    }
    else
    {
      G0 = gp[0 * ROW];
      G1 = gp[1 * ROW];
      G2 = gp[2 * ROW];
      G3 = gp[3 * ROW];
      G4 = gp[4 * ROW];
      G5 = gp[5 * ROW];
    }
    T gx = 0, gy, gz;
    if constexpr (IDENTITY)
    {
      gy = ay[q0];
      gz = az[q0];
#pragma unroll
      for (int i = 0; i < ND; ++i)
        gx += tab.dphi[q0 * ND + i] * ax[i];
    }
    else
    {
      gy = 0;
      gz = 0;
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        gx += tab.dphi[q0 * ND + i] * ax[i];
        gy += tab.phi[q0 * ND + i] * ay[i];
        gz += tab.phi[q0 * ND + i] * az[i];
      }
    }
    // reference src/laplacian_gpu.hpp:269-273
    // SCALE = false: the caller has already multiplied the cell's input by kappa (the operator is
    // linear in kappa), which saves three multiplies per quadrature point
    T fx = G0 * gx + G1 * gy + G2 * gz;
    T fy = G1 * gx + G3 * gy + G4 * gz;
    T fz = G2 * gx + G4 * gy + G5 * gz;
    if constexpr (SCALE)
    {
      fx *= kappa;
      fy *= kappa;
      fz *= kappa;
    }
    if constexpr (IDENTITY)
    {
      by[q0] = fy;
      bz[q0] = fz;
#pragma unroll
      for (int i = 0; i < ND; ++i)
        bx[i] += tab.dphi[q0 * ND + i] * fx;
    }
    else
    {
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        bx[i] += tab.dphi[q0 * ND + i] * fx;
        by[i] += tab.phi[q0 * ND + i] * fy;
        bz[i] += tab.phi[q0 * ND + i] * fz;
      }
    }
  }
}

// Backward X phase of lane (c, i): transposed y/z contractions of b_x, b_y, b_z -> out (dof plane i).
template <typename T, int ND, int NQ, int RL, int CUBE, bool IDENTITY>
BD_PHASE_FN void slab_backward_x(const SlabTables<T, ND, NQ>& tab, const T* __restrict__ src0,
                                                T (&out)[ND * ND])
{
#pragma unroll
  for (int t = 0; t < ND * ND; ++t)
    out[t] = 0;
#pragma unroll
  for (int q2 = 0; q2 < NQ; ++q2)
  {
    T e1[ND], e2[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      e1[j] = 0;
      e2[j] = 0;
    }
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      const T* src = src0 + q1 * (ND * RL) + q2;
      const T bx = src[0], by = src[CUBE], bz = src[2 * CUBE];
      if constexpr (IDENTITY)
      {
        e1[q1] += bx;
        e2[q1] += bz;
#pragma unroll
        for (int j = 0; j < ND; ++j)
          e1[j] += tab.dphi[q1 * ND + j] * by;
      }
      else
      {
#pragma unroll
        for (int j = 0; j < ND; ++j)
        {
          e1[j] += tab.phi[q1 * ND + j] * bx;
          e1[j] += tab.dphi[q1 * ND + j] * by;
          e2[j] += tab.phi[q1 * ND + j] * bz;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      if constexpr (IDENTITY)
      {
        out[j * ND + q2] += e1[j];
#pragma unroll
        for (int k = 0; k < ND; ++k)
          out[j * ND + k] += tab.dphi[q2 * ND + k] * e2[j];
      }
      else
      {
#pragma unroll
        for (int k = 0; k < ND; ++k)
        {
          out[j * ND + k] += tab.phi[q2 * ND + k] * e1[j];
          out[j * ND + k] += tab.dphi[q2 * ND + k] * e2[j];
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Plain kernel: G loaded straight from global memory (kept as the fallback for configurations the
// asynchronous-copy kernel does not cover, and for comparison: BD_STIFFNESS_VARIANT=slab).
// ------------------------------------------------------------------------------------------------
template <typename T, int P, int NQ, bool IDENTITY>
__global__ void __launch_bounds__(SlabConfig<T, P, NQ>::THREADS)
    stiffness_slab_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ Gs,
                          const T* __restrict__ cell_constants,
                          const std::int32_t* __restrict__ dofmap,
                          const std::int32_t* __restrict__ cells, int ncells,
                          const std::int8_t* __restrict__ bc,
                          const __grid_constant__ SlabTables<T, P + 1, NQ> tab)
{
  using C = SlabConfig<T, P, NQ>;
  constexpr int ND = C::ND, CW = C::CW, RL = C::RL, CUBE = C::CUBE;
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  T* smem = reinterpret_cast<T*>(smem_raw);

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / NQ;        // cell within the warp's batch
  const int m = lane - c * NQ;    // team member: dof plane i (X phases) / point plane q2 (Z phase)
  const bool lane_ok = c < CW;
  T* cube = smem + warp * (3 * CUBE);
  const int rbase = (lane_ok ? c : 0) * NQ;

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * C::WARPS;
  for (int batch = blockIdx.x * C::WARPS + warp; batch < nbatches; batch += warps_total)
  {
    const int pos = batch * CW + c;
    const bool active = lane_ok && pos < ncells;
    const bool xlane = active && m < ND; // lanes that own a dof plane

    // gather (reference :153-170)
    std::int32_t dof[ND2];
    T ul[ND2];
    if (xlane)
    {
      const std::int32_t* row = dofmap + static_cast<std::int64_t>(cells[pos]) * ND3 + m * ND2;
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        dof[t] = row[t];
      std::int8_t flag[ND2];
#pragma unroll
      for (int t = 0; t < ND2; ++t) // loads first, stores after (see stiffness.cuh)
      {
        ul[t] = u[dof[t]];
        flag[t] = bc[dof[t]];
      }
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        if (flag[t])
        {
          y[dof[t]] = ul[t]; // b[dof] = u[dof] on constrained dofs
          dof[t] = -1;
          ul[t] = 0;
        }
      slab_forward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab, ul, cube + m * RL + rbase);
    }
    __syncwarp();

    if (active)
    {
      const T kappa = cell_constants[pos];
      const T* Gb = Gs + static_cast<std::int64_t>(batch) * C::G_BATCH + rbase + m;
      // fully unrolled for small NQ; a real loop otherwise (code size), only addresses depend on q1
#pragma unroll(NQ <= 5 ? NQ : 1)
      for (int q1 = 0; q1 < NQ; ++q1)
      {
        T ax[ND], ay[ND], az[ND], bx[ND], by[ND], bz[ND];
        T* src = cube + (q1 * ND) * RL + rbase + m;
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          ax[i] = src[i * RL];
          ay[i] = src[i * RL + CUBE];
          az[i] = src[i * RL + 2 * CUBE];
          bx[i] = 0;
          by[i] = 0;
          bz[i] = 0;
        }
        slab_z_points<T, ND, NQ, C::ROW, IDENTITY, NQ, true>(
            tab, ax, ay, az, bx, by, bz, Gb + static_cast<std::int64_t>(q1) * (NQ * 6 * C::ROW), kappa, 0);
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

    if (xlane)
    {
      T out[ND2];
      slab_backward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab, cube + m * RL + rbase, out);
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        if (dof[t] >= 0)
          atomicAdd(y + dof[t], out[t]); // result unused -> RED.E.ADD (reference :424-425)
    }
    __syncwarp(); // the cube is rewritten by the next batch's forward phase
  }
}

// ------------------------------------------------------------------------------------------------
// Asynchronous-copy kernel: the G stream is decoupled from the math.
//
// ncu on the plain kernel (profiles/r1_v2_*) shows it purely latency bound: 11.6 of 14.9 cycles
// between issues are "long scoreboard" (waiting for G / gather loads) at 12 warps per SM.  Here
// every warp keeps a ring of NSTAGE shared-memory slots and one elected lane streams the batch's G
// through it with 1-D bulk asynchronous copies (cp.async.bulk -> UBLKCP, completion on an
// mbarrier).  One slot = Q0C consecutive q0 rows of one q1 slice (contiguous in the slab layout);
// NSTAGE-1 slots are always in flight and the stream runs on across batch boundaries, so the next
// batch's first slots arrive while the current batch is still in its backward phase.
//
// The kernel reads a *masked dofmap*: the cell dofmap with the BC marker of each dof folded into
// bit 31 (built once per operator), which removes the dependent byte gather bc[dof] from the hot
// loop.  With PREFETCH the dof indices and input values of the next batch are loaded into
// registers one phase ahead; without it (large ND, register budget) the indices are simply
// re-read before the scatter (they are L1/L2 hits).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ std::uint32_t smem_addr(const void* p)
{
  return static_cast<std::uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(std::uint64_t* bar, std::uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(std::uint64_t* bar, std::uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void* dst, const void* src, std::uint32_t bytes,
                                              std::uint64_t* bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar))
               : "memory");
}
// G is read exactly once per operator application: mark its lines evict-first in L2 so that the
// 126 MB L2 keeps the re-used data (input/output vector planes shared by neighbouring cells, dofmap).
__device__ __forceinline__ std::uint64_t l2_evict_first_policy()
{
  std::uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_copy_g2s_hint(void* dst, const void* src, std::uint32_t bytes,
                                                   std::uint64_t* bar, std::uint64_t policy)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, "
               "[%3], %4;" ::"r"(smem_addr(dst)),
               "l"(src), "r"(bytes), "r"(smem_addr(bar)), "l"(policy)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(std::uint64_t* bar, std::uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred P1;\n"
               "LAB_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
               "@P1 bra DONE;\n"
               "bra LAB_WAIT;\n"
               "DONE:\n"
               "}" ::"r"(smem_addr(bar)),
               "r"(parity)
               : "memory");
}

constexpr std::int32_t kBcBit = static_cast<std::int32_t>(0x80000000u);

template <typename T, int P, int NQ, int NSTAGE, int Q0C, int WARPS_>
struct SlabTmaConfig
{
  using Base = SlabConfig<T, P, NQ>;
  static_assert(NQ % Q0C == 0, "Q0C must divide NQ");
  static constexpr int NCH = NQ / Q0C;                            // slots per q1 slice
  static constexpr int SLOT = Q0C * 6 * Base::ROW;                // scalars per slot
  static constexpr int SLOT_BYTES = SLOT * int(sizeof(T));
  static constexpr int CUBE_BYTES = (3 * Base::CUBE * int(sizeof(T)) + 15) / 16 * 16;
  static constexpr int WARP_SMEM = CUBE_BYTES + NSTAGE * SLOT_BYTES + 64; // + mbarriers
  static constexpr int MAX_SMEM = 226 * 1024;
  static constexpr int WARPS = WARPS_;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_SMEM;
  static constexpr bool SUPPORTED = (SLOT_BYTES % 16 == 0) && (CTA_SMEM <= MAX_SMEM);
};

template <typename T, int P, int NQ, bool IDENTITY, int NSTAGE, int Q0C, int WARPS_, bool PREFETCH>
__global__ void __launch_bounds__(WARPS_ * 32)
    stiffness_slab_tma_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ Gs,
                              const T* __restrict__ cell_constants,
                              const std::int32_t* __restrict__ mdofmap,
                              const std::int32_t* __restrict__ cells, int ncells,
                              const __grid_constant__ SlabTables<T, P + 1, NQ> tab)
{
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
      if constexpr (NQ >= kSlabRolledFromNQ && !IDENTITY)
        slab_forward_x_rolled<T, ND, NQ, RL, CUBE>(tab, ul, cube + m * RL + rbase);
      else
        slab_forward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab, ul, cube + m * RL + rbase);
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
          slab_z_points<T, ND, NQ, ROW, IDENTITY, Q0C, false, false>(tab, ax, ay, az, bx, by, bz, gs, kappa,
                                                                     ch * Q0C);
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
      if constexpr (NQ >= kSlabRolledFromNQ && !IDENTITY)
        slab_backward_x_rolled<T, ND, NQ, RL, CUBE>(tab, cube + m * RL + rbase, out);
      else
        slab_backward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab, cube + m * RL + rbase, out);
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

// Masked dofmap: bit 31 of every entry = BC marker of that dof (setup, once per operator).
__global__ void __launch_bounds__(256)
    mask_dofmap_kernel(std::int64_t n, const std::int32_t* __restrict__ dofmap,
                       const std::int8_t* __restrict__ bc, std::int32_t* __restrict__ out);

// Slab layout -> reference layout G[pos][comp][nq^3] (tests / bd_laplacian_geometry only).
template <typename T>
__global__ void __launch_bounds__(256)
    slab_g_to_reference_kernel(int nq, const T* __restrict__ Gs, T* __restrict__ G, std::int64_t ncells)
{
  const int nq3 = nq * nq * nq;
  const std::int64_t total = ncells * 6 * nq3;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
       idx < total; idx += stride)
  {
    const std::int64_t pos = idx / (6 * nq3);
    const int rem = static_cast<int>(idx - pos * (6 * nq3));
    const int comp = rem / nq3, q = rem - comp * nq3;
    const int q0 = q / (nq * nq), q1 = (q / nq) % nq, q2 = q % nq;
    G[idx] = Gs[slab_g_index(nq, pos, comp, q0, q1, q2)];
  }
}

} // namespace bd
