// "Pencil" sum-factorised stiffness action for sm_100a: the operator kernel for the quadrature
// sizes where neither the register slabs (stiffness_slab.cuh) nor the DMMA tiles
// (stiffness_mma8.cuh) fit -- nq = 9 (degree 7, qmode 1) in fp64 and nq = 8 in fp32 arithmetic.
//
// Same mathematics as stiffness_operator_gpu<T,P,Q> of the reference (src/laplacian_gpu.hpp:91-426)
// with the re-association of stiffness_slab.cuh (a_x = Phi_y Phi_z u, a_y = DPhi_y Phi_z u,
// a_z = Phi_y DPhi_z u, then the x contractions, G, and the same chain transposed).  What differs
// is the decomposition: NQ*NQ threads own one cell and a thread always holds one *pencil* (a line
// of nd or nq values along the direction being contracted) in registers:
//
//   phase 1  thread (i, j)    u[i][j][:]              -> Pz, Dz [i][j][qz]          (z contraction)
//   phase 2  thread (i, qz)   Pz, Dz [i][:][qz]       -> A, B, C [i][qy][qz]        (y contraction)
//   phase 3  thread (qy, qz)  A, B, C [:][qy][qz]     -> gradient at the nq points (qx, qy, qz),
//                             G (streamed from HBM, one coalesced row of nq^2 scalars per
//                             component and qx), transposed x contraction -> A', B', C'
//   phase 4  thread (i, qz)   A', B', C' [i][:][qz]   -> Pz', Dz' [i][j][qz]        (y transposed)
//   phase 5  thread (i, j)    Pz', Dz' [i][j][:]      -> y[i][j][:], scatter-add    (z transposed)
//
// Every multiply-add takes its matrix entry from the constant bank and its vector entry from a
// register; shared memory is touched once per value and phase (one LDS/STS per ~nd FMAs), and the
// per-thread state is 3 pencils in and 3 out (no spills at nd = 8 in fp64, where a whole dof plane
// per thread -- the slab kernel -- needs 128 registers for the plane alone).  The price is four
// CTA barriers per cell; the phase 1/5 and 2/4 pairs use the same thread <-> location mapping, so
// the buffers can be reused across cells without further barriers.
//
// G is read in the reference layout [pos][comp][q0][q1][q2] (src/geometry_gpu.hpp:24-119): every
// thread copies the six entries of its own point of a q0 row with cp.async (L2 evict-first hint)
// into a private slot of a shared-memory ring of RD rows, RD - 1 rows ahead of their use.  The rows
// of all the cells a CTA visits form one stream, so with RD = NQ + 1 a row is requested a whole
// cell before it is needed (while the previous cell is still in phase 3) and the HBM latency is
// covered by phases 4, 5, 1 and 2.  Nobody else reads a thread's slots: the ring needs no barrier,
// only cp.async.wait_group.
//
// fp64: the outer loop of every phase is rolled.  DFMA cannot take a constant-bank operand on
// sm_100 (the table entry comes through a uniform register, one LDCU per use); fully unrolled, the
// compiler parks table entries in ordinary registers and spills.  fp32 FFMA reads c[][] directly,
// so there everything is unrolled.
#pragma once

#include <cstdint>

#include "stiffness_slab.cuh"

namespace bd
{
// Strides (in scalars) of the phase 1/2 exchange buffer S1[v][i][j][qz] = v*ND*a + i*a + j*b + qz,
// chosen offline (brute force over paddings) so that both the (i,j)-lane and the (i,qz)-lane access
// patterns are free of bank conflicts (or as close as a linear layout gets).
struct PencilStrides
{
  int a, b;
};
constexpr PencilStrides pencil_s1_strides(int nd, int nq, bool is_double)
{
  if (is_double)
  {
    if (nd == 8 && nq == 9) return {89, 10};
    if (nd == 7 && nq == 8) return {63, 9};
    if (nd == 6 && nq == 7) return {42, 7};
    if (nd == 5 && nq == 6) return {38, 7};
    if (nd == 4 && nq == 5) return {20, 5};
    if (nd == 8 && nq == 8) return {72, 9};
    if (nd == 7 && nq == 7) return {63, 9};
    if (nd == 6 && nq == 6) return {54, 9};
  }
  else
  {
    if (nd == 8 && nq == 9) return {105, 12};
    if (nd == 7 && nq == 8) return {72, 9};
    if (nd == 6 && nq == 7) return {42, 7};
    if (nd == 5 && nq == 6) return {45, 9};
    if (nd == 4 && nq == 5) return {24, 5};
    if (nd == 8 && nq == 8) return {72, 9};
    if (nd == 7 && nq == 7) return {71, 10};
    if (nd == 6 && nq == 6) return {42, 7};
  }
  const int b = nq | 1;
  return {nd * b, b};
}

template <typename T, int P, int NQ, int CPB, int RD_>
struct PencilConfig
{
  static constexpr int RD = RD_; // rows of G in the ring (2 .. NQ + 1)
  static constexpr int ND = P + 1;
  static constexpr int TPC = NQ * NQ; // threads per cell
  static constexpr int THREADS = (CPB * TPC + 31) / 32 * 32;
  static constexpr int BANKS = sizeof(T) == 8 ? 16 : 32;
  static constexpr PencilStrides S1 = pencil_s1_strides(ND, NQ, sizeof(T) == 8);
  static constexpr int S1I = S1.a;      // stride of i
  static constexpr int S1J = S1.b;      // stride of j
  static constexpr int S1V = ND * S1I;  // one of the two S1 arrays (Pz, Dz)
  // S2[v][i][qy*NQ + qz]: the (qy,qz) lanes are contiguous; the (i,qz) lanes see i*S2I + qz, which is
  // conflict free when S2I = NQ (mod banks)
  static constexpr int S2I = TPC + ((NQ - TPC % BANKS) % BANKS + BANKS) % BANKS;
  static constexpr int S2V = ND * S2I;
  static constexpr int RING = RD * 6 * TPC; // G ring: [slot][comp][qy*NQ + qz]
  static constexpr int CELL_SCALARS = (2 * S1V + 3 * S2V + RING + 3) / 4 * 4;
  static constexpr int CTA_SMEM = CPB * CELL_SCALARS * int(sizeof(T));
  static constexpr bool SUPPORTED = CTA_SMEM <= 226 * 1024 && THREADS <= 1024 && RD >= 2 && RD <= NQ + 1;
};

// One G entry (read once per operator application: evict-first in L2) from global to shared memory.
template <typename T>
__device__ __forceinline__ void cp_async_scalar(std::uint32_t dst_smem, const T* src, std::uint64_t policy)
{
  static_assert(sizeof(T) == 4 || sizeof(T) == 8, "cp.async.ca copies 4, 8 or 16 bytes");
  if constexpr (sizeof(T) == 8)
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"(dst_smem), "l"(src),
                 "l"(policy)
                 : "memory");
  else
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(dst_smem), "l"(src),
                 "l"(policy)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait()
{
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// ND consecutive dof indices of one (i, j) pencil, vectorised where the row alignment allows.
template <int ND>
__device__ __forceinline__ void pencil_load_dofs(const std::int32_t* row, std::int32_t (&d)[ND])
{
  if (!row)
  {
#pragma unroll
    for (int k = 0; k < ND; ++k)
      d[k] = kBcBit; // inactive pencil: looks constrained, is never gathered or scattered
    return;
  }
  if constexpr (ND % 4 == 0)
  {
#pragma unroll
    for (int v = 0; v < ND / 4; ++v)
    {
      const int4 q = __ldg(reinterpret_cast<const int4*>(row) + v);
      d[4 * v] = q.x, d[4 * v + 1] = q.y, d[4 * v + 2] = q.z, d[4 * v + 3] = q.w;
    }
  }
  else if constexpr (ND % 2 == 0)
  {
#pragma unroll
    for (int v = 0; v < ND / 2; ++v)
    {
      const int2 q = __ldg(reinterpret_cast<const int2*>(row) + v);
      d[2 * v] = q.x, d[2 * v + 1] = q.y;
    }
  }
  else
  {
#pragma unroll
    for (int k = 0; k < ND; ++k)
      d[k] = __ldg(row + k);
  }
}

// fp32 tables for the FFMA2 variant (pair_fma, stiffness_slab_f2.cuh: a scalar from a register times a
// pair of table entries from a uniform register, added to a register pair).  Forward phases pair two
// neighbouring quadrature points, backward phases two neighbouring dofs; odd extents are zero padded.
template <int ND, int NQ>
struct PencilPairTables
{
  static constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
  float2 fphi[ND * NQP];  // [k][qp] = (Phi[2qp][k],  Phi[2qp+1][k])
  float2 fdphi[ND * NQP]; // [k][qp] = (DPhi[2qp][k], DPhi[2qp+1][k])
  float2 bphi[NQ * NDP];  // [q][kp] = (Phi[q][2kp],  Phi[q][2kp+1])
  float2 bdphi[NQ * NDP]; // [q][kp] = (DPhi[q][2kp], DPhi[q][2kp+1])
};
// sum_k row[k] * v[k].  fp64: two partial sums -- the rolled loops leave only a few independent DFMA
// chains per warp, and with 9-12 warps per SM their latency shows (ncu: "wait" is the top stall).
template <typename T, int N>
__device__ __forceinline__ T pencil_dot(const T* __restrict__ row, const T (&v)[N])
{
  if constexpr (sizeof(T) == 8 && N >= 4)
  {
    T s0 = row[0] * v[0], s1 = row[1] * v[1];
#pragma unroll
    for (int k = 2; k + 1 < N; k += 2)
    {
      s0 += row[k] * v[k];
      s1 += row[k + 1] * v[k + 1];
    }
    if constexpr (N % 2 == 1)
      s0 += row[N - 1] * v[N - 1];
    return s0 + s1;
  }
  else
  {
    T s = 0;
#pragma unroll
    for (int k = 0; k < N; ++k)
      s += row[k] * v[k];
    return s;
  }
}

// Minimum CTAs per SM for __launch_bounds__.  Without it the compiler spends registers on scheduling
// freedom and halves the occupancy.  fp64 (one cell of nq = 9, 96 threads): 4 CTAs = 12 warps, <= 168
// registers; fp32: <= 128 registers, as far as shared memory lets that many CTAs in.
template <typename T, int P, int NQ, int CPB, int RD>
constexpr int pencil_min_ctas()
{
  using C = PencilConfig<T, P, NQ, CPB, RD>;
  const int by_regs = sizeof(T) == 8 ? 65536 / (168 * C::THREADS) : 65536 / (128 * C::THREADS);
  const int by_smem = (220 * 1024) / (C::CTA_SMEM + 1024);
  const int n = by_regs < by_smem ? by_regs : by_smem;
  return n < 1 ? 1 : n;
}

// STREAM: the G rows of consecutive cells form one stream through the ring (rows are requested across
// the cell boundary); otherwise the ring restarts with every cell (rows 0 .. RD-2 requested at the top of
// the cell), which needs fewer registers.
// PAIRED (fp32 only): the contractions run on FFMA2 with the tables of `ptab`.
template <typename T, int P, int NQ, int CPB, int RD, bool UPF, bool STREAM, bool PAIRED>
__global__ void __launch_bounds__(PencilConfig<T, P, NQ, CPB, RD>::THREADS, pencil_min_ctas<T, P, NQ, CPB, RD>())
    stiffness_pencil_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ G,
                            const T* __restrict__ cell_constants,
                            const std::int32_t* __restrict__ mdofmap,
                            const std::int32_t* __restrict__ cells, int ncells,
                            const __grid_constant__ SlabTables<T, P + 1, NQ> tab,
                            const __grid_constant__ PencilPairTables<PAIRED ? P + 1 : 1, PAIRED ? NQ : 1> ptab)
{
  static_assert(!PAIRED || sizeof(T) == 4, "FFMA2 is an fp32 instruction");
  using C = PencilConfig<T, P, NQ, CPB, RD>;
  constexpr int NQP = (NQ + 1) / 2, NDP = (P + 2) / 2;
  constexpr int ND = C::ND, TPC = C::TPC, NQ3 = TPC * NQ, ND2 = ND * ND, ND3 = ND2 * ND;
  constexpr int S1I = C::S1I, S1J = C::S1J, S1V = C::S1V, S2I = C::S2I, S2V = C::S2V;
  constexpr int OUTER = sizeof(T) == 8 ? 1 : NQ; // unroll factor of the outer loop of each phase
  static_assert(RD >= 2 && RD <= NQ + 1 && (STREAM || RD <= NQ), "ring depth");
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int tid = threadIdx.x;
  const int c = tid / TPC; // cell slot of this thread
  const int rem = tid - c * TPC;
  const bool slot_ok = c < CPB;
  T* S1 = reinterpret_cast<T*>(smem_raw) + (slot_ok ? c : 0) * C::CELL_SCALARS;
  T* S2 = S1 + 2 * S1V;
  // phases 1 and 5: pencil (i1, j1, :) of the dofs
  const bool role1 = slot_ok && rem < ND2;
  const int i1 = role1 ? rem / ND : 0, j1 = role1 ? rem % ND : 0;
  // phases 2 and 4: pencil (i2, :, z2)
  const bool role2 = slot_ok && rem < ND * NQ;
  const int i2 = role2 ? rem / NQ : 0, z2 = role2 ? rem % NQ : 0;
  T* const s1_own = S1 + i1 * S1I + j1 * S1J; // + v*S1V + qz
  T* const s1_col = S1 + i2 * S1I + z2;       // + v*S1V + j*S1J
  T* const s2_col = S2 + i2 * S2I + z2;       // + v*S2V + qy*NQ
  T* const s2_own = S2 + rem;                 // + v*S2V + i*S2I   (phase 3: point (:, qy, qz))
  T* const ring = S2 + 3 * S2V + rem;         // + (slot*6 + comp)*TPC   (phase 3, private to the thread)
  constexpr int SLOT = 6 * TPC;

  const int nbatches = (ncells + CPB - 1) / CPB;
  const std::uint64_t g_policy = l2_evict_first_policy();
  auto row_of = [&](int batch) -> const std::int32_t* {
    const int pos = batch * CPB + c;
    if (!(role1 && batch < nbatches && pos < ncells))
      return nullptr;
    return mdofmap + static_cast<std::int64_t>(cells[pos]) * ND3 + (i1 * ND + j1) * ND;
  };

  // G rows: one stream over all the cells this thread visits; row g lives in slot g % RD
  auto g_of = [&](int batch) -> const T* {
    const int pos = batch * CPB + c;
    if (!(slot_ok && batch < nbatches && pos < ncells))
      return nullptr;
    return G + static_cast<std::int64_t>(pos) * (6 * NQ3) + rem;
  };
  // The destination is a per-thread shared-memory address in an ordinary register, advanced slot by
  // slot.  (Derived from the rolled loop counter instead, ptxas 12.9 folds the slot offset into a
  // uniform-register operand of LDGSTS next to the cache-hint descriptor and emits an encoding that
  // faults -- seen for nd = 5, nq = 6 in fp64 -- hence the opaque asm in the rolled variant.)
  constexpr std::uint32_t SLOT_BYTES = SLOT * sizeof(T);
  const std::uint32_t ring_base = smem_addr(ring);
  auto issue_row = [&](std::uint32_t dst, const T* gcell, int qx) {
    const T* src = gcell + qx * TPC;
#pragma unroll
    for (int comp = 0; comp < 6; ++comp)
      cp_async_scalar(dst + comp * TPC * std::uint32_t(sizeof(T)), src + comp * NQ3, g_policy);
  };
  std::uint32_t ring_wr = ring_base; // slot of the next row to request
  const T* ring_rd = ring;           // slot of the next row to consume
  const T* g_cur = g_of(blockIdx.x);
  if (STREAM && g_cur)
  {
    // rows 0 .. RD-2 of the first cell, one commit group per row
#pragma unroll
    for (int r = 0; r < RD - 1; ++r)
    {
      issue_row(ring_wr, g_cur, r);
      cp_async_commit();
      ring_wr += SLOT_BYTES;
    }
  }

  std::int32_t dof[ND];
  T uv[ND];
  {
    const std::int32_t* row = row_of(blockIdx.x);
    pencil_load_dofs<ND>(row, dof);
    if constexpr (UPF)
    {
#pragma unroll
      for (int k = 0; k < ND; ++k)
        uv[k] = row ? u[dof[k] & 0x7fffffff] : T(0);
    }
  }

  for (int batch = blockIdx.x; batch < nbatches; batch += gridDim.x)
  {
    const int pos = batch * CPB + c;
    const bool active = slot_ok && pos < ncells;
    const bool act1 = role1 && active;
    const bool act2 = role2 && active;
    // next batch: the address of its dof row is needed right after phase 1, its G rows in phase 3
    const std::int32_t* next_row = row_of(batch + gridDim.x);
    const T* g_next = STREAM ? g_of(batch + gridDim.x) : nullptr;
    if constexpr (!STREAM)
    {
      // ring restarted per cell: slot of row qx is qx % RD (static in the unrolled variant, derived
      // from the loop counter in the rolled one -- tests/test_abi.py checks the SASS for the faulting
      // LDGSTS form described above)
      g_cur = g_of(batch);
      if (g_cur)
      {
#pragma unroll
        for (int r = 0; r < RD - 1; ++r)
        {
          issue_row(ring_base + r * SLOT_BYTES, g_cur, r);
          cp_async_commit();
        }
      }
      ring_wr = ring_base + (RD - 1) * SLOT_BYTES; // row RD-1 goes into slot RD-1 (used when RD != 3)
    }

    // ---- phase 1: gather, BC masking (reference :153-170), z contraction
    if (act1)
    {
      const T kappa = cell_constants[pos];
      if constexpr (!UPF)
      {
#pragma unroll
        for (int k = 0; k < ND; ++k)
          uv[k] = u[dof[k] & 0x7fffffff];
      }
#pragma unroll
      for (int k = 0; k < ND; ++k)
      {
        if (dof[k] < 0)
        {
          y[dof[k] & 0x7fffffff] = uv[k]; // b[dof] = u[dof]; the dof contributes 0 and receives nothing
          uv[k] = 0;
        }
        uv[k] *= kappa; // kappa folded into the input: y = A(kappa) u = A(1) (kappa u)
      }
      if constexpr (PAIRED)
      {
        float2 pz2[NQP], dz2[NQP];
#pragma unroll
        for (int qp = 0; qp < NQP; ++qp)
          pz2[qp] = make_float2(0.f, 0.f), dz2[qp] = make_float2(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < ND; ++k)
#pragma unroll
          for (int qp = 0; qp < NQP; ++qp)
          {
            pz2[qp] = pair_fma(uv[k], ptab.fphi[k * NQP + qp], pz2[qp]);
            dz2[qp] = pair_fma(uv[k], ptab.fdphi[k * NQP + qp], dz2[qp]);
          }
#pragma unroll
        for (int qz = 0; qz < NQ; ++qz)
        {
          s1_own[qz] = pair_get(pz2, qz);
          s1_own[S1V + qz] = pair_get(dz2, qz);
        }
      }
      else
      {
#pragma unroll OUTER
        for (int qz = 0; qz < NQ; ++qz)
        {
          s1_own[qz] = pencil_dot<T, ND>(tab.phi + qz * ND, uv);
          s1_own[S1V + qz] = pencil_dot<T, ND>(tab.dphi + qz * ND, uv);
        }
      }
    }
    std::int32_t ndof[ND];
    pencil_load_dofs<ND>(next_row, ndof);
    __syncthreads();

    // ---- phase 2: y contraction
    if (act2)
    {
      T pzv[ND], dzv[ND];
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        pzv[j] = s1_col[j * S1J];
        dzv[j] = s1_col[S1V + j * S1J];
      }
      if constexpr (PAIRED)
      {
        float2 a2[NQP], b2[NQP], c2[NQP];
#pragma unroll
        for (int qp = 0; qp < NQP; ++qp)
          a2[qp] = make_float2(0.f, 0.f), b2[qp] = make_float2(0.f, 0.f), c2[qp] = make_float2(0.f, 0.f);
#pragma unroll
        for (int j = 0; j < ND; ++j)
#pragma unroll
          for (int qp = 0; qp < NQP; ++qp)
          {
            a2[qp] = pair_fma(pzv[j], ptab.fphi[j * NQP + qp], a2[qp]);
            b2[qp] = pair_fma(pzv[j], ptab.fdphi[j * NQP + qp], b2[qp]);
            c2[qp] = pair_fma(dzv[j], ptab.fphi[j * NQP + qp], c2[qp]);
          }
#pragma unroll
        for (int qy = 0; qy < NQ; ++qy)
        {
          s2_col[qy * NQ] = pair_get(a2, qy);
          s2_col[S2V + qy * NQ] = pair_get(b2, qy);
          s2_col[2 * S2V + qy * NQ] = pair_get(c2, qy);
        }
      }
      else
      {
#pragma unroll OUTER
        for (int qy = 0; qy < NQ; ++qy)
        {
          s2_col[qy * NQ] = pencil_dot<T, ND>(tab.phi + qy * ND, pzv);
          s2_col[S2V + qy * NQ] = pencil_dot<T, ND>(tab.dphi + qy * ND, pzv);
          s2_col[2 * S2V + qy * NQ] = pencil_dot<T, ND>(tab.phi + qy * ND, dzv);
        }
      }
    }
    __syncthreads();

    // ---- phase 3: x contraction, G (reference :239-273), transposed x contraction
    if (active)
    {
      T A[ND], B[ND], Cz[ND], oA[ND], oB[ND], oC[ND];
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        A[i] = s2_own[i * S2I];
        B[i] = s2_own[S2V + i * S2I];
        Cz[i] = s2_own[2 * S2V + i * S2I];
        oA[i] = 0, oB[i] = 0, oC[i] = 0;
      }
      // PAIRED: the gradient at all nq points of the line first (needs no G), outputs as pairs
      float2 gx2[PAIRED ? NQP : 1], gy2[PAIRED ? NQP : 1], gz2[PAIRED ? NQP : 1];
      float2 oA2[PAIRED ? NDP : 1], oB2[PAIRED ? NDP : 1], oC2[PAIRED ? NDP : 1];
      if constexpr (PAIRED)
      {
#pragma unroll
        for (int qp = 0; qp < NQP; ++qp)
          gx2[qp] = make_float2(0.f, 0.f), gy2[qp] = make_float2(0.f, 0.f), gz2[qp] = make_float2(0.f, 0.f);
#pragma unroll
        for (int ip = 0; ip < NDP; ++ip)
          oA2[ip] = make_float2(0.f, 0.f), oB2[ip] = make_float2(0.f, 0.f), oC2[ip] = make_float2(0.f, 0.f);
#pragma unroll
        for (int i = 0; i < ND; ++i)
#pragma unroll
          for (int qp = 0; qp < NQP; ++qp)
          {
            gx2[qp] = pair_fma(A[i], ptab.fdphi[i * NQP + qp], gx2[qp]);
            gy2[qp] = pair_fma(B[i], ptab.fphi[i * NQP + qp], gy2[qp]);
            gz2[qp] = pair_fma(Cz[i], ptab.fphi[i * NQP + qp], gz2[qp]);
          }
      }
#pragma unroll OUTER
      for (int qx = 0; qx < NQ; ++qx)
      {
        // row g of the stream is commit group g, and RD - 1 + g groups have been committed so far
        cp_async_wait<RD - 2>();
        const T* gr = STREAM ? ring_rd : ring + (qx % RD) * SLOT;
        const T G0 = gr[0], G1 = gr[TPC], G2 = gr[2 * TPC], G3 = gr[3 * TPC], G4 = gr[4 * TPC],
                G5 = gr[5 * TPC];
        if constexpr (STREAM)
          ring_rd = (ring_rd == ring + (RD - 1) * SLOT) ? ring : ring_rd + SLOT;
        T gx, gy, gz;
        if constexpr (PAIRED)
          gx = pair_get(gx2, qx), gy = pair_get(gy2, qx), gz = pair_get(gz2, qx);
        else
        {
          gx = pencil_dot<T, ND>(tab.dphi + qx * ND, A);
          gy = pencil_dot<T, ND>(tab.phi + qx * ND, B);
          gz = pencil_dot<T, ND>(tab.phi + qx * ND, Cz);
        }
        const T fx = G0 * gx + G1 * gy + G2 * gz;
        const T fy = G1 * gx + G3 * gy + G4 * gz;
        const T fz = G2 * gx + G4 * gy + G5 * gz;
        // request row g + RD - 1 (of this cell or of the next one) into the slot consumed in the
        // previous iteration: its values went into last iteration's fx, fy, fz
        if constexpr (STREAM)
        {
          if constexpr (OUTER == 1)
            asm volatile("" : "+r"(ring_wr));
          if (qx + RD - 1 < NQ)
            issue_row(ring_wr, g_cur, qx + RD - 1);
          else if (g_next)
            issue_row(ring_wr, g_next, qx + RD - 1 - NQ);
          ring_wr = (ring_wr == ring_base + (RD - 1) * SLOT_BYTES) ? ring_base : ring_wr + SLOT_BYTES;
        }
        else if constexpr (RD == 3)
        {
          // the shipped fp64 configuration: slot from the loop counter (its SASS is checked, see above)
          if (qx + RD - 1 < NQ)
            issue_row(ring_base + ((qx + RD - 1) % RD) * SLOT_BYTES, g_cur, qx + RD - 1);
        }
        else
        {
          // other depths: running destination address in an ordinary register (for RD = 5 ptxas emits the
          // faulting LDGSTS form when the slot is derived from the loop counter)
          if constexpr (OUTER == 1)
            asm volatile("" : "+r"(ring_wr));
          if (qx + RD - 1 < NQ)
            issue_row(ring_wr, g_cur, qx + RD - 1);
          ring_wr = (ring_wr == ring_base + (RD - 1) * SLOT_BYTES) ? ring_base : ring_wr + SLOT_BYTES;
        }
        cp_async_commit(); // also when empty: keeps the group count uniform
        if constexpr (PAIRED)
        {
#pragma unroll
          for (int ip = 0; ip < NDP; ++ip)
          {
            oA2[ip] = pair_fma(fx, ptab.bdphi[qx * NDP + ip], oA2[ip]);
            oB2[ip] = pair_fma(fy, ptab.bphi[qx * NDP + ip], oB2[ip]);
            oC2[ip] = pair_fma(fz, ptab.bphi[qx * NDP + ip], oC2[ip]);
          }
        }
        else
        {
#pragma unroll
          for (int i = 0; i < ND; ++i)
          {
            oA[i] += tab.dphi[qx * ND + i] * fx;
            oB[i] += tab.phi[qx * ND + i] * fy;
            oC[i] += tab.phi[qx * ND + i] * fz;
          }
        }
      }
      // the thread overwrites exactly the entries it read: no barrier between the read and the write
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        s2_own[i * S2I] = PAIRED ? T(pair_get(oA2, i)) : oA[i];
        s2_own[S2V + i * S2I] = PAIRED ? T(pair_get(oB2, i)) : oB[i];
        s2_own[2 * S2V + i * S2I] = PAIRED ? T(pair_get(oC2, i)) : oC[i];
      }
    }
    // next cell's input values: in flight during phases 4 and 5 (registers are free again)
    T nuv[UPF ? ND : 1];
    if constexpr (UPF)
    {
#pragma unroll
      for (int k = 0; k < ND; ++k)
        nuv[k] = next_row ? u[ndof[k] & 0x7fffffff] : T(0);
    }
    __syncthreads();

    // ---- phase 4: transposed y contraction
    if (act2)
    {
      T pz[ND], dz[ND];
      float2 pz2[PAIRED ? NDP : 1], dz2[PAIRED ? NDP : 1];
#pragma unroll
      for (int j = 0; j < ND; ++j)
        pz[j] = 0, dz[j] = 0;
      if constexpr (PAIRED)
      {
#pragma unroll
        for (int jp = 0; jp < NDP; ++jp)
          pz2[jp] = make_float2(0.f, 0.f), dz2[jp] = make_float2(0.f, 0.f);
      }
#pragma unroll OUTER
      for (int qy = 0; qy < NQ; ++qy)
      {
        const T a = s2_col[qy * NQ], b = s2_col[S2V + qy * NQ], cz = s2_col[2 * S2V + qy * NQ];
        if constexpr (PAIRED)
        {
#pragma unroll
          for (int jp = 0; jp < NDP; ++jp)
          {
            pz2[jp] = pair_fma(a, ptab.bphi[qy * NDP + jp], pz2[jp]);
            pz2[jp] = pair_fma(b, ptab.bdphi[qy * NDP + jp], pz2[jp]);
            dz2[jp] = pair_fma(cz, ptab.bphi[qy * NDP + jp], dz2[jp]);
          }
        }
        else
        {
#pragma unroll
          for (int j = 0; j < ND; ++j)
          {
            pz[j] += tab.phi[qy * ND + j] * a;
            pz[j] += tab.dphi[qy * ND + j] * b;
            dz[j] += tab.phi[qy * ND + j] * cz;
          }
        }
      }
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        s1_col[j * S1J] = PAIRED ? T(pair_get(pz2, j)) : pz[j];
        s1_col[S1V + j * S1J] = PAIRED ? T(pair_get(dz2, j)) : dz[j];
      }
    }
    __syncthreads();

    // ---- phase 5: transposed z contraction, scatter-add (reference :424-425)
    if (act1)
    {
      T out[ND];
      float2 out2[PAIRED ? NDP : 1];
#pragma unroll
      for (int k = 0; k < ND; ++k)
        out[k] = 0;
      if constexpr (PAIRED)
      {
#pragma unroll
        for (int kp = 0; kp < NDP; ++kp)
          out2[kp] = make_float2(0.f, 0.f);
      }
#pragma unroll OUTER
      for (int qz = 0; qz < NQ; ++qz)
      {
        const T pz = s1_own[qz], dz = s1_own[S1V + qz];
        if constexpr (PAIRED)
        {
#pragma unroll
          for (int kp = 0; kp < NDP; ++kp)
          {
            out2[kp] = pair_fma(pz, ptab.bphi[qz * NDP + kp], out2[kp]);
            out2[kp] = pair_fma(dz, ptab.bdphi[qz * NDP + kp], out2[kp]);
          }
        }
        else
        {
#pragma unroll
          for (int k = 0; k < ND; ++k)
          {
            out[k] += tab.phi[qz * ND + k] * pz;
            out[k] += tab.dphi[qz * ND + k] * dz;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < ND; ++k)
        if (dof[k] >= 0)
          red_add(y + dof[k], PAIRED ? T(pair_get(out2, k)) : out[k]);
    }
    // Phases 1 and 5 touch the same S1 entries from the same thread, so the next cell's phase 1 may
    // start at once; its phase 2 writes S2 only after the barrier that follows phase 1, which every
    // thread reaches after its phase 4 reads of S2.  The ring is private to the thread.
#pragma unroll
    for (int k = 0; k < ND; ++k)
    {
      dof[k] = ndof[k];
      if constexpr (UPF)
        uv[k] = nuv[k];
    }
    if constexpr (STREAM)
      g_cur = g_next;
  }
}
} // namespace bd
