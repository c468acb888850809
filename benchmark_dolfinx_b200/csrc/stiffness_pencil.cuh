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
// into a private slot of a small shared-memory ring, RD - 1 rows ahead of their use; nobody else
// reads the slot, so the ring needs no barrier, only cp.async.wait_group.
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

// Cells per CTA: fill the warps of a 96..256-thread CTA as completely as possible.
constexpr int pencil_default_cpb(int nq)
{
  return nq >= 9 ? 1 : (nq == 8 ? 2 : (nq == 7 ? 5 : (nq == 6 ? 7 : (256 / (nq * nq)))));
}

constexpr int kPencilRing = 3; // q0 rows of G in flight per cell

template <typename T, int P, int NQ, int CPB>
struct PencilConfig
{
  static constexpr int RD = kPencilRing;
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
  static constexpr bool SUPPORTED = CTA_SMEM <= 226 * 1024 && THREADS <= 1024;
};

// One G entry (read once per operator application: evict-first in L2) from global to shared memory.
template <typename T>
__device__ __forceinline__ void cp_async_scalar(T* dst, const T* src, std::uint64_t policy)
{
  static_assert(sizeof(T) == 4 || sizeof(T) == 8, "cp.async.ca copies 4, 8 or 16 bytes");
  if constexpr (sizeof(T) == 8)
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 8, %2;" ::"r"(smem_addr(dst)), "l"(src),
                 "l"(policy)
                 : "memory");
  else
    asm volatile("cp.async.ca.shared.global.L2::cache_hint [%0], [%1], 4, %2;" ::"r"(smem_addr(dst)), "l"(src),
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

template <typename T, int P, int NQ, int CPB>
__global__ void __launch_bounds__(PencilConfig<T, P, NQ, CPB>::THREADS)
    stiffness_pencil_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ G,
                            const T* __restrict__ cell_constants,
                            const std::int32_t* __restrict__ mdofmap,
                            const std::int32_t* __restrict__ cells, int ncells,
                            const __grid_constant__ SlabTables<T, P + 1, NQ> tab)
{
  using C = PencilConfig<T, P, NQ, CPB>;
  constexpr int ND = C::ND, TPC = C::TPC, NQ3 = TPC * NQ, ND2 = ND * ND, ND3 = ND2 * ND;
  constexpr int S1I = C::S1I, S1J = C::S1J, S1V = C::S1V, S2I = C::S2I, S2V = C::S2V, RD = C::RD;
  constexpr int OUTER = sizeof(T) == 8 ? 1 : NQ; // unroll factor of the outer loop of each phase
  static_assert(RD >= 2 && RD <= NQ, "ring depth");
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

  const int nbatches = (ncells + CPB - 1) / CPB;
  const std::uint64_t g_policy = l2_evict_first_policy();
  auto row_of = [&](int batch) -> const std::int32_t* {
    const int pos = batch * CPB + c;
    if (!(role1 && batch < nbatches && pos < ncells))
      return nullptr;
    return mdofmap + static_cast<std::int64_t>(cells[pos]) * ND3 + (i1 * ND + j1) * ND;
  };

  std::int32_t dof[ND];
  pencil_load_dofs<ND>(row_of(blockIdx.x), dof);

  for (int batch = blockIdx.x; batch < nbatches; batch += gridDim.x)
  {
    const int pos = batch * CPB + c;
    const bool active = slot_ok && pos < ncells;
    const bool act1 = role1 && active;
    const bool act2 = role2 && active;
    // next batch's dof row: its address is needed right after phase 1
    const std::int32_t* next_row = row_of(batch + gridDim.x);

    // the first RD - 1 q0 rows of G travel while phases 1 and 2 run (one commit group per row)
    const T* Gc = G + static_cast<std::int64_t>(active ? pos : 0) * (6 * NQ3) + rem;
    auto issue_row = [&](int qx) {
      T* dst = ring + (qx % RD) * (6 * TPC);
      const T* src = Gc + qx * TPC;
#pragma unroll
      for (int comp = 0; comp < 6; ++comp)
        cp_async_scalar(dst + comp * TPC, src + comp * NQ3, g_policy);
    };
    if (active)
    {
#pragma unroll
      for (int r = 0; r < RD - 1; ++r)
      {
        issue_row(r);
        cp_async_commit();
      }
    }

    // ---- phase 1: gather, BC masking (reference :153-170), z contraction
    if (act1)
    {
      const T kappa = cell_constants[pos];
      T uv[ND];
#pragma unroll
      for (int k = 0; k < ND; ++k)
        uv[k] = u[dof[k] & 0x7fffffff];
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
#pragma unroll OUTER
      for (int qz = 0; qz < NQ; ++qz)
      {
        T pz = 0, dz = 0;
#pragma unroll
        for (int k = 0; k < ND; ++k)
        {
          pz += tab.phi[qz * ND + k] * uv[k];
          dz += tab.dphi[qz * ND + k] * uv[k];
        }
        s1_own[qz] = pz;
        s1_own[S1V + qz] = dz;
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
#pragma unroll OUTER
      for (int qy = 0; qy < NQ; ++qy)
      {
        T a = 0, b = 0, cz = 0;
#pragma unroll
        for (int j = 0; j < ND; ++j)
        {
          a += tab.phi[qy * ND + j] * pzv[j];
          b += tab.dphi[qy * ND + j] * pzv[j];
          cz += tab.phi[qy * ND + j] * dzv[j];
        }
        s2_col[qy * NQ] = a;
        s2_col[S2V + qy * NQ] = b;
        s2_col[2 * S2V + qy * NQ] = cz;
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
#pragma unroll OUTER
      for (int qx = 0; qx < NQ; ++qx)
      {
        // RD - 1 + qx groups have been committed; row qx is group qx
        cp_async_wait<RD - 2>();
        const T* gr = ring + (qx % RD) * (6 * TPC);
        const T G0 = gr[0], G1 = gr[TPC], G2 = gr[2 * TPC], G3 = gr[3 * TPC], G4 = gr[4 * TPC],
                G5 = gr[5 * TPC];
        T gx = 0, gy = 0, gz = 0;
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          gx += tab.dphi[qx * ND + i] * A[i];
          gy += tab.phi[qx * ND + i] * B[i];
          gz += tab.phi[qx * ND + i] * Cz[i];
        }
        const T fx = G0 * gx + G1 * gy + G2 * gz;
        const T fy = G1 * gx + G3 * gy + G4 * gz;
        const T fz = G2 * gx + G4 * gy + G5 * gz;
        // refill the slot read in the previous iteration (its values have been consumed above: gx..fz
        // depend on this iteration's loads, which were issued after them)
        if (qx + RD - 1 < NQ)
          issue_row(qx + RD - 1);
        cp_async_commit(); // also when empty: keeps the group count uniform
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          oA[i] += tab.dphi[qx * ND + i] * fx;
          oB[i] += tab.phi[qx * ND + i] * fy;
          oC[i] += tab.phi[qx * ND + i] * fz;
        }
      }
      // the thread overwrites exactly the entries it read: no barrier between the read and the write
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        s2_own[i * S2I] = oA[i];
        s2_own[S2V + i * S2I] = oB[i];
        s2_own[2 * S2V + i * S2I] = oC[i];
      }
    }
    __syncthreads();

    // ---- phase 4: transposed y contraction
    if (act2)
    {
      T pz[ND], dz[ND];
#pragma unroll
      for (int j = 0; j < ND; ++j)
        pz[j] = 0, dz[j] = 0;
#pragma unroll OUTER
      for (int qy = 0; qy < NQ; ++qy)
      {
        const T a = s2_col[qy * NQ], b = s2_col[S2V + qy * NQ], cz = s2_col[2 * S2V + qy * NQ];
#pragma unroll
        for (int j = 0; j < ND; ++j)
        {
          pz[j] += tab.phi[qy * ND + j] * a;
          pz[j] += tab.dphi[qy * ND + j] * b;
          dz[j] += tab.phi[qy * ND + j] * cz;
        }
      }
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        s1_col[j * S1J] = pz[j];
        s1_col[S1V + j * S1J] = dz[j];
      }
    }
    __syncthreads();

    // ---- phase 5: transposed z contraction, scatter-add (reference :424-425)
    if (act1)
    {
      T out[ND];
#pragma unroll
      for (int k = 0; k < ND; ++k)
        out[k] = 0;
#pragma unroll OUTER
      for (int qz = 0; qz < NQ; ++qz)
      {
        const T pz = s1_own[qz], dz = s1_own[S1V + qz];
#pragma unroll
        for (int k = 0; k < ND; ++k)
        {
          out[k] += tab.phi[qz * ND + k] * pz;
          out[k] += tab.dphi[qz * ND + k] * dz;
        }
      }
#pragma unroll
      for (int k = 0; k < ND; ++k)
        if (dof[k] >= 0)
          red_add(y + dof[k], out[k]);
    }
    // Phases 1 and 5 touch the same S1 entries from the same thread, so the next cell's phase 1 may
    // start at once; its phase 2 writes S2 only after the barrier that follows phase 1, which every
    // thread reaches after its phase 4 reads of S2.  The ring is private to the thread.
#pragma unroll
    for (int k = 0; k < ND; ++k)
      dof[k] = ndof[k];
  }
}
} // namespace bd
