// fp32 variant of the pencil kernel (stiffness_pencil.cuh) for quadrature sizes whose per-cell G block is a multiple
// of 16 bytes (nq = 8: 12 288 B): the stored-G kernel for Q6 qmode 1 fp32 (42.6 -> 50.3 GDoF/s against the plain pencil
// kernel, profiles/r2_experiments_ab.txt; parity in tests/test_gpu_operator.py like every other kernel).
//
// Two things differ from stiffness_pencil_kernel<float, ...>, both taken from the ncu source page of
// that kernel at Q6 fp32 (profiles/r1_pencil_final_ncu_*_q6f32.*: 24 % of the stall samples are long
// scoreboard, concentrated on three instructions):
//
//  * G: the cell's whole G block ([comp][q0][q1][q2], contiguous in the reference layout) arrives
//    with ONE bulk asynchronous copy (cp.async.bulk -> UBLKCP, completion on an mbarrier) issued by
//    thread 0 right after the barrier that follows phase 3 of the previous cell, instead of 48
//    4-byte LDGSTS per thread through L1 (9.7 % of the samples sat on the first use of a G value,
//    although it had been requested a whole cell earlier: the per-SM capacity for outstanding
//    4-byte copies, not DRAM, was the limit);
//  * the cell index and kappa of a cell are loaded one cell ahead (6.7 % of the samples sat on the
//    two instructions that first use them).
//
// Everything else -- thread <-> pencil mapping, shared-memory strides, phases, FFMA2 option -- is
// the pencil kernel's; one cell per CTA.
#pragma once

#include "stiffness_pencil.cuh"

namespace bd
{
template <int P, int NQ>
struct PencilBulkConfig
{
  using B = PencilConfig<float, P, NQ, 1, 2>;
  static constexpr int ND = P + 1, TPC = NQ * NQ, THREADS = (TPC + 31) / 32 * 32;
  static constexpr int S1I = B::S1I, S1J = B::S1J, S1V = B::S1V, S2I = B::S2I, S2V = B::S2V;
  static constexpr int GOFF = (2 * S1V + 3 * S2V + 3) / 4 * 4; // floats; the G block starts 16-byte aligned
  static constexpr int GCELL = 6 * NQ * NQ * NQ;               // floats per cell
  static constexpr int GBYTES = GCELL * 4;
  static constexpr int BAR_OFF = (GOFF + GCELL) * 4; // bytes
  static constexpr int CTA_SMEM = BAR_OFF + 16;
  static constexpr bool SUPPORTED = GBYTES % 16 == 0 && CTA_SMEM <= 226 * 1024;
  // <= 128 registers, as far as shared memory lets that many CTAs in
  static constexpr int BY_REGS = 65536 / (128 * THREADS), BY_SMEM = (220 * 1024) / (CTA_SMEM + 1024);
  static constexpr int MIN_CTAS = BY_REGS < BY_SMEM ? (BY_REGS < 1 ? 1 : BY_REGS) : (BY_SMEM < 1 ? 1 : BY_SMEM);
};

template <int P, int NQ, bool PAIRED>
__global__ void __launch_bounds__(PencilBulkConfig<P, NQ>::THREADS, PencilBulkConfig<P, NQ>::MIN_CTAS)
    stiffness_pencil_bulk_kernel(const float* __restrict__ u, float* __restrict__ y, const float* __restrict__ G,
                                 const float* __restrict__ cell_constants,
                                 const std::int32_t* __restrict__ mdofmap,
                                 const std::int32_t* __restrict__ cells, int ncells,
                                 const __grid_constant__ SlabTables<float, P + 1, NQ> tab,
                                 const __grid_constant__ PencilPairTables<PAIRED ? P + 1 : 1, PAIRED ? NQ : 1> ptab)
{
  using C = PencilBulkConfig<P, NQ>;
  static_assert(C::SUPPORTED, "the per-cell G block must be a multiple of 16 bytes");
  constexpr int ND = C::ND, TPC = C::TPC, ND2 = ND * ND, ND3 = ND2 * ND;
  constexpr int S1I = C::S1I, S1J = C::S1J, S1V = C::S1V, S2I = C::S2I, S2V = C::S2V;
  constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* const S1 = reinterpret_cast<float*>(smem_raw);
  float* const S2 = S1 + 2 * S1V;
  float* const gbuf = S1 + C::GOFF;
  std::uint64_t* const bar = reinterpret_cast<std::uint64_t*>(smem_raw + C::BAR_OFF);

  const int rem = threadIdx.x;
  const bool slot_ok = rem < TPC;
  const bool role1 = rem < ND2; // phases 1 and 5: pencil (i1, j1, :) of the dofs
  const int i1 = role1 ? rem / ND : 0, j1 = role1 ? rem % ND : 0;
  const bool role2 = rem < ND * NQ; // phases 2 and 4: pencil (i2, :, z2)
  const int i2 = role2 ? rem / NQ : 0, z2 = role2 ? rem % NQ : 0;
  float* const s1_own = S1 + i1 * S1I + j1 * S1J;
  float* const s1_col = S1 + i2 * S1I + z2;
  float* const s2_col = S2 + i2 * S2I + z2;
  float* const s2_own = S2 + (slot_ok ? rem : 0);
  const float* const g_own = gbuf + (slot_ok ? rem : 0); // + (comp*NQ + q0)*TPC

  if (threadIdx.x == 0)
  {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const std::uint64_t g_policy = l2_evict_first_policy();
  auto request_g = [&](int pos) { // thread 0 only
    mbar_expect_tx(bar, C::GBYTES);
    bulk_copy_g2s_hint(gbuf, G + static_cast<std::int64_t>(pos) * C::GCELL, C::GBYTES, bar, g_policy);
  };
  const int stride = gridDim.x;
  int pos = blockIdx.x;
  if (threadIdx.x == 0 && pos < ncells)
    request_g(pos);

  auto cell_of = [&](int p) -> std::int32_t { return (role1 && p < ncells) ? cells[p] : -1; };
  auto kappa_of = [&](int p) -> float { return (role1 && p < ncells) ? cell_constants[p] : 0.f; };
  auto row_from = [&](std::int32_t cell) -> const std::int32_t* {
    return cell >= 0 ? mdofmap + static_cast<std::int64_t>(cell) * ND3 + (i1 * ND + j1) * ND : nullptr;
  };
  std::int32_t cell_n1 = cell_of(pos + stride); // cell index of the next cell: loaded a cell ahead
  float kappa = kappa_of(pos);
  std::int32_t dof[ND];
  float uv[ND];
  {
    const std::int32_t* row = row_from(cell_of(pos));
    pencil_load_dofs<ND>(row, dof);
#pragma unroll
    for (int k = 0; k < ND; ++k)
      uv[k] = row ? u[dof[k] & 0x7fffffff] : 0.f;
  }

  std::uint32_t parity = 0;
  for (; pos < ncells; pos += stride)
  {
    const std::int32_t* next_row = row_from(cell_n1);
    const std::int32_t cell_n2 = cell_of(pos + 2 * stride);
    const float kappa_next = kappa_of(pos + stride);

    // ---- phase 1: BC masking (reference :153-170), z contraction
    if (role1)
    {
#pragma unroll
      for (int k = 0; k < ND; ++k)
      {
        if (dof[k] < 0)
        {
          y[dof[k] & 0x7fffffff] = uv[k]; // b[dof] = u[dof]; the dof contributes 0 and receives nothing
          uv[k] = 0;
        }
        uv[k] *= kappa;
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
#pragma unroll
        for (int qz = 0; qz < NQ; ++qz)
        {
          s1_own[qz] = pencil_dot<float, ND>(tab.phi + qz * ND, uv);
          s1_own[S1V + qz] = pencil_dot<float, ND>(tab.dphi + qz * ND, uv);
        }
      }
    }
    std::int32_t ndof[ND];
    pencil_load_dofs<ND>(next_row, ndof);
    __syncthreads();

    // ---- phase 2: y contraction
    if (role2)
    {
      float pzv[ND], dzv[ND];
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
#pragma unroll
        for (int qy = 0; qy < NQ; ++qy)
        {
          s2_col[qy * NQ] = pencil_dot<float, ND>(tab.phi + qy * ND, pzv);
          s2_col[S2V + qy * NQ] = pencil_dot<float, ND>(tab.dphi + qy * ND, pzv);
          s2_col[2 * S2V + qy * NQ] = pencil_dot<float, ND>(tab.phi + qy * ND, dzv);
        }
      }
    }
    __syncthreads();

    // ---- phase 3: x contraction, G (reference :239-273), transposed x contraction
    if (slot_ok)
    {
      float A[ND], B[ND], Cz[ND];
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        A[i] = s2_own[i * S2I];
        B[i] = s2_own[S2V + i * S2I];
        Cz[i] = s2_own[2 * S2V + i * S2I];
      }
      float oA[ND], oB[ND], oC[ND];
      float2 gx2[PAIRED ? NQP : 1], gy2[PAIRED ? NQP : 1], gz2[PAIRED ? NQP : 1];
      float2 oA2[PAIRED ? NDP : 1], oB2[PAIRED ? NDP : 1], oC2[PAIRED ? NDP : 1];
#pragma unroll
      for (int i = 0; i < ND; ++i)
        oA[i] = 0, oB[i] = 0, oC[i] = 0;
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
      mbar_wait(bar, parity); // this cell's G block has landed
#pragma unroll
      for (int qx = 0; qx < NQ; ++qx)
      {
        const float* gr = g_own + qx * TPC;
        const float G0 = gr[0], G1 = gr[NQ * TPC], G2 = gr[2 * NQ * TPC], G3 = gr[3 * NQ * TPC],
                    G4 = gr[4 * NQ * TPC], G5 = gr[5 * NQ * TPC];
        float gx, gy, gz;
        if constexpr (PAIRED)
          gx = pair_get(gx2, qx), gy = pair_get(gy2, qx), gz = pair_get(gz2, qx);
        else
        {
          gx = pencil_dot<float, ND>(tab.dphi + qx * ND, A);
          gy = pencil_dot<float, ND>(tab.phi + qx * ND, B);
          gz = pencil_dot<float, ND>(tab.phi + qx * ND, Cz);
        }
        const float fx = G0 * gx + G1 * gy + G2 * gz;
        const float fy = G1 * gx + G3 * gy + G4 * gz;
        const float fz = G2 * gx + G4 * gy + G5 * gz;
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
        s2_own[i * S2I] = PAIRED ? pair_get(oA2, i) : oA[i];
        s2_own[S2V + i * S2I] = PAIRED ? pair_get(oB2, i) : oB[i];
        s2_own[2 * S2V + i * S2I] = PAIRED ? pair_get(oC2, i) : oC[i];
      }
    }
    // next cell's input values: in flight during phases 4 and 5
    float nuv[ND];
#pragma unroll
    for (int k = 0; k < ND; ++k)
      nuv[k] = next_row ? u[ndof[k] & 0x7fffffff] : 0.f;
    __syncthreads();
    // every thread is done with the G block: the next cell's may overwrite it (lands during phases 4, 5, 1, 2)
    parity ^= 1;
    if (threadIdx.x == 0 && pos + stride < ncells)
      request_g(pos + stride);

    // ---- phase 4: transposed y contraction
    if (role2)
    {
      float pz[ND], dz[ND];
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
#pragma unroll
      for (int qy = 0; qy < NQ; ++qy)
      {
        const float a = s2_col[qy * NQ], b = s2_col[S2V + qy * NQ], cz = s2_col[2 * S2V + qy * NQ];
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
        s1_col[j * S1J] = PAIRED ? pair_get(pz2, j) : pz[j];
        s1_col[S1V + j * S1J] = PAIRED ? pair_get(dz2, j) : dz[j];
      }
    }
    __syncthreads();

    // ---- phase 5: transposed z contraction, scatter-add (reference :424-425)
    if (role1)
    {
      float out[ND];
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
#pragma unroll
      for (int qz = 0; qz < NQ; ++qz)
      {
        const float pz = s1_own[qz], dz = s1_own[S1V + qz];
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
          red_add(y + dof[k], PAIRED ? pair_get(out2, k) : out[k]);
    }
    // buffer reuse across cells: as in stiffness_pencil_kernel (phases 1/5 and 2/4 share their mapping)
#pragma unroll
    for (int k = 0; k < ND; ++k)
    {
      dof[k] = ndof[k];
      uv[k] = nuv[k];
    }
    cell_n1 = cell_n2;
    kappa = kappa_next;
  }
}
} // namespace bd
