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

template <typename T>
__device__ __forceinline__ T ld_stream(const T* p)
{
  return __ldg(p);
}

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
  // element (k3, q1, i, row r): cube[k3*CUBE + (q1*ND + i)*RL + r], r = c*NQ + q2
  const int rbase = (lane_ok ? c : 0) * NQ;

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * C::WARPS;
  for (int batch = blockIdx.x * C::WARPS + warp; batch < nbatches; batch += warps_total)
  {
    const int pos = batch * CW + c;
    const bool active = lane_ok && pos < ncells;
    const bool xlane = active && m < ND; // lanes that own a dof plane

    // ------------------------------------------------------------------ gather (X layout)
    std::int32_t dof[ND2];
    T ul[ND2];
    if (xlane)
    {
      const std::int32_t* row = dofmap + static_cast<std::int64_t>(cells[pos]) * ND3 + m * ND2;
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        dof[t] = row[t];
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        const T v = u[dof[t]];
        if (bc[dof[t]])
        {
          y[dof[t]] = v; // b[dof] = u[dof] on constrained dofs (reference :163-166)
          dof[t] = -1;
          ul[t] = 0;
        }
        else
          ul[t] = v;
      }
    }
    else
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        dof[t] = -1;
        ul[t] = 0;
      }
    }

    // ------------------------------------------------------------------ forward X phase
    // stream over q2: column of Phi_z u / DPhi_z u, then the y contraction, then store
    if (xlane)
    {
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
          T* dst = cube + (q1 * ND + m) * RL + rbase + q2;
          dst[0] = ax;
          dst[CUBE] = ay;
          dst[2 * CUBE] = az;
        }
      }
    }
    __syncwarp();

    // ------------------------------------------------------------------ Z phase: x contractions + G
    if (active)
    {
      const T kappa = cell_constants[pos];
      const T* Gb = Gs + static_cast<std::int64_t>(batch) * C::G_BATCH + rbase + m;
      // fully unrolled for small NQ; a real loop otherwise (code size), only addresses depend on q1
#pragma unroll(NQ <= 5 ? NQ : 1)
      for (int q1 = 0; q1 < NQ; ++q1)
      {
        T ax[ND], ay[ND], az[ND];
        T* src = cube + (q1 * ND) * RL + rbase + m;
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          ax[i] = src[i * RL];
          ay[i] = src[i * RL + CUBE];
          az[i] = src[i * RL + 2 * CUBE];
        }
        T bx[ND], by[ND], bz[ND];
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          bx[i] = 0;
          by[i] = 0;
          bz[i] = 0;
        }
#pragma unroll
        for (int q0 = 0; q0 < NQ; ++q0)
        {
          const T* g = Gb + (static_cast<std::int64_t>(q1) * NQ + q0) * (6 * C::ROW);
          const T G0 = ld_stream(g + 0 * C::ROW);
          const T G1 = ld_stream(g + 1 * C::ROW);
          const T G2 = ld_stream(g + 2 * C::ROW);
          const T G3 = ld_stream(g + 3 * C::ROW);
          const T G4 = ld_stream(g + 4 * C::ROW);
          const T G5 = ld_stream(g + 5 * C::ROW);
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
          // reference :269-273
          const T fx = kappa * (G0 * gx + G1 * gy + G2 * gz);
          const T fy = kappa * (G1 * gx + G3 * gy + G4 * gz);
          const T fz = kappa * (G2 * gx + G4 * gy + G5 * gz);
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

    // ------------------------------------------------------------------ backward X phase + scatter
    if (xlane)
    {
      T out[ND2];
#pragma unroll
      for (int t = 0; t < ND2; ++t)
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
          const T* src = cube + (q1 * ND + m) * RL + rbase + q2;
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
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        if (dof[t] >= 0)
          atomicAdd(y + dof[t], out[t]); // result unused -> RED.E.ADD (reference :424-425)
    }
    __syncwarp(); // the cube is rewritten by the next batch's forward phase
  }
}

// ---------------------------------------------------------------------------------------------
// Asynchronous-copy variant: the G stream is decoupled from the math.
//
// ncu on the plain slab kernel (profiles/r1_v2_*) shows it purely latency bound: 11.6 of 14.9
// cycles between issues are "long scoreboard" (waiting for G / gather loads) at 12 warps per SM.
// Here every warp keeps a ring of NSTAGE shared-memory slots and one elected lane streams the
// batch's G through it with 1-D bulk asynchronous copies (cp.async.bulk -> UBLKCP, completion on
// an mbarrier): one slot = one q1 slice of the batch (NQ*6*ROW scalars, contiguous in the slab
// layout), NSTAGE-1 slices are always in flight, and the stream runs on across batch boundaries,
// so the next batch's first slices arrive while the current batch is still in its backward phase.
// The dof indices and input values of the next batch are prefetched into registers the same way.
// ---------------------------------------------------------------------------------------------
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

template <typename T, int P, int NQ, int NSTAGE>
struct SlabTmaConfig
{
  using Base = SlabConfig<T, P, NQ>;
  static constexpr int SLICE = NQ * 6 * Base::ROW;                 // scalars per q1 slice of a batch
  static constexpr int SLICE_BYTES = SLICE * int(sizeof(T));
  static constexpr int CUBE_BYTES = (3 * Base::CUBE * int(sizeof(T)) + 15) / 16 * 16;
  static constexpr int WARP_SMEM = CUBE_BYTES + NSTAGE * SLICE_BYTES + 64; // + mbarriers
  static constexpr int MAX_SMEM = 225 * 1024;
  static constexpr int WARPS = WARP_SMEM * 2 + 2048 <= MAX_SMEM ? 2 : 1;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_SMEM;
  static constexpr bool SUPPORTED = (SLICE_BYTES % 16 == 0) && (WARP_SMEM + 1024 <= MAX_SMEM);
};

template <typename T, int P, int NQ, bool IDENTITY, int NSTAGE>
__global__ void __launch_bounds__(SlabTmaConfig<T, P, NQ, NSTAGE>::THREADS)
    stiffness_slab_tma_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ Gs,
                              const T* __restrict__ cell_constants,
                              const std::int32_t* __restrict__ dofmap,
                              const std::int32_t* __restrict__ cells, int ncells,
                              const std::int8_t* __restrict__ bc,
                              const __grid_constant__ SlabTables<T, P + 1, NQ> tab)
{
  using C = SlabConfig<T, P, NQ>;
  using TC = SlabTmaConfig<T, P, NQ, NSTAGE>;
  constexpr int ND = C::ND, CW = C::CW, RL = C::RL, CUBE = C::CUBE, ROW = C::ROW;
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / NQ;
  const int m = lane - c * NQ;
  const bool lane_ok = c < CW;
  unsigned char* wbase = smem_raw + warp * TC::WARP_SMEM;
  T* cube = reinterpret_cast<T*>(wbase);
  T* ring = reinterpret_cast<T*>(wbase + TC::CUBE_BYTES);
  std::uint64_t* bars = reinterpret_cast<std::uint64_t*>(wbase + TC::CUBE_BYTES + NSTAGE * TC::SLICE_BYTES);
  const int rbase = (lane_ok ? c : 0) * NQ;

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * TC::WARPS;
  const int first = blockIdx.x * TC::WARPS + warp;
  const int my_batches = first < nbatches ? (nbatches - first + warps_total - 1) / warps_total : 0;
  const int total_slices = my_batches * NQ;

  if (lane == 0)
  {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s)
      mbar_init(bars + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();

  // slice s of this warp's stream = (batch first + (s / NQ) * warps_total, q1 = s % NQ)
  auto issue = [&](int s) {
    const int k = s / NQ, q1 = s - k * NQ;
    const std::int64_t batch = first + static_cast<std::int64_t>(k) * warps_total;
    const T* src = Gs + batch * C::G_BATCH + static_cast<std::int64_t>(q1) * TC::SLICE;
    const int st = s % NSTAGE;
    mbar_expect_tx(bars + st, TC::SLICE_BYTES);
    bulk_copy_g2s(ring + st * TC::SLICE, src, TC::SLICE_BYTES, bars + st);
  };
  if (lane == 0)
  {
    for (int s = 0; s < NSTAGE - 1 && s < total_slices; ++s)
      issue(s);
  }

  // dof indices / inputs of the current batch (prefetched one batch ahead)
  std::int32_t dof[ND2];
  T ul[ND2];
  std::int8_t flag[ND2];
  auto load_dofs = [&](int batch, std::int32_t (&d)[ND2]) {
    const int pos = batch * CW + c;
    const bool x = lane_ok && pos < ncells && m < ND && batch < nbatches;
    if (x)
    {
      const std::int32_t* row = dofmap + static_cast<std::int64_t>(cells[pos]) * ND3 + m * ND2;
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        d[t] = row[t];
    }
    else
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        d[t] = -1;
    }
  };
  auto load_values = [&](const std::int32_t (&d)[ND2], T (&v)[ND2], std::int8_t (&f)[ND2]) {
#pragma unroll
    for (int t = 0; t < ND2; ++t)
    {
      if (d[t] >= 0)
      {
        v[t] = u[d[t]];
        f[t] = bc[d[t]];
      }
      else
      {
        v[t] = 0;
        f[t] = 0;
      }
    }
  };
  if (my_batches > 0)
  {
    load_dofs(first, dof);
    load_values(dof, ul, flag);
  }

  int slice = 0;
  for (int k = 0; k < my_batches; ++k)
  {
    const int batch = first + k * warps_total;
    const int pos = batch * CW + c;
    const bool active = lane_ok && pos < ncells;
    const bool xlane = active && m < ND;

    // BC masking of the gathered values (reference :163-169)
#pragma unroll
    for (int t = 0; t < ND2; ++t)
    {
      if (flag[t])
      {
        y[dof[t]] = ul[t];
        dof[t] = -1;
        ul[t] = 0;
      }
    }

    // ------------------------------------------------------------------ forward X phase
    if (xlane)
    {
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
            for (int kk = 0; kk < ND; ++kk)
              s += tab.dphi[q2 * ND + kk] * ul[j * ND + kk];
            dz[j] = s;
          }
          else
          {
            T s = 0, d = 0;
#pragma unroll
            for (int kk = 0; kk < ND; ++kk)
            {
              s += tab.phi[q2 * ND + kk] * ul[j * ND + kk];
              d += tab.dphi[q2 * ND + kk] * ul[j * ND + kk];
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
          T* dst = cube + (q1 * ND + m) * RL + rbase + q2;
          dst[0] = ax;
          dst[CUBE] = ay;
          dst[2 * CUBE] = az;
        }
      }
    }
    __syncwarp();

    // next batch's dof indices: in flight during the whole middle phase
    std::int32_t ndof[ND2];
    load_dofs(batch + warps_total, ndof);

    // ------------------------------------------------------------------ Z phase
    const T kappa = active ? cell_constants[pos] : T(0);
#pragma unroll 1
    for (int q1 = 0; q1 < NQ; ++q1, ++slice)
    {
      // refill the slot consumed one step ago with the slice NSTAGE-1 ahead
      if (lane == 0 && slice + NSTAGE - 1 < total_slices)
        issue(slice + NSTAGE - 1);
      const int st = slice % NSTAGE;
      mbar_wait(bars + st, (slice / NSTAGE) & 1);
      if (active)
      {
        const T* gs = ring + st * TC::SLICE + rbase + m;
        T ax[ND], ay[ND], az[ND];
        T* src = cube + (q1 * ND) * RL + rbase + m;
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          ax[i] = src[i * RL];
          ay[i] = src[i * RL + CUBE];
          az[i] = src[i * RL + 2 * CUBE];
        }
        T bx[ND], by[ND], bz[ND];
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          bx[i] = 0;
          by[i] = 0;
          bz[i] = 0;
        }
#pragma unroll
        for (int q0 = 0; q0 < NQ; ++q0)
        {
          const T* g = gs + q0 * (6 * ROW);
          const T G0 = g[0 * ROW], G1 = g[1 * ROW], G2 = g[2 * ROW];
          const T G3 = g[3 * ROW], G4 = g[4 * ROW], G5 = g[5 * ROW];
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
          const T fx = kappa * (G0 * gx + G1 * gy + G2 * gz);
          const T fy = kappa * (G1 * gx + G3 * gy + G4 * gz);
          const T fz = kappa * (G2 * gx + G4 * gy + G5 * gz);
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
#pragma unroll
        for (int i = 0; i < ND; ++i)
        {
          src[i * RL] = bx[i];
          src[i * RL + CUBE] = by[i];
          src[i * RL + 2 * CUBE] = bz[i];
        }
      }
      __syncwarp(); // every lane is done with slot `st` before it is refilled
    }

    // next batch's input values: in flight during the backward phase
    T nul[ND2];
    std::int8_t nflag[ND2];
    load_values(ndof, nul, nflag);

    // ------------------------------------------------------------------ backward X phase + scatter
    if (xlane)
    {
      T out[ND2];
#pragma unroll
      for (int t = 0; t < ND2; ++t)
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
          const T* src = cube + (q1 * ND + m) * RL + rbase + q2;
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
            for (int kk = 0; kk < ND; ++kk)
              out[j * ND + kk] += tab.dphi[q2 * ND + kk] * e2[j];
          }
          else
          {
#pragma unroll
            for (int kk = 0; kk < ND; ++kk)
            {
              out[j * ND + kk] += tab.phi[q2 * ND + kk] * e1[j];
              out[j * ND + kk] += tab.dphi[q2 * ND + kk] * e2[j];
            }
          }
        }
      }
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        if (dof[t] >= 0)
          atomicAdd(y + dof[t], out[t]);
    }
    __syncwarp(); // the cube is rewritten by the next batch's forward phase
#pragma unroll
    for (int t = 0; t < ND2; ++t)
    {
      dof[t] = ndof[t];
      ul[t] = nul[t];
      flag[t] = nflag[t];
    }
  }
}

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
