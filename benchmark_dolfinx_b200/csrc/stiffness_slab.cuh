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
  extern __shared__ __align__(16) unsigned char smem_raw[];
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
