// fp64 tensor-core (DMMA) stiffness action for 8 quadrature points per direction (degree 6 with
// qmode 1, degree 7 with qmode 0) and, with one padded lane group, for 7 (degree 5 with qmode 1,
// degree 6 with qmode 0), sm_100a.
//
// Why: for nq = 8 the slab kernel is limited by per-thread state (49-64 values per dof plane ->
// spills, 4 warps per SM) and reaches ~20 % of the HBM roofline (profiles/r1_v4_*), while the
// contractions are compute-heavy (~50k FMA per cell).  On B200 mma.sync.m8n8k4.f64 runs at the full
// fp64 rate (37 TFLOP/s measured, scripts/microbench/fp64_rates.cu) with 1/8 of the instructions
// and no shared-memory operand traffic, and 8 x 8 x 8 maps onto it exactly.
//
// One warp owns one cell.  A cell tensor T[x][y][z] (8 x 8 x 8) lives in the warp's registers in
// the layout of the MMA accumulator: lane = 4*x + t holds T[x][y][2t], T[x][y][2t+1] for y = 0..7
// (16 doubles per lane).  With that single layout the three kinds of 1-D contraction are:
//   z (column index):  the tile T[:, y, :] is directly a valid A operand -- its two accumulator
//                      registers are the two k-steps when the constant B operand is permuted
//                      accordingly -- so D = T[:, y, :] * M^T is 2 MMAs per tile, no data movement;
//   y (tile index):    thread-local, 64 FMAs per register pair with the matrix in the constant bank;
//   x (row index):     the tile is needed as the B operand, i.e. transposed across the warp:
//                      4 shuffles per tile, then D = M * T[:, y, :] is 2 MMAs per tile.
// All of interpolation, gradient, G, transposed gradient and transposed interpolation
// (reference src/laplacian_gpu.hpp:175-420) are done in registers; shared memory only holds the
// ring of asynchronous bulk copies (UBLKCP) that streams G, one y-tile (3 KB) per slot.
// G is stored per cell as [y][comp][s][lane] (s = z & 1), i.e. in accumulator order.
#pragma once

#include <cstdint>

#include "stiffness_slab.cuh"

namespace bd
{
struct Mma8Tables
{
  double phi[64];  // Phi[q][i], 8 x 8, column 7 zero when nd = 7 (identity when nd = 8, qmode 0)
  double dphi[64]; // dphi1[q][r], 8 x 8
};

// Storage type TS may be float: the vectors and G are then read/written as float and converted, the
// arithmetic stays fp64 on the tensor cores (there is no fp32 MMA short of TF32, and this fp64 path is
// faster than fp32 SIMT contractions for nq = 8 while being more accurate than fp32 arithmetic).
// One ring slot = one y-tile of G.  nq = 8: 6 components x 2 registers x 32 lanes in accumulator
// order.  nq = 7 (one padded lane group): only the 7 x 7 real points of the tile are stored,
// [comp][q0][q2], rounded up to a multiple of 4 scalars so every slot is a 16-byte multiple.
__host__ __device__ constexpr int mma8_slot_scalars(int nq)
{
  return nq == 8 ? 6 * 2 * 32 : (6 * nq * nq + 3) / 4 * 4;
}
template <typename TS, int NQ>
struct Mma8Ring
{
  static constexpr int STAGES = sizeof(TS) == 8 ? 4 : 8; // ~12 KB per warp either way
  static constexpr int SLOT = mma8_slot_scalars(NQ);
  static constexpr int SLOT_BYTES = SLOT * int(sizeof(TS));
  static constexpr int WARP_SMEM = STAGES * SLOT_BYTES + 128;
};
// Scalars of G per cell and position of (cell pos, comp, q0, q1, q2) in the DMMA G layout.
__host__ __device__ constexpr std::int64_t mma8_cell_scalars(int nq) { return std::int64_t(nq) * mma8_slot_scalars(nq); }
__host__ __device__ inline std::int64_t mma8_g_index(int nq, std::int64_t pos, int comp, int q0, int q1, int q2)
{
  const std::int64_t base = pos * mma8_cell_scalars(nq) + static_cast<std::int64_t>(q1) * mma8_slot_scalars(nq);
  if (nq == 8)
    return base + (comp * 2 + (q2 & 1)) * 32 + 4 * q0 + (q2 >> 1);
  return base + comp * nq * nq + q0 * nq + q2;
}

__device__ __forceinline__ void dmma(double& d0, double& d1, double a, double b)
{
  asm("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
      : "+d"(d0), "+d"(d1)
      : "d"(a), "d"(b));
}

// Contraction over the column index z: out[x][z'] = sum_z t[x][z] M(z', z), bf[s] = M(g, 2t+s).
__device__ __forceinline__ void mma8_col(double& o0, double& o1, double t0, double t1, const double (&bf)[2])
{
  o0 = 0;
  o1 = 0;
  dmma(o0, o1, t0, bf[0]);
  dmma(o0, o1, t1, bf[1]);
}

// Contraction over the row index x: out[x'][z] = sum_x M(x', x) t[x][z], af[s] = M(g, 4s+t).
__device__ __forceinline__ void mma8_row(double& o0, double& o1, double t0, double t1, const double (&af)[2],
                                         int src0, int src1, bool odd)
{
  // B operand of k-step s: lane (g, t) needs t[4s + t][g], held by lane 4*(4s+t) + (g >> 1) in
  // register (g & 1)
  const double a0 = __shfl_sync(0xffffffffu, t0, src0), a1 = __shfl_sync(0xffffffffu, t1, src0);
  const double c0 = __shfl_sync(0xffffffffu, t0, src1), c1 = __shfl_sync(0xffffffffu, t1, src1);
  o0 = 0;
  o1 = 0;
  dmma(o0, o1, af[0], odd ? a1 : a0);
  dmma(o0, o1, af[1], odd ? c1 : c0);
}

// Thread-local contraction over the tile index y: out[y'] = sum_y M[y'*8 + y] in[y] (TRANSPOSE: M[y*8 + y']).
template <bool TRANSPOSE, bool ACCUMULATE>
__device__ __forceinline__ void mma8_tile(const double (&M)[64], const double (&in)[8][2], double (&out)[8][2])
{
#pragma unroll
  for (int yo = 0; yo < 8; ++yo)
  {
    double s0 = ACCUMULATE ? out[yo][0] : 0.0, s1 = ACCUMULATE ? out[yo][1] : 0.0;
#pragma unroll
    for (int yi = 0; yi < 8; ++yi)
    {
      const double m = TRANSPOSE ? M[yi * 8 + yo] : M[yo * 8 + yi];
      s0 += m * in[yi][0];
      s1 += m * in[yi][1];
    }
    out[yo][0] = s0;
    out[yo][1] = s1;
  }
}

// PREFETCH: the next cell's dof indices are loaded while this cell is computed.  (Prefetching the input values
// too -- 16 more live doubles -- spilled and ran 2.6x slower: profiles/r2_experiments_ab.txt.)
template <typename TS, int P, int NQ, bool IDENTITY, int kMma8Warps, bool PREFETCH>
__global__ void __launch_bounds__(kMma8Warps * 32)
    stiffness_mma8_kernel(const TS* __restrict__ u, TS* __restrict__ y, const TS* __restrict__ Gf,
                          const TS* __restrict__ cell_constants,
                          const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                          int ncells, const __grid_constant__ Mma8Tables tab)
{
  constexpr int ND = P + 1;
  static_assert((NQ == 7 || NQ == 8) && ND <= NQ && IDENTITY == (ND == NQ), "nq = 7 or 8");
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  using Ring = Mma8Ring<TS, NQ>;
  constexpr int NSTAGE = Ring::STAGES;
  constexpr int SLOT = Ring::SLOT;
  constexpr int kMma8SlotBytes = Ring::SLOT_BYTES;
  constexpr int kMma8WarpSmem = Ring::WARP_SMEM;
  constexpr std::int64_t kMma8CellG = mma8_cell_scalars(NQ);
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  unsigned char* wbase = smem_raw + warp * kMma8WarpSmem;
  TS* ring = reinterpret_cast<TS*>(wbase);
  std::uint64_t* bars = reinterpret_cast<std::uint64_t*>(wbase + NSTAGE * kMma8SlotBytes);

  // constant operand fragments
  double phi_col[2], phiT_col[2], d_col[2], dT_col[2]; // B operands: M(g, 2t+s) / M^T(g, 2t+s)
  double phi_row[2], phiT_row[2], d_row[2], dT_row[2]; // A operands: M(g, 4s+t) / M^T(g, 4s+t)
#pragma unroll
  for (int s = 0; s < 2; ++s)
  {
    phi_col[s] = tab.phi[g * 8 + 2 * t + s];
    phiT_col[s] = tab.phi[(2 * t + s) * 8 + g];
    d_col[s] = tab.dphi[g * 8 + 2 * t + s];
    dT_col[s] = tab.dphi[(2 * t + s) * 8 + g];
    phi_row[s] = tab.phi[g * 8 + 4 * s + t];
    phiT_row[s] = tab.phi[(4 * s + t) * 8 + g];
    d_row[s] = tab.dphi[g * 8 + 4 * s + t];
    dT_row[s] = tab.dphi[(4 * s + t) * 8 + g];
  }
  const int src0 = 4 * t + (g >> 1), src1 = 4 * (4 + t) + (g >> 1);
  const bool odd = g & 1;

  const int warps_total = gridDim.x * kMma8Warps;
  const int first = blockIdx.x * kMma8Warps + warp;
  const int my_cells = first < ncells ? (ncells - first + warps_total - 1) / warps_total : 0;
  const int total_slots = my_cells * NQ;

  if (lane == 0)
  {
#pragma unroll
    for (int s = 0; s < NSTAGE; ++s)
      mbar_init(bars + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  const std::uint64_t g_policy = l2_evict_first_policy();
  auto issue = [&](int s) {
    const int k = s / NQ, tile = s - k * NQ;
    const std::int64_t pos = first + static_cast<std::int64_t>(k) * warps_total;
    const TS* src = Gf + pos * kMma8CellG + static_cast<std::int64_t>(tile) * SLOT;
    const int st = s % NSTAGE;
    mbar_expect_tx(bars + st, kMma8SlotBytes);
    bulk_copy_g2s_hint(ring + st * SLOT, src, kMma8SlotBytes, bars + st, g_policy);
  };
  if (lane == 0)
    for (int s = 0; s < NSTAGE - 1 && s < total_slots; ++s)
      issue(s);

  // dof indices of lane (g, t): (x = g, y, z = 2t + s); entries outside the ND^3 dofs look like
  // constrained dofs that are never touched
  auto load_dofs = [&](std::int32_t cell, std::int32_t (&d)[8][2]) {
    const std::int32_t* row = mdofmap + static_cast<std::int64_t>(cell) * ND3;
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
#pragma unroll
      for (int s = 0; s < 2; ++s)
      {
        const int z = 2 * t + s;
        const bool valid = (ND == 8) || (g < ND && yy < ND && z < ND);
        d[yy][s] = valid ? row[(g * ND + yy) * ND + z] : kBcBit;
      }
  };
  std::int32_t dof[8][2], ndof[8][2];
  // cell index and kappa are loaded one / two cells ahead (first-use stalls otherwise)
  std::int32_t cell_cur = my_cells > 0 ? cells[first] : 0;
  std::int32_t cell_next = my_cells > 1 ? cells[first + warps_total] : 0;
  double kappa = my_cells > 0 ? static_cast<double>(cell_constants[first]) : 0.0;
  if (PREFETCH && my_cells > 0)
    load_dofs(cell_cur, ndof);
  auto load_values = [&](const std::int32_t (&d)[8][2], double (&v)[8][2]) {
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
#pragma unroll
      for (int s = 0; s < 2; ++s)
      {
        const int z = 2 * t + s;
        const bool valid = (ND == 8) || (g < ND && yy < ND && z < ND);
        v[yy][s] = valid ? static_cast<double>(u[d[yy][s] & 0x7fffffff]) : 0.0;
      }
  };

  int slot = 0;
  for (int k = 0; k < my_cells; ++k)
  {
    const int pos = first + k * warps_total;
    const std::int32_t cell_next2 = k + 2 < my_cells ? cells[pos + 2 * warps_total] : 0;
    const double kappa_next = k + 1 < my_cells ? static_cast<double>(cell_constants[pos + warps_total]) : 0.0;

    // ---- gather: T[x = g][y][z = 2t + s], zero outside the ND^3 dofs and on constrained dofs
    double T[8][2];
    if constexpr (PREFETCH)
    {
#pragma unroll
      for (int yy = 0; yy < 8; ++yy)
#pragma unroll
        for (int s = 0; s < 2; ++s)
          dof[yy][s] = ndof[yy][s];
    }
    else
      load_dofs(cell_cur, dof);
    // all 16 loads are issued before the first use (a store placed between them would serialise the
    // round trips: the warp stalls in order on the store's operands)
    load_values(dof, T);
    // the next cell's indices travel while this cell is computed
    if (PREFETCH && k + 1 < my_cells)
      load_dofs(cell_next, ndof);
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
#pragma unroll
      for (int s = 0; s < 2; ++s)
      {
        const int z = 2 * t + s;
        const bool valid = (ND == 8) || (g < ND && yy < ND && z < ND);
        if (valid && dof[yy][s] < 0)
        {
          y[dof[yy][s] & 0x7fffffff] = static_cast<TS>(T[yy][s]); // b[dof] = u[dof] on constrained dofs (reference :163-166)
          T[yy][s] = 0;
        }
      }

    // ---- interpolation to the quadrature points (reference :175-211)
    if constexpr (!IDENTITY)
    {
      double W[8][2];
#pragma unroll
      for (int yy = 0; yy < 8; ++yy)
        mma8_col(W[yy][0], W[yy][1], T[yy][0], T[yy][1], phi_col); // z
      mma8_tile<false, false>(tab.phi, W, T);                      // y
#pragma unroll
      for (int yy = 0; yy < 8; ++yy)
      {
        double o0, o1;
        mma8_row(o0, o1, T[yy][0], T[yy][1], phi_row, src0, src1, odd); // x
        T[yy][0] = o0;
        T[yy][1] = o1;
      }
    }

    // ---- gradient, G, transposed gradient (reference :239-314), tile by tile
    double F[8][2]; // g_y, then f_y
    mma8_tile<false, false>(tab.dphi, T, F);
    double R[8][2];
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
    {
      if (yy >= NQ)
      {
        // padded tile (nq = 7): no quadrature points, nothing to add
        F[yy][0] = F[yy][1] = 0;
        R[yy][0] = R[yy][1] = 0;
        continue;
      }
      if (lane == 0 && slot + NSTAGE - 1 < total_slots)
        issue(slot + NSTAGE - 1);
      double gx0, gx1, gz0, gz1;
      mma8_row(gx0, gx1, T[yy][0], T[yy][1], d_row, src0, src1, odd);
      mma8_col(gz0, gz1, T[yy][0], T[yy][1], d_col);
      const int st = slot % NSTAGE;
      mbar_wait(bars + st, (slot / NSTAGE) & 1);
      double fx[2], fz[2];
#pragma unroll
      for (int s = 0; s < 2; ++s)
      {
        double G0, G1, G2, G3, G4, G5;
        if constexpr (NQ == 8)
        {
          const TS* gs = ring + st * SLOT + lane; // [comp][s][lane]
          G0 = gs[(0 * 2 + s) * 32], G1 = gs[(1 * 2 + s) * 32], G2 = gs[(2 * 2 + s) * 32];
          G3 = gs[(3 * 2 + s) * 32], G4 = gs[(4 * 2 + s) * 32], G5 = gs[(5 * 2 + s) * 32];
        }
        else
        {
          // [comp][q0 = g][q2 = 2t + s]; lanes of the padded group read nothing and get G = 0
          const bool real = g < NQ && 2 * t + s < NQ;
          const TS* gs = ring + st * SLOT + (real ? g * NQ + 2 * t + s : 0);
          G0 = real ? static_cast<double>(gs[0 * NQ * NQ]) : 0.0;
          G1 = real ? static_cast<double>(gs[1 * NQ * NQ]) : 0.0;
          G2 = real ? static_cast<double>(gs[2 * NQ * NQ]) : 0.0;
          G3 = real ? static_cast<double>(gs[3 * NQ * NQ]) : 0.0;
          G4 = real ? static_cast<double>(gs[4 * NQ * NQ]) : 0.0;
          G5 = real ? static_cast<double>(gs[5 * NQ * NQ]) : 0.0;
        }
        const double gx = s ? gx1 : gx0, gy = F[yy][s], gz = s ? gz1 : gz0;
        fx[s] = kappa * (G0 * gx + G1 * gy + G2 * gz);
        F[yy][s] = kappa * (G1 * gx + G3 * gy + G4 * gz);
        fz[s] = kappa * (G2 * gx + G4 * gy + G5 * gz);
      }
      __syncwarp(); // every lane is done with slot `st` before it is refilled
      ++slot;
      double rx0, rx1, rz0, rz1;
      mma8_row(rx0, rx1, fx[0], fx[1], dT_row, src0, src1, odd);
      mma8_col(rz0, rz1, fz[0], fz[1], dT_col);
      R[yy][0] = rx0 + rz0;
      R[yy][1] = rx1 + rz1;
    }
    mma8_tile<true, true>(tab.dphi, F, R);

    // ---- transposed interpolation (reference :388-420)
    if constexpr (!IDENTITY)
    {
      double W[8][2];
#pragma unroll
      for (int yy = 0; yy < 8; ++yy)
        mma8_col(W[yy][0], W[yy][1], R[yy][0], R[yy][1], phiT_col); // z
      mma8_tile<true, false>(tab.phi, W, R);                       // y
#pragma unroll
      for (int yy = 0; yy < 8; ++yy)
      {
        double o0, o1;
        mma8_row(o0, o1, R[yy][0], R[yy][1], phiT_row, src0, src1, odd); // x
        R[yy][0] = o0;
        R[yy][1] = o1;
      }
    }

    // ---- scatter-add (reference :424-425)
#pragma unroll
    for (int yy = 0; yy < 8; ++yy)
#pragma unroll
      for (int s = 0; s < 2; ++s)
        if (dof[yy][s] >= 0)
          atomicAdd(y + dof[yy][s], static_cast<TS>(R[yy][s]));
    cell_cur = cell_next;
    cell_next = cell_next2;
    kappa = kappa_next;
  }
}

// Geometry tensor in accumulator order from the reference layout (tests) and back.
template <typename T>
__global__ void __launch_bounds__(256)
    mma8_g_to_reference_kernel(int nq, const T* __restrict__ Gf, T* __restrict__ G, std::int64_t ncells)
{
  const int nq3 = nq * nq * nq;
  const std::int64_t total = ncells * 6 * nq3;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
       idx += stride)
  {
    const std::int64_t pos = idx / (6 * nq3);
    const int rem = static_cast<int>(idx - pos * (6 * nq3));
    const int comp = rem / nq3, q = rem % nq3;
    G[idx] = Gf[mma8_g_index(nq, pos, comp, q / (nq * nq), (q / nq) % nq, q % nq)];
  }
}

} // namespace bd
