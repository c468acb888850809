// Dof-space stiffness action on meshes of parallelepipeds for degrees 4..7 (nd = 5..8): the algorithm of
// stiffness_affine.cuh (y = sum over the 1-D operators M, S, C weighted by the six entries of G_cell), organised for
// planes that do not fit the register file in fp64.
//
//   * The gathered dof plane u[i][.][.] of lane (c, i) lives in shared memory (U[j*nd+k][lane]), filled with
//     asynchronous copies (cp.async: no registers, no scoreboard) straight from the input vector.
//   * The work proceeds in nd rounds, one output column k1 per round: z contraction of the whole plane for that
//     column (4 nd values per row kept in registers), y contraction row by row (8 sums), exchange of the column's
//     4 x nd values per lane through a small shared-memory buffer, x contraction for (., j, k1), scatter of nd values.
//     k is a spectator of the x contraction, so nothing larger than a column is ever exchanged or accumulated.
//   * All loops over table rows are real loops (dynamic index into the constant bank), see stiffness_affine.cuh.
// 34 nd^4 flop per cell whatever the quadrature (qmode only changes M, S, C): at degree 6 / qmode 1 that is 81.6k
// against 96.2k algorithmic / 118.8k executed by the DMMA kernel, at degree 7 / qmode 1 139k against 330k.
#pragma once

#include "stiffness_affine.cuh"

namespace bd
{
template <typename T, int P, int WARPS_, bool COMPACT, bool DIAG = false>
struct AffineColConfig
{
  using X = AffineConfig<T, P, WARPS_, 1, COMPACT, DIAG ? 2 : 4>; // exchange buffer of one column
  static constexpr int ND = P + 1, CW = 32 / ND, ND2 = ND * ND;
  static constexpr int U_BYTES = ND2 * 32 * int(sizeof(T));
  static constexpr int X_BYTES = X::WARP_BYTES;
  static constexpr int WARP_BYTES = U_BYTES + X_BYTES;
  static constexpr int WARPS = WARPS_;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_BYTES;
};

template <typename T>
__device__ __forceinline__ void cp_async_scalar(T* dst_smem, const T* src)
{
  if constexpr (sizeof(T) == 8)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_addr(dst_smem)), "l"(src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr(dst_smem)), "l"(src) : "memory");
}

// DIAG: G_cell diagonal (cells with mutually orthogonal edges), see stiffness_affine.cuh.
template <typename T, int P, int WARPS_, int MINB, bool COMPACT, bool DIAG = false>
__global__ void __launch_bounds__(WARPS_ * 32, MINB)
    stiffness_affine_col_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ xgeom,
                                const std::int32_t* __restrict__ gdofmap, const T* __restrict__ cell_constants,
                                const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                                int ncells, const __grid_constant__ AffineTables<T, P + 1> tab)
{
  using CC = AffineColConfig<T, P, WARPS_, COMPACT, DIAG>;
  using AC = typename CC::X;
  constexpr int ND = CC::ND, CW = CC::CW, ND2 = CC::ND2, ND3 = ND2 * ND, NDP = AffineTables<T, ND>::NDP;
  constexpr int SI = AC::SI, SJ = AC::SJ, PL = AC::PL;
  // row loops four iterations at a time (measured: fp64 Q6 35.4 / 38.6 / 40.3 / 40.6 GDoF/s with 1 / 2 / 4 / all, but all
  // spills at nd = 8 and loses at nd = 5, 6; profiles/r2_affine_col_ab.txt); with DIAG in fp64 at nd = 6, 7 all of them
  // (Q5 61.8 -> 63.7, Q6 53.2 -> 56.5; nd = 5: 62.3 -> 60.8 and fp32 unchanged or worse; profiles/r2_brick_ab.txt)
  constexpr int kColUnroll = (DIAG && sizeof(T) == 8 && (ND == 6 || ND == 7)) ? ND : 4;
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / ND;
  const int m = lane - c * ND;
  const bool lane_ok = c < CW;
  T* U = reinterpret_cast<T*>(smem_raw + warp * CC::WARP_BYTES) + lane; // element t of this lane's plane: U[t * 32]
  T* buf = reinterpret_cast<T*>(smem_raw + warp * CC::WARP_BYTES + CC::U_BYTES)
           + (lane_ok ? (c / AC::CH) * AC::HS + (c % AC::CH) : 0);

  T* yy = y + (static_cast<unsigned>(ncells) >> 31); // computed register value, see stiffness_affine.cuh
  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * CC::WARPS;
  const int first = blockIdx.x * CC::WARPS + warp;

  // the cell index heads the chain cell -> dofmap row -> values: requested one batch ahead
  std::int32_t cell_ahead = (lane_ok && first < nbatches && first * CW + c < ncells) ? cells[first * CW + c] : -1;
  for (int batch = first; batch < nbatches; batch += warps_total)
  {
    const int pos = batch * CW + c;
    const std::int32_t cell = cell_ahead;
    {
      const int nb = batch + warps_total, npos = nb * CW + c;
      cell_ahead = (lane_ok && nb < nbatches && npos < ncells) ? cells[npos] : -1;
    }
    const T kappa = cell >= 0 ? cell_constants[pos] : T(0);
    const std::int32_t* row = cell >= 0 ? mdofmap + static_cast<std::int64_t>(cell) * ND3 : nullptr;

    // ---- gather of the plane (i = m, ., .) into shared memory; bit t of `bcmask` = dof t is constrained
    std::uint64_t bcmask = 0;
    if (row)
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        const std::int32_t d = row[m * ND2 + t];
        cp_async_scalar(U + t * 32, u + (d & 0x7fffffff));
        bcmask |= static_cast<std::uint64_t>(d < 0) << t;
      }
    }
    else
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        U[t * 32] = T(0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    // the next batch's part of the dofmap row and its vertex ids towards L2 (no register, no scoreboard): their
    // loads at the top of the next batch then find them on chip
    if (cell_ahead >= 0)
    {
      const char* nrow = reinterpret_cast<const char*>(mdofmap + static_cast<std::int64_t>(cell_ahead) * ND3 + m * ND2);
#pragma unroll
      for (int b = 0; b < ND2 * 4 + 127; b += 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(nrow + (b < ND2 * 4 ? b : ND2 * 4 - 4)));
      if (m == 0)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(gdofmap + static_cast<std::int64_t>(cell_ahead) * 8));
    }

    // ---- geometry: Gc = kappa K K^T / det J from three edge vectors (lanes without a cell: unit cube)
    T Gc[6];
    {
      T X[4][3];
#pragma unroll
      for (int v = 0; v < 4; ++v)
      {
        X[v][0] = T(v == 1), X[v][1] = T(v == 2), X[v][2] = T(v == 3);
      }
      if (cell >= 0)
      {
        const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cell) * 8;
        const int2 lo = __ldg(reinterpret_cast<const int2*>(gd));
        const std::int32_t v0 = lo.x, v3 = lo.y, v2 = __ldg(gd + 2), v1 = __ldg(gd + 4); // vertices 0, 1(c), 2(b), 4(a)
        const std::int32_t vv[4] = {v0, v1, v2, v3};
#pragma unroll
        for (int v = 0; v < 4; ++v)
        {
          const T* xk = xgeom + 3 * static_cast<std::int64_t>(vv[v]);
          X[v][0] = __ldg(xk), X[v][1] = __ldg(xk + 1), X[v][2] = __ldg(xk + 2);
        }
      }
      T a[3], b[3], cc[3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        a[i] = X[1][i] - X[0][i];
        b[i] = X[2][i] - X[0][i];
        cc[i] = X[3][i] - X[0][i];
      }
      otf_cell_g(a, b, cc, kappa, Gc);
    }
    [[maybe_unused]] const T G00 = Gc[0], G01 = Gc[1], G02 = Gc[2], G11 = Gc[3], G12 = Gc[4], G22 = Gc[5];

    asm volatile("cp.async.wait_group 0;" ::: "memory");
    // BC masking (reference :163-169): b[dof] = u[dof], the dof contributes 0 and receives nothing (rare path)
    while (bcmask)
    {
      const int t = __ffsll(static_cast<long long>(bcmask)) - 1;
      bcmask &= bcmask - 1;
      const std::int32_t d = row[m * ND2 + t];
      y[d & 0x7fffffff] = U[t * 32];
      U[t * 32] = T(0);
    }
    __syncwarp();

    // ---- nd rounds, one output column k1 each
#pragma unroll 1
    for (int k1 = 0; k1 < ND; ++k1)
    {
      // z contraction of the whole plane for column k1: (M_z u, S_z u, C_z u, C^T_z u)[j][k1].  Unrolled over k: the
      // four table rows of k1 (4 nd entries) are fetched once per round.
      constexpr int NDX = DIAG ? 1 : ND, NDPX = DIAG ? 1 : NDP; // the C / C^T pieces do not exist with DIAG
      T mz[ND], sz[ND], cz[NDX], ctz[NDX];
#pragma unroll
      for (int j = 0; j < ND; ++j)
        mz[j] = sz[j] = 0;
#pragma unroll
      for (int j = 0; j < NDX; ++j)
        cz[j] = ctz[j] = 0;
      {
        T Mr[NDP], Sr[NDP], Cr[NDPX], Ctr[NDPX];
        load_table_row<T, NDP>(tab.M, k1, Mr);
        load_table_row<T, NDP>(tab.S, k1, Sr);
        if constexpr (!DIAG)
        {
          load_table_row<T, NDP>(tab.C, k1, Cr);
          load_table_row<T, NDP>(tab.CT, k1, Ctr);
        }
#pragma unroll
        for (int k = 0; k < ND; ++k)
        {
#pragma unroll
          for (int j = 0; j < ND; ++j)
          {
            const T v = U[(j * ND + k) * 32];
            mz[j] = fma(Mr[k], v, mz[j]);
            sz[j] = fma(Sr[k], v, sz[j]);
            if constexpr (!DIAG)
            {
              cz[j] = fma(Cr[k], v, cz[j]);
              ctz[j] = fma(Ctr[k], v, ctz[j]);
            }
          }
        }
      }
      // y contraction, one output row j1 at a time, and the Gc combinations (see stiffness_affine.cuh)
#pragma unroll kColUnroll
      for (int j1 = 0; j1 < ND; ++j1)
      {
        T p1 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0, s6 = 0, s7 = 0;
        T Mr[NDP], Sr[NDP], Cr[NDPX], Ctr[NDPX];
        load_table_row<T, NDP>(tab.M, j1, Mr);
        load_table_row<T, NDP>(tab.S, j1, Sr);
        if constexpr (!DIAG)
        {
          load_table_row<T, NDP>(tab.C, j1, Cr);
          load_table_row<T, NDP>(tab.CT, j1, Ctr);
        }
#pragma unroll
        for (int j = 0; j < ND; ++j)
        {
          const T Mj = Mr[j], Sj = Sr[j];
          p1 = fma(Mj, mz[j], p1);
          s1 = fma(Sj, mz[j], s1);
          s2 = fma(Mj, sz[j], s2);
          if constexpr (!DIAG)
          {
            const T Cj = Cr[j], Ctj = Ctr[j];
            s3 = fma(Cj, ctz[j], fma(Ctj, cz[j], s3));
            s4 = fma(Ctj, mz[j], s4);
            s5 = fma(Mj, ctz[j], s5);
            s6 = fma(Cj, mz[j], s6);
            s7 = fma(Mj, cz[j], s7);
          }
        }
        if (lane_ok)
        {
          T* dst = buf + j1 * SJ + m * SI;
          dst[0] = G00 * p1; // for S_x
          if constexpr (DIAG)
            dst[PL] = fma(G11, s1, G22 * s2); // for M_x
          else
          {
            dst[PL] = fma(G11, s1, fma(G22, s2, G12 * s3)); // for M_x
            dst[2 * PL] = fma(G01, s4, G02 * s5);           // for C_x
            dst[3 * PL] = fma(G01, s6, G02 * s7);           // for C^T_x
          }
        }
      }
      __syncwarp();

      // x contraction for (., j = m, k1) and scatter.  Four accumulators per output (one per operator) instead of one
      // nested chain: 4 nd independent multiply-adds per input plane i instead of nd chains of depth 4.
      if (lane_ok)
      {
        // scatter indices first: dof (i1, j = m, k1) at srow[i1 * ND2]; in flight during the contraction
        std::int32_t sd[ND];
        if (row)
        {
          const std::int32_t* srow = row + m * ND + k1;
#pragma unroll
          for (int i1 = 0; i1 < ND; ++i1)
            sd[i1] = srow[i1 * ND2];
        }
        else
        {
#pragma unroll
          for (int i1 = 0; i1 < ND; ++i1)
            sd[i1] = kBcBit;
        }
        T oS[ND], oM[ND], oC[NDX], oT[NDX];
#pragma unroll
        for (int i1 = 0; i1 < ND; ++i1)
          oS[i1] = oM[i1] = 0;
#pragma unroll
        for (int i1 = 0; i1 < NDX; ++i1)
          oC[i1] = oT[i1] = 0;
#pragma unroll kColUnroll
        for (int i = 0; i < ND; ++i)
        {
          const T* src = buf + m * SJ + i * SI;
          const T q1 = src[0], q2 = src[PL];
          // column i of S, M, C and row i of C, all read as rows (S, M symmetric; CT = C transposed)
          T Sr[NDP], Mr[NDP], Ctr[NDPX], Cr[NDPX];
          load_table_row<T, NDP>(tab.S, i, Sr);
          load_table_row<T, NDP>(tab.M, i, Mr);
#pragma unroll
          for (int i1 = 0; i1 < ND; ++i1)
          {
            oS[i1] = fma(Sr[i1], q1, oS[i1]);
            oM[i1] = fma(Mr[i1], q2, oM[i1]);
          }
          if constexpr (!DIAG)
          {
            const T q3 = src[2 * PL], q4 = src[3 * PL];
            load_table_row<T, NDP>(tab.CT, i, Ctr);
            load_table_row<T, NDP>(tab.C, i, Cr);
#pragma unroll
            for (int i1 = 0; i1 < ND; ++i1)
            {
              oC[i1] = fma(Ctr[i1], q3, oC[i1]);
              oT[i1] = fma(Cr[i1], q4, oT[i1]);
            }
          }
        }
        // a constrained dof receives nothing (reference :424-425): it is sent +0.0, so that the scatter is nd
        // unconditional REDs instead of nd branch regions (see stiffness_affine.cuh)
#pragma unroll
        for (int i1 = 0; i1 < ND; ++i1)
        {
          T o = oS[i1] + oM[i1];
          if constexpr (!DIAG)
            o += oC[i1] + oT[i1];
          atomicAdd(yy + (sd[i1] & 0x7fffffff), sd[i1] < 0 ? T(0) : o);
        }
      }
      __syncwarp(); // the exchange buffer is rewritten by the next round
    }
    // the next batch's gather overwrites U: every lane is done with it (last read in the last round's z contraction,
    // which precedes the __syncwarp above)
  }
}
} // namespace bd
