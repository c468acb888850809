// Stiffness action on meshes of parallelepipeds, evaluated in dof space (degree <= 3).
//
// On an affine cell the geometry factor of the reference (src/geometry_gpu.hpp:113-130) is G(q) = w_q0 w_q1 w_q2 Gc
// with a constant 3x3 matrix Gc, so the quadrature sums of stiffness_operator_gpu (src/laplacian_gpu.hpp:175-420)
//     y = sum_{d,e} Gc[d][e] B_d^T W B_e u,     B_0 = DPhi_x (x) Phi_y (x) Phi_z, ... (DPhi = dphi1 * phi0)
// factor direction by direction into three nd x nd matrices built once on the host from the same tables:
//     M = Phi^T W Phi,   S = DPhi^T W DPhi,   C = DPhi^T W Phi          (W = diag of the 1-D weights)
//     y = Gc00 S_x M_y M_z u + Gc11 M_x S_y M_z u + Gc22 M_x M_y S_z u
//       + Gc01 (C_x C^T_y M_z + C^T_x C_y M_z) u + Gc02 (C_x M_y C^T_z + C^T_x M_y C_z) u
//       + Gc12 (M_x C_y C^T_z + M_x C^T_y C_z) u.
// Same terms as the quadrature loop, summed in another order (differences at rounding level; same tests, same
// tolerances).  Grouped by the x operator it takes 13 nd^3 multiply-adds per dof plane for the y / z part and
// 4 nd^3 for x: 4352 per Q3 cell instead of the ~9300 the quadrature-space kernel executes, and the quadrature
// size drops out of the kernel altogether (qmode only changes M, S, C).
//
// Execution: a team of ND lanes owns a cell, a warp owns 32 / ND cells, every lane of a full warp works in both
// phases.  Phase A, lane (c, i): the dof plane u[i][.][.] -> four planes P1..P4[i][.][.] (y and z contracted, Gc
// folded in).  Exchange through shared memory (conflict-free strides found at compile time).  Phase B, lane
// (c, j): x contraction for the column (., j, .), then scatter-add.  Only __syncwarp between the phases.
// General (non-affine) cells never reach this kernel (the operator checks every cell at construction).
#pragma once

#include "stiffness_slab_otf.cuh"

namespace bd
{
// Rows are padded to a multiple of 16 bytes and C is also stored transposed, so that every use in the kernels reads a
// contiguous, aligned row: the compiler then fetches two (fp64) or four (fp32) entries per constant-bank load.  M and S
// are exactly symmetric (the host mirrors them), so their columns are their rows.
template <typename T, int ND>
struct alignas(16) AffineTables
{
  static constexpr int NDP = sizeof(T) == 8 ? (ND + 1) / 2 * 2 : (ND + 3) / 4 * 4;
  T M[ND * NDP];  // M[a][b] = sum_q w_q Phi[q][a] Phi[q][b]
  T S[ND * NDP];  // S[a][b] = sum_q w_q DPhi[q][a] DPhi[q][b]
  T C[ND * NDP];  // C[a][b] = sum_q w_q DPhi[q][a] Phi[q][b]
  T CT[ND * NDP]; // CT[a][b] = C[b][a]
};

// One padded table row (NDP entries, 16-byte aligned) with 16-byte constant-bank loads.
template <typename T, int NDP>
__device__ __forceinline__ void load_table_row(const T* __restrict__ table, int r, T (&out)[NDP])
{
  if constexpr (sizeof(T) == 8)
  {
    const double2* p = reinterpret_cast<const double2*>(table + r * NDP);
#pragma unroll
    for (int v = 0; v < NDP / 2; ++v)
    {
      const double2 x = p[v];
      out[2 * v] = x.x, out[2 * v + 1] = x.y;
    }
  }
  else
  {
    const float4* p = reinterpret_cast<const float4*>(table + r * NDP);
#pragma unroll
    for (int v = 0; v < NDP / 4; ++v)
    {
      const float4 x = p[v];
      out[4 * v] = x.x, out[4 * v + 1] = x.y, out[4 * v + 2] = x.z, out[4 * v + 3] = x.w;
    }
  }
}

// Unroll factor of the row loops of stiffness_affine_kernel (measured, profiles/r2_affine_ab.txt): two iterations at a time
// in fp64 (Q3 48.4 -> 50.7 GDoF/s; four spill: 33.6), everything in fp32 and at nd = 2 (Q4 fp32 70.2 -> 76.2, Q1 16.2 -> 18.3).
constexpr int affine_unroll(bool is_double, int nd, bool diag = false) { return (!is_double || nd == 2 || diag) ? 4 : 2; }
struct AffineStrides
{
  int ch, si, sj; // cells per group of the cell index, strides of i and j
};

// Layout of the exchange buffer.  Element (plane p, i, j, k, cell c) sits at
//     p*PL + (c / CH)*HS + k*SK + j*SJ + i*SI + (c % CH),      SK = nd*SJ, HS = KCH*SK, PL = ceil(CW / CH)*HS.
// Phase A writes with lanes (c, i) [fixed p, j, k], phase B reads with lanes (c, j) [fixed p, i, k]; both must be
// bank-conflict free within each group of lanes the hardware serves together (16 lanes for 8-byte words, 32 for
// 4-byte words).  With CH = CW the cell index is simply the fastest one; splitting it (CH = cells per 16-lane group)
// needs less padding when the groups coincide with whole cells (nd = 4 in fp64: 20.5 KB per warp instead of 26).
constexpr bool affine_strides_ok(int nd, bool is_double, int ch, int si, int sj, int kch)
{
  const int cw = 32 / nd;
  const int nbanks = is_double ? 16 : 32, group = is_double ? 16 : 32;
  const int hs = kch * nd * sj;
  for (int pass = 0; pass < 2; ++pass)
    for (int g0 = 0; g0 < 32; g0 += group)
    {
      bool used[32] = {};
      for (int l = g0; l < g0 + group; ++l)
      {
        const int c = l / nd, m = l % nd;
        if (c >= cw)
          continue;
        const int b = ((c / ch) * hs + m * (pass == 0 ? si : sj) + (c % ch)) % nbanks;
        if (used[b])
          return false;
        used[b] = true;
      }
    }
  return true;
}
constexpr AffineStrides affine_strides(int nd, bool is_double, int kch)
{
  const int cw = 32 / nd;
  AffineStrides best{cw, cw, nd * cw};
  int best_size = 1 << 30;
  for (int ch = 1; ch <= cw; ++ch)
  {
    if (cw % ch != 0)
      continue;
    for (int si = ch; si < ch + 33; ++si)
      for (int sj = nd * si; sj < nd * si + 33; ++sj)
        if (affine_strides_ok(nd, is_double, ch, si, sj, kch))
        {
          const int size = (cw / ch) * kch * nd * sj;
          if (size < best_size)
          {
            best_size = size;
            best = AffineStrides{ch, si, sj};
          }
          break; // larger sj only grows the buffer for this (ch, si)
        }
  }
  return best;
}

// COMPACT: no padding at all (cell index fastest, then i, then j): the phase-B reads then have 2-way bank conflicts for
// some nd, which costs little where the exchange is a small part of the work (column-round kernel) and buys occupancy.
// NPLANES: planes exchanged between the phases (4: for S_x, M_x, C_x, C^T_x; 2 when G_cell is diagonal).
template <typename T, int P, int WARPS_, int KCH_, bool COMPACT = false, int NPLANES = 4>
struct AffineConfig
{
  static constexpr int ND = P + 1;
  static constexpr int KCH = KCH_; // output columns k per exchange round (divides ND)
  static_assert(ND % KCH_ == 0, "KCH must divide nd");
  static constexpr int CW = 32 / ND;
  static constexpr AffineStrides ST = COMPACT ? AffineStrides{32 / (P + 1), 32 / (P + 1), (P + 1) * (32 / (P + 1))}
                                              : affine_strides(ND, sizeof(T) == 8, KCH_);
  static constexpr int CH = ST.ch, SI = ST.si, SJ = ST.sj, SK = ND * SJ, HS = KCH * SK, PL = (CW / CH) * HS;
  static constexpr int WARP_BYTES = (NPLANES * PL * int(sizeof(T)) + 15) / 16 * 16;
  static constexpr int WARPS = WARPS_;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_BYTES;
};

// DIAG: every cell has mutually orthogonal edges (bricks; the operator checks at construction), so G_cell is diagonal and
// only  y = Gc00 S_x M_y M_z u + Gc11 M_x S_y M_z u + Gc22 M_x M_y S_z u  is left: 7 instead of 17 applications of an
// nd x nd matrix per dof plane (14 nd^4 flop per cell), two exchanged planes instead of four.
// SDOF_EARLY: the scatter indices are requested at the top of the batch instead of between the phases (with DIAG phase
// B is too short to cover them: Q3 fp64 62.9 -> 69.3 GDoF/s).
template <typename T, int P, int WARPS_, int MINB, int KCH, bool DIAG = false, bool SDOF_EARLY = false>
__global__ void __launch_bounds__(WARPS_ * 32, MINB)
    stiffness_affine_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ xgeom,
                            const std::int32_t* __restrict__ gdofmap, const T* __restrict__ cell_constants,
                            const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                            int ncells, const __grid_constant__ AffineTables<T, P + 1> tab)
{
  using AC = AffineConfig<T, P, WARPS_, KCH, false, DIAG ? 2 : 4>;
  constexpr int ND = AC::ND, CW = AC::CW, SI = AC::SI, SJ = AC::SJ, SK = AC::SK, PL = AC::PL;
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND, NDP = AffineTables<T, ND>::NDP;
  constexpr bool DEEP = sizeof(T) == 8;
  constexpr int kAffineUnroll = affine_unroll(sizeof(T) == 8, ND, DIAG);
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / ND;
  const int m = lane - c * ND;
  const bool lane_ok = c < CW;
  T* buf = reinterpret_cast<T*>(smem_raw + warp * AC::WARP_BYTES)
           + (lane_ok ? (c / AC::CH) * AC::HS + (c % AC::CH) : 0);

  // The output pointer as a computed register value (ncells is never negative): left to itself ptxas re-reads the
  // parameter from the constant bank before every atomic (nd^2 LDC per batch on the long scoreboard).
  T* yy = y + (static_cast<unsigned>(ncells) >> 31);
  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * AC::WARPS;
  const int first = blockIdx.x * AC::WARPS + warp;
  const int my_batches = first < nbatches ? (nbatches - first + warps_total - 1) / warps_total : 0;

  auto cell_of = [&](int batch) -> std::int32_t {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cells[pos] : -1;
  };
  auto kappa_of = [&](int batch) -> T {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cell_constants[pos] : T(0);
  };
  // phase A gathers the plane (i = m, ., .): entries m*ND2 + j*ND + k of the cell's row
  auto load_dofs = [&](std::int32_t cell, std::int32_t (&d)[ND2]) {
    const std::int32_t* row = cell >= 0 ? mdofmap + static_cast<std::int64_t>(cell) * ND3 + m * ND2 : nullptr;
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      d[t] = row ? row[t] : kBcBit; // lanes without a cell look like constrained dofs
  };
  auto load_values = [&](std::int32_t cell, const std::int32_t (&d)[ND2], T (&v)[ND2]) {
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      v[t] = cell >= 0 ? u[d[t] & 0x7fffffff] : T(0);
  };
  // the four vertices that span a parallelepiped: v = 0, 4, 2, 1 of the tensor-product order
  // (Loads of exactly the entries that are used: with 16-byte loads the unused components are dead registers, the
  // compiler reuses them at once and the next instruction then waits for the load to land -- 15 % of the stall
  // samples of the first version sat on such a write-after-write hazard.)
  auto load_vertex_ids = [&](std::int32_t cell, std::int32_t (&v)[4]) {
    if (cell >= 0)
    {
      const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cell) * 8;
      const int2 lo = __ldg(reinterpret_cast<const int2*>(gd));
      v[0] = lo.x, v[3] = lo.y, v[2] = __ldg(gd + 2), v[1] = __ldg(gd + 4);
    }
    else
      v[0] = v[1] = v[2] = v[3] = -1;
  };
  auto load_vertices = [&](const std::int32_t (&v)[4], T (&X)[4][3]) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
      if (v[k] >= 0)
      {
        const T* xk = xgeom + 3 * static_cast<std::int64_t>(v[k]);
        X[k][0] = __ldg(xk), X[k][1] = __ldg(xk + 1), X[k][2] = __ldg(xk + 2);
      }
      else
      {
        X[k][0] = T(k == 1), X[k][1] = T(k == 2), X[k][2] = T(k == 3); // lanes without a cell: unit cube
      }
    }
  };
  // Gc = kappa K K^T / det J from the three edge vectors, order 00, 10, 20, 11, 21, 22
  auto cell_g = [&](const T (&X)[4][3], T kap, T (&G)[6]) {
    T a[3], b[3], cc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
      a[i] = X[1][i] - X[0][i];
      b[i] = X[2][i] - X[0][i];
      cc[i] = X[3][i] - X[0][i];
    }
    otf_cell_g(a, b, cc, kap, G);
  };

  // Software pipeline over the warp's batches: dof and vertex indices are requested two batches ahead, input values
  // and vertex coordinates one batch ahead (at the top of the batch, so that they have the whole batch to arrive:
  // requested after phase A, as in the first version, 15 % of the stall samples sat on their first use).
  std::int32_t dof[ND2], ndof[ND2], vid[4];
  T ul[ND2];
  std::int32_t cell_cur = my_batches > 0 ? cell_of(first) : -1;
  std::int32_t cell_next = my_batches > 1 ? cell_of(first + warps_total) : -1;
  // DEEP: the cell index itself is requested three batches ahead -- it heads the chain cell -> dofmap row -> values, and
  // read where it is needed its latency is exposed once per batch (18.6 % of the stall samples of the diagonal-G kernel)
  [[maybe_unused]] std::int32_t cell_next2_reg = (DEEP && my_batches > 2) ? cell_of(first + 2 * warps_total) : -1;
  T kappa = my_batches > 0 ? kappa_of(first) : T(0);
  load_vertex_ids(cell_cur, vid);
  load_dofs(cell_cur, dof);
  if constexpr (DEEP)
    load_dofs(cell_next, ndof);
  load_values(cell_cur, dof, ul);
  T Gc[6]; // geometry of the current batch's cell, computed one batch ahead
  {
    T X[4][3];
    load_vertices(vid, X);
    cell_g(X, kappa, Gc);
  }
  if constexpr (DEEP)
    load_vertex_ids(cell_next, vid);

  for (int kb = 0; kb < my_batches; ++kb)
  {
    const int batch = first + kb * warps_total;
    std::int32_t cell_next2, cell_next3 = -1;
    if constexpr (DEEP)
    {
      cell_next2 = cell_next2_reg;
      cell_next3 = kb + 3 < my_batches ? cell_of(batch + 3 * warps_total) : -1;
    }
    else
      cell_next2 = kb + 2 < my_batches ? cell_of(batch + 2 * warps_total) : -1;
    const T kappa_next = kb + 1 < my_batches ? kappa_of(batch + warps_total) : T(0);

    [[maybe_unused]] const T G00 = Gc[0], G01 = Gc[1], G02 = Gc[2], G11 = Gc[3], G12 = Gc[4], G22 = Gc[5];

    // BC masking (reference :163-169): b[dof] = u[dof], the dof contributes 0 and receives nothing
    if (cell_cur >= 0)
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        if (dof[t] < 0)
        {
          y[dof[t] & 0x7fffffff] = ul[t];
          ul[t] = 0;
        }
    }
    // next batch's input values and vertices (their indices arrived during the previous batch), and the indices of
    // the batch after it
    // (DEEP = fp64.  In fp32 the extra index set costs more than the earlier request gains -- Q4 fp32 70.2 vs 58.8
    // GDoF/s -- so there the values and vertices are requested after phase A and the indices one batch ahead.)
    T nul[ND2];
    T Xn[4][3];
    std::int32_t n2dof[DEEP ? ND2 : 1];
    if constexpr (DEEP)
    {
      load_values(cell_next, ndof, nul);
      load_vertices(vid, Xn);
      load_dofs(cell_next2, n2dof);
      load_vertex_ids(cell_next2, vid);
    }
    else
    {
      load_dofs(cell_next, ndof);
      load_vertex_ids(cell_next, vid);
    }

    // ---- The two phases, in rounds of KCH output columns k (k is a spectator of the x contraction, so the exchange
    // buffer only ever holds KCH columns: 4 planes x KCH x nd^2 values per cell instead of 4 nd^3, which is what
    // lets 16 one-warp CTAs share an SM).
    //
    // Phase A (y and z contractions of the plane, one output column k1 at a time).  The loops over j1 and, for
    // KCH > 1, k1 are real loops on purpose: unrolled, the 3 nd^2 table entries (more than the uniform register
    // file holds in fp64) end up parked in ordinary registers next to 4 nd^2 accumulators and the kernel spills at
    // 255 registers (first version: 30.6 GDoF/s at Q3, no better than the quadrature-space kernel).  Rolled, a
    // column needs the plane, 4 nd z values and 8 sums; table rows come from the constant bank by dynamic index.
    T out[ND2];
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      out[t] = 0;
    std::int32_t sdof[ND2]; // scatter indices of phase B, [i1][k]
    if constexpr (SDOF_EARLY)
    {
      const std::int32_t* row = cell_cur >= 0 ? mdofmap + static_cast<std::int64_t>(cell_cur) * ND3 + m * ND : nullptr;
#pragma unroll
      for (int i1 = 0; i1 < ND; ++i1)
#pragma unroll
        for (int k = 0; k < ND; ++k)
          sdof[i1 * ND + k] = row ? row[i1 * ND2 + k] : kBcBit;
    }
#pragma unroll
    for (int r = 0; r < ND / KCH; ++r)
    {
      if (r > 0)
        __syncwarp(); // the previous round's phase B is done with the buffer
#pragma unroll 1
      for (int kk = 0; kk < KCH; ++kk)
      {
        const int k1 = r * KCH + kk;
        T mz[ND], sz[ND], cz[DIAG ? 1 : ND], ctz[DIAG ? 1 : ND]; // (M_z u, S_z u, C_z u, C^T_z u)[j][k1]
        {
          T Mr[NDP], Sr[NDP], Cr[DIAG ? 1 : NDP], Ctr[DIAG ? 1 : NDP];
          load_table_row<T, NDP>(tab.M, k1, Mr);
          load_table_row<T, NDP>(tab.S, k1, Sr);
          if constexpr (!DIAG)
          {
            load_table_row<T, NDP>(tab.C, k1, Cr);
            load_table_row<T, NDP>(tab.CT, k1, Ctr);
          }
#pragma unroll
          for (int j = 0; j < ND; ++j)
          {
            T a0 = 0, a1 = 0, a2 = 0, a3 = 0;
#pragma unroll
            for (int k = 0; k < ND; ++k)
            {
              const T v = ul[j * ND + k];
              a0 = fma(Mr[k], v, a0);
              a1 = fma(Sr[k], v, a1);
              if constexpr (!DIAG)
              {
                a2 = fma(Cr[k], v, a2);  // C_z
                a3 = fma(Ctr[k], v, a3); // C^T_z
              }
            }
            mz[j] = a0, sz[j] = a1;
            if constexpr (!DIAG)
              cz[j] = a2, ctz[j] = a3;
          }
        }
#pragma unroll kAffineUnroll
        for (int j1 = 0; j1 < ND; ++j1)
        {
          T p1 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, s5 = 0, s6 = 0, s7 = 0;
          T Mr[NDP], Sr[NDP], Cr[DIAG ? 1 : NDP], Ctr[DIAG ? 1 : NDP];
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
            p1 = fma(Mj, mz[j], p1); // M_y M_z
            s1 = fma(Sj, mz[j], s1); // S_y M_z
            s2 = fma(Mj, sz[j], s2); // M_y S_z
            if constexpr (!DIAG)
            {
              const T Cj = Cr[j], Ctj = Ctr[j];
              s3 = fma(Cj, ctz[j], fma(Ctj, cz[j], s3)); // C_y C^T_z + C^T_y C_z
              s4 = fma(Ctj, mz[j], s4);                  // C^T_y M_z
              s5 = fma(Mj, ctz[j], s5);                  // M_y C^T_z
              s6 = fma(Cj, mz[j], s6);                   // C_y M_z
              s7 = fma(Mj, cz[j], s7);                   // M_y C_z
            }
          }
          if (lane_ok)
          {
            T* dst = buf + kk * SK + j1 * SJ + m * SI;
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
      }
      __syncwarp();

      if (r == 0)
      {
        // this lane's scatter indices (column (., j = m, .): a second look at the cell's row, which the team has
        // just read): in flight during phase B
        if constexpr (!DEEP)
        {
          load_values(cell_next, ndof, nul);
          load_vertices(vid, Xn);
        }
        if constexpr (!SDOF_EARLY)
        {
          const std::int32_t* row =
              cell_cur >= 0 ? mdofmap + static_cast<std::int64_t>(cell_cur) * ND3 + m * ND : nullptr;
#pragma unroll
          for (int i1 = 0; i1 < ND; ++i1)
#pragma unroll
            for (int k = 0; k < ND; ++k)
              sdof[i1 * ND + k] = row ? row[i1 * ND2 + k] : kBcBit;
        }
      }

      // Phase B: x contraction for the column j = m and the round's k columns (loop over the input plane i rolled)
      if (lane_ok)
      {
#pragma unroll kAffineUnroll
        for (int i = 0; i < ND; ++i)
        {
          T q1[KCH], q2[KCH], q3[DIAG ? 1 : KCH], q4[DIAG ? 1 : KCH];
#pragma unroll
          for (int kk = 0; kk < KCH; ++kk)
          {
            const T* src = buf + kk * SK + m * SJ + i * SI;
            q1[kk] = src[0], q2[kk] = src[PL];
            if constexpr (!DIAG)
              q3[kk] = src[2 * PL], q4[kk] = src[3 * PL];
          }
          // column i of S, M, C and row i of C: all read as rows (S, M symmetric; CT = C transposed)
          T Sr[NDP], Mr[NDP], Ctr[DIAG ? 1 : NDP], Cr[DIAG ? 1 : NDP];
          load_table_row<T, NDP>(tab.S, i, Sr);
          load_table_row<T, NDP>(tab.M, i, Mr);
          if constexpr (!DIAG)
          {
            load_table_row<T, NDP>(tab.CT, i, Ctr);
            load_table_row<T, NDP>(tab.C, i, Cr);
          }
#pragma unroll
          for (int i1 = 0; i1 < ND; ++i1)
          {
            const T Si = Sr[i1], Mi = Mr[i1];
#pragma unroll
            for (int kk = 0; kk < KCH; ++kk)
            {
              T& o = out[i1 * ND + r * KCH + kk];
              if constexpr (DIAG)
                o = fma(Si, q1[kk], fma(Mi, q2[kk], o));
              else
                o = fma(Si, q1[kk], fma(Mi, q2[kk], fma(Ctr[i1], q3[kk], fma(Cr[i1], q4[kk], o))));
            }
          }
        }
      }
    }
    if (lane_ok)
    {
      // A constrained dof receives nothing (reference :424-425): it is sent +0.0 instead, so that the scatter is nd^2
      // unconditional REDs -- written as `if (d >= 0) atomicAdd(...)` each one is its own branch region (BSSY / BRA
      // / RED / BSYNC).  Constrained dofs are boundary dofs: few, and y[d] + 0.0 = y[d].
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        atomicAdd(yy + (sdof[t] & 0x7fffffff), sdof[t] < 0 ? T(0) : out[t]);
    }
    __syncwarp(); // the buffer is rewritten by the next batch's phase A

#pragma unroll
    for (int t = 0; t < ND2; ++t)
    {
      dof[t] = ndof[t];
      if constexpr (DEEP)
        ndof[t] = n2dof[t];
      ul[t] = nul[t];
    }
    cell_cur = cell_next;
    cell_next = cell_next2;
    cell_next2_reg = cell_next3;
    kappa = kappa_next;
    cell_g(Xn, kappa, Gc);
  }
}
} // namespace bd
