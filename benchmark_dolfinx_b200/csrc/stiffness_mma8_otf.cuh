// DMMA stiffness kernel (stiffness_mma8.cuh) with the geometry factor computed on the fly: no G array, no ring.
//
// Same register layout and contraction scheme as stiffness_mma8_kernel; only the G stage differs.  The warp's
// cell is affine or not as a whole (one cell per warp), so the choice is a warp-uniform branch:
//   * affine cell: G(q) = w_q Gc, six scalars computed once per cell from the vertices (see
//     stiffness_slab_otf.cuh for the reasoning and the tolerance);
//   * general cell: the trilinear map's coefficients are reduced per y-tile (xi1 is the same for the whole warp in
//     a tile) by lanes 0..7 into a small per-warp table in shared memory; a point then costs 6 FMA for J plus
//     adj(J), det J and f = K (K^T g) w / det J.
// Quadrature weights are applied per point (w[q0] w[q2] per lane, w[q1] per tile): this kernel runs the
// reference's chain interpolate -> gradient -> G -> gradient^T -> interpolate^T, and the transposed gradient mixes
// the point index before the transposed interpolation sees it, so the weights cannot ride on a table as they do
// in the re-associated slab kernel.  Padded points of the nq = 7 case have weight 0.  Geometry arithmetic is fp64
// also when vectors and coordinates are stored as float.
#pragma once

#include "stiffness_mma8.cuh"
#include "stiffness_slab_otf.cuh"

namespace bd
{
struct Mma8OtfTables
{
  Mma8Tables base;
  double w[8];     // 1-D weights, 0 beyond nq
  double pts[8];   // 1-D points, 0.5 beyond nq
};

// doubles of shared memory per warp: Gc (6, padded to 8), then [tile][U0(3), U1(3), V0(3)] padded to 10
constexpr int kMma8OtfWarpDoubles = 8 + 8 * 10;

// Two doubles from shared memory, re-read at every use: keeping the cell's six G entries in shared memory instead
// of 12 registers is what lets this kernel fit the register budget of the stored-G kernel (3 CTAs of 4 warps per
// SM); `volatile` stops the compiler from hoisting the loads back into registers.
__device__ __forceinline__ double2 lds_f64x2(const double* p)
{
  double2 v;
  asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(smem_addr(p)) : "memory");
  return v;
}

template <typename TS, int P, int NQ, bool IDENTITY, int kMma8Warps, bool AFFINE_ONLY, int MINB = 1, bool PREFETCH = true,
          int FENCE_EVERY = 1>
__global__ void __launch_bounds__(kMma8Warps * 32, MINB)
    stiffness_mma8_otf_kernel(const TS* __restrict__ u, TS* __restrict__ y, const TS* __restrict__ xgeom,
                              const std::int32_t* __restrict__ gdofmap, const TS* __restrict__ cell_constants,
                              const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                              int ncells, const __grid_constant__ Mma8OtfTables otab)
{
  constexpr int ND = P + 1;
  static_assert((NQ == 7 || NQ == 8) && ND <= NQ && IDENTITY == (ND == NQ), "nq = 7 or 8");
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  const Mma8Tables& tab = otab.base;
  __shared__ __align__(16) double geo_smem[kMma8Warps * (AFFINE_ONLY ? 8 : kMma8OtfWarpDoubles)];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  __shared__ std::uint64_t fence_bars[kMma8Warps];
  std::uint64_t* fbar = fence_bars + warp;
  if (lane == 0)
    sched_fence_init(fbar);
  __syncwarp();
  double* gcs = geo_smem + warp * (AFFINE_ONLY ? 8 : kMma8OtfWarpDoubles); // the cell's kappa K K^T / det J
  [[maybe_unused]] double* geo = gcs + 8;                                  // per-tile pieces of J (general cells)

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
  // lane (g, t) owns the points (q0 = g, q1 = tile, q2 = 2t + s)
  const double wl[2] = {otab.w[g] * otab.w[2 * t], otab.w[g] * otab.w[2 * t + 1]};
  [[maybe_unused]] const double xi0 = otab.pts[g];
  [[maybe_unused]] const double xi2[2] = {otab.pts[2 * t], otab.pts[2 * t + 1]};

  const int warps_total = gridDim.x * kMma8Warps;
  const int first = blockIdx.x * kMma8Warps + warp;
  const int my_cells = first < ncells ? (ncells - first + warps_total - 1) / warps_total : 0;

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
  constexpr int NV = AFFINE_ONLY ? 4 : 8;
  auto load_vertex_ids = [&](std::int32_t cell, std::int32_t (&v)[NV]) {
    if constexpr (AFFINE_ONLY)
    {
      // vertices 0, 4, 2, 1 with loads of exactly those entries (see stiffness_slab_otf.cuh)
      const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cell) * 8;
      const int2 lo = __ldg(reinterpret_cast<const int2*>(gd));
      v[0] = lo.x, v[3] = lo.y, v[2] = __ldg(gd + 2), v[1] = __ldg(gd + 4);
    }
    else
    {
      const int4* gd = reinterpret_cast<const int4*>(gdofmap + static_cast<std::int64_t>(cell) * 8);
      const int4 lo = __ldg(gd), hi = __ldg(gd + 1);
      v[0] = lo.x, v[1] = lo.y, v[2] = lo.z, v[3] = lo.w;
      v[4] = hi.x, v[5] = hi.y, v[6] = hi.z, v[7] = hi.w;
    }
  };

  std::int32_t dof[8][2], ndof[8][2];
  std::int32_t cell_cur = my_cells > 0 ? cells[first] : 0;
  std::int32_t cell_next = my_cells > 1 ? cells[first + warps_total] : 0;
  double kappa = my_cells > 0 ? static_cast<double>(cell_constants[first]) : 0.0;
  std::int32_t vid[NV];
  if (PREFETCH && my_cells > 0)
  {
    load_dofs(cell_cur, ndof);
    load_vertex_ids(cell_cur, vid);
  }

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
    {
      load_dofs(cell_cur, dof);
      load_vertex_ids(cell_cur, vid);
    }
    load_values(dof, T);
    // the cell's vertices (every lane the same addresses: one broadcast transaction each)
    double X[NV][3];
#pragma unroll
    for (int v = 0; v < NV; ++v)
    {
      const TS* xk = xgeom + 3 * static_cast<std::int64_t>(vid[v]);
      X[v][0] = static_cast<double>(__ldg(xk));
      X[v][1] = static_cast<double>(__ldg(xk + 1));
      X[v][2] = static_cast<double>(__ldg(xk + 2));
    }
    // the next cell's indices travel while this cell is computed
    if (PREFETCH && k + 1 < my_cells)
    {
      load_dofs(cell_next, ndof);
      load_vertex_ids(cell_next, vid);
    }
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

    // ---- geometry of the cell
    double Gc[6];
    [[maybe_unused]] double col1[2][3] = {{0, 0, 0}, {0, 0, 0}};
    bool affine = true;
    if constexpr (AFFINE_ONLY)
    {
      double a[3], b[3], c[3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        a[i] = X[1][i] - X[0][i];
        b[i] = X[2][i] - X[0][i];
        c[i] = X[3][i] - X[0][i];
      }
      otf_cell_g(a, b, c, kappa, Gc);
      __syncwarp(); // every lane has finished the previous cell's G stage
      if (lane < 6)
        gcs[lane] = Gc[lane == 0 ? 0 : lane == 1 ? 1 : lane == 2 ? 2 : lane == 3 ? 3 : lane == 4 ? 4 : 5];
      __syncwarp();
    }
    else
    {
      TrilinearMap<double> map;
      map.from_vertices(X);
      affine = map.affine(static_cast<double>(otf_affine_eps<TS>()));
      __syncwarp(); // every lane has finished the previous cell's G stage
      if (affine)
      {
        otf_cell_g(map.c100, map.c010, map.c001, kappa, Gc);
        if (lane < 6)
          gcs[lane] = Gc[lane == 0 ? 0 : lane == 1 ? 1 : lane == 2 ? 2 : lane == 3 ? 3 : lane == 4 ? 4 : 5];
        __syncwarp();
      }
      else
      {
        if (lane < 8)
        {
          // tile `lane`: xi1 = pts[lane];  col0 = U0 + U1 xi2,  col2 = V0 + U1 xi0
          const double x1 = otab.pts[lane];
          double* gt = geo + lane * 10;
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            gt[i] = map.c100[i] + map.c110[i] * x1;
            gt[3 + i] = map.c101[i] + map.c111[i] * x1;
            gt[6 + i] = map.c001[i] + map.c011[i] * x1;
          }
        }
#pragma unroll
        for (int s = 0; s < 2; ++s)
#pragma unroll
          for (int i = 0; i < 3; ++i)
            col1[s][i] = (map.c010[i] + map.c110[i] * xi0) + (map.c011[i] + map.c111[i] * xi0) * xi2[s];
        __syncwarp();
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
      if (yy % FENCE_EVERY == 0)
        sched_fence(fbar); // see sched_fence
      double gx0, gx1, gz0, gz1;
      mma8_row(gx0, gx1, T[yy][0], T[yy][1], d_row, src0, src1, odd);
      mma8_col(gz0, gz1, T[yy][0], T[yy][1], d_col);
      double fx[2], fz[2];
      const double wy = otab.w[yy];
      if (AFFINE_ONLY || affine)
      {
        const double2 g01 = lds_f64x2(gcs), g23 = lds_f64x2(gcs + 2), g45 = lds_f64x2(gcs + 4);
#pragma unroll
        for (int s = 0; s < 2; ++s)
        {
          const double gx = s ? gx1 : gx0, gy = F[yy][s], gz = s ? gz1 : gz0;
          double f0 = g01.x * gx + g01.y * gy + g23.x * gz;
          double f1 = g01.y * gx + g23.y * gy + g45.x * gz;
          double f2 = g23.x * gx + g45.x * gy + g45.y * gz;
          const double wq = wl[s] * wy;
          fx[s] = f0 * wq;
          F[yy][s] = f1 * wq;
          fz[s] = f2 * wq;
        }
      }
      else
      {
        if constexpr (!AFFINE_ONLY)
        {
          const double* gt = geo + yy * 10;
          double U0[3], U1[3], col2[3];
#pragma unroll
          for (int i = 0; i < 3; ++i)
          {
            U0[i] = gt[i];
            U1[i] = gt[3 + i];
            col2[i] = gt[6 + i] + U1[i] * xi0;
          }
#pragma unroll
          for (int s = 0; s < 2; ++s)
          {
            const double gx = s ? gx1 : gx0, gy = F[yy][s], gz = s ? gz1 : gz0;
            double col0[3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
              col0[i] = U0[i] + U1[i] * xi2[s];
            double K[3][3], det;
            otf_adjugate(col0, col1[s], col2, K, det);
            const double sc = kappa * (wl[s] * wy) * otf_rcp(det);
            double f0, f1, f2;
            otf_apply_kkt(K, sc, gx, gy, gz, f0, f1, f2);
            fx[s] = f0;
            F[yy][s] = f1;
            fz[s] = f2;
          }
        }
      }
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
} // namespace bd
