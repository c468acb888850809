// Slab stiffness kernel with the geometry factor computed on the fly (no G array in memory).
//
// The reference precomputes G = K K^T w / det J at every quadrature point (geometry_computation_gpu,
// src/geometry_gpu.hpp:89-130) and streams it back in stiffness_operator_gpu (src/laplacian_gpu.hpp:256-273):
// 6 T bytes per point, ~90 % of the operator's memory traffic.  Here the same quantity is rebuilt inside the
// stiffness kernel from the 8 cell vertices (24 T + 32 bytes per cell instead of 6 T nq^3):
//
//   * The coordinate map of a hexahedron is trilinear,
//       x(xi) = c000 + c100 xi0 + c010 xi1 + c001 xi2 + c110 xi0 xi1 + c101 xi0 xi2 + c011 xi1 xi2 + c111 xi0 xi1 xi2,
//     so the columns of J are  dx/dxi0 = c100 + c110 xi1 + c101 xi2 + c111 xi1 xi2  (and cyclically); with the
//     lane's xi2 and the slice's xi1 hoisted, a point costs 6 FMA for J, then adj(J), det J and
//     f = K (K^T g) / det J  -- about 58 multiply-adds per point against 9 with a stored G.
//   * For a parallelepiped (every cell of an unperturbed box mesh) the mixed coefficients vanish, J is constant
//     and G(q) = w_q G_cell with six scalars per cell: the G stage costs what it costs with a stored G, and the
//     kernel no longer streams anything but the vectors and the dofmap.  A batch takes this shortcut when every
//     one of its cells is affine to within the rounding of the stored coordinates (decided in the kernel from the
//     vertices it has loaded anyway); the AFFINE_ONLY instantiation is used when the operator has verified at
//     construction that every cell of the mesh is.
//   * Quadrature weights: w_q1 w_q2 are folded into the tables of the transposed y/z contractions (rows of Phi /
//     DPhi scaled by the 1-D weight), w_q0 is applied per point (a second, weighted pair of tables in the middle
//     phase costs ~80 registers: with no constant-bank operand for DFMA the compiler keeps the entries of a
//     rolled loop in registers); kappa goes into G_cell resp. 1/det J.
//
// Everything else (teams of NQ lanes, register slabs, exchange cube, masked dofmap, prefetch of the next batch's
// dof indices and values) is the slab kernel of stiffness_slab.cuh; there is no ring and no mbarrier.
#pragma once

#include "stiffness_slab.cuh"

namespace bd
{
template <typename T, int ND, int NQ>
struct SlabOtfTables
{
  SlabTables<T, ND, NQ> fwd; // Phi, DPhi
  SlabTables<T, ND, NQ> bwd; // rows scaled by the 1-D quadrature weight (unscaled when Phi is the identity)
  T w[NQ];                   // 1-D weights
  T pts[NQ];                 // 1-D quadrature points on [0, 1]
};

// Tolerance of the affine test: a mixed coefficient counts as zero when it is below the rounding noise of the
// coordinates it was computed from.
template <typename T>
__host__ __device__ constexpr T otf_affine_eps()
{
  return sizeof(T) == 8 ? T(8 * 2.220446049250313e-16) : T(8 * 1.1920929e-07f);
}

// Monomial coefficients of the trilinear map from the 8 vertices X[v], v = 4a + 2b + c (tensor-product order,
// x slowest: the order of geometry_dofmap, src/laplacian.hpp:111-113).
template <typename T>
struct TrilinearMap
{
  T c100[3], c010[3], c001[3], c110[3], c101[3], c011[3], c111[3];
  T scale; // max |coordinate| of the cell's vertices

  __device__ __forceinline__ void from_vertices(const T (&X)[8][3])
  {
    T s = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
      c100[i] = X[4][i] - X[0][i];
      c010[i] = X[2][i] - X[0][i];
      c001[i] = X[1][i] - X[0][i];
      c110[i] = (X[6][i] - X[4][i]) - c010[i];
      c101[i] = (X[5][i] - X[4][i]) - c001[i];
      c011[i] = (X[3][i] - X[2][i]) - c001[i];
      c111[i] = ((X[7][i] - X[6][i]) - (X[5][i] - X[4][i])) - ((X[3][i] - X[2][i]) - c001[i]);
#pragma unroll
      for (int v = 0; v < 8; ++v)
        s = fmax(s, fabs(X[v][i]));
    }
    scale = s;
  }
  __device__ __forceinline__ bool affine() const { return affine(otf_affine_eps<T>()); }
  // `eps`: relative rounding level of the *stored* coordinates (they may be float while T is double)
  __device__ __forceinline__ bool affine(T eps) const
  {
    const T tol = eps * scale;
    T m = 0;
#pragma unroll
    for (int i = 0; i < 3; ++i)
      m = fmax(fmax(m, fabs(c110[i])), fmax(fmax(fabs(c101[i]), fabs(c011[i])), fabs(c111[i])));
    return m <= tol;
  }
};

template <typename T>
__device__ __forceinline__ T otf_rcp(T x)
{
  if constexpr (sizeof(T) == 8)
    return __drcp_rn(x);
  else
    return __frcp_rn(x);
}

// adj(J), det J for J given by its columns (J[i][0] = a[i], J[i][1] = b[i], J[i][2] = c[i]); same expressions as
// the reference (src/geometry_gpu.hpp:100-110).
template <typename T>
__device__ __forceinline__ void otf_adjugate(const T (&a)[3], const T (&b)[3], const T (&c)[3], T (&K)[3][3], T& det)
{
  K[0][0] = b[1] * c[2] - c[1] * b[2];
  K[0][1] = -b[0] * c[2] + c[0] * b[2];
  K[0][2] = b[0] * c[1] - c[0] * b[1];
  K[1][0] = -a[1] * c[2] + c[1] * a[2];
  K[1][1] = a[0] * c[2] - c[0] * a[2];
  K[1][2] = -a[0] * c[1] + c[0] * a[1];
  K[2][0] = a[1] * b[2] - b[1] * a[2];
  K[2][1] = -a[0] * b[2] + b[0] * a[2];
  K[2][2] = a[0] * b[1] - b[0] * a[1];
  det = a[0] * K[0][0] - b[0] * K[1][0] + c[0] * K[2][0];
}

// Six entries of kappa K K^T / det J in the reference's order (00, 10, 20, 11, 21, 22; src/geometry_gpu.hpp:113-130).
template <typename T>
__device__ __forceinline__ void otf_cell_g(const T (&a)[3], const T (&b)[3], const T (&c)[3], T kappa, T (&Gc)[6])
{
  T K[3][3], det;
  otf_adjugate(a, b, c, K, det);
  const T s = kappa * otf_rcp(det);
  Gc[0] = (K[0][0] * K[0][0] + K[0][1] * K[0][1] + K[0][2] * K[0][2]) * s;
  Gc[1] = (K[1][0] * K[0][0] + K[1][1] * K[0][1] + K[1][2] * K[0][2]) * s;
  Gc[2] = (K[2][0] * K[0][0] + K[2][1] * K[0][1] + K[2][2] * K[0][2]) * s;
  Gc[3] = (K[1][0] * K[1][0] + K[1][1] * K[1][1] + K[1][2] * K[1][2]) * s;
  Gc[4] = (K[2][0] * K[1][0] + K[2][1] * K[1][1] + K[2][2] * K[1][2]) * s;
  Gc[5] = (K[2][0] * K[2][0] + K[2][1] * K[2][1] + K[2][2] * K[2][2]) * s;
}

// f = s K (K^T g): the action of G = s K K^T (G[i][j] = s sum_k K[i][k] K[j][k]) without forming G.
template <typename T>
__device__ __forceinline__ void otf_apply_kkt(const T (&K)[3][3], T s, T gx, T gy, T gz, T& fx, T& fy, T& fz)
{
  const T h0 = (K[0][0] * gx + K[1][0] * gy + K[2][0] * gz) * s;
  const T h1 = (K[0][1] * gx + K[1][1] * gy + K[2][1] * gz) * s;
  const T h2 = (K[0][2] * gx + K[1][2] * gy + K[2][2] * gz) * s;
  fx = K[0][0] * h0 + K[0][1] * h1 + K[0][2] * h2;
  fy = K[1][0] * h0 + K[1][1] * h1 + K[1][2] * h2;
  fz = K[2][0] * h0 + K[2][1] * h1 + K[2][2] * h2;
}

// Z phase of lane (c, q2) for one q1: x contractions, G from the cell geometry, transposed x contractions.
// AFFINE: G = (w) Gc.  Otherwise J(q) from the hoisted column pieces:
//   col0 (constant along q0), col1 = B0 + A1 xi0, col2 = C0 + C1 xi0.
// `wrow` = w_q1 w_q2 when those weights are applied here (IDENTITY: no transposed interpolation to carry them).
template <typename T, int ND, int NQ, bool IDENTITY, bool AFFINE>
__device__ __forceinline__ void slab_z_points_otf(const SlabOtfTables<T, ND, NQ>& tab, const T (&ax)[ND],
                                                  const T (&ay)[ND], const T (&az)[ND], T (&bx)[ND], T (&by)[ND],
                                                  T (&bz)[ND], const T (&Gc)[6], const T (&col0)[3],
                                                  const T (&B0)[3], const T (&A1)[3], const T (&C0)[3],
                                                  const T (&C1)[3], T kappa, T wrow)
{
#pragma unroll
  for (int q0 = 0; q0 < NQ; ++q0)
  {
    T gx = 0, gy, gz;
    if constexpr (IDENTITY)
    {
      gy = ay[q0];
      gz = az[q0];
#pragma unroll
      for (int i = 0; i < ND; ++i)
        gx += tab.fwd.dphi[q0 * ND + i] * ax[i];
    }
    else
    {
      gy = 0;
      gz = 0;
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        gx += tab.fwd.dphi[q0 * ND + i] * ax[i];
        gy += tab.fwd.phi[q0 * ND + i] * ay[i];
        gz += tab.fwd.phi[q0 * ND + i] * az[i];
      }
    }
    T fx, fy, fz;
    if constexpr (AFFINE)
    {
      // reference src/laplacian_gpu.hpp:269-273 with G = w_q Gc
      fx = Gc[0] * gx + Gc[1] * gy + Gc[2] * gz;
      fy = Gc[1] * gx + Gc[3] * gy + Gc[4] * gz;
      fz = Gc[2] * gx + Gc[4] * gy + Gc[5] * gz;
      const T wq = IDENTITY ? wrow * tab.w[q0] : tab.w[q0];
      fx *= wq;
      fy *= wq;
      fz *= wq;
    }
    else
    {
      const T x0 = tab.pts[q0];
      T col1[3], col2[3];
#pragma unroll
      for (int i = 0; i < 3; ++i)
      {
        col1[i] = B0[i] + A1[i] * x0;
        col2[i] = C0[i] + C1[i] * x0;
      }
      T K[3][3], det;
      otf_adjugate(col0, col1, col2, K, det);
      const T s = kappa * otf_rcp(det) * (IDENTITY ? wrow * tab.w[q0] : tab.w[q0]);
      otf_apply_kkt(K, s, gx, gy, gz, fx, fy, fz);
    }
    if constexpr (IDENTITY)
    {
      by[q0] = fy;
      bz[q0] = fz;
#pragma unroll
      for (int i = 0; i < ND; ++i)
        bx[i] += tab.fwd.dphi[q0 * ND + i] * fx;
    }
    else
    {
#pragma unroll
      for (int i = 0; i < ND; ++i)
      {
        bx[i] += tab.fwd.dphi[q0 * ND + i] * fx;
        by[i] += tab.fwd.phi[q0 * ND + i] * fy;
        bz[i] += tab.fwd.phi[q0 * ND + i] * fz;
      }
    }
  }
}

// Scheduling fence.  The stored-G kernels wait on an mbarrier in every y-tile / G slot, and ptxas does not move
// instructions across that spin loop: the unrolled tiles stay in program order and the DMMA kernel needs 164
// registers.  Take the wait away (nothing to wait for here) and the tiles become one basic block that ptxas
// interleaves until 255 registers and 400+ bytes of spills; neither __syncwarp, nor an empty asm with a memory
// clobber, nor a branch on a register brings the order back (tried, compile-only).  So the on-the-fly kernels keep one
// mbarrier per warp whose first phase is completed once at start-up; waiting for that phase returns at once.
__device__ __forceinline__ void sched_fence_init(std::uint64_t* bar)
{
  mbar_init(bar, 1);
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void sched_fence(std::uint64_t* bar) { mbar_wait(bar, 0); }

template <typename T, int P, int NQ, int WARPS_>
struct SlabOtfConfig
{
  using Base = SlabConfig<T, P, NQ>;
  static constexpr int CUBE_BYTES = (3 * Base::CUBE * int(sizeof(T)) + 15) / 16 * 16;
  static constexpr int WARPS = WARPS_;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * CUBE_BYTES + WARPS * 8; // + one mbarrier per warp (sched_fence)
  static constexpr bool SUPPORTED = CTA_SMEM <= 226 * 1024;
};

#ifndef BD_SLAB_OTF_FENCE
#define BD_SLAB_OTF_FENCE 1
#endif
template <typename T, int P, int NQ, bool IDENTITY, bool AFFINE_ONLY, int WARPS_, bool PREFETCH, int MINB = 1>
__global__ void __launch_bounds__(WARPS_ * 32, MINB)
    stiffness_slab_otf_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ xgeom,
                              const std::int32_t* __restrict__ gdofmap, const T* __restrict__ cell_constants,
                              const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                              int ncells, const __grid_constant__ SlabOtfTables<T, P + 1, NQ> tab)
{
  using C = SlabConfig<T, P, NQ>;
  using OC = SlabOtfConfig<T, P, NQ, WARPS_>;
  constexpr int ND = C::ND, CW = C::CW, RL = C::RL, CUBE = C::CUBE;
  constexpr int ND2 = ND * ND, ND3 = ND2 * ND;
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / NQ;
  const int m = lane - c * NQ;
  const bool lane_ok = c < CW;
  T* cube = reinterpret_cast<T*>(smem_raw + warp * OC::CUBE_BYTES);
  const int rbase = (lane_ok ? c : 0) * NQ;
  std::uint64_t* fbar = reinterpret_cast<std::uint64_t*>(smem_raw + OC::WARPS * OC::CUBE_BYTES) + warp;
  if (lane == 0)
    sched_fence_init(fbar);
  __syncwarp();

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * OC::WARPS;
  const int first = blockIdx.x * OC::WARPS + warp;
  const int my_batches = first < nbatches ? (nbatches - first + warps_total - 1) / warps_total : 0;

  // cell of this lane's team in a batch (-1: no cell); every lane of the team gets it (the Z phase needs the geometry)
  auto cell_of = [&](int batch) -> std::int32_t {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cells[pos] : -1;
  };
  auto kappa_of = [&](int batch) -> T {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cell_constants[pos] : T(0);
  };
  auto row_from = [&](std::int32_t cell) -> const std::int32_t* {
    return (cell >= 0 && m < ND) ? mdofmap + static_cast<std::int64_t>(cell) * ND3 + m * ND2 : nullptr;
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
  // vertex indices of a cell: all 8 (general) or the 4 that span a parallelepiped (v = 0, 4, 2, 1)
  constexpr int NV = AFFINE_ONLY ? 4 : 8;
  auto load_vertex_ids = [&](std::int32_t cell, std::int32_t (&v)[NV]) {
    if (cell >= 0)
    {
      if constexpr (AFFINE_ONLY)
      {
        // exactly the four entries that are used: the unused components of a 16-byte load are dead registers,
        // the compiler reuses them at once and the next instruction then waits for the load to land (ncu: 8 % of
        // the stall samples of the first version sat on that write-after-write hazard)
        const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cell) * 8;
        const int2 lo = __ldg(reinterpret_cast<const int2*>(gd));
        v[0] = lo.x;
        v[3] = lo.y;
        v[2] = __ldg(gd + 2);
        v[1] = __ldg(gd + 4);
      }
      else
      {
        const int4* gd = reinterpret_cast<const int4*>(gdofmap + static_cast<std::int64_t>(cell) * 8);
        const int4 lo = __ldg(gd), hi = __ldg(gd + 1);
        v[0] = lo.x, v[1] = lo.y, v[2] = lo.z, v[3] = lo.w;
        v[4] = hi.x, v[5] = hi.y, v[6] = hi.z, v[7] = hi.w;
      }
    }
    else
    {
#pragma unroll
      for (int k = 0; k < NV; ++k)
        v[k] = -1;
    }
  };
  // lanes without a cell get the unit cube: finite, well conditioned, never used
  auto load_vertices = [&](const std::int32_t (&v)[NV], T (&X)[NV][3]) {
#pragma unroll
    for (int k = 0; k < NV; ++k)
    {
      if (v[k] >= 0)
      {
        const T* xk = xgeom + 3 * static_cast<std::int64_t>(v[k]);
        X[k][0] = __ldg(xk);
        X[k][1] = __ldg(xk + 1);
        X[k][2] = __ldg(xk + 2);
      }
      else
      {
        const int vv = AFFINE_ONLY ? (k == 1 ? 4 : (k == 2 ? 2 : (k == 3 ? 1 : 0))) : k;
        X[k][0] = T((vv >> 2) & 1);
        X[k][1] = T((vv >> 1) & 1);
        X[k][2] = T(vv & 1);
      }
    }
  };

  std::int32_t dof[ND2];
  T ul[ND2];
  std::int32_t cell_cur = my_batches > 0 ? cell_of(first) : -1;
  std::int32_t cell_next = my_batches > 1 ? cell_of(first + warps_total) : -1;
  T kappa = my_batches > 0 ? kappa_of(first) : T(0);
  std::int32_t vid[NV];
  load_vertex_ids(cell_cur, vid);
  if constexpr (PREFETCH)
  {
    const std::int32_t* row = row_from(cell_cur);
    load_dofs(row, dof);
    load_values(row, dof, ul);
  }

  for (int k = 0; k < my_batches; ++k)
  {
    const int batch = first + k * warps_total;
    const int pos = batch * CW + c;
    const bool active = lane_ok && pos < ncells;
    const bool xlane = active && m < ND;
    const std::int32_t* row = row_from(cell_cur);
    const std::int32_t cell_next2 = k + 2 < my_batches ? cell_of(batch + 2 * warps_total) : -1;
    const T kappa_next = k + 1 < my_batches ? kappa_of(batch + warps_total) : T(0);

    // vertices of this batch's cells: in flight during the forward phase
    T X[NV][3];
    load_vertices(vid, X);

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
      }
      if constexpr (NQ >= kSlabRolledFromNQ && !IDENTITY)
        slab_forward_x_rolled<T, ND, NQ, RL, CUBE>(tab.fwd, ul, cube + m * RL + rbase);
      else
        slab_forward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab.fwd, ul, cube + m * RL + rbase);
    }
    __syncwarp();

    // next batch's dof and vertex indices: in flight during the whole middle phase
    std::int32_t ndof[PREFETCH ? ND2 : 1];
    const std::int32_t* nrow = nullptr;
    if constexpr (PREFETCH)
    {
      nrow = row_from(cell_next);
      load_dofs(nrow, ndof);
    }
    load_vertex_ids(cell_next, vid);

    // ---- geometry of the lane's cell
    T Gc[6] = {0, 0, 0, 0, 0, 0};
    T col0[3] = {0, 0, 0}, B0[3] = {0, 0, 0}, A0[3] = {0, 0, 0}, A1[3] = {0, 0, 0};
    [[maybe_unused]] TrilinearMap<T> map;
    bool batch_affine = true;
    if constexpr (AFFINE_ONLY)
    {
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
    else
    {
      map.from_vertices(X);
      batch_affine = __all_sync(0xffffffffu, map.affine());
      if (batch_affine)
        otf_cell_g(map.c100, map.c010, map.c001, kappa, Gc);
      else
      {
        const T x2 = tab.pts[m];
#pragma unroll
        for (int i = 0; i < 3; ++i)
        {
          A0[i] = map.c100[i] + map.c101[i] * x2;
          A1[i] = map.c110[i] + map.c111[i] * x2;
          B0[i] = map.c010[i] + map.c011[i] * x2;
        }
      }
    }
    [[maybe_unused]] const T w2 = tab.w[m];

    // Middle phase (all lanes run it when the padded row is at least 32 wide, see stiffness_slab.cuh)
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
#if BD_SLAB_OTF_FENCE
      sched_fence(fbar);
#endif
      const T wrow = IDENTITY ? tab.w[q1] * w2 : T(1);
      if (zactive)
      {
        if (AFFINE_ONLY || batch_affine)
          slab_z_points_otf<T, ND, NQ, IDENTITY, true>(tab, ax, ay, az, bx, by, bz, Gc, col0, B0, A1, col0, col0,
                                                       kappa, wrow);
        else
        {
          if constexpr (!AFFINE_ONLY)
          {
            const T x1 = tab.pts[q1];
            T C0[3], C1[3];
#pragma unroll
            for (int i = 0; i < 3; ++i)
            {
              col0[i] = A0[i] + A1[i] * x1;
              C0[i] = map.c001[i] + map.c011[i] * x1;
              C1[i] = map.c101[i] + map.c111[i] * x1;
            }
            slab_z_points_otf<T, ND, NQ, IDENTITY, false>(tab, ax, ay, az, bx, by, bz, Gc, col0, B0, A1, C0, C1,
                                                          kappa, wrow);
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

    // next batch's input values: in flight during the backward phase
    T nul[PREFETCH ? ND2 : 1];
    if constexpr (PREFETCH)
      load_values(nrow, ndof, nul);

    if (xlane)
    {
      T out[ND2];
      if constexpr (NQ >= kSlabRolledFromNQ && !IDENTITY)
        slab_backward_x_rolled<T, ND, NQ, RL, CUBE>(tab.bwd, cube + m * RL + rbase, out);
      else
        slab_backward_x<T, ND, NQ, RL, CUBE, IDENTITY>(tab.bwd, cube + m * RL + rbase, out);
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
          const std::int32_t d = row[t];
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

// Setup: count[0] = cells of a list that are not affine by the kernel's own criterion; count[1] = cells that are not
// affine with mutually orthogonal edges (bricks, in any orientation), i.e. whose G_cell = kappa K K^T / det J has
// off-diagonal entries above rounding level (8 eps of the largest diagonal entry; exactly 0 on an axis-aligned box
// mesh).  Where count[1] = 0 the dof-space kernels drop the three mixed terms (stiffness_affine.cuh, DIAG).
template <typename T>
__global__ void __launch_bounds__(256)
    count_nonaffine_kernel(const T* __restrict__ xgeom, const std::int32_t* __restrict__ gdofmap,
                           const std::int32_t* __restrict__ cells, std::int64_t ncells,
                           unsigned long long* __restrict__ count)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  unsigned local = 0, local_no = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < ncells; i += stride)
  {
    const std::int32_t* gd = gdofmap + static_cast<std::int64_t>(cells[i]) * 8;
    T X[8][3];
#pragma unroll
    for (int v = 0; v < 8; ++v)
    {
      const T* xk = xgeom + 3 * static_cast<std::int64_t>(gd[v]);
      X[v][0] = xk[0];
      X[v][1] = xk[1];
      X[v][2] = xk[2];
    }
    TrilinearMap<T> map;
    map.from_vertices(X);
    const bool aff = map.affine();
    T Gc[6];
    otf_cell_g(map.c100, map.c010, map.c001, T(1), Gc);
    const T diag = fmax(fabs(Gc[0]), fmax(fabs(Gc[3]), fabs(Gc[5])));
    const T off = fmax(fabs(Gc[1]), fmax(fabs(Gc[2]), fabs(Gc[4])));
    local += aff ? 0u : 1u;
    local_no += (aff && off <= otf_affine_eps<T>() * diag) ? 0u : 1u; // NaN (degenerate cell) counts as not orthogonal
  }
  local = __reduce_add_sync(0xffffffffu, local);
  local_no = __reduce_add_sync(0xffffffffu, local_no);
  if ((threadIdx.x & 31) == 0 && local)
    atomicAdd(count, static_cast<unsigned long long>(local));
  if ((threadIdx.x & 31) == 0 && local_no)
    atomicAdd(count + 1, static_cast<unsigned long long>(local_no));
}
} // namespace bd
