// Dof-space stiffness action on meshes of parallelepipeds (degree <= 3, degree 4 in fp32): the arithmetic of
// stiffness_affine.cuh (phases A and B, the 1-D operators M, S, C, the DIAG variant for bricks) behind a gather
// pipeline that runs on asynchronous copies instead of registers.
//
// Why: with the diagonal-G variant a batch is short, and the register pipeline of stiffness_affine_kernel (indices two
// batches ahead, values one: 3 x nd^2 index registers + 2 x nd^2 values) is what limits it -- 196 registers = 2 warps
// per scheduler, 39 % of the stall samples on the long scoreboard (loads that share a scoreboard with the prefetches
// issued just before).  Here, per warp in shared memory:
//   IB[3]  the dofmap rows of the batch in flight, the next one and the one after (cp.async, 16-byte pieces at nd = 4),
//   UB     the gathered input values of the next batch (cp.async, one per dof, addresses read from IB),
//   X      the exchange buffer between the phases (as before).
// No load of the pipeline touches a register or a scoreboard; one cp.async.wait_group + __syncwarp per batch.  The
// scatter indices (the same row, seen by the lane that owns column j instead of plane i) are read back from IB
// transposed, so the row is fetched from global memory once.  The cells' vertex coordinates take the same route (12
// scalars per cell, shared out over the lanes of the team).  What is left in registers: the plane and the accumulators.
#pragma once

#include "stiffness_affine_col.cuh"

namespace bd
{
// KCH: output columns k per exchange round (divides nd): the exchange buffer holds KCH columns of the planes.
template <typename T, int P, int WARPS_, bool DIAG, int KCH>
struct AffinePipeConfig
{
  using X = AffineConfig<T, P, WARPS_, KCH, false, DIAG ? 2 : 4>;
  static constexpr int ND = P + 1, ND2 = ND * ND, CW = 32 / ND;
  // nd = 4: a lane's part of the row is 64 aligned bytes = four 16-byte pieces [j][0..3]; piece v of lane l sits at
  // int4 slot v * 33 + l (33: the transposed read of the scatter indices is then conflict-free as well)
  static constexpr bool IDX16 = ND == 4;
  static constexpr int X_BYTES = X::WARP_BYTES;
  static constexpr int U_BYTES = ND2 * 32 * int(sizeof(T));
  static constexpr int I_INTS = IDX16 ? (ND2 / 4) * 33 * 4 : ND2 * 32;
  static constexpr int I_BYTES = (I_INTS * 4 + 15) / 16 * 16;
  static constexpr int V_BYTES = (2 * 12 * CW * int(sizeof(T)) + 15) / 16 * 16; // vertex coordinates, two batches
  static constexpr int WARP_BYTES = X_BYTES + U_BYTES + 3 * I_BYTES + V_BYTES;
  static constexpr int WARPS = WARPS_;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int CTA_SMEM = WARPS * WARP_BYTES;
};

__device__ __forceinline__ void cp_async_16(void* dst_smem, const void* src)
{
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr(dst_smem)), "l"(src) : "memory");
}

template <typename T, int P, int WARPS_, int MINB, bool DIAG, int KCH>
__global__ void __launch_bounds__(WARPS_ * 32, MINB)
    stiffness_affine_pipe_kernel(const T* __restrict__ u, T* __restrict__ y, const T* __restrict__ xgeom,
                                 const std::int32_t* __restrict__ gdofmap, const T* __restrict__ cell_constants,
                                 const std::int32_t* __restrict__ mdofmap, const std::int32_t* __restrict__ cells,
                                 const std::int8_t* __restrict__ merge, int ncells,
                                 const __grid_constant__ AffineTables<T, P + 1> tab)
{
  using PC = AffinePipeConfig<T, P, WARPS_, DIAG, KCH>;
  using AC = typename PC::X;
  constexpr int ND = PC::ND, CW = PC::CW, ND2 = PC::ND2, ND3 = ND2 * ND, NDP = AffineTables<T, ND>::NDP;
  constexpr int SI = AC::SI, SJ = AC::SJ, SK = AC::SK, PL = AC::PL;
  constexpr bool IDX16 = PC::IDX16;
  constexpr int kAffineUnroll = affine_unroll(sizeof(T) == 8, ND, DIAG);
  constexpr int NDX = DIAG ? 1 : ND, NDPX = DIAG ? 1 : NDP; // the C / C^T pieces do not exist with DIAG
  extern __shared__ __align__(128) unsigned char smem_raw[];

  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  const int c = lane / ND;
  const int m = lane - c * ND;
  const bool lane_ok = c < CW;
  unsigned char* wbase = smem_raw + warp * PC::WARP_BYTES;
  T* buf = reinterpret_cast<T*>(wbase) + (lane_ok ? (c / AC::CH) * AC::HS + (c % AC::CH) : 0);
  T* UB = reinterpret_cast<T*>(wbase + PC::X_BYTES) + lane; // value t of this lane's plane: UB[t * 32]
  std::int32_t* ib_cur = reinterpret_cast<std::int32_t*>(wbase + PC::X_BYTES + PC::U_BYTES);
  std::int32_t* ib_n1 = ib_cur + PC::I_BYTES / 4;
  std::int32_t* ib_n2 = ib_n1 + PC::I_BYTES / 4;
  // vertex coordinates of the cell's four spanning vertices, [vertex][component]; the team shares one copy
  T* vb_cur = reinterpret_cast<T*>(wbase + PC::X_BYTES + PC::U_BYTES + 3 * PC::I_BYTES) + 12 * (lane_ok ? c : 0);
  T* vb_n1 = vb_cur + 12 * CW;
  // The output pointer as a computed register value (ncells is never negative): left to itself ptxas re-reads the
  // parameter from the constant bank before every atomic (nd^2 LDC per batch, 9 % of the stall samples waiting for
  // them).
  T* yy = y + (static_cast<unsigned>(ncells) >> 31);

  const int nbatches = (ncells + CW - 1) / CW;
  const int warps_total = gridDim.x * PC::WARPS;
  const int first = blockIdx.x * PC::WARPS + warp;
  const int my_batches = first < nbatches ? (nbatches - first + warps_total - 1) / warps_total : 0;

  auto cell_of = [&](int batch) -> std::int32_t {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cells[pos] : -1;
  };
  auto kappa_of = [&](int batch) -> T {
    const int pos = batch * CW + c;
    return (lane_ok && batch < nbatches && pos < ncells) ? cell_constants[pos] : T(0);
  };
  // does this lane's cell hand its k = nd - 1 face to the next cell of the batch? (build_merge_flags; none without)
  auto merge_of = [&](int batch) -> int {
    const int pos = batch * CW + c;
    return (merge && lane_ok && batch < nbatches && pos < ncells) ? merge[pos] : 0;
  };
  // slot of entry t = j * nd + k of lane l's plane in an index buffer
  auto islot = [&](int t, int l) -> int { return IDX16 ? ((t / 4) * 33 + l) * 4 + (t % 4) : t * 32 + l; };

  // the lane's part of the cell's dofmap row (plane i = m) -> index buffer, asynchronously
  auto request_indices = [&](std::int32_t* ib, std::int32_t cell) {
    if (cell >= 0)
    {
      const std::int32_t* row = mdofmap + static_cast<std::int64_t>(cell) * ND3 + m * ND2;
      if constexpr (IDX16)
      {
#pragma unroll
        for (int v = 0; v < ND2 / 4; ++v)
          cp_async_16(ib + (v * 33 + lane) * 4, row + 4 * v);
      }
      else
      {
#pragma unroll
        for (int t = 0; t < ND2; ++t)
          cp_async_scalar(ib + islot(t, lane), row + t);
      }
    }
    else
    {
      // lanes without a cell look like constrained dofs (nothing gathered, nothing scattered)
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        ib[islot(t, lane)] = kBcBit;
    }
  };
  // input values of the plane -> UB, asynchronously, addresses from an index buffer that has arrived; returns the mask
  // of constrained dofs (bit t)
  auto request_values = [&](const std::int32_t* ib, std::int32_t cell) -> std::uint32_t {
    std::uint32_t bcmask = 0;
    if (cell >= 0)
    {
      // all index reads first: the copies are volatile asm with a memory clobber, reads placed between them would be
      // serialised with them (each read followed at once by its consumer)
      std::int32_t dd[ND2];
      if constexpr (IDX16)
      {
#pragma unroll
        for (int v = 0; v < ND2 / 4; ++v)
        {
          const int4 d = *reinterpret_cast<const int4*>(ib + (v * 33 + lane) * 4);
          dd[4 * v] = d.x, dd[4 * v + 1] = d.y, dd[4 * v + 2] = d.z, dd[4 * v + 3] = d.w;
        }
      }
      else
      {
#pragma unroll
        for (int t = 0; t < ND2; ++t)
          dd[t] = ib[islot(t, lane)];
      }
#pragma unroll
      for (int t = 0; t < ND2; ++t)
      {
        cp_async_scalar(UB + t * 32, u + (dd[t] & 0x7fffffff));
        bcmask |= static_cast<std::uint32_t>(dd[t] < 0) << t;
      }
    }
    else
    {
#pragma unroll
      for (int t = 0; t < ND2; ++t)
        UB[t * 32] = T(0);
    }
    return bcmask;
  };
  // the four vertices that span a parallelepiped: v = 0, 4, 2, 1 of the tensor-product order (exact-width loads, see
  // stiffness_affine.cuh)
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
  // the 12 coordinates of those vertices -> vertex buffer, asynchronously, shared out over the lanes of the team
  // (lanes without a cell: unit cube)
  auto request_vertices = [&](T* vb, const std::int32_t (&v)[4]) {
    if (!lane_ok)
      return;
#pragma unroll
    for (int s0 = 0; s0 < 12; s0 += ND)
    {
      const int s = s0 + m;
      if (s < 12)
      {
        const int k = s / 3, comp = s - 3 * k;
        const std::int32_t vk = k == 0 ? v[0] : k == 1 ? v[1] : k == 2 ? v[2] : v[3];
        if (vk >= 0)
          cp_async_scalar(vb + s, xgeom + 3 * static_cast<std::int64_t>(vk) + comp);
        else
          vb[s] = T(comp + 1 == k);
      }
    }
  };
  // Gc = kappa K K^T / det J from the three edge vectors in a vertex buffer that has arrived
  auto cell_g = [&](const T* vb, T kap, T (&G)[6]) {
    T X[12];
#pragma unroll
    for (int s = 0; s < 12; ++s)
      X[s] = vb[s];
    T a[3], b[3], cc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i)
    {
      a[i] = X[3 + i] - X[i];
      b[i] = X[6 + i] - X[i];
      cc[i] = X[9 + i] - X[i];
    }
    otf_cell_g(a, b, cc, kap, G);
  };

  // ---- prologue: rows of the first two batches, values of the first, geometry of the first
  std::int32_t cell_cur = cell_of(first);
  std::int32_t cell_n1 = cell_of(first + warps_total);
  std::int32_t cell_n2 = cell_of(first + 2 * warps_total);
  std::int32_t vid[4];
  load_vertex_ids(cell_cur, vid);
  request_indices(ib_cur, cell_cur);
  request_indices(ib_n1, cell_n1);
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp();
  std::uint32_t bcmask = request_values(ib_cur, cell_cur);
  request_vertices(vb_cur, vid);
  asm volatile("cp.async.commit_group;" ::: "memory");
  load_vertex_ids(cell_n1, vid);
  T kappa = kappa_of(first);
  int mflag = merge_of(first);

  for (int kb = 0; kb < my_batches; ++kb)
  {
    const int batch = first + kb * warps_total;
    // the cell index heads the chain cell -> row -> values: three batches ahead
    const std::int32_t cell_n3 = cell_of(batch + 3 * warps_total);
    const T kappa_next = kappa_of(batch + warps_total);
    const int mflag_next = merge_of(batch + warps_total);
    // this batch's values and vertices and the next batch's row have arrived (every lane's copies: the row is read
    // transposed, the vertices by the whole team)
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    T Gc[6];
    cell_g(vb_cur, kappa, Gc);
    [[maybe_unused]] const T G00 = Gc[0], G01 = Gc[1], G02 = Gc[2], G11 = Gc[3], G12 = Gc[4], G22 = Gc[5];

    // BC masking (reference :163-169): b[dof] = u[dof], the dof contributes 0 and receives nothing (rare path)
    while (bcmask)
    {
      const int t = __ffs(static_cast<int>(bcmask)) - 1;
      bcmask &= bcmask - 1;
      const std::int32_t d = ib_cur[islot(t, lane)];
      y[d & 0x7fffffff] = UB[t * 32];
      UB[t * 32] = T(0);
    }
    T ul[ND2];
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      ul[t] = UB[t * 32];

    // next batch's values, the row of the batch after it, next batch's vertices, vertex ids of the batch after it
    const std::uint32_t bcmask_next = request_values(ib_n1, cell_n1);
    request_indices(ib_n2, cell_n2);
    request_vertices(vb_n1, vid);
    asm volatile("cp.async.commit_group;" ::: "memory");
    load_vertex_ids(cell_n2, vid);

    // ---- Phase A (y and z contractions of the plane, one output column k1 at a time; rolled loops with dynamic
    // table rows, see stiffness_affine.cuh)
    T out[ND2];
#pragma unroll
    for (int t = 0; t < ND2; ++t)
      out[t] = 0;
#pragma unroll
    for (int r = 0; r < ND / KCH; ++r)
    {
    if (r > 0)
      __syncwarp(); // the previous round's phase B is done with the buffer
#pragma unroll 1
    for (int kk = 0; kk < KCH; ++kk)
    {
      const int k1 = r * KCH + kk;
      T mz[ND], sz[ND], cz[NDX], ctz[NDX]; // (M_z u, S_z u, C_z u, C^T_z u)[j][k1]
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
              a2 = fma(Cr[k], v, a2);
              a3 = fma(Ctr[k], v, a3);
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

    // ---- Phase B: x contraction for the column j = m (loop over the input plane i rolled), then scatter-add
    if (lane_ok)
    {
#pragma unroll kAffineUnroll
      for (int i = 0; i < ND; ++i)
      {
        constexpr int KX = DIAG ? 1 : KCH;
        T q1[KCH], q2[KCH], q3[KX], q4[KX];
#pragma unroll
        for (int k = 0; k < KCH; ++k)
        {
          const T* src = buf + k * SK + m * SJ + i * SI;
          q1[k] = src[0], q2[k] = src[PL];
          if constexpr (!DIAG)
            q3[k] = src[2 * PL], q4[k] = src[3 * PL];
        }
        // column i of S, M, C and row i of C: all read as rows (S, M symmetric; CT = C transposed)
        T Sr[NDP], Mr[NDP], Ctr[NDPX], Cr[NDPX];
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
          for (int k = 0; k < KCH; ++k)
          {
            T& o = out[i1 * ND + r * KCH + k];
            if constexpr (DIAG)
              o = fma(Si, q1[k], fma(Mi, q2[k], o));
            else
              o = fma(Si, q1[k], fma(Mi, q2[k], fma(Ctr[i1], q3[k], fma(Cr[i1], q4[k], o))));
          }
        }
      }
    }
    } // rounds
    {
      // Consecutive cells of the list that share a whole face k = nd - 1 / k = 0 (found at construction: on the box
      // mesh every pair along z): the first cell hands its face values to the second, lane (c, j) -> lane (c + 1, j),
      // and skips those atomics.  The atomic unit is what bounds this kernel (plain stores instead, a timing
      // experiment: 3.94 -> 3.33 ms at Q3 fp64; merged: 3.46 ms).
      constexpr unsigned team_mask = CW * ND == 32 ? 0xffffffffu : ((1u << (CW * ND % 32)) - 1u);
      if (lane_ok)
      {
        const int from_prev = __shfl_up_sync(team_mask, mflag, ND);
#pragma unroll
        for (int i1 = 0; i1 < ND; ++i1)
        {
          const T face = __shfl_up_sync(team_mask, out[i1 * ND + ND - 1], ND);
          if (c > 0 && from_prev)
            out[i1 * ND] += face;
        }
      }
    }
    if (lane_ok)
    {
      // scatter indices: dof (i1, j = m, k) is entry (m, k) of the part of the row that lane (c, i1) fetched
#pragma unroll
      for (int i1 = 0; i1 < ND; ++i1)
      {
        std::int32_t sd[ND];
        if constexpr (IDX16)
        {
          const int4 d = *reinterpret_cast<const int4*>(ib_cur + (m * 33 + c * ND + i1) * 4);
          sd[0] = d.x, sd[1] = d.y, sd[2] = d.z, sd[3] = d.w;
        }
        else
        {
#pragma unroll
          for (int k = 0; k < ND; ++k)
            sd[k] = ib_cur[islot(m * ND + k, c * ND + i1)];
        }
        // A constrained dof receives nothing (reference :424-425): it is sent +0.0 instead, so that the scatter is
        // nd^2 unconditional REDs -- written as `if (d >= 0) atomicAdd(...)` each one is its own branch region
        // (BSSY / BRA / RED / BSYNC).  Constrained dofs are boundary dofs: few, and y[d] + 0.0 = y[d].
#pragma unroll
        for (int k = 0; k < ND; ++k)
          if (k < ND - 1 || !mflag)
            atomicAdd(yy + (sd[k] & 0x7fffffff), sd[k] < 0 ? T(0) : out[i1 * ND + k]);
      }
    }
    __syncwarp(); // the exchange buffer and this batch's row buffer are rewritten from the top of the next batch on

    std::int32_t* tmp = ib_cur;
    ib_cur = ib_n1;
    ib_n1 = ib_n2;
    ib_n2 = tmp;
    cell_cur = cell_n1;
    cell_n1 = cell_n2;
    cell_n2 = cell_n3;
    bcmask = bcmask_next;
    kappa = kappa_next;
    mflag = mflag_next;
    T* vtmp = vb_cur;
    vb_cur = vb_n1;
    vb_n1 = vtmp;
  }
  // copies requested for batches this warp does not have (plain stores only: cell < 0) -- nothing is in flight
}
} // namespace bd
