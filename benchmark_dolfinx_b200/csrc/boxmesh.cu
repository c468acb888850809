// Structured stand-in for the DOLFINx host setup, generated on the device.
//
// The reference builds its mesh, dofmap, ghost layer, boundary markers and RHS with DOLFINx
// (src/mesh.cpp:26-114,190-218, src/main.cpp:54-102, src/laplacian_solver.cpp:100-109); those
// take minutes at 300M dofs per GPU (examples/slurm.out:94-114) and DOLFINx is not available
// here.  This unit produces the same *arrays* for one rank of a block partition of the
// compute_mesh_size box, by closed-form index arithmetic in kernels:
//   cell_dofmap [ncells][(P+1)^3]  tensor-product local order, x slowest
//   geometry_dofmap [ncells][8]    vertex v = a*4 + b*2 + c
//   xgeom [nnodes][3]              unit cube (+ optional x perturbation)
//   bc_marker [ndofs]              all lattice points on the cube surface
//   lcells / bcells                src/mesh.cpp:154-188
//   halo lists                     owner->ghost, ghosts ordered by (owner rank, global id)
// Numbering: owned dofs first (lexicographic in the owned point box), ghosts after; a lattice
// plane shared by two blocks belongs to the upper block; ghost cells are exactly the cells
// of other ranks that touch an owned dof (the lower neighbours), so no reverse scatter is
// needed (same idea as the reference's ghost layer, src/mesh.hpp:42-44).
#include <algorithm>
#include <cmath>
#include <vector>

#include "common.cuh"
#include "vector_ops.cuh"

namespace bd
{
struct RankGeom
{
  int b[3];
  std::int64_t c0[3], c1[3], e0[3]; // owned cell range, extended (ghost layer) lower bound
  std::int64_t lo[3], hi[3];        // owned lattice points [lo, hi)
};

static RankGeom rank_geom(const std::int64_t n[3], int P, const int pgrid[3], int rank)
{
  RankGeom g;
  g.b[0] = rank / (pgrid[1] * pgrid[2]);
  g.b[1] = (rank / pgrid[2]) % pgrid[1];
  g.b[2] = rank % pgrid[2];
  for (int d = 0; d < 3; ++d)
  {
    g.c0[d] = (g.b[d] * n[d]) / pgrid[d];
    g.c1[d] = ((g.b[d] + 1) * n[d]) / pgrid[d];
    g.lo[d] = g.c0[d] * P;
    g.hi[d] = g.c1[d] * P + (g.b[d] == pgrid[d] - 1 ? 1 : 0);
    g.e0[d] = g.c0[d] > 0 ? g.c0[d] - 1 : 0;
  }
  return g;
}

struct BoxParams
{
  int P, nd;
  std::int64_t n[3];      // global cells
  std::int64_t NP[3];     // global lattice extents
  std::int64_t c0[3], oc[3]; // owned cell block origin / extents
  std::int64_t v0[3], nv[3]; // local vertex box
  std::int64_t NV[3];
  // zone tables for the local dof numbering: per dim, zone 0 = below, 1 = owned, 2 = above
  std::int64_t zlo[3][3], zsz[3][3];
  std::int64_t zone_offset[27]; // first local index of the zone's dofs
  std::int64_t ncells_owned, ncells;
  double perturb; // amplitude of the x perturbation (0 = none)
};

__device__ __forceinline__ void cell_coords(const BoxParams& p, const std::int32_t* ghost_cells,
                                            std::int64_t l, std::int64_t c[3])
{
  if (l < p.ncells_owned)
  {
    c[2] = p.c0[2] + l % p.oc[2];
    c[1] = p.c0[1] + (l / p.oc[2]) % p.oc[1];
    c[0] = p.c0[0] + l / (p.oc[1] * p.oc[2]);
  }
  else
  {
    const std::int32_t* g = ghost_cells + 3 * (l - p.ncells_owned);
    c[0] = g[0];
    c[1] = g[1];
    c[2] = g[2];
  }
}

__device__ __forceinline__ std::int32_t local_dof(const BoxParams& p, const std::int64_t g[3])
{
  int z[3];
#pragma unroll
  for (int d = 0; d < 3; ++d)
    z[d] = g[d] < p.zlo[d][1] ? 0 : (g[d] < p.zlo[d][1] + p.zsz[d][1] ? 1 : 2);
  const int code = z[0] * 9 + z[1] * 3 + z[2];
  const std::int64_t idx
      = ((g[0] - p.zlo[0][z[0]]) * p.zsz[1][z[1]] + (g[1] - p.zlo[1][z[1]])) * p.zsz[2][z[2]]
        + (g[2] - p.zlo[2][z[2]]);
  return static_cast<std::int32_t>(p.zone_offset[code] + idx);
}

__global__ void __launch_bounds__(256)
    dofmap_kernel(const __grid_constant__ BoxParams p, const std::int32_t* __restrict__ ghost_cells,
                  std::int32_t* __restrict__ dofmap, std::int8_t* __restrict__ bc)
{
  const int nd = p.nd, nd3 = nd * nd * nd;
  const std::int64_t total = p.ncells * nd3;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
       idx < total; idx += stride)
  {
    const std::int64_t l = idx / nd3;
    const int t = static_cast<int>(idx - l * nd3);
    std::int64_t c[3];
    cell_coords(p, ghost_cells, l, c);
    const int a[3] = {t / (nd * nd), (t / nd) % nd, t % nd};
    std::int64_t g[3];
    bool surface = false;
#pragma unroll
    for (int d = 0; d < 3; ++d)
    {
      g[d] = c[d] * p.P + a[d];
      surface = surface || g[d] == 0 || g[d] == p.NP[d] - 1;
    }
    const std::int32_t dof = local_dof(p, g);
    dofmap[idx] = dof;
    bc[dof] = surface ? 1 : 0; // every sharing cell writes the same value
  }
}

__global__ void __launch_bounds__(256)
    geometry_dofmap_kernel(const __grid_constant__ BoxParams p,
                           const std::int32_t* __restrict__ ghost_cells,
                           std::int32_t* __restrict__ gdofmap)
{
  const std::int64_t total = p.ncells * 8;
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
       idx < total; idx += stride)
  {
    const std::int64_t l = idx >> 3;
    const int v = static_cast<int>(idx & 7);
    std::int64_t c[3];
    cell_coords(p, ghost_cells, l, c);
    const std::int64_t vi = c[0] + ((v >> 2) & 1) - p.v0[0];
    const std::int64_t vj = c[1] + ((v >> 1) & 1) - p.v0[1];
    const std::int64_t vk = c[2] + (v & 1) - p.v0[2];
    gdofmap[idx] = static_cast<std::int32_t>((vi * p.nv[1] + vj) * p.nv[2] + vk);
  }
}

__device__ __forceinline__ double perturb_unit(std::uint64_t gid)
{
  // splitmix64 of the global vertex id: uniform in [-1, 1), identical on every rank
  std::uint64_t z = (gid + 42ull) * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return static_cast<double>(z >> 11) * (2.0 / 9007199254740992.0) - 1.0;
}

template <typename T>
__global__ void __launch_bounds__(256) xgeom_kernel(const __grid_constant__ BoxParams p, T* __restrict__ x)
{
  const std::int64_t total = p.nv[0] * p.nv[1] * p.nv[2];
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t idx = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
       idx < total; idx += stride)
  {
    const std::int64_t vk = p.v0[2] + idx % p.nv[2];
    const std::int64_t vj = p.v0[1] + (idx / p.nv[2]) % p.nv[1];
    const std::int64_t vi = p.v0[0] + idx / (p.nv[1] * p.nv[2]);
    double xv = static_cast<double>(vi) / static_cast<double>(p.n[0]);
    if (p.perturb != 0.0)
    {
      const std::uint64_t gid = static_cast<std::uint64_t>((vi * p.NV[1] + vj) * p.NV[2] + vk);
      xv += perturb_unit(gid) * p.perturb;
    }
    x[3 * idx + 0] = static_cast<T>(xv);
    x[3 * idx + 1] = static_cast<T>(static_cast<double>(vj) / static_cast<double>(p.n[1]));
    x[3 * idx + 2] = static_cast<T>(static_cast<double>(vk) / static_cast<double>(p.n[2]));
  }
}

// ---- RHS: b = sum_cells Phi^T (w detJ Phi f), f interpolated at the physical dof nodes.
// One CTA per cell, double precision throughout, runtime (nd, nq); setup code, not a hot path.
constexpr int kRhsThreads = 128;
struct RhsTables
{
  double phi[9 * 8];  // [nq][nd]
  double nodes[8];    // dof nodes
};

template <typename T>
__global__ void __launch_bounds__(kRhsThreads)
    rhs_kernel(int nd, int nq, std::int64_t ncells, const T* __restrict__ xgeom,
               const std::int32_t* __restrict__ gdofmap, const std::int32_t* __restrict__ dofmap,
               const double* __restrict__ weights, const double* __restrict__ dphi,
               const __grid_constant__ RhsTables tab, double* __restrict__ b)
{
  extern __shared__ double smem[];
  const int nd3 = nd * nd * nd, nq3 = nq * nq * nq;
  double* bufA = smem;        // nq^3
  double* bufB = bufA + nq3;  // nq^3
  double* xc = bufB + nq3;    // 24
  const int tid = threadIdx.x;
  for (std::int64_t cell = blockIdx.x; cell < ncells; cell += gridDim.x)
  {
    __syncthreads();
    if (tid < 24)
      xc[tid] = static_cast<double>(xgeom[3 * static_cast<std::int64_t>(gdofmap[cell * 8 + tid / 3]) + tid % 3]);
    __syncthreads();
    // f at the dof nodes (src/main.cpp:81-92) -> bufB[nd^3]
    for (int t = tid; t < nd3; t += kRhsThreads)
    {
      const int i = t / (nd * nd), j = (t / nd) % nd, k = t % nd;
      const double L[3][2] = {{1 - tab.nodes[i], tab.nodes[i]},
                              {1 - tab.nodes[j], tab.nodes[j]},
                              {1 - tab.nodes[k], tab.nodes[k]}};
      double x = 0, y = 0;
      for (int v = 0; v < 8; ++v)
      {
        const double N = L[0][(v >> 2) & 1] * L[1][(v >> 1) & 1] * L[2][v & 1];
        x += N * xc[3 * v];
        y += N * xc[3 * v + 1];
      }
      const double dx = (x - 0.5) * (x - 0.5), dy = (y - 0.5) * (y - 0.5);
      bufB[t] = 1000 * exp(-(dx + dy) / 0.02);
    }
    __syncthreads();
    // up: x, y, z
    for (int t = tid; t < nq * nd * nd; t += kRhsThreads)
    {
      const int q = t / (nd * nd), r = t % (nd * nd);
      double s = 0;
      for (int i = 0; i < nd; ++i)
        s += tab.phi[q * nd + i] * bufB[i * nd * nd + r];
      bufA[t] = s;
    }
    __syncthreads();
    for (int t = tid; t < nq * nq * nd; t += kRhsThreads)
    {
      const int q = t / (nq * nd), r = (t / nd) % nq, k = t % nd;
      double s = 0;
      for (int j = 0; j < nd; ++j)
        s += tab.phi[r * nd + j] * bufA[(q * nd + j) * nd + k];
      bufB[t] = s;
    }
    __syncthreads();
    for (int t = tid; t < nq3; t += kRhsThreads)
    {
      const int qr = t / nq, sidx = t % nq;
      double s = 0;
      for (int k = 0; k < nd; ++k)
        s += tab.phi[sidx * nd + k] * bufB[qr * nd + k];
      // scale by w * detJ at this point
      double J[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
      for (int v = 0; v < 8; ++v)
        for (int jj = 0; jj < 3; ++jj)
        {
          const double d = dphi[(static_cast<std::int64_t>(jj) * nq3 + t) * 8 + v];
          J[0][jj] += xc[3 * v] * d;
          J[1][jj] += xc[3 * v + 1] * d;
          J[2][jj] += xc[3 * v + 2] * d;
        }
      const double detJ = J[0][0] * (J[1][1] * J[2][2] - J[1][2] * J[2][1])
                          - J[0][1] * (J[1][0] * J[2][2] - J[1][2] * J[2][0])
                          + J[0][2] * (J[1][0] * J[2][1] - J[1][1] * J[2][0]);
      bufA[t] = s * weights[t] * detJ;
    }
    __syncthreads();
    // down: x, y, z
    for (int t = tid; t < nd * nq * nq; t += kRhsThreads)
    {
      const int i = t / (nq * nq), r = t % (nq * nq);
      double s = 0;
      for (int q = 0; q < nq; ++q)
        s += tab.phi[q * nd + i] * bufA[q * nq * nq + r];
      bufB[t] = s;
    }
    __syncthreads();
    for (int t = tid; t < nd * nd * nq; t += kRhsThreads)
    {
      const int i = t / (nd * nq), j = (t / nq) % nd, sidx = t % nq;
      double s = 0;
      for (int r = 0; r < nq; ++r)
        s += tab.phi[r * nd + j] * bufB[(i * nq + r) * nq + sidx];
      bufA[t] = s;
    }
    __syncthreads();
    for (int t = tid; t < nd3; t += kRhsThreads)
    {
      const int ij = t / nd, k = t % nd;
      double s = 0;
      for (int sidx = 0; sidx < nq; ++sidx)
        s += tab.phi[sidx * nd + k] * bufA[ij * nq + sidx];
      atomicAdd(b + dofmap[cell * nd3 + t], s);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
    rhs_finish_kernel(std::int64_t n_owned, std::int64_t n_total, const double* __restrict__ acc,
                      const std::int8_t* __restrict__ bc, T* __restrict__ b)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n_total;
       i += stride)
    b[i] = (i < n_owned && !bc[i]) ? static_cast<T>(acc[i]) : T(0); // bc.set (src/laplacian_solver.cpp:105)
}
} // namespace bd

struct bd_boxmesh
{
  int scalar_bits = 64, P = 1, rank = 0, nranks = 1;
  int pgrid[3] = {1, 1, 1};
  std::int64_t n[3] = {1, 1, 1};
  bd::BoxParams params{};
  std::int64_t ncells_owned = 0, ncells = 0, ndofs_owned = 0, nghost = 0, nnodes = 0;
  bd::DeviceArray<std::int32_t> cell_dofmap, geometry_dofmap, lcells, bcells, ghost_cells;
  bd::DeviceArray<std::int8_t> bc_marker;
  bd::DeviceArray<char> xgeom;
  std::vector<std::int32_t> send_ranks, recv_ranks, send_indices;
  std::vector<std::int64_t> send_counts, recv_counts;
  std::int64_t n_lcells = 0, n_bcells = 0;
};

using namespace bd;

extern "C" int bd_boxmesh_create(int scalar_bits, const int64_t n_in[3], int degree, int rank,
                                 int nranks, const int32_t pgrid_in[3], double geom_perturb_fact,
                                 bd_boxmesh** out)
{
  BD_REQUIRE(out && n_in && pgrid_in, "bd_boxmesh_create: null argument");
  *out = nullptr;
  BD_REQUIRE(scalar_bits == 32 || scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  if (degree < 1 || degree > 7)
    return fail(BD_ERR_UNSUPPORTED, "Unsupported degree [operator]");
  BD_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
  BD_REQUIRE(pgrid_in[0] * pgrid_in[1] * pgrid_in[2] == nranks, "partition grid does not match nranks");
  for (int d = 0; d < 3; ++d)
    BD_REQUIRE(n_in[d] >= pgrid_in[d] && n_in[d] >= 1, "fewer cells than ranks in a direction");

  auto m = new bd_boxmesh;
  auto bail = [&](int rc) {
    delete m;
    return rc;
  };
  m->scalar_bits = scalar_bits;
  m->P = degree;
  m->rank = rank;
  m->nranks = nranks;
  const int P = degree, nd = P + 1;
  const int pgrid[3] = {pgrid_in[0], pgrid_in[1], pgrid_in[2]};
  for (int d = 0; d < 3; ++d)
  {
    m->n[d] = n_in[d];
    m->pgrid[d] = pgrid[d];
  }
  const RankGeom g = rank_geom(m->n, P, pgrid, rank);
  BoxParams& p = m->params;
  p.P = P;
  p.nd = nd;
  p.perturb = geom_perturb_fact != 0.0 ? geom_perturb_fact * 1.0 / static_cast<double>(m->n[0]) : 0.0;
  std::int64_t ext[3];
  for (int d = 0; d < 3; ++d)
  {
    p.n[d] = m->n[d];
    p.NP[d] = m->n[d] * P + 1;
    p.NV[d] = m->n[d] + 1;
    p.c0[d] = g.c0[d];
    p.oc[d] = g.c1[d] - g.c0[d];
    p.v0[d] = g.e0[d];
    p.nv[d] = g.c1[d] - g.e0[d] + 1;
    ext[d] = g.c1[d] - g.e0[d];
    // zones: below [e0*P, lo), owned [lo, hi), above [hi, c1*P + 1)
    p.zlo[d][0] = g.e0[d] * P;
    p.zsz[d][0] = g.lo[d] - g.e0[d] * P;
    p.zlo[d][1] = g.lo[d];
    p.zsz[d][1] = g.hi[d] - g.lo[d];
    p.zlo[d][2] = g.hi[d];
    p.zsz[d][2] = g.c1[d] * P + 1 - g.hi[d];
  }
  m->ncells_owned = p.oc[0] * p.oc[1] * p.oc[2];
  m->ncells = ext[0] * ext[1] * ext[2];
  p.ncells_owned = m->ncells_owned;
  p.ncells = m->ncells;
  m->ndofs_owned = p.zsz[0][1] * p.zsz[1][1] * p.zsz[2][1];
  m->nnodes = p.nv[0] * p.nv[1] * p.nv[2];
  // ghost zones ordered by zone code == ascending owner rank; offsets + receive plan
  std::int64_t off = m->ndofs_owned;
  for (int code = 0; code < 27; ++code)
  {
    const int z[3] = {code / 9, (code / 3) % 3, code % 3};
    if (code == 13)
    {
      p.zone_offset[code] = 0;
      continue;
    }
    p.zone_offset[code] = off;
    const std::int64_t cnt = p.zsz[0][z[0]] * p.zsz[1][z[1]] * p.zsz[2][z[2]];
    if (cnt > 0)
    {
      const int nb[3] = {g.b[0] + z[0] - 1, g.b[1] + z[1] - 1, g.b[2] + z[2] - 1};
      m->recv_ranks.push_back((nb[0] * pgrid[1] + nb[1]) * pgrid[2] + nb[2]);
      m->recv_counts.push_back(cnt);
    }
    off += cnt;
  }
  m->nghost = off - m->ndofs_owned;
  if (off > 0x7fffffffLL || m->ncells * nd * nd * nd > 0x7fffffffLL * 4ll)
    return bail(fail(BD_ERR_INVALID_ARGUMENT, "local mesh too large for 32-bit dof indices"));

  // ---- send lists: for every neighbour t, the points of its ghost zone -t that I own,
  //      in lexicographic (= ascending global id) order, as my local owned indices
  for (int code = 0; code < 27; ++code)
  {
    if (code == 13)
      continue;
    const int t[3] = {code / 9 - 1, (code / 3) % 3 - 1, code % 3 - 1};
    const int rb[3] = {g.b[0] + t[0], g.b[1] + t[1], g.b[2] + t[2]};
    bool valid = true;
    for (int d = 0; d < 3; ++d)
      valid = valid && rb[d] >= 0 && rb[d] < pgrid[d];
    if (!valid)
      continue;
    const int rrank = (rb[0] * pgrid[1] + rb[1]) * pgrid[2] + rb[2];
    const RankGeom rg = rank_geom(m->n, P, pgrid, rrank);
    std::int64_t blo[3], bhi[3]; // global sub-box the receiver ghosts from me
    bool empty = false;
    for (int d = 0; d < 3; ++d)
    {
      if (t[d] == 1) // I am below the receiver: its zone 0
      {
        blo[d] = rg.e0[d] * P;
        bhi[d] = rg.lo[d];
      }
      else if (t[d] == 0)
      {
        blo[d] = rg.lo[d];
        bhi[d] = rg.hi[d];
      }
      else // I am above the receiver: its zone 2
      {
        blo[d] = rg.hi[d];
        bhi[d] = rg.c1[d] * P + 1;
      }
      empty = empty || bhi[d] <= blo[d];
    }
    if (empty)
      continue;
    std::int64_t cnt = 0;
    for (std::int64_t x = blo[0]; x < bhi[0]; ++x)
      for (std::int64_t y = blo[1]; y < bhi[1]; ++y)
        for (std::int64_t z = blo[2]; z < bhi[2]; ++z)
        {
          const std::int64_t li = ((x - g.lo[0]) * p.zsz[1][1] + (y - g.lo[1])) * p.zsz[2][1] + (z - g.lo[2]);
          m->send_indices.push_back(static_cast<std::int32_t>(li));
          ++cnt;
        }
    m->send_ranks.push_back(rrank);
    m->send_counts.push_back(cnt);
  }
  // send lists must be in ascending receiver rank to pair with the receivers' segments
  // (zone codes of t are lexicographic in the receiver's block coordinates, hence ascending).

  // ---- cell lists (host; surface-sized except lcells)
  std::vector<std::int32_t> ghost_cells, lcells, bcells;
  {
    const bool up[3] = {g.b[0] < pgrid[0] - 1, g.b[1] < pgrid[1] - 1, g.b[2] < pgrid[2] - 1};
    lcells.reserve(static_cast<size_t>(m->ncells_owned));
    std::int64_t l = 0;
    for (std::int64_t i = 0; i < p.oc[0]; ++i)
      for (std::int64_t j = 0; j < p.oc[1]; ++j)
        for (std::int64_t k = 0; k < p.oc[2]; ++k, ++l)
        {
          const bool touches_ghost = (up[0] && i == p.oc[0] - 1) || (up[1] && j == p.oc[1] - 1)
                                     || (up[2] && k == p.oc[2] - 1);
          (touches_ghost ? bcells : lcells).push_back(static_cast<std::int32_t>(l));
        }
    for (std::int64_t i = g.e0[0]; i < g.c1[0]; ++i)
      for (std::int64_t j = g.e0[1]; j < g.c1[1]; ++j)
        for (std::int64_t k = g.e0[2]; k < g.c1[2]; ++k)
          if (i < g.c0[0] || j < g.c0[1] || k < g.c0[2])
          {
            ghost_cells.push_back(static_cast<std::int32_t>(i));
            ghost_cells.push_back(static_cast<std::int32_t>(j));
            ghost_cells.push_back(static_cast<std::int32_t>(k));
            bcells.push_back(static_cast<std::int32_t>(l++));
          }
  }
  m->n_lcells = static_cast<std::int64_t>(lcells.size());
  m->n_bcells = static_cast<std::int64_t>(bcells.size());
  int rc;
#define TRY(expr)                                                                                  \
  if ((rc = (expr)) != BD_SUCCESS)                                                                 \
    return bail(rc);
  TRY(m->ghost_cells.assign(ghost_cells.data(), ghost_cells.size(), false));
  TRY(m->lcells.assign(lcells.data(), lcells.size(), false));
  TRY(m->bcells.assign(bcells.data(), bcells.size(), false));
  TRY(m->cell_dofmap.alloc(static_cast<size_t>(m->ncells) * nd * nd * nd));
  TRY(m->geometry_dofmap.alloc(static_cast<size_t>(m->ncells) * 8));
  TRY(m->bc_marker.alloc(static_cast<size_t>(m->ndofs_owned + m->nghost)));
  TRY(m->xgeom.alloc(static_cast<size_t>(m->nnodes) * 3 * (scalar_bits / 8)));
#undef TRY
  const int grid = sm_count() * 8;
  dofmap_kernel<<<grid, 256>>>(p, m->ghost_cells.ptr, m->cell_dofmap.ptr, m->bc_marker.ptr);
  geometry_dofmap_kernel<<<grid, 256>>>(p, m->ghost_cells.ptr, m->geometry_dofmap.ptr);
  if (scalar_bits == 64)
    xgeom_kernel<double><<<grid, 256>>>(p, reinterpret_cast<double*>(m->xgeom.ptr));
  else
    xgeom_kernel<float><<<grid, 256>>>(p, reinterpret_cast<float*>(m->xgeom.ptr));
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess)
    return bail(fail(BD_ERR_CUDA, "box mesh generation failed: %s", cudaGetErrorString(e)));
  *out = m;
  return BD_SUCCESS;
}

extern "C" int bd_boxmesh_get(bd_boxmesh* m, bd_boxmesh_info* info)
{
  BD_REQUIRE(m && info, "bd_boxmesh_get: null argument");
  info->ncells_owned = m->ncells_owned;
  info->ncells = m->ncells;
  info->ndofs_owned = m->ndofs_owned;
  info->nghost = m->nghost;
  info->nnodes = m->nnodes;
  info->n_lcells = m->n_lcells;
  info->n_bcells = m->n_bcells;
  info->ndofs_global = m->params.NP[0] * m->params.NP[1] * m->params.NP[2];
  info->ncells_global = m->n[0] * m->n[1] * m->n[2];
  info->cell_dofmap = m->cell_dofmap.ptr;
  info->geometry_dofmap = m->geometry_dofmap.ptr;
  info->xgeom = m->xgeom.ptr;
  info->bc_marker = m->bc_marker.ptr;
  info->lcells = m->lcells.ptr;
  info->bcells = m->bcells.ptr;
  info->n_send = static_cast<int32_t>(m->send_ranks.size());
  info->n_recv = static_cast<int32_t>(m->recv_ranks.size());
  info->send_ranks = m->send_ranks.data();
  info->send_counts = m->send_counts.data();
  info->send_indices_host = m->send_indices.data();
  info->recv_ranks = m->recv_ranks.data();
  info->recv_counts = m->recv_counts.data();
  return BD_SUCCESS;
}

extern "C" int bd_boxmesh_destroy(bd_boxmesh* m)
{
  delete m;
  return BD_SUCCESS;
}

extern "C" int bd_boxmesh_assemble_rhs(bd_boxmesh* m, int qmode, int use_gauss, void* b_dev,
                                       void* stream_)
{
  BD_REQUIRE(m && b_dev, "bd_boxmesh_assemble_rhs: null argument");
  BD_REQUIRE(qmode == 0 || qmode == 1, "Invalid qmode.");
  cudaStream_t stream = as_stream(stream_);
  const int nd = m->P + 1, nq = m->P + 1 + qmode, nq3 = nq * nq * nq;
  std::vector<double> phi(nq * nd), w(nq3), dphi(3 * static_cast<size_t>(nq3) * 8), nodes(nd);
  BD_CHECK(bd_tables(64, m->P, qmode, use_gauss, phi.data(), nullptr, w.data(), dphi.data(),
                     nodes.data(), nullptr, nullptr));
  RhsTables tab{};
  std::copy(phi.begin(), phi.end(), tab.phi);
  std::copy(nodes.begin(), nodes.end(), tab.nodes);
  DeviceArray<double> w_d, dphi_d, acc;
  BD_CHECK(w_d.assign(w.data(), w.size(), false));
  BD_CHECK(dphi_d.assign(dphi.data(), dphi.size(), false));
  const std::int64_t ntot = m->ndofs_owned + m->nghost;
  BD_CHECK(acc.alloc(static_cast<size_t>(ntot)));
  BD_CUDA_CHECK(cudaMemsetAsync(acc.ptr, 0, static_cast<size_t>(ntot) * sizeof(double), stream));
  const size_t smem = (2 * static_cast<size_t>(nq3) + 24) * sizeof(double);
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 16;
  const int grid = static_cast<int>(m->ncells < cap ? m->ncells : cap);
  if (m->scalar_bits == 64)
    rhs_kernel<double><<<grid, kRhsThreads, smem, stream>>>(
        nd, nq, m->ncells, reinterpret_cast<const double*>(m->xgeom.ptr), m->geometry_dofmap.ptr,
        m->cell_dofmap.ptr, w_d.ptr, dphi_d.ptr, tab, acc.ptr);
  else
    rhs_kernel<float><<<grid, kRhsThreads, smem, stream>>>(
        nd, nq, m->ncells, reinterpret_cast<const float*>(m->xgeom.ptr), m->geometry_dofmap.ptr,
        m->cell_dofmap.ptr, w_d.ptr, dphi_d.ptr, tab, acc.ptr);
  BD_CUDA_CHECK(cudaGetLastError());
  const int g2 = vec_grid(ntot);
  if (m->scalar_bits == 64)
    rhs_finish_kernel<double><<<g2, 256, 0, stream>>>(m->ndofs_owned, ntot, acc.ptr, m->bc_marker.ptr,
                                                     static_cast<double*>(b_dev));
  else
    rhs_finish_kernel<float><<<g2, 256, 0, stream>>>(m->ndofs_owned, ntot, acc.ptr, m->bc_marker.ptr,
                                                    static_cast<float*>(b_dev));
  BD_CUDA_CHECK(cudaGetLastError());
  BD_CUDA_CHECK(cudaStreamSynchronize(stream)); // temporaries are freed on return
  return BD_SUCCESS;
}
