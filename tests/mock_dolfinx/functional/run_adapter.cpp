// Runs benchmark_dolfinx_b200/host/dolfinx_adapter.hpp for real, on functional stand-ins of DOLFINx / Basix / MPI
// (this directory): ranks are threads, meshes are the library's structured stand-in (bd_boxmesh_*).
//
//   run_adapter lists <nranks> <ndofs_global> <degree> [scramble]
//       every rank builds its IndexMap and calls compute_halo_lists (the MPI_Alltoall(v) part of the halo plan) --
//       no NCCL, so any number of ranks runs on one GPU.  Checks, for every rank and neighbour, that entry k of the
//       block owner o sends to rank r carries the global index r expects at the ghost position it lands on; with
//       `scramble` the ghosts of the IndexMap are presented in a shuffled order (unpack path).
//   run_adapter solve <nranks> <ndofs_global> <degree> <qmode> <cg_iterations>
//       constructor, apply and cg_solve through the adapter, one GPU per rank (needs nranks GPUs for nranks > 1);
//       prints JSON with the global norms of u (the assembled right-hand side), y = A u and the CG iterate.
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>
#include <string>
#include <thread>

#include <dolfinx_adapter.hpp>

namespace
{
struct RankMesh
{
  bd_boxmesh* h = nullptr;
  bd_boxmesh_info info{};
  std::vector<std::int32_t> cell_dofmap, geometry_dofmap, lcells, bcells;
  std::vector<std::int8_t> bc;
  std::vector<double> x;
};

void ck(int rc)
{
  if (rc != BD_SUCCESS)
  {
    std::fprintf(stderr, "bd error: %s\n", bd_last_error());
    std::exit(2);
  }
}
template <typename V>
void d2h(std::vector<V>& dst, const void* src, size_t n)
{
  dst.resize(n);
  if (n && cudaMemcpy(dst.data(), src, n * sizeof(V), cudaMemcpyDeviceToHost) != cudaSuccess)
  {
    std::fprintf(stderr, "cudaMemcpy failed\n");
    std::exit(2);
  }
}

void partition(int nranks, std::int32_t p[3])
{
  const int t[9][3] = {{1, 1, 1}, {1, 1, 1}, {2, 1, 1}, {0, 0, 0}, {2, 2, 1}, {0, 0, 0}, {0, 0, 0}, {0, 0, 0}, {2, 2, 2}};
  if (nranks > 8 || t[nranks][0] == 0)
  {
    std::fprintf(stderr, "nranks must be 1, 2, 4 or 8\n");
    std::exit(2);
  }
  std::copy(t[nranks], t[nranks] + 3, p);
}

RankMesh build_mesh(int rank, int nranks, std::int64_t ndofs, int degree, int device)
{
  cudaSetDevice(device);
  RankMesh m;
  std::int64_t n[3];
  ck(bd_compute_mesh_size(ndofs, degree, n));
  std::int32_t p[3];
  partition(nranks, p);
  ck(bd_boxmesh_create(64, n, degree, rank, nranks, p, 0.0, &m.h));
  ck(bd_boxmesh_get(m.h, &m.info));
  const bd_boxmesh_info& i = m.info;
  const int nd3 = (degree + 1) * (degree + 1) * (degree + 1);
  d2h(m.cell_dofmap, i.cell_dofmap, static_cast<size_t>(i.ncells) * nd3);
  d2h(m.geometry_dofmap, i.geometry_dofmap, static_cast<size_t>(i.ncells) * 8);
  d2h(m.lcells, i.lcells, static_cast<size_t>(i.n_lcells));
  d2h(m.bcells, i.bcells, static_cast<size_t>(i.n_bcells));
  d2h(m.bc, i.bc_marker, static_cast<size_t>(i.ndofs_owned + i.nghost));
  d2h(m.x, i.xgeom, static_cast<size_t>(i.nnodes) * 3);
  return m;
}

// Global numbering: owned dofs of rank r are offset_r .. offset_r + n_owned_r; the ghosts of rank r from owner o are,
// in the stand-in's order, the entries o lists for destination r (bd_boxmesh_info.send_indices_host).
void fill_index_maps(const std::vector<RankMesh>& meshes, std::vector<dolfinx::common::IndexMap>& maps, MPI_Comm comm)
{
  const int nr = static_cast<int>(meshes.size());
  std::vector<std::int64_t> offset(nr + 1, 0);
  for (int r = 0; r < nr; ++r)
    offset[r + 1] = offset[r] + meshes[r].info.ndofs_owned;
  maps.resize(nr);
  for (int r = 0; r < nr; ++r)
  {
    auto& M = maps[r];
    M.n_local = static_cast<std::int32_t>(meshes[r].info.ndofs_owned);
    M.offset = offset[r];
    M.communicator = comm;
    const bd_boxmesh_info& ir = meshes[r].info;
    for (int a = 0; a < ir.n_recv; ++a)
    {
      const int o = ir.recv_ranks[a];
      const bd_boxmesh_info& io = meshes[o].info;
      std::int64_t off = 0;
      int b = 0;
      for (; b < io.n_send && io.send_ranks[b] != r; ++b)
        off += io.send_counts[b];
      if (b == io.n_send || io.send_counts[b] != ir.recv_counts[a])
      {
        std::fprintf(stderr, "stand-in meshes disagree on the %d -> %d message\n", o, r);
        std::exit(2);
      }
      for (std::int64_t k = 0; k < io.send_counts[b]; ++k)
      {
        M.ghost_ids.push_back(offset[o] + io.send_indices_host[off + k]);
        M.ghost_owners.push_back(o);
      }
    }
    if (static_cast<std::int64_t>(M.ghost_ids.size()) != ir.nghost)
    {
      std::fprintf(stderr, "ghost count mismatch on rank %d\n", r);
      std::exit(2);
    }
  }
}

int mode_lists(int nranks, std::int64_t ndofs, int degree, bool scramble)
{
  std::vector<RankMesh> meshes;
  for (int r = 0; r < nranks; ++r)
    meshes.push_back(build_mesh(r, nranks, ndofs, degree, 0));
  bd_mock_mpi_ctx ctx(nranks);
  std::vector<dolfinx::common::IndexMap> maps;
  fill_index_maps(meshes, maps, &ctx);
  if (scramble)
    for (int r = 0; r < nranks; ++r)
    {
      std::vector<int> perm(maps[r].ghost_ids.size());
      std::iota(perm.begin(), perm.end(), 0);
      std::mt19937 gen(17 + r);
      std::shuffle(perm.begin(), perm.end(), gen);
      auto ids = maps[r].ghost_ids;
      auto ow = maps[r].ghost_owners;
      for (size_t i = 0; i < perm.size(); ++i)
      {
        maps[r].ghost_ids[i] = ids[perm[i]];
        maps[r].ghost_owners[i] = ow[perm[i]];
      }
    }
  std::vector<benchdolfinx::b200::HaloLists> lists(nranks);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; ++r)
    th.emplace_back([&, r] {
      bd_mock_mpi_rank = r;
      lists[r] = benchdolfinx::b200::compute_halo_lists(maps[r]);
    });
  for (auto& t : th)
    t.join();
  long long checked = 0, bad = 0;
  for (int r = 0; r < nranks; ++r)
  {
    const auto& Lr = lists[r];
    // without scrambling the plan must be the stand-in's own (same neighbours, counts and indices, receives in place)
    if (!scramble)
    {
      const bd_boxmesh_info& i = meshes[r].info;
      std::int64_t tot = 0;
      for (int a = 0; a < i.n_send; ++a)
        tot += i.send_counts[a];
      const bool same = Lr.ordered && static_cast<int>(Lr.send_ranks.size()) == i.n_send
                        && static_cast<int>(Lr.recv_ranks.size()) == i.n_recv
                        && std::equal(Lr.send_ranks.begin(), Lr.send_ranks.end(), i.send_ranks)
                        && std::equal(Lr.send_counts.begin(), Lr.send_counts.end(), i.send_counts)
                        && std::equal(Lr.recv_ranks.begin(), Lr.recv_ranks.end(), i.recv_ranks)
                        && std::equal(Lr.recv_counts.begin(), Lr.recv_counts.end(), i.recv_counts)
                        && static_cast<std::int64_t>(Lr.send_idx.size()) == tot
                        && std::equal(Lr.send_idx.begin(), Lr.send_idx.end(), i.send_indices_host);
      if (!same)
        ++bad;
    }
    // semantics, any ghost order: what owner o sends as entry k of its block for r is the global index that r expects
    // at the ghost position the entry lands on
    std::int64_t recv_off = 0;
    for (size_t a = 0; a < Lr.recv_ranks.size(); ++a)
    {
      const int o = Lr.recv_ranks[a];
      const auto& Lo = lists[o];
      std::int64_t off = 0;
      size_t b = 0;
      for (; b < Lo.send_ranks.size() && Lo.send_ranks[b] != r; ++b)
        off += Lo.send_counts[b];
      if (b == Lo.send_ranks.size() || Lo.send_counts[b] != Lr.recv_counts[a])
      {
        ++bad;
        continue;
      }
      for (std::int64_t k = 0; k < Lr.recv_counts[a]; ++k, ++checked)
      {
        const std::int64_t sent = maps[o].offset + Lo.send_idx[off + k];
        const std::int32_t lands = Lr.order[recv_off + k];
        if (Lo.send_idx[off + k] < 0 || Lo.send_idx[off + k] >= maps[o].n_local || maps[r].ghost_ids[lands] != sent)
          ++bad;
      }
      recv_off += Lr.recv_counts[a];
    }
    if (recv_off != static_cast<std::int64_t>(maps[r].ghost_ids.size()))
      ++bad;
  }
  std::printf("{\"mode\": \"lists\", \"nranks\": %d, \"scramble\": %d, \"entries_checked\": %lld, \"bad\": %lld}\n", nranks,
              int(scramble), checked, bad);
  return bad == 0 ? 0 : 1;
}

// The reference's device vector alias, reduced to what the adapter touches: array().data().get().
struct DeviceVector
{
  using value_type = double;
  struct Arr
  {
    double* p = nullptr;
    struct Ptr
    {
      double* q;
      double* get() const { return q; }
    };
    Ptr data() const { return {p}; }
  } a;
  Arr& array() { return a; }
  const Arr& array() const { return a; }
};

int mode_solve(int nranks, std::int64_t ndofs, int degree, int qmode, int cg_its)
{
  int ndev = 0;
  cudaGetDeviceCount(&ndev);
  if (ndev < nranks)
  {
    std::printf("{\"mode\": \"solve\", \"skipped\": \"needs %d GPUs, %d present\"}\n", nranks, ndev);
    return 77;
  }
  std::vector<RankMesh> meshes;
  for (int r = 0; r < nranks; ++r)
    meshes.push_back(build_mesh(r, nranks, ndofs, degree, r));
  bd_mock_mpi_ctx ctx(nranks);
  std::vector<dolfinx::common::IndexMap> maps;
  fill_index_maps(meshes, maps, &ctx);
  std::vector<double> unorm(nranks), ynorm(nranks), xnorm(nranks);
  std::vector<int> its(nranks, 0), failed(nranks, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; ++r)
    th.emplace_back([&, r] {
      bd_mock_mpi_rank = r;
      cudaSetDevice(r);
      try
      {
        const RankMesh& m = meshes[r];
        dolfinx::fem::FunctionSpace<double> V;
        *V.dm->index_map = maps[r];
        V.dm->cell_dofs = m.cell_dofmap;
        V.msh->geom.coords = m.x;
        V.msh->geom.cell_nodes = m.geometry_dofmap;
        V.msh->topo->cell_map->n_local = static_cast<std::int32_t>(m.info.ncells_owned);
        V.msh->topo->cell_map->ghost_ids.resize(static_cast<size_t>(m.info.ncells - m.info.ncells_owned));
        V.msh->topo->cell_map->ghost_owners.resize(V.msh->topo->cell_map->ghost_ids.size());
        V.msh->topo->cell_map->communicator = &ctx;
        V.msh->topo->lcells = m.lcells;
        V.msh->topo->bcells = m.bcells;
        dolfinx::fem::DirichletBC<double> bc;
        bc.marker = m.bc;
        benchdolfinx::b200::MatFreeLaplacianGPU<double> op(V, bc, degree, qmode, 2.0, basix::quadrature::type::gll);
        const size_t n = static_cast<size_t>(m.info.ndofs_owned + m.info.nghost);
        DeviceVector u, y, x;
        cudaMalloc(&u.a.p, n * 8);
        cudaMalloc(&y.a.p, n * 8);
        cudaMalloc(&x.a.p, n * 8);
        cudaMemset(y.a.p, 0, n * 8);
        cudaMemset(x.a.p, 0, n * 8);
        ck(bd_boxmesh_assemble_rhs(m.h, qmode, 0, u.a.p, nullptr)); // u := b (src/laplacian_solver.cpp:100-109)
        op.apply(u, y);
        ck(bd_norm(op.comm(), 64, m.info.ndofs_owned, u.a.p, &unorm[r]));
        ck(bd_norm(op.comm(), 64, m.info.ndofs_owned, y.a.p, &ynorm[r]));
        if (cg_its > 0)
        {
          its[r] = benchdolfinx::b200::cg_solve(op, x, u, cg_its, 0.0);
          ck(bd_norm(op.comm(), 64, m.info.ndofs_owned, x.a.p, &xnorm[r]));
        }
        cudaFree(u.a.p);
        cudaFree(y.a.p);
        cudaFree(x.a.p);
      }
      catch (const std::exception& e)
      {
        std::fprintf(stderr, "rank %d: %s\n", r, e.what());
        failed[r] = 1;
        std::exit(3); // a failed rank must not leave the others waiting in a collective
      }
    });
  for (auto& t : th)
    t.join();
  std::printf("{\"mode\": \"solve\", \"nranks\": %d, \"u_norm\": %.17g, \"y_norm\": %.17g, \"x_norm\": %.17g, \"iterations\": %d}\n",
              nranks, unorm[0], ynorm[0], xnorm[0], its[0]);
  return 0;
}
} // namespace

int main(int argc, char** argv)
{
  if (argc >= 5 && !std::strcmp(argv[1], "lists"))
    return mode_lists(std::atoi(argv[2]), std::atoll(argv[3]), std::atoi(argv[4]), argc > 5 && !std::strcmp(argv[5], "scramble"));
  if (argc >= 7 && !std::strcmp(argv[1], "solve"))
    return mode_solve(std::atoi(argv[2]), std::atoll(argv[3]), std::atoi(argv[4]), std::atoi(argv[5]), std::atoi(argv[6]));
  std::fprintf(stderr, "usage: run_adapter lists <nranks> <ndofs> <degree> [scramble] | solve <nranks> <ndofs> <degree> <qmode> <cg_its>\n");
  return 2;
}
