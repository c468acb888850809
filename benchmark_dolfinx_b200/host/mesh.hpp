// Host-side mirror of the reference's src/mesh.hpp: mesh sizing and the (stand-in) box mesh.
#pragma once

#include <array>
#include <cstdint>

#include "vector.hpp"

namespace benchdolfinx
{
/// src/mesh.cpp:117-152
inline std::array<std::int64_t, 3> compute_mesh_size(std::int64_t ndofs_global, int degree)
{
  std::array<std::int64_t, 3> nx;
  check(bd_compute_mesh_size(ndofs_global, degree, nx.data()));
  return nx;
}

inline std::array<int, 3> partition_grid(int nranks)
{
  switch (nranks)
  {
  case 1: return {1, 1, 1};
  case 2: return {2, 1, 1};
  case 4: return {2, 2, 1};
  case 8: return {2, 2, 2};
  default: return {nranks, 1, 1};
  }
}

/// What create_mesh + create_functionspace + DirichletBC give the operator in the reference
/// (src/mesh.cpp:190-218, src/main.cpp:54-102), for one rank of the block partition.
template <typename T>
class BoxMesh
{
public:
  BoxMesh(std::array<std::int64_t, 3> n, int degree, int rank, int nranks, double geom_perturb_fact,
          bd_comm* comm)
      : degree(degree)
  {
    const auto pg = partition_grid(nranks);
    const std::int32_t pgrid[3] = {pg[0], pg[1], pg[2]};
    check(bd_boxmesh_create(8 * sizeof(T), n.data(), degree, rank, nranks, pgrid, geom_perturb_fact, &handle));
    check(bd_boxmesh_get(handle, &info));
    map.n_owned = info.ndofs_owned;
    map.n_ghost = info.nghost;
    map.size_global_ = info.ndofs_global;
    map.comm_ = comm;
    if (comm && nranks > 1)
      check(bd_halo_create(comm, 8 * sizeof(T), info.ndofs_owned, info.nghost, info.n_send, info.send_ranks,
                           info.send_counts, info.send_indices_host, info.n_recv, info.recv_ranks,
                           info.recv_counts, &map.halo));
  }
  ~BoxMesh()
  {
    bd_halo_destroy(map.halo);
    bd_boxmesh_destroy(handle);
  }
  BoxMesh(const BoxMesh&) = delete;
  int degree;
  bd_boxmesh* handle = nullptr;
  bd_boxmesh_info info{};
  IndexMap map;
};
} // namespace benchdolfinx
