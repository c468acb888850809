// DOLFINx v0.10 adapter: the reference's GPU operator class and CG entry point on top of the
// benchdolfinx_b200 C ABI.  Meant to be dropped next to the reference's sources and included in
// place of its laplacian.hpp GPU class (src/laplacian.hpp:89-447) and cg.hpp (src/cg.hpp:89-169).
//
// STATUS: DOLFINx, Basix and MPI are not installable in the build environment of this repository.  The header is
// (a) type-checked against declaration-only mock-ups (tests/mock_dolfinx/, tests/test_adapter_compiles.py) and
// (b) EXECUTED on the GPU against functional stand-ins (tests/mock_dolfinx/functional/: DOLFINx classes backed by the
// structured stand-in mesh, Basix calls backed by bd_tables, MPI as threads-as-ranks): constructor, halo lists,
// apply and cg_solve reproduce the reference's known answers (tests/test_gpu_adapter.py).  It has not met the real
// libraries.
//
// What it does, in the reference's terms:
//   * constructor  = src/laplacian.hpp:102-227: lcells/bcells (compute_boundary_cells), BC markers,
//     quadrature + 1-D tables from Basix, device copies and the G precompute (inside
//     bd_laplacian_create);
//   * halo plan    = what dolfinx::common::Scatterer + pack/unpack functors do for
//     Vector::scatter_fwd_begin/end (src/vector.hpp:95-149), rebuilt from the IndexMap: one
//     MPI_Alltoall(v) at construction tells every owner which of its dofs to send where; ghosts
//     need not be ordered by owner (bd_halo_set_unpack);
//   * apply        = src/laplacian.hpp:355-403;
//   * cg_solve     = src/cg.hpp:89-169 for this operator type.
#pragma once

#include <algorithm>
#include <cmath>
#include <limits>
#include <array>
#include <cstdint>
#include <functional>
#include <numeric>
#include <span>
#include <stdexcept>
#include <vector>

#include <basix/finite-element.h>
#include <basix/interpolation.h>
#include <basix/quadrature.h>
#include <dolfinx/common/IndexMap.h>
#include <dolfinx/fem/DirichletBC.h>
#include <dolfinx/fem/FunctionSpace.h>
#include <cuda_runtime.h>
#include <mpi.h>

#include "benchdolfinx_b200.h"
#include "mesh.hpp" // the reference's: benchdolfinx::compute_boundary_cells (src/mesh.hpp)

namespace benchdolfinx::b200
{
inline void check(int rc)
{
  if (rc != BD_SUCCESS)
    throw std::runtime_error(bd_last_error());
}

/// What every owner sends to whom, and where received entries land: the host-side part of the forward-scatter plan
/// (what dolfinx::common::Scatterer derives from an IndexMap), separated from the NCCL objects so that it can be
/// checked on its own.
struct HaloLists
{
  std::vector<std::int32_t> send_ranks, recv_ranks; ///< neighbour ranks, ascending
  std::vector<std::int64_t> send_counts, recv_counts;
  std::vector<std::int32_t> send_idx; ///< local (owned) indices to send, concatenated by destination rank
  std::vector<std::int32_t> order;    ///< message entry i (receives concatenated by source rank) lands at ghost order[i]
  bool ordered = true;                ///< order is the identity: receives can land in the ghost segment directly
};

/// One MPI_Alltoall + one MPI_Alltoallv: every rank tells the owners of its ghosts which global indices it wants.
inline HaloLists compute_halo_lists(const dolfinx::common::IndexMap& map)
{
  MPI_Comm comm = map.comm();
  int size = 1;
  MPI_Comm_size(comm, &size);
  HaloLists L;
  const std::int64_t offset = map.local_range()[0];
  std::span<const std::int64_t> ghosts = map.ghosts();
  std::span<const int> owners = map.owners();
  const std::size_t ng = ghosts.size();
  // message order: ghosts grouped by owner rank (stable), position i of the messages is ghost order[i]
  L.order.resize(ng);
  std::iota(L.order.begin(), L.order.end(), 0);
  std::stable_sort(L.order.begin(), L.order.end(), [&](int a, int b) { return owners[a] < owners[b]; });
  L.ordered = std::is_sorted(L.order.begin(), L.order.end());
  std::vector<int> want_count(size, 0);
  for (std::size_t i = 0; i < ng; ++i)
    ++want_count[owners[i]];
  std::vector<int> want_displ(size + 1, 0);
  std::partial_sum(want_count.begin(), want_count.end(), want_displ.begin() + 1);
  std::vector<std::int64_t> want(ng);
  for (std::size_t i = 0; i < ng; ++i)
    want[i] = ghosts[L.order[i]];
  // owners learn what to send
  std::vector<int> give_count(size, 0);
  MPI_Alltoall(want_count.data(), 1, MPI_INT, give_count.data(), 1, MPI_INT, comm);
  std::vector<int> give_displ(size + 1, 0);
  std::partial_sum(give_count.begin(), give_count.end(), give_displ.begin() + 1);
  std::vector<std::int64_t> give(give_displ[size]);
  MPI_Alltoallv(want.data(), want_count.data(), want_displ.data(), MPI_INT64_T, give.data(), give_count.data(),
                give_displ.data(), MPI_INT64_T, comm);
  for (int r = 0; r < size; ++r)
  {
    if (give_count[r] > 0)
    {
      L.send_ranks.push_back(r);
      L.send_counts.push_back(give_count[r]);
    }
    if (want_count[r] > 0)
    {
      L.recv_ranks.push_back(r);
      L.recv_counts.push_back(want_count[r]);
    }
  }
  L.send_idx.resize(give.size());
  for (std::size_t i = 0; i < give.size(); ++i)
    L.send_idx[i] = static_cast<std::int32_t>(give[i] - offset);
  return L;
}

/// NCCL communicator + forward-scatter plan for one dolfinx::common::IndexMap.
template <typename T>
class HaloPlan
{
public:
  HaloPlan(const dolfinx::common::IndexMap& map)
  {
    MPI_Comm comm = map.comm();
    int rank = 0, size = 1;
    MPI_Comm_rank(comm, &rank);
    MPI_Comm_size(comm, &size);
    char id[BD_UNIQUE_ID_BYTES] = {0};
    if (size > 1 && rank == 0)
      check(bd_comm_unique_id(id));
    if (size > 1)
      MPI_Bcast(id, BD_UNIQUE_ID_BYTES, MPI_BYTE, 0, comm);
    check(bd_comm_create(size > 1 ? id : nullptr, rank, size, &_comm));
    if (size == 1)
      return;
    const HaloLists L = compute_halo_lists(map);
    check(bd_halo_create(_comm, 8 * sizeof(T), map.size_local(), static_cast<std::int64_t>(map.num_ghosts()),
                         static_cast<int>(L.send_ranks.size()), L.send_ranks.data(), L.send_counts.data(),
                         L.send_idx.data(), static_cast<int>(L.recv_ranks.size()), L.recv_ranks.data(),
                         L.recv_counts.data(), &_halo));
    if (!L.ordered)
      check(bd_halo_set_unpack(_halo, L.order.data()));
  }
  ~HaloPlan()
  {
    bd_halo_destroy(_halo);
    bd_comm_destroy(_comm);
  }
  HaloPlan(const HaloPlan&) = delete;
  bd_comm* comm() const { return _comm; }
  bd_halo* halo() const { return _halo; }

private:
  bd_comm* _comm = nullptr;
  bd_halo* _halo = nullptr;
};

template <typename T>
class MatFreeLaplacianGPU
{
public:
  using value_type = T;

  /// Same signature as the reference constructor (src/laplacian.hpp:102-104).
  MatFreeLaplacianGPU(const dolfinx::fem::FunctionSpace<T>& V, const dolfinx::fem::DirichletBC<T>& bc, int degree,
                      int qmode, T constant, basix::quadrature::type quad_type)
      : _plan(*V.dofmap()->index_map)
  {
    auto mesh = V.mesh();
    auto dofmap = V.dofmap();
    auto [lcells, bcells] = benchdolfinx::compute_boundary_cells(*mesh->topology(), *dofmap);

    const std::int32_t ndofs = dofmap->index_map->size_local() + dofmap->index_map->num_ghosts();
    std::vector<std::int8_t> marker(ndofs, 0);
    bc.mark_dofs(marker);

    // quadrature degree and collocated element variant (src/laplacian.hpp:123-139)
    basix::element::lagrange_variant variant;
    int qdegree;
    const int p = degree + qmode;
    if (quad_type == basix::quadrature::type::gauss_jacobi)
    {
      variant = basix::element::lagrange_variant::gl_warped;
      qdegree = 2 * p;
    }
    else if (quad_type == basix::quadrature::type::gll)
    {
      variant = basix::element::lagrange_variant::gll_warped;
      qdegree = (p > 2) ? 2 * p - 2 : 2 * p - 1;
    }
    else
      throw std::runtime_error("Unsupported quadrature type for mat-free operator");

    auto [Gpoints, Gweights] = basix::quadrature::make_quadrature<T>(
        quad_type, basix::cell::type::hexahedron, basix::polyset::type::standard, qdegree);
    const auto& cmap = mesh->geometry().cmap();
    std::array<std::size_t, 4> phi_shape = cmap.tabulate_shape(1, Gweights.size());
    std::vector<T> phi_b(phi_shape[0] * phi_shape[1] * phi_shape[2] * phi_shape[3]);
    cmap.tabulate(1, Gpoints, {Gweights.size(), 3}, phi_b);
    std::vector<T> dphi_geometry(phi_b.begin() + phi_b.size() / 4, phi_b.end()); // derivatives only

    basix::FiniteElement<T> element0
        = basix::create_element<T>(basix::element::family::P, basix::cell::type::interval, degree,
                                   basix::element::lagrange_variant::gll_warped,
                                   basix::element::dpc_variant::unset, false);
    auto [points, weights] = basix::quadrature::make_quadrature<T>(
        quad_type, basix::cell::type::interval, basix::polyset::type::standard, qdegree);
    const int nq = static_cast<int>(weights.size());
    if (Gweights.size() != static_cast<std::size_t>(nq) * nq * nq)
      throw std::runtime_error("3D and 1D weight mismatch");
    basix::FiniteElement<T> element1
        = basix::create_element<T>(basix::element::family::P, basix::cell::type::interval, nq - 1, variant,
                                   basix::element::dpc_variant::unset, true);
    auto [phi0, shape_I] = basix::compute_interpolation_operator(element0, element1);
    const T eps = std::numeric_limits<T>::epsilon();
    bool identity = shape_I[0] == shape_I[1];
    for (std::size_t i = 0; i < shape_I[0]; ++i)
      for (std::size_t j = 0; j < shape_I[1]; ++j)
      {
        T& v = phi0[i * shape_I[1] + j];
        if (std::abs(v) < 5 * eps)
          v = 0;
        if (std::abs(v - (i == j ? T(1) : T(0))) > 5 * eps)
          identity = false;
      }
    if (qmode == 0 && !identity)
      throw std::runtime_error("Expecting identity matrix for qmode=0");
    auto [table, tshape] = element1.tabulate(1, points, {weights.size(), 1});
    std::vector<T> dphi1(table.begin() + table.size() / 2, table.end());

    bd_laplacian_desc d{};
    d.scalar_bits = 8 * sizeof(T);
    d.degree = degree;
    d.nq = nq;
    d.is_identity = identity;
    d.kappa = constant;
    const int tdim = mesh->topology()->dim();
    d.ncells = mesh->topology()->index_map(tdim)->size_local() + mesh->topology()->index_map(tdim)->num_ghosts();
    d.ndofs = ndofs;
    d.nnodes = mesh->geometry().x().size() / 3;
    d.n_lcells = lcells.size();
    d.n_bcells = bcells.size();
    d.cell_dofmap = dofmap->map().data_handle();
    d.xgeom = mesh->geometry().x().data();
    d.geometry_dofmap = mesh->geometry().dofmap().data_handle();
    d.bc_marker = marker.data();
    d.lcells = lcells.data();
    d.bcells = bcells.data();
    d.phi0 = phi0.data();
    d.dphi1 = dphi1.data();
    d.weights = Gweights.data();
    d.dphi_geometry = dphi_geometry.data();
    d.flags = 0; // host arrays: the handle keeps its own device copies
    check(bd_laplacian_create(&d, &_op));
  }
  ~MatFreeLaplacianGPU() { bd_laplacian_destroy(_op); }
  MatFreeLaplacianGPU(const MatFreeLaplacianGPU&) = delete;

  /// Same signature as src/laplacian.hpp:355-356; Vector is the reference's device vector alias
  /// (la::Vector with thrust::device_vector storage, src/vector.hpp:88-90).
  template <typename Vector>
  void apply(Vector& in, Vector& out)
  {
    check(bd_laplacian_apply_halo(_op, _plan.halo(), in.array().data().get(), out.array().data().get(), nullptr));
    if (cudaStreamSynchronize(nullptr) != cudaSuccess) // the reference's device_synchronize(), :347
      throw std::runtime_error("CUDA failure in MatFreeLaplacianGPU::apply");
  }

  bd_laplacian* handle() const { return _op; }
  bd_halo* halo() const { return _plan.halo(); }
  bd_comm* comm() const { return _plan.comm(); }

private:
  HaloPlan<T> _plan;
  bd_laplacian* _op = nullptr;
};

/// cg_solve (src/cg.hpp:89-169) for the operator above: the fused, device-resident loop.
template <typename T, typename Vector>
int cg_solve(MatFreeLaplacianGPU<T>& A, Vector& x, const Vector& b, int max_iter, T rtol = 0)
{
  int its = 0;
  check(bd_cg_solve(A.handle(), A.halo(), A.comm(), x.array().data().get(), b.array().data().get(), max_iter,
                    static_cast<double>(rtol), &its, nullptr));
  if (cudaStreamSynchronize(nullptr) != cudaSuccess)
    throw std::runtime_error("CUDA failure in cg_solve");
  return its;
}
} // namespace benchdolfinx::b200
