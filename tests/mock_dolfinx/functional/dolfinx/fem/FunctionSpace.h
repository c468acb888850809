// Functional stand-ins for the DOLFINx classes the adapter reads (mesh, geometry, dofmap, function space): host
// copies of the structured stand-in mesh's arrays (test infrastructure).
#pragma once
#include <array>
#include <cstdint>
#include <memory>
#include <span>
#include <stdexcept>
#include <vector>
#include <dolfinx/common/IndexMap.h>
#include "benchdolfinx_b200.h"
namespace dolfinx
{
template <typename V>
struct mock_mdspan
{
  const std::vector<V>* v = nullptr;
  const V* data_handle() const { return v->data(); }
  std::size_t size() const { return v->size(); }
};
namespace mesh
{
class Topology
{
public:
  std::shared_ptr<common::IndexMap> cell_map = std::make_shared<common::IndexMap>();
  std::vector<std::int32_t> lcells, bcells; // what compute_boundary_cells returns for this mesh
  int dim() const { return 3; }
  std::shared_ptr<const common::IndexMap> index_map(int) const { return cell_map; }
};
} // namespace mesh
namespace fem
{
// Trilinear hexahedron, vertices in tensor-product order (v = 4a + 2b + c): values and first derivatives at points.
template <typename T>
class CoordinateElement
{
public:
  std::array<std::size_t, 4> tabulate_shape(std::size_t nd, std::size_t npts) const { return {nd == 1 ? 4u : 1u, npts, 8, 1}; }
  void tabulate(int nd, std::span<const T> x, std::array<std::size_t, 2> shape, std::span<T> out) const
  {
    if (nd != 1 || shape[1] != 3)
      throw std::runtime_error("mock dolfinx: tabulate(1, points[npts][3])");
    const std::size_t npts = shape[0];
    for (std::size_t q = 0; q < npts; ++q)
      for (int v = 0; v < 8; ++v)
      {
        const int abc[3] = {(v >> 2) & 1, (v >> 1) & 1, v & 1};
        T val[3], der[3];
        for (int d = 0; d < 3; ++d)
        {
          const T xi = x[3 * q + d];
          val[d] = abc[d] ? xi : T(1) - xi;
          der[d] = abc[d] ? T(1) : T(-1);
        }
        out[(0 * npts + q) * 8 + v] = val[0] * val[1] * val[2];
        out[(1 * npts + q) * 8 + v] = der[0] * val[1] * val[2];
        out[(2 * npts + q) * 8 + v] = val[0] * der[1] * val[2];
        out[(3 * npts + q) * 8 + v] = val[0] * val[1] * der[2];
      }
  }
};
class DofMap
{
public:
  std::shared_ptr<common::IndexMap> index_map = std::make_shared<common::IndexMap>();
  std::vector<std::int32_t> cell_dofs;
  mock_mdspan<std::int32_t> map() const { return {&cell_dofs}; }
};
} // namespace fem
namespace mesh
{
template <typename T>
class Geometry
{
public:
  std::vector<T> coords;
  std::vector<std::int32_t> cell_nodes;
  fem::CoordinateElement<T> element;
  std::span<const T> x() const { return coords; }
  mock_mdspan<std::int32_t> dofmap() const { return {&cell_nodes}; }
  const fem::CoordinateElement<T>& cmap() const { return element; }
};
template <typename T>
class Mesh
{
public:
  std::shared_ptr<Topology> topo = std::make_shared<Topology>();
  Geometry<T> geom;
  std::shared_ptr<const Topology> topology() const { return topo; }
  const Geometry<T>& geometry() const { return geom; }
};
} // namespace mesh
namespace fem
{
template <typename T>
class FunctionSpace
{
public:
  std::shared_ptr<mesh::Mesh<T>> msh = std::make_shared<mesh::Mesh<T>>();
  std::shared_ptr<DofMap> dm = std::make_shared<DofMap>();
  std::shared_ptr<const mesh::Mesh<T>> mesh() const { return msh; }
  std::shared_ptr<const DofMap> dofmap() const { return dm; }
};
} // namespace fem
} // namespace dolfinx
