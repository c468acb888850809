#pragma once
#include <array>
#include <cstdint>
#include <memory>
#include <span>
#include <vector>
#include <dolfinx/common/IndexMap.h>
namespace dolfinx
{
template <typename V>
struct mock_mdspan
{
  const V* data_handle() const;
  std::size_t size() const;
};
namespace mesh
{
class Topology
{
public:
  int dim() const;
  std::shared_ptr<const common::IndexMap> index_map(int) const;
};
} // namespace mesh
namespace fem
{
template <typename T>
class CoordinateElement
{
public:
  std::array<std::size_t, 4> tabulate_shape(std::size_t, std::size_t) const;
  void tabulate(int, std::span<const T>, std::array<std::size_t, 2>, std::span<T>) const;
};
class DofMap
{
public:
  std::shared_ptr<const common::IndexMap> index_map;
  mock_mdspan<std::int32_t> map() const;
};
} // namespace fem
namespace mesh
{
template <typename T>
class Geometry
{
public:
  std::span<const T> x() const;
  mock_mdspan<std::int32_t> dofmap() const;
  const fem::CoordinateElement<T>& cmap() const;
};
template <typename T>
class Mesh
{
public:
  std::shared_ptr<const Topology> topology() const;
  const Geometry<T>& geometry() const;
};
} // namespace mesh
namespace fem
{
template <typename T>
class FunctionSpace
{
public:
  std::shared_ptr<const mesh::Mesh<T>> mesh() const;
  std::shared_ptr<const DofMap> dofmap() const;
};
} // namespace fem
} // namespace dolfinx
