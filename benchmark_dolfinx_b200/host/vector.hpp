// Device vector + vector functions: host-side mirror of the reference's src/vector.hpp on top of
// the C ABI.  Owned entries first, ghosts after (dolfinx::la::Vector layout, src/vector.hpp:88-90).
#pragma once

#include <cuda_runtime.h>

#include <cmath>
#include <cstdint>
#include <stdexcept>
#include <string>

#include "../../include/benchdolfinx_b200.h"

namespace benchdolfinx
{
inline void check(int rc)
{
  if (rc != BD_SUCCESS)
    throw std::runtime_error(bd_last_error());
}
inline void check_cuda(cudaError_t e)
{
  if (e != cudaSuccess)
    throw std::runtime_error(std::string("CUDA reports ") + cudaGetErrorString(e));
}

/// Parallel layout of a vector: the part of dolfinx::common::IndexMap the path uses.
struct IndexMap
{
  std::int64_t n_owned = 0, n_ghost = 0, size_global_ = 0;
  bd_comm* comm_ = nullptr; // may be null on one rank
  bd_halo* halo = nullptr;  // forward-scatter plan, may be null on one rank
  std::int64_t size_local() const { return n_owned; }
  std::int64_t num_ghosts() const { return n_ghost; }
  std::int64_t size_global() const { return size_global_; }
  bd_comm* comm() const { return comm_; }
};

template <typename T>
class Vector
{
public:
  using value_type = T;
  Vector(const IndexMap* map, int bs = 1) : _map(map), _bs(bs)
  {
    const std::size_t n = static_cast<std::size_t>(map->n_owned + map->n_ghost);
    check_cuda(cudaMalloc(&_data, (n > 0 ? n : 1) * sizeof(T)));
    check_cuda(cudaMemset(_data, 0, (n > 0 ? n : 1) * sizeof(T)));
  }
  ~Vector() { cudaFree(_data); }
  Vector(const Vector&) = delete;
  Vector& operator=(const Vector&) = delete;
  T* array() { return _data; }
  const T* array() const { return _data; }
  const IndexMap* index_map() const { return _map; }
  int bs() const { return _bs; }
  std::int64_t size() const { return _map->n_owned + _map->n_ghost; }
  /// owner -> ghost update (scatter_fwd_begin/end with pack/unpack functors in the reference)
  void scatter_fwd()
  {
    if (_map->halo)
    {
      check(bd_halo_begin(_map->halo, _data, nullptr));
      check(bd_halo_end(_map->halo, nullptr));
    }
  }

private:
  const IndexMap* _map;
  int _bs;
  T* _data = nullptr;
};

/// src/vector.hpp:159-176
template <typename U>
double inner_product(const U& a, const U& b)
{
  double r = 0;
  check(bd_inner_product(a.index_map()->comm(), 8 * sizeof(typename U::value_type),
                         a.bs() * a.index_map()->size_local(), a.array(), b.array(), &r));
  return r;
}
/// src/vector.hpp:196-218 (l2)
template <typename U>
double norm(const U& a)
{
  return std::sqrt(inner_product(a, a));
}
/// r = alpha * x + y over the owned entries (src/vector.hpp:228-240)
template <typename U, typename S>
void axpy(U& r, S alpha, const U& x, const U& y)
{
  check(bd_axpy(8 * sizeof(typename U::value_type), x.index_map()->size_local(), static_cast<double>(alpha),
                x.array(), y.array(), r.array(), nullptr));
}
/// src/vector.hpp:288-292
template <typename U, typename S>
void set_value(U& x, S value)
{
  check(bd_set_value(8 * sizeof(typename U::value_type), x.size(), static_cast<double>(value), x.array(), nullptr));
}
} // namespace benchdolfinx
