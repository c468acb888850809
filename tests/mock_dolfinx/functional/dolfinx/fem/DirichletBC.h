#pragma once
#include <algorithm>
#include <cstdint>
#include <span>
#include <vector>
#include "FunctionSpace.h"
namespace dolfinx::fem
{
template <typename T>
class DirichletBC
{
public:
  std::vector<std::int8_t> marker;
  void mark_dofs(std::span<std::int8_t> m) const { std::copy(marker.begin(), marker.end(), m.begin()); }
};
} // namespace dolfinx::fem
