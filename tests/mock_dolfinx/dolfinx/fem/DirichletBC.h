#pragma once
#include <cstdint>
#include <span>
#include "FunctionSpace.h"
namespace dolfinx::fem
{
template <typename T>
class DirichletBC
{
public:
  void mark_dofs(std::span<std::int8_t>) const;
};
} // namespace dolfinx::fem
