#pragma once
#include <array>
#include <cstdint>
#include <mpi.h>
#include <span>
namespace dolfinx::common
{
class IndexMap
{
public:
  std::int32_t size_local() const;
  std::int32_t num_ghosts() const;
  std::array<std::int64_t, 2> local_range() const;
  std::span<const std::int64_t> ghosts() const;
  std::span<const int> owners() const;
  MPI_Comm comm() const;
};
} // namespace dolfinx::common
