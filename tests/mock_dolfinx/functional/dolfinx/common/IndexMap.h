// Functional stand-in for dolfinx::common::IndexMap: plain data, filled by the test driver (test infrastructure).
#pragma once
#include <array>
#include <cstdint>
#include <mpi.h>
#include <span>
#include <vector>
namespace dolfinx::common
{
class IndexMap
{
public:
  std::int32_t n_local = 0;
  std::int64_t offset = 0;
  std::vector<std::int64_t> ghost_ids;
  std::vector<int> ghost_owners;
  MPI_Comm communicator = nullptr;
  std::int32_t size_local() const { return n_local; }
  std::int32_t num_ghosts() const { return static_cast<std::int32_t>(ghost_ids.size()); }
  std::array<std::int64_t, 2> local_range() const { return {offset, offset + n_local}; }
  std::span<const std::int64_t> ghosts() const { return ghost_ids; }
  std::span<const int> owners() const { return ghost_owners; }
  MPI_Comm comm() const { return communicator; }
};
} // namespace dolfinx::common
