// mock of the reference's src/mesh.hpp (compute_boundary_cells, src/mesh.cpp:154-188)
#pragma once
#include <array>
#include <cstdint>
#include <vector>
#include <dolfinx/fem/FunctionSpace.h>
namespace benchdolfinx
{
std::array<std::vector<std::int32_t>, 2> compute_boundary_cells(const dolfinx::mesh::Topology&,
                                                                const dolfinx::fem::DofMap&);
}
