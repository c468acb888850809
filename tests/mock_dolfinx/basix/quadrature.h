#pragma once
#include <array>
#include <vector>
namespace basix
{
namespace cell { enum class type { interval, hexahedron }; }
namespace polyset { enum class type { standard }; }
namespace quadrature
{
enum class type { gauss_jacobi, gll };
template <typename T>
std::array<std::vector<T>, 2> make_quadrature(type, cell::type, polyset::type, int);
} // namespace quadrature
} // namespace basix
