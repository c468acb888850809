#pragma once
#include <array>
#include <span>
#include <stdexcept>
#include <utility>
#include <vector>
#include "quadrature.h"
namespace basix
{
namespace element
{
enum class family { P };
enum class lagrange_variant { gll_warped, gl_warped };
enum class dpc_variant { unset };
} // namespace element
// 1-D Lagrange element of a degree on GLL (gll_warped) or Gauss (gl_warped) nodes.
template <typename T>
class FiniteElement
{
public:
  int degree = 1;
  element::lagrange_variant variant = element::lagrange_variant::gll_warped;
  // tabulate(1, points): [value | d/dx][npts][ndofs][1]; only used by the adapter with the element's own nodes
  std::pair<std::vector<T>, std::array<std::size_t, 4>> tabulate(int nd, std::span<const T> x,
                                                                 std::array<std::size_t, 2> shape) const
  {
    const int n = degree + 1;
    if (nd != 1 || static_cast<int>(shape[0]) != n || static_cast<int>(x.size()) != n)
      throw std::runtime_error("mock basix: tabulate is implemented at the element's own nodes only");
    std::vector<T> phi0(static_cast<size_t>(n) * n), dphi1(static_cast<size_t>(n) * n),
        w(static_cast<size_t>(n) * n * n), dg(static_cast<size_t>(3) * n * n * n * 8);
    std::vector<double> nodes(n), pts(n);
    int ident = 0;
    if (bd_tables(8 * sizeof(T), degree, 0, variant == element::lagrange_variant::gl_warped, phi0.data(), dphi1.data(),
                  w.data(), dg.data(), nodes.data(), pts.data(), &ident) != BD_SUCCESS)
      throw std::runtime_error(bd_last_error());
    std::vector<T> out(static_cast<size_t>(2) * n * n, T(0));
    for (int q = 0; q < n; ++q)
      out[static_cast<size_t>(q) * n + q] = T(1);             // values: identity at the own nodes
    std::copy(dphi1.begin(), dphi1.end(), out.begin() + static_cast<size_t>(n) * n); // derivatives [q][i]
    return {out, {2, static_cast<std::size_t>(n), static_cast<std::size_t>(n), 1}};
  }
};
template <typename T>
FiniteElement<T> create_element(element::family, cell::type, int degree, element::lagrange_variant v, element::dpc_variant,
                                bool)
{
  FiniteElement<T> e;
  e.degree = degree;
  e.variant = v;
  return e;
}
} // namespace basix
