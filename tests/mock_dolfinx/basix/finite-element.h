#pragma once
#include <array>
#include <span>
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
template <typename T>
class FiniteElement
{
public:
  std::pair<std::vector<T>, std::array<std::size_t, 4>> tabulate(int, std::span<const T>,
                                                                 std::array<std::size_t, 2>) const;
};
template <typename T>
FiniteElement<T> create_element(element::family, cell::type, int, element::lagrange_variant, element::dpc_variant,
                                bool);
} // namespace basix
