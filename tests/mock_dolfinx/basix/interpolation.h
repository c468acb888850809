#pragma once
#include "finite-element.h"
namespace basix
{
template <typename T>
std::pair<std::vector<T>, std::array<std::size_t, 2>> compute_interpolation_operator(const FiniteElement<T>&,
                                                                                      const FiniteElement<T>&);
}
