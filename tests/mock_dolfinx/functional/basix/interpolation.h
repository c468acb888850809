#pragma once
#include "finite-element.h"
namespace basix
{
// Interpolation matrix from element0 (degree P on GLL nodes) to the nodes of element1: phi0[nq][P+1].
template <typename T>
std::pair<std::vector<T>, std::array<std::size_t, 2>> compute_interpolation_operator(const FiniteElement<T>& e0,
                                                                                      const FiniteElement<T>& e1)
{
  const int nd = e0.degree + 1, nq = e1.degree + 1, qmode = nq - nd;
  if (qmode != 0 && qmode != 1)
    throw std::runtime_error("mock basix: interpolation for qmode 0 / 1 only");
  std::vector<T> phi0(static_cast<size_t>(nq) * nd), dphi1(static_cast<size_t>(nq) * nq),
      w(static_cast<size_t>(nq) * nq * nq), dg(static_cast<size_t>(3) * nq * nq * nq * 8);
  std::vector<double> nodes(nd), pts(nq);
  int ident = 0;
  if (bd_tables(8 * sizeof(T), e0.degree, qmode, e1.variant == element::lagrange_variant::gl_warped, phi0.data(),
                dphi1.data(), w.data(), dg.data(), nodes.data(), pts.data(), &ident) != BD_SUCCESS)
    throw std::runtime_error(bd_last_error());
  return {phi0, {static_cast<std::size_t>(nq), static_cast<std::size_t>(nd)}};
}
} // namespace basix
