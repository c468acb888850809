// Host-side mirror of cg_solve (reference src/cg.hpp:89-169).
#pragma once

#include <type_traits>

#include "laplacian.hpp"

namespace benchdolfinx
{
/// Generic duck-typed version: same recurrence, update order and stopping rule as the reference;
/// used for operators other than the matrix-free one (the CSR check operator).
template <typename Operator, typename Vector, typename S>
int cg_solve(Operator& A, Vector& x, const Vector& b, int max_iter, S rtol = 0.0)
{
  using T = typename Vector::value_type;
  static_assert(std::is_same<typename Operator::value_type, T>(), "Type mismatch between Vector and Operator");
  Vector r(x.index_map(), x.bs()), y(x.index_map(), x.bs()), p(x.index_map(), x.bs());
  A.apply(x, y);
  axpy(r, -1, y, b);
  axpy(p, 0.0, r, r);
  double rnorm0 = inner_product(p, r), rnorm = rnorm0;
  const double rtol2 = static_cast<double>(rtol) * static_cast<double>(rtol);
  int k = 0;
  while (k < max_iter)
  {
    ++k;
    A.apply(p, y); // includes the forward scatter of p
    const double alpha = rnorm / inner_product(p, y);
    axpy(x, alpha, p, x);
    axpy(r, -alpha, y, r);
    const double rnorm_new = inner_product(r, r);
    const double beta = rnorm_new / rnorm;
    rnorm = rnorm_new;
    if (rnorm / rnorm0 < rtol2)
      break;
    axpy(p, beta, p, r);
  }
  return k;
}

/// Matrix-free operator: the whole solve is the library's fused, device-resident loop.
template <typename T, typename S>
int cg_solve(MatFreeLaplacianGPU<T>& A, Vector<T>& x, const Vector<T>& b, int max_iter, S rtol = 0.0)
{
  int its = 0;
  check(bd_cg_solve(A.handle(), A.halo(), A.comm(), x.array(), b.array(), max_iter, static_cast<double>(rtol),
                    &its, nullptr));
  return its;
}
} // namespace benchdolfinx
