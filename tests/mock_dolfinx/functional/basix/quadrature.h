// Functional stand-in for the Basix calls of the adapter, backed by the library's own table generator (bd_tables):
// Gauss-Lobatto-Legendre / Gauss-Legendre rules on [0,1] in ascending order (test infrastructure).
#pragma once
#include <array>
#include <stdexcept>
#include <vector>
#include "benchdolfinx_b200.h"
namespace basix
{
namespace cell { enum class type { interval, hexahedron }; }
namespace polyset { enum class type { standard }; }
namespace quadrature
{
enum class type { gauss_jacobi, gll };
// number of 1-D points for the quadrature degree the reference asks for (src/laplacian.hpp:123-139)
inline int mock_points(type t, int qdegree)
{
  if (t == type::gauss_jacobi)
    return qdegree / 2 + 1;                       // qdegree = 2p
  return qdegree % 2 ? (qdegree + 1) / 2 + 1 : qdegree / 2 + 2; // 2p-1 (p <= 2) or 2p-2
}
template <typename T>
struct mock_rule
{
  std::vector<double> pts;
  std::vector<T> w3; // nq^3 weights
};
template <typename T>
mock_rule<T> mock_make_rule(type t, int nq)
{
  const int bits = 8 * sizeof(T);
  mock_rule<T> r;
  r.pts.resize(nq);
  r.w3.resize(static_cast<size_t>(nq) * nq * nq);
  std::vector<T> phi0(static_cast<size_t>(nq) * nq), dphi1(static_cast<size_t>(nq) * nq),
      dg(static_cast<size_t>(3) * nq * nq * nq * 8);
  std::vector<double> nodes(nq);
  int ident = 0;
  if (bd_tables(bits, nq - 1, 0, t == type::gauss_jacobi, phi0.data(), dphi1.data(), r.w3.data(), dg.data(), nodes.data(),
                r.pts.data(), &ident) != BD_SUCCESS)
    throw std::runtime_error(bd_last_error());
  return r;
}
// returns {points (npts x dim, row-major), weights}
template <typename T>
std::array<std::vector<T>, 2> make_quadrature(type t, cell::type ct, polyset::type, int qdegree)
{
  const int nq = mock_points(t, qdegree);
  mock_rule<T> r = mock_make_rule<T>(t, nq);
  std::array<std::vector<T>, 2> out;
  if (ct == cell::type::interval)
  {
    out[0].assign(r.pts.begin(), r.pts.end());
    out[1].resize(nq);
    double wsum = 0;
    for (T w : r.w3)
      wsum += w;
    for (int a = 0; a < nq; ++a)
    {
      double s = 0;
      for (int b = 0; b < nq * nq; ++b)
        s += r.w3[static_cast<size_t>(a) * nq * nq + b];
      out[1][a] = static_cast<T>(s / wsum); // weights of the 1-D rule sum to 1, so do the 3-D ones
    }
    return out;
  }
  out[0].resize(static_cast<size_t>(nq) * nq * nq * 3);
  for (int q0 = 0; q0 < nq; ++q0)
    for (int q1 = 0; q1 < nq; ++q1)
      for (int q2 = 0; q2 < nq; ++q2)
      {
        const size_t q = (static_cast<size_t>(q0) * nq + q1) * nq + q2;
        out[0][3 * q + 0] = static_cast<T>(r.pts[q0]);
        out[0][3 * q + 1] = static_cast<T>(r.pts[q1]);
        out[0][3 * q + 2] = static_cast<T>(r.pts[q2]);
      }
  out[1] = r.w3;
  return out;
}
} // namespace quadrature
} // namespace basix
