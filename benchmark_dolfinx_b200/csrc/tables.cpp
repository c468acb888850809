// 1-D tables of the operator, computed on the host in extended precision.
//
// Replaces what the reference constructor obtains from Basix v0.10 (src/laplacian.hpp:123-212):
// the quadrature rule (GLL, or Gauss-Legendre with --use_gauss), the interpolation table phi0
// from the degree-P GLL element to the element collocated with the quadrature points, its
// derivative table dphi1, the tensor-product weights and the derivative tabulation of the
// trilinear coordinate element.  nq = P + 1 + qmode in every case the reference supports
// (q_map, :125-134).  1-D nodes are in ascending order.
#include <cmath>
#include <limits>
#include <vector>

#include "common.cuh"

namespace
{
using ld = long double;

// P_n(x) and P_n'(x), valid at +-1
void legendre(int n, ld x, ld& p, ld& dp)
{
  ld p0 = 1, d0 = 0;
  if (n == 0)
  {
    p = p0;
    dp = d0;
    return;
  }
  ld p1 = x, d1 = 1;
  for (int k = 2; k <= n; ++k)
  {
    ld p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k;
    ld d2 = d0 + (2 * k - 1) * p1;
    p0 = p1;
    p1 = p2;
    d0 = d1;
    d1 = d2;
  }
  p = p1;
  dp = d1;
}

void gll(int n, std::vector<ld>& x, std::vector<ld>& w)
{
  const int N = n - 1;
  const ld pi = acosl(-1.0L);
  x.resize(n);
  w.resize(n);
  for (int i = 0; i < n; ++i)
  {
    ld xi = -cosl(pi * i / N);
    if (i > 0 && i < N)
    {
      for (int it = 0; it < 100; ++it)
      {
        // q = P_{N+1} - P_{N-1} vanishes at the interior GLL nodes
        ld pa, da, pb, db;
        legendre(N + 1, xi, pa, da);
        legendre(N - 1, xi, pb, db);
        ld dx = (pa - pb) / (da - db);
        xi -= dx;
        if (fabsl(dx) < 1e-19L)
          break;
      }
    }
    else
      xi = (i == 0) ? -1 : 1;
    x[i] = xi;
  }
  for (int i = 0; i < n; ++i)
  {
    ld p, dp;
    legendre(N, x[i], p, dp);
    w[i] = 2 / (N * (N + 1) * p * p);
  }
  for (int i = 0; i < n; ++i)
  {
    x[i] = (x[i] + 1) / 2;
    w[i] /= 2;
  }
}

void gauss_legendre(int n, std::vector<ld>& x, std::vector<ld>& w)
{
  const ld pi = acosl(-1.0L);
  x.resize(n);
  w.resize(n);
  for (int i = 0; i < n; ++i)
  {
    ld xi = -cosl(pi * (i + 0.75L) / (n + 0.5L));
    for (int it = 0; it < 100; ++it)
    {
      ld p, dp;
      legendre(n, xi, p, dp);
      ld dx = p / dp;
      xi -= dx;
      if (fabsl(dx) < 1e-19L)
        break;
    }
    ld p, dp;
    legendre(n, xi, p, dp);
    x[i] = (xi + 1) / 2;
    w[i] = 1 / ((1 - xi * xi) * dp * dp);
  }
}

template <typename T>
int tables_impl(int degree, int qmode, int use_gauss, T* phi0, T* dphi1, T* weights, T* dphi_geometry,
                double* dof_nodes, double* qpts, int* is_identity)
{
  const int nd = degree + 1, nq = degree + 1 + qmode;
  std::vector<ld> xd, wd, xq, wq;
  gll(nd, xd, wd);
  if (use_gauss)
    gauss_legendre(nq, xq, wq);
  else
    gll(nq, xq, wq);
  const T eps = std::numeric_limits<T>::epsilon();
  bool identity = (nd == nq);
  for (int q = 0; q < nq; ++q)
    for (int i = 0; i < nd; ++i)
    {
      ld v = 1;
      for (int j = 0; j < nd; ++j)
        if (j != i)
          v *= (xq[q] - xd[j]) / (xd[i] - xd[j]);
      T t = static_cast<T>(v);
      if (std::abs(t) < 5 * eps) // src/laplacian.hpp:188-193
        t = 0;
      if (phi0)
        phi0[q * nd + i] = t;
      if (nd == nq && std::abs(t - (q == i ? T(1) : T(0))) > 5 * eps) // matrix_is_identity, :34-54
        identity = false;
    }
  if (is_identity)
    *is_identity = identity ? 1 : 0;
  if (dphi1)
  {
    std::vector<ld> bw(nq, 1);
    for (int i = 0; i < nq; ++i)
      for (int j = 0; j < nq; ++j)
        if (j != i)
          bw[i] /= (xq[i] - xq[j]);
    for (int q = 0; q < nq; ++q)
    {
      ld diag = 0;
      for (int i = 0; i < nq; ++i)
        if (i != q)
        {
          ld v = (bw[i] / bw[q]) / (xq[q] - xq[i]);
          dphi1[q * nq + i] = static_cast<T>(v);
          diag -= v;
        }
      dphi1[q * nq + q] = static_cast<T>(diag);
    }
  }
  const int nq3 = nq * nq * nq;
  if (weights)
    for (int a = 0; a < nq; ++a)
      for (int b = 0; b < nq; ++b)
        for (int c = 0; c < nq; ++c)
          weights[(a * nq + b) * nq + c] = static_cast<T>(wq[a] * wq[b] * wq[c]);
  if (dphi_geometry)
    for (int d = 0; d < 3; ++d)
      for (int a = 0; a < nq; ++a)
        for (int b = 0; b < nq; ++b)
          for (int c = 0; c < nq; ++c)
          {
            const int q = (a * nq + b) * nq + c;
            const ld pt[3] = {xq[a], xq[b], xq[c]};
            for (int v = 0; v < 8; ++v)
            {
              const int bit[3] = {(v >> 2) & 1, (v >> 1) & 1, v & 1};
              ld val = 1;
              for (int e = 0; e < 3; ++e)
              {
                if (e == d)
                  val *= bit[e] ? 1 : -1;
                else
                  val *= bit[e] ? pt[e] : 1 - pt[e];
              }
              dphi_geometry[(static_cast<size_t>(d) * nq3 + q) * 8 + v] = static_cast<T>(val);
            }
          }
  if (dof_nodes)
    for (int i = 0; i < nd; ++i)
      dof_nodes[i] = static_cast<double>(xd[i]);
  if (qpts)
    for (int i = 0; i < nq; ++i)
      qpts[i] = static_cast<double>(xq[i]);
  return BD_SUCCESS;
}
} // namespace

extern "C" int bd_tables(int scalar_bits, int degree, int qmode, int use_gauss, void* phi0, void* dphi1,
                         void* weights, void* dphi_geometry, double* dof_nodes, double* qpts,
                         int* is_identity)
{
  BD_REQUIRE(scalar_bits == 32 || scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  if (degree < 1 || degree > 7)
    return bd::fail(BD_ERR_UNSUPPORTED, "Unsupported degree [operator]");
  BD_REQUIRE(qmode == 0 || qmode == 1, "Invalid qmode.");
  if (scalar_bits == 64)
    return tables_impl<double>(degree, qmode, use_gauss, static_cast<double*>(phi0),
                               static_cast<double*>(dphi1), static_cast<double*>(weights),
                               static_cast<double*>(dphi_geometry), dof_nodes, qpts, is_identity);
  return tables_impl<float>(degree, qmode, use_gauss, static_cast<float*>(phi0),
                            static_cast<float*>(dphi1), static_cast<float*>(weights),
                            static_cast<float*>(dphi_geometry), dof_nodes, qpts, is_identity);
}

extern "C" int bd_compute_mesh_size(int64_t ndofs_global, int degree, int64_t nx_out[3])
{
  // src/mesh.cpp:117-152: same start value, search window, loop order and strict '<'
  BD_REQUIRE(nx_out && degree >= 1 && ndofs_global >= 1, "bd_compute_mesh_size: bad argument");
  const double nx_approx = (std::pow(static_cast<double>(ndofs_global), 1. / 3.) - 1) / degree;
  const int64_t n0 = static_cast<int64_t>(nx_approx + 0.5);
  nx_out[0] = nx_out[1] = nx_out[2] = n0;
  auto misfit_of = [&](int64_t a, int64_t b, int64_t c) {
    int64_t m = (a * degree + 1) * (b * degree + 1) * (c * degree + 1) - ndofs_global;
    return m < 0 ? -m : m;
  };
  int64_t best = misfit_of(n0, n0, n0);
  const int64_t lo = n0 - 5 > 1 ? n0 - 5 : 1;
  for (int64_t a = lo; a < n0 + 6; ++a)
    for (int64_t b = lo; b < n0 + 6; ++b)
      for (int64_t c = lo; c < n0 + 6; ++c)
      {
        const int64_t m = misfit_of(a, b, c);
        if (m < best)
        {
          best = m;
          nx_out[0] = a;
          nx_out[1] = b;
          nx_out[2] = c;
        }
      }
  return BD_SUCCESS;
}
