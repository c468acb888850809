// C-ABI shim around the reference's own, UNMODIFIED CPU kernels (oracle; test
// infrastructure, never part of the product).
//
// The two headers below are compiled from where they lie under the reference
// tree (-I$(REF_SRC), see oracle/Makefile); nothing is copied into this repo.
//   stiffness_operator_cpu<T,P,Q>   /root/reference/src/laplacian_cpu.hpp:57-146  (qmode 0 only)
//   geometry_computation_cpu<T,Q>   /root/reference/src/geometry_cpu.hpp:25-112
// The three <basix/...> includes of laplacian_cpu.hpp:8-10 are satisfied by the
// empty headers in oracle/stubs/ (the kernels use nothing from Basix).
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <span>
#include <vector>

#include "geometry_cpu.hpp"
#include "laplacian_cpu.hpp"

namespace
{
template <typename T, int P>
void stiffness(const void* u, const void* constants, void* b, const void* G, const void* phi0,
               const void* dphi1, const std::int32_t* dofmap, const int* entities, int n,
               const std::int8_t* bc)
{
  benchdolfinx::stiffness_operator_cpu<T, P, P + 1>(
      static_cast<const T*>(u), static_cast<const T*>(constants), static_cast<T*>(b),
      static_cast<const T*>(G), static_cast<const T*>(phi0), static_cast<const T*>(dphi1), dofmap,
      entities, n, bc, true);
}

template <typename T, int Q>
void geometry(const void* xgeom, std::size_t nx, void* G, const std::int32_t* gdofmap,
              std::size_t ngd, const void* dphi, const void* weights, const int* entities, int n)
{
  constexpr std::size_t nq = Q * Q * Q;
  benchdolfinx::geometry_computation_cpu<T, Q>(
      std::span<const T>(static_cast<const T*>(xgeom), nx),
      std::span<T>(static_cast<T*>(G), std::size_t(n) * nq * 6),
      std::span<const std::int32_t>(gdofmap, ngd),
      std::span<const T>(static_cast<const T*>(dphi), 3 * nq * 8),
      std::span<const T>(static_cast<const T*>(weights), nq), std::span<const int>(entities, n));
}
} // namespace

extern "C"
{
// returns 0 on success, 1 for an unsupported (scalar, degree)
int ref_stiffness_operator_cpu(int scalar_bits, int degree, const void* u, const void* constants,
                               void* b, const void* G, const void* phi0, const void* dphi1,
                               const std::int32_t* dofmap, const int* entities, int n,
                               const std::int8_t* bc)
{
#define CASE(P)                                                                                    \
  case P:                                                                                          \
    if (scalar_bits == 64)                                                                         \
      stiffness<double, P>(u, constants, b, G, phi0, dphi1, dofmap, entities, n, bc);              \
    else                                                                                           \
      stiffness<float, P>(u, constants, b, G, phi0, dphi1, dofmap, entities, n, bc);               \
    return 0;
  if (scalar_bits != 32 && scalar_bits != 64)
    return 1;
  switch (degree)
  {
    CASE(1) CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7)
  default:
    return 1;
  }
#undef CASE
}

int ref_geometry_computation_cpu(int scalar_bits, int nq1d, const void* xgeom, std::size_t nx,
                                 void* G, const std::int32_t* gdofmap, std::size_t ngd,
                                 const void* dphi, const void* weights, const int* entities, int n)
{
#define CASE(Q)                                                                                    \
  case Q:                                                                                          \
    if (scalar_bits == 64)                                                                         \
      geometry<double, Q>(xgeom, nx, G, gdofmap, ngd, dphi, weights, entities, n);                 \
    else                                                                                           \
      geometry<float, Q>(xgeom, nx, G, gdofmap, ngd, dphi, weights, entities, n);                  \
    return 0;
  if (scalar_bits != 32 && scalar_bits != 64)
    return 1;
  switch (nq1d)
  {
    CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8) CASE(9)
  default:
    return 1;
  }
#undef CASE
}
}
