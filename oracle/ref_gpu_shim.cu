// C-ABI shim around the reference's own, UNMODIFIED GPU kernels, recompiled for sm_100a
// (oracle; test / baseline infrastructure, never part of the product).
//
// The headers are compiled from where they lie under the reference tree (-I$(REF_SRC), see
// oracle/Makefile); nothing is copied into this repo.
//   stiffness_operator_gpu<T,P,Q>     /root/reference/src/laplacian_gpu.hpp:91-426
//   geometry_computation_gpu<T,Q>     /root/reference/src/geometry_gpu.hpp:26-132
// Launch shapes are the reference's (src/laplacian.hpp:298-306, :249-266): one CTA per cell,
// block (Q,Q,Q) resp. Q^3 threads, default stream.  Used by tests (a second, GPU-side checker)
// and by bench.py as "the kernel to beat" on the same box.
#define USE_CUDA 1
#include <cassert>
#include <cstdint>
#include <cuda_runtime.h>

#include "laplacian_gpu.hpp"

namespace
{
template <typename T, int P, int Q>
int stiffness(const void* u, const void* constants, void* b, const void* G, const void* phi0,
              const void* dphi1, const std::int32_t* dofmap, const int* entities, int n,
              const std::int8_t* bc, bool identity)
{
  dim3 block(Q, Q, Q), grid(n);
  benchdolfinx::stiffness_operator_gpu<T, P, Q><<<grid, block>>>(
      static_cast<const T*>(u), static_cast<const T*>(constants), static_cast<T*>(b),
      static_cast<const T*>(G), static_cast<const T*>(phi0), static_cast<const T*>(dphi1), dofmap,
      entities, n, bc, identity);
  return cudaGetLastError() == cudaSuccess ? 0 : 2;
}

template <typename T, int Q>
int geometry(const void* xgeom, void* G, const std::int32_t* gdofmap, const void* dphi,
             const void* weights, const int* entities, int n)
{
  benchdolfinx::geometry_computation_gpu<T, Q><<<n, Q * Q * Q>>>(
      static_cast<const T*>(xgeom), static_cast<T*>(G), gdofmap, static_cast<const T*>(dphi),
      static_cast<const T*>(weights), entities, n);
  return cudaGetLastError() == cudaSuccess ? 0 : 2;
}
} // namespace

extern "C"
{
// all pointers are device pointers; returns 0 on success, 1 unsupported, 2 launch failure
int ref_stiffness_operator_gpu(int scalar_bits, int degree, int nq, const void* u, const void* constants,
                               void* b, const void* G, const void* phi0, const void* dphi1,
                               const std::int32_t* dofmap, const int* entities, int n,
                               const std::int8_t* bc, int identity)
{
  if (n == 0)
    return 0;
#define CASE(P, Q)                                                                                 \
  if (degree == P && nq == Q)                                                                      \
    return scalar_bits == 64                                                                       \
               ? stiffness<double, P, Q>(u, constants, b, G, phi0, dphi1, dofmap, entities, n, bc, identity) \
               : stiffness<float, P, Q>(u, constants, b, G, phi0, dphi1, dofmap, entities, n, bc, identity);
  CASE(1, 2) CASE(2, 3) CASE(3, 4) CASE(4, 5) CASE(5, 6) CASE(6, 7) CASE(7, 8)
  CASE(1, 3) CASE(2, 4) CASE(3, 5) CASE(4, 6) CASE(5, 7) CASE(6, 8) CASE(7, 9)
#undef CASE
  return 1;
}

int ref_geometry_computation_gpu(int scalar_bits, int nq, const void* xgeom, void* G,
                                 const std::int32_t* gdofmap, const void* dphi, const void* weights,
                                 const int* entities, int n)
{
  if (n == 0)
    return 0;
#define CASE(Q)                                                                                    \
  if (nq == Q)                                                                                     \
    return scalar_bits == 64 ? geometry<double, Q>(xgeom, G, gdofmap, dphi, weights, entities, n)  \
                             : geometry<float, Q>(xgeom, G, gdofmap, dphi, weights, entities, n);
  CASE(2) CASE(3) CASE(4) CASE(5) CASE(6) CASE(7) CASE(8) CASE(9)
#undef CASE
  return 1;
}
}
