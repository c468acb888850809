// Shared helpers for the benchdolfinx_b200 CUDA library (sm_100a only).
#pragma once

#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <string>

#include "../../include/benchdolfinx_b200.h"

namespace bd
{
// Thread-local message of the last failure; returned by bd_last_error().
std::string& last_error_slot();
int fail(int code, const char* fmt, ...);

#define BD_CUDA_CHECK(expr)                                                                        \
  do                                                                                               \
  {                                                                                                \
    cudaError_t bd_err_ = (expr);                                                                  \
    if (bd_err_ != cudaSuccess)                                                                    \
      return ::bd::fail(BD_ERR_CUDA, "%s:%d: CUDA reports %s for `%s`", __FILE__, __LINE__,        \
                        cudaGetErrorString(bd_err_), #expr);                                       \
  } while (0)

#define BD_CHECK(expr)                                                                             \
  do                                                                                               \
  {                                                                                                \
    int bd_rc_ = (expr);                                                                           \
    if (bd_rc_ != BD_SUCCESS)                                                                      \
      return bd_rc_;                                                                               \
  } while (0)

#define BD_REQUIRE(cond, ...)                                                                      \
  do                                                                                               \
  {                                                                                                \
    if (!(cond))                                                                                   \
      return ::bd::fail(BD_ERR_INVALID_ARGUMENT, __VA_ARGS__);                                     \
  } while (0)

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

// Number of SMs of the current device (cached per device).
int sm_count();

// B200: 148 SMs.  Grids of the streaming kernels are sized as a multiple of this.
constexpr int kB200SMs = 148;

// Test hook (bd_debug_set_max_grid): upper bound on the grid of every persistent stiffness kernel, 0 = none.
// With a grid of 2-3 CTAs even a few-thousand-cell test mesh drives every warp through many batches, i.e.
// through the steady state of the rings / prefetch rotations that the benchmark sizes spend all their time in.
int debug_max_grid();
inline std::int64_t clamp_grid(std::int64_t want, std::int64_t resident)
{
  std::int64_t g = want < resident ? want : resident;
  const int cap = debug_max_grid();
  if (cap > 0 && g > cap)
    g = cap;
  return g < 1 ? 1 : g;
}

template <typename T>
struct scalar_traits;
template <>
struct scalar_traits<float>
{
  static constexpr int bits = 32;
};
template <>
struct scalar_traits<double>
{
  static constexpr int bits = 64;
};

// Owned-or-borrowed device array.
template <typename T>
struct DeviceArray
{
  T* ptr = nullptr;
  size_t count = 0;
  bool owned = false;
  ~DeviceArray() { release(); }
  DeviceArray() = default;
  DeviceArray(const DeviceArray&) = delete;
  DeviceArray& operator=(const DeviceArray&) = delete;
  void release()
  {
    if (owned && ptr)
      cudaFree(ptr);
    ptr = nullptr;
    count = 0;
    owned = false;
  }
  int alloc(size_t n)
  {
    release();
    if (n == 0)
      return BD_SUCCESS;
    BD_CUDA_CHECK(cudaMalloc(&ptr, n * sizeof(T)));
    count = n;
    owned = true;
    return BD_SUCCESS;
  }
  // Copy from a host or device pointer (or borrow a device pointer).
  int assign(const T* src, size_t n, bool borrow_if_device)
  {
    release();
    if (n == 0)
      return BD_SUCCESS;
    if (!src)
      return fail(BD_ERR_INVALID_ARGUMENT, "null array of %zu entries", n);
    cudaPointerAttributes attr{};
    cudaError_t e = cudaPointerGetAttributes(&attr, src);
    if (e != cudaSuccess)
    {
      cudaGetLastError();
      attr.type = cudaMemoryTypeUnregistered;
    }
    const bool on_device = attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
    if (on_device && borrow_if_device)
    {
      ptr = const_cast<T*>(src);
      count = n;
      owned = false;
      return BD_SUCCESS;
    }
    BD_CHECK(alloc(n));
    BD_CUDA_CHECK(cudaMemcpy(ptr, src, n * sizeof(T), cudaMemcpyDefault));
    return BD_SUCCESS;
  }
};

} // namespace bd
