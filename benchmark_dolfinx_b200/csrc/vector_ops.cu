// Vector kernels: replace the thrust calls of the reference's src/vector.hpp and src/cg.hpp
// (thrust::transform axpy, thrust::inner_product, thrust::fill, pack_gpu / unpack_gpu).
// All of them are pure HBM streams: grid-stride loops over a grid that is a multiple of the SM
// count, dot products reduced with warp shuffles, per-CTA partials summed in a fixed order by a
// second tiny kernel (bitwise reproducible, result stays in device memory).
#include "vector_ops.cuh"

#include <mutex>

namespace bd
{
template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    axpy_kernel(std::int64_t n, T alpha, const T* __restrict__ x, const T* __restrict__ y,
                T* __restrict__ r)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
    r[i] = x[i] * alpha + y[i]; // same expression as the reference functor (vx * alpha + vy)
}

template <typename T>
__global__ void __launch_bounds__(kVecBlock) fill_kernel(std::int64_t n, T value, T* __restrict__ x)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
    x[i] = value;
}

template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    dot_partial_kernel(std::int64_t n, const T* __restrict__ a, const T* __restrict__ b,
                       double* __restrict__ partials)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  double s = 0;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
    s += static_cast<double>(a[i]) * static_cast<double>(b[i]);
  s = block_sum(s);
  if (threadIdx.x == 0)
    partials[blockIdx.x] = s;
}

__global__ void __launch_bounds__(kVecBlock)
    reduce_partials_kernel(int nparts, const double* __restrict__ partials,
                           double* __restrict__ result)
{
  double s = 0;
  for (int i = threadIdx.x; i < nparts; i += kVecBlock)
    s += partials[i];
  s = block_sum(s);
  if (threadIdx.x == 0)
    result[0] = s;
}

int launch_reduce_partials(int nparts, const double* partials, double* result, cudaStream_t stream)
{
  reduce_partials_kernel<<<1, kVecBlock, 0, stream>>>(nparts, partials, result);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    pack_kernel(std::int64_t n, const std::int32_t* __restrict__ idx, const T* __restrict__ in,
                T* __restrict__ out)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
    out[i] = in[idx[i]];
}

template <typename T>
__global__ void __launch_bounds__(kVecBlock)
    unpack_kernel(std::int64_t n, const std::int32_t* __restrict__ idx, const T* __restrict__ in,
                  T* __restrict__ out)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += stride)
    out[idx[i]] = in[i];
}

template <typename T>
int dot_impl(std::int64_t n, const void* a, const void* b, double* result, double* scratch,
             cudaStream_t stream)
{
  const int grid = vec_grid(n);
  dot_partial_kernel<T><<<grid, kVecBlock, 0, stream>>>(n, static_cast<const T*>(a),
                                                        static_cast<const T*>(b), scratch);
  BD_CUDA_CHECK(cudaGetLastError());
  return launch_reduce_partials(grid, scratch, result, stream);
}
} // namespace bd

using namespace bd;

#define BD_BITS_CHECK(bits)                                                                        \
  BD_REQUIRE((bits) == 32 || (bits) == 64, "Invalid float size. Must be 32 or 64.")

extern "C" int bd_axpy(int scalar_bits, int64_t n, double alpha, const void* x, const void* y,
                       void* r, void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  if (n == 0)
    return BD_SUCCESS;
  BD_REQUIRE(x && y && r, "bd_axpy: null pointer argument");
  const int grid = vec_grid(n);
  if (scalar_bits == 64)
    axpy_kernel<double><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, alpha, static_cast<const double*>(x), static_cast<const double*>(y),
        static_cast<double*>(r));
  else
    axpy_kernel<float><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, static_cast<float>(alpha), static_cast<const float*>(x), static_cast<const float*>(y),
        static_cast<float*>(r));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

extern "C" int bd_set_value(int scalar_bits, int64_t n, double value, void* x, void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  if (n == 0)
    return BD_SUCCESS;
  BD_REQUIRE(x, "bd_set_value: null pointer argument");
  if (value == 0.0)
  {
    BD_CUDA_CHECK(cudaMemsetAsync(x, 0, static_cast<size_t>(n) * (scalar_bits / 8), as_stream(stream)));
    return BD_SUCCESS;
  }
  const int grid = vec_grid(n);
  if (scalar_bits == 64)
    fill_kernel<double><<<grid, kVecBlock, 0, as_stream(stream)>>>(n, value, static_cast<double*>(x));
  else
    fill_kernel<float><<<grid, kVecBlock, 0, as_stream(stream)>>>(n, static_cast<float>(value),
                                                                 static_cast<float*>(x));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

extern "C" size_t bd_dot_scratch_bytes(void) { return sizeof(double) * kMaxPartials; }

extern "C" int bd_dot(int scalar_bits, int64_t n, const void* a, const void* b, double* result_dev,
                      void* scratch_dev, void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  BD_REQUIRE(result_dev && scratch_dev, "bd_dot: null result or scratch pointer");
  BD_REQUIRE(n == 0 || (a && b), "bd_dot: null vector pointer");
  return scalar_bits == 64 ? dot_impl<double>(n, a, b, result_dev, static_cast<double*>(scratch_dev),
                                              as_stream(stream))
                           : dot_impl<float>(n, a, b, result_dev, static_cast<double*>(scratch_dev),
                                             as_stream(stream));
}

extern "C" int bd_dot_host(int scalar_bits, int64_t n, const void* a, const void* b,
                           double* result_host)
{
  BD_REQUIRE(result_host, "bd_dot_host: null result pointer");
  double* buf = nullptr;
  BD_CUDA_CHECK(cudaMalloc(&buf, bd_dot_scratch_bytes() + sizeof(double)));
  int rc = bd_dot(scalar_bits, n, a, b, buf + kMaxPartials, buf, nullptr);
  if (rc == BD_SUCCESS)
  {
    cudaError_t e = cudaMemcpy(result_host, buf + kMaxPartials, sizeof(double), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess)
      rc = fail(BD_ERR_CUDA, "bd_dot_host: %s", cudaGetErrorString(e));
  }
  cudaFree(buf);
  return rc;
}

extern "C" int bd_pack(int scalar_bits, int64_t n, const int32_t* idx, const void* in, void* out,
                       void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  if (n == 0)
    return BD_SUCCESS;
  BD_REQUIRE(idx && in && out, "bd_pack: null pointer argument");
  const int grid = vec_grid(n, 1);
  if (scalar_bits == 64)
    pack_kernel<double><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, idx, static_cast<const double*>(in), static_cast<double*>(out));
  else
    pack_kernel<float><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, idx, static_cast<const float*>(in), static_cast<float*>(out));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

extern "C" int bd_unpack(int scalar_bits, int64_t n, const int32_t* idx, const void* in, void* out,
                         void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  if (n == 0)
    return BD_SUCCESS;
  BD_REQUIRE(idx && in && out, "bd_unpack: null pointer argument");
  const int grid = vec_grid(n, 1);
  if (scalar_bits == 64)
    unpack_kernel<double><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, idx, static_cast<const double*>(in), static_cast<double*>(out));
  else
    unpack_kernel<float><<<grid, kVecBlock, 0, as_stream(stream)>>>(
        n, idx, static_cast<const float*>(in), static_cast<float*>(out));
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// ---------------------------------------------------------------------------------------------
// bd_dot_static: the same two-stage dot product, for callers that must not allocate.
//
// The per-CTA partial sums live in a file-scope device array instead of caller-provided scratch,
// so the call needs no cudaMalloc (which would also synchronise the whole device in the middle
// of an asynchronous solver loop) and works when the application has already committed all of the
// GPU's memory.  Every GPU that loads the library holds these 16 KB until the process exits.
// ---------------------------------------------------------------------------------------------
namespace
{
static __device__ __align__(256) double g_static_dot_partials[bd::kMaxPartials];

// In host code `g_static_dot_partials` is only a stand-in for the symbol, not a device address;
// the address a kernel can use is looked up once per device, with that device current.
std::mutex g_static_mutex;
double* g_static_partials_of_device[64] = {nullptr};

int static_partials(double** out)
{
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64)
    return bd::fail(BD_ERR_INVALID_ARGUMENT, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_partials_of_device[dev])
  {
    double* ptr = nullptr;
    BD_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_static_dot_partials));
    g_static_partials_of_device[dev] = ptr;
  }
  *out = g_static_partials_of_device[dev];
  return BD_SUCCESS;
}
} // namespace

extern "C" int bd_dot_static(int scalar_bits, int64_t n, const void* a, const void* b,
                             double* result_dev, void* stream)
{
  BD_BITS_CHECK(scalar_bits);
  BD_REQUIRE(n >= 0, "negative length");
  if (n > BD_DOT_STATIC_MAX_N)
    return fail(BD_ERR_INVALID_ARGUMENT,
                "bd_dot_static: %lld entries exceed the supported maximum of %lld (use bd_dot with "
                "caller-provided scratch for longer vectors)",
                (long long)n, (long long)BD_DOT_STATIC_MAX_N);
  BD_REQUIRE(result_dev, "bd_dot_static: null result pointer");
  BD_REQUIRE(n == 0 || (a && b), "bd_dot_static: null vector pointer");
  double* partials = nullptr;
  BD_CHECK(static_partials(&partials));
  // Nothing in the array has to start at zero: the first kernel overwrites partials[0, grid) and
  // the second reads exactly those entries, so whatever the previous call left behind is ignored.
  return scalar_bits == 64 ? dot_impl<double>(n, a, b, result_dev, partials, as_stream(stream))
                           : dot_impl<float>(n, a, b, result_dev, partials, as_stream(stream));
}
