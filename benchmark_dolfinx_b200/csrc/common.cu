#include "common.cuh"

#include <atomic>
#include <cstring>

namespace bd
{
static std::atomic<int> g_debug_max_grid{0};
int debug_max_grid() { return g_debug_max_grid.load(std::memory_order_relaxed); }

std::string& last_error_slot()
{
  static thread_local std::string slot;
  return slot;
}

int fail(int code, const char* fmt, ...)
{
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error_slot() = buf;
  return code;
}

int sm_count()
{
  static thread_local int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64)
    return kB200SMs;
  if (cached[dev] == 0)
  {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kB200SMs;
    cached[dev] = n;
  }
  return cached[dev];
}
} // namespace bd

extern "C" int bd_debug_set_max_grid(int max_ctas)
{
  BD_REQUIRE(max_ctas >= 0, "bd_debug_set_max_grid: negative grid cap");
  bd::g_debug_max_grid.store(max_ctas, std::memory_order_relaxed);
  return BD_SUCCESS;
}

extern "C" const char* bd_last_error(void) { return bd::last_error_slot().c_str(); }

extern "C" int bd_device_information(char* buf, size_t n)
{
  BD_REQUIRE(buf && n > 0, "bd_device_information: null buffer");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
  {
    cudaGetLastError();
    snprintf(buf, n, "benchdolfinx_b200: no CUDA device (there is no CPU fallback)\n");
    return bd::fail(BD_ERR_NO_DEVICE, "no usable CUDA device: %s",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  int dev = 0;
  BD_CUDA_CHECK(cudaGetDevice(&dev));
  cudaDeviceProp prop{};
  BD_CUDA_CHECK(cudaGetDeviceProperties(&prop, dev));
  int driver = 0, runtime = 0;
  cudaDriverGetVersion(&driver);
  cudaRuntimeGetVersion(&runtime);
  snprintf(buf, n,
           "Device Information:\n"
           "\tDevice: %s (device %d of %d)\n"
           "\tCompute capability: %d.%d, SMs: %d\n"
           "\tGlobal memory: %.1f GB, L2: %.1f MB, shared memory per SM: %zu KB\n"
           "\tCUDA driver %d, runtime %d\n",
           prop.name, dev, count, prop.major, prop.minor, prop.multiProcessorCount,
           prop.totalGlobalMem / 1e9, prop.l2CacheSize / 1e6, prop.sharedMemPerMultiprocessor / 1024,
           driver, runtime);
  return BD_SUCCESS;
}
