// Live measurement of the fp64 arithmetic peaks of the current device, for roofline denominators:
// dependent-chain-free DFMA and DMMA (mma.sync m8n8k4 f64) loops, 8 warps per CTA, 4 CTAs per SM.
// (The same loops as scripts/microbench/fp64_rates.cu.)  No counterpart in the reference.
#include "common.cuh"

namespace bd
{
__global__ void __launch_bounds__(256) peak_dfma_kernel(double* out, int iters, double a, double b)
{
  double x[8];
  for (int i = 0; i < 8; ++i)
    x[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      x[i] = x[i] * a + b;
  }
  double s = 0;
  for (int i = 0; i < 8; ++i)
    s += x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) peak_dmma_kernel(double* out, int iters, double a, double b)
{
  double c[4][2];
  for (int i = 0; i < 4; ++i)
    c[i][0] = c[i][1] = threadIdx.x;
  for (int it = 0; it < iters; ++it)
  {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1])
                   : "d"(a), "d"(b));
  }
  double s = 0;
  for (int i = 0; i < 4; ++i)
    s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
} // namespace bd

extern "C" int bd_fp64_peak(double* dfma_tflops, double* dmma_tflops)
{
  BD_REQUIRE(dfma_tflops && dmma_tflops, "bd_fp64_peak: null argument");
  const int sms = bd::sm_count();
  const int grid = sms * 4, block = 256, iters = 20000;
  double* out = nullptr;
  BD_CUDA_CHECK(cudaMalloc(&out, sizeof(double) * grid * block));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best[2] = {0, 0};
  for (int which = 0; which < 2; ++which)
    for (int rep = 0; rep < 3; ++rep)
    {
      cudaEventRecord(e0);
      if (which == 0)
        bd::peak_dfma_kernel<<<grid, block>>>(out, iters, 1.0000001, 1e-9);
      else
        bd::peak_dmma_kernel<<<grid, block>>>(out, iters, 1.0000001, 1e-9);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms = 0;
      cudaEventElapsedTime(&ms, e0, e1);
      // FMAs: DFMA 8 per thread and iteration; DMMA 4 instructions x 8*8*4 = 256 FMAs per warp and iteration
      const double fmas = which == 0 ? static_cast<double>(grid) * block * iters * 8
                                     : static_cast<double>(grid) * (block / 32) * iters * 4 * 256;
      const double tf = ms > 0 ? 2 * fmas / ms / 1e9 : 0;
      if (rep > 0 && tf > best[which])
        best[which] = tf;
    }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  const cudaError_t e = cudaGetLastError();
  cudaFree(out);
  if (e != cudaSuccess)
    return bd::fail(BD_ERR_CUDA, "bd_fp64_peak: %s", cudaGetErrorString(e));
  *dfma_tflops = best[0];
  *dmma_tflops = best[1];
  return BD_SUCCESS;
}
