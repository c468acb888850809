// Microbenchmark: fp64 FMA (DFMA) vs fp64 tensor-core (DMMA m8n8k4) throughput on one GPU.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a fp64_rates.cu -o fp64_rates && ./fp64_rates
#include <cstdio>
#include <cuda_runtime.h>

__global__ void dfma_kernel(double* out, int iters, double a, double b)
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

__global__ void dmma_kernel(double* out, int iters, double a, double b)
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

int main()
{
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps = 1; warps <= 8; warps *= 2)
  {
    for (int which = 0; which < 2; ++which)
    {
      dim3 grid(sms * 4), block(32 * warps);
      for (int rep = 0; rep < 2; ++rep)
      {
        cudaEventRecord(e0);
        if (which == 0)
          dfma_kernel<<<grid, block>>>(out, iters, 1.0000001, 1e-9);
        else
          dmma_kernel<<<grid, block>>>(out, iters, 1.0000001, 1e-9);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      double fmas = which == 0 ? (double)grid.x * block.x * iters * 8 : (double)grid.x * warps * iters * 4 * 256;
      printf("%s warps/CTA=%d CTAs/SM=4: %.2f ms, %.2f TFLOP/s (fp64)\n", which == 0 ? "DFMA" : "DMMA", warps, ms,
             2 * fmas / ms / 1e9);
    }
  }
  printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
