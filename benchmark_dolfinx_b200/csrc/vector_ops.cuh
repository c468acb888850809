// Device helpers shared by the vector kernels and the fused CG kernels.
#pragma once

#include "common.cuh"

namespace bd
{
constexpr int kVecBlock = 256;       // threads per CTA of the streaming vector kernels
constexpr int kMaxPartials = 2048;   // upper bound on CTAs of a reduction (148 SMs x 8 = 1184)

// CTAs for a streaming pass over n entries: a multiple of the SM count, capped.
inline int vec_grid(std::int64_t n, int per_thread = 4)
{
  const std::int64_t want = (n + static_cast<std::int64_t>(kVecBlock) * per_thread - 1)
                            / (static_cast<std::int64_t>(kVecBlock) * per_thread);
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
  std::int64_t g = want < cap ? want : cap;
  if (g < 1)
    g = 1;
  if (g > kMaxPartials)
    g = kMaxPartials;
  return static_cast<int>(g);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Sum over the CTA (blockDim.x == kVecBlock); result valid in thread 0.  Fixed order.
__device__ __forceinline__ double block_sum(double v)
{
  __shared__ double warp_part[kVecBlock / 32];
  v = warp_sum(v);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads(); // protect warp_part against a previous use
  if (lane == 0)
    warp_part[w] = v;
  __syncthreads();
  double s = 0;
  if (w == 0)
  {
    s = lane < kVecBlock / 32 ? warp_part[lane] : 0.0;
    s = warp_sum(s);
  }
  return s;
}

// Second stage of every reduction: sums `nparts` per-CTA partials in a fixed order.
int launch_reduce_partials(int nparts, const double* partials, double* result, cudaStream_t stream);
} // namespace bd
