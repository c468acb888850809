// Two fp32 multiply-adds per instruction (sm_100 FFMA2: a scalar from a register times a pair of table
// entries from a uniform register, added to a register pair).  Used by the fp32 pencil kernels.
#pragma once

#include <cmath>
#include <cstdint>

#include <cuda_runtime.h>

#define BD_PHASE_FN __device__ __forceinline__

namespace bd
{
// acc + s * t, both halves
BD_PHASE_FN float2 pair_fma(float s, float2 t, float2 acc)
{
#ifdef __CUDA_ARCH__
  return __ffma2_rn(make_float2(s, s), t, acc); // SASS: FFMA2 R, R.F32, UR.F32x2, R.F32x2
#else
  return make_float2(fmaf(s, t.x, acc.x), fmaf(s, t.y, acc.y));
#endif
}
// acc + a * t, half by half
BD_PHASE_FN float2 pair_mul_fma(float2 a, float2 t, float2 acc)
{
#ifdef __CUDA_ARCH__
  return __ffma2_rn(a, t, acc);
#else
  return make_float2(fmaf(a.x, t.x, acc.x), fmaf(a.y, t.y, acc.y));
#endif
}
template <int N>
BD_PHASE_FN float pair_get(const float2 (&v)[N], int i)
{
  return (i & 1) ? v[i >> 1].y : v[i >> 1].x;
}

} // namespace bd
