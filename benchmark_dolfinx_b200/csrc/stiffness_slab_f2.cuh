// fp32 phases of the slab kernel on FFMA2 (sm_100: two fp32 FMAs per instruction and issue slot).
//
// The fp32 slab kernel is instruction-issue bound (ncu: issue slots 70 % busy, DRAM 65 %, FMA pipe
// 39 %, profiles/r1_final_ncu_summary_f32_q3.txt), and more than half of its instructions are FFMA.
// These variants of slab_forward_x / slab_z_points / slab_backward_x (stiffness_slab.cuh) compute
// exactly the same sums in the same order, two outputs per instruction:
//
//   forward, z contraction   (pz, dz)[j][q2, q2+1]  += ul[j][k]        * (Phi|DPhi)[q2, q2+1][k]     scalar x table pair
//   forward, y contraction   (ax, ay, az)[q1][q2, q2+1] += (Phi|DPhi)[q1][j] * (pz|dz)[j][q2, q2+1]  duplicated entry x data pair
//   middle,  x contraction   (gx, gy, gz)[q0, q0+1]  += (ax|ay|az)[i]   * (DPhi|Phi)[q0, q0+1][i]     scalar x table pair
//   middle,  x transposed    (bx, by, bz)[i, i+1]    += (fx|fy|fz)      * (DPhi|Phi)[q0][i, i+1]      scalar x table pair
//   backward, y transposed   (e1, e2)[j, j+1]        += (bx|by|bz)      * (Phi|DPhi)[q1][j, j+1]      scalar x table pair
//   backward, z transposed   out[j][k, k+1]          += (e1|e2)[j]      * (Phi|DPhi)[q2][k, k+1]      scalar x table pair
//
// The G stage stays scalar.  Odd extents are padded with zero table entries.  Selected with
// BD_SLAB_F2=1 (fp32, qmode 1, nq <= 5); tests/test_host_phases.py runs these functions on the host
// against the scalar ones.
#pragma once

#include <cmath>
#include <cstdint>

#include <cuda_runtime.h>

#ifdef BD_HOST_TEST
#define BD_PHASE_FN __host__ __device__ inline
#else
#define BD_PHASE_FN __device__ __forceinline__
#endif

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

// Tables of the FFMA2 phases, derived on the host from SlabTables (Phi = phi, DPhi = dphi).
template <int ND, int NQ>
struct SlabPairTables
{
  static constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
  float2 qphi[ND * NQP];  // [k][qp] = (Phi[2qp][k],  Phi[2qp+1][k])      pairs of quadrature points
  float2 qdphi[ND * NQP]; // [k][qp] = (DPhi[2qp][k], DPhi[2qp+1][k])
  float2 dphi2[NQ * NDP]; // [q][kp] = (DPhi[q][2kp], DPhi[q][2kp+1])     pairs of dofs
  float2 phi2[NQ * NDP];  // [q][kp] = (Phi[q][2kp],  Phi[q][2kp+1])
  float2 dup_phi[NQ * ND];  // [q][k] = (Phi[q][k],  Phi[q][k])           duplicated entries
  float2 dup_dphi[NQ * ND]; // [q][k] = (DPhi[q][k], DPhi[q][k])
};

template <int ND, int NQ>
inline void fill_slab_pair_tables(SlabPairTables<ND, NQ>& pt, const float* phi, const float* dphi)
{
  constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
  auto at = [](const float* t, int q, int k) { return (q < NQ && k < ND) ? t[q * ND + k] : 0.f; };
  for (int k = 0; k < ND; ++k)
    for (int qp = 0; qp < NQP; ++qp)
    {
      pt.qphi[k * NQP + qp] = make_float2(at(phi, 2 * qp, k), at(phi, 2 * qp + 1, k));
      pt.qdphi[k * NQP + qp] = make_float2(at(dphi, 2 * qp, k), at(dphi, 2 * qp + 1, k));
    }
  for (int q = 0; q < NQ; ++q)
  {
    for (int kp = 0; kp < NDP; ++kp)
    {
      pt.phi2[q * NDP + kp] = make_float2(at(phi, q, 2 * kp), at(phi, q, 2 * kp + 1));
      pt.dphi2[q * NDP + kp] = make_float2(at(dphi, q, 2 * kp), at(dphi, q, 2 * kp + 1));
    }
    for (int k = 0; k < ND; ++k)
    {
      pt.dup_phi[q * ND + k] = make_float2(phi[q * ND + k], phi[q * ND + k]);
      pt.dup_dphi[q * ND + k] = make_float2(dphi[q * ND + k], dphi[q * ND + k]);
    }
  }
}

// Forward X phase (same contract as slab_forward_x, non-identity interpolation).
template <int ND, int NQ, int RL, int CUBE>
BD_PHASE_FN void slab_forward_x_f2(const SlabPairTables<ND, NQ>& pt, const float (&ul)[ND * ND],
                                   float* __restrict__ dst0)
{
  constexpr int NQP = (NQ + 1) / 2;
#pragma unroll
  for (int qp = 0; qp < NQP; ++qp)
  {
    // z contraction for the points q2 = 2qp, 2qp+1
    float2 pz[ND], dz[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      float2 s = make_float2(0.f, 0.f), d = make_float2(0.f, 0.f);
#pragma unroll
      for (int k = 0; k < ND; ++k)
      {
        s = pair_fma(ul[j * ND + k], pt.qphi[k * NQP + qp], s);
        d = pair_fma(ul[j * ND + k], pt.qdphi[k * NQP + qp], d);
      }
      pz[j] = s;
      dz[j] = d;
    }
    // y contraction
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      float2 ax = make_float2(0.f, 0.f), ay = ax, az = ax;
#pragma unroll
      for (int j = 0; j < ND; ++j)
      {
        ax = pair_mul_fma(pt.dup_phi[q1 * ND + j], pz[j], ax);
        ay = pair_mul_fma(pt.dup_dphi[q1 * ND + j], pz[j], ay);
        az = pair_mul_fma(pt.dup_phi[q1 * ND + j], dz[j], az);
      }
      float* dst = dst0 + q1 * (ND * RL) + 2 * qp;
      dst[0] = ax.x;
      dst[CUBE] = ay.x;
      dst[2 * CUBE] = az.x;
      if (2 * qp + 1 < NQ)
      {
        dst[1] = ax.y;
        dst[CUBE + 1] = ay.y;
        dst[2 * CUBE + 1] = az.y;
      }
    }
  }
}

// Z phase for one q1 and all nq points q0 (same contract as slab_z_points<..., Q0N = NQ, SCALE = false>
// with Q0B = 0): b* are overwritten, not accumulated.
template <int ND, int NQ, int ROW>
BD_PHASE_FN void slab_z_points_f2(const SlabPairTables<ND, NQ>& pt, const float (&ax)[ND],
                                  const float (&ay)[ND], const float (&az)[ND], float (&bx)[ND],
                                  float (&by)[ND], float (&bz)[ND], const float* __restrict__ g)
{
  constexpr int NQP = (NQ + 1) / 2, NDP = (ND + 1) / 2;
  float2 gx[NQP], gy[NQP], gz[NQP];
#pragma unroll
  for (int qp = 0; qp < NQP; ++qp)
    gx[qp] = make_float2(0.f, 0.f), gy[qp] = make_float2(0.f, 0.f), gz[qp] = make_float2(0.f, 0.f);
#pragma unroll
  for (int i = 0; i < ND; ++i)
#pragma unroll
    for (int qp = 0; qp < NQP; ++qp)
    {
      gx[qp] = pair_fma(ax[i], pt.qdphi[i * NQP + qp], gx[qp]);
      gy[qp] = pair_fma(ay[i], pt.qphi[i * NQP + qp], gy[qp]);
      gz[qp] = pair_fma(az[i], pt.qphi[i * NQP + qp], gz[qp]);
    }
  float2 ox[NDP], oy[NDP], oz[NDP];
#pragma unroll
  for (int ip = 0; ip < NDP; ++ip)
    ox[ip] = make_float2(0.f, 0.f), oy[ip] = make_float2(0.f, 0.f), oz[ip] = make_float2(0.f, 0.f);
#pragma unroll
  for (int q0 = 0; q0 < NQ; ++q0)
  {
    const float* gp = g + q0 * (6 * ROW);
    const float G0 = gp[0 * ROW], G1 = gp[1 * ROW], G2 = gp[2 * ROW], G3 = gp[3 * ROW], G4 = gp[4 * ROW],
                G5 = gp[5 * ROW];
    const float vx = pair_get(gx, q0), vy = pair_get(gy, q0), vz = pair_get(gz, q0);
    // reference src/laplacian_gpu.hpp:269-273 (kappa is already in the input)
    const float fx = G0 * vx + G1 * vy + G2 * vz;
    const float fy = G1 * vx + G3 * vy + G4 * vz;
    const float fz = G2 * vx + G4 * vy + G5 * vz;
#pragma unroll
    for (int ip = 0; ip < NDP; ++ip)
    {
      ox[ip] = pair_fma(fx, pt.dphi2[q0 * NDP + ip], ox[ip]);
      oy[ip] = pair_fma(fy, pt.phi2[q0 * NDP + ip], oy[ip]);
      oz[ip] = pair_fma(fz, pt.phi2[q0 * NDP + ip], oz[ip]);
    }
  }
#pragma unroll
  for (int i = 0; i < ND; ++i)
  {
    bx[i] = pair_get(ox, i);
    by[i] = pair_get(oy, i);
    bz[i] = pair_get(oz, i);
  }
}

// Backward X phase (same contract as slab_backward_x, non-identity interpolation).
template <int ND, int NQ, int RL, int CUBE>
BD_PHASE_FN void slab_backward_x_f2(const SlabPairTables<ND, NQ>& pt, const float* __restrict__ src0,
                                    float (&out)[ND * ND])
{
  constexpr int NDP = (ND + 1) / 2;
  float2 o[ND * NDP];
#pragma unroll
  for (int t = 0; t < ND * NDP; ++t)
    o[t] = make_float2(0.f, 0.f);
#pragma unroll
  for (int q2 = 0; q2 < NQ; ++q2)
  {
    float2 e1[NDP], e2[NDP];
#pragma unroll
    for (int jp = 0; jp < NDP; ++jp)
      e1[jp] = make_float2(0.f, 0.f), e2[jp] = make_float2(0.f, 0.f);
#pragma unroll
    for (int q1 = 0; q1 < NQ; ++q1)
    {
      const float* src = src0 + q1 * (ND * RL) + q2;
      const float bx = src[0], by = src[CUBE], bz = src[2 * CUBE];
#pragma unroll
      for (int jp = 0; jp < NDP; ++jp)
      {
        e1[jp] = pair_fma(bx, pt.phi2[q1 * NDP + jp], e1[jp]);
        e1[jp] = pair_fma(by, pt.dphi2[q1 * NDP + jp], e1[jp]);
        e2[jp] = pair_fma(bz, pt.phi2[q1 * NDP + jp], e2[jp]);
      }
    }
#pragma unroll
    for (int j = 0; j < ND; ++j)
    {
      const float a = pair_get(e1, j), b = pair_get(e2, j);
#pragma unroll
      for (int kp = 0; kp < NDP; ++kp)
      {
        o[j * NDP + kp] = pair_fma(a, pt.phi2[q2 * NDP + kp], o[j * NDP + kp]);
        o[j * NDP + kp] = pair_fma(b, pt.dphi2[q2 * NDP + kp], o[j * NDP + kp]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < ND; ++j)
#pragma unroll
    for (int k = 0; k < ND; ++k)
      out[j * ND + k] = (k & 1) ? o[j * NDP + (k >> 1)].y : o[j * NDP + (k >> 1)].x;
}
} // namespace bd
