/* Plain-C restatement of the reference's matrix-free Laplacian cell loop (oracle;
 * TEST INFRASTRUCTURE, never linked into the product).
 *
 * Used (a) as a second, independent checker next to oracle/laplacian.py and
 * (b) as the timed "port" CPU baseline of bench.py when oracle/_ref (the
 * reference's own CPU kernels) is unavailable, and for qmode 1, which the
 * reference CPU kernel does not support (src/laplacian_cpu.hpp:73-74).
 *
 * Follows, stage by stage:
 *   gather + BC masking        src/laplacian_cpu.hpp:81-94   (= laplacian_gpu.hpp:153-170)
 *   interpolation nd -> nq     src/laplacian_gpu.hpp:175-211 (skipped when identity)
 *   reference gradient         src/laplacian_cpu.hpp:115-123 (= laplacian_gpu.hpp:239-251)
 *   G, kappa                   src/laplacian_cpu.hpp:107-126 (= laplacian_gpu.hpp:256-273)
 *   transposed gradient        src/laplacian_cpu.hpp:130-140 (= laplacian_gpu.hpp:307-314)
 *   interpolation^T nq -> nd   src/laplacian_gpu.hpp:388-420
 *   scatter                    src/laplacian_cpu.hpp:141-143 (= laplacian_gpu.hpp:424-425)
 *   geometry                   src/geometry_cpu.hpp:43-110
 * G is indexed with 64-bit offsets (the reference CPU code uses int, :107).
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define MAXQ 9

#define DEFINE_PORT(T, SUFFIX)                                                                     \
  static void sweep_##SUFFIX(const T* M, int rows, int cols, int transpose, const T* in, T* out,   \
                             int n0, int n1, int n2, int axis)                                     \
  {                                                                                                \
    /* out[..a'..] = sum_a M[a'][a] in[..a..] along `axis` (M^T if transpose); in is n0*n1*n2 */  \
    int nin[3] = {n0, n1, n2};                                                                     \
    int nout[3] = {n0, n1, n2};                                                                    \
    int no = transpose ? cols : rows, ni = transpose ? rows : cols;                                \
    nout[axis] = no;                                                                               \
    for (int i = 0; i < nout[0]; ++i)                                                              \
      for (int j = 0; j < nout[1]; ++j)                                                            \
        for (int k = 0; k < nout[2]; ++k)                                                          \
        {                                                                                          \
          int o[3] = {i, j, k};                                                                    \
          T acc = 0;                                                                               \
          for (int a = 0; a < ni; ++a)                                                             \
          {                                                                                        \
            int s[3] = {i, j, k};                                                                  \
            s[axis] = a;                                                                           \
            T m = transpose ? M[a * cols + o[axis]] : M[o[axis] * cols + a];                       \
            acc += m * in[(s[0] * nin[1] + s[1]) * nin[2] + s[2]];                                 \
          }                                                                                        \
          out[(i * nout[1] + j) * nout[2] + k] = acc;                                              \
        }                                                                                          \
  }                                                                                                \
                                                                                                   \
  void port_stiffness_##SUFFIX(int nd, int nq, int identity, const T* u, const T* constants,       \
                               T* b, const T* G, const T* phi0, const T* dphi1,                    \
                               const int32_t* dofmap, const int32_t* entities, int n_entities,     \
                               const int8_t* bc)                                                   \
  {                                                                                                \
    const int nd3 = nd * nd * nd, nq3 = nq * nq * nq;                                              \
    T s0[MAXQ * MAXQ * MAXQ], s1[MAXQ * MAXQ * MAXQ], s2[MAXQ * MAXQ * MAXQ];                      \
    T fx[MAXQ * MAXQ * MAXQ], fy[MAXQ * MAXQ * MAXQ], fz[MAXQ * MAXQ * MAXQ];                      \
    for (int c = 0; c < n_entities; ++c)                                                           \
    {                                                                                              \
      const int32_t* dofs = dofmap + (size_t)entities[c] * nd3;                                    \
      for (int i = 0; i < nd3; ++i)                                                                \
      {                                                                                            \
        int32_t d = dofs[i];                                                                       \
        if (bc[d])                                                                                 \
        {                                                                                          \
          b[d] = u[d];                                                                             \
          s0[i] = 0;                                                                               \
        }                                                                                          \
        else                                                                                       \
          s0[i] = u[d];                                                                            \
      }                                                                                            \
      T* uq = s0;                                                                                  \
      if (!identity)                                                                               \
      {                                                                                            \
        sweep_##SUFFIX(phi0, nq, nd, 0, s0, s1, nd, nd, nd, 0);                                    \
        sweep_##SUFFIX(phi0, nq, nd, 0, s1, s2, nq, nd, nd, 1);                                    \
        sweep_##SUFFIX(phi0, nq, nd, 0, s2, s0, nq, nq, nd, 2);                                    \
      }                                                                                            \
      sweep_##SUFFIX(dphi1, nq, nq, 0, uq, fx, nq, nq, nq, 0);                                     \
      sweep_##SUFFIX(dphi1, nq, nq, 0, uq, fy, nq, nq, nq, 1);                                     \
      sweep_##SUFFIX(dphi1, nq, nq, 0, uq, fz, nq, nq, nq, 2);                                     \
      const T* Gc = G + (size_t)c * 6 * nq3;                                                       \
      const T coeff = constants[c];                                                                \
      for (int q = 0; q < nq3; ++q)                                                                \
      {                                                                                            \
        T vx = fx[q], vy = fy[q], vz = fz[q];                                                      \
        T G0 = Gc[q], G1 = Gc[nq3 + q], G2 = Gc[2 * nq3 + q], G3 = Gc[3 * nq3 + q],                \
          G4 = Gc[4 * nq3 + q], G5 = Gc[5 * nq3 + q];                                              \
        fx[q] = coeff * (G0 * vx + G1 * vy + G2 * vz);                                             \
        fy[q] = coeff * (G1 * vx + G3 * vy + G4 * vz);                                             \
        fz[q] = coeff * (G2 * vx + G4 * vy + G5 * vz);                                             \
      }                                                                                            \
      sweep_##SUFFIX(dphi1, nq, nq, 1, fx, s0, nq, nq, nq, 0);                                     \
      sweep_##SUFFIX(dphi1, nq, nq, 1, fy, s1, nq, nq, nq, 1);                                     \
      sweep_##SUFFIX(dphi1, nq, nq, 1, fz, s2, nq, nq, nq, 2);                                     \
      for (int q = 0; q < nq3; ++q)                                                                \
        s0[q] += s1[q] + s2[q];                                                                    \
      T* r = s0;                                                                                   \
      if (!identity)                                                                               \
      {                                                                                            \
        sweep_##SUFFIX(phi0, nq, nd, 1, s0, s1, nq, nq, nq, 0);                                    \
        sweep_##SUFFIX(phi0, nq, nd, 1, s1, s2, nd, nq, nq, 1);                                    \
        sweep_##SUFFIX(phi0, nq, nd, 1, s2, s0, nd, nd, nq, 2);                                    \
      }                                                                                            \
      for (int i = 0; i < nd3; ++i)                                                                \
      {                                                                                            \
        int32_t d = dofs[i];                                                                       \
        if (!bc[d])                                                                                \
          b[d] += r[i];                                                                            \
      }                                                                                            \
    }                                                                                              \
  }                                                                                                \
                                                                                                   \
  void port_geometry_##SUFFIX(int nq, const T* xgeom, T* G, const int32_t* gdofmap,                \
                              const T* dphi, const T* weights, const int32_t* entities,            \
                              int n_entities)                                                      \
  {                                                                                                \
    const int nq3 = nq * nq * nq;                                                                  \
    for (int c = 0; c < n_entities; ++c)                                                           \
    {                                                                                              \
      T x[8][3];                                                                                   \
      for (int v = 0; v < 8; ++v)                                                                  \
        for (int j = 0; j < 3; ++j)                                                                \
          x[v][j] = xgeom[3 * (size_t)gdofmap[(size_t)entities[c] * 8 + v] + j];                   \
      for (int q = 0; q < nq3; ++q)                                                                \
      {                                                                                            \
        T J[3][3];                                                                                 \
        for (int i = 0; i < 3; ++i)                                                                \
          for (int j = 0; j < 3; ++j)                                                              \
          {                                                                                        \
            J[i][j] = 0;                                                                           \
            for (int k = 0; k < 8; ++k)                                                            \
              J[i][j] += x[k][i] * dphi[((size_t)j * nq3 + q) * 8 + k];                            \
          }                                                                                        \
        T K[3][3] = {{J[1][1] * J[2][2] - J[1][2] * J[2][1], -J[0][1] * J[2][2] + J[0][2] * J[2][1],\
                      J[0][1] * J[1][2] - J[0][2] * J[1][1]},                                      \
                     {-J[1][0] * J[2][2] + J[1][2] * J[2][0], J[0][0] * J[2][2] - J[0][2] * J[2][0],\
                      -J[0][0] * J[1][2] + J[0][2] * J[1][0]},                                     \
                     {J[1][0] * J[2][1] - J[1][1] * J[2][0], -J[0][0] * J[2][1] + J[0][1] * J[2][0],\
                      J[0][0] * J[1][1] - J[0][1] * J[1][0]}};                                     \
        T detJ = J[0][0] * K[0][0] - J[0][1] * K[1][0] + J[0][2] * K[2][0];                        \
        T* g = G + (size_t)c * 6 * nq3 + q;                                                        \
        T s = weights[q] / detJ;                                                                   \
        g[0] = (K[0][0] * K[0][0] + K[0][1] * K[0][1] + K[0][2] * K[0][2]) * s;                    \
        g[nq3] = (K[1][0] * K[0][0] + K[1][1] * K[0][1] + K[1][2] * K[0][2]) * s;                  \
        g[2 * nq3] = (K[2][0] * K[0][0] + K[2][1] * K[0][1] + K[2][2] * K[0][2]) * s;              \
        g[3 * nq3] = (K[1][0] * K[1][0] + K[1][1] * K[1][1] + K[1][2] * K[1][2]) * s;              \
        g[4 * nq3] = (K[2][0] * K[1][0] + K[2][1] * K[1][1] + K[2][2] * K[1][2]) * s;              \
        g[5 * nq3] = (K[2][0] * K[2][0] + K[2][1] * K[2][1] + K[2][2] * K[2][2]) * s;              \
      }                                                                                            \
    }                                                                                              \
  }

DEFINE_PORT(double, f64)
DEFINE_PORT(float, f32)
