// Host-side check of the FFMA2 phase functions (csrc/stiffness_slab_f2.cuh) against the scalar phase
// functions of the slab kernel (csrc/stiffness_slab.cuh): same inputs, same outputs.  Compiled with
// -DBD_HOST_TEST, which makes the per-lane phase functions __host__ __device__; nothing runs on a GPU.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "stiffness_slab.cuh"

using namespace bd;

static std::mt19937 rng(12345);
static float rnd() { return std::uniform_real_distribution<float>(-1.f, 1.f)(rng); }

static double rel_err(const std::vector<float>& a, const std::vector<float>& b)
{
  double num = 0, den = 0;
  for (size_t i = 0; i < a.size(); ++i)
  {
    num += (double(a[i]) - b[i]) * (double(a[i]) - b[i]);
    den += double(b[i]) * b[i];
  }
  return std::sqrt(num / (den > 0 ? den : 1));
}

template <int ND, int NQ>
int check()
{
  constexpr int RL = 37, CUBE = NQ * ND * RL, ROW = 32;
  SlabTables<float, ND, NQ> tab;
  for (int i = 0; i < NQ * ND; ++i)
    tab.phi[i] = rnd(), tab.dphi[i] = rnd();
  SlabPairTables<ND, NQ> pt;
  fill_slab_pair_tables<ND, NQ>(pt, tab.phi, tab.dphi);
  int bad = 0;

  // forward X
  {
    float ul[ND * ND];
    for (float& v : ul)
      v = rnd();
    std::vector<float> a(3 * CUBE, 7.f), b(3 * CUBE, 7.f);
    slab_forward_x<float, ND, NQ, RL, CUBE, false>(tab, ul, a.data() + 3);
    slab_forward_x_f2<ND, NQ, RL, CUBE>(pt, ul, b.data() + 3);
    const double e = rel_err(b, a);
    std::printf("nd=%d nq=%d forward  rel err %.2e\n", ND, NQ, e);
    bad += !(e < 2e-6);
  }
  // Z points
  {
    float ax[ND], ay[ND], az[ND], bx[ND], by[ND], bz[ND], cx[ND], cy[ND], cz[ND];
    for (int i = 0; i < ND; ++i)
      ax[i] = rnd(), ay[i] = rnd(), az[i] = rnd(), bx[i] = by[i] = bz[i] = 0.f, cx[i] = cy[i] = cz[i] = 9.f;
    std::vector<float> g(NQ * 6 * ROW);
    for (float& v : g)
      v = rnd();
    slab_z_points<float, ND, NQ, ROW, false, NQ, false, false>(tab, ax, ay, az, bx, by, bz, g.data() + 5, 1.f, 0);
    slab_z_points_f2<ND, NQ, ROW>(pt, ax, ay, az, cx, cy, cz, g.data() + 5);
    std::vector<float> a, b;
    for (int i = 0; i < ND; ++i)
    {
      a.insert(a.end(), {bx[i], by[i], bz[i]});
      b.insert(b.end(), {cx[i], cy[i], cz[i]});
    }
    const double e = rel_err(b, a);
    std::printf("nd=%d nq=%d z points rel err %.2e\n", ND, NQ, e);
    bad += !(e < 2e-6);
  }
  // backward X
  {
    std::vector<float> cube(3 * CUBE);
    for (float& v : cube)
      v = rnd();
    float o1[ND * ND], o2[ND * ND];
    slab_backward_x<float, ND, NQ, RL, CUBE, false>(tab, cube.data() + 3, o1);
    slab_backward_x_f2<ND, NQ, RL, CUBE>(pt, cube.data() + 3, o2);
    const double e = rel_err(std::vector<float>(o2, o2 + ND * ND), std::vector<float>(o1, o1 + ND * ND));
    std::printf("nd=%d nq=%d backward rel err %.2e\n", ND, NQ, e);
    bad += !(e < 2e-6);
  }
  return bad;
}

int main()
{
  const int bad = check<2, 3>() + check<3, 4>() + check<4, 5>() + check<3, 5>();
  std::printf(bad ? "FAILED\n" : "OK\n");
  return bad ? 1 : 0;
}
