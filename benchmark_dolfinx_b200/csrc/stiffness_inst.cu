// Instantiation unit: compiled four times with -DBD_INST_T={double,float} -DBD_INST_QMODE={0,1}.
#include "stiffness_launch.cuh"

#ifndef BD_INST_T
#error "define BD_INST_T, BD_INST_SUFFIX and BD_INST_QMODE"
#endif

#define BD_CAT_(a, b, c) a##b##_q##c
#define BD_CAT(a, b, c) BD_CAT_(a, b, c)

#define BD_CAT_OTF_(a, b, c) a##b##_q##c
#define BD_CAT_OTF(a, b, c) BD_CAT_OTF_(a, b, c)

namespace bd
{
#ifdef BD_INST_OTF
// ---- on-the-fly geometry (slab kernel, nq <= 6)
#if BD_INST_QMODE == 0
#define BD_OTF_CASE(P)                                                                             \
  case P:                                                                                          \
    return launch_stiffness_slab_otf<BD_INST_T, P, P + 1, true>(a);
#else
#define BD_OTF_CASE(P)                                                                             \
  case P:                                                                                          \
    return launch_stiffness_slab_otf<BD_INST_T, P, P + 2, false>(a);
#endif
int BD_CAT_OTF(launch_stiffness_otf_, BD_INST_SUFFIX, BD_INST_QMODE)(int degree, const StiffnessArgs& a)
{
  switch (degree)
  {
    BD_OTF_CASE(1)
    BD_OTF_CASE(2)
    BD_OTF_CASE(3)
    BD_OTF_CASE(4)
#if BD_INST_QMODE == 0
    BD_OTF_CASE(5)
#endif
  default:
    return fail(BD_ERR_UNSUPPORTED, "on-the-fly geometry: the slab kernel covers nq <= 6");
  }
}
#else
#if BD_INST_QMODE == 0
#define BD_CASE(P)                                                                                 \
  case P:                                                                                          \
    return !a.slab ? launch_stiffness<BD_INST_T, P, P + 1, true>(a)                                \
           : a.g_layout == 0 ? launch_stiffness_pencil<BD_INST_T, P, P + 1, true>(a)               \
                             : launch_stiffness_slab<BD_INST_T, P, P + 1, true>(a);
#else
#define BD_CASE(P)                                                                                 \
  case P:                                                                                          \
    return !a.slab ? launch_stiffness<BD_INST_T, P, P + 2, false>(a)                               \
           : a.g_layout == 0 ? launch_stiffness_pencil<BD_INST_T, P, P + 2, false>(a)              \
                             : launch_stiffness_slab<BD_INST_T, P, P + 2, false>(a);
#endif

int BD_CAT(launch_stiffness_, BD_INST_SUFFIX, BD_INST_QMODE)(int degree, const StiffnessArgs& a)
{
  switch (degree)
  {
    BD_CASE(1)
    BD_CASE(2)
    BD_CASE(3)
    BD_CASE(4)
    BD_CASE(5)
    BD_CASE(6)
    BD_CASE(7)
  default:
    return fail(BD_ERR_UNSUPPORTED, "Unsupported degree [operator]");
  }
}
#endif // BD_INST_OTF
} // namespace bd
