// Instantiations of the column-round dof-space kernel (stiffness_affine_col.cuh): degree 4..7, qmode 1, both scalar types.
#include "stiffness_launch.cuh"

namespace bd
{
int launch_stiffness_affine_col_any(int scalar_bits, int degree, const StiffnessArgs& a)
{
#define BD_COL_CASE(P)                                                                             \
  case P:                                                                                          \
    if (a.g_diagonal)                                                                              \
      return scalar_bits == 64                                                                     \
                 ? launch_stiffness_affine_col<double, P, P + 2, false, 1, affine_col_compact(64, P), true>(a) \
                 : launch_stiffness_affine_col<float, P, P + 2, false, 1, affine_col_compact(32, P), true>(a); \
    return scalar_bits == 64                                                                       \
               ? launch_stiffness_affine_col<double, P, P + 2, false, 1, affine_col_compact(64, P)>(a) \
               : launch_stiffness_affine_col<float, P, P + 2, false, 1, affine_col_compact(32, P)>(a);
  switch (degree)
  {
    BD_COL_CASE(4)
    BD_COL_CASE(5)
    BD_COL_CASE(6)
    BD_COL_CASE(7)
  default:
    return fail(BD_ERR_UNSUPPORTED, "the column-round dof-space kernel covers degree 4..7");
  }
#undef BD_COL_CASE
}
} // namespace bd
