// Dispatch + C entry point for the stiffness kernel.
#include "stiffness_launch.cuh"
#include "stiffness_mma8.cuh"

#include <cstdlib>
#include <cstring>

namespace bd
{
int stiffness_variant()
{
  static const int v = [] {
    const char* e = std::getenv("BD_STIFFNESS_VARIANT");
    if (!e)
      return 0;
    if (!std::strcmp(e, "slab"))
      return 1;
    if (!std::strcmp(e, "tma"))
      return 2;
    if (!std::strcmp(e, "pencil"))
      return 3;
    return 0;
  }();
  return v;
}

__global__ void __launch_bounds__(256)
    mask_dofmap_kernel(std::int64_t n, const std::int32_t* __restrict__ dofmap,
                       const std::int8_t* __restrict__ bc, std::int32_t* __restrict__ out)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  for (std::int64_t i = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride)
  {
    const std::int32_t d = dofmap[i];
    out[i] = bc[d] ? (d | kBcBit) : d;
  }
}

int build_masked_dofmap(std::int64_t n, const std::int32_t* dofmap, const std::int8_t* bc,
                        std::int32_t* out, cudaStream_t stream)
{
  if (n == 0)
    return BD_SUCCESS;
  const std::int64_t want = (n + 1023) / 1024;
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
  mask_dofmap_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, stream>>>(n, dofmap, bc, out);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

// flags[p] = 1 when the cells at positions p and p + 1 of the list are in the same batch of `cw` cells and the whole
// face k = nd - 1 of the first coincides, dof by dof (BC bit included), with the face k = 0 of the second.
__global__ void __launch_bounds__(256)
    merge_flags_kernel(std::int64_t ncells, int nd, int cw, const std::int32_t* __restrict__ mdofmap,
                       const std::int32_t* __restrict__ cells, std::int8_t* __restrict__ flags)
{
  const std::int64_t stride = static_cast<std::int64_t>(gridDim.x) * blockDim.x;
  const int nd2 = nd * nd, nd3 = nd2 * nd;
  for (std::int64_t p = static_cast<std::int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; p < ncells; p += stride)
  {
    bool same = p + 1 < ncells && (p % cw) != cw - 1;
    if (same)
    {
      const std::int32_t* ra = mdofmap + static_cast<std::int64_t>(cells[p]) * nd3;
      const std::int32_t* rb = mdofmap + static_cast<std::int64_t>(cells[p + 1]) * nd3;
      for (int ij = 0; ij < nd2 && same; ++ij)
        same = ra[ij * nd + nd - 1] == rb[ij * nd];
    }
    flags[p] = same ? 1 : 0;
  }
}

int build_merge_flags(std::int64_t ncells, int nd, int cw, const std::int32_t* mdofmap, const std::int32_t* cells,
                      std::int8_t* flags, cudaStream_t stream)
{
  if (ncells == 0)
    return BD_SUCCESS;
  const std::int64_t want = (ncells + 255) / 256;
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
  merge_flags_kernel<<<static_cast<int>(want < cap ? want : cap), 256, 0, stream>>>(ncells, nd, cw, mdofmap, cells,
                                                                                   flags);
  BD_CUDA_CHECK(cudaGetLastError());
  return BD_SUCCESS;
}

int preferred_g_layout(int scalar_bits, int degree, int nq)
{
  // pencil kernel (reference layout): see pencil_instantiated() in stiffness_launch.cuh
  const bool pencil = nq == degree + 2 && (nq == 9 || (scalar_bits == 32 && (nq == 7 || nq == 8)));
  if (pencil && stiffness_variant() != 1 && stiffness_variant() != 2)
    return 0;
  if ((nq == 8 || nq == 7) && stiffness_variant() != 1)
    return 2;
  return 1;
}

std::int64_t g_layout_size(int layout, int nq, std::int64_t ncells)
{
  if (layout == 2)
    return ncells * mma8_cell_scalars(nq);
  if (layout == 1)
    return slab_g_size(nq, ncells);
  return ncells * 6 * nq * nq * nq;
}

int stiffness_dispatch(int scalar_bits, int degree, int nq, bool identity, const StiffnessArgs& a)
{
  if (a.geom_mode != 0)
  {
    // geometry on the fly: slab kernel for nq <= 6, DMMA kernel for nq = 7, 8
    if (scalar_bits != 32 && scalar_bits != 64)
      return fail(BD_ERR_INVALID_ARGUMENT, "Invalid float size. Must be 32 or 64.");
    if (a.geom_mode == 1 && affine_col_used(scalar_bits, degree, identity))
      return launch_stiffness_affine_col_any(scalar_bits, degree, a);
    if (!otf_supported(degree, nq) || identity != (nq == degree + 1))
      return fail(BD_ERR_UNSUPPORTED, "on-the-fly geometry is not available for degree %d, nq %d", degree, nq);
    if (nq >= 7)
      return launch_stiffness_mma8_otf(scalar_bits, degree, identity, a);
    if (identity)
      return scalar_bits == 64 ? launch_stiffness_otf_f64_q0(degree, a) : launch_stiffness_otf_f32_q0(degree, a);
    return scalar_bits == 64 ? launch_stiffness_otf_f64_q1(degree, a) : launch_stiffness_otf_f32_q1(degree, a);
  }
  if (a.g_layout == 2)
  {
    if (nq != 8 && nq != 7)
      return fail(BD_ERR_INVALID_ARGUMENT, "the DMMA layout needs nq = 7 or 8");
    return launch_stiffness_mma8(scalar_bits, degree, identity, a);
  }
  if (scalar_bits != 32 && scalar_bits != 64)
    return fail(BD_ERR_INVALID_ARGUMENT, "Invalid float size. Must be 32 or 64.");
  if (degree < 1 || degree > 7)
    return fail(BD_ERR_UNSUPPORTED, "Unsupported degree [operator]");
  if (nq == degree + 1)
  {
    if (!identity)
      return fail(BD_ERR_UNSUPPORTED, "Expecting identity matrix for qmode=0");
    return scalar_bits == 64 ? launch_stiffness_f64_q0(degree, a) : launch_stiffness_f32_q0(degree, a);
  }
  if (nq == degree + 2)
  {
    if (identity)
      return fail(BD_ERR_UNSUPPORTED, "identity interpolation with nq = degree + 2");
    return scalar_bits == 64 ? launch_stiffness_f64_q1(degree, a) : launch_stiffness_f32_q1(degree, a);
  }
  return fail(BD_ERR_UNSUPPORTED, "Unsupported nq");
}
} // namespace bd

extern "C" int bd_stiffness_apply(int scalar_bits, int degree, int nq, int is_identity,
                                  const void* in, const void* cell_constants, void* out,
                                  const void* G, const void* phi0_host, const void* dphi1_host,
                                  const int32_t* cell_dofmap, const int32_t* cells, int64_t ncells,
                                  const int8_t* bc_marker, void* stream)
{
  BD_REQUIRE(ncells >= 0, "negative cell count");
  if (ncells == 0)
    return BD_SUCCESS;
  BD_REQUIRE(in && out && G && cell_constants && phi0_host && dphi1_host && cell_dofmap && cells
                 && bc_marker,
             "bd_stiffness_apply: null pointer argument");
  bd::StiffnessArgs a{in,     out,       G,         cell_constants, cell_dofmap, cells,
                      ncells, bc_marker, phi0_host, dphi1_host,     bd::as_stream(stream)};
  return bd::stiffness_dispatch(scalar_bits, degree, nq, is_identity != 0, a);
}
