// bd_laplacian: replaces the host class MatFreeLaplacianGPU<T> of the reference
// (src/laplacian.hpp:89-447): owns device copies of the mesh data, precomputes the geometry
// tensor at construction and applies the operator cell list by cell list.
#include "operator.cuh"
#include "stiffness_launch.cuh"

#include <cmath>
#include <cstring>

using namespace bd;

namespace
{
int check_device_available()
{
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
  {
    cudaGetLastError();
    return fail(BD_ERR_NO_DEVICE, "no usable CUDA device (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  return BD_SUCCESS;
}

// Recover the 1-D quadrature rule from the 3-D tables the caller supplied (weights[nq^3], dphi_geometry
// [3][nq^3][8]) and check that they have the structure the on-the-fly kernels assume: tensor-product weights,
// trilinear coordinate element with vertex v = 4a + 2b + c, point q = (q0 nq + q1) nq + q2 (the reference gets
// both from Basix, src/laplacian.hpp:144-169).  Leaves op->w1d / op->pts1d empty when the tables do not fit.
template <typename T>
void recover_1d_rule(bd_laplacian* op, const std::vector<char>& wbuf, const std::vector<char>& dbuf)
{
  const int nq = op->nq, nq2 = nq * nq, nq3 = nq2 * nq;
  const T* w3 = reinterpret_cast<const T*>(wbuf.data());
  const T* dphi = reinterpret_cast<const T*>(dbuf.data());
  auto D = [&](int d, int q, int v) { return static_cast<double>(dphi[(static_cast<size_t>(d) * nq3 + q) * 8 + v]); };
  std::vector<double> w(nq), x(nq);
  double wsum = 0;
  for (int q = 0; q < nq3; ++q)
    wsum += static_cast<double>(w3[q]);
  for (int a = 0; a < nq; ++a)
  {
    double s = 0;
    for (int r = 0; r < nq2; ++r)
      s += static_cast<double>(w3[a * nq2 + r]);
    w[a] = s / (wsum > 0 ? std::cbrt(wsum) * std::cbrt(wsum) : 1.0); // sum over the other two directions
    // d/dxi1 of vertex (a=1,b=1,c=*) summed over c = xi0 at the point; take a point with q0 = a
    x[a] = D(1, a * nq2, 6) + D(1, a * nq2, 7);
  }
  const double tol = sizeof(T) == 8 ? 1e-12 : 2e-5;
  bool ok = wsum > 0;
  for (int q0 = 0; q0 < nq && ok; ++q0)
    for (int q1 = 0; q1 < nq && ok; ++q1)
      for (int q2 = 0; q2 < nq && ok; ++q2)
      {
        const int q = (q0 * nq + q1) * nq + q2;
        if (std::fabs(static_cast<double>(w3[q]) - w[q0] * w[q1] * w[q2]) > tol * wsum)
          ok = false;
        const double xi[3] = {x[q0], x[q1], x[q2]};
        for (int v = 0; v < 8 && ok; ++v)
        {
          const int abc[3] = {(v >> 2) & 1, (v >> 1) & 1, v & 1};
          for (int d = 0; d < 3 && ok; ++d)
          {
            double e = 1;
            for (int k = 0; k < 3; ++k)
              e *= (k == d) ? (abc[k] ? 1.0 : -1.0) : (abc[k] ? xi[k] : 1.0 - xi[k]);
            if (std::fabs(D(d, q, v) - e) > tol)
              ok = false;
          }
        }
      }
  if (ok)
  {
    op->w1d = w;
    op->pts1d = x;
  }
}

template <typename T>
int count_nonaffine(bd_laplacian* op)
{
  const std::int64_t n = op->n_lcells + op->n_bcells;
  op->n_nonaffine = op->n_nonorthogonal = 0;
  if (n == 0)
    return BD_SUCCESS;
  unsigned long long* cnt = nullptr;
  BD_CUDA_CHECK(cudaMalloc(&cnt, 2 * sizeof(unsigned long long)));
  cudaMemset(cnt, 0, 2 * sizeof(unsigned long long));
  const std::int64_t want = (n + 255) / 256;
  const std::int64_t cap = static_cast<std::int64_t>(sm_count()) * 8;
  count_nonaffine_kernel<T><<<static_cast<int>(want < cap ? want : cap), 256>>>(
      reinterpret_cast<const T*>(op->xgeom.ptr), op->geometry_dofmap.ptr, op->cell_list.ptr, n, cnt);
  unsigned long long h[2] = {0, 0};
  const cudaError_t e = cudaMemcpy(h, cnt, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(cnt);
  if (e != cudaSuccess)
    return fail(BD_ERR_CUDA, "affine-cell count failed: %s", cudaGetErrorString(e));
  op->n_nonaffine = static_cast<std::int64_t>(h[0]);
  op->n_nonorthogonal = static_cast<std::int64_t>(h[1]);
  return BD_SUCCESS;
}

// G for the whole list in the layout of the stored-G kernels (no-op for the on-the-fly variants).
int precompute_geometry(bd_laplacian* op)
{
  const size_t ts = op->scalar_size();
  op->g_layout = preferred_g_layout(op->scalar_bits, op->degree, op->nq);
  op->g_bcells_offset = g_layout_size(op->g_layout, op->nq, op->n_lcells);
  const std::int64_t g_total = op->g_bcells_offset + g_layout_size(op->g_layout, op->nq, op->n_bcells);
  BD_CHECK(op->G.alloc(static_cast<size_t>(g_total > 0 ? g_total : 1) * ts));
  BD_CHECK(geometry_compute_layout(op->scalar_bits, op->nq, op->xgeom.ptr, op->geometry_dofmap.ptr,
                                   op->dphi_geometry.ptr, op->weights.ptr, op->cell_list.ptr, op->n_lcells,
                                   op->G.ptr, nullptr, op->g_layout));
  BD_CHECK(geometry_compute_layout(op->scalar_bits, op->nq, op->xgeom.ptr, op->geometry_dofmap.ptr,
                                   op->dphi_geometry.ptr, op->weights.ptr, op->cell_list.ptr + op->n_lcells,
                                   op->n_bcells, op->G.ptr + static_cast<size_t>(op->g_bcells_offset) * ts,
                                   nullptr, op->g_layout));
  return BD_SUCCESS;
}

template <typename T>
int fill_constants(bd_laplacian* op)
{
  // _cell_constants(num_cells, constant)  (src/laplacian.hpp:105); read by list position
  const std::int64_t n = op->n_lcells + op->n_bcells;
  BD_CHECK(op->cell_constants.alloc(static_cast<size_t>(n > 0 ? n : 1) * sizeof(T)));
  return bd_set_value(op->scalar_bits, n, op->kappa, op->cell_constants.ptr, nullptr);
}
} // namespace

extern "C" int bd_laplacian_create(const bd_laplacian_desc* d, bd_laplacian** out)
{
  BD_REQUIRE(d && out, "bd_laplacian_create: null argument");
  *out = nullptr;
  BD_CHECK(check_device_available());
  BD_REQUIRE(d->scalar_bits == 32 || d->scalar_bits == 64, "Invalid float size. Must be 32 or 64.");
  if (d->degree < 1 || d->degree > 7)
    return fail(BD_ERR_UNSUPPORTED, "Unsupported degree [operator]");
  if (d->nq != d->degree + 1 && d->nq != d->degree + 2)
    return fail(BD_ERR_UNSUPPORTED, "Unsupported nq");
  if (d->nq == d->degree + 1 && !d->is_identity)
    return fail(BD_ERR_UNSUPPORTED, "Expecting identity matrix for qmode=0");
  if (d->nq == d->degree + 2 && d->is_identity)
    return fail(BD_ERR_UNSUPPORTED, "identity interpolation with nq = degree + 2");
  BD_REQUIRE(d->ncells >= 0 && d->ndofs >= 0 && d->nnodes >= 0 && d->n_lcells >= 0 && d->n_bcells >= 0,
             "negative size in descriptor");
  BD_REQUIRE(d->n_lcells + d->n_bcells <= d->ncells, "lcells + bcells exceed ncells");
  BD_REQUIRE(d->ndofs <= 0x7fffffffLL && d->ncells <= 0x7fffffffLL,
             "local dof / cell counts must fit int32 (as in the reference dofmap)");
  const std::int64_t nd3 = static_cast<std::int64_t>(d->degree + 1) * (d->degree + 1) * (d->degree + 1);
  BD_REQUIRE(d->ncells * nd3 <= 0x7fffffffLL * 4ll, "dofmap too large");
  BD_REQUIRE(d->phi0 && d->dphi1 && d->weights && d->dphi_geometry, "null table pointer");

  auto op = new bd_laplacian;
  auto guard_fail = [&](int rc) {
    delete op;
    return rc;
  };
  op->scalar_bits = d->scalar_bits;
  op->degree = d->degree;
  op->nq = d->nq;
  op->is_identity = d->is_identity != 0;
  op->kappa = d->kappa;
  op->ncells = d->ncells;
  op->ndofs = d->ndofs;
  op->nnodes = d->nnodes;
  op->n_lcells = d->n_lcells;
  op->n_bcells = d->n_bcells;
  if (cudaGetDevice(&op->device) != cudaSuccess)
    return guard_fail(fail(BD_ERR_CUDA, "cudaGetDevice failed"));
  const bool borrow = (d->flags & BD_FLAG_BORROW_DEVICE_ARRAYS) != 0;
  const size_t ts = op->scalar_size();
  const int nq3 = d->nq * d->nq * d->nq;
  int rc;
#define TRY(expr)                                                                                  \
  if ((rc = (expr)) != BD_SUCCESS)                                                                 \
    return guard_fail(rc);
  TRY(op->cell_dofmap.assign(d->cell_dofmap, static_cast<size_t>(d->ncells * nd3), borrow));
  TRY(op->geometry_dofmap.assign(d->geometry_dofmap, static_cast<size_t>(d->ncells) * 8, borrow));
  TRY(op->bc_marker.assign(d->bc_marker, static_cast<size_t>(d->ndofs), borrow));
  TRY(op->xgeom.assign(static_cast<const char*>(d->xgeom), static_cast<size_t>(d->nnodes) * 3 * ts, borrow));
  TRY(op->weights.assign(static_cast<const char*>(d->weights), static_cast<size_t>(nq3) * ts, false));
  TRY(op->dphi_geometry.assign(static_cast<const char*>(d->dphi_geometry),
                               static_cast<size_t>(3) * nq3 * 8 * ts, false));
  // [lcells..., bcells...] in one list: G is indexed by position in it (src/laplacian.hpp:215-223)
  const std::int64_t nlist = d->n_lcells + d->n_bcells;
  TRY(op->cell_list.alloc(static_cast<size_t>(nlist > 0 ? nlist : 1)));
  if (d->n_lcells > 0)
  {
    if (!d->lcells)
      return guard_fail(fail(BD_ERR_INVALID_ARGUMENT, "null lcells"));
    if (cudaMemcpy(op->cell_list.ptr, d->lcells, d->n_lcells * sizeof(std::int32_t), cudaMemcpyDefault)
        != cudaSuccess)
      return guard_fail(fail(BD_ERR_CUDA, "copy of lcells failed: %s", cudaGetErrorString(cudaGetLastError())));
  }
  if (d->n_bcells > 0)
  {
    if (!d->bcells)
      return guard_fail(fail(BD_ERR_INVALID_ARGUMENT, "null bcells"));
    if (cudaMemcpy(op->cell_list.ptr + d->n_lcells, d->bcells, d->n_bcells * sizeof(std::int32_t),
                   cudaMemcpyDefault)
        != cudaSuccess)
      return guard_fail(fail(BD_ERR_CUDA, "copy of bcells failed: %s", cudaGetErrorString(cudaGetLastError())));
  }
  // tables: host copies (kernel launch parameter)
  auto copy_table = [&](std::vector<char>& dst, const void* src, size_t bytes) -> int {
    dst.resize(bytes);
    BD_CUDA_CHECK(cudaMemcpy(dst.data(), src, bytes, cudaMemcpyDefault));
    return BD_SUCCESS;
  };
  TRY(copy_table(op->phi0, d->phi0, static_cast<size_t>(d->nq) * (d->degree + 1) * ts));
  TRY(copy_table(op->dphi1, d->dphi1, static_cast<size_t>(d->nq) * d->nq * ts));
  TRY(d->scalar_bits == 64 ? fill_constants<double>(op) : fill_constants<float>(op));
  TRY(op->masked_dofmap.alloc(static_cast<size_t>(d->ncells * nd3 > 0 ? d->ncells * nd3 : 1)));
  TRY(build_masked_dofmap(d->ncells * nd3, op->cell_dofmap.ptr, op->bc_marker.ptr, op->masked_dofmap.ptr,
                          nullptr));
  // Geometry variant.  On the fly needs the 1-D rule behind the caller's 3-D tables and a kernel for this nq.
  {
    std::vector<char> wbuf, dbuf;
    TRY(copy_table(wbuf, d->weights, static_cast<size_t>(nq3) * ts));
    TRY(copy_table(dbuf, d->dphi_geometry, static_cast<size_t>(3) * nq3 * 8 * ts));
    if (d->scalar_bits == 64)
      recover_1d_rule<double>(op, wbuf, dbuf);
    else
      recover_1d_rule<float>(op, wbuf, dbuf);
  }
  const bool want_otf = (d->flags & BD_FLAG_GEOMETRY_ONTHEFLY) != 0;
  const bool want_pre = (d->flags & BD_FLAG_GEOMETRY_PRECOMPUTED) != 0;
  if (want_otf && want_pre)
    return guard_fail(fail(BD_ERR_INVALID_ARGUMENT, "both geometry flags set"));
  // a kernel for general cells exists for nq <= 8; on meshes of parallelepipeds the dof-space kernels also cover
  // qmode 1 at every degree (nq = 9 included)
  const bool tables_ok = !op->w1d.empty();
  const bool otf_general = tables_ok && otf_supported(op->degree, op->nq);
  const bool otf_affine = tables_ok && (otf_general || affine_dofspace_exists(op->degree, op->is_identity));
  if (otf_affine && !want_pre)
    TRY(d->scalar_bits == 64 ? count_nonaffine<double>(op) : count_nonaffine<float>(op));
  if (want_otf && !(op->n_nonaffine == 0 ? otf_affine : otf_general))
    return guard_fail(fail(BD_ERR_UNSUPPORTED,
                           "on-the-fly geometry needs tensor-product quadrature weights, the trilinear coordinate "
                           "element in tensor-product vertex order and, for cells that are not parallelepipeds, "
                           "nq <= 8 (nq = %d, %lld such cells)", op->nq, (long long)op->n_nonaffine));
  // default: the shortcut kernels when every cell is a parallelepiped and they are the measured-faster choice for
  // this (scalar type, degree, nq) (otf_preferred); a stored G otherwise -- for general cells rebuilding G per point
  // loses to streaming it (Q3 fp64 18.6 vs 20.1, Q6 26.8 vs 35.0, Q3 fp32 26.0 vs 34.8 GDoF/s)
  if (want_otf)
    op->geom_mode = op->n_nonaffine == 0 ? 1 : 2;
  else if (!want_pre && otf_affine && op->n_nonaffine == 0
           && otf_preferred(op->scalar_bits, op->degree, op->nq, op->is_identity))
    op->geom_mode = 1;
  else
    op->geom_mode = 0;
  // every cell a brick (orthogonal edges, any orientation -- the benchmark's own unperturbed box mesh): G_cell is
  // diagonal and the dof-space kernels drop the three mixed terms
  op->g_diagonal = op->geom_mode == 1 && op->n_nonorthogonal == 0 && (d->flags & BD_FLAG_GEOMETRY_FULL_TENSOR) == 0
                   && (affine_kernel_used(op->scalar_bits, op->degree, op->is_identity)
                       || affine_col_used(op->scalar_bits, op->degree, op->is_identity));
  if (op->g_diagonal && affine_kernel_used(op->scalar_bits, op->degree, op->is_identity))
  {
    // shared z faces of consecutive listed cells, per launch segment (batches of 32 / nd cells start at the segment)
    const int nd = op->degree + 1;
    TRY(op->merge_flags.alloc(static_cast<size_t>(op->n_lcells + op->n_bcells > 0 ? op->n_lcells + op->n_bcells : 1)));
    TRY(build_merge_flags(op->n_lcells, nd, 32 / nd, op->masked_dofmap.ptr, op->cell_list.ptr, op->merge_flags.ptr,
                          nullptr));
    TRY(build_merge_flags(op->n_bcells, nd, 32 / nd, op->masked_dofmap.ptr, op->cell_list.ptr + op->n_lcells,
                          op->merge_flags.ptr + op->n_lcells, nullptr));
  }
  // Geometry precompute over the whole list (compute_geometry, src/laplacian.hpp:238-272), written
  // straight into the layout the stiffness kernel streams; lcells and bcells are separate segments
  // because they are separate launches (the bcells pointer offset of src/laplacian.hpp:329).
  if (op->geom_mode == 0)
    TRY(precompute_geometry(op));
  if (cudaDeviceSynchronize() != cudaSuccess) // device_synchronize()  (src/laplacian.hpp:224)
    return guard_fail(fail(BD_ERR_CUDA, "geometry precompute failed: %s",
                           cudaGetErrorString(cudaGetLastError())));
#undef TRY
  *out = op;
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_destroy(bd_laplacian* op)
{
  if (op)
  {
    for (cudaEvent_t e : op->prof_events)
      cudaEventDestroy(e);
    for (cudaEvent_t e : op->batch_events)
      if (e)
        cudaEventDestroy(e);
    for (cudaStream_t st : op->batch_streams)
      if (st)
        cudaStreamDestroy(st);
  }
  delete op;
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_geometry(bd_laplacian* op, const void** G_dev, int64_t* n_entries)
{
  BD_REQUIRE(op && G_dev && n_entries, "bd_laplacian_geometry: null argument");
  const std::int64_t nlist = op->n_lcells + op->n_bcells;
  const size_t ts = op->scalar_size();
  const size_t bytes = static_cast<size_t>(nlist) * op->g_cell_stride() * ts;
  if (op->geom_mode != 0 && (op->G_reference.count != bytes || bytes == 0))
  {
    // on-the-fly operator: G exists only on request (tests, the assembled-CSR check), in the reference layout
    BD_CHECK(op->G_reference.alloc(bytes > 0 ? bytes : 1));
    BD_CHECK(geometry_compute_layout(op->scalar_bits, op->nq, op->xgeom.ptr, op->geometry_dofmap.ptr,
                                     op->dphi_geometry.ptr, op->weights.ptr, op->cell_list.ptr, nlist,
                                     op->G_reference.ptr, nullptr, 0));
    BD_CUDA_CHECK(cudaStreamSynchronize(nullptr));
  }
  else if (op->G_reference.count != bytes || bytes == 0)
  {
    BD_CHECK(op->G_reference.alloc(bytes > 0 ? bytes : 1));
    BD_CHECK(layout_g_to_reference(op->scalar_bits, op->nq, op->G.ptr, op->G_reference.ptr, op->n_lcells,
                                   nullptr, op->g_layout));
    BD_CHECK(layout_g_to_reference(op->scalar_bits, op->nq,
                                   op->G.ptr + static_cast<size_t>(op->g_bcells_offset) * ts,
                                   op->G_reference.ptr + static_cast<size_t>(op->n_lcells) * op->g_cell_stride() * ts,
                                   op->n_bcells, nullptr, op->g_layout));
    BD_CUDA_CHECK(cudaStreamSynchronize(nullptr));
  }
  *G_dev = op->G_reference.ptr;
  *n_entries = nlist * op->g_cell_stride();
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_apply_part(bd_laplacian* op, int part, const void* in_dev, void* out_dev,
                                       void* stream)
{
  BD_REQUIRE(op && in_dev && out_dev, "bd_laplacian_apply: null argument");
  BD_REQUIRE(part == 0 || part == 1, "part must be 0 (zero + lcells) or 1 (bcells)");
  BD_REQUIRE(in_dev != out_dev, "in and out must not alias");
  const size_t ts = op->scalar_size();
  StiffnessArgs a{};
  a.u = in_dev;
  a.y = out_dev;
  a.dofmap = op->cell_dofmap.ptr;
  a.bc = op->bc_marker.ptr;
  a.phi0_host = op->phi0.data();
  a.dphi1_host = op->dphi1.data();
  a.stream = as_stream(stream);
  a.slab = true;
  a.g_layout = op->g_layout;
  a.mdofmap = op->masked_dofmap.ptr;
  a.geom_mode = op->geom_mode;
  a.g_diagonal = op->g_diagonal;
  a.xgeom = op->xgeom.ptr;
  a.gdofmap = op->geometry_dofmap.ptr;
  a.w1d = op->w1d.data();
  a.pts1d = op->pts1d.data();
  if (part == 0)
  {
    // set_value(out, 0) over owned + ghost entries (src/laplacian.hpp:359)
    BD_CUDA_CHECK(cudaMemsetAsync(out_dev, 0, static_cast<size_t>(op->ndofs) * ts, a.stream));
    a.G = op->G.ptr;
    a.cell_constants = op->cell_constants.ptr;
    a.cells = op->cell_list.ptr;
    a.merge = op->merge_flags.ptr;
    a.ncells = op->n_lcells;
  }
  else
  {
    // geometry_ptr += 6 * Q^3 * lcells.size()  (src/laplacian.hpp:329)
    a.G = op->G.ptr + static_cast<size_t>(op->g_bcells_offset) * ts;
    a.cell_constants = op->cell_constants.ptr + static_cast<size_t>(op->n_lcells) * ts;
    a.cells = op->cell_list.ptr + op->n_lcells;
    a.merge = op->merge_flags.ptr ? op->merge_flags.ptr + op->n_lcells : nullptr;
    a.ncells = op->n_bcells;
  }
  if (!op->profile || a.ncells == 0)
    return stiffness_dispatch(op->scalar_bits, op->degree, op->nq, op->is_identity, a);
  // bracket exactly the kernel launch with events on the launching stream
  if (op->prof_used + 2 > op->prof_events.size())
  {
    for (int i = 0; i < 2; ++i)
    {
      cudaEvent_t e;
      BD_CUDA_CHECK(cudaEventCreate(&e));
      op->prof_events.push_back(e);
    }
  }
  BD_CUDA_CHECK(cudaEventRecord(op->prof_events[op->prof_used], a.stream));
  const int rc = stiffness_dispatch(op->scalar_bits, op->degree, op->nq, op->is_identity, a);
  BD_CUDA_CHECK(cudaEventRecord(op->prof_events[op->prof_used + 1], a.stream));
  op->prof_used += 2;
  return rc;
}

extern "C" int bd_laplacian_describe(bd_laplacian* op, char* buf, size_t n)
{
  BD_REQUIRE(op && buf && n > 0, "bd_laplacian_describe: null argument");
  const char* t = op->scalar_bits == 64 ? "double" : "float";
  const char* kernel = op->g_layout == 2 ? "bd::stiffness_mma8_kernel (fp64 DMMA m8n8k4 arithmetic, UBLKCP G ring)"
                       : op->g_layout == 0 ? "bd::stiffness_pencil_kernel (register pencils, streamed G rows)"
                       : (stiffness_variant() == 1 ? "bd::stiffness_slab_kernel (register slabs, LDG G)"
                                                   : "bd::stiffness_slab_tma_kernel (register slabs, UBLKCP G ring)");
  if (op->geom_mode != 0)
  {
    kernel = (op->geom_mode == 1 && affine_col_used(op->scalar_bits, op->degree, op->is_identity))
                 ? "bd::stiffness_affine_col_kernel (dof-space 1-D operators M, S, C, column rounds; geometry on the fly)"
             : op->nq >= 7 ? "bd::stiffness_mma8_otf_kernel (fp64 DMMA m8n8k4 arithmetic, geometry on the fly)"
             : (op->g_diagonal && op->degree == 3 && affine_kernel_used(op->scalar_bits, op->degree, op->is_identity))
                 ? "bd::stiffness_affine_pipe_kernel (dof-space 1-D operators M, S; asynchronous gather, shared faces "
                   "merged; geometry on the fly)"
             : (op->geom_mode == 1 && affine_kernel_used(op->scalar_bits, op->degree, op->is_identity))
                 ? "bd::stiffness_affine_kernel (dof-space 1-D operators M, S, C; geometry on the fly)"
                 : "bd::stiffness_slab_otf_kernel (register slabs, geometry on the fly)";
    snprintf(buf, n, "%s, T=%s, P=%d, nq=%d, %s", kernel, t, op->degree, op->nq,
             op->g_diagonal       ? "all cells bricks (orthogonal edges): G(q) = w_q diag(G_cell)"
             : op->geom_mode == 1 ? "all cells affine: G(q) = w_q G_cell"
                                  : "general cells: J(q) from the 8 vertices");
    return BD_SUCCESS;
  }
  snprintf(buf, n, "%s, T=%s, P=%d, nq=%d, G layout %d", kernel, t, op->degree, op->nq, op->g_layout);
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_geometry_mode(bd_laplacian* op, int* mode, int64_t* n_nonaffine)
{
  BD_REQUIRE(op && mode && n_nonaffine, "bd_laplacian_geometry_mode: null argument");
  *mode = op->g_diagonal ? 3 : op->geom_mode;
  *n_nonaffine = op->n_nonaffine;
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_profile(bd_laplacian* op, int enable)
{
  BD_REQUIRE(op, "bd_laplacian_profile: null operator");
  op->profile = enable != 0;
  op->prof_used = 0;
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_profile_read(bd_laplacian* op, double* total_ms, int64_t* launches)
{
  BD_REQUIRE(op && total_ms && launches, "bd_laplacian_profile_read: null argument");
  double tot = 0;
  for (size_t i = 0; i + 1 < op->prof_used; i += 2)
  {
    BD_CUDA_CHECK(cudaEventSynchronize(op->prof_events[i + 1]));
    float ms = 0;
    BD_CUDA_CHECK(cudaEventElapsedTime(&ms, op->prof_events[i], op->prof_events[i + 1]));
    tot += ms;
  }
  *total_ms = tot;
  *launches = static_cast<int64_t>(op->prof_used / 2);
  op->prof_used = 0;
  return BD_SUCCESS;
}

extern "C" int bd_laplacian_apply(bd_laplacian* op, const void* in_dev, void* out_dev, void* stream)
{
  BD_CHECK(bd_laplacian_apply_part(op, 0, in_dev, out_dev, stream));
  return bd_laplacian_apply_part(op, 1, in_dev, out_dev, stream);
}

extern "C" int bd_laplacian_apply_host(bd_laplacian* op, bd_halo* halo, const void* in_host,
                                       void* out_host)
{
  BD_REQUIRE(op && in_host && out_host, "bd_laplacian_apply_host: null argument");
  const size_t bytes = static_cast<size_t>(op->ndofs) * op->scalar_size();
  if (op->stage_in.count != bytes)
  {
    BD_CHECK(op->stage_in.alloc(bytes));
    BD_CHECK(op->stage_out.alloc(bytes));
  }
  BD_CUDA_CHECK(cudaMemcpyAsync(op->stage_in.ptr, in_host, bytes, cudaMemcpyHostToDevice, nullptr));
  BD_CHECK(bd_laplacian_apply_halo(op, halo, op->stage_in.ptr, op->stage_out.ptr, nullptr));
  BD_CUDA_CHECK(cudaMemcpyAsync(out_host, op->stage_out.ptr, bytes, cudaMemcpyDeviceToHost, nullptr));
  BD_CUDA_CHECK(cudaStreamSynchronize(nullptr));
  return BD_SUCCESS;
}

// Pipelined host-vector entry point: H2D of vector i+1, the operator on vector i and D2H of vector
// i-1 run concurrently on three streams with two staging pairs, so a sequence of host vectors costs
// max(H2D, apply, D2H) per vector instead of their sum (PCIe is full duplex).
extern "C" int bd_laplacian_apply_host_batch(bd_laplacian* op, bd_halo* halo, int nvec,
                                             const void* const* in_host, void* const* out_host)
{
  BD_REQUIRE(op && nvec >= 0, "bd_laplacian_apply_host_batch: bad argument");
  if (nvec == 0)
    return BD_SUCCESS;
  BD_REQUIRE(in_host && out_host, "bd_laplacian_apply_host_batch: null argument");
  const size_t bytes = static_cast<size_t>(op->ndofs) * op->scalar_size();
  if (op->stage_in.count != bytes)
  {
    BD_CHECK(op->stage_in.alloc(bytes));
    BD_CHECK(op->stage_out.alloc(bytes));
  }
  if (op->stage_in2.count != bytes)
  {
    BD_CHECK(op->stage_in2.alloc(bytes));
    BD_CHECK(op->stage_out2.alloc(bytes));
  }
  if (!op->batch_streams[0])
  {
    for (auto& st : op->batch_streams)
      BD_CUDA_CHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    for (auto& ev : op->batch_events)
      BD_CUDA_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  }
  cudaStream_t h2d = op->batch_streams[0], comp = op->batch_streams[1], d2h = op->batch_streams[2];
  char* sin[2] = {op->stage_in.ptr, op->stage_in2.ptr};
  char* sout[2] = {op->stage_out.ptr, op->stage_out2.ptr};
  cudaEvent_t* ev_in = op->batch_events;       // [2] input staged
  cudaEvent_t* ev_done = op->batch_events + 2; // [2] operator finished (input slot free, output ready)
  cudaEvent_t* ev_out = op->batch_events + 4;  // [2] output copied (output slot free)
  for (int i = 0; i < nvec; ++i)
  {
    const int s = i & 1;
    BD_REQUIRE(in_host[i] && out_host[i], "bd_laplacian_apply_host_batch: null vector %d", i);
    if (i >= 2)
      BD_CUDA_CHECK(cudaStreamWaitEvent(h2d, ev_done[s], 0));
    BD_CUDA_CHECK(cudaMemcpyAsync(sin[s], in_host[i], bytes, cudaMemcpyHostToDevice, h2d));
    BD_CUDA_CHECK(cudaEventRecord(ev_in[s], h2d));
    BD_CUDA_CHECK(cudaStreamWaitEvent(comp, ev_in[s], 0));
    if (i >= 2)
      BD_CUDA_CHECK(cudaStreamWaitEvent(comp, ev_out[s], 0));
    BD_CHECK(bd_laplacian_apply_halo(op, halo, sin[s], sout[s], comp));
    BD_CUDA_CHECK(cudaEventRecord(ev_done[s], comp));
    BD_CUDA_CHECK(cudaStreamWaitEvent(d2h, ev_done[s], 0));
    BD_CUDA_CHECK(cudaMemcpyAsync(out_host[i], sout[s], bytes, cudaMemcpyDeviceToHost, d2h));
    BD_CUDA_CHECK(cudaEventRecord(ev_out[s], d2h));
  }
  BD_CUDA_CHECK(cudaStreamSynchronize(d2h));
  BD_CUDA_CHECK(cudaStreamSynchronize(comp));
  return BD_SUCCESS;
}
