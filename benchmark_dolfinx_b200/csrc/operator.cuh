// Host-side operator object behind the opaque bd_laplacian handle.
#pragma once

#include <vector>

#include "common.cuh"

struct bd_laplacian
{
  int scalar_bits = 64;
  int degree = 0, nq = 0;
  bool is_identity = false;
  double kappa = 0;
  std::int64_t ncells = 0, ndofs = 0, nnodes = 0, n_lcells = 0, n_bcells = 0;
  int device = 0;
  // device data (owned or borrowed)
  bd::DeviceArray<std::int32_t> cell_dofmap, geometry_dofmap, cell_list; // cell_list = [lcells.., bcells..]
  bd::DeviceArray<std::int8_t> bc_marker;
  bd::DeviceArray<std::int32_t> masked_dofmap; // cell_dofmap with the BC marker in bit 31
  bd::DeviceArray<char> xgeom, weights, dphi_geometry, G, cell_constants;
  // 1-D tables on the host (they travel to the kernel as a launch parameter)
  std::vector<char> phi0, dphi1;
  // staging for bd_laplacian_apply_host
  bd::DeviceArray<char> stage_in, stage_out;
  // second staging pair + streams/events of the pipelined bd_laplacian_apply_host_batch
  bd::DeviceArray<char> stage_in2, stage_out2;
  cudaStream_t batch_streams[3] = {nullptr, nullptr, nullptr}; // h2d, compute, d2h
  cudaEvent_t batch_events[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  // optional per-launch timing of the stiffness kernels (bd_laplacian_profile)
  bool profile = false;
  std::vector<cudaEvent_t> prof_events; // pairs (start, stop), consumed by bd_laplacian_profile_read
  size_t prof_used = 0;
  // work vectors + scalars of bd_cg_solve, kept between calls
  bd::DeviceArray<char> cg_work;
  size_t scalar_size() const { return static_cast<size_t>(scalar_bits / 8); }
  // G is stored in the slab layout (stiffness_slab.cuh): lcells segment, then bcells segment,
  // each rounded up to whole warp batches
  bool slab = true;
  int g_layout = 1; // 0 reference order (pencil kernel), 1 slab, 2 DMMA accumulator order (see stiffness_launch.cuh)
  std::int64_t g_bcells_offset = 0; // scalars before the bcells segment
  bd::DeviceArray<char> G_reference; // reference-layout copy, built on demand (bd_laplacian_geometry)
  // Geometry variant (bd_laplacian_desc.flags): 0 = G precomputed and streamed, 1 = computed
  // on the fly and every cell of both lists is affine (shortcut kernel), 2 = on the fly, general cells.
  int geom_mode = 0;
  std::int64_t n_nonaffine = -1;      // cells of [lcells, bcells] that are not parallelepipeds (-1: not counted)
  std::int64_t n_nonorthogonal = -1;  // ... that are not parallelepipeds with mutually orthogonal edges (bricks)
  bd::DeviceArray<std::int8_t> merge_flags; // per list position, see build_merge_flags (dof-space pipe kernel)
  bool g_diagonal = false;            // geom_mode 1, n_nonorthogonal == 0 and a dof-space kernel serves this configuration
  std::vector<double> w1d, pts1d;     // 1-D rule recovered from weights / dphi_geometry (empty: not tensor-product)
  std::int64_t g_cell_stride() const { return 6ll * nq * nq * nq; }
};
