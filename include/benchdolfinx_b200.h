/* benchdolfinx_b200 -- C ABI of the B200-native matrix-free Poisson hot path.
 *
 * Drop-in boundary for benchmark-dolfinx's GPU operator path: the sum-factorised
 * Q_k hexahedral Laplacian apply(in, out), its geometry-factor precompute, and the
 * CG / vector kernels around it.  Plain pointers and sizes only; no C++, CUDA or
 * torch types cross this boundary (`stream` arguments are a `cudaStream_t` passed
 * as `void*`, NULL = the legacy default stream).
 *
 * Every entry point cites the reference interface it replaces (paths relative to
 * the reference tree, file:line).  All functions return BD_SUCCESS (0) or a
 * negative error code; bd_last_error() returns the message of the last failure on
 * the calling thread.  No exceptions cross the ABI (the reference throws
 * std::runtime_error / exit(1)s: src/laplacian.hpp:378,400, src/util.hpp:36-45).
 *
 * Scalars: `scalar_bits` is 32 (float) or 64 (double) -- the reference's
 * `--float` option (src/main.cpp:150-151,277-288).  Vectors follow the
 * dolfinx::la::Vector layout the reference uses (src/vector.hpp:88-90): owned
 * entries first, ghost entries after.
 */
#ifndef BENCHDOLFINX_B200_H
#define BENCHDOLFINX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BD_SUCCESS 0
#define BD_ERR_INVALID_ARGUMENT (-1) /* bad size / null pointer / inconsistent descriptor */
#define BD_ERR_UNSUPPORTED (-2)      /* "Unsupported degree [operator]" / "Unsupported nq" */
#define BD_ERR_CUDA (-3)             /* a CUDA runtime call failed */
#define BD_ERR_NCCL (-4)             /* NCCL missing or a NCCL call failed */
#define BD_ERR_NO_DEVICE (-5)        /* no usable CUDA device: there is no CPU fallback */

/* Message of the last error on this thread ("" if none). Never NULL. */
const char* bd_last_error(void);

/* Test hook, no counterpart in the reference (its kernels launch one CTA per cell,
 * src/laplacian.hpp:298-306, so they have no persistent steady state to exercise): caps the
 * grid of every persistent stiffness kernel at `max_ctas` CTAs (0 = no cap, the default).
 * With 2-3 CTAs a test mesh of a few thousand cells drives each warp through many batches,
 * which is where the benchmark sizes spend all their time.  Process-wide; not for production. */
int bd_debug_set_max_grid(int max_ctas);

/* Measured fp64 arithmetic peaks of the current device in TFLOP/s (plain DFMA and tensor-core DMMA m8n8k4
 * loops without dependent chains, ~50 ms): roofline denominators for the compute-bound variants.  No
 * counterpart in the reference. */
int bd_fp64_peak(double* dfma_tflops, double* dmma_tflops);

/* Library / device banner (replaces get_device_information, src/util.cpp:10-52).
 * Writes at most `n` bytes including the terminator. */
int bd_device_information(char* buf, size_t n);

/* ------------------------------------------------------------------------- *
 * Operator: replaces MatFreeLaplacianGPU<T> (src/laplacian.hpp:89-447)
 * ------------------------------------------------------------------------- */
typedef struct bd_laplacian bd_laplacian; /* opaque handle */

/* Everything the reference constructor (src/laplacian.hpp:102-227) extracts from
 * the DOLFINx FunctionSpace / DirichletBC / Basix, as raw arrays.  Each pointer
 * may be a host or a device pointer (detected); the handle keeps its own device
 * copy unless BD_FLAG_BORROW_DEVICE_ARRAYS is set, in which case *device* arrays
 * are used in place and must outlive the handle. */
typedef struct bd_laplacian_desc
{
  int32_t scalar_bits;  /* 32 | 64 */
  int32_t degree;       /* P, 1..7                      (src/laplacian.hpp:361-400) */
  int32_t nq;           /* 1-D quadrature points, P+1 or P+2 (_op_nq, :173) */
  int32_t is_identity;  /* phi0 is the identity         (_is_identity, :194) */
  double kappa;         /* coefficient on all cells     (_cell_constants, :105) */
  int64_t ncells;       /* owned + ghost cells */
  int64_t ndofs;        /* owned + ghost dofs */
  int64_t nnodes;       /* geometry nodes */
  int64_t n_lcells;     /* cells with no ghost dof      (_lcells, :116-121) */
  int64_t n_bcells;     /* boundary + ghost cells       (_bcells) */
  const int32_t* cell_dofmap;     /* [ncells][ (P+1)^3 ]  (_cell_dofmap, :106-108) */
  const void* xgeom;              /* T [nnodes][3]        (_xgeom, :109-110) */
  const int32_t* geometry_dofmap; /* [ncells][8], tp order (_geometry_dofmap, :111-113) */
  const int8_t* bc_marker;        /* [ndofs]              (_bc_marker, :114) */
  const int32_t* lcells;          /* [n_lcells] */
  const int32_t* bcells;          /* [n_bcells] */
  const void* phi0;               /* T [nq][P+1]          (_phi0_const, :206-207) */
  const void* dphi1;              /* T [nq][nq]           (_dphi1_const, :210-212) */
  const void* weights;            /* T [nq^3]             (_g_weights, :158) */
  const void* dphi_geometry;      /* T [3][nq^3][8]       (_dphi_geometry, :157) */
  uint32_t flags;
} bd_laplacian_desc;

#define BD_FLAG_BORROW_DEVICE_ARRAYS 1u
/* Geometry variant of the operator (reference: geometry_computation_gpu fills G once, src/laplacian.hpp:238-272,
 * and stiffness_operator_gpu streams it, src/laplacian_gpu.hpp:256-273).  Default (neither flag): G is computed on
 * the fly inside the stiffness kernel when every cell of the mesh is a parallelepiped (G(q) = w_q G_cell, no G
 * array at all), precomputed and streamed otherwise.  The flags force one or the other; on-the-fly needs the
 * tensor-product quadrature / trilinear coordinate element the reference uses and nq <= 8, else
 * BD_ERR_UNSUPPORTED. */
#define BD_FLAG_GEOMETRY_ONTHEFLY 2u
#define BD_FLAG_GEOMETRY_PRECOMPUTED 4u
/* On meshes whose cells all have mutually orthogonal edges (bricks, e.g. the benchmark's unperturbed box) G_cell is
 * diagonal and the on-the-fly dof-space kernels skip its three off-diagonal terms.  This flag keeps the full
 * symmetric G_cell (for A/B measurements and tests of the general parallelepiped kernel on a box mesh). */
#define BD_FLAG_GEOMETRY_FULL_TENSOR 8u

/* Constructor: copies/borrows the arrays, precomputes G over [lcells..., bcells...]
 * (compute_geometry, src/laplacian.hpp:214-224,238-272) on the current device. */
int bd_laplacian_create(const bd_laplacian_desc* desc, bd_laplacian** out);
int bd_laplacian_destroy(bd_laplacian* op);

/* out = A in for device vectors of length ndofs (owned + ghost); replaces
 * MatFreeLaplacianGPU::apply (src/laplacian.hpp:355-403) for a vector whose ghost
 * entries are already up to date: zero `out` (:359), stiffness kernel on lcells
 * then bcells (:291-345).  Asynchronous on `stream` (the reference ends with
 * device_synchronize, :347; call cudaStreamSynchronize yourself if you need it). */
int bd_laplacian_apply(bd_laplacian* op, const void* in_dev, void* out_dev, void* stream);

/* Same operator, split as the reference overlaps it with the halo exchange
 * (impl_operator, src/laplacian.hpp:281-349): part 0 zeroes `out` and runs the
 * lcells kernel, part 1 runs the bcells kernel. */
int bd_laplacian_apply_part(bd_laplacian* op, int part, const void* in_dev, void* out_dev,
                            void* stream);

/* End-to-end entry point for HOST vectors of length ndofs: H2D copy of `in_host`, forward
 * scatter (if `halo` is not NULL) + apply, D2H copy of the result, synchronous.  Pinned host
 * memory gives full PCIe rate.  (The reference benchmark does the H2D at
 * src/laplacian_solver.cpp:108-109 and reads results back at :130-131.) */
struct bd_halo;
int bd_laplacian_apply_host(bd_laplacian* op, struct bd_halo* halo, const void* in_host,
                            void* out_host);

/* The same for a sequence of host vectors (e.g. several right-hand sides), pipelined: the H2D copy
 * of vector i+1, the operator on vector i and the D2H copy of vector i-1 overlap on three streams
 * with two device staging pairs.  out_host[i] = A in_host[i]; synchronous on return. */
int bd_laplacian_apply_host_batch(bd_laplacian* op, struct bd_halo* halo, int nvec,
                                  const void* const* in_host, void* const* out_host);

/* Per-launch device timing of the stiffness kernels, for the roofline report: while enabled,
 * every kernel launch of bd_laplacian_apply* is bracketed by CUDA events on the launching stream
 * (no synchronisation); bd_laplacian_profile_read waits for them, returns the summed kernel time
 * and the number of launches since the last read, and resets the counters.  (The reference only
 * has dolfinx::common::Timer scopes around whole phases, src/laplacian_solver.cpp:90,197.) */
int bd_laplacian_profile(bd_laplacian* op, int enable);
/* Which geometry variant the operator uses (no counterpart in the reference, which always precomputes):
 * *mode = 0 precomputed G, 1 on the fly with every cell affine, 2 on the fly with general cells, 3 on the fly with
 * every cell a brick (affine with orthogonal edges: diagonal G_cell, dof-space kernels only);
 * *n_nonaffine = number of listed cells that are not parallelepipeds (to the rounding of their coordinates). */
int bd_laplacian_geometry_mode(bd_laplacian* op, int* mode, int64_t* n_nonaffine);
/* One-line description of the stiffness kernel variant this operator launches (for reports). */
int bd_laplacian_describe(bd_laplacian* op, char* buf, size_t n);
int bd_laplacian_profile_read(bd_laplacian* op, double* total_ms, int64_t* launches);

/* Read-only access to the precomputed geometry tensor, device pointer to
 * T [n_lcells + n_bcells][6][nq^3] (_g_entity, src/laplacian.hpp:429). */
int bd_laplacian_geometry(bd_laplacian* op, const void** G_dev, int64_t* n_entries);

/* Geometry-factor kernel on its own; replaces geometry_computation_gpu<T,Q>
 * (src/geometry_gpu.hpp:26-132).  All pointers are device pointers. */
int bd_geometry_compute(int scalar_bits, int nq, const void* xgeom, const int32_t* geometry_dofmap,
                        const void* dphi_geometry, const void* weights, const int32_t* cells,
                        int64_t ncells, void* G_out, void* stream);

/* Stiffness kernel on its own for one cell list; replaces the launch of
 * stiffness_operator_gpu<T,P,Q> (src/laplacian_gpu.hpp:91-426; launch sites
 * src/laplacian.hpp:298-306,334-342).  `G` is indexed by position in `cells`,
 * `cell_constants` likewise.  Accumulates into `out` (no zeroing).  Device pointers. */
int bd_stiffness_apply(int scalar_bits, int degree, int nq, int is_identity, const void* in,
                       const void* cell_constants, void* out, const void* G, const void* phi0_host,
                       const void* dphi1_host, const int32_t* cell_dofmap, const int32_t* cells,
                       int64_t ncells, const int8_t* bc_marker, void* stream);

/* ------------------------------------------------------------------------- *
 * Vector kernels: replace src/vector.hpp and detail:: of src/cg.hpp.
 * `n` counts owned entries (size_local), as in the reference (src/vector.hpp:164).
 * ------------------------------------------------------------------------- */
/* r = alpha*x + y  (axpy, src/vector.hpp:228-240; detail::axpy src/cg.hpp:21-43) */
int bd_axpy(int scalar_bits, int64_t n, double alpha, const void* x, const void* y, void* r,
            void* stream);
/* x[0:n] = value (set_value, src/vector.hpp:288-292) */
int bd_set_value(int scalar_bits, int64_t n, double value, void* x, void* stream);
/* Bytes of device scratch bd_dot / bd_cg_solve need for vectors of any length. */
size_t bd_dot_scratch_bytes(void);
/* *result_dev (a double in device memory) = sum_i a[i]*b[i] over n entries
 * (local part of inner_product, src/vector.hpp:159-176 / src/cg.hpp:51-79, without
 * the MPI_Allreduce).  Products are accumulated in double for both scalar types,
 * in a fixed order (bitwise reproducible run to run).  `scratch_dev` must hold
 * bd_dot_scratch_bytes() bytes and must not be shared by overlapping calls. */
int bd_dot(int scalar_bits, int64_t n, const void* a, const void* b, double* result_dev,
           void* scratch_dev, void* stream);
/* Allocation-free variant of bd_dot.  It has no counterpart in the reference (whose
 * thrust::inner_product allocates its own temporary storage on every call).  It exists for host
 * codes that cannot let a library allocate on the device: a solver loop kept fully asynchronous
 * (cudaMalloc / cudaFree synchronise the device), or a run that has already handed all of the
 * GPU's memory to its own pools, where a late cudaMalloc can simply fail.  The per-CTA partial
 * sums go to a fixed 16 KB array in the library's device image (one copy on every GPU that loads
 * the library, held until exit) instead of caller scratch; the call itself performs no
 * cudaMalloc / cudaFree and calls no allocating routine.
 *   - Vectors of up to BD_DOT_STATIC_MAX_N entries; longer ones are rejected with
 *     BD_ERR_INVALID_ARGUMENT (use bd_dot).
 *   - The first call on each device looks up the array's device address (that device must be
 *     current); make that first call before starting any stream capture.  CUDA loads device
 *     code lazily: if no vector kernel of this library (bd_axpy, bd_dot, bd_set_value, ...) has
 *     run on the device yet, that first call is also the one that loads the code and places the
 *     array, so it has to happen while the device still has a little free memory -- i.e. before
 *     the application hands the rest of it to its own pools.  Later calls take nothing.
 *   - All calls on a device share the array: two calls whose kernels can overlap on the GPU
 *     (different streams, or different host threads) race on it and give wrong sums; calls on
 *     one stream are safe.  The array keeps the previous call's partial sums; none of it needs
 *     clearing, every entry that is read is first overwritten by the same call.
 *   - Same kernels, grid and summation order as bd_dot: results are bitwise identical. */
#define BD_DOT_STATIC_MAX_N 2147483647LL
int bd_dot_static(int scalar_bits, int64_t n, const void* a, const void* b, double* result_dev,
                  void* stream);
/* Host-returning convenience: local dot, synchronous (allocates its own scratch). */
int bd_dot_host(int scalar_bits, int64_t n, const void* a, const void* b, double* result_host);
/* out[i] = in[idx[i]]   (pack_gpu, src/vector.hpp:31-41) */
int bd_pack(int scalar_bits, int64_t n, const int32_t* idx, const void* in, void* out,
            void* stream);
/* out[idx[i]] = in[i]   (unpack_gpu, src/vector.hpp:52-62) */
int bd_unpack(int scalar_bits, int64_t n, const int32_t* idx, const void* in, void* out,
              void* stream);

/* ------------------------------------------------------------------------- *
 * Inter-GPU: replaces the CUDA-aware MPI of the reference (DOLFINx Scatterer driven
 * by Vector::scatter_fwd_begin/end, src/laplacian.hpp:286,320, src/cg.hpp:133-136;
 * MPI_Allreduce, src/vector.hpp:173, src/cg.hpp:76) by NCCL over NVLink.
 * One communicator per rank (= per GPU), any process/thread layout.
 * ------------------------------------------------------------------------- */
typedef struct bd_comm bd_comm;
#define BD_UNIQUE_ID_BYTES 128
/* Rank 0 creates the id and distributes the 128 bytes by any means. */
int bd_comm_unique_id(void* id_out);
int bd_comm_create(const void* id, int rank, int nranks, bd_comm** out);
int bd_comm_destroy(bd_comm* comm);

/* Forward-scatter plan (owner -> ghost).  Ghost entries are ordered by source
 * rank, so each receive lands directly in the ghost segment (no unpack kernel);
 * send entries are packed with bd_pack's kernel into an internal buffer. */
typedef struct bd_halo bd_halo;
int bd_halo_create(bd_comm* comm, int scalar_bits, int64_t n_owned, int64_t n_ghost, int n_send,
                   const int32_t* send_ranks, const int64_t* send_counts,
                   const int32_t* send_indices_host, int n_recv, const int32_t* recv_ranks,
                   const int64_t* recv_counts, bd_halo** out);
/* For ghost segments that are NOT ordered by source rank (DOLFINx makes no such promise): message
 * entry i (messages concatenated in recv_ranks order) belongs at ghost position ghost_positions[i];
 * the plan then receives into an internal buffer and scatters with the unpack kernel (unpack_gpu,
 * src/vector.hpp:52-62, get_unpack_fn :125-149).  NULL switches back to in-place receives. */
int bd_halo_set_unpack(bd_halo* halo, const int32_t* ghost_positions_host);
int bd_halo_destroy(bd_halo* halo);
/* x[n_owned:] <- owners' values.  begin: pack + post sends/receives on the plan's
 * own stream after the work already queued on `stream`; end: make `stream` wait. */
int bd_halo_begin(bd_halo* halo, void* x_dev, void* stream);
int bd_halo_end(bd_halo* halo, void* stream);
/* Optional timing of the exchange inside bd_laplacian_apply_halo (no counterpart in the reference, whose
 * scatter is synchronous, src/vector.hpp:117,146): while enabled, every application records CUDA events around
 * the pack, the NCCL send/recv group and the interior-cell kernel.  bd_halo_profile_read sums, over the
 * applications since the last read: pack_ms (pack kernel, on the caller's stream), exchange_ms (the NCCL group on
 * the plan's stream) and exposed_ms (how long the ghost values arrived AFTER the interior-cell kernel had
 * finished, i.e. the part of the exchange that was not hidden). */
int bd_halo_profile(bd_halo* halo, int enable);
int bd_halo_profile_read(bd_halo* halo, double* pack_ms, double* exchange_ms, double* exposed_ms,
                         int64_t* applications);

/* out = A in including the halo exchange of `in`, overlapped with the lcells
 * kernel exactly as impl_operator does (src/laplacian.hpp:281-349).  `halo` may
 * be NULL (single rank).  `in_dev` is modified in its ghost segment only. */
int bd_laplacian_apply_halo(bd_laplacian* op, bd_halo* halo, void* in_dev, void* out_dev,
                            void* stream);

/* ------------------------------------------------------------------------- *
 * CG: replaces cg_solve (src/cg.hpp:89-169): unpreconditioned, same update order,
 * `max_iter` iterations when rtol == 0.  x and b are device vectors of length
 * n_owned + n_ghost; `halo`/`comm` may be NULL on a single rank.  All scalars stay
 * on the device; the call is asynchronous on `stream` unless rtol > 0 (which needs
 * the residual on the host each iteration, as the reference does).
 * Returns the number of iterations in *iterations.
 * ------------------------------------------------------------------------- */
int bd_cg_solve(bd_laplacian* op, bd_halo* halo, bd_comm* comm, void* x_dev, const void* b_dev,
                int max_iter, double rtol, int* iterations, void* stream);

/* Jacobi preconditioning.  The reference carries the scaffolding only: the inverse diagonal of the
 * *assembled* matrix (compute_diag_inv, src/csr.hpp:79-107; MatrixOperator::get_diag_inverse, :154-162),
 * pointwise_mult (src/vector.hpp:266-285) and the remark "p <- beta*p + M^-1(r)" in cg_solve
 * (src/cg.hpp:165).
 * bd_laplacian_diagonal_inverse: diag_inv_dev[i] = 1 / A_ii for the operator's ndofs entries, computed
 * matrix-free from the operator's tables and geometry (1 on constrained rows, as the assembled matrix after
 * set_diagonal, src/laplacian_solver.cpp:182-184).  Owned entries are complete; ghost entries hold partial
 * sums and must not be used.  Synchronous (allocates and frees a temporary of ndofs doubles). */
int bd_laplacian_diagonal_inverse(bd_laplacian* op, void* diag_inv_dev, void* stream);
/* w[i] = x[i] * y[i] over n entries (pointwise_mult, src/vector.hpp:266-285). */
int bd_pointwise_mult(int scalar_bits, int64_t n, const void* x, const void* y, void* w, void* stream);
/* CG with z = D^-1 r: p0 = z0, alpha = (r.z)/(p.Ap), beta = (r.z)_new/(r.z), p = beta p + z; stops when
 * (r.z)/(r.z)_0 < rtol^2 (the reference's test on its own `rnorm`, src/cg.hpp:162, which is r.z once p0 = z0).
 * Arguments as bd_cg_solve plus the inverse diagonal. */
int bd_cg_solve_jacobi(bd_laplacian* op, bd_halo* halo, bd_comm* comm, void* x_dev, const void* b_dev,
                       const void* diag_inv_dev, int max_iter, double rtol, int* iterations, void* stream);

/* Global inner product / l2 norm over owned entries, synchronous, host result
 * (inner_product + MPI_Allreduce, src/vector.hpp:159-176; norm, :196-218). */
int bd_inner_product(bd_comm* comm, int scalar_bits, int64_t n_owned, const void* a, const void* b,
                     double* result_host);
int bd_norm(bd_comm* comm, int scalar_bits, int64_t n_owned, const void* a, double* result_host);

/* ------------------------------------------------------------------------- *
 * Assembled-CSR check path (--mat_comp): replaces MatrixOperator<T> (src/csr.hpp:116-233) and the
 * CPU assembly that feeds it (src/laplacian_solver.cpp:160-191).  The matrix is assembled on the
 * device from the operator's own mesh data and geometry tensor, without sum factorisation: BC rows
 * and columns dropped, unit diagonal on BC rows, owned rows only (n_owned of them).
 * A correctness check, not a performance path.
 * ------------------------------------------------------------------------- */
typedef struct bd_csr bd_csr;
int bd_csr_assemble(bd_laplacian* op, int64_t n_owned, bd_csr** out);
int bd_csr_destroy(bd_csr* A);
int bd_csr_info(bd_csr* A, int64_t* nrows, int64_t* ncols, int64_t* nnz);
/* out = A in  (MatrixOperator::apply, src/csr.hpp:174-217): zero `out` (owned + ghost), then the
 * owned rows; `in` needs current ghost values. */
int bd_csr_apply(bd_csr* A, const void* in_dev, void* out_dev, void* stream);

/* ------------------------------------------------------------------------- *
 * Stand-in for the DOLFINx host setup (mesh, dofmap, BCs, RHS), generated on the
 * device for one rank of a block partition of the compute_mesh_size box
 * (src/mesh.cpp:117-152,190-218; src/main.cpp:54-102; src/laplacian_solver.cpp:100-109).
 * ------------------------------------------------------------------------- */
typedef struct bd_boxmesh bd_boxmesh;
typedef struct bd_boxmesh_info
{
  int64_t ncells_owned, ncells, ndofs_owned, nghost, nnodes, n_lcells, n_bcells;
  int64_t ndofs_global, ncells_global;
  const int32_t* cell_dofmap;     /* device */
  const int32_t* geometry_dofmap; /* device */
  const void* xgeom;              /* device, T[nnodes][3] */
  const int8_t* bc_marker;        /* device */
  const int32_t* lcells;          /* device */
  const int32_t* bcells;          /* device */
  int32_t n_send, n_recv;
  const int32_t* send_ranks;        /* host [n_send] */
  const int64_t* send_counts;       /* host [n_send] */
  const int32_t* send_indices_host; /* host, concatenated */
  const int32_t* recv_ranks;        /* host [n_recv] */
  const int64_t* recv_counts;       /* host [n_recv] */
} bd_boxmesh_info;

/* src/mesh.cpp:117-152 */
int bd_compute_mesh_size(int64_t ndofs_global, int degree, int64_t nx_out[3]);
int bd_boxmesh_create(int scalar_bits, const int64_t n[3], int degree, int rank, int nranks,
                      const int32_t pgrid[3], double geom_perturb_fact, bd_boxmesh** out);
int bd_boxmesh_get(bd_boxmesh* mesh, bd_boxmesh_info* info);
int bd_boxmesh_destroy(bd_boxmesh* mesh);
/* b = M f with BC rows zeroed on owned dofs, ghost entries 0: assemble_vector(L) + bc.set
 * (src/laplacian_solver.cpp:100-105) with f = 1000 exp(-((x-.5)^2+(y-.5)^2)/0.02) interpolated at
 * the physical dof nodes (src/main.cpp:81-92) and the operator's quadrature (qmode, use_gauss).
 * b_dev has ndofs_owned + nghost entries of the mesh's scalar type.  Synchronous. */
int bd_boxmesh_assemble_rhs(bd_boxmesh* mesh, int qmode, int use_gauss, void* b_dev, void* stream);

/* 1-D tables the reference takes from Basix (src/laplacian.hpp:123-212), host
 * arrays of `scalar_bits` type: phi0[nq*(P+1)], dphi1[nq*nq], weights[nq^3],
 * dphi_geometry[3*nq^3*8]; dof_nodes[P+1] and qpts[nq] are always double.
 * Returns is_identity in *is_identity. */
int bd_tables(int scalar_bits, int degree, int qmode, int use_gauss, void* phi0, void* dphi1,
              void* weights, void* dphi_geometry, double* dof_nodes, double* qpts,
              int* is_identity);

#ifdef __cplusplus
}
#endif
#endif /* BENCHDOLFINX_B200_H */
