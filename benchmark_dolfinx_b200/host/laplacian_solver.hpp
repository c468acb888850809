// Host-side mirror of the reference's src/laplacian_solver.{hpp,cpp}: the timed benchmark body.
#pragma once

#include <chrono>
#include <iostream>
#include <memory>

#include "cg.hpp"
#include "csr.hpp"

namespace benchdolfinx
{
/// src/laplacian_solver.hpp:13-20
struct BenchmarkResults
{
  double mat_free_time;
  double unorm;
  double ynorm;
  double znorm;
  double enorm;
};

/// src/laplacian_solver.cpp:66-230.  `barrier` synchronises the ranks (threads) around the timed region.
template <typename T, typename Barrier>
BenchmarkResults laplace_action_gpu(BoxMesh<T>& mesh, int degree, int qmode, T kappa, int nreps,
                                    bool use_gauss, bool matrix_comparison, bool use_cg, int rank,
                                    Barrier&& barrier)
{
  Vector<T> u(&mesh.map, 1), y(&mesh.map, 1);
  MatFreeLaplacianGPU<T> op(mesh, degree, qmode, kappa, use_gauss);
  // u := b = assemble_vector(L), BC rows zeroed (src/laplacian_solver.cpp:100-109), then a full scatter (:111-114)
  check(bd_boxmesh_assemble_rhs(mesh.handle, qmode, use_gauss, u.array(), nullptr));
  u.scatter_fwd();
  check_cuda(cudaDeviceSynchronize());
  BenchmarkResults b_results = {0, 0, 0, 0, 0};

  barrier();
  auto start = std::chrono::high_resolution_clock::now();
  if (use_cg)
    cg_solve(op, y, u, nreps, T(0.0));
  else
  {
    for (int i = 0; i < nreps; ++i)
      op.apply(u, y);
  }
  check_cuda(cudaDeviceSynchronize());
  barrier();
  auto stop = std::chrono::high_resolution_clock::now();
  std::chrono::duration<double> duration = stop - start;

  const double unorm = norm(u), ynorm = norm(y);
  b_results.unorm = unorm;
  b_results.ynorm = ynorm;
  b_results.mat_free_time = duration.count();
  if (rank == 0)
  {
    const std::int64_t ndofs_global = mesh.map.size_global();
    std::string comp_type = use_cg ? "CG" : "Action";
    std::cout << "Computation time (" << comp_type << "): " << duration.count() << "s" << std::endl;
    std::cout << "Computation rate (Gdofs/s): " << ndofs_global * nreps / (1e9 * duration.count()) << std::endl;
    std::cout << "Norm of u = " << unorm << std::endl;
    std::cout << "Norm of y = " << ynorm << std::endl;
  }

  if (matrix_comparison)
  {
    MatrixOperator<T> mat_op(op, mesh.map);
    Vector<T> z(&mesh.map, 1);
    if (use_cg)
      cg_solve(mat_op, z, u, nreps, T(0.0));
    else
    {
      for (int i = 0; i < nreps; ++i)
        mat_op.apply(u, z);
    }
    Vector<T> e(&mesh.map, 1);
    axpy(e, T{-1}, y, z);
    const double znorm = norm(z), enorm = norm(e);
    b_results.znorm = znorm;
    b_results.enorm = enorm;
    if (rank == 0)
    {
      std::cout << "Norm of u = " << unorm << std::endl;
      std::cout << "Norm of z = " << znorm << std::endl;
      std::cout << "Norm of error = " << enorm << std::endl;
      std::cout << "Relative norm of error = " << enorm / znorm << std::endl;
    }
  }
  return b_results;
}
} // namespace benchdolfinx
