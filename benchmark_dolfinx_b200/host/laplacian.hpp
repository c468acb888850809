// Host-side mirror of MatFreeLaplacianGPU<T> (reference src/laplacian.hpp:89-447) on the C ABI.
#pragma once

#include <vector>

#include "mesh.hpp"

namespace benchdolfinx
{
template <typename T>
class MatFreeLaplacianGPU
{
public:
  using value_type = T;

  /// Same argument meaning as src/laplacian.hpp:102-104; the FunctionSpace / DirichletBC pair is the
  /// stand-in BoxMesh, `use_gauss` replaces basix::quadrature::type.
  MatFreeLaplacianGPU(const BoxMesh<T>& mesh, int degree, int qmode, T constant, bool use_gauss)
      : _halo(mesh.map.halo), _comm(mesh.map.comm_)
  {
    const int nd = degree + 1, nq = degree + 1 + qmode, nq3 = nq * nq * nq;
    std::vector<T> phi0(nq * nd), dphi1(nq * nq), weights(nq3), dphi(3 * static_cast<std::size_t>(nq3) * 8);
    int is_identity = 0;
    check(bd_tables(8 * sizeof(T), degree, qmode, use_gauss, phi0.data(), dphi1.data(), weights.data(),
                    dphi.data(), nullptr, nullptr, &is_identity));
    if (qmode == 0 && !is_identity)
      throw std::runtime_error("Expecting identity matrix for qmode=0"); // src/laplacian.hpp:197-198
    bd_laplacian_desc d{};
    d.scalar_bits = 8 * sizeof(T);
    d.degree = degree;
    d.nq = nq;
    d.is_identity = is_identity;
    d.kappa = constant;
    d.ncells = mesh.info.ncells;
    d.ndofs = mesh.info.ndofs_owned + mesh.info.nghost;
    d.nnodes = mesh.info.nnodes;
    d.n_lcells = mesh.info.n_lcells;
    d.n_bcells = mesh.info.n_bcells;
    d.cell_dofmap = mesh.info.cell_dofmap;
    d.xgeom = mesh.info.xgeom;
    d.geometry_dofmap = mesh.info.geometry_dofmap;
    d.bc_marker = mesh.info.bc_marker;
    d.lcells = mesh.info.lcells;
    d.bcells = mesh.info.bcells;
    d.phi0 = phi0.data();
    d.dphi1 = dphi1.data();
    d.weights = weights.data();
    d.dphi_geometry = dphi.data();
    d.flags = BD_FLAG_BORROW_DEVICE_ARRAYS;
    check(bd_laplacian_create(&d, &_op));
  }
  ~MatFreeLaplacianGPU() { bd_laplacian_destroy(_op); }
  MatFreeLaplacianGPU(const MatFreeLaplacianGPU&) = delete;

  /// out = A in, including the forward scatter of `in` (src/laplacian.hpp:355-403, 281-349)
  template <typename Vector>
  void apply(Vector& in, Vector& out)
  {
    check(bd_laplacian_apply_halo(_op, _halo, in.array(), out.array(), nullptr));
  }

  bd_laplacian* handle() const { return _op; }
  bd_halo* halo() const { return _halo; }
  bd_comm* comm() const { return _comm; }

private:
  bd_laplacian* _op = nullptr;
  bd_halo* _halo;
  bd_comm* _comm;
};
} // namespace benchdolfinx
