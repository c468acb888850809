// Host-side mirror of MatrixOperator<T> (reference src/csr.hpp:116-233): the assembled operator of the
// --mat_comp check, assembled on the device from the matrix-free operator's own data.
#pragma once

#include "laplacian.hpp"

namespace benchdolfinx
{
template <typename T>
class MatrixOperator
{
public:
  using value_type = T;
  MatrixOperator(const MatFreeLaplacianGPU<T>& op, const IndexMap& map) : _halo(op.halo())
  {
    check(bd_csr_assemble(op.handle(), map.size_local(), &_A));
  }
  ~MatrixOperator() { bd_csr_destroy(_A); }
  MatrixOperator(const MatrixOperator&) = delete;
  /// y = A x (src/csr.hpp:174-217)
  template <typename Vector>
  void apply(Vector& x, Vector& y)
  {
    x.scatter_fwd();
    check(bd_csr_apply(_A, x.array(), y.array(), nullptr));
  }
  std::int64_t nnz() const
  {
    std::int64_t n = 0;
    check(bd_csr_info(_A, nullptr, nullptr, &n));
    return n;
  }

private:
  bd_csr* _A = nullptr;
  bd_halo* _halo;
};
} // namespace benchdolfinx
