// Instantiates the adapter for both scalar types against the mock-ups (syntax / type check only).
#include <dolfinx_adapter.hpp>

template <typename T>
struct DeviceArrayMock
{
  struct Ptr
  {
    T* get() const;
  };
  Ptr data() const;
};
template <typename T>
struct VectorMock
{
  using value_type = T;
  DeviceArrayMock<T>& array();
  const DeviceArrayMock<T>& array() const;
};

template <typename T>
int use(const dolfinx::fem::FunctionSpace<T>& V, const dolfinx::fem::DirichletBC<T>& bc, VectorMock<T>& x,
        VectorMock<T>& y)
{
  benchdolfinx::b200::MatFreeLaplacianGPU<T> op(V, bc, 3, 1, T(2), basix::quadrature::type::gll);
  op.apply(x, y);
  return benchdolfinx::b200::cg_solve(op, y, x, 10, T(0));
}
template int use<double>(const dolfinx::fem::FunctionSpace<double>&, const dolfinx::fem::DirichletBC<double>&,
                         VectorMock<double>&, VectorMock<double>&);
template int use<float>(const dolfinx::fem::FunctionSpace<float>&, const dolfinx::fem::DirichletBC<float>&,
                        VectorMock<float>&, VectorMock<float>&);
