"""--mat_comp on the GPU: matrix-free operator vs the device-assembled CSR operator
(src/laplacian_solver.cpp:151-227) and the CSR operator vs the oracle's scipy assembly."""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import assemble_csr, assemble_rhs
from oracle.tables import Tables

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("degree,qmode,perturb,dtype,tol", [
    (3, 0, 0.0, np.float64, 1e-13), (3, 1, 0.0, np.float64, 1e-13), (2, 1, 0.2, np.float64, 1e-13),
    (1, 1, 0.2, np.float64, 1e-13), (5, 1, 0.1, np.float64, 1e-13), (6, 1, 0.0, np.float64, 1e-13),
    (3, 1, 0.2, np.float32, 1e-5)])
def test_csr_operator_matches_matrix_free_and_oracle(degree, qmode, perturb, dtype, tol):
    import torch
    import benchmark_dolfinx_b200 as bd
    mesh = create_box_mesh(compute_mesh_size(3000, degree), degree, geom_perturb_fact=perturb)
    t = Tables(degree, qmode)
    u = assemble_rhs(mesh, t) + np.random.default_rng(3).uniform(-1, 1, mesh.ndofs_local)
    z_ref = assemble_csr(mesh, t) @ u
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    A = bd.MatrixOperator(op)
    assert A.shape == (mesh.ndofs_local, mesh.ndofs_local) and A.nnz > 0
    ud = torch.from_numpy(u.astype(dtype)).cuda()
    y, z = torch.zeros_like(ud), torch.full_like(ud, 3.0)
    op.apply(ud, y)
    A.apply(ud, z)
    torch.cuda.synchronize()
    zh, yh = z.cpu().numpy().astype(np.float64), y.cpu().numpy().astype(np.float64)
    assert np.linalg.norm(zh - z_ref) <= tol * np.linalg.norm(z_ref)
    assert np.linalg.norm(zh - yh) <= tol * np.linalg.norm(zh)     # "Relative norm of error"


def test_mat_comp_baseline_config_1():
    """BASELINE config 1: --mat_comp --ndofs_global=10000 --degree=3 fp64 (matrix-free vs CSR),
    plus the CG variant of the same check."""
    import benchmark_dolfinx_b200 as bd
    for qmode, ynorm in ((0, 1.1590666300856363), (1, 1.1415775083651323)):
        mesh = bd.BoxMesh(bd.compute_mesh_size(10000, 3), 3)
        res, *_ = bd.laplace_action_gpu(mesh, 3, qmode, 2.0, 1, matrix_comparison=True)
        assert abs(res.ynorm - ynorm) <= 1e-12 * ynorm
        assert abs(res.znorm - res.ynorm) <= 1e-13 * res.ynorm
        assert res.enorm / res.znorm < 1e-13                      # README.md:155-156 expects ~1e-15
    mesh = bd.BoxMesh(bd.compute_mesh_size(10000, 3), 3)
    res, *_ = bd.laplace_action_gpu(mesh, 3, 1, 2.0, 10, use_cg=True, matrix_comparison=True)
    assert abs(res.ynorm - 129.5221707929043) <= 1e-11 * res.ynorm
    assert res.enorm / res.znorm < 1e-11
