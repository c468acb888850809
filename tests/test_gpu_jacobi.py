"""Jacobi-preconditioned CG: matrix-free diagonal against the diagonal of the oracle's assembled matrix
(what the reference's compute_diag_inv reads, src/csr.hpp:79-107), pointwise_mult (src/vector.hpp:266-285) and
the preconditioned solve against the oracle's restatement of "p <- beta*p + M^-1(r)" (src/cg.hpp:165)."""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian, assemble_csr, assemble_rhs, cg_solve, pcg_solve_jacobi
from oracle.tables import Tables as OracleTables

pytestmark = pytest.mark.gpu
RTOL = {np.float64: 1e-12, np.float32: 1e-5}


def _rel(a, b):
    return np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / np.linalg.norm(b.astype(np.float64))


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("degree,qmode,perturb", [(1, 1, 0.2), (2, 0, 0.2), (3, 1, 0.2), (3, 0, 0.0), (4, 1, 0.0),
                                                  (6, 1, 0.2), (7, 1, 0.2)])
def test_matrix_free_diagonal_matches_assembled_matrix(degree, qmode, perturb, dtype):
    import benchmark_dolfinx_b200 as bd
    mesh = create_box_mesh(compute_mesh_size(1500, degree), degree, geom_perturb_fact=perturb)
    t = OracleTables(degree, qmode)
    A = assemble_csr(mesh, t, kappa=2.0)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    dinv = op.get_diag_inverse().cpu().numpy()
    assert _rel(dinv, 1.0 / A.diagonal()) < RTOL[dtype]
    assert np.all(dinv[mesh.bc_marker != 0] == 1.0)


def test_pointwise_mult():
    import torch
    import benchmark_dolfinx_b200 as bd
    for dt in (torch.float64, torch.float32):
        x = torch.rand(100_003, dtype=dt, device="cuda")
        y = torch.rand(100_003, dtype=dt, device="cuda")
        w = torch.zeros_like(x)
        bd.pointwise_mult(w, x, y, 90_000)
        torch.cuda.synchronize()
        assert torch.equal(w[:90_000], x[:90_000] * y[:90_000]) and float(w[90_000:].abs().max()) == 0.0


@pytest.mark.parametrize("geometry", ["precomputed", "onthefly"])
@pytest.mark.parametrize("degree,qmode", [(3, 1), (2, 0), (6, 1)])
def test_jacobi_cg_matches_oracle(degree, qmode, geometry):
    import torch
    import benchmark_dolfinx_b200 as bd
    mesh = create_box_mesh(compute_mesh_size(6000, degree), degree, geom_perturb_fact=0.2)
    t = OracleTables(degree, qmode)
    oracle = OracleLaplacian(mesh, t, kappa=2.0, dtype=np.float64)
    b = assemble_rhs(mesh, t)
    dinv_ref = 1.0 / assemble_csr(mesh, t, kappa=2.0).diagonal()
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, geometry=geometry)
    dinv = op.get_diag_inverse()
    bd_ = torch.from_numpy(b).cuda()
    # fixed iteration count
    x_ref = np.zeros_like(b)
    its_ref = pcg_solve_jacobi(oracle.apply, x_ref, b, dinv_ref, 12)
    x = torch.zeros_like(bd_)
    its = bd.cg_solve(op, x, bd_, 12, diag_inv=dinv)
    torch.cuda.synchronize()
    assert its == its_ref == 12 and _rel(x.cpu().numpy(), x_ref) < 1e-11
    # tolerance stop: same iteration count as the oracle, and no more iterations than plain CG needs
    x_ref = np.zeros_like(b)
    its_ref = pcg_solve_jacobi(oracle.apply, x_ref, b, dinv_ref, 500, rtol=1e-8)
    x = torch.zeros_like(bd_)
    its = bd.cg_solve(op, x, bd_, 500, 1e-8, diag_inv=dinv)
    torch.cuda.synchronize()
    assert its == its_ref and _rel(x.cpu().numpy(), x_ref) < 1e-8
    its_plain = cg_solve(oracle.apply, np.zeros_like(b), b, 500, rtol=1e-8)
    assert its <= its_plain
