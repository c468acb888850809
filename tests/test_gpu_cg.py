"""cg_solve through the C ABI vs the oracle restatement of src/cg.hpp:89-169."""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian, assemble_csr, assemble_rhs, cg_solve
from oracle.tables import Tables

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("ndofs,degree,qmode,its,perturb", [
    (1000, 3, 0, 10, 0.0), (10000, 3, 1, 50, 0.0), (5000, 6, 1, 20, 0.0), (3000, 2, 1, 30, 0.2),
    (3000, 1, 0, 25, 0.2), (4000, 4, 1, 15, 0.1), (4000, 5, 0, 12, 0.0), (4000, 7, 1, 8, 0.2),
])
def test_cg_matches_oracle_fp64(ndofs, degree, qmode, its, perturb):
    import torch
    import benchmark_dolfinx_b200 as bd
    n = compute_mesh_size(ndofs, degree)
    mesh = create_box_mesh(n, degree, geom_perturb_fact=perturb)
    t = Tables(degree, qmode)
    oracle = OracleLaplacian(mesh, t)
    b = assemble_rhs(mesh, t)
    x_ref = np.zeros_like(b)
    k_ref = cg_solve(oracle.apply, x_ref, b, its)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
    bd_ = torch.from_numpy(b).cuda()
    x = torch.zeros_like(bd_)
    k = bd.cg_solve(op, x, bd_, its)
    torch.cuda.synchronize()
    assert k == k_ref == its                       # rtol = 0: exactly max_iter iterations
    assert _rel(x.cpu().numpy(), x_ref) < 1e-12    # north_star fp64 tolerance


@pytest.mark.parametrize("args,ynorm", [
    ((1000, 3, 0, 10), 46.17746829199996), ((10000, 3, 1, 10), 129.5221707929043),
    ((10000, 3, 1, 50), 167.59244725655134), ((5000, 6, 1, 20), 90.15797406544854)])
def test_cg_known_answers(args, ynorm):
    import torch
    import benchmark_dolfinx_b200 as bd
    ndofs, degree, qmode, its = args
    n = bd.compute_mesh_size(ndofs, degree)
    mesh = bd.BoxMesh(n, degree)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
    u = mesh.assemble_rhs(qmode)
    y = torch.zeros_like(u)
    assert bd.cg_solve(op, y, u, its) == its
    assert abs(bd.norm(y) - ynorm) <= 1e-11 * ynorm


def test_cg_fp32():
    import torch
    import benchmark_dolfinx_b200 as bd
    degree, qmode, its = 3, 1, 12
    mesh = create_box_mesh(compute_mesh_size(8000, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    oracle = OracleLaplacian(mesh, t)
    b = assemble_rhs(mesh, t)
    x_ref = np.zeros_like(b)
    cg_solve(oracle.apply, x_ref, b, its)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=np.float32)
    x = torch.zeros(len(b), dtype=torch.float32, device="cuda")
    assert bd.cg_solve(op, x, torch.from_numpy(b.astype(np.float32)).cuda(), its) == its
    assert _rel(x.cpu().numpy(), x_ref) < 1e-5     # north_star fp32 tolerance


def test_cg_rtol_stops_at_the_same_iteration():
    """src/cg.hpp:162-163: break when rnorm/rnorm0 < rtol^2."""
    import torch
    import benchmark_dolfinx_b200 as bd
    degree, qmode = 2, 1
    mesh = create_box_mesh(compute_mesh_size(3000, degree), degree)
    t = Tables(degree, qmode)
    oracle = OracleLaplacian(mesh, t)
    b = assemble_rhs(mesh, t)
    for rtol in (1e-2, 1e-6):
        x_ref = np.zeros_like(b)
        k_ref = cg_solve(oracle.apply, x_ref, b, 500, rtol=rtol)
        op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
        x = torch.zeros(len(b), dtype=torch.float64, device="cuda")
        k = bd.cg_solve(op, x, torch.from_numpy(b).cuda(), 500, rtol=rtol)
        assert k == k_ref and k < 500
        assert _rel(x.cpu().numpy(), x_ref) < 1e-10


def test_mat_comp_check():
    """BASELINE config 1: --mat_comp --ndofs_global=10000 --degree=3: matrix-free (GPU) vs the
    assembled CSR operator (src/laplacian_solver.cpp:151-227); README.md:155-156 expects ~1e-15."""
    import torch
    import benchmark_dolfinx_b200 as bd
    degree = 3
    for qmode in (0, 1):
        mesh = create_box_mesh(compute_mesh_size(10000, degree), degree)
        t = Tables(degree, qmode)
        u = assemble_rhs(mesh, t)
        z = assemble_csr(mesh, t) @ u
        op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
        ud = torch.from_numpy(u).cuda()
        yd = torch.zeros_like(ud)
        op.apply(ud, yd)
        y = yd.cpu().numpy()
        enorm, znorm = np.linalg.norm(y - z), np.linalg.norm(z)
        assert enorm / znorm < 1e-13
        assert np.isclose(np.linalg.norm(y), znorm, rtol=1e-13)
