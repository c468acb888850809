"""Parity of the CUDA operator (through the C ABI) with the oracle on seeded inputs."""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian, assemble_rhs
from oracle.tables import Tables as OracleTables

pytestmark = pytest.mark.gpu

RTOL = {np.float64: 1e-12, np.float32: 1e-5}   # BASELINE.json north_star tolerances


def _rel(a, b):
    return np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / np.linalg.norm(b.astype(np.float64))


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("qmode", [0, 1])
@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5, 6, 7])
@pytest.mark.parametrize("perturb", [0.0, 0.2])
def test_apply_matches_oracle(degree, qmode, dtype, perturb):
    import torch
    import benchmark_dolfinx_b200 as bd
    n = compute_mesh_size(4000, degree)
    mesh = create_box_mesh(n, degree, geom_perturb_fact=perturb)
    ot = OracleTables(degree, qmode)
    oracle = OracleLaplacian(mesh, ot, kappa=2.0, dtype=np.float64)
    rng = np.random.default_rng(1234)
    u = rng.uniform(-1, 1, mesh.ndofs_local)
    y_ref = oracle.apply(u.copy())

    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    G = op.geometry()
    assert _rel(G, oracle.G) < RTOL[dtype]
    ud = torch.from_numpy(u.astype(dtype)).cuda()
    yd = torch.full_like(ud, 7.0)     # apply must overwrite, not accumulate
    op.apply(ud, yd)
    torch.cuda.synchronize()
    assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]
    # host-buffer entry point
    yh = np.zeros(mesh.ndofs_local, dtype=dtype)
    op.apply_host(u.astype(dtype), yh)
    assert _rel(yh, y_ref) < RTOL[dtype]


def test_reference_goldens():
    """src/test_output.py:19 and examples/mat_comp-16.json through the CUDA path."""
    import torch
    import benchmark_dolfinx_b200 as bd
    for ndofs, degree, qmode, unorm, ynorm in [
            (1000, 3, 0, None, 9.912865833415553),
            (100000, 3, 1, 0.6895773850559623, 0.14150257625641838)]:
        n = compute_mesh_size(ndofs, degree)
        mesh = create_box_mesh(n, degree)
        u = assemble_rhs(mesh, OracleTables(degree, qmode))
        op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
        ud = torch.from_numpy(u).cuda()
        yd = torch.zeros_like(ud)
        op.apply(ud, yd)
        assert abs(bd.norm(yd) - ynorm) / ynorm < 1e-12
        if unorm is not None:
            assert abs(bd.norm(ud) - unorm) / unorm < 1e-12


def test_host_batch_entry_point_matches_single_calls():
    """bd_laplacian_apply_host_batch (pipelined copies) == bd_laplacian_apply_host, vector by vector."""
    import benchmark_dolfinx_b200 as bd
    degree, qmode = 3, 1
    mesh = create_box_mesh(compute_mesh_size(30000, degree), degree, geom_perturb_fact=0.2)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0)
    rng = np.random.default_rng(5)
    ins = [rng.uniform(-1, 1, mesh.ndofs_local) for _ in range(5)]
    outs = [np.zeros(mesh.ndofs_local) for _ in range(5)]
    op.apply_host_batch(ins, outs)
    for a, b in zip(ins, outs):
        ref = np.zeros_like(a)
        op.apply_host(a, ref)
        assert _rel(b, ref) < 1e-14
    op.apply_host_batch([], [])
