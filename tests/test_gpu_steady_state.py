"""Steady state of the persistent stiffness kernels against the oracle.

At the benchmark sizes every warp of the slab / DMMA / pencil kernels processes hundreds of batches: the G
ring runs on across batch boundaries, the prefetched dof indices / values rotate into place, the cell index
and kappa are loaded one and two batches ahead.  On a test mesh of a few thousand cells a full grid gives every
warp at most one batch, so that code never runs.  `bd_debug_set_max_grid` caps the grid: with 1 or 3 CTAs each
warp walks over many batches, including a ragged last one, and the result is compared with the numpy oracle
(reference math: src/laplacian_gpu.hpp:153-425) for every degree x qmode x scalar type.
"""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian
from oracle.tables import Tables as OracleTables

pytestmark = pytest.mark.gpu

RTOL = {np.float64: 1e-12, np.float32: 1e-5}   # BASELINE.json north_star tolerances


def _rel(a, b):
    return np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / np.linalg.norm(b.astype(np.float64))


@pytest.fixture
def grid_cap():
    import benchmark_dolfinx_b200 as bd
    yield bd.debug_set_max_grid
    bd.debug_set_max_grid(0)


_ORACLE_CACHE = {}


def _oracle_case(degree, qmode):
    key = (degree, qmode)
    if key not in _ORACLE_CACHE:
        # 20000 dofs: 64 (Q7) to 17000 (Q1) cells; odd cell counts leave the last batch ragged
        mesh = create_box_mesh(compute_mesh_size(20000, degree), degree, geom_perturb_fact=0.2)
        oracle = OracleLaplacian(mesh, OracleTables(degree, qmode), kappa=2.0, dtype=np.float64)
        u = np.random.default_rng(100 * degree + qmode).uniform(-1, 1, mesh.ndofs_local)
        _ORACLE_CACHE[key] = (mesh, u, oracle.apply(u.copy()))
    return _ORACLE_CACHE[key]


@pytest.mark.parametrize("cap", [1, 3])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("qmode", [0, 1])
@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5, 6, 7])
def test_capped_grid_matches_oracle(degree, qmode, dtype, cap, grid_cap):
    import torch
    import benchmark_dolfinx_b200 as bd
    mesh, u, y_ref = _oracle_case(degree, qmode)
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    ud = torch.from_numpy(u.astype(dtype)).cuda()
    grid_cap(cap)
    for _ in range(2):                     # second call: rings and mbarrier phases start from a used state
        yd = torch.full_like(ud, 7.0)
        op.apply(ud, yd)
        torch.cuda.synchronize()
        assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]
    grid_cap(0)
    yd = torch.full_like(ud, 7.0)
    op.apply(ud, yd)
    torch.cuda.synchronize()
    assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]


@pytest.mark.parametrize("cap", [2])
@pytest.mark.parametrize("degree,qmode", [(3, 1), (6, 1), (2, 0)])
def test_capped_grid_two_parts(degree, qmode, cap, grid_cap):
    """lcells and bcells are separate launches with separate G segments: split the list of a single-rank
    mesh artificially so that both launches run their steady state."""
    import torch
    import benchmark_dolfinx_b200 as bd
    mesh, u, y_ref = _oracle_case(degree, qmode)

    class Split:
        pass
    m = Split()
    for k in ("cell_dofmap", "xgeom", "geometry_dofmap", "bc_marker"):
        setattr(m, k, getattr(mesh, k))
    allc = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int32)
    cut = (2 * len(allc)) // 3 + 1
    m.lcells, m.bcells = allc[:cut], allc[cut:]
    op = bd.MatFreeLaplacianGPU(m, degree, qmode, 2.0)
    ud = torch.from_numpy(u).cuda()
    yd = torch.full_like(ud, -3.0)
    grid_cap(cap)
    op.apply(ud, yd)
    torch.cuda.synchronize()
    assert _rel(yd.cpu().numpy(), y_ref) < 1e-12
