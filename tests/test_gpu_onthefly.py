"""Geometry on the fly (csrc/stiffness_slab_otf.cuh, csrc/stiffness_mma8_otf.cuh) against the oracle.

The reference always precomputes G (geometry_computation_gpu, src/geometry_gpu.hpp:89-130) and streams it
(src/laplacian_gpu.hpp:256-273).  The on-the-fly variants rebuild it in the stiffness kernel from the cell
vertices: the general trilinear path and the shortcut for parallelepipeds (G(q) = w_q G_cell).  Same oracle, same
tolerances as every other operator test; grids capped so that each warp also runs its steady state.
"""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian
from oracle.tables import Tables as OracleTables

pytestmark = pytest.mark.gpu

RTOL = {np.float64: 1e-12, np.float32: 1e-5}


def _rel(a, b):
    return np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / np.linalg.norm(b.astype(np.float64))


class _Mesh:
    """Oracle mesh with replaced vertex coordinates."""

    def __init__(self, mesh, xgeom):
        for k in ("cell_dofmap", "geometry_dofmap", "bc_marker", "lcells", "bcells", "ndofs_local", "ncells"):
            setattr(self, k, getattr(mesh, k))
        self.ndofs_owned = mesh.ndofs_local
        self.xgeom = xgeom


def _case(degree, qmode, kind):
    """kind: box (axis-aligned bricks; box-full: the same mesh, diagonal-G shortcut switched off), rotated (the box
    under a rotation: bricks only to rounding), sheared (affine map of the box: parallelepipeds), perturbed (general
    hexahedra), mixed (a few perturbed vertices in an otherwise sheared mesh)."""
    mesh = create_box_mesh(compute_mesh_size(12000, degree), degree,
                           geom_perturb_fact=0.2 if kind == "perturbed" else 0.0)
    x = mesh.xgeom.copy()
    if kind in ("sheared", "mixed"):
        A = np.array([[1.0, 0.3, -0.2], [0.1, 0.9, 0.25], [-0.15, 0.2, 1.1]])
        x = x @ A.T + np.array([0.3, -0.1, 0.2])
    if kind == "rotated":
        a, b = 0.7, -0.4
        Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        Rx = np.array([[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]])
        x = x @ (Rz @ Rx).T + np.array([0.3, -0.1, 0.2])
    if kind == "mixed":
        rng = np.random.default_rng(7)
        h = 1.0 / max(compute_mesh_size(12000, degree))
        pick = rng.choice(len(x), size=max(1, len(x) // 15), replace=False)
        x[pick] += rng.uniform(-0.15 * h, 0.15 * h, (len(pick), 3))
    m = _Mesh(mesh, x)
    oracle = OracleLaplacian(m, OracleTables(degree, qmode), kappa=2.0, dtype=np.float64)
    u = np.random.default_rng(10 * degree + qmode).uniform(-1, 1, mesh.ndofs_local)
    return m, u, oracle


_CACHE = {}


def _cached(degree, qmode, kind):
    kind = "box" if kind == "box-full" else kind
    key = (degree, qmode, kind)
    if key not in _CACHE:
        m, u, oracle = _case(degree, qmode, kind)
        _CACHE[key] = (m, u, oracle.apply(u.copy()), oracle.G)
    return _CACHE[key]


@pytest.fixture
def grid_cap():
    import benchmark_dolfinx_b200 as bd
    yield bd.debug_set_max_grid
    bd.debug_set_max_grid(0)


_PQ = [(p, q) for p in range(1, 8) for q in (0, 1)]


@pytest.mark.parametrize("kind", ["box", "box-full", "rotated", "sheared", "perturbed", "mixed"])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("degree,qmode", _PQ)
def test_onthefly_matches_oracle(degree, qmode, dtype, kind, grid_cap):
    import torch
    import benchmark_dolfinx_b200 as bd
    m, u, y_ref, G_ref = _cached(degree, qmode, kind)
    if degree + 1 + qmode == 9 and kind in ("perturbed", "mixed"):
        # nq = 9 has a kernel for parallelepiped cells only (the dof-space one): general cells must be refused
        with pytest.raises(bd.BenchDolfinxError):
            bd.MatFreeLaplacianGPU(m, degree, qmode, 2.0, dtype=dtype, geometry="onthefly")
        return
    op = bd.MatFreeLaplacianGPU(m, degree, qmode, 2.0, dtype=dtype,
                                geometry="onthefly-full" if kind == "box-full" else "onthefly")
    mode, nonaffine = op.geometry_mode()
    if kind == "box":
        # bricks: the dof-space kernels (qmode 1) use the diagonal of G_cell only
        assert (mode, nonaffine) == ("onthefly-brick" if qmode == 1 else "onthefly-affine", 0)
        assert ("bricks" in op.describe()) == (qmode == 1)
    if kind in ("box-full", "sheared") and dtype == np.float64:
        assert (mode, nonaffine) == ("onthefly-affine", 0)
    if kind == "rotated" and dtype == np.float64:
        assert mode in ("onthefly-brick", "onthefly-affine") and nonaffine == 0
    if kind in ("perturbed", "mixed"):
        assert mode == "onthefly-general" and nonaffine > 0
    assert "on the fly" in op.describe()
    ud = torch.from_numpy(u.astype(dtype)).cuda()
    for cap in (2, 0):
        grid_cap(cap)
        for _ in range(2):
            yd = torch.full_like(ud, 7.0)
            op.apply(ud, yd)
            torch.cuda.synchronize()
            assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]
    # G is not stored; on request it is computed in the reference's layout
    assert _rel(op.geometry(), G_ref) < RTOL[dtype]


def test_default_choice_and_forced_modes():
    import torch
    import benchmark_dolfinx_b200 as bd
    for kind, expect in (("box", "onthefly-brick"), ("sheared", "onthefly-affine"), ("perturbed", "precomputed")):
        m, u, y_ref, _ = _cached(3, 1, kind)
        op = bd.MatFreeLaplacianGPU(m, 3, 1, 2.0)
        assert op.geometry_mode()[0] == expect
        if kind == "box":
            assert "affine_pipe" in op.describe()   # nd = 4 bricks: asynchronous gather, merged faces
        pre = bd.MatFreeLaplacianGPU(m, 3, 1, 2.0, geometry="precomputed")
        assert pre.geometry_mode()[0] == "precomputed"
        ud = torch.from_numpy(u).cuda()
        y1, y2 = torch.zeros_like(ud), torch.zeros_like(ud)
        op.apply(ud, y1)
        pre.apply(ud, y2)
        torch.cuda.synchronize()
        assert _rel(y1.cpu().numpy(), y_ref) < 1e-12 and _rel(y2.cpu().numpy(), y_ref) < 1e-12
    # every degree has a dof-space kernel for parallelepipeds at qmode 1 (nq = 9 included) ...
    m7 = create_box_mesh(compute_mesh_size(3000, 7), 7)
    op7 = bd.MatFreeLaplacianGPU(m7, 7, 1, 2.0)
    assert op7.geometry_mode()[0] == "onthefly-brick" and "affine_col" in op7.describe()
    assert bd.MatFreeLaplacianGPU(m7, 7, 1, 2.0, geometry="auto-full").geometry_mode()[0] == "onthefly-affine"
    # ... while qmode 0 keeps the quadrature-space kernels (Phi = I makes them the cheaper ones): DMMA at nq = 7
    m6 = create_box_mesh(compute_mesh_size(3000, 6), 6)
    assert "affine_col" in bd.MatFreeLaplacianGPU(m6, 6, 1, 2.0).describe()
    assert "affine_col" in bd.MatFreeLaplacianGPU(m6, 6, 1, 2.0, dtype=np.float32).describe()
    assert "mma8_otf" in bd.MatFreeLaplacianGPU(m6, 6, 0, 2.0).describe()


def test_onthefly_cg_and_csr_check():
    """CG through the on-the-fly operator equals the oracle's CG; the assembled-CSR check still works (G on demand)."""
    import torch
    import benchmark_dolfinx_b200 as bd
    from oracle.laplacian import cg_solve
    m, u, _, _ = _cached(3, 1, "sheared")
    oracle = OracleLaplacian(m, OracleTables(3, 1), kappa=2.0, dtype=np.float64)
    op = bd.MatFreeLaplacianGPU(m, 3, 1, 2.0)
    assert op.geometry_mode()[0] == "onthefly-affine"
    b = u.copy()
    b[m.bc_marker != 0] = 0
    x_ref = np.zeros_like(b)
    its_ref = cg_solve(oracle.apply, x_ref, b, 8)
    bd_, x = torch.from_numpy(b).cuda(), torch.zeros(len(b), dtype=torch.float64, device="cuda")
    its = bd.cg_solve(op, x, bd_, 8)
    torch.cuda.synchronize()
    assert its == its_ref and _rel(x.cpu().numpy(), x_ref) < 1e-12
    A = bd.MatrixOperator(op)
    y, z = torch.zeros_like(bd_), torch.zeros_like(bd_)
    op.apply(bd_, y)
    A.apply(bd_, z)
    torch.cuda.synchronize()
    assert _rel(y.cpu().numpy(), z.cpu().numpy()) < 1e-12


@pytest.mark.parametrize("order", ["reversed", "shuffled", "split"])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_brick_kernel_with_reordered_cell_lists(order, dtype, grid_cap):
    """The brick kernel at nd = 4 merges the faces that consecutive cells of the list share (found at construction from
    the dofmap).  Lists in which consecutive cells are not neighbours, are neighbours the other way round, or are split
    over the two launches (lcells / bcells) must give the same result: the action does not depend on the list order."""
    import torch
    import benchmark_dolfinx_b200 as bd
    m, u, y_ref, _ = _cached(3, 1, "box")
    cells = np.concatenate([m.lcells, m.bcells]).astype(np.int32)
    rng = np.random.default_rng(3)
    if order == "reversed":
        lc, bc = cells[::-1].copy(), cells[:0].copy()
    elif order == "shuffled":
        lc, bc = rng.permutation(cells).astype(np.int32), cells[:0].copy()
    else:  # a ragged cut in the middle of a z line, both parts in mesh order
        cut = len(cells) // 2 + 3
        lc, bc = cells[:cut].copy(), cells[cut:].copy()
    mm = _Mesh(m, m.xgeom)
    mm.lcells, mm.bcells = lc, bc
    op = bd.MatFreeLaplacianGPU(mm, 3, 1, 2.0, dtype=dtype)
    assert op.geometry_mode()[0] == "onthefly-brick" and "affine_pipe" in op.describe()
    ud = torch.from_numpy(u.astype(dtype)).cuda()
    for cap in (2, 0):
        grid_cap(cap)
        yd = torch.full_like(ud, 7.0)
        op.apply(ud, yd)
        torch.cuda.synchronize()
        assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]
