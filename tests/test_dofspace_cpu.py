"""The dof-space form of the action that the GPU kernels for parallelepipeds / bricks execute (oracle/dofspace.py,
csrc/stiffness_affine*.cuh) against the restatement of the reference's quadrature loop (oracle/laplacian.py), on CPU."""
import numpy as np
import pytest

from oracle import dofspace
from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian
from oracle.tables import Tables


def _rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


class _Mesh:
    def __init__(self, mesh, xgeom):
        for k in ("cell_dofmap", "geometry_dofmap", "bc_marker", "lcells", "bcells", "ndofs_local", "ncells"):
            setattr(self, k, getattr(mesh, k))
        self.xgeom = xgeom


def _mesh(degree, kind):
    m = create_box_mesh(compute_mesh_size(1500, degree), degree)
    x = m.xgeom.copy()
    if kind == "sheared":
        x = x @ np.array([[1.0, 0.3, -0.2], [0.1, 0.9, 0.25], [-0.15, 0.2, 1.1]]).T + np.array([0.3, -0.1, 0.2])
    if kind == "rotated":
        a, b = 0.7, -0.4
        Rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
        Rx = np.array([[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]])
        x = x @ (Rz @ Rx).T
    return _Mesh(m, x)


@pytest.mark.parametrize("kind", ["box", "rotated", "sheared"])
@pytest.mark.parametrize("degree,qmode", [(1, 1), (2, 0), (3, 1), (4, 1), (6, 1)])
def test_dofspace_form_equals_quadrature_loop(degree, qmode, kind):
    m = _mesh(degree, kind)
    t = Tables(degree, qmode)
    u = np.random.default_rng(degree).uniform(-1, 1, m.ndofs_local)
    y_ref = OracleLaplacian(m, t, kappa=2.0).apply(u.copy())
    cells = np.concatenate([m.lcells, m.bcells])
    assert dofspace.is_affine(m.xgeom, m.geometry_dofmap, cells).all()
    brick = dofspace.is_brick(dofspace.cell_G(m.xgeom, m.geometry_dofmap, cells)).all()
    # exact on the axis-aligned box; a rotated box carries rounding of the size of its coordinates (not of its edges)
    # in every vertex, which the 8-eps criterion may or may not let through: both kernels are correct for it
    assert brick if kind == "box" else (not brick if kind == "sheared" else True)
    assert _rel(dofspace.apply(u, m, t, 2.0), y_ref) < 1e-13
    # bricks: the three mixed terms vanish (to rounding on the rotated box); on sheared cells they do not
    err_diag = _rel(dofspace.apply(u, m, t, 2.0, diagonal_only=True), y_ref)
    if brick or kind == "rotated":
        assert err_diag < 1e-12
    else:
        assert err_diag > 1e-3


def test_perturbed_cells_are_not_affine():
    m = create_box_mesh(compute_mesh_size(1500, 2), 2, geom_perturb_fact=0.2)
    cells = np.concatenate([m.lcells, m.bcells])
    assert not dofspace.is_affine(m.xgeom, m.geometry_dofmap, cells).all()
