"""numpy restatement of the dof-space form of the stiffness action on meshes of parallelepipeds.

TEST INFRASTRUCTURE (like everything under oracle/): imported by tests/ only.  This is NOT the reference's algorithm
but the re-association of it that the GPU kernels csrc/stiffness_affine*.cuh execute; it is
kept here so that the CPU test suite pins the algebra against the restatement of the reference
(oracle/laplacian.py, src/laplacian_gpu.hpp:175-420) without a GPU.

On an affine cell the geometry factor of the reference (src/geometry_gpu.hpp:113-130) is
G(q) = w_q0 w_q1 w_q2 Gc with a constant symmetric 3x3 matrix Gc = K K^T / det J, K = adj(J), so the quadrature
sums factor direction by direction into three nd x nd matrices
    M = Phi^T W Phi,   S = DPhi^T W DPhi,   C = DPhi^T W Phi        (DPhi = dphi1 @ phi0, W = diag of the 1-D weights)
    y = Gc00 S.M.M u + Gc11 M.S.M u + Gc22 M.M.S u
      + Gc01 (C.C^T.M + C^T.C.M) u + Gc02 (C.M.C^T + C^T.M.C) u + Gc12 (M.C.C^T + M.C^T.C) u
(a.b.c = a acting along x, b along y, c along z).  On bricks (mutually orthogonal edges) Gc is diagonal and only the
first line is left.
"""
from __future__ import annotations

import numpy as np


def matrices(tables):
    """M, S, C from the operator's own 1-D tables (fill_affine_tables, csrc/stiffness_launch.cuh)."""
    phi = tables.phi0.astype(np.float64)
    dphi = tables.dphi1.astype(np.float64) @ phi
    w = tables.qwts.astype(np.float64)
    M = phi.T @ (w[:, None] * phi)
    S = dphi.T @ (w[:, None] * dphi)
    C = dphi.T @ (w[:, None] * phi)
    return M, S, C


def cell_G(xgeom, geometry_dofmap, cells):
    """Gc[c, 3, 3] = K K^T / det J from the three edge vectors of each cell (vertices 0, 4, 2, 1 of the
    tensor-product order), otf_cell_g in csrc/stiffness_slab_otf.cuh; same expressions as src/geometry_gpu.hpp:100-130."""
    X = xgeom[geometry_dofmap[cells]].astype(np.float64)           # [nc, 8, 3]
    J = np.stack([X[:, 4] - X[:, 0], X[:, 2] - X[:, 0], X[:, 1] - X[:, 0]], axis=2)   # columns a, b, c
    K = np.linalg.inv(J) * np.linalg.det(J)[:, None, None]         # adj(J)
    # The reference's own "detJ" (src/geometry_gpu.hpp:110, src/geometry_cpu.hpp:89): the middle term enters with a
    # minus sign although K[1][0] already carries the cofactor's sign.  It equals det J whenever J[0][1] K[1][0] = 0 --
    # every mesh the reference builds (only x coordinates are ever perturbed, src/mesh.cpp:199-206) -- and differs on
    # sheared or rotated cells.  Kernels and oracle keep the reference's expression: parity is with the reference.
    det = J[:, 0, 0] * K[:, 0, 0] - J[:, 0, 1] * K[:, 1, 0] + J[:, 0, 2] * K[:, 2, 0]
    return np.einsum("cik,cjk->cij", K, K) / det[:, None, None]


def is_affine(xgeom, geometry_dofmap, cells, eps=8 * np.finfo(np.float64).eps):
    """The kernels' criterion (TrilinearMap::affine): mixed coefficients below eps * max|x| of the cell."""
    X = xgeom[geometry_dofmap[cells]].astype(np.float64)
    c110 = (X[:, 6] - X[:, 4]) - (X[:, 2] - X[:, 0])
    c101 = (X[:, 5] - X[:, 4]) - (X[:, 1] - X[:, 0])
    c011 = (X[:, 3] - X[:, 2]) - (X[:, 1] - X[:, 0])
    c111 = ((X[:, 7] - X[:, 6]) - (X[:, 5] - X[:, 4])) - ((X[:, 3] - X[:, 2]) - (X[:, 1] - X[:, 0]))
    mixed = np.max(np.abs(np.stack([c110, c101, c011, c111], axis=1)), axis=(1, 2))
    return mixed <= eps * np.max(np.abs(X), axis=(1, 2))


def is_brick(Gc, eps=8 * np.finfo(np.float64).eps):
    """count_nonaffine_kernel's second criterion: off-diagonal entries of Gc at rounding level."""
    diag = np.max(np.abs(np.stack([Gc[:, 0, 0], Gc[:, 1, 1], Gc[:, 2, 2]], axis=1)), axis=1)
    off = np.max(np.abs(np.stack([Gc[:, 0, 1], Gc[:, 0, 2], Gc[:, 1, 2]], axis=1)), axis=1)
    return off <= eps * diag


def apply(u, mesh, tables, kappa=2.0, diagonal_only=False):
    """y = A u over [lcells, bcells] in the dof-space form; BC handling as the reference
    (src/laplacian_gpu.hpp:163-169, 424-425)."""
    nd = tables.nd
    cells = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int64)
    M, S, C = matrices(tables)
    Gc = kappa * cell_G(mesh.xgeom, mesh.geometry_dofmap, cells)
    dm = mesh.cell_dofmap[cells]
    bc = mesh.bc_marker[dm] != 0
    y = np.zeros_like(u, dtype=np.float64)
    y[dm[bc]] = u[dm[bc]]
    ul = np.where(bc, 0, u[dm]).reshape(-1, nd, nd, nd)

    def tp(a, b, c, v):
        return np.einsum("ai,bj,ck,cnijk->cnabk" if False else "ai,bj,dk,nijk->nabd", a, b, c, v)

    g = lambda i, j: Gc[:, i, j][:, None, None, None]
    r = g(0, 0) * tp(S, M, M, ul) + g(1, 1) * tp(M, S, M, ul) + g(2, 2) * tp(M, M, S, ul)
    if not diagonal_only:
        r = r + g(0, 1) * (tp(C, C.T, M, ul) + tp(C.T, C, M, ul))
        r = r + g(0, 2) * (tp(C, M, C.T, ul) + tp(C.T, M, C, ul))
        r = r + g(1, 2) * (tp(M, C, C.T, ul) + tp(M, C.T, C, ul))
    r = r.reshape(len(cells), -1)
    np.add.at(y, dm[~bc], r[~bc])
    return y
