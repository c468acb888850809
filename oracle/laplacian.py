"""numpy restatement of the reference's matrix-free Laplacian path (oracle).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and the
CPU-baseline legs of bench.py.  Every function cites the reference code it follows.
"""
from __future__ import annotations

import numpy as np


# --------------------------------------------------------------------------- geometry
def jacobian(xgeom, geometry_dofmap, dphi_geometry, cells):
    """J[c, q, i, j] = sum_k x_k[i] * dphi[j, q, k]   (src/geometry_gpu.hpp:89-97)."""
    coords = xgeom[geometry_dofmap[cells]]                 # [nc, 8, 3]
    return np.einsum("cki,jqk->cqij", coords, dphi_geometry)


def geometry_G(xgeom, geometry_dofmap, dphi_geometry, weights, cells):
    """G[c, 6, nq^3]: K = adj(J), G = K K^T w / detJ, components (00,10,20,11,21,22).

    Follows src/geometry_gpu.hpp:100-130 == src/geometry_cpu.hpp:79-109.
    """
    J = jacobian(xgeom, geometry_dofmap, dphi_geometry, cells)
    K = np.empty_like(J)
    K[..., 0, 0] = J[..., 1, 1] * J[..., 2, 2] - J[..., 1, 2] * J[..., 2, 1]
    K[..., 0, 1] = -J[..., 0, 1] * J[..., 2, 2] + J[..., 0, 2] * J[..., 2, 1]
    K[..., 0, 2] = J[..., 0, 1] * J[..., 1, 2] - J[..., 0, 2] * J[..., 1, 1]
    K[..., 1, 0] = -J[..., 1, 0] * J[..., 2, 2] + J[..., 1, 2] * J[..., 2, 0]
    K[..., 1, 1] = J[..., 0, 0] * J[..., 2, 2] - J[..., 0, 2] * J[..., 2, 0]
    K[..., 1, 2] = -J[..., 0, 0] * J[..., 1, 2] + J[..., 0, 2] * J[..., 1, 0]
    K[..., 2, 0] = J[..., 1, 0] * J[..., 2, 1] - J[..., 1, 1] * J[..., 2, 0]
    K[..., 2, 1] = -J[..., 0, 0] * J[..., 2, 1] + J[..., 0, 1] * J[..., 2, 0]
    K[..., 2, 2] = J[..., 0, 0] * J[..., 1, 1] - J[..., 0, 1] * J[..., 1, 0]
    detJ = J[..., 0, 0] * K[..., 0, 0] - J[..., 0, 1] * K[..., 1, 0] + J[..., 0, 2] * K[..., 2, 0]
    s = weights[None, :] / detJ
    pairs = [(0, 0), (1, 0), (2, 0), (1, 1), (2, 1), (2, 2)]
    G = np.empty((J.shape[0], 6, J.shape[1]), dtype=J.dtype)
    for n, (a, b) in enumerate(pairs):
        G[:, n, :] = (K[..., a, 0] * K[..., b, 0] + K[..., a, 1] * K[..., b, 1]
                      + K[..., a, 2] * K[..., b, 2]) * s
    return G, detJ


# --------------------------------------------------------------------------- operator
def stiffness_apply(u, y, G, kappa, cell_dofmap, cells, bc_marker, phi0, dphi1, is_identity):
    """y += A_cells u for one cell list; G indexed by list position.

    Follows src/laplacian_gpu.hpp:153-170 (gather, BC), :175-211 (interpolation),
    :239-251 (gradient), :256-273 (G, kappa), :307-314 (gradient^T), :388-420
    (interpolation^T), :424-425 (scatter); for qmode 0 identical to
    src/laplacian_cpu.hpp:77-145.
    """
    nq = dphi1.shape[0]
    nd = phi0.shape[1]
    dm = cell_dofmap[cells]                                  # [nc, nd^3]
    bc = bc_marker[dm] != 0
    # b[dof] = u[dof] for BC dofs of the visited cells
    y[dm[bc]] = u[dm[bc]]
    ul = np.where(bc, 0, u[dm]).reshape(-1, nd, nd, nd)
    if not is_identity:
        ul = np.einsum("qi,cijk->cqjk", phi0, ul)
        ul = np.einsum("qj,cijk->ciqk", phi0, ul)
        ul = np.einsum("qk,cijk->cijq", phi0, ul)
    gx = np.einsum("qi,cijk->cqjk", dphi1, ul)
    gy = np.einsum("qj,cijk->ciqk", dphi1, ul)
    gz = np.einsum("qk,cijk->cijq", dphi1, ul)
    Gc = G.reshape(len(cells), 6, nq, nq, nq)
    G0, G1, G2, G3, G4, G5 = (Gc[:, i] for i in range(6))
    fx = kappa * (G0 * gx + G1 * gy + G2 * gz)
    fy = kappa * (G1 * gx + G3 * gy + G4 * gz)
    fz = kappa * (G2 * gx + G4 * gy + G5 * gz)
    r = (np.einsum("qi,cqjk->cijk", dphi1, fx) + np.einsum("qj,ciqk->cijk", dphi1, fy)
         + np.einsum("qk,cijq->cijk", dphi1, fz))
    if not is_identity:
        r = np.einsum("qi,cqjk->cijk", phi0, r)
        r = np.einsum("qj,ciqk->cijk", phi0, r)
        r = np.einsum("qk,cijq->cijk", phi0, r)
    r = r.reshape(len(cells), -1)
    np.add.at(y, dm[~bc], r[~bc])
    return y


class OracleLaplacian:
    """CPU twin of MatFreeLaplacianGPU (src/laplacian.hpp:89-447) on stand-in arrays."""

    def __init__(self, mesh, tables, kappa=2.0, dtype=np.float64):
        self.mesh, self.t, self.kappa, self.dtype = mesh, tables, dtype(kappa), np.dtype(dtype)
        self.cells = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int32)
        self.xgeom = mesh.xgeom.astype(dtype)
        self.phi0 = tables.phi0.astype(dtype)
        self.dphi1 = tables.dphi1.astype(dtype)
        self.G, _ = geometry_G(self.xgeom, mesh.geometry_dofmap, tables.dphi_geometry.astype(dtype),
                               tables.weights3d.astype(dtype), self.cells)
        self.G = self.G.astype(dtype)

    def apply(self, u, y=None):
        """src/laplacian.hpp:355-403 with the halo already done by the caller."""
        if y is None:
            y = np.empty_like(u)
        y[:] = 0                                             # set_value(out, 0)  :359
        return stiffness_apply(u, y, self.G, self.kappa, self.mesh.cell_dofmap, self.cells,
                               self.mesh.bc_marker, self.phi0, self.dphi1, self.t.is_identity)


# --------------------------------------------------------------------------- RHS
def source_f(x, y):
    """src/main.cpp:81-92."""
    return 1000.0 * np.exp(-((x - 0.5) ** 2 + (y - 0.5) ** 2) / 0.02)


def assemble_rhs(mesh, tables, dtype=np.float64):
    """b = sum_cells Phi^T (w detJ Phi f), BC rows zeroed; ghost entries left 0.

    FFCx form inner(w0, v) * dx (src/poisson64.py:66) with the operator's
    quadrature (src/forms.cpp:142-156); bc.set (src/laplacian_solver.cpp:105).
    f is interpolated at the physical dof nodes (src/main.cpp:81-92).
    """
    t = tables
    nd, nq = t.nd, t.nq
    cells = np.arange(mesh.ncells)
    coords = mesh.xgeom[mesh.geometry_dofmap[cells]].reshape(-1, 2, 2, 2, 3)
    L = np.stack([1 - t.dof_nodes, t.dof_nodes], axis=1)     # [nd, 2]
    xn = np.einsum("ia,jb,kc,nabcx->nijkx", L, L, L, coords)  # physical dof-node coordinates
    f = source_f(xn[..., 0], xn[..., 1])
    _, detJ = geometry_G(mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry, t.weights3d, cells)
    phi = t.phi0.astype(np.float64)
    fq = np.einsum("qi,cijk->cqjk", phi, f)
    fq = np.einsum("qj,cijk->ciqk", phi, fq)
    fq = np.einsum("qk,cijk->cijq", phi, fq)
    fq = fq * (t.weights3d[None, :] * detJ).reshape(-1, nq, nq, nq)
    r = np.einsum("qi,cqjk->cijk", phi, fq)
    r = np.einsum("qj,ciqk->cijk", phi, r)
    r = np.einsum("qk,cijq->cijk", phi, r)
    b = np.zeros(mesh.ndofs_local, dtype=np.float64)
    np.add.at(b, mesh.cell_dofmap.ravel(), r.reshape(-1))
    b[mesh.bc_marker != 0] = 0
    b[mesh.ndofs_owned:] = 0
    return b.astype(dtype)


# --------------------------------------------------------------------------- assembled CSR
def assemble_csr(mesh, tables, kappa=2.0):
    """Assembled stiffness matrix with BC rows/cols dropped and unit BC diagonal.

    Restates assemble_matrix(a, {bc}) + set_diagonal(1.0)
    (src/laplacian_solver.cpp:182-184).  Single-rank meshes only.
    """
    import scipy.sparse as sp
    t = tables
    nd, nq = t.nd, t.nq
    n3 = nd ** 3
    cells = np.arange(mesh.ncells)
    G, _ = geometry_G(mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry, t.weights3d, cells)
    phi = t.phi0.astype(np.float64)
    dq = t.dphi1.astype(np.float64) @ phi                    # d/dx of dof basis at quad points
    # B[d][q, i] dense [nq^3, nd^3]
    def tp(a, b, c):
        return np.einsum("qi,rj,sk->qrsijk", a, b, c).reshape(nq ** 3, n3)
    B = [tp(dq, phi, phi), tp(phi, dq, phi), tp(phi, phi, dq)]
    idx = {(0, 0): 0, (1, 0): 1, (0, 1): 1, (2, 0): 2, (0, 2): 2, (1, 1): 3, (2, 1): 4, (1, 2): 4, (2, 2): 5}
    Ae = np.zeros((mesh.ncells, n3, n3))
    for d in range(3):
        for e in range(3):
            Ae += np.einsum("qi,cq,qj->cij", B[d], kappa * G[:, idx[(d, e)], :], B[e], optimize=True)
    dm = mesh.cell_dofmap
    rows = np.repeat(dm, n3, axis=1).ravel()
    cols = np.tile(dm, (1, n3)).ravel()
    vals = Ae.ravel().copy()
    bc = mesh.bc_marker != 0
    vals[bc[rows] | bc[cols]] = 0
    n = mesh.ndofs_local
    A = sp.coo_matrix((vals, (rows, cols)), shape=(n, n)).tocsr()
    A = A + sp.diags(bc.astype(np.float64))
    A.eliminate_zeros()
    return A.tocsr()


# --------------------------------------------------------------------------- CG
def cg_solve(apply, x, b, max_iter, rtol=0.0, n_owned=None, dot=None, scatter=None):
    """Unpreconditioned CG, src/cg.hpp:89-169 (same update order, rtol test and count).

    `apply(p, y)` computes y = A p; `dot(a, b)` is the global inner product over
    owned entries; `scatter(p)` refreshes ghost entries.
    """
    n = len(x) if n_owned is None else n_owned
    if dot is None:
        dot = lambda a, c: float(np.dot(a[:n], c[:n]))
    y = np.zeros_like(x)
    apply(x, y)
    r = np.zeros_like(x)
    r[:n] = -1 * y[:n] + b[:n]
    p = np.zeros_like(x)
    p[:n] = 0.0 * r[:n] + r[:n]
    rnorm0 = dot(p, r)
    rnorm = rnorm0
    rtol2 = rtol * rtol
    k = 0
    while k < max_iter:
        k += 1
        if scatter is not None:
            scatter(p)
        apply(p, y)
        alpha = rnorm / dot(p, y)
        x[:n] = p[:n] * alpha + x[:n]
        r[:n] = y[:n] * (-alpha) + r[:n]
        rnorm_new = dot(r, r)
        beta = rnorm_new / rnorm
        rnorm = rnorm_new
        if rnorm / rnorm0 < rtol2:
            break
        p[:n] = p[:n] * beta + r[:n]
    return k


def pcg_solve_jacobi(apply, x, b, diag_inv, max_iter, rtol=0.0, n_owned=None, dot=None, scatter=None):
    """Jacobi-preconditioned CG: the reference's cg_solve (src/cg.hpp:89-169) with the step its comment at :165
    names, "p <- beta*p + M^-1(r)", M^-1 = the inverse diagonal (compute_diag_inv, src/csr.hpp:79-107) applied
    with pointwise_mult (src/vector.hpp:266-285).  With p0 = z0 the reference's `rnorm = inner_product(p, r)`
    (:112) is r.z, which is what alpha, beta and the rtol test use."""
    n = len(x) if n_owned is None else n_owned
    if dot is None:
        dot = lambda a, c: float(np.dot(a[:n], c[:n]))
    y = np.zeros_like(x)
    apply(x, y)
    r = np.zeros_like(x)
    r[:n] = -1 * y[:n] + b[:n]
    p = np.zeros_like(x)
    z = np.zeros_like(x)
    p[:n] = diag_inv[:n] * r[:n]
    rz0 = dot(p, r)
    rz = rz0
    rtol2 = rtol * rtol
    k = 0
    while k < max_iter:
        k += 1
        if scatter is not None:
            scatter(p)
        apply(p, y)
        alpha = rz / dot(p, y)
        x[:n] = p[:n] * alpha + x[:n]
        r[:n] = y[:n] * (-alpha) + r[:n]
        z[:n] = diag_inv[:n] * r[:n]
        rz_new = dot(r, z)
        beta = rz_new / rz
        rz = rz_new
        if rz / rz0 < rtol2:
            break
        p[:n] = p[:n] * beta + z[:n]
    return k
