"""1-D tables for the sum-factorised Q_k hex Laplacian (oracle; test infrastructure).

Restates, without Basix, what the reference's operator constructor obtains from
Basix v0.10.0 (not vendored under /root/reference; pinned in README.md:84):

* quadrature rule selection            src/laplacian.hpp:123-146
* coordinate-element derivative table  src/laplacian.hpp:148-157
* 1-D dof element (GLL, degree P)      src/laplacian.hpp:161-164
* 1-D "collocated" element + phi0      src/laplacian.hpp:179-194
* dphi1 (derivative half of tabulate)  src/laplacian.hpp:201,210-212

All nodes are in ascending order on [0, 1].  Any consistent 1-D ordering gives
identical operator results; Basix's own ordering only matters for
a bitwise DOLFINx drop-in, where the tables come from Basix anyway.
"""
from __future__ import annotations

import numpy as np

LD = np.longdouble


def gll_nodes_weights(n: int):
    """n-point Gauss-Lobatto-Legendre rule on [0, 1] (ascending)."""
    if n < 2:
        raise ValueError("GLL needs at least 2 points")
    N = n - 1
    # Chebyshev-Gauss-Lobatto initial guess, Newton on (1-x^2) P_N'(x)
    x = -np.cos(np.pi * np.arange(n, dtype=LD) / N)
    for _ in range(100):
        # q(x) = P_{N+1} - P_{N-1} has the GLL nodes as roots (prop. to (1-x^2)P_N')
        pN1, dN1 = _legendre_safe(N + 1, x)
        pNm, dNm = _legendre_safe(N - 1, x)
        q = pN1 - pNm
        dq = dN1 - dNm
        dx = q / dq
        x = x - dx
        if np.max(np.abs(dx)) < 1e-19:
            break
    x[0], x[-1] = -1, 1
    pN, _ = _legendre_safe(N, x)
    w = 2 / (N * (N + 1) * pN * pN)
    return ((x + 1) / 2).astype(LD), (w / 2).astype(LD)


def _legendre_safe(n: int, x):
    """P_n and P_n' valid also at x = +-1."""
    x = np.asarray(x, dtype=LD)
    p0 = np.ones_like(x)
    d0 = np.zeros_like(x)
    if n == 0:
        return p0, d0
    p1 = x.copy()
    d1 = np.ones_like(x)
    for k in range(2, n + 1):
        p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k
        d2 = d0 + (2 * k - 1) * p1
        p0, p1 = p1, p2
        d0, d1 = d1, d2
    return p1, d1


def gl_nodes_weights(n: int):
    """n-point Gauss-Legendre rule on [0, 1] (ascending)."""
    x = np.polynomial.legendre.leggauss(n)[0].astype(LD)
    for _ in range(5):  # polish in extended precision
        p, dp = _legendre_safe(n, x)
        x = x - p / dp
    _, dp = _legendre_safe(n, x)
    w = 2 / ((1 - x * x) * dp * dp)
    return ((x + 1) / 2).astype(LD), (w / 2).astype(LD)


def lagrange_eval(nodes, x):
    """L[q, i] = l_i(x_q) for the Lagrange basis on `nodes`."""
    nodes = np.asarray(nodes, dtype=LD)
    x = np.asarray(x, dtype=LD)
    n = len(nodes)
    L = np.ones((len(x), n), dtype=LD)
    for i in range(n):
        for j in range(n):
            if j != i:
                L[:, i] *= (x - nodes[j]) / (nodes[i] - nodes[j])
    return L


def lagrange_deriv_at_nodes(nodes):
    """D[q, i] = l_i'(x_q) for the Lagrange basis on `nodes`, at the nodes."""
    nodes = np.asarray(nodes, dtype=LD)
    n = len(nodes)
    bw = np.ones(n, dtype=LD)  # barycentric weights
    for i in range(n):
        for j in range(n):
            if j != i:
                bw[i] /= (nodes[i] - nodes[j])
    D = np.zeros((n, n), dtype=LD)
    for q in range(n):
        for i in range(n):
            if i != q:
                D[q, i] = (bw[i] / bw[q]) / (nodes[q] - nodes[i])
        D[q, q] = -np.sum(D[q, :])
    return D


class Tables:
    """Everything the operator constructor derives from (degree, qmode, rule).

    Attributes (float64 unless `dtype` says otherwise):
      nd, nq           1-D dof / quadrature point counts
      dof_nodes        [nd]   GLL nodes of the degree-P element
      qpts, qwts       [nq]   1-D quadrature rule on [0, 1]
      phi0             [nq, nd]  l_i^{dof nodes}(x_q), |.| < 5 eps flushed to 0
      dphi1            [nq, nq]  d/dx l_i^{quad nodes}(x_q)
      weights3d        [nq^3]    tensor-product weights, q = q0*nq^2 + q1*nq + q2
      dphi_geometry    [3, nq^3, 8] derivative of the trilinear vertex basis,
                       vertex v = a*4 + b*2 + c (x slowest)
      is_identity      phi0 is the identity (qmode 0 with GLL)
    """

    def __init__(self, degree: int, qmode: int, use_gauss: bool = False, dtype=np.float64):
        if degree < 1 or degree > 7:
            raise ValueError("Unsupported degree [operator]")
        if qmode not in (0, 1):
            raise ValueError("Invalid qmode.")
        self.degree, self.qmode, self.use_gauss = degree, qmode, use_gauss
        self.nd = degree + 1
        # src/laplacian.hpp:125-134 (q_map) + Basix point counts => nq = P+1+qmode
        self.nq = degree + 1 + qmode
        dof_nodes, _ = gll_nodes_weights(self.nd)
        if use_gauss:
            qpts, qwts = gl_nodes_weights(self.nq)
        else:
            qpts, qwts = gll_nodes_weights(self.nq)
        phi0 = lagrange_eval(dof_nodes, qpts)
        eps = np.finfo(dtype).eps
        phi0 = phi0.astype(dtype)
        phi0[np.abs(phi0) < 5 * eps] = 0  # src/laplacian.hpp:188-193
        # src/laplacian.hpp:34-54 (matrix_is_identity)
        self.is_identity = bool(
            phi0.shape[0] == phi0.shape[1]
            and np.all(np.abs(phi0 - np.eye(phi0.shape[0])) <= 5 * eps))
        if qmode == 0 and not self.is_identity:
            raise RuntimeError("Expecting identity matrix for qmode=0")  # :197-198
        self.dof_nodes = dof_nodes.astype(np.float64)
        self.qpts = qpts.astype(np.float64)
        self.qwts = qwts.astype(np.float64)
        self.phi0 = np.ascontiguousarray(phi0)
        self.dphi1 = np.ascontiguousarray(lagrange_deriv_at_nodes(qpts).astype(dtype))
        w = qwts
        self.weights3d = np.ascontiguousarray(
            (w[:, None, None] * w[None, :, None] * w[None, None, :]).reshape(-1).astype(dtype))
        self.dphi_geometry = np.ascontiguousarray(self._dphi_geometry(qpts).astype(dtype))
        self.dtype = np.dtype(dtype)

    def _dphi_geometry(self, qpts):
        nq = len(qpts)
        L = np.stack([1 - qpts, qpts], axis=1)          # [nq, 2]  value of L_a at point
        dL = np.stack([-np.ones(nq, dtype=LD), np.ones(nq, dtype=LD)], axis=1)
        out = np.zeros((3, nq, nq, nq, 2, 2, 2), dtype=LD)
        for d in range(3):
            f = [L, L, L]
            f[d] = dL
            out[d] = np.einsum("ia,jb,kc->ijkabc", f[0], f[1], f[2])
        return out.reshape(3, nq ** 3, 8)
