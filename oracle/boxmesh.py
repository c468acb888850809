"""Structured stand-in for the DOLFINx host setup (oracle; test infrastructure).

DOLFINx v0.10 (mesh, dofmap, IndexMap, Scatterer) is not available, so the arrays
the operator consumes are synthesised for a block
partition of the `compute_mesh_size` box.  This numpy version is deliberately
written by brute force (global ids, sorting) so that it is an independent check
of the closed-form device generator in the product.

Follows:
  compute_mesh_size              src/mesh.cpp:117-152
  ghost layer (one cell deep)    src/mesh.cpp:26-114 (here: only the cells that
                                 touch an owned dof)
  tp-ordered geometry dofmap     src/mesh.cpp:89-100,209-215
  x-perturbation                 src/mesh.cpp:199-207 (amplitude identical; the
                                 mt19937(42) stream in DOLFINx vertex order is
                                 not reproducible without DOLFINx, a counter-
                                 based hash of the global vertex id is used)
  lcells / bcells split          src/mesh.cpp:154-188
  boundary dofs (whole surface)  src/main.cpp:94-102
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


def compute_mesh_size(ndofs_global: int, degree: int):
    """src/mesh.cpp:117-152, including loop order and the strict '<'."""
    nx_approx = (float(ndofs_global) ** (1.0 / 3.0) - 1) / degree
    n0 = int(nx_approx + 0.5)
    nx = [n0, n0, n0]
    best = abs((n0 * degree + 1) ** 3 - ndofs_global)
    lo = max(1, n0 - 5)
    for a in range(lo, n0 + 6):
        for b in range(lo, n0 + 6):
            for c in range(lo, n0 + 6):
                misfit = abs((a * degree + 1) * (b * degree + 1) * (c * degree + 1) - ndofs_global)
                if misfit < best:
                    best = misfit
                    nx = [a, b, c]
    return nx


def partition_grid(nranks: int):
    """Block-partition shape used by the stand-in."""
    table = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
    if nranks in table:
        return table[nranks]
    # generic: factor greedily into three near-equal factors
    p = [1, 1, 1]
    n = nranks
    f = 2
    fs = []
    while n > 1:
        while n % f == 0:
            fs.append(f)
            n //= f
        f += 1
    for f in sorted(fs, reverse=True):
        p[int(np.argmin(p))] *= f
    return tuple(sorted(p, reverse=True))


_SPLITMIX_GAMMA = np.uint64(0x9E3779B97F4A7C15)


def perturb_unit(gid: np.ndarray) -> np.ndarray:
    """Counter-based uniform in [-1, 1) from a global vertex id (splitmix64)."""
    with np.errstate(over="ignore"):
        z = (gid.astype(np.uint64) + np.uint64(42)) * _SPLITMIX_GAMMA
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0


@dataclass
class BoxMesh:
    """Arrays for one rank of a px*py*pz block partition of an n0*n1*n2 hex box."""
    n: tuple
    degree: int
    rank: int
    nranks: int
    pgrid: tuple
    ncells_owned: int = 0
    ncells: int = 0                 # owned + ghost
    ndofs_owned: int = 0
    nghost: int = 0
    ndofs_global: int = 0
    ncells_global: int = 0
    cell_dofmap: np.ndarray = None      # int32 [ncells, nd^3]
    geometry_dofmap: np.ndarray = None  # int32 [ncells, 8]
    xgeom: np.ndarray = None            # float64 [nnodes, 3]
    bc_marker: np.ndarray = None        # int8 [ndofs_owned + nghost]
    lcells: np.ndarray = None           # int32
    bcells: np.ndarray = None           # int32
    dof_global: np.ndarray = None       # int64 [ndofs_owned + nghost] global lattice id
    ghost_owner: np.ndarray = None      # int32 [nghost]
    # halo plan: receive segments are contiguous in ghost order
    recv_ranks: list = field(default_factory=list)
    recv_counts: list = field(default_factory=list)
    send_ranks: list = field(default_factory=list)
    send_indices: list = field(default_factory=list)   # list of int32 arrays (local owned idx)

    @property
    def ndofs_local(self):
        return self.ndofs_owned + self.nghost


def _block_range(n, p, b):
    return (b * n) // p, ((b + 1) * n) // p


def _rank_geometry(n, P, pgrid, rank):
    b = [rank // (pgrid[1] * pgrid[2]), (rank // pgrid[2]) % pgrid[1], rank % pgrid[2]]
    g = []
    for d in range(3):
        c0, c1 = _block_range(n[d], pgrid[d], b[d])
        lo = c0 * P
        hi = c1 * P + (1 if b[d] == pgrid[d] - 1 else 0)
        e0 = max(c0 - 1, 0)
        g.append(dict(c0=c0, c1=c1, lo=lo, hi=hi, e0=e0))
    return b, g


def _owner_rank_of_points(n, P, pgrid, gx, gy, gz):
    """Owner rank of global lattice points (lower interface plane belongs to the upper block)."""
    ob = []
    for d, gcoord in enumerate((gx, gy, gz)):
        starts = np.array([_block_range(n[d], pgrid[d], b)[0] * P for b in range(pgrid[d])])
        ob.append(np.searchsorted(starts, gcoord, side="right") - 1)
    return (ob[0] * pgrid[1] + ob[1]) * pgrid[2] + ob[2]


def create_box_mesh(n, degree, rank=0, nranks=1, geom_perturb_fact=0.0, pgrid=None) -> BoxMesh:
    n = tuple(int(v) for v in n)
    P = degree
    nd = P + 1
    pgrid = tuple(pgrid) if pgrid is not None else partition_grid(nranks)
    assert pgrid[0] * pgrid[1] * pgrid[2] == nranks
    b, g = _rank_geometry(n, P, pgrid, rank)
    NP = [n[d] * P + 1 for d in range(3)]          # global lattice extents
    NV = [n[d] + 1 for d in range(3)]

    m = BoxMesh(n=n, degree=P, rank=rank, nranks=nranks, pgrid=pgrid)
    m.ndofs_global = NP[0] * NP[1] * NP[2]
    m.ncells_global = n[0] * n[1] * n[2]

    # ---- cells: owned (lexicographic in block) then ghost (ascending global id)
    ext = [np.arange(g[d]["e0"], g[d]["c1"]) for d in range(3)]
    CI, CJ, CK = np.meshgrid(*ext, indexing="ij")
    ci, cj, ck = CI.ravel(), CJ.ravel(), CK.ravel()
    owned = (ci >= g[0]["c0"]) & (cj >= g[1]["c0"]) & (ck >= g[2]["c0"])
    order = np.concatenate([np.nonzero(owned)[0], np.nonzero(~owned)[0]])
    ci, cj, ck = ci[order], cj[order], ck[order]
    m.ncells_owned = int(owned.sum())
    m.ncells = len(ci)

    # ---- dof points referenced by local cells, owner and local numbering
    a = np.arange(nd)
    A, B, C = np.meshgrid(a, a, a, indexing="ij")   # local index a*nd^2 + b*nd + c, x slowest
    gx = (ci[:, None] * P + A.ravel()[None, :])
    gy = (cj[:, None] * P + B.ravel()[None, :])
    gz = (ck[:, None] * P + C.ravel()[None, :])
    gid = (gx.astype(np.int64) * NP[1] + gy) * NP[2] + gz        # [ncells, nd^3]
    uniq = np.unique(gid)
    ux = uniq // (NP[1] * NP[2])
    uy = (uniq // NP[2]) % NP[1]
    uz = uniq % NP[2]
    owner = _owner_rank_of_points(n, P, pgrid, ux, uy, uz)
    mine = owner == rank
    own_ids = uniq[mine]                                        # ascending global id == lexicographic
    gh_ids = uniq[~mine]
    gh_owner = owner[~mine]
    gorder = np.lexsort((gh_ids, gh_owner))                     # by (owner rank, global id)
    gh_ids, gh_owner = gh_ids[gorder], gh_owner[gorder]
    m.ndofs_owned = len(own_ids)
    m.nghost = len(gh_ids)
    m.dof_global = np.concatenate([own_ids, gh_ids]).astype(np.int64)
    m.ghost_owner = gh_owner.astype(np.int32)
    sorter = np.argsort(m.dof_global, kind="stable")
    pos = np.searchsorted(m.dof_global, gid.ravel(), sorter=sorter)
    m.cell_dofmap = sorter[pos].reshape(gid.shape).astype(np.int32)

    # ---- BC markers: every lattice point on the cube surface (src/main.cpp:99-102)
    dx = m.dof_global // (NP[1] * NP[2])
    dy = (m.dof_global // NP[2]) % NP[1]
    dz = m.dof_global % NP[2]
    m.bc_marker = ((dx == 0) | (dx == NP[0] - 1) | (dy == 0) | (dy == NP[1] - 1)
                   | (dz == 0) | (dz == NP[2] - 1)).astype(np.int8)

    # ---- geometry: local vertex box [e0, c1] per dim, lexicographic numbering
    v0 = [g[d]["e0"] for d in range(3)]
    nv = [g[d]["c1"] - g[d]["e0"] + 1 for d in range(3)]
    VI, VJ, VK = np.meshgrid(*[np.arange(v0[d], v0[d] + nv[d]) for d in range(3)], indexing="ij")
    vgid = (VI.astype(np.int64) * NV[1] + VJ) * NV[2] + VK
    x = np.stack([VI / n[0], VJ / n[1], VK / n[2]], axis=-1).reshape(-1, 3).astype(np.float64)
    if geom_perturb_fact != 0.0:
        x[:, 0] += perturb_unit(vgid.ravel()) * (geom_perturb_fact * 1.0 / n[0])
    m.xgeom = np.ascontiguousarray(x)
    gd = np.empty((m.ncells, 8), dtype=np.int32)
    for av in range(2):
        for bv in range(2):
            for cv in range(2):
                gd[:, av * 4 + bv * 2 + cv] = (((ci + av - v0[0]) * nv[1] + (cj + bv - v0[1])) * nv[2]
                                               + (ck + cv - v0[2]))
    m.geometry_dofmap = gd

    # ---- lcells / bcells (src/mesh.cpp:154-188)
    has_ghost = (m.cell_dofmap >= m.ndofs_owned).any(axis=1)
    has_ghost[m.ncells_owned:] = True
    m.lcells = np.nonzero(~has_ghost)[0].astype(np.int32)
    m.bcells = np.nonzero(has_ghost)[0].astype(np.int32)

    # ---- halo plan
    if m.nghost:
        rr, cnt = np.unique(m.ghost_owner, return_counts=True)
        m.recv_ranks, m.recv_counts = [int(v) for v in rr], [int(v) for v in cnt]
    # send lists: for every other rank, the points it ghosts that I own, ascending global id
    if nranks > 1:
        for other in range(nranks):
            if other == rank:
                continue
            og = _ghost_ids_of_rank(n, P, pgrid, other)
            ids = og[0][og[1] == rank]
            if len(ids):
                m.send_ranks.append(other)
                m.send_indices.append(np.searchsorted(own_ids, ids).astype(np.int32))
    return m


def _ghost_ids_of_rank(n, P, pgrid, rank):
    """(sorted ghost global ids, owners) of `rank`, brute force."""
    b, g = _rank_geometry(n, P, pgrid, rank)
    NP = [n[d] * P + 1 for d in range(3)]
    rng = [np.arange(g[d]["e0"] * P, g[d]["c1"] * P + 1) for d in range(3)]
    X, Y, Z = np.meshgrid(*rng, indexing="ij")
    gx, gy, gz = X.ravel(), Y.ravel(), Z.ravel()
    owner = _owner_rank_of_points(n, P, pgrid, gx, gy, gz)
    gid = (gx.astype(np.int64) * NP[1] + gy) * NP[2] + gz
    keep = owner != rank
    gid, owner = gid[keep], owner[keep]
    o = np.argsort(gid)
    return gid[o], owner[o]
