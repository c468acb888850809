"""Pin the CPU oracle: reference known-answer values, the reference's own CPU kernels
(oracle/_ref, built from /root/reference/src) and the plain-C port.  CPU only."""
import os

import numpy as np
import pytest

from oracle import native
from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian, assemble_csr, assemble_rhs, cg_solve
from oracle.tables import Tables

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _run(ndofs, degree, qmode, use_gauss=False, cg=0, perturb=0.0):
    n = compute_mesh_size(ndofs, degree)
    mesh = create_box_mesh(n, degree, geom_perturb_fact=perturb)
    t = Tables(degree, qmode, use_gauss)
    op = OracleLaplacian(mesh, t)
    u = assemble_rhs(mesh, t)
    if cg:
        y = np.zeros_like(u)
        cg_solve(op.apply, y, u, cg)
    else:
        y = op.apply(u)
    return mesh, u, y


def test_compute_mesh_size_goldens():
    # examples/mat_comp-16.json, examples/Q3-300M.json, examples/Q6-500M.json
    for ndofs, degree, ncells, ndofs_out in [
            (100000, 3, 3468, 100048),
            (64 * 300_000_000, 3, 710505627, 19205158720),
            (64 * 500_000_000, 6, 148022664, 32003126875)]:
        n = compute_mesh_size(ndofs, degree)
        assert n[0] * n[1] * n[2] == ncells
        assert (n[0] * degree + 1) * (n[1] * degree + 1) * (n[2] * degree + 1) == ndofs_out
    assert compute_mesh_size(1000, 3) == [3, 3, 3]
    assert compute_mesh_size(10000, 3) == [5, 8, 8]


def test_ci_known_answer():
    """src/test_output.py:15-19: --ndofs=1000 --degree=3 --qmode=0 --nreps=1 --float=64."""
    mesh, u, y = _run(1000, 3, 0)
    assert mesh.ndofs_global == 1000
    assert np.isclose(np.linalg.norm(y), 9.912865833415553, rtol=1e-13, atol=0)


def test_example_mat_comp_16():
    """examples/mat_comp-16.json: Q3, qmode 1, GLL, 100048 dofs."""
    mesh, u, y = _run(100000, 3, 1)
    assert mesh.ncells_global == 3468 and mesh.ndofs_global == 100048
    assert np.isclose(np.linalg.norm(u), 0.6895773850559623, rtol=1e-13, atol=0)
    assert np.isclose(np.linalg.norm(y), 0.14150257625641838, rtol=1e-13, atol=0)


# SURVEY.md appendix B (derived with the restatement that reproduces the two goldens above)
@pytest.mark.parametrize("args,unorm,ynorm", [
    ((10000, 3, 0), 2.2794459800498785, 1.1590666300856363),
    ((10000, 3, 1), 2.2790421417877087, 1.1415775083651323),
    ((1000, 1, 0), 6.185029717368478, 2.2734429258669544),
    ((1000, 2, 1), 8.355331483208953, 11.93069742770147),
    ((5000, 6, 1), 4.736123103582485, 6.299722546620158),
    ((5000, 7, 1), 2.5629538656265947, 1.1524308636572587),
    ((1000, 3, 1, True), 7.880191187816367, 10.037510387732647),
    ((1000, 3, 0, False, 10), 7.879507448215885, 46.17746829199996),
    ((10000, 3, 1, False, 50), 2.2790421417877087, 167.59244725655134),
])
def test_secondary_regression_values(args, unorm, ynorm):
    mesh, u, y = _run(*args)
    assert np.isclose(np.linalg.norm(u), unorm, rtol=1e-12, atol=0)
    assert np.isclose(np.linalg.norm(y), ynorm, rtol=1e-11, atol=0)


@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5, 6, 7])
@pytest.mark.parametrize("perturb", [0.0, 0.2])
@pytest.mark.parametrize("strict", [False, True])
def test_oracle_matches_reference_cpu_kernels(degree, perturb, strict):
    """numpy restatement vs the reference's unmodified stiffness_operator_cpu /
    geometry_computation_cpu (qmode 0, all they support)."""
    ref = native.ref_lib(strict)
    if ref is None:
        pytest.skip("oracle/_ref not built (reference sources absent)")
    n = compute_mesh_size(3000, degree)
    mesh = create_box_mesh(n, degree, geom_perturb_fact=perturb)
    t = Tables(degree, 0)
    op = OracleLaplacian(mesh, t)
    u = np.random.default_rng(1234).uniform(-1, 1, mesh.ndofs_local)
    y = op.apply(u)
    G = native.ref_geometry(ref, t.nq, mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry,
                            t.weights3d, op.cells)
    assert np.abs(G - op.G).max() <= 1e-13 * np.abs(op.G).max()
    y2 = np.zeros_like(u)
    native.ref_stiffness(ref, degree, u, y2, G, np.full(mesh.ncells, 2.0), t.phi0, t.dphi1,
                         mesh.cell_dofmap, op.cells, mesh.bc_marker)
    assert np.linalg.norm(y2 - y) <= 1e-13 * np.linalg.norm(y)


def test_reference_cpu_kernels_fp32():
    ref = native.ref_lib(False)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    degree = 3
    mesh = create_box_mesh(compute_mesh_size(3000, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, 0)
    op = OracleLaplacian(mesh, t)
    u = np.random.default_rng(5).uniform(-1, 1, mesh.ndofs_local)
    y = op.apply(u)
    f = np.float32
    G = native.ref_geometry(ref, t.nq, mesh.xgeom.astype(f), mesh.geometry_dofmap,
                            t.dphi_geometry.astype(f), t.weights3d.astype(f), op.cells)
    y2 = np.zeros(mesh.ndofs_local, dtype=f)
    native.ref_stiffness(ref, degree, u.astype(f), y2, G, np.full(mesh.ncells, 2.0, dtype=f),
                         t.phi0.astype(f), t.dphi1.astype(f), mesh.cell_dofmap, op.cells,
                         mesh.bc_marker)
    assert np.linalg.norm(y2 - y) <= 1e-5 * np.linalg.norm(y)


@pytest.mark.parametrize("degree,qmode", [(1, 1), (2, 1), (3, 1), (3, 0), (4, 0), (6, 1), (7, 1)])
def test_oracle_matches_c_port(degree, qmode):
    port = native.port_lib()
    if port is None:
        pytest.skip("oracle/libport.so not built")
    mesh = create_box_mesh(compute_mesh_size(3000, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    op = OracleLaplacian(mesh, t)
    u = np.random.default_rng(1).uniform(-1, 1, mesh.ndofs_local)
    y = op.apply(u)
    G = native.port_geometry(port, t.nq, mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry,
                             t.weights3d, op.cells)
    assert np.abs(G - op.G).max() <= 1e-13 * np.abs(op.G).max()
    y2 = np.zeros_like(u)
    native.port_stiffness(port, t.nd, t.nq, t.is_identity, u, y2, G, np.full(mesh.ncells, 2.0),
                          t.phi0, t.dphi1, mesh.cell_dofmap, op.cells, mesh.bc_marker)
    assert np.linalg.norm(y2 - y) <= 1e-13 * np.linalg.norm(y)


@pytest.mark.parametrize("degree,qmode,perturb", [(3, 0, 0.0), (3, 1, 0.2), (2, 1, 0.2), (5, 1, 0.1)])
def test_matrix_free_vs_assembled_csr(degree, qmode, perturb):
    """--mat_comp self-check (src/laplacian_solver.cpp:151-227); README expects ~1e-15."""
    mesh = create_box_mesh(compute_mesh_size(2000, degree), degree, geom_perturb_fact=perturb)
    t = Tables(degree, qmode)
    op = OracleLaplacian(mesh, t)
    u = assemble_rhs(mesh, t) + np.random.default_rng(3).uniform(-1, 1, mesh.ndofs_local)
    z = assemble_csr(mesh, t) @ u
    y = op.apply(u)
    assert np.linalg.norm(y - z) / np.linalg.norm(z) < 1e-13


def test_golden_fixtures_match_oracle():
    """tests/golden/*.npz were produced by tests/golden/make_golden.py from the reference's own
    CPU kernels (oracle/_ref); the restatement must reproduce them without /root/reference."""
    files = sorted(f for f in os.listdir(GOLDEN) if f.endswith(".npz"))
    assert files, "no golden fixtures committed"
    for f in files:
        d = np.load(os.path.join(GOLDEN, f))
        degree, qmode, perturb = int(d["degree"]), int(d["qmode"]), float(d["perturb"])
        mesh = create_box_mesh([int(v) for v in d["n"]], degree, geom_perturb_fact=perturb)
        t = Tables(degree, qmode)
        op = OracleLaplacian(mesh, t)
        y = op.apply(d["u"].copy())
        assert np.linalg.norm(y - d["y"]) <= 1e-13 * np.linalg.norm(d["y"]), f
        assert np.isclose(np.linalg.norm(y), float(d["y_norm"]), rtol=1e-13, atol=0)


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_partitioned_oracle_equals_single_rank(nranks):
    """The block-partitioned stand-in (ghost layer, owned-first numbering, halo lists) gives the
    same operator as the single-rank mesh."""
    degree, qmode = 2, 1
    n = [5, 4, 6]
    t = Tables(degree, qmode)
    m1 = create_box_mesh(n, degree, geom_perturb_fact=0.2)
    u1 = np.random.default_rng(7).uniform(-1, 1, m1.ndofs_local)
    y1 = OracleLaplacian(m1, t).apply(u1)
    g2l = {int(g): i for i, g in enumerate(m1.dof_global)}
    meshes = [create_box_mesh(n, degree, rank=r, nranks=nranks, geom_perturb_fact=0.2) for r in range(nranks)]
    assert sum(m.ndofs_owned for m in meshes) == m1.ndofs_global
    vecs = []
    for m in meshes:
        v = np.zeros(m.ndofs_local)
        own = m.dof_global[:m.ndofs_owned]
        v[:m.ndofs_owned] = u1[[g2l[int(g)] for g in own]]
        vecs.append(v)
    # forward scatter through the send/recv lists
    for r, m in enumerate(meshes):
        off = m.ndofs_owned
        for src, cnt in zip(m.recv_ranks, m.recv_counts):
            ms = meshes[src]
            k = ms.send_ranks.index(r)
            assert len(ms.send_indices[k]) == cnt
            vecs[r][off:off + cnt] = vecs[src][ms.send_indices[k]]
            off += cnt
        assert off == m.ndofs_local
    for r, m in enumerate(meshes):
        y = OracleLaplacian(m, t).apply(vecs[r])
        own = m.dof_global[:m.ndofs_owned]
        ref = y1[[g2l[int(g)] for g in own]]
        assert np.linalg.norm(y[:m.ndofs_owned] - ref) <= 1e-13 * np.linalg.norm(y1)
        # lcells touch no ghost dof; bcells = the rest incl. every ghost cell (src/mesh.cpp:154-188)
        assert (m.cell_dofmap[m.lcells] < m.ndofs_owned).all()
        assert len(m.lcells) + len(m.bcells) == m.ncells


def test_oracle_jacobi_pcg_solves_the_assembled_system():
    """The oracle's Jacobi-preconditioned CG (the checker of tests/test_gpu_jacobi.py) against a direct solve of the
    oracle's assembled matrix, and never slower than plain CG in iterations."""
    import scipy.sparse.linalg as spla
    from oracle.laplacian import assemble_csr, assemble_rhs, cg_solve, pcg_solve_jacobi
    for degree, qmode, perturb in ((2, 1, 0.2), (3, 0, 0.0)):
        mesh = create_box_mesh(compute_mesh_size(2000, degree), degree, geom_perturb_fact=perturb)
        t = Tables(degree, qmode)
        A = assemble_csr(mesh, t, kappa=2.0)
        b = assemble_rhs(mesh, t)
        op = OracleLaplacian(mesh, t, kappa=2.0)
        x = np.zeros_like(b)
        its = pcg_solve_jacobi(op.apply, x, b, 1.0 / A.diagonal(), 500, rtol=1e-10)
        x_direct = spla.spsolve(A.tocsc(), b)
        assert np.linalg.norm(x - x_direct) / np.linalg.norm(x_direct) < 1e-8
        its_plain = cg_solve(op.apply, np.zeros_like(b), b, 500, rtol=1e-10)
        assert 0 < its <= its_plain
