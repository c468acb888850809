"""Device-generated stand-in mesh (closed-form kernels) vs the brute-force numpy oracle."""
import numpy as np
import pytest

from oracle.boxmesh import create_box_mesh
from oracle.laplacian import assemble_rhs
from oracle.tables import Tables

pytestmark = pytest.mark.gpu

CASES = [  # n, degree, nranks
    ([3, 3, 3], 3, 1), ([5, 8, 8], 3, 1), ([4, 5, 6], 1, 2), ([5, 4, 6], 2, 4), ([4, 4, 4], 3, 8),
    ([3, 5, 7], 4, 2), ([2, 2, 3], 7, 1), ([5, 3, 2], 2, 2), ([9, 9, 9], 1, 8), ([2, 3, 4], 6, 4),
    ([7, 2, 2], 2, 3), ([6, 6, 5], 2, 12),
]


@pytest.mark.parametrize("n,degree,nranks", CASES)
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_boxmesh_matches_oracle(n, degree, nranks, dtype):
    import benchmark_dolfinx_b200 as bd
    for rank in range(nranks):
        o = create_box_mesh(n, degree, rank=rank, nranks=nranks, geom_perturb_fact=0.2)
        m = bd.BoxMesh(n, degree, rank, nranks, geom_perturb_fact=0.2, dtype=dtype)
        assert (m.ncells_owned, m.ncells, m.ndofs_owned, m.nghost) == (o.ncells_owned, o.ncells,
                                                                       o.ndofs_owned, o.nghost)
        assert (m.ndofs_global, m.ncells_global) == (o.ndofs_global, o.ncells_global)
        h = m.host_arrays()
        assert np.array_equal(h["cell_dofmap"], o.cell_dofmap)
        assert np.array_equal(h["geometry_dofmap"], o.geometry_dofmap)
        assert np.array_equal(h["bc_marker"], o.bc_marker)
        assert np.array_equal(h["lcells"], o.lcells) and np.array_equal(h["bcells"], o.bcells)
        tol = 1e-15 if dtype == np.float64 else 1e-7
        assert np.abs(h["xgeom"] - o.xgeom).max() <= tol
        assert list(h["recv_ranks"]) == o.recv_ranks and list(h["recv_counts"]) == o.recv_counts
        assert list(h["send_ranks"]) == o.send_ranks
        assert list(h["send_counts"]) == [len(v) for v in o.send_indices]
        if o.send_indices:
            assert np.array_equal(h["send_indices"], np.concatenate(o.send_indices))


@pytest.mark.parametrize("ndofs,degree,qmode,use_gauss,unorm", [
    (1000, 3, 0, False, 7.879507448215885),          # SURVEY appendix B (first row = CI case)
    (100000, 3, 1, False, 0.6895773850559623),       # examples/mat_comp-16.json u_norm
    (1000, 1, 1, False, 5.214132462541599),
    (5000, 6, 1, False, 4.736123103582485),
    (1000, 3, 1, True, 7.880191187816367),
])
def test_rhs_matches_reference_u_norm(ndofs, degree, qmode, use_gauss, unorm):
    import benchmark_dolfinx_b200 as bd
    n = bd.compute_mesh_size(ndofs, degree)
    m = bd.BoxMesh(n, degree)
    u = m.assemble_rhs(qmode, use_gauss)
    assert abs(bd.norm(u) - unorm) <= 1e-12 * unorm


@pytest.mark.parametrize("degree,qmode,nranks", [(2, 1, 1), (3, 0, 2), (4, 1, 4)])
def test_rhs_matches_oracle_entrywise(degree, qmode, nranks):
    import benchmark_dolfinx_b200 as bd
    n = [4, 3, 5]
    for rank in range(nranks):
        o = create_box_mesh(n, degree, rank=rank, nranks=nranks, geom_perturb_fact=0.2)
        b_ref = assemble_rhs(o, Tables(degree, qmode))
        m = bd.BoxMesh(n, degree, rank, nranks, geom_perturb_fact=0.2)
        b = m.assemble_rhs(qmode).cpu().numpy()
        assert np.linalg.norm(b - b_ref) <= 1e-13 * np.linalg.norm(b_ref)
        assert (b[o.bc_marker != 0] == 0).all() and (b[o.ndofs_owned:] == 0).all()
