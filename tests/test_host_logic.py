"""Host-side logic of the product (tables, mesh sizing, byte model) against the oracle. CPU only."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5, 6, 7])
@pytest.mark.parametrize("qmode,use_gauss", [(0, False), (1, False), (1, True)])
@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_tables_match_oracle(degree, qmode, use_gauss, dtype):
    import benchmark_dolfinx_b200 as bd
    from oracle.tables import Tables
    t = bd.Tables(degree, qmode, use_gauss, dtype)
    o = Tables(degree, qmode, use_gauss, dtype)
    eps = np.finfo(dtype).eps
    assert t.nq == o.nq == degree + 1 + qmode and t.is_identity == o.is_identity
    assert np.abs(t.phi0 - o.phi0).max() <= 4 * eps
    assert np.abs(t.dphi1 - o.dphi1).max() <= 64 * eps * np.abs(o.dphi1).max()
    assert np.abs(t.weights3d - o.weights3d).max() <= 4 * eps
    assert np.abs(t.dphi_geometry - o.dphi_geometry).max() <= 4 * eps
    assert np.abs(t.dof_nodes - o.dof_nodes).max() <= 1e-15
    assert np.abs(t.qpts - o.qpts).max() <= 1e-15
    # partition of unity / exactness properties, independent of the oracle
    assert np.abs(t.phi0.sum(axis=1) - 1).max() <= 16 * eps
    assert np.abs(t.dphi1.sum(axis=1)).max() <= 256 * eps * np.abs(t.dphi1).max()
    assert abs(t.weights3d.sum() - 1) <= 64 * eps


def test_compute_mesh_size_matches_reference_examples():
    import benchmark_dolfinx_b200 as bd
    from oracle.boxmesh import compute_mesh_size
    assert bd.compute_mesh_size(100000, 3) == [12, 17, 17]                   # examples/mat_comp-16.json
    assert bd.compute_mesh_size(64 * 300_000_000, 3) == [887, 893, 897]      # examples/Q3-300M.json
    assert bd.compute_mesh_size(64 * 500_000_000, 6) == [524, 529, 534]      # examples/Q6-500M.json
    for nd, p in [(1000, 3), (500, 3), (10000, 3), (1, 1), (27, 2), (50_000_000, 1), (50_000_000, 2),
                  (300_000_000, 3), (500_000_000, 6), (2_400_000_000, 3), (1_000_000_000, 6)]:
        assert bd.compute_mesh_size(nd, p) == compute_mesh_size(nd, p), (nd, p)


def test_partition_grid():
    import benchmark_dolfinx_b200 as bd
    from oracle.boxmesh import partition_grid
    for n in (1, 2, 3, 4, 6, 8, 12, 16):
        g = bd.partition_grid(n)
        assert g[0] * g[1] * g[2] == n and tuple(g) == tuple(partition_grid(n))


def test_byte_model_matches_survey():
    """algorithmic bytes per dof (G precomputed, fp64, qmode 1)."""
    import bench
    for degree, ndofs, expected in [(3, 300_000_000, 247.7), (6, 500_000_000, 136.7),
                                    (1, 50_000_000, 1334.2), (2, 50_000_000, 411.3)]:
        from oracle.boxmesh import compute_mesh_size
        n = compute_mesh_size(ndofs, degree)
        nc = n[0] * n[1] * n[2]
        nd = (n[0] * degree + 1) * (n[1] * degree + 1) * (n[2] * degree + 1)
        b = bench.algorithmic_bytes_per_apply(nd, nc, degree, degree + 2, 8) / nd
        assert abs(b - expected) < 0.15, (degree, b)
    assert bench.algorithmic_flops_per_cell(3, 5, 1) == 14630
    assert bench.algorithmic_flops_per_cell(6, 8, 1) == 96224


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: the package must not import, link or execute it."""
    pkg = os.path.join(ROOT, "benchmark_dolfinx_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "oracle/" not in text and "libport" not in text and "libref_cpu" not in text, f
