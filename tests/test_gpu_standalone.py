"""The stand-alone replacements of the two reference kernel launches, called through the C ABI with
the reference's own data layout: bd_geometry_compute (geometry_computation_gpu, src/geometry_gpu.hpp)
and bd_stiffness_apply (stiffness_operator_gpu, src/laplacian_gpu.hpp), plus edge cases of the
operator's cell lists (empty lists, lists that do not fill a warp batch)."""
import numpy as np
import pytest

from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.laplacian import OracleLaplacian, stiffness_apply
from oracle.tables import Tables

pytestmark = pytest.mark.gpu
RTOL = {np.float64: 1e-12, np.float32: 1e-5}


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("degree,qmode", [(1, 0), (1, 1), (2, 1), (3, 0), (3, 1), (4, 1), (5, 0), (6, 1), (7, 0), (7, 1)])
def test_geometry_and_stiffness_entry_points(degree, qmode, dtype):
    import torch
    from benchmark_dolfinx_b200 import _lib
    lib, check = _lib.lib, _lib.check
    bits = 64 if dtype == np.float64 else 32
    mesh = create_box_mesh(compute_mesh_size(2500, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    oracle = OracleLaplacian(mesh, t)
    # an arbitrary sub-list in scrambled order: G and kappa are indexed by list position
    rng = np.random.default_rng(11)
    cells = rng.permutation(mesh.ncells)[: max(1, (2 * mesh.ncells) // 3)].astype(np.int32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    xg, gd = dev(mesh.xgeom.astype(dtype)), dev(mesh.geometry_dofmap)
    dphi, w = dev(t.dphi_geometry.astype(dtype)), dev(t.weights3d.astype(dtype))
    cd = dev(cells)
    dm_d, bc_d = dev(mesh.cell_dofmap), dev(mesh.bc_marker)   # keep alive: launches are asynchronous
    G = torch.zeros(len(cells) * 6 * t.nq ** 3, dtype=xg.dtype, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    check(lib.bd_geometry_compute(bits, t.nq, xg.data_ptr(), gd.data_ptr(), dphi.data_ptr(), w.data_ptr(),
                                  cd.data_ptr(), len(cells), G.data_ptr(), s))
    from oracle.laplacian import geometry_G
    G_ref, _ = geometry_G(mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry, t.weights3d, cells)
    torch.cuda.synchronize()
    assert _rel(G.cpu().numpy(), G_ref.ravel()) < RTOL[dtype]

    u = rng.uniform(-1, 1, mesh.ndofs_local)
    y_ref = np.full(mesh.ndofs_local, 0.5)                       # accumulates, does not zero
    stiffness_apply(u, y_ref, G_ref, 2.0, mesh.cell_dofmap, cells, mesh.bc_marker, t.phi0, t.dphi1, t.is_identity)
    ud, yd = dev(u.astype(dtype)), torch.full((mesh.ndofs_local,), 0.5, dtype=xg.dtype, device="cuda")
    consts = torch.full((len(cells),), 2.0, dtype=xg.dtype, device="cuda")
    phi0, dphi1 = np.ascontiguousarray(t.phi0.astype(dtype)), np.ascontiguousarray(t.dphi1.astype(dtype))
    check(lib.bd_stiffness_apply(bits, degree, t.nq, int(t.is_identity), ud.data_ptr(), consts.data_ptr(),
                                 yd.data_ptr(), G.data_ptr(), phi0.ctypes.data, dphi1.ctypes.data,
                                 dm_d.data_ptr(), cd.data_ptr(), len(cells),
                                 bc_d.data_ptr(), s))
    torch.cuda.synchronize()
    assert _rel(yd.cpu().numpy(), y_ref) < RTOL[dtype]
    # empty list: nothing happens
    check(lib.bd_stiffness_apply(bits, degree, t.nq, int(t.is_identity), ud.data_ptr(), consts.data_ptr(),
                                 yd.data_ptr(), G.data_ptr(), phi0.ctypes.data, dphi1.ctypes.data,
                                 dm_d.data_ptr(), cd.data_ptr(), 0, bc_d.data_ptr(), s))


@pytest.mark.parametrize("degree,qmode", [(3, 1), (6, 1), (5, 1), (2, 0)])
@pytest.mark.parametrize("split", ["all_l", "all_b", "one_l", "ragged"])
def test_operator_cell_list_edge_cases(degree, qmode, split):
    """lcells / bcells of any sizes (empty, one cell, not a multiple of the warp batch) give the same
    operator: the split only decides which launch a cell belongs to."""
    import torch
    import benchmark_dolfinx_b200 as bd

    class M:
        pass
    mesh = create_box_mesh(compute_mesh_size(2500, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    u = np.random.default_rng(2).uniform(-1, 1, mesh.ndofs_local)
    y_ref = OracleLaplacian(mesh, t).apply(u.copy())
    allc = np.arange(mesh.ncells, dtype=np.int32)
    k = {"all_l": mesh.ncells, "all_b": 0, "one_l": 1, "ragged": (mesh.ncells * 2) // 3 + 1}[split]
    m = M()
    for name in ("cell_dofmap", "xgeom", "geometry_dofmap", "bc_marker", "ndofs_owned"):
        setattr(m, name, getattr(mesh, name))
    m.lcells, m.bcells = allc[:k], allc[k:]
    op = bd.MatFreeLaplacianGPU(m, degree, qmode, 2.0)
    ud = torch.from_numpy(u).cuda()
    yd = torch.zeros_like(ud)
    op.apply(ud, yd)
    torch.cuda.synchronize()
    assert _rel(yd.cpu().numpy(), y_ref) < 1e-12
    assert _rel(op.geometry(), OracleLaplacian(mesh, t).G) < 1e-12
