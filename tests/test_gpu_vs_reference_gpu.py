"""Our CUDA path against the reference's OWN GPU kernels (stiffness_operator_gpu,
geometry_computation_gpu: unmodified sources recompiled for sm_100a into oracle/_ref/libref_gpu.so)
on the same device arrays: same G, same operator result, every degree / qmode / scalar type."""
import numpy as np
import pytest

from oracle import native
from oracle.boxmesh import compute_mesh_size, create_box_mesh
from oracle.tables import Tables

pytestmark = pytest.mark.gpu
RTOL = {np.float64: 1e-12, np.float32: 1e-5}


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / np.linalg.norm(b)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("qmode", [0, 1])
@pytest.mark.parametrize("degree", [1, 2, 3, 4, 5, 6, 7])
def test_operator_matches_reference_gpu_kernels(degree, qmode, dtype):
    import torch
    import benchmark_dolfinx_b200 as bd
    ref = native.ref_gpu_lib()
    if ref is None:
        pytest.skip("oracle/_ref/libref_gpu.so not built (reference sources absent at build time)")
    bits = 64 if dtype == np.float64 else 32
    mesh = create_box_mesh(compute_mesh_size(4000, degree), degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    cells = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int32)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    xg, gd, dm, bc, cd = (dev(mesh.xgeom.astype(dtype)), dev(mesh.geometry_dofmap), dev(mesh.cell_dofmap),
                          dev(mesh.bc_marker), dev(cells))
    dphi, w = dev(t.dphi_geometry.astype(dtype)), dev(t.weights3d.astype(dtype))
    phi0, dphi1 = dev(t.phi0.astype(dtype)), dev(t.dphi1.astype(dtype))
    nq3 = t.nq ** 3
    G_ref = torch.zeros(len(cells) * 6 * nq3, dtype=xg.dtype, device="cuda")
    native.ref_gpu_geometry(ref, t.nq, xg.data_ptr(), G_ref.data_ptr(), gd.data_ptr(), dphi.data_ptr(),
                            w.data_ptr(), cd.data_ptr(), len(cells), bits)
    u = np.random.default_rng(1234).uniform(-1, 1, mesh.ndofs_local).astype(dtype)
    ud = dev(u)
    y_ref = torch.zeros_like(ud)
    consts = torch.full((len(cells),), 2.0, dtype=xg.dtype, device="cuda")
    native.ref_gpu_stiffness(ref, degree, t.nq, t.is_identity, ud.data_ptr(), y_ref.data_ptr(), G_ref.data_ptr(),
                             consts.data_ptr(), phi0.data_ptr(), dphi1.data_ptr(), dm.data_ptr(), cd.data_ptr(),
                             len(cells), bc.data_ptr(), bits)
    torch.cuda.synchronize()
    op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
    assert _rel(op.geometry().ravel(), G_ref.cpu().numpy()) < RTOL[dtype]
    y = torch.full_like(ud, 9.0)
    op.apply(ud, y)
    torch.cuda.synchronize()
    assert _rel(y.cpu().numpy(), y_ref.cpu().numpy()) < RTOL[dtype]
