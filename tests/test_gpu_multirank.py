"""The partitioned operator on ONE GPU: every rank's mesh/operator is built in turn, ghost values
are filled with the halo plan (host-side emulation of the exchange; NCCL itself needs >1 GPU and
is covered by bench.py --gpus N), and the owned results must equal the single-rank operator."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _global_ids(bd, mesh, n, degree):
    """Global lattice id of every local dof, from the device dofmap (cell-local tensor order)."""
    from oracle.boxmesh import create_box_mesh
    o = create_box_mesh(n, degree, rank=mesh.rank, nranks=mesh.nranks)
    return o.dof_global


@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("degree,qmode", [(2, 1), (3, 0), (6, 1), (7, 1)])  # slab, slab, DMMA, pencil kernels
def test_partitioned_apply_equals_single_rank(nranks, degree, qmode):
    import torch
    import benchmark_dolfinx_b200 as bd
    n = [5, 4, 6]
    m1 = bd.BoxMesh(n, degree, geom_perturb_fact=0.2)
    g1 = _global_ids(bd, m1, n, degree)
    u1 = np.sin(0.37 * g1.astype(np.float64)) + 0.25
    op1 = bd.MatFreeLaplacianGPU(m1, degree, qmode, 2.0)
    y1d = torch.zeros(m1.ndofs_local, dtype=torch.float64, device="cuda")
    op1.apply(torch.from_numpy(u1).cuda(), y1d)
    y1 = y1d.cpu().numpy()
    pos = {int(g): i for i, g in enumerate(g1)}
    meshes = [bd.BoxMesh(n, degree, r, nranks, geom_perturb_fact=0.2) for r in range(nranks)]
    gids = [_global_ids(bd, m, n, degree) for m in meshes]
    hosts = [m.host_arrays() for m in meshes]
    vecs = []
    for m, g in zip(meshes, gids):
        v = np.zeros(m.ndofs_local)
        v[:m.ndofs_owned] = np.sin(0.37 * g[:m.ndofs_owned].astype(np.float64)) + 0.25
        vecs.append(v)
    for r, (m, h) in enumerate(zip(meshes, hosts)):      # forward scatter via the plans
        off = m.ndofs_owned
        for src, cnt in zip(h["recv_ranks"], h["recv_counts"]):
            hs = hosts[src]
            k = list(hs["send_ranks"]).index(r)
            start = int(hs["send_counts"][:k].sum())
            idx = hs["send_indices"][start:start + int(hs["send_counts"][k])]
            assert len(idx) == cnt
            vecs[r][off:off + cnt] = vecs[src][idx]
            off += cnt
    total = 0
    for r, (m, g) in enumerate(zip(meshes, gids)):
        assert np.allclose(vecs[r], np.sin(0.37 * g.astype(np.float64)) + 0.25)
        op = bd.MatFreeLaplacianGPU(m, degree, qmode, 2.0)
        yd = torch.zeros(m.ndofs_local, dtype=torch.float64, device="cuda")
        op.apply(torch.from_numpy(vecs[r]).cuda(), yd)
        y = yd.cpu().numpy()[:m.ndofs_owned]
        ref = y1[[pos[int(v)] for v in g[:m.ndofs_owned]]]
        assert np.linalg.norm(y - ref) <= 1e-13 * np.linalg.norm(y1)
        total += m.ndofs_owned
    assert total == m1.ndofs_global


def test_single_rank_comm_and_halo_are_no_ops():
    import torch
    import benchmark_dolfinx_b200 as bd
    comm = bd.Comm(0, 1)
    mesh = bd.BoxMesh([3, 3, 3], 2)
    halo = bd.Halo(comm, mesh)
    op = bd.MatFreeLaplacianGPU(mesh, 2, 1, 2.0, halo=halo)
    op0 = bd.MatFreeLaplacianGPU(mesh, 2, 1, 2.0)
    u = mesh.assemble_rhs(1)
    y, y0 = torch.zeros_like(u), torch.zeros_like(u)
    op.apply(u, y)
    op0.apply(u, y0)
    x, x0 = torch.zeros_like(u), torch.zeros_like(u)
    assert bd.cg_solve(op, x, u, 7, comm=comm) == bd.cg_solve(op0, x0, u, 7) == 7
    torch.cuda.synchronize()
    assert torch.equal(x, x0) or (x - x0).abs().max() <= 1e-12 * x0.abs().max()
    assert (y - y0).abs().max() <= 1e-13 * y0.abs().max()
    assert abs(bd.norm(u, mesh.ndofs_owned, comm) - bd.norm(u)) == 0
