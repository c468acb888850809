"""Multi-rank host logic on CPU: world_size-2 and -4 `gloo` process groups exchange halos with
the stand-in's send/receive plan and reduce dot products, and must reproduce the single-rank
operator and CG (the NCCL path of the product uses the same plan)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, degree, qmode, cg_its, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    from oracle.boxmesh import create_box_mesh
    from oracle.laplacian import OracleLaplacian, cg_solve
    from oracle.tables import Tables
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mesh = create_box_mesh(n, degree, rank=rank, nranks=world, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    op = OracleLaplacian(mesh, t)
    no = mesh.ndofs_owned

    def scatter_fwd(x):
        reqs, bufs = [], []
        for dst, idx in zip(mesh.send_ranks, mesh.send_indices):
            reqs.append(dist.isend(torch.from_numpy(np.ascontiguousarray(x[idx])), dst))
        off = no
        for src, cnt in zip(mesh.recv_ranks, mesh.recv_counts):
            buf = torch.empty(cnt, dtype=torch.float64)
            reqs.append(dist.irecv(buf, src))
            bufs.append((off, cnt, buf))
            off += cnt
        for r in reqs:
            r.wait()
        for off, cnt, buf in bufs:
            x[off:off + cnt] = buf.numpy()

    def dot(a, b):
        v = torch.tensor([float(np.dot(a[:no], b[:no]))], dtype=torch.float64)
        dist.all_reduce(v)
        return float(v.item())

    # deterministic global vector: value depends on the global lattice id only
    u = np.sin(0.37 * mesh.dof_global.astype(np.float64)) + 0.25
    u[no:] = 0
    scatter_fwd(u)
    assert np.allclose(u[no:], np.sin(0.37 * mesh.dof_global[no:].astype(np.float64)) + 0.25)
    y = op.apply(u.copy())
    x = np.zeros_like(u)

    def apply(p, out):
        scatter_fwd(p)
        op.apply(p, out)

    its = cg_solve(apply, x, np.where(mesh.bc_marker != 0, 0, u), cg_its, n_owned=no, dot=dot)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), gid=mesh.dof_global[:no], y=y[:no], x=x[:no],
             its=its, unorm=np.sqrt(dot(u, u)))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_gloo_partitioned_apply_and_cg(world, tmp_path):
    import torch.multiprocessing as mp

    from oracle.boxmesh import create_box_mesh
    from oracle.laplacian import OracleLaplacian, cg_solve
    from oracle.tables import Tables
    n, degree, qmode, cg_its = [4, 5, 3], 2, 1, 8
    mp.spawn(_worker, args=(world, _free_port(), n, degree, qmode, cg_its, str(tmp_path)), nprocs=world,
             join=True)
    m1 = create_box_mesh(n, degree, geom_perturb_fact=0.2)
    t = Tables(degree, qmode)
    op = OracleLaplacian(m1, t)
    u = np.sin(0.37 * m1.dof_global.astype(np.float64)) + 0.25
    y1 = op.apply(u.copy())
    x1 = np.zeros_like(u)
    its1 = cg_solve(op.apply, x1, np.where(m1.bc_marker != 0, 0, u), cg_its)
    pos = {int(g): i for i, g in enumerate(m1.dof_global)}
    seen = 0
    for r in range(world):
        d = np.load(tmp_path / f"rank{r}.npz")
        idx = [pos[int(g)] for g in d["gid"]]
        assert np.linalg.norm(d["y"] - y1[idx]) <= 1e-13 * np.linalg.norm(y1)
        assert np.linalg.norm(d["x"] - x1[idx]) <= 1e-11 * np.linalg.norm(x1)
        assert int(d["its"]) == its1
        assert np.isclose(float(d["unorm"]), np.linalg.norm(u), rtol=1e-14)
        seen += len(idx)
    assert seen == m1.ndofs_global
