"""Worker for tests/test_gpu_nccl.py: run under torchrun on N GPUs of one node.

Every rank builds its block of a small global mesh, the ranks exchange halos and reduce dot
products through the library's NCCL path, and the *global* norms must equal the reference's own
known answers (examples/mat_comp-16.json) which were produced on other
partitionings -- norms are partition independent."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    import benchmark_dolfinx_b200 as bd
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    comm = bd.Comm.from_torch_distributed()
    out = []
    cases = [  # ndofs_global, degree, qmode, cg iterations, expected u_norm, expected y_norm
        (100000, 3, 1, 0, 0.6895773850559623, 0.14150257625641838),   # examples/mat_comp-16.json
        (10000, 3, 1, 50, 2.2790421417877087, 167.59244725655134),
        (5000, 6, 1, 20, 4.736123103582485, 90.15797406544854),
        (10000, 3, 0, 0, 2.2794459800498785, 1.1590666300856363),
    ]
    for ndofs, degree, qmode, its, unorm, ynorm in cases:
        n = bd.compute_mesh_size(ndofs, degree)
        pg = bd.partition_grid(world)
        if any(n[d] < pg[d] for d in range(3)):
            continue
        mesh = bd.BoxMesh(n, degree, rank, world)
        res, op, u, y = bd.laplace_action_gpu(mesh, degree, qmode, 2.0, max(its, 1), use_cg=its > 0, comm=comm)
        out.append(dict(case=[ndofs, degree, qmode, its], unorm=res.unorm, ynorm=res.ynorm,
                        unorm_ref=unorm, ynorm_ref=ynorm))
        assert abs(res.unorm - unorm) <= 1e-12 * unorm, (rank, out[-1])
        assert abs(res.ynorm - ynorm) <= 1e-11 * ynorm, (rank, out[-1])
    # both geometry variants through the halo path, with the exchange timing switched on
    n = bd.compute_mesh_size(100000, 3)
    mesh = bd.BoxMesh(n, 3, rank, world)
    halo = bd.Halo(comm, mesh)
    u = mesh.assemble_rhs(1)
    norms = {}
    for geo in ("precomputed", "onthefly", "onthefly-full"):
        op = bd.MatFreeLaplacianGPU(mesh, 3, 1, 2.0, halo=halo, geometry=geo)
        y = torch.zeros_like(u)
        halo.profile(True)
        for _ in range(3):
            op.apply(u, y)
        torch.cuda.synchronize()
        rec = halo.profile_read()
        halo.profile(False)
        assert rec["applications"] == 3 and rec["exchange_ms"] >= 0 and rec["exposed_ms"] >= 0, (rank, rec)
        norms[geo] = bd.norm(y, mesh.ndofs_owned, comm)
        assert abs(norms[geo] - 0.14150257625641838) <= 1e-12 * 0.14150257625641838, (rank, geo, norms)
        # Jacobi-preconditioned CG: every rank must report the same iteration count and a converged residual
        dinv = op.get_diag_inverse()
        x = torch.zeros_like(u)
        its = bd.cg_solve(op, x, u, 200, 1e-8, comm, diag_inv=dinv)
        t = torch.tensor([its], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert int(t.item()) == its and 0 < its < 200, (rank, its)
    # ghosts in scrambled order: the same exchange through the unpack path (bd_halo_set_unpack)
    import numpy as np
    mesh = bd.BoxMesh(bd.compute_mesh_size(20000, 3), 3, rank, world)
    halo = bd.Halo(comm, mesh)
    x = torch.zeros(mesh.ndofs_local, dtype=torch.float64, device="cuda")
    x[:mesh.ndofs_owned] = torch.arange(mesh.ndofs_owned, dtype=torch.float64, device="cuda") + 1000.0 * rank
    ref = x.clone()
    halo.scatter_fwd(ref)
    perm = np.random.default_rng(rank).permutation(mesh.nghost).astype(np.int32)
    halo.set_unpack(perm)
    y = x.clone()
    halo.scatter_fwd(y)
    torch.cuda.synchronize()
    expect = ref.clone()
    expect[mesh.ndofs_owned + torch.from_numpy(perm.astype(np.int64)).cuda()] = ref[mesh.ndofs_owned:]
    assert torch.equal(y, expect), rank
    halo.set_unpack(None)
    if rank == 0:
        print("MULTI_GPU_CHECK " + json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
