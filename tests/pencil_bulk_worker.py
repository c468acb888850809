"""Parity of the experimental bulk-copy pencil kernel (BD_PENCIL_BULK=1: fp32, degree 6, qmode 1) with the
oracle, with and without FFMA2 (BD_PENCIL_PAIRED).  One JSON line per case; exit code 0 iff all pass."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle.boxmesh import compute_mesh_size, create_box_mesh  # noqa: E402
from oracle.laplacian import OracleLaplacian  # noqa: E402
from oracle.tables import Tables as OracleTables  # noqa: E402


def main():
    import torch
    import benchmark_dolfinx_b200 as bd
    ok = True
    degree = 6
    for ndofs in (4000, 300000, 1500000):   # 1.5M dofs: more cells than resident CTAs (persistent loop)
        mesh = create_box_mesh(compute_mesh_size(ndofs, degree), degree, geom_perturb_fact=0.2)
        oracle = OracleLaplacian(mesh, OracleTables(degree, 1), kappa=2.0, dtype=np.float64)
        u = np.random.default_rng(ndofs).uniform(-1, 1, mesh.ndofs_local)
        y_ref = oracle.apply(u.copy())
        op = bd.MatFreeLaplacianGPU(mesh, degree, 1, 2.0, dtype=np.float32)
        ud = torch.from_numpy(u.astype(np.float32)).cuda()
        yd = torch.full_like(ud, 7.0)
        errs = []
        for _ in range(2):
            op.apply(ud, yd)
            torch.cuda.synchronize()
            y = yd.cpu().numpy().astype(np.float64)
            errs.append(float(np.linalg.norm(y - y_ref) / np.linalg.norm(y_ref)))
        good = max(errs) < 1e-5
        ok = ok and good
        print(json.dumps({"ndofs": int(mesh.ndofs_local), "ncells": int(mesh.ncells), "err": errs, "ok": good}),
              flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
