"""Generic worker for kernel variants selected by environment variables (latched per process): parity of the
operator with the oracle for the cases given on the command line as degree:qmode:dtype:ndofs (dtype f64|f32).
One JSON line per case; exit code 0 iff every case is within tolerance."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle.boxmesh import compute_mesh_size, create_box_mesh  # noqa: E402
from oracle.laplacian import OracleLaplacian  # noqa: E402
from oracle.tables import Tables as OracleTables  # noqa: E402

RTOL = {"f64": 1e-12, "f32": 1e-5}


def main():
    import torch
    import benchmark_dolfinx_b200 as bd
    ok = True
    for spec in sys.argv[1:]:
        degree, qmode, name, ndofs = spec.split(":")
        degree, qmode, ndofs = int(degree), int(qmode), int(ndofs)
        dtype = np.float64 if name == "f64" else np.float32
        mesh = create_box_mesh(compute_mesh_size(ndofs, degree), degree, geom_perturb_fact=0.2)
        oracle = OracleLaplacian(mesh, OracleTables(degree, qmode), kappa=2.0, dtype=np.float64)
        u = np.random.default_rng(degree * 1000 + qmode).uniform(-1, 1, mesh.ndofs_local)
        y_ref = oracle.apply(u.copy())
        op = bd.MatFreeLaplacianGPU(mesh, degree, qmode, 2.0, dtype=dtype)
        ud = torch.from_numpy(u.astype(dtype)).cuda()
        yd = torch.full_like(ud, 7.0)
        errs = []
        for _ in range(2):
            op.apply(ud, yd)
            torch.cuda.synchronize()
            y = yd.cpu().numpy().astype(np.float64)
            errs.append(float(np.linalg.norm(y - y_ref) / np.linalg.norm(y_ref)))
        good = max(errs) < RTOL[name]
        ok = ok and good
        print(json.dumps({"case": spec, "ncells": int(mesh.ncells), "err": errs, "kernel": op.describe(),
                          "ok": good}), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
