"""Generate tests/golden/*.npz from the reference's own CPU kernels.

Run in the build container (needs oracle/_ref, i.e. /root/reference/src):
    python tests/golden/make_golden.py
Each fixture holds a seeded input vector u and y = A u computed by the UNMODIFIED
stiffness_operator_cpu / geometry_computation_cpu of the reference (strict-IEEE build), driven
with the stand-in mesh arrays and tables of oracle/.  qmode 0 only (all the reference CPU kernel
supports, src/laplacian_cpu.hpp:73-74).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import native  # noqa: E402
from oracle.boxmesh import compute_mesh_size, create_box_mesh  # noqa: E402
from oracle.laplacian import assemble_rhs  # noqa: E402
from oracle.tables import Tables  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = [  # (name, ndofs, degree, perturb, input)
    ("ci_q3_1000_rhs", 1000, 3, 0.0, "rhs"),
    ("q1_800_random_perturbed", 800, 1, 0.2, "random"),
    ("q2_900_random_perturbed", 900, 2, 0.2, "random"),
    ("q3_1500_random_perturbed", 1500, 3, 0.2, "random"),
    ("q5_1500_random", 1500, 5, 0.0, "random"),
    ("q7_1500_random_perturbed", 1500, 7, 0.1, "random"),
]


def main():
    ref = native.ref_lib(strict=True)
    if ref is None:
        raise SystemExit("oracle/_ref is not built: run `make -C oracle ref` where /root/reference exists")
    for name, ndofs, degree, perturb, kind in CASES:
        n = compute_mesh_size(ndofs, degree)
        mesh = create_box_mesh(n, degree, geom_perturb_fact=perturb)
        t = Tables(degree, 0)
        cells = np.concatenate([mesh.lcells, mesh.bcells]).astype(np.int32)
        if kind == "rhs":
            u = assemble_rhs(mesh, t)
        else:
            u = np.random.default_rng(1234).uniform(-1, 1, mesh.ndofs_local)
        G = native.ref_geometry(ref, t.nq, mesh.xgeom, mesh.geometry_dofmap, t.dphi_geometry,
                                t.weights3d, cells)
        y = np.zeros_like(u)
        native.ref_stiffness(ref, degree, u, y, G, np.full(mesh.ncells, 2.0), t.phi0, t.dphi1,
                             mesh.cell_dofmap, cells, mesh.bc_marker)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), n=np.array(n), degree=degree, qmode=0,
                            perturb=perturb, kappa=2.0, u=u, y=y, y_norm=np.linalg.norm(y))
        print(name, n, mesh.ndofs_global, np.linalg.norm(y))


if __name__ == "__main__":
    main()
