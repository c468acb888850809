"""CPU oracle for the matrix-free Poisson hot path of benchmark-dolfinx.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline / reference
legs of ``bench.py`` may import, link or execute it, and only as the checker
(or as the timed CPU baseline), never as the thing shipped or measured as the
GPU path.  The product (``benchmark_dolfinx_b200``) never imports this package.

Contents
--------
``tables``     1-D GLL / Gauss-Legendre nodes, the interpolation table ``phi0``,
               the derivative table ``dphi1`` and the trilinear coordinate
               element tabulation -- restating what the reference obtains from
               Basix v0.10.0 (un-vendored; call sites
               ``src/laplacian.hpp:144-212``).
``boxmesh``    ``compute_mesh_size`` (``src/mesh.cpp:117-152``) and a structured
               stand-in for the DOLFINx box mesh / dofmap / ghost layer
               (``src/mesh.cpp:26-114,154-218``).
``laplacian``  numpy restatement of ``geometry_computation_{cpu,gpu}``,
               ``stiffness_operator_{cpu,gpu}`` (qmode 0 and 1), RHS assembly,
               the assembled-CSR check and ``cg_solve``.
``port.c``     the same operator in plain C (used as the timed "port" CPU
               baseline when ``oracle/_ref`` is unavailable).
``ref_shim.cpp`` + ``Makefile``  compile the reference's own, unmodified
               ``laplacian_cpu.hpp`` / ``geometry_cpu.hpp`` from where they lie
               under ``/root/reference/src`` into ``oracle/_ref/`` (never copied).

Parity pinning: the restatement is checked against every known-answer value the
reference holds for this path (``src/test_output.py:19`` and
``examples/mat_comp-16.json``; mesh sizing goldens in ``examples/*.json``) and
against ``oracle/_ref`` (the reference's own CPU kernels) in ``tests/``.
CG results, fp32, degree != 3, ``--use_gauss`` and perturbed geometry have no
reference golden: for those, parity is pinned only against ``oracle/_ref``
(qmode 0) and the matrix-free-vs-assembled-CSR self-consistency check.
"""
