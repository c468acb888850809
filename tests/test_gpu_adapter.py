"""The DOLFINx adapter header (benchmark_dolfinx_b200/host/dolfinx_adapter.hpp) EXECUTED, not just type-checked.

DOLFINx, Basix and MPI cannot be installed here, so the header runs against functional stand-ins
(tests/mock_dolfinx/functional/): DOLFINx classes holding the arrays of the library's structured stand-in mesh,
the Basix calls answered by bd_tables, MPI as threads-as-ranks.  What is checked:
  * `compute_halo_lists` (the MPI_Alltoall(v) part of the halo plan, src/vector.hpp:95-149 in the reference's terms):
    2, 4 and 8 ranks on one GPU (no NCCL involved) -- every entry an owner sends lands on the ghost position that
    expects that global index, also with the IndexMap's ghosts in a shuffled order (unpack path);
  * constructor -> apply -> cg_solve through the adapter reproduce the reference's known answers
    (src/test_output.py:19, examples/mat_comp-16.json, and the oracle-pinned CG value), on 1 rank and -- when the box
    has the GPUs -- on 2 ranks over NCCL.
"""
import json
import os
import shutil
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def driver(tmp_path_factory):
    d = tmp_path_factory.mktemp("adapter")
    func = os.path.join(ROOT, "tests", "mock_dolfinx", "functional")
    shutil.copy(os.path.join(ROOT, "benchmark_dolfinx_b200", "host", "dolfinx_adapter.hpp"), d)
    shutil.copy(os.path.join(func, "mesh.hpp"), d)          # as the header sits next to the reference's src/mesh.hpp
    exe = str(d / "run_adapter")
    libdir = os.path.join(ROOT, "benchmark_dolfinx_b200")
    cmd = ["g++", "-std=c++20", "-O1", "-Wall", "-Werror", f"-I{d}", f"-I{func}", f"-I{os.path.join(ROOT, 'include')}",
           "-I/usr/local/cuda/include", "-o", exe, os.path.join(func, "run_adapter.cpp"), f"-L{libdir}",
           "-lbenchdolfinx_b200", "-L/usr/local/cuda/lib64", "-lcudart", "-lpthread", f"-Wl,-rpath,{libdir}"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-4000:]
    return exe


def _run(exe, *args, ok=(0,)):
    r = subprocess.run([exe, *map(str, args)], capture_output=True, text=True, timeout=600)
    assert r.returncode in ok, (r.stdout[-2000:], r.stderr[-2000:])
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    return r.returncode, lines[-1]


@pytest.mark.parametrize("scramble", [False, True])
@pytest.mark.parametrize("nranks,degree", [(2, 3), (4, 2), (8, 3), (8, 1)])
def test_halo_lists_from_index_map(driver, nranks, degree, scramble):
    _, out = _run(driver, "lists", nranks, 30000, degree, *(["scramble"] if scramble else []))
    assert out["bad"] == 0 and out["entries_checked"] > 0


def test_adapter_reproduces_reference_known_answers_one_rank(driver):
    _, out = _run(driver, "solve", 1, 1000, 3, 0, 0)                 # src/test_output.py:19
    assert abs(out["y_norm"] - 9.912865833415553) <= 1e-12 * 9.912865833415553
    _, out = _run(driver, "solve", 1, 100000, 3, 1, 0)               # examples/mat_comp-16.json
    assert abs(out["u_norm"] - 0.6895773850559623) <= 1e-12 * 0.6895773850559623
    assert abs(out["y_norm"] - 0.14150257625641838) <= 1e-12 * 0.14150257625641838
    _, out = _run(driver, "solve", 1, 10000, 3, 1, 50)               # CG value pinned by the oracle
    assert out["iterations"] == 50 and abs(out["x_norm"] - 167.59244725655134) <= 1e-11 * 167.59244725655134


def test_adapter_two_ranks_over_nccl(driver):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _, out = _run(driver, "solve", 2, 100000, 3, 1, 0)
    assert abs(out["y_norm"] - 0.14150257625641838) <= 1e-12 * 0.14150257625641838
    _, out = _run(driver, "solve", 2, 10000, 3, 1, 50)
    assert out["iterations"] == 50 and abs(out["x_norm"] - 167.59244725655134) <= 1e-11 * 167.59244725655134
