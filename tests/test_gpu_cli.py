"""The bench_dolfinx command line (reference src/main.cpp) end to end, checked exactly as the
reference's CI checks it (src/test_output.py:15-19, .github/workflows/ci.yml:105)."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "benchmark_dolfinx_b200", "bench_dolfinx")


def _run(args, tmp_path):
    out = tmp_path / "a.json"
    r = subprocess.run([EXE] + args + ["--json", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    return json.load(open(out)), r.stdout


def test_ci_command_and_test_output_py(tmp_path):
    data, stdout = _run(["--ndofs=1000", "--degree=3", "--qmode=0", "--nreps=1", "--mat_comp", "--float=64"],
                        tmp_path)
    # src/test_output.py:15-19
    assert data["output"]["ndofs_global"] == 1000
    assert np.isclose(data["output"]["y_norm"], data["output"]["z_norm"])
    assert np.isclose(data["output"]["y_norm"], 9.912865833415553)
    assert abs(data["output"]["y_norm"] - 9.912865833415553) < 1e-12 * 9.912865833415553
    # same keys as the reference's JSON (examples/mat_comp-16.json)
    assert set(data["input"]) == {"p", "mpi_size", "ndofs_local_requested", "nreps", "scalar_size", "use_gauss",
                                  "mat_comp", "qmode", "cg"}
    assert set(data["output"]) == {"ncells_global", "ndofs_global", "mat_free_time", "u_norm", "y_norm", "z_norm",
                                   "gdof_per_second"}
    for line in ("Computation time (Action):", "Computation rate (Gdofs/s):", "Norm of u =", "Norm of y =",
                 "Norm of z =", "Norm of error =", "Relative norm of error ="):
        assert line in stdout


def test_example_mat_comp_16(tmp_path):
    """examples/mat_comp-16.json: --nreps=1 --mat_comp --ndofs_global=100000 --degree=3 (qmode 1)."""
    data, _ = _run(["--nreps=1", "--mat_comp", "--ndofs_global=100000", "--degree=3"], tmp_path)
    o = data["output"]
    assert (o["ncells_global"], o["ndofs_global"]) == (3468, 100048)
    assert abs(o["u_norm"] - 0.6895773850559623) < 1e-12
    assert abs(o["y_norm"] - 0.14150257625641838) < 1e-13
    assert abs(o["z_norm"] - 0.14150257625641852) < 1e-13


def test_cg_fp32_and_errors(tmp_path):
    data, stdout = _run(["--ndofs_global=10000", "--degree=3", "--cg", "--nreps=10"], tmp_path)
    assert data["input"]["cg"] is True and "Computation time (CG):" in stdout
    assert abs(data["output"]["y_norm"] - 129.5221707929043) < 1e-9
    data, _ = _run(["--ndofs_global=10000", "--degree=3", "--float=32", "--nreps=2"], tmp_path)
    assert abs(data["output"]["y_norm"] - 1.1415775083651323) < 1e-4
    for bad in (["--float=16"], ["--qmode=2"], ["--degree=9"], ["--platform=cpu"], ["--ndofs=10", "--ndofs_global=10"],
                ["--degree=3", "--qmode=0", "--use_gauss"]):
        r = subprocess.run([EXE] + bad, capture_output=True, text=True, timeout=120)
        assert r.returncode != 0, bad
