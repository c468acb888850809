"""The FFMA2 phase functions of the fp32 slab kernel (csrc/stiffness_slab_f2.cuh) against the scalar phase
functions, on the host: tests/host/phases_check.cu is compiled with -DBD_HOST_TEST (the per-lane phase
functions become __host__ __device__) and run on the CPU.  No GPU involved."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ffma2_phases_equal_scalar_phases_on_host(tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "phases_check")
    cmd = [nvcc, "-std=c++17", "-O1", "-DBD_HOST_TEST", "--expt-relaxed-constexpr",
           "-gencode", "arch=compute_100a,code=sm_100a",
           "-I", os.path.join(ROOT, "benchmark_dolfinx_b200", "csrc"), "-I", os.path.join(ROOT, "include"),
           "-o", exe, os.path.join(ROOT, "tests", "host", "phases_check.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip().endswith("OK"), r.stdout[-3000:]
