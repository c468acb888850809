"""The DOLFINx adapter header (benchmark_dolfinx_b200/host/dolfinx_adapter.hpp) type-checks against
API mock-ups of MPI / Basix / DOLFINx v0.10 (none of which can be installed here).  CPU only.
The header is copied next to the mock of the reference's mesh.hpp, as it would sit in the
reference's src/ directory."""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_adapter_instantiates_for_float_and_double(tmp_path):
    mock = os.path.join(ROOT, "tests", "mock_dolfinx")
    shutil.copy(os.path.join(ROOT, "benchmark_dolfinx_b200", "host", "dolfinx_adapter.hpp"), tmp_path)
    shutil.copy(os.path.join(mock, "mesh.hpp"), tmp_path)
    shutil.copy(os.path.join(mock, "use_adapter.cpp"), tmp_path)
    cuda_inc = "/usr/local/cuda/include"
    cmd = ["g++", "-std=c++20", "-fsyntax-only", "-Wall", "-Werror", f"-I{tmp_path}", f"-I{mock}",
           f"-I{os.path.join(ROOT, 'include')}", f"-I{cuda_inc}", os.path.join(tmp_path, "use_adapter.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-4000:]
