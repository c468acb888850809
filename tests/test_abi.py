"""The C-ABI library loads and exports every symbol include/*.h declares (CPU only: no compute)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "benchdolfinx_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(bd_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_functions():
    names = declared_functions()
    assert "bd_laplacian_apply" in names and "bd_cg_solve" in names and len(names) >= 30


def test_library_exports_every_declared_symbol():
    from benchmark_dolfinx_b200 import _lib
    exported = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True,
                              text=True, check=True).stdout
    exported = set(re.findall(r" T (bd_[a-z0-9_]+)", exported))
    for name in declared_functions():
        assert name in exported, f"{name} declared in the header but not exported"
        assert hasattr(_lib.lib, name)
        assert name in _lib.PROTOTYPES, f"{name} has no ctypes prototype"
    assert set(_lib.PROTOTYPES) <= set(declared_functions())


def test_no_torch_types_in_the_abi():
    text = open(HEADER).read()
    assert "torch" not in text.lower().replace("no c++, cuda or\n * torch types", "")
    assert "at::" not in text and "#include <cuda" not in text


def test_struct_layouts_match_the_header():
    from benchmark_dolfinx_b200 import _lib
    # compile a tiny C program printing sizeof/offsetof and compare with ctypes
    src = r'''
#include <stdio.h>
#include <stddef.h>
#include "benchdolfinx_b200.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu\n", sizeof(bd_laplacian_desc), offsetof(bd_laplacian_desc, kappa),
         offsetof(bd_laplacian_desc, cell_dofmap), offsetof(bd_laplacian_desc, flags),
         sizeof(bd_boxmesh_info));
  return 0; }'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), c, "-o", exe], check=True)
        out = subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()
    D = _lib.LaplacianDesc
    assert [int(v) for v in out] == [ctypes.sizeof(D), D.kappa.offset, D.cell_dofmap.offset,
                                     D.flags.offset, ctypes.sizeof(_lib.BoxMeshInfo)]


def test_fails_loudly_without_a_gpu():
    """No CPU fallback: constructing the operator without a device is an error, not a slow path."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import benchmark_dolfinx_b200 as bd
    from oracle.boxmesh import create_box_mesh
    mesh = create_box_mesh([2, 2, 2], 2)
    with pytest.raises(bd.BenchDolfinxError) as e:
        bd.MatFreeLaplacianGPU(mesh, 2, 1, 2.0)
    assert e.value.code == -5 and "no CPU fallback" in str(e.value)
    with pytest.raises(bd.BenchDolfinxError):
        bd.device_information()


def test_argument_errors_mirror_the_reference():
    """src/laplacian.hpp:378,400 ("Unsupported degree [operator]", "Unsupported nq"),
    src/main.cpp:212-213,226-227 (float size, qmode)."""
    import benchmark_dolfinx_b200 as bd
    from benchmark_dolfinx_b200 import _lib
    with pytest.raises(bd.BenchDolfinxError, match="Unsupported degree"):
        bd.Tables(8, 1)
    with pytest.raises(bd.BenchDolfinxError, match="Invalid qmode"):
        bd.Tables(3, 2)
    with pytest.raises(ValueError, match="Invalid float size"):
        bd.Tables(3, 1, dtype=np.float16)
    with pytest.raises(bd.BenchDolfinxError, match="Expecting identity matrix for qmode=0"):
        bd.Tables(3, 0, use_gauss=True)          # src/laplacian.hpp:197-198
    rc = _lib.lib.bd_stiffness_apply(64, 9, 10, 0, 1, 1, 1, 1, 1, 1, 1, 1, 5, 1, None)
    assert rc == -2 and b"Unsupported degree" in _lib.lib.bd_last_error()
    rc = _lib.lib.bd_stiffness_apply(64, 3, 7, 0, 1, 1, 1, 1, 1, 1, 1, 1, 5, 1, None)
    assert rc == -2 and b"Unsupported nq" in _lib.lib.bd_last_error()
    rc = _lib.lib.bd_stiffness_apply(16, 3, 4, 1, 1, 1, 1, 1, 1, 1, 1, 1, 5, 1, None)
    assert rc == -1 and b"Invalid float size" in _lib.lib.bd_last_error()


def test_no_ldgsts_with_uniform_register_offset():
    """ptxas 12.9 can fold a rolled-loop slot offset into a uniform-register operand of LDGSTS (cp.async
    with an L2 cache hint) and emit `LDGSTS [R+UR0+imm], desc[UR1]`, which faults at run time (seen in
    csrc/stiffness_pencil.cuh for nd = 5, nq = 6, fp64).  The shipped library must not contain that form."""
    import shutil
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    from benchmark_dolfinx_b200 import _lib
    sass = subprocess.run([cuobjdump, "-sass", _lib.LIB_PATH], capture_output=True, text=True, timeout=600).stdout
    copies = [l for l in sass.splitlines() if "LDGSTS" in l]
    assert copies, "expected cp.async copies in the pencil kernel"
    bad = [l.strip() for l in copies if re.search(r"\[R\d+\+UR\d+", l)]
    assert not bad, bad[:4]
