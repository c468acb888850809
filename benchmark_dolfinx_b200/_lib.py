"""ctypes binding of libbenchdolfinx_b200.so (the C ABI declared in include/benchdolfinx_b200.h).

The library is the product; there is no CPU fallback.  Importing this module fails
loudly when the shared library has not been built (``python -c "import
__graft_entry__ as g; g.build()"`` or ``make -C benchmark_dolfinx_b200``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbenchdolfinx_b200.so")

BD_SUCCESS = 0
BD_FLAG_BORROW_DEVICE_ARRAYS = 1
BD_FLAG_GEOMETRY_ONTHEFLY = 2
BD_FLAG_GEOMETRY_PRECOMPUTED = 4
BD_FLAG_GEOMETRY_FULL_TENSOR = 8
BD_UNIQUE_ID_BYTES = 128


class BenchDolfinxError(RuntimeError):
    """Raised for every non-zero status of the C ABI (mirrors the reference's std::runtime_error)."""

    def __init__(self, code, message):
        super().__init__(f"[bd error {code}] {message}")
        self.code = code


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build the CUDA extension first "
        "(make -C benchmark_dolfinx_b200, or __graft_entry__.build()). "
        "benchmark_dolfinx_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_i8p = C.POINTER(C.c_int8)


class LaplacianDesc(C.Structure):
    _fields_ = [
        ("scalar_bits", C.c_int32), ("degree", C.c_int32), ("nq", C.c_int32),
        ("is_identity", C.c_int32), ("kappa", C.c_double),
        ("ncells", C.c_int64), ("ndofs", C.c_int64), ("nnodes", C.c_int64),
        ("n_lcells", C.c_int64), ("n_bcells", C.c_int64),
        ("cell_dofmap", C.c_void_p), ("xgeom", C.c_void_p), ("geometry_dofmap", C.c_void_p),
        ("bc_marker", C.c_void_p), ("lcells", C.c_void_p), ("bcells", C.c_void_p),
        ("phi0", C.c_void_p), ("dphi1", C.c_void_p), ("weights", C.c_void_p),
        ("dphi_geometry", C.c_void_p), ("flags", C.c_uint32),
    ]


class BoxMeshInfo(C.Structure):
    _fields_ = [
        ("ncells_owned", C.c_int64), ("ncells", C.c_int64), ("ndofs_owned", C.c_int64),
        ("nghost", C.c_int64), ("nnodes", C.c_int64), ("n_lcells", C.c_int64),
        ("n_bcells", C.c_int64), ("ndofs_global", C.c_int64), ("ncells_global", C.c_int64),
        ("cell_dofmap", C.c_void_p), ("geometry_dofmap", C.c_void_p), ("xgeom", C.c_void_p),
        ("bc_marker", C.c_void_p), ("lcells", C.c_void_p), ("bcells", C.c_void_p),
        ("n_send", C.c_int32), ("n_recv", C.c_int32),
        ("send_ranks", C.c_void_p), ("send_counts", C.c_void_p), ("send_indices_host", C.c_void_p),
        ("recv_ranks", C.c_void_p), ("recv_counts", C.c_void_p),
    ]


# name -> (restype, argtypes).  Must list every function include/benchdolfinx_b200.h declares
# (tests/test_abi.py checks the two against each other).
_V = C.c_void_p
PROTOTYPES = {
    "bd_last_error": (C.c_char_p, []),
    "bd_debug_set_max_grid": (C.c_int, [C.c_int]),
    "bd_fp64_peak": (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "bd_device_information": (C.c_int, [C.c_char_p, C.c_size_t]),
    "bd_laplacian_create": (C.c_int, [C.POINTER(LaplacianDesc), C.POINTER(_V)]),
    "bd_laplacian_destroy": (C.c_int, [_V]),
    "bd_laplacian_apply": (C.c_int, [_V, _V, _V, _V]),
    "bd_laplacian_apply_part": (C.c_int, [_V, C.c_int, _V, _V, _V]),
    "bd_laplacian_apply_host": (C.c_int, [_V, _V, _V, _V]),
    "bd_laplacian_apply_host_batch": (C.c_int, [_V, _V, C.c_int, _V, _V]),
    "bd_laplacian_profile": (C.c_int, [_V, C.c_int]),
    "bd_laplacian_geometry_mode": (C.c_int, [_V, C.POINTER(C.c_int), C.POINTER(C.c_int64)]),
    "bd_laplacian_describe": (C.c_int, [_V, C.c_char_p, C.c_size_t]),
    "bd_laplacian_profile_read": (C.c_int, [_V, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "bd_laplacian_geometry": (C.c_int, [_V, C.POINTER(_V), C.POINTER(C.c_int64)]),
    "bd_geometry_compute": (C.c_int, [C.c_int, C.c_int, _V, _V, _V, _V, _V, C.c_int64, _V, _V]),
    "bd_stiffness_apply": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _V, _V, _V, _V, _V, _V, _V,
                                     _V, C.c_int64, _V, _V]),
    "bd_axpy": (C.c_int, [C.c_int, C.c_int64, C.c_double, _V, _V, _V, _V]),
    "bd_set_value": (C.c_int, [C.c_int, C.c_int64, C.c_double, _V, _V]),
    "bd_dot_scratch_bytes": (C.c_size_t, []),
    "bd_dot": (C.c_int, [C.c_int, C.c_int64, _V, _V, _V, _V, _V]),
    "bd_dot_static": (C.c_int, [C.c_int, C.c_int64, _V, _V, _V, _V]),
    "bd_dot_host": (C.c_int, [C.c_int, C.c_int64, _V, _V, C.POINTER(C.c_double)]),
    "bd_pack": (C.c_int, [C.c_int, C.c_int64, _V, _V, _V, _V]),
    "bd_unpack": (C.c_int, [C.c_int, C.c_int64, _V, _V, _V, _V]),
    "bd_comm_unique_id": (C.c_int, [_V]),
    "bd_comm_create": (C.c_int, [_V, C.c_int, C.c_int, C.POINTER(_V)]),
    "bd_comm_destroy": (C.c_int, [_V]),
    "bd_halo_create": (C.c_int, [_V, C.c_int, C.c_int64, C.c_int64, C.c_int, _V, _V, _V, C.c_int, _V,
                                 _V, C.POINTER(_V)]),
    "bd_halo_set_unpack": (C.c_int, [_V, _V]),
    "bd_halo_destroy": (C.c_int, [_V]),
    "bd_halo_begin": (C.c_int, [_V, _V, _V]),
    "bd_halo_end": (C.c_int, [_V, _V]),
    "bd_halo_profile": (C.c_int, [_V, C.c_int]),
    "bd_halo_profile_read": (C.c_int, [_V, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                       C.POINTER(C.c_int64)]),
    "bd_laplacian_apply_halo": (C.c_int, [_V, _V, _V, _V, _V]),
    "bd_cg_solve": (C.c_int, [_V, _V, _V, _V, _V, C.c_int, C.c_double, C.POINTER(C.c_int), _V]),
    "bd_laplacian_diagonal_inverse": (C.c_int, [_V, _V, _V]),
    "bd_pointwise_mult": (C.c_int, [C.c_int, C.c_int64, _V, _V, _V, _V]),
    "bd_cg_solve_jacobi": (C.c_int, [_V, _V, _V, _V, _V, _V, C.c_int, C.c_double, C.POINTER(C.c_int), _V]),
    "bd_inner_product": (C.c_int, [_V, C.c_int, C.c_int64, _V, _V, C.POINTER(C.c_double)]),
    "bd_norm": (C.c_int, [_V, C.c_int, C.c_int64, _V, C.POINTER(C.c_double)]),
    "bd_csr_assemble": (C.c_int, [_V, C.c_int64, C.POINTER(_V)]),
    "bd_csr_destroy": (C.c_int, [_V]),
    "bd_csr_info": (C.c_int, [_V, c_i64p, c_i64p, c_i64p]),
    "bd_csr_apply": (C.c_int, [_V, _V, _V, _V]),
    "bd_compute_mesh_size": (C.c_int, [C.c_int64, C.c_int, c_i64p]),
    "bd_boxmesh_create": (C.c_int, [C.c_int, c_i64p, C.c_int, C.c_int, C.c_int, c_i32p, C.c_double,
                                    C.POINTER(_V)]),
    "bd_boxmesh_get": (C.c_int, [_V, C.POINTER(BoxMeshInfo)]),
    "bd_boxmesh_destroy": (C.c_int, [_V]),
    "bd_boxmesh_assemble_rhs": (C.c_int, [_V, C.c_int, C.c_int, _V, _V]),
    "bd_tables": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, _V, _V, _V, _V, _V, _V,
                            C.POINTER(C.c_int)]),
}

for _name, (_res, _args) in PROTOTYPES.items():
    _fn = getattr(lib, _name)          # AttributeError here == the library lacks a declared symbol
    _fn.restype = _res
    _fn.argtypes = _args


def check(rc):
    if rc != BD_SUCCESS:
        raise BenchDolfinxError(rc, lib.bd_last_error().decode(errors="replace"))
