"""ctypes access to the oracle's native checkers (TEST INFRASTRUCTURE).

* ``ref``    -- oracle/_ref/libref_cpu{,_strict}.so: the reference's own unmodified CPU
               kernels (src/laplacian_cpu.hpp, src/geometry_cpu.hpp) behind oracle/ref_shim.cpp
* ``port``   -- oracle/libport.so: the plain-C restatement oracle/port.c
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


def ref_lib(strict=False):
    return _load(os.path.join(_HERE, "_ref", "libref_cpu_strict.so" if strict else "libref_cpu.so"))


def ref_gpu_lib():
    """oracle/_ref/libref_gpu.so: the reference's unmodified GPU kernels (src/laplacian_gpu.hpp,
    src/geometry_gpu.hpp) recompiled for sm_100a behind oracle/ref_gpu_shim.cu."""
    return _load(os.path.join(_HERE, "_ref", "libref_gpu.so"))


def port_lib():
    return _load(os.path.join(_HERE, "libport.so"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _bits(dtype):
    return 64 if np.dtype(dtype) == np.float64 else 32


def ref_geometry(lib, nq, xgeom, geometry_dofmap, dphi, weights, cells):
    """geometry_computation_cpu<T,Q> (src/geometry_cpu.hpp:25-112); returns G[ncells, 6, nq^3]."""
    dt = xgeom.dtype
    G = np.zeros((len(cells), 6, nq ** 3), dtype=dt)
    cells = np.ascontiguousarray(cells, dtype=np.int32)
    rc = lib.ref_geometry_computation_cpu(
        C.c_int(_bits(dt)), C.c_int(nq), _p(xgeom), C.c_size_t(xgeom.size), _p(G),
        _p(geometry_dofmap), C.c_size_t(geometry_dofmap.size), _p(dphi), _p(weights), _p(cells),
        C.c_int(len(cells)))
    if rc != 0:
        raise RuntimeError("reference geometry kernel: unsupported configuration")
    return G


def ref_stiffness(lib, degree, u, y, G, constants, phi0, dphi1, cell_dofmap, cells, bc_marker):
    """stiffness_operator_cpu<T,P,P+1> (src/laplacian_cpu.hpp:57-146), qmode 0 only; y += A u."""
    cells = np.ascontiguousarray(cells, dtype=np.int32)
    rc = lib.ref_stiffness_operator_cpu(
        C.c_int(_bits(u.dtype)), C.c_int(degree), _p(u), _p(constants), _p(y), _p(G), _p(phi0),
        _p(dphi1), _p(cell_dofmap), _p(cells), C.c_int(len(cells)), _p(bc_marker))
    if rc != 0:
        raise RuntimeError("reference stiffness kernel: unsupported configuration")
    return y


def port_stiffness(lib, nd, nq, identity, u, y, G, constants, phi0, dphi1, cell_dofmap, cells,
                   bc_marker):
    fn = lib.port_stiffness_f64 if u.dtype == np.float64 else lib.port_stiffness_f32
    cells = np.ascontiguousarray(cells, dtype=np.int32)
    fn.restype = None
    fn(C.c_int(nd), C.c_int(nq), C.c_int(int(identity)), _p(u), _p(constants), _p(y), _p(G),
       _p(phi0), _p(dphi1), _p(cell_dofmap), _p(cells), C.c_int(len(cells)), _p(bc_marker))
    return y


def port_geometry(lib, nq, xgeom, geometry_dofmap, dphi, weights, cells):
    dt = xgeom.dtype
    fn = lib.port_geometry_f64 if dt == np.float64 else lib.port_geometry_f32
    fn.restype = None
    G = np.zeros((len(cells), 6, nq ** 3), dtype=dt)
    cells = np.ascontiguousarray(cells, dtype=np.int32)
    fn(C.c_int(nq), _p(xgeom), _p(G), _p(geometry_dofmap), _p(dphi), _p(weights), _p(cells),
       C.c_int(len(cells)))
    return G


def ref_gpu_stiffness(lib, degree, nq, identity, u, y, G, constants, phi0, dphi1, cell_dofmap, cells, ncells,
                      bc_marker, bits):
    """Launch stiffness_operator_gpu<T,P,Q> (src/laplacian_gpu.hpp:91-426) with the reference's launch
    shape; every array argument is a device pointer (int); y += A u on the default stream."""
    rc = lib.ref_stiffness_operator_gpu(
        C.c_int(bits), C.c_int(degree), C.c_int(nq), C.c_void_p(u), C.c_void_p(constants), C.c_void_p(y),
        C.c_void_p(G), C.c_void_p(phi0), C.c_void_p(dphi1), C.c_void_p(cell_dofmap), C.c_void_p(cells),
        C.c_int(ncells), C.c_void_p(bc_marker), C.c_int(int(identity)))
    if rc != 0:
        raise RuntimeError(f"reference GPU stiffness kernel failed with {rc}")


def ref_gpu_geometry(lib, nq, xgeom, G, geometry_dofmap, dphi, weights, cells, ncells, bits):
    """Launch geometry_computation_gpu<T,Q> (src/geometry_gpu.hpp:26-132); device pointers."""
    rc = lib.ref_geometry_computation_gpu(
        C.c_int(bits), C.c_int(nq), C.c_void_p(xgeom), C.c_void_p(G), C.c_void_p(geometry_dofmap),
        C.c_void_p(dphi), C.c_void_p(weights), C.c_void_p(cells), C.c_int(ncells))
    if rc != 0:
        raise RuntimeError(f"reference GPU geometry kernel failed with {rc}")
