"""benchmark_dolfinx_b200 -- B200-native matrix-free Poisson hot path of benchmark-dolfinx.

Host-side mirror (Python) of the reference's operator interface for this path, sitting on the
C ABI of ``libbenchdolfinx_b200.so`` (``include/benchdolfinx_b200.h``):

===============================  ==========================================================
here                             reference
===============================  ==========================================================
``MatFreeLaplacianGPU.apply``    ``MatFreeLaplacianGPU<T>::apply``   src/laplacian.hpp:355-403
``cg_solve``                     ``cg_solve``                        src/cg.hpp:89-169
``axpy / inner_product / norm    ``benchdolfinx::axpy ...``          src/vector.hpp:159-292
/ set_value``
``BoxMesh``                      ``create_mesh`` + DOLFINx dofmap    src/mesh.cpp:190-218
``compute_mesh_size``            ``compute_mesh_size``               src/mesh.cpp:117-152
``Tables``                       Basix tables in the constructor     src/laplacian.hpp:123-212
``laplace_action_gpu``           ``laplace_action_gpu``              src/laplacian_solver.cpp:66-230
===============================  ==========================================================

Vectors are ``torch`` CUDA tensors (device memory + streams only; every kernel is ours).
There is no CPU fallback: without the CUDA library the import fails, without a GPU the
constructors raise ``BenchDolfinxError``.
"""
from __future__ import annotations

import ctypes as C
import math
import time
from dataclasses import dataclass

import numpy as np

from . import _lib
from ._lib import BenchDolfinxError, check, lib

__all__ = [
    "BenchDolfinxError", "Tables", "BoxMesh", "Comm", "Halo", "MatFreeLaplacianGPU", "MatrixOperator",
    "cg_solve",
    "axpy", "inner_product", "norm", "set_value", "pointwise_mult", "compute_mesh_size", "partition_grid",
    "laplace_action_gpu", "BenchmarkResults", "device_information",
]


def _bits(dtype) -> int:
    """32 or 64 for a numpy / python / torch floating type (the reference's --float)."""
    name = str(dtype)
    if name.startswith("torch."):
        name = name[6:]
    else:
        try:
            name = np.dtype(dtype).name
        except TypeError:
            pass
    if name == "float64":
        return 64
    if name == "float32":
        return 32
    raise ValueError("Invalid float size. Must be 32 or 64.")


def _np_dtype(bits):
    return np.float64 if bits == 64 else np.float32


def _torch_dtype(bits):
    import torch
    return torch.float64 if bits == 64 else torch.float32


def _ptr(a):
    """Raw pointer of a numpy array (host), a torch tensor (device) or an int."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data
    return a.data_ptr()


def _stream_ptr(stream):
    if stream is None:
        import torch
        return torch.cuda.current_stream().cuda_stream
    return stream if isinstance(stream, int) else stream.cuda_stream


def device_information() -> str:
    buf = C.create_string_buffer(1024)
    check(lib.bd_device_information(buf, 1024))
    return buf.value.decode()


def debug_set_max_grid(max_ctas: int) -> None:
    """Test hook: cap the grid of the persistent stiffness kernels (0 = no cap)."""
    check(lib.bd_debug_set_max_grid(int(max_ctas)))


def fp64_peak():
    """(DFMA, DMMA) TFLOP/s measured on the current device."""
    a, b = C.c_double(), C.c_double()
    check(lib.bd_fp64_peak(C.byref(a), C.byref(b)))
    return a.value, b.value


def compute_mesh_size(ndofs_global: int, degree: int):
    """src/mesh.cpp:117-152."""
    out = (C.c_int64 * 3)()
    check(lib.bd_compute_mesh_size(int(ndofs_global), int(degree), out))
    return [int(out[0]), int(out[1]), int(out[2])]


def partition_grid(nranks: int):
    """Block partition of the box over the ranks (one rank per GPU)."""
    table = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}
    if nranks in table:
        return table[nranks]
    p, n, f, fs = [1, 1, 1], nranks, 2, []
    while n > 1:
        while n % f == 0:
            fs.append(f)
            n //= f
        f += 1
    for f in sorted(fs, reverse=True):
        p[p.index(min(p))] *= f
    return tuple(sorted(p, reverse=True))


class Tables:
    """1-D tables of the operator (what the reference takes from Basix, src/laplacian.hpp:123-212)."""

    def __init__(self, degree: int, qmode: int, use_gauss: bool = False, dtype=np.float64):
        bits = _bits(dtype)
        self.degree, self.qmode, self.use_gauss = degree, qmode, bool(use_gauss)
        nd, nq = degree + 1, degree + 1 + qmode
        dt = _np_dtype(bits)
        self.nd, self.nq, self.dtype = nd, nq, np.dtype(dt)
        self.phi0 = np.zeros((nq, max(nd, 1)), dtype=dt)
        self.dphi1 = np.zeros((nq, nq), dtype=dt)
        self.weights3d = np.zeros(nq ** 3, dtype=dt)
        self.dphi_geometry = np.zeros((3, nq ** 3, 8), dtype=dt)
        self.dof_nodes = np.zeros(nd, dtype=np.float64)
        self.qpts = np.zeros(nq, dtype=np.float64)
        ident = C.c_int(0)
        check(lib.bd_tables(bits, degree, qmode, int(use_gauss), _ptr(self.phi0), _ptr(self.dphi1),
                            _ptr(self.weights3d), _ptr(self.dphi_geometry), _ptr(self.dof_nodes),
                            _ptr(self.qpts), C.byref(ident)))
        self.is_identity = bool(ident.value)
        if qmode == 0 and not self.is_identity:
            raise BenchDolfinxError(-2, "Expecting identity matrix for qmode=0")  # src/laplacian.hpp:197-198


_cudart = None


def _memcpy_to_host(dst: np.ndarray, src_ptr: int):
    global _cudart
    if _cudart is None:
        import torch  # makes sure libcudart is loaded
        _cudart = C.CDLL("libcudart.so.12", mode=C.RTLD_GLOBAL)
        _cudart.cudaMemcpy.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        _cudart.cudaMemcpy.restype = C.c_int
    rc = _cudart.cudaMemcpy(dst.ctypes.data, src_ptr, dst.nbytes, 2)
    if rc != 0:
        raise BenchDolfinxError(-3, f"cudaMemcpy D2H failed with {rc}")
    return dst


class BoxMesh:
    """Device-resident structured stand-in for the DOLFINx mesh / dofmap / BC setup of one rank."""

    def __init__(self, n, degree, rank=0, nranks=1, geom_perturb_fact=0.0, dtype=np.float64, pgrid=None):
        self.bits = _bits(dtype)
        self.degree, self.rank, self.nranks = int(degree), int(rank), int(nranks)
        self.n = tuple(int(v) for v in n)
        self.pgrid = tuple(pgrid) if pgrid is not None else partition_grid(nranks)
        self._h = C.c_void_p()
        n_arr = (C.c_int64 * 3)(*self.n)
        p_arr = (C.c_int32 * 3)(*self.pgrid)
        check(lib.bd_boxmesh_create(self.bits, n_arr, self.degree, self.rank, self.nranks, p_arr,
                                    float(geom_perturb_fact), C.byref(self._h)))
        self.info = _lib.BoxMeshInfo()
        check(lib.bd_boxmesh_get(self._h, C.byref(self.info)))
        i = self.info
        self.ncells_owned, self.ncells = i.ncells_owned, i.ncells
        self.ndofs_owned, self.nghost = i.ndofs_owned, i.nghost
        self.ndofs_local = i.ndofs_owned + i.nghost
        self.nnodes, self.n_lcells, self.n_bcells = i.nnodes, i.n_lcells, i.n_bcells
        self.ndofs_global, self.ncells_global = i.ndofs_global, i.ncells_global

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # module globals may be gone at interpreter exit
            lib.bd_boxmesh_destroy(h)

    # -- host copies (tests / small meshes)
    def _host(self, ptr, shape, dtype):
        out = np.empty(shape, dtype=dtype)
        return _memcpy_to_host(out, ptr) if out.size else out

    def host_arrays(self):
        nd3 = (self.degree + 1) ** 3
        i = self.info
        def hlist(ptr, count, ct):
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(count,)).copy() if count else np.zeros(0, dtype=ct)
        send_counts = hlist(i.send_counts, i.n_send, C.c_int64)
        return dict(
            cell_dofmap=self._host(i.cell_dofmap, (self.ncells, nd3), np.int32),
            geometry_dofmap=self._host(i.geometry_dofmap, (self.ncells, 8), np.int32),
            xgeom=self._host(i.xgeom, (self.nnodes, 3), _np_dtype(self.bits)),
            bc_marker=self._host(i.bc_marker, (self.ndofs_local,), np.int8),
            lcells=self._host(i.lcells, (self.n_lcells,), np.int32),
            bcells=self._host(i.bcells, (self.n_bcells,), np.int32),
            send_ranks=hlist(i.send_ranks, i.n_send, C.c_int32), send_counts=send_counts,
            send_indices=hlist(i.send_indices_host, int(send_counts.sum()), C.c_int32),
            recv_ranks=hlist(i.recv_ranks, i.n_recv, C.c_int32),
            recv_counts=hlist(i.recv_counts, i.n_recv, C.c_int64),
        )

    def assemble_rhs(self, qmode, use_gauss=False, stream=None):
        """u := b = M f with BC rows zeroed (src/laplacian_solver.cpp:100-109)."""
        import torch
        b = torch.zeros(self.ndofs_local, dtype=_torch_dtype(self.bits), device="cuda")
        check(lib.bd_boxmesh_assemble_rhs(self._h, int(qmode), int(use_gauss), b.data_ptr(),
                                          _stream_ptr(stream)))
        return b


class Comm:
    """One NCCL communicator per rank; replaces the MPI communicator of the reference."""

    def __init__(self, rank=0, nranks=1, unique_id: bytes | None = None):
        self.rank, self.nranks = rank, nranks
        self._h = C.c_void_p()
        buf = C.create_string_buffer(unique_id, _lib.BD_UNIQUE_ID_BYTES) if unique_id is not None else None
        check(lib.bd_comm_create(buf, rank, nranks, C.byref(self._h)))

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(_lib.BD_UNIQUE_ID_BYTES)
        check(lib.bd_comm_unique_id(buf))
        return buf.raw

    @classmethod
    def from_torch_distributed(cls):
        """Bootstrap over an initialised torch.distributed process group (id broadcast only)."""
        import torch
        import torch.distributed as dist
        rank, nranks = dist.get_rank(), dist.get_world_size()
        if nranks == 1:
            return cls(0, 1, None)
        obj = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        return cls(rank, nranks, obj[0])

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # module globals may be gone at interpreter exit
            lib.bd_comm_destroy(h)


class Halo:
    """Forward scatter plan (owner -> ghost) for the vectors of one mesh."""

    def __init__(self, comm: Comm, mesh, dtype=None):
        self.comm = comm
        bits = mesh.bits if dtype is None else _bits(dtype)
        self._h = C.c_void_p()
        if isinstance(mesh, BoxMesh):
            i = mesh.info
            check(lib.bd_halo_create(comm._h, bits, mesh.ndofs_owned, mesh.nghost, i.n_send,
                                     i.send_ranks, i.send_counts, i.send_indices_host, i.n_recv,
                                     i.recv_ranks, i.recv_counts, C.byref(self._h)))
        else:  # oracle-style mesh object with python lists
            sr = np.asarray(mesh.send_ranks, dtype=np.int32)
            sc = np.asarray([len(v) for v in mesh.send_indices], dtype=np.int64)
            si = (np.concatenate(mesh.send_indices).astype(np.int32) if len(mesh.send_indices)
                  else np.zeros(0, np.int32))
            rr = np.asarray(mesh.recv_ranks, dtype=np.int32)
            rc = np.asarray(mesh.recv_counts, dtype=np.int64)
            check(lib.bd_halo_create(comm._h, bits, mesh.ndofs_owned, mesh.nghost, len(sr), _ptr(sr),
                                     _ptr(sc), _ptr(si), len(rr), _ptr(rr), _ptr(rc), C.byref(self._h)))

    def set_unpack(self, ghost_positions):
        """Message entry i lands at ghost position ghost_positions[i] (ghosts in arbitrary order)."""
        if ghost_positions is None:
            check(lib.bd_halo_set_unpack(self._h, None))
        else:
            gp = np.ascontiguousarray(ghost_positions, dtype=np.int32)
            check(lib.bd_halo_set_unpack(self._h, _ptr(gp)))

    def begin(self, x, stream=None):
        check(lib.bd_halo_begin(self._h, x.data_ptr(), _stream_ptr(stream)))

    def end(self, stream=None):
        check(lib.bd_halo_end(self._h, _stream_ptr(stream)))

    def scatter_fwd(self, x, stream=None):
        self.begin(x, stream)
        self.end(stream)

    def profile(self, enable: bool):
        """Time pack / exchange / exposed wait of every operator application that uses this plan."""
        check(lib.bd_halo_profile(self._h, int(enable)))

    def profile_read(self):
        """dict(pack_ms, exchange_ms, exposed_ms, applications) summed since the last read."""
        a, b, c, n = C.c_double(), C.c_double(), C.c_double(), C.c_int64()
        check(lib.bd_halo_profile_read(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(n)))
        return {"pack_ms": a.value, "exchange_ms": b.value, "exposed_ms": c.value, "applications": n.value}

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # module globals may be gone at interpreter exit
            lib.bd_halo_destroy(h)


class MatFreeLaplacianGPU:
    """Matrix-free Q_k Laplacian, same constructor meaning and ``apply(in, out)`` as the reference
    class (src/laplacian.hpp:102-104,355-356); the FunctionSpace/DirichletBC pair is replaced by a
    mesh object carrying the raw arrays (``BoxMesh`` on the device, or any object with numpy
    ``cell_dofmap, xgeom, geometry_dofmap, bc_marker, lcells, bcells`` attributes)."""

    def __init__(self, mesh, degree: int, qmode: int, constant: float, use_gauss: bool = False,
                 dtype=np.float64, halo: Halo | None = None, tables: Tables | None = None,
                 geometry: str = "auto"):
        """``geometry``: "precomputed" = G stored and streamed as the reference does
        (src/laplacian.hpp:238-272), "onthefly" = G rebuilt inside the stiffness kernel from the cell
        vertices, "auto" (default) = on the fly when every cell is a parallelepiped, precomputed otherwise.
        "auto-full" / "onthefly-full": the same, but without the diagonal-G shortcut the dof-space kernels take on
        meshes of bricks (cells with mutually orthogonal edges, e.g. the unperturbed box)."""
        if geometry not in ("auto", "precomputed", "onthefly", "auto-full", "onthefly-full"):
            raise BenchDolfinxError(-1, "geometry must be 'auto', 'precomputed', 'onthefly', 'auto-full' or "
                                        "'onthefly-full'")
        self.bits = _bits(dtype)
        self.value_type = _np_dtype(self.bits)
        self.degree, self.qmode = int(degree), int(qmode)
        if qmode not in (0, 1):
            raise BenchDolfinxError(-1, "Invalid qmode.")
        self.tables = tables if tables is not None else Tables(degree, qmode, use_gauss, self.value_type)
        t = self.tables
        d = _lib.LaplacianDesc()
        d.scalar_bits, d.degree, d.nq = self.bits, self.degree, t.nq
        d.is_identity, d.kappa = int(t.is_identity), float(constant)
        self._keep = []
        if isinstance(mesh, BoxMesh):
            if mesh.bits != self.bits:
                raise BenchDolfinxError(-1, "mesh and operator scalar types differ")
            i = mesh.info
            d.ncells, d.ndofs, d.nnodes = mesh.ncells, mesh.ndofs_local, mesh.nnodes
            d.n_lcells, d.n_bcells = mesh.n_lcells, mesh.n_bcells
            d.cell_dofmap, d.xgeom, d.geometry_dofmap = i.cell_dofmap, i.xgeom, i.geometry_dofmap
            d.bc_marker, d.lcells, d.bcells = i.bc_marker, i.lcells, i.bcells
            d.flags = _lib.BD_FLAG_BORROW_DEVICE_ARRAYS
            self._keep.append(mesh)
            self.n_owned = mesh.ndofs_owned
        else:
            arrs = dict(cell_dofmap=np.ascontiguousarray(mesh.cell_dofmap, dtype=np.int32),
                        xgeom=np.ascontiguousarray(mesh.xgeom, dtype=self.value_type),
                        geometry_dofmap=np.ascontiguousarray(mesh.geometry_dofmap, dtype=np.int32),
                        bc_marker=np.ascontiguousarray(mesh.bc_marker, dtype=np.int8),
                        lcells=np.ascontiguousarray(mesh.lcells, dtype=np.int32),
                        bcells=np.ascontiguousarray(mesh.bcells, dtype=np.int32))
            self._keep.append(arrs)
            d.ncells, d.ndofs = arrs["cell_dofmap"].shape[0], arrs["bc_marker"].shape[0]
            d.nnodes = arrs["xgeom"].shape[0]
            d.n_lcells, d.n_bcells = len(arrs["lcells"]), len(arrs["bcells"])
            for k, v in arrs.items():
                setattr(d, k, _ptr(v))
            d.flags = 0
            self.n_owned = getattr(mesh, "ndofs_owned", d.ndofs)
        d.flags |= {"auto": 0, "onthefly": _lib.BD_FLAG_GEOMETRY_ONTHEFLY,
                    "precomputed": _lib.BD_FLAG_GEOMETRY_PRECOMPUTED,
                    "auto-full": _lib.BD_FLAG_GEOMETRY_FULL_TENSOR,
                    "onthefly-full": _lib.BD_FLAG_GEOMETRY_ONTHEFLY | _lib.BD_FLAG_GEOMETRY_FULL_TENSOR}[geometry]
        d.phi0, d.dphi1 = _ptr(t.phi0), _ptr(t.dphi1)
        d.weights, d.dphi_geometry = _ptr(t.weights3d), _ptr(t.dphi_geometry)
        self.ndofs = int(d.ndofs)
        self.halo = halo
        self._h = C.c_void_p()
        check(lib.bd_laplacian_create(C.byref(d), C.byref(self._h)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # module globals may be gone at interpreter exit
            lib.bd_laplacian_destroy(h)

    def _check_vec(self, v):
        if not v.is_cuda or v.numel() != self.ndofs or _bits(v.dtype) != self.bits or not v.is_contiguous():
            raise BenchDolfinxError(-1, "vector must be a contiguous CUDA tensor of the operator's "
                                        f"scalar type with {self.ndofs} entries (owned + ghost)")

    def apply(self, in_, out, stream=None):
        """out = A in (src/laplacian.hpp:355-403), including the forward scatter of ``in`` when the
        operator was given a halo plan.  Asynchronous on the current torch stream."""
        self._check_vec(in_)
        self._check_vec(out)
        halo = self.halo._h if self.halo is not None else None
        check(lib.bd_laplacian_apply_halo(self._h, halo, in_.data_ptr(), out.data_ptr(),
                                          _stream_ptr(stream)))

    def apply_host(self, in_host: np.ndarray, out_host: np.ndarray):
        """Host-buffer entry point: H2D, (forward scatter +) apply, D2H; synchronous."""
        assert in_host.dtype == self.value_type and out_host.dtype == self.value_type
        assert in_host.size == self.ndofs and out_host.size == self.ndofs
        halo = self.halo._h if self.halo is not None else None
        check(lib.bd_laplacian_apply_host(self._h, halo, _ptr(in_host), _ptr(out_host)))

    def apply_host_batch(self, ins, outs):
        """outs[i] = A ins[i] for lists of host vectors; copies and operator are pipelined."""
        n = len(ins)
        assert n == len(outs)
        for a, b in zip(ins, outs):
            assert a.dtype == self.value_type and b.dtype == self.value_type
            assert a.size == self.ndofs and b.size == self.ndofs
        pin = (C.c_void_p * n)(*[_ptr(a) for a in ins])
        pout = (C.c_void_p * n)(*[_ptr(b) for b in outs])
        halo = self.halo._h if self.halo is not None else None
        check(lib.bd_laplacian_apply_host_batch(self._h, halo, n, pin, pout))

    def geometry_mode(self):
        """("precomputed" | "onthefly-affine" | "onthefly-general" | "onthefly-brick", number of non-affine cells
        or -1).  "onthefly-brick": every cell affine with orthogonal edges, dof-space kernel with diagonal G_cell."""
        mode, n = C.c_int(), C.c_int64()
        check(lib.bd_laplacian_geometry_mode(self._h, C.byref(mode), C.byref(n)))
        return ("precomputed", "onthefly-affine", "onthefly-general", "onthefly-brick")[mode.value], n.value

    def get_diag_inverse(self, diag_inv=None, stream=None):
        """1 / diag(A) as a device vector (MatrixOperator::get_diag_inverse, src/csr.hpp:154-162, computed
        matrix-free here; 1 on constrained rows).  Ghost entries are partial sums: use the owned part only."""
        import torch
        if diag_inv is None:
            diag_inv = torch.empty(self.ndofs, dtype=_torch_dtype(self.bits), device="cuda")
        self._check_vec(diag_inv)
        check(lib.bd_laplacian_diagonal_inverse(self._h, diag_inv.data_ptr(), _stream_ptr(stream)))
        return diag_inv

    def describe(self) -> str:
        buf = C.create_string_buffer(256)
        check(lib.bd_laplacian_describe(self._h, buf, 256))
        return buf.value.decode()

    def profile(self, enable: bool):
        """Bracket every stiffness-kernel launch with CUDA events (roofline measurement)."""
        check(lib.bd_laplacian_profile(self._h, int(enable)))

    def profile_read(self):
        """(summed kernel milliseconds, launches) since the last read."""
        ms, n = C.c_double(), C.c_int64()
        check(lib.bd_laplacian_profile_read(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def geometry(self) -> np.ndarray:
        """Host copy of the precomputed G, [n_lcells + n_bcells][6][nq^3]."""
        p, n = C.c_void_p(), C.c_int64()
        check(lib.bd_laplacian_geometry(self._h, C.byref(p), C.byref(n)))
        nq3 = self.tables.nq ** 3
        out = np.empty((n.value // (6 * nq3), 6, nq3), dtype=self.value_type)
        return _memcpy_to_host(out, p.value) if out.size else out


class MatrixOperator:
    """Assembled-CSR operator for the --mat_comp check (src/csr.hpp:116-233), assembled on the
    device from a MatFreeLaplacianGPU's own data; same ``apply(in, out)`` duck type."""

    def __init__(self, op: MatFreeLaplacianGPU):
        self.value_type, self.bits, self.ndofs, self.n_owned = op.value_type, op.bits, op.ndofs, op.n_owned
        self.halo = op.halo
        self._keep = op
        self._h = C.c_void_p()
        check(lib.bd_csr_assemble(op._h, int(op.n_owned), C.byref(self._h)))
        nr, nc, nnz = C.c_int64(), C.c_int64(), C.c_int64()
        check(lib.bd_csr_info(self._h, C.byref(nr), C.byref(nc), C.byref(nnz)))
        self.shape, self.nnz = (nr.value, nc.value), nnz.value

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:          # module globals may be gone at interpreter exit
            lib.bd_csr_destroy(h)

    _check_vec = MatFreeLaplacianGPU._check_vec

    def apply(self, in_, out, stream=None):
        self._check_vec(in_)
        self._check_vec(out)
        if self.halo is not None:
            self.halo.scatter_fwd(in_, stream)
        check(lib.bd_csr_apply(self._h, in_.data_ptr(), out.data_ptr(), _stream_ptr(stream)))


# ------------------------------------------------------------------ vector functions (src/vector.hpp)
def axpy(r, alpha, x, y, n=None, stream=None):
    """r = alpha*x + y over the first n (owned) entries (src/vector.hpp:228-240)."""
    n = r.numel() if n is None else n
    check(lib.bd_axpy(_bits(r.dtype), n, float(alpha), x.data_ptr(), y.data_ptr(), r.data_ptr(),
                      _stream_ptr(stream)))


def set_value(x, value, stream=None):
    """src/vector.hpp:288-292."""
    check(lib.bd_set_value(_bits(x.dtype), x.numel(), float(value), x.data_ptr(), _stream_ptr(stream)))


def inner_product(a, b, n=None, comm: Comm | None = None) -> float:
    """Global inner product over owned entries (src/vector.hpp:159-176)."""
    import torch
    torch.cuda.current_stream().synchronize()
    n = a.numel() if n is None else n
    out = C.c_double()
    check(lib.bd_inner_product(comm._h if comm is not None else None, _bits(a.dtype), n, a.data_ptr(),
                               b.data_ptr(), C.byref(out)))
    return out.value


def norm(a, n=None, comm: Comm | None = None) -> float:
    """l2 norm over owned entries (src/vector.hpp:196-218)."""
    return math.sqrt(inner_product(a, a, n, comm))


def pointwise_mult(w, x, y, n=None, stream=None):
    """w = x * y over the first n (owned) entries (src/vector.hpp:266-285)."""
    n = w.numel() if n is None else n
    check(lib.bd_pointwise_mult(_bits(w.dtype), n, x.data_ptr(), y.data_ptr(), w.data_ptr(), _stream_ptr(stream)))


def cg_solve(A: MatFreeLaplacianGPU, x, b, max_iter: int, rtol: float = 0.0, comm: Comm | None = None,
             stream=None, diag_inv=None) -> int:
    """Unpreconditioned CG (src/cg.hpp:89-169); returns the iteration count.  For the matrix-free
    operator the whole solve is the fused device loop of bd_cg_solve; any other duck-typed operator
    (the CSR check operator) runs the same recurrence through the vector functions."""
    A._check_vec(x)
    A._check_vec(b)
    if comm is None and getattr(A, "halo", None) is not None:
        comm = A.halo.comm            # the sums must run over the ranks the halo plan spans
    if not isinstance(A, MatFreeLaplacianGPU):
        if diag_inv is not None:
            raise BenchDolfinxError(-1, "Jacobi preconditioning is implemented for the matrix-free operator")
        return _cg_solve_generic(A, x, b, max_iter, rtol, comm)
    its = C.c_int(0)
    halo = A.halo._h if A.halo is not None else None
    if diag_inv is not None:     # Jacobi: "p <- beta*p + M^-1(r)" (src/cg.hpp:165)
        A._check_vec(diag_inv)
        check(lib.bd_cg_solve_jacobi(A._h, halo, comm._h if comm is not None else None, x.data_ptr(), b.data_ptr(),
                                     diag_inv.data_ptr(), int(max_iter), float(rtol), C.byref(its),
                                     _stream_ptr(stream)))
        return its.value
    check(lib.bd_cg_solve(A._h, halo, comm._h if comm is not None else None, x.data_ptr(), b.data_ptr(),
                          int(max_iter), float(rtol), C.byref(its), _stream_ptr(stream)))
    return its.value


def _cg_solve_generic(A, x, b, max_iter, rtol, comm):
    import torch
    n = A.n_owned
    y, r, p = torch.zeros_like(x), torch.zeros_like(x), torch.zeros_like(x)
    A.apply(x, y)
    axpy(r, -1.0, y, b, n)
    axpy(p, 0.0, r, r, n)
    rnorm0 = rnorm = inner_product(p, r, n, comm)
    k = 0
    while k < max_iter:
        k += 1
        A.apply(p, y)
        alpha = rnorm / inner_product(p, y, n, comm)
        axpy(x, alpha, p, x, n)
        axpy(r, -alpha, y, r, n)
        rnorm_new = inner_product(r, r, n, comm)
        beta = rnorm_new / rnorm
        rnorm = rnorm_new
        if rnorm / rnorm0 < rtol * rtol:
            break
        axpy(p, beta, p, r, n)
    return k


# ------------------------------------------------------------------ benchmark harness
@dataclass
class BenchmarkResults:
    """src/laplacian_solver.hpp:13-20."""
    mat_free_time: float = 0.0
    unorm: float = 0.0
    ynorm: float = 0.0
    znorm: float = 0.0
    enorm: float = 0.0


def laplace_action_gpu(mesh: BoxMesh, degree, qmode, kappa, nreps, use_gauss=False, use_cg=False,
                       comm: Comm | None = None, dtype=np.float64, verbose=False, matrix_comparison=False):
    """The timed benchmark body of the reference (src/laplacian_solver.cpp:66-149): build the
    operator, u := assembled RHS, then ``nreps`` operator applications or ``nreps`` CG iterations,
    wall-clocked, followed by the norms of u and y."""
    import torch
    bits = _bits(dtype)
    halo = Halo(comm, mesh) if (comm is not None and comm.nranks > 1) else None
    op = MatFreeLaplacianGPU(mesh, degree, qmode, kappa, use_gauss, dtype, halo=halo)
    u = mesh.assemble_rhs(qmode, use_gauss)
    y = torch.zeros_like(u)
    if halo is not None:
        halo.scatter_fwd(u)                                      # src/laplacian_solver.cpp:111-114
    torch.cuda.synchronize()
    start = time.perf_counter()
    if use_cg:
        cg_solve(op, y, u, nreps, 0.0, comm)
    else:
        for _ in range(nreps):
            op.apply(u, y)
    torch.cuda.synchronize()
    res = BenchmarkResults(mat_free_time=time.perf_counter() - start)
    res.unorm = norm(u, mesh.ndofs_owned, comm)
    res.ynorm = norm(y, mesh.ndofs_owned, comm)
    if verbose and (comm is None or comm.rank == 0):
        kind = "CG" if use_cg else "Action"
        print(f"Computation time ({kind}): {res.mat_free_time}s")
        print(f"Computation rate (Gdofs/s): {mesh.ndofs_global * nreps / (1e9 * res.mat_free_time)}")
        print(f"Norm of u = {res.unorm}")
        print(f"Norm of y = {res.ynorm}")
    if matrix_comparison:                                        # src/laplacian_solver.cpp:151-227
        mat_op = MatrixOperator(op)
        z = torch.zeros_like(u)
        if use_cg:
            cg_solve(mat_op, z, u, nreps, 0.0, comm)
        else:
            for _ in range(nreps):
                mat_op.apply(u, z)
        e = torch.zeros_like(u)
        axpy(e, -1.0, y, z, mesh.ndofs_owned)
        res.znorm = norm(z, mesh.ndofs_owned, comm)
        res.enorm = norm(e, mesh.ndofs_owned, comm)
        if verbose and (comm is None or comm.rank == 0):
            print(f"Norm of u = {res.unorm}")
            print(f"Norm of z = {res.znorm}")
            print(f"Norm of error = {res.enorm}")
            print(f"Relative norm of error = {res.enorm / res.znorm}")
    return res, op, u, y
