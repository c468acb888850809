"""Vector kernels (axpy, dot, norm, set_value, pack, unpack) through the C ABI vs numpy."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = [0, 1, 2, 31, 32, 33, 255, 256, 257, 1000, 65537, 1_000_003, 5_000_000]


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("n", SIZES)
def test_axpy_dot_norm(n, dtype):
    import torch
    import benchmark_dolfinx_b200 as bd
    rng = np.random.default_rng(n + 1)
    x, y = rng.uniform(-1, 1, n).astype(dtype), rng.uniform(-1, 1, n).astype(dtype)
    xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    rd = torch.zeros_like(xd)
    bd.axpy(rd, 0.75, xd, yd)
    torch.cuda.synchronize()
    # the kernel fuses the multiply-add, numpy rounds twice: allow one rounding of the product
    err = np.abs(rd.cpu().numpy().astype(np.float64) - (x.astype(np.float64) * 0.75 + y.astype(np.float64)))
    assert (err <= 2 * np.finfo(dtype).eps * (np.abs(x) * 0.75 + np.abs(y)) + 1e-300).all()
    ref = float(np.dot(x.astype(np.float64), y.astype(np.float64)))
    got = bd.inner_product(xd, yd)
    tol = 1e-13 * max(1.0, np.sqrt(n)) if dtype == np.float64 else 1e-12 * max(1.0, n)
    assert abs(got - ref) <= tol * max(1.0, abs(ref)) + 1e-300
    assert abs(bd.norm(xd) - np.linalg.norm(x.astype(np.float64))) <= 1e-12 * max(1.0, np.sqrt(n))
    # owned-entries-only semantics (size_local, src/vector.hpp:164)
    if n >= 2:
        k = n // 2
        assert abs(bd.inner_product(xd, yd, k) - float(np.dot(x[:k].astype(np.float64), y[:k].astype(np.float64)))) \
            <= tol * max(1.0, abs(ref))


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_dot_is_reproducible_and_device_resident(dtype):
    import torch
    import benchmark_dolfinx_b200 as bd
    from benchmark_dolfinx_b200 import _lib
    n = 3_000_001
    x = torch.from_numpy(np.random.default_rng(0).uniform(-1, 1, n).astype(dtype)).cuda()
    scratch = torch.empty(_lib.lib.bd_dot_scratch_bytes() // 8, dtype=torch.float64, device="cuda")
    res = torch.zeros(2, dtype=torch.float64, device="cuda")
    bits = 64 if dtype == np.float64 else 32
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib.bd_dot(bits, n, x.data_ptr(), x.data_ptr(), res.data_ptr(), scratch.data_ptr(), s))
    _lib.check(_lib.lib.bd_dot(bits, n, x.data_ptr(), x.data_ptr(), res.data_ptr() + 8, scratch.data_ptr(), s))
    torch.cuda.synchronize()
    r = res.cpu().numpy()
    assert r[0] == r[1]                                         # bitwise, fixed summation order
    assert abs(r[0] - float(np.dot(x.cpu().numpy().astype(np.float64), x.cpu().numpy().astype(np.float64)))) <= 1e-10 * r[0]
    out = C.c_double()
    _lib.check(_lib.lib.bd_dot_host(bits, n, x.data_ptr(), x.data_ptr(), C.byref(out)))
    assert out.value == r[0]


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_set_value_pack_unpack(dtype):
    import torch
    import benchmark_dolfinx_b200 as bd
    from benchmark_dolfinx_b200 import _lib
    n = 100_003
    x = torch.empty(n, dtype=torch.float64 if dtype == np.float64 else torch.float32, device="cuda")
    bd.set_value(x, 3.5)
    torch.cuda.synchronize()
    assert (x.cpu().numpy() == dtype(3.5)).all()
    bd.set_value(x, 0.0)
    torch.cuda.synchronize()
    assert (x.cpu().numpy() == 0).all()
    rng = np.random.default_rng(3)
    v = rng.uniform(-1, 1, n).astype(dtype)
    idx = rng.permutation(n)[:40_001].astype(np.int32)
    vd, idxd = torch.from_numpy(v).cuda(), torch.from_numpy(idx).cuda()
    packed = torch.zeros(len(idx), dtype=vd.dtype, device="cuda")
    bits = 64 if dtype == np.float64 else 32
    s = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib.bd_pack(bits, len(idx), idxd.data_ptr(), vd.data_ptr(), packed.data_ptr(), s))
    torch.cuda.synchronize()
    assert np.array_equal(packed.cpu().numpy(), v[idx])            # out[i] = in[idx[i]]
    out = torch.zeros(n, dtype=vd.dtype, device="cuda")
    _lib.check(_lib.lib.bd_unpack(bits, len(idx), idxd.data_ptr(), packed.data_ptr(), out.data_ptr(), s))
    torch.cuda.synchronize()
    ref = np.zeros(n, dtype=dtype)
    ref[idx] = v[idx]                                              # out[idx[i]] = in[i]
    assert np.array_equal(out.cpu().numpy(), ref)


def test_fp64_peak_measurement():
    """bd_fp64_peak (roofline denominators of bench.py): plausible numbers for a B200-class device."""
    import benchmark_dolfinx_b200 as bd
    dfma, dmma = bd.fp64_peak()
    assert 5.0 < dfma < 100.0 and 5.0 < dmma < 100.0
