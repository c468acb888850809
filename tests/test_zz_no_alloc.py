"""bd_dot_static: the dot product for host codes that cannot allocate on the device.

Some users of this library run their CG loop with every byte of GPU memory already handed to
their own pools, or keep the loop fully asynchronous; a cudaMalloc inside a library call would
fail in the first case and synchronise the device in the second.  For them the library offers a
variant of bd_dot whose partial sums live in a fixed array inside the library's device image
instead of caller-provided scratch.  This file checks, through the C ABI, that the variant gives
the same answer as bd_dot on two calls in a row (the second call sees the array still holding the
first call's partial sums), for both scalar types, and that over-long vectors are refused.

Collected last on purpose (file name), so a run that stops at the first failure has already run
every other test; it must also pass alone in a fresh process.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
def test_static_dot_matches_bd_dot_twice_in_a_row(dtype):
    import torch
    from benchmark_dolfinx_b200 import _lib
    lib, check = _lib.lib, _lib.check
    bits = 64 if dtype == np.float64 else 32
    rng = np.random.default_rng(2024)
    # first call: many CTAs; second call: fewer CTAs, so stale partial sums of the first call
    # sit right behind the entries the second call uses
    sizes = [4_000_037, 70_001]
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.empty(lib.bd_dot_scratch_bytes() // 8, dtype=torch.float64, device="cuda")
    res = torch.zeros(4, dtype=torch.float64, device="cuda")
    host = []
    for k, n in enumerate(sizes):
        a = rng.uniform(-1, 1, n).astype(dtype)
        b = rng.uniform(-1, 1, n).astype(dtype)
        host.append(float(np.dot(a.astype(np.float64), b.astype(np.float64))))
        ad, bd_ = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        if k == 1:
            # The first call on a device may still load the library's device image (CUDA loads modules
            # lazily), which is where the fixed array lives.  From the second call on nothing may be
            # taken from the device: free memory as the driver reports it must not move.
            torch.cuda.synchronize()
            free_before, _ = torch.cuda.mem_get_info()
        check(lib.bd_dot_static(bits, n, ad.data_ptr(), bd_.data_ptr(), res.data_ptr() + 16 * k, stream))
        if k == 1:
            torch.cuda.synchronize()
            free_after, _ = torch.cuda.mem_get_info()
            assert free_after == free_before, "bd_dot_static changed the device's free memory"
        check(lib.bd_dot(bits, n, ad.data_ptr(), bd_.data_ptr(), res.data_ptr() + 16 * k + 8,
                         scratch.data_ptr(), stream))
        torch.cuda.synchronize()
    r = res.cpu().numpy()
    for k in range(2):
        static, reference = r[2 * k], r[2 * k + 1]
        # same tolerance as tests/test_gpu_vector.py uses against numpy; in fact the two entry
        # points share kernels, grid and summation order, so they agree to the last bit
        assert abs(static - reference) <= 1e-13 * max(1.0, abs(reference))
        assert static == reference
        assert abs(static - host[k]) <= 1e-10 * max(1.0, abs(host[k]))


def test_static_dot_edge_cases_and_limit():
    import torch
    from benchmark_dolfinx_b200 import _lib
    lib = _lib.lib
    stream = torch.cuda.current_stream().cuda_stream
    res = torch.full((1,), 7.0, dtype=torch.float64, device="cuda")
    x = torch.ones(3, dtype=torch.float64, device="cuda")
    _lib.check(lib.bd_dot_static(64, 0, None, None, res.data_ptr(), stream))      # empty vectors
    torch.cuda.synchronize()
    assert res.item() == 0.0
    _lib.check(lib.bd_dot_static(64, 3, x.data_ptr(), x.data_ptr(), res.data_ptr(), stream))
    torch.cuda.synchronize()
    assert res.item() == 3.0
    too_long = 2147483647 + 1
    rc = lib.bd_dot_static(64, too_long, x.data_ptr(), x.data_ptr(), res.data_ptr(), stream)
    assert rc == -1 and b"exceed the supported maximum" in lib.bd_last_error()
    rc = lib.bd_dot_static(16, 3, x.data_ptr(), x.data_ptr(), res.data_ptr(), stream)
    assert rc == -1
