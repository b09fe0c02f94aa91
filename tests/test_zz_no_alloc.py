"""rtb_dev_inclusive_scan_u32_static (GPU box): the allocation-free variant of the device scan.

No counterpart in the reference; the entry point was added
for callers that cannot call cudaMalloc (CUDA graph capture, serving loops).  Its scratch is a
file-scope device array inside librtb200.so (csrc/sort.cu), n <= RTB_SCAN_STATIC_MAX = 2^26.
Checked through the C ABI against the existing entry point rtb_dev_inclusive_scan_u32 (pool
scratch) and against numpy, bit for bit (integer work: tolerance 0), on calls in a row that share
the array.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import pyrtb

pytestmark = pytest.mark.gpu
RTB_ERR_LIMIT = -4
STATIC_MAX = 1 << 26


@pytest.fixture(scope="module")
def lib():
    L = pyrtb.lib()
    assert pyrtb.device_count() >= 1, "no CUDA device (there is no CPU fallback)"
    for f in (L.rtb_dev_inclusive_scan_u32, L.rtb_dev_inclusive_scan_u32_static):
        f.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        f.restype = C.c_int
    return L


def _both(L, host):
    """host: uint32 array.  Returns (existing entry point's result, static entry point's result)."""
    import torch
    torch.cuda.set_device(0)
    src = torch.from_numpy(host.view(np.int32)).cuda()
    a, b = src.clone(), src.clone()
    torch.cuda.synchronize()
    assert L.rtb_dev_inclusive_scan_u32(a.data_ptr(), host.size, None) == 0, L.rtb_last_error()
    assert L.rtb_dev_inclusive_scan_u32_static(b.data_ptr(), host.size, None) == 0, L.rtb_last_error()
    torch.cuda.synchronize()
    return a.cpu().numpy().view(np.uint32), b.cpu().numpy().view(np.uint32)


def test_static_scan_equals_the_pool_scan_on_calls_in_a_row(lib):
    rng = np.random.default_rng(0x5CA7)
    # 3 levels of tile sums, then 1 level, then 2: the later calls find the earlier calls' sums in the array
    for n in (5_000_000, 1_000, 300_000, 5_000_000, 2_048, 2_049, 1):
        host = rng.integers(0, 1 << 20, n, dtype=np.uint32)
        want = np.cumsum(host.astype(np.uint64)).astype(np.uint32)      # the scan is modulo 2^32
        pool, static = _both(lib, host)
        assert np.array_equal(pool, want), n
        assert np.array_equal(static, pool), n


def test_static_scan_wraps_like_the_pool_scan(lib):
    host = np.full(3_000_000, 0xFFFFFFF0, np.uint32)
    pool, static = _both(lib, host)
    assert np.array_equal(static, pool)
    assert np.array_equal(static, np.cumsum(host.astype(np.uint64)).astype(np.uint32))


def test_static_scan_at_its_maximum_size(lib):
    host = np.ones(STATIC_MAX, np.uint32)
    pool, static = _both(lib, host)
    assert np.array_equal(static, pool)
    assert static[0] == 1 and static[-1] == STATIC_MAX and np.array_equal(static[::4099], np.arange(1, STATIC_MAX + 1, 4099, dtype=np.uint32))


def test_static_scan_rejects_sizes_beyond_its_scratch(lib):
    import torch
    buf = torch.zeros(16, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    rc = lib.rtb_dev_inclusive_scan_u32_static(buf.data_ptr(), STATIC_MAX + 1, None)      # rejected before any access
    assert rc == RTB_ERR_LIMIT
    msg = lib.rtb_last_error().decode()
    assert "RTB_SCAN_STATIC_MAX" in msg and str(STATIC_MAX) in msg, msg
    assert lib.rtb_dev_inclusive_scan_u32_static(buf.data_ptr(), 0, None) == 0
    assert lib.rtb_dev_inclusive_scan_u32_static(None, 5, None) == -1
