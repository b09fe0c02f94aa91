"""Known-answer vectors of the reference's manual harness and benchmark asserts, replayed
on the device sort/scan (gpu/test.cu:288-305 prefix sum of 1..15; :308-340 sort of 0..384;
gpu/bench.cu:249-251 sortedness of 1 M random keys; :335,372 scan total == sum)."""
import numpy as np
import pytest

import pyrtb

pytestmark = pytest.mark.gpu


def test_prefix_sum_1_to_15(gpu):
    out = pyrtb.inclusive_scan(np.arange(1, 16, dtype=np.uint64))
    assert out.tolist() == [n * (n + 1) // 2 for n in range(1, 16)] and out[-1] == 120


def test_sort_0_to_384_is_identity(gpu):
    keys = np.arange(385, dtype=np.uint32)
    k, v = pyrtb.radix_sort_pairs(keys, np.arange(385, dtype=np.uint32))
    assert np.array_equal(k, keys) and np.array_equal(v, keys)


@pytest.mark.parametrize("n", [0, 1, 31, 4096, 4097, 100000, 1000000])
def test_scan_matches_numpy(gpu, n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 8, n).astype(np.uint32)       # rand() % 8, PREFIX_SUM_COUNT = 100000 in bench.cu
    out = pyrtb.inclusive_scan(a)
    assert np.array_equal(out, np.cumsum(a, dtype=np.uint64).astype(np.uint32))
    if n:
        assert out[-1] == a.sum()
    b = rng.integers(0, 2**40, n).astype(np.uint64)
    assert np.array_equal(pyrtb.inclusive_scan(b), np.cumsum(b, dtype=np.uint64))


@pytest.mark.parametrize("n", [0, 1, 2, 33, 4095, 4096, 4097, 12289, 1000000])
def test_radix_sort_random_keys_stable(gpu, n):
    rng = np.random.default_rng(n + 1)
    keys = rng.integers(0, 2**63, n).astype(np.uint64)
    if n > 100:
        keys[rng.integers(0, n, n // 3)] = keys[0]        # many duplicates: stability matters
    vals = np.arange(n, dtype=np.uint32)
    k, v = pyrtb.radix_sort_pairs(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k, keys[order])
    assert np.array_equal(v, order.astype(np.uint32))
    k32 = rng.integers(0, 2**32, n).astype(np.uint32)
    s32, _ = pyrtb.radix_sort_pairs(k32, vals)
    assert np.array_equal(s32, np.sort(k32))
