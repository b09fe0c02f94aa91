"""Integer tricks of the projection passes, replayed on the CPU (no GPU, no library call).

* lightgrid.cuh: an entry's rectangle is four 7-bit bytes x0, 127 - x1, y0, 127 - y1; a ray's point
  is the bytes a = floor(qx + .01), 127 - ceil(qx - .01), c, 127 - d with bit 7 set in each, and
  "the point is inside the rectangle" is ((P - r) & 0x80808080) == 0x80808080.
* raster.cuh: k / w for k < 1024, w <= 32 is (k * c_rast_rw[w]) >> 16 with the table in the source."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "raytracing-gpu_b200", "csrc")


def _pack_entry(x0, x1, y0, y1):
    return (x0 | ((127 - x1) << 8) | (y0 << 16) | ((127 - y1) << 24)).astype(np.uint32)


def _pack_point(qx, qy):
    b = lambda v: np.clip(v, 0, 127).astype(np.uint32)
    a, bb = b(np.floor(qx + 0.01)), b(np.ceil(qx - 0.01))
    c, d = b(np.floor(qy + 0.01)), b(np.ceil(qy - 0.01))
    return ((a | ((127 - bb) << 8) | (c << 16) | ((127 - d) << 24)) | np.uint32(0x80808080)).astype(np.uint32), a, bb, c, d


def test_four_byte_rectangle_compare_is_the_four_comparisons():
    rng = np.random.default_rng(11)
    n = 2_000_000
    x0 = rng.integers(0, 128, n).astype(np.uint32); x1 = rng.integers(0, 128, n).astype(np.uint32)
    y0 = rng.integers(0, 128, n).astype(np.uint32); y1 = rng.integers(0, 128, n).astype(np.uint32)
    # points anywhere in the cell, many of them exactly on byte values and on the cell's border
    qx = np.where(rng.random(n) < 0.3, rng.integers(0, 128, n).astype(np.float64), rng.uniform(0, 127, n))
    qy = np.where(rng.random(n) < 0.3, rng.integers(0, 128, n).astype(np.float64), rng.uniform(0, 127, n))
    r = _pack_entry(x0, x1, y0, y1)
    P, a, b, c, d = _pack_point(qx, qy)
    got = ((P - r) & np.uint32(0x80808080)) == np.uint32(0x80808080)        # uint32 wrap-around as on the device
    want = (x0 <= a) & (x1 >= b) & (y0 <= c) & (y1 >= d)
    assert np.array_equal(got, want)
    # and those are the float comparisons of the first version: inside the rectangle grown by 0.01
    flt = ~((qx < x0 - 0.01) | (qx > x1 + 0.01) | (qy < y0 - 0.01) | (qy > y1 + 0.01))
    assert np.array_equal(want, flt)
    assert 0.02 < want.mean() < 0.5


def test_empty_and_full_rectangles():
    q = np.array([0.0, 63.5, 127.0])
    P, *_ = _pack_point(q, q)
    full = _pack_entry(np.uint32([0]), np.uint32([127]), np.uint32([0]), np.uint32([127]))
    assert (((P - full) & np.uint32(0x80808080)) == np.uint32(0x80808080)).all()
    empty = _pack_entry(np.uint32([100]), np.uint32([20]), np.uint32([0]), np.uint32([127]))     # x0 > x1
    assert not (((P - empty) & np.uint32(0x80808080)) == np.uint32(0x80808080)).any()


def test_item_width_reciprocal_table():
    src = open(os.path.join(CSRC, "raster.cuh")).read()
    m = re.search(r"c_rast_rw\[33\]\s*=\s*\{([^}]*)\}", src)
    assert m, "table not found"
    rw = [int(x) for x in m.group(1).split(",")]
    assert len(rw) == 33
    k = np.arange(1024, dtype=np.uint64)
    for w in range(1, 33):
        assert np.array_equal((k * np.uint64(rw[w])) >> np.uint64(16), k // np.uint64(w)), w
        assert int(k[-1]) * rw[w] < 2 ** 32, "the product stays in 32 bits"
