"""The repo's .svati loader (host/parser.cpp) pinned to the REFERENCE's own loader (CPU, build container).

oracle/_ref/libref_parser.so is the reference's gpu/parser.cpp:64-112 + gpu/parse_obj.cpp:15-62 +
gpu/scene.cu:146-218 compiled here as host C++ (oracle/Makefile, oracle/refparser_driver.cpp).  Every
shipped scene and a generated file with odd number spellings must give byte-identical vertex / normal /
material / light / camera arrays through both loaders, and the committed golden fixtures (what the GPU
box tests with) must hold exactly the arrays the reference reads -- so the parity tests no longer rest
on our own parser.  Needs /root/reference (or the prebuilt library plus the scene files): skipped otherwise."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_py
import pyrtb
from util import golden_names, load_golden

SCENES = os.path.join(oracle_py.REFERENCE, "tests")
pytestmark = pytest.mark.skipif(not os.path.isdir(SCENES), reason="needs the reference checkout (scene files)")


def _ref():
    path = os.path.join(oracle_py.ORACLE_DIR, "_ref", "libref_parser.so")
    if not os.path.exists(path):
        oracle_py._build("_ref/libref_parser.so")
    L = C.CDLL(path)
    L.refparser_parse.restype = C.c_void_p
    L.refparser_parse.argtypes = [C.c_char_p]
    L.refparser_counts.argtypes = [C.c_void_p, C.c_void_p]
    L.refparser_copy.argtypes = [C.c_void_p] * 7
    return L


def ref_parse(path):
    L = _ref()
    h = L.refparser_parse(os.fsencode(path))
    cnt = np.zeros(3, np.uint64)
    L.refparser_counts(h, cnt.ctypes.data)
    n_obj, n_tri, n_light = (int(x) for x in cnt)
    v = np.zeros((n_tri, 3, 3), np.float32); n = np.zeros((n_tri, 3, 3), np.float32)
    tc = np.zeros(n_obj, np.uint32); m = np.zeros((n_obj, 13), np.float32)
    li = np.zeros((n_light, 7), np.float32); cam = np.zeros(12, np.float64)
    L.refparser_copy(h, v.ctypes.data, n.ctypes.data, tc.ctypes.data, m.ctypes.data, li.ctypes.data, cam.ctypes.data)
    return dict(vertex=v, normal=n, tri_counts=tc.astype(np.int64), materials=m, lights=li, camera=cam)


def _same(a, b, what):
    for k in ("vertex", "normal", "materials", "lights"):
        x, y = np.ascontiguousarray(a[k], np.float32).copy(), np.ascontiguousarray(b[k], np.float32).copy()
        if k == "lights":
            # parse_a_light (gpu/parser.cpp:26-35) never sets an ambient light's v: the reference holds stack
            # garbage there, the repo's loader zeroes it; nothing reads it (gpu/light.cu:59-67)
            for arr in (x, y):
                arr[arr[:, 0] == pyrtb.AMBIENT, 4:7] = 0
        assert x.shape == y.shape, (what, k, x.shape, y.shape)
        assert np.array_equal(x.view(np.uint32), y.view(np.uint32)), (what, k)
    assert np.array_equal(np.asarray(a["tri_counts"], np.int64), np.asarray(b["tri_counts"], np.int64)), what
    ca, cb = np.asarray(a["camera"], np.float64), np.asarray(b["camera"], np.float64)
    assert np.array_equal(ca.astype(np.float32).view(np.uint32), cb.astype(np.float32).view(np.uint32)), (what, "camera")


@pytest.mark.parametrize("name", sorted(f[:-6] for f in os.listdir(SCENES) if f.endswith(".svati")) if os.path.isdir(SCENES) else [])
def test_every_shipped_scene_parses_like_the_reference(name):
    path = os.path.join(SCENES, name + ".svati")
    want = ref_parse(path)
    hs = pyrtb.HostScene.parse(path)
    try:
        _same(hs.arrays(), want, name)
    finally:
        hs.close()
    assert name in golden_names()
    _, z = load_golden(name)
    _same({k: z[k] for k in ("vertex", "normal", "tri_counts", "materials", "lights", "camera")}, want, name + " (golden fixture)")


ODD = """camera 31 17 0.5 -.25 +3. 1e0 0.E0 -0.0 0 -1E+0 1e-45 70.000001
a_light .1 0.20000000298023223876953125 3e-1
d_light 1 1 1 5e-1 -1.0e+0 0.5000000000000000000000000000001
p_light 0.99999997 1.00000006 16777217 1e38 -1e-38 3.4028234e38
object 6
Ka 0.1 0.1 0.1
Kd 0.333333333333333333333 0.6666667 1.
Ks 0 0 0
Ns 1.5e1
Ni 1
Nr 0.5
d 1
v 1e-3 2.5E-3 -7.e-1
v 00012.50 -0007 +.125
v 123456789 0.000000001234 9.87654321e-5
v 1.17549435e-38 1.4e-45 0.1
v 16777216 16777217 16777218
v 0.30000001192092896 0.3 0.29999998
vn 0 1 0
vn 0 1. 0
vn 0 1e0 0
vn -0 -1 -0
vn 1e-1 1e-2 1e-3
vn 3.0 4.0 12.0
object 3
Kd 1 0 0
v 0 0 0
v 1 0 0
v 0 1 0
vn 0 0 1
vn 0 0 1
vn 0 0 1
"""


def test_odd_number_spellings_parse_like_the_reference(tmp_path):
    path = os.path.join(tmp_path, "odd.svati")
    open(path, "w").write(ODD)
    want = ref_parse(path)
    hs = pyrtb.HostScene.parse(path)
    try:
        got = hs.arrays()
        _same(got, want, "odd.svati")
        assert got["vertex"].shape == (3, 3, 3) and list(got["tri_counts"]) == [2, 1]
    finally:
        hs.close()
