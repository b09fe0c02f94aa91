"""ctypes view of the CHECKER libraries (test infrastructure only):
oracle/liboracle.so (our CPU restatement) and, when built, oracle/_ref/libref_gpusem.so
(the reference's gpu/*.cu shader sources compiled for the host) and
oracle/_ref/libref_cpu.so (the reference's cpu/*.c sources)."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REFERENCE = "/root/reference"


def _build(target):
    subprocess.run(["make", "-C", ORACLE_DIR, target], check=True, stdout=subprocess.DEVNULL)


_oracle = None


def oracle():
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            _build("liboracle.so")
        L = C.CDLL(path)
        L.orc_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_int]
        L.orc_render_samples.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int]
        L.orc_trace_rays.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int]
        L.orc_primary_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        for f in ("orc_init_color", "orc_color_mul", "orc_color_mults", "orc_color_add"):
            getattr(L, f).restype = C.c_uint32
        L.orc_init_color.argtypes = [C.c_float] * 3
        L.orc_color_mul.argtypes = [C.c_uint32, C.c_float]
        L.orc_color_mults.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_color_add.argtypes = [C.c_uint32, C.c_uint32]
        _oracle = L
    return _oracle


def have_reference():
    return os.path.isdir(os.path.join(REFERENCE, "gpu"))


def ref_lib(name):
    """name: 'gpusem' or 'cpu'.  Returns None when neither the built library nor the
    reference checkout is available (the GPU box)."""
    path = os.path.join(ORACLE_DIR, "_ref", "libref_%s.so" % name)
    if not os.path.exists(path):
        if not have_reference():
            return None
        _build("_ref/libref_%s.so" % name)
    L = C.CDLL(path)
    if name == "gpusem":
        L.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_collide.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
        L.ref_primary_hits.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
    else:
        L.refcpu_scene_create.restype = C.c_void_p
        L.refcpu_scene_create.argtypes = [C.c_void_p]
        L.refcpu_scene_destroy.argtypes = [C.c_void_p]
        L.refcpu_collide.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.refcpu_trace_samples.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p]
    return L


def orc_render(hs, aliasing=3, want_ids=True, want_hires=False, nthreads=0):
    """hs: pyrtb.HostScene (camera at OUTPUT resolution).  Returns dict(rgba, tri, obj, hires, counts)."""
    s = hs.scaled(aliasing)
    w, h = hs.width, hs.height
    W, H = w * aliasing, h * aliasing
    rgba = np.zeros((h, w, 4), np.uint8)
    tri = np.zeros((H, W), np.int32) if want_ids else None
    obj = np.zeros((H, W), np.int32) if want_ids else None
    hires = np.zeros((H, W, 4), np.uint8) if want_hires else None
    counts = np.zeros(3, np.uint64)
    rc = oracle().orc_render(C.byref(s), aliasing, w, h, rgba.ctypes.data, tri.ctypes.data if want_ids else None,
                             obj.ctypes.data if want_ids else None, hires.ctypes.data if want_hires else None,
                             counts.ctypes.data, nthreads)
    assert rc == 0
    return dict(rgba=rgba, tri=tri, obj=obj, hires=hires, counts=counts)


def orc_render_samples(hs, xy, aliasing=3, nthreads=0):
    s = hs.scaled(aliasing)
    xy = np.ascontiguousarray(xy, np.int32).reshape(-1, 2)
    n = xy.shape[0]
    tri = np.zeros(n, np.int32)
    obj = np.zeros(n, np.int32)
    col = np.zeros((n, 4), np.uint8)
    counts = np.zeros(3, np.uint64)
    rc = oracle().orc_render_samples(C.byref(s), xy.ctypes.data, n, tri.ctypes.data, obj.ctypes.data, col.ctypes.data,
                                     counts.ctypes.data, nthreads)
    assert rc == 0
    return dict(tri=tri, obj=obj, color=col, counts=counts)


def orc_trace_rays(hs, rays, any_hit=False, nthreads=0):
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    n = rays.shape[0]
    tri = np.zeros(n, np.int32)
    dist = np.zeros(n, np.float32)
    hit = np.zeros((n, 6), np.float32)
    rc = oracle().orc_trace_rays(C.byref(hs.s), int(any_hit), rays.ctypes.data, n, tri.ctypes.data, dist.ctypes.data,
                                 hit.ctypes.data, nthreads)
    assert rc == 0
    return tri, dist, hit


def orc_primary_hits(hs, aliasing=3, nthreads=0):
    s = hs.scaled(aliasing)
    W, H = hs.width * aliasing, hs.height * aliasing
    hit = np.zeros((H, W, 6), np.float32)
    tri = np.zeros((H, W), np.int32)
    assert oracle().orc_primary_hits(C.byref(s), hit.ctypes.data, tri.ctypes.data, nthreads) == 0
    return hit, tri


def ref_render(hs, aliasing=3, want_hires=False, nthreads=0):
    L = ref_lib("gpusem")
    s = hs.scaled(aliasing)
    w, h = hs.width, hs.height
    rgba = np.zeros((h, w, 4), np.uint8)
    hires = np.zeros((h * aliasing, w * aliasing, 4), np.uint8) if want_hires else None
    rc = L.ref_render(C.byref(s), aliasing, w, h, rgba.ctypes.data, hires.ctypes.data if want_hires else None, nthreads)
    assert rc == 0
    return rgba, hires


def ref_primary_hits(hs, aliasing=3, nthreads=0):
    L = ref_lib("gpusem")
    s = hs.scaled(aliasing)
    W, H = hs.width * aliasing, hs.height * aliasing
    hit = np.zeros((H, W, 6), np.float32)
    assert L.ref_primary_hits(C.byref(s), hit.ctypes.data, nthreads) == 0
    return hit


def ref_collide(hs, rays, nthreads=0):
    L = ref_lib("gpusem")
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    hit = np.zeros((rays.shape[0], 6), np.float32)
    assert L.ref_collide(C.byref(hs.s), rays.ctypes.data, rays.shape[0], hit.ctypes.data, nthreads) == 0
    return hit


def refcpu_collide(hs, rays):
    L = ref_lib("cpu")
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    hit = np.zeros((rays.shape[0], 6), np.float32)
    h = L.refcpu_scene_create(C.byref(hs.s))
    try:
        assert L.refcpu_collide(h, rays.ctypes.data, rays.shape[0], hit.ctypes.data) == 0
    finally:
        L.refcpu_scene_destroy(h)
    return hit
