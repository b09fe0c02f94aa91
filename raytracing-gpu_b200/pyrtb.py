"""ctypes view of librtb200.so (include/rtb200.h, include/rt_host.h).

Harness-side only: the product is the C ABI and the `rt` CLI; this module lets the
tests and bench.py drive exactly those entry points.  It never computes anything
itself and it fails loudly when the shared library has not been built.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(_HERE, "librtb200.so")   # env override: A/B builds in tools/

AMBIENT, DIRECTIONAL, POINT = 0, 1, 2
TRACER_TREE, TRACER_BRUTE = 0, 1
BUILD_DEFAULT, BUILD_NO_TREE = 0, 1
RTB_OK, RTB_ERR_ARG, RTB_ERR_CUDA, RTB_ERR_NOGPU, RTB_ERR_LIMIT = 0, -1, -2, -3, -4


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]


class Light(C.Structure):
    _fields_ = [("type", C.c_int), ("r", C.c_float), ("g", C.c_float), ("b", C.c_float), ("v", Vec3)]


class Camera(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("position", Vec3), ("u", Vec3), ("v", Vec3),
                ("fov", C.c_float)]


class TrianglesLayout(C.Structure):
    _fields_ = [("vertex", C.POINTER(Vec3)), ("normal", C.POINTER(Vec3))]


class Object(C.Structure):
    _fields_ = [("triangles", TrianglesLayout), ("triangle_count", C.c_uint32), ("ka", Vec3), ("kd", Vec3),
                ("ks", Vec3), ("ns", C.c_float), ("ni", C.c_float), ("nr", C.c_float), ("d", C.c_float)]


class SceneStruct(C.Structure):
    _fields_ = [("vertex", C.POINTER(Vec3)), ("normal", C.POINTER(Vec3)), ("objects", C.POINTER(Object)),
                ("object_count", C.c_size_t), ("lights", C.POINTER(Light)), ("light_count", C.c_size_t),
                ("camera", Camera), ("triangle_count", C.c_size_t), ("aabbs", C.c_void_p), ("octree", C.c_void_p)]


class FrameOpts(C.Structure):
    _fields_ = [("aliasing", C.c_int), ("tracer", C.c_int), ("rank", C.c_int), ("nranks", C.c_int),
                ("count_rays", C.c_int), ("frame_kernel", C.c_int), ("primary", C.c_int), ("reserved", C.c_int)]


class Stats(C.Structure):
    _fields_ = [("triangles", C.c_uint64), ("objects", C.c_uint64), ("lights", C.c_uint64), ("nodes", C.c_uint64),
                ("leaves", C.c_uint64), ("max_depth", C.c_uint64), ("max_leaf_tris", C.c_uint64),
                ("tree_bytes", C.c_uint64), ("triangle_bytes", C.c_uint64), ("upload_ms", C.c_double),
                ("build_ms", C.c_double), ("frame_ms", C.c_double), ("rays_primary", C.c_uint64),
                ("rays_shadow", C.c_uint64), ("rays_bounce", C.c_uint64), ("node_visits", C.c_uint64),
                ("tri_tests", C.c_uint64), ("kernel_launches", C.c_uint64), ("build_launches", C.c_uint64),
                ("key_mode", C.c_uint64), ("mode_cost", C.c_uint64), ("builder", C.c_uint64)]


assert C.sizeof(Vec3) == 12 and C.sizeof(Light) == 28 and C.sizeof(Camera) == 48
assert C.sizeof(Object) == 72 and C.sizeof(SceneStruct) == 120

# every symbol include/rtb200.h and include/rt_host.h declare with C linkage
ABI_SYMBOLS = [
    "rtb_last_error", "rtb_render", "rtb_render_multi", "rtb_scene_create", "rtb_scene_destroy", "rtb_scene_set_camera",
    "rtb_scene_render", "rtb_scene_render_device", "rtb_scene_sync", "rtb_scene_hit_ids", "rtb_scene_trace_rays",
    "rtb_scene_stats", "rtb_radix_sort_pairs_u32", "rtb_radix_sort_pairs_u64", "rtb_inclusive_scan_u32",
    "rtb_inclusive_scan_u64", "rtb_dev_radix_sort_pairs_u64", "rtb_dev_inclusive_scan_u32", "rtb_dev_inclusive_scan_u32_static", "rtb_scene_dump_tree", "rtb_scene_raster_state", "rtb_scene_profile", "rtb_scene_kernel_ms", "rtb_scene_exchange_create", "rtb_scene_exchange_attach", "rtb_dev_download_async", "rtb_dev_download_wait",
    "rtb_device_count", "rtb_version", "rtb_scene_root", "rtb_set_margin_k", "rtb_set_schedule", "rtb_warmup", "rtb_scene_timer_start",
    "rtb_scene_timer_stop", "rtb_host_alloc", "rtb_host_free", "rtb_dev_alloc", "rtb_dev_free", "rtb_dev_memset",
    "rtb_dev_download", "rtb_scene_mem_alloc", "rtb_scene_mem_free", "rtb_ipc_export", "rtb_ipc_open", "rtb_ipc_close",
    "rtb_parse_svati", "rtb_free_scene", "rtb_scene_reserve", "rtb_scene_mem_pageable", "rtb_write_png", "rtb_cli_main", "rtb_synth_terrain", "rtb_write_svati",
]


class RtbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("rtb200 error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("librtb200.so is not built (run `make -C raytracing-gpu_b200` or "
                               "__graft_entry__.build()); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        L.rtb_last_error.restype = C.c_char_p
        L.rtb_version.restype = C.c_char_p
        L.rtb_host_alloc.restype = C.c_void_p
        L.rtb_host_alloc.argtypes = [C.c_size_t]
        L.rtb_host_free.argtypes = [C.c_void_p]
        L.rtb_scene_create.argtypes = [C.POINTER(SceneStruct), C.c_int, C.c_uint, C.POINTER(C.c_void_p)]
        L.rtb_scene_destroy.argtypes = [C.c_void_p]
        L.rtb_scene_destroy.restype = None
        L.rtb_scene_set_camera.argtypes = [C.c_void_p, C.POINTER(Camera)]
        L.rtb_scene_render.argtypes = [C.c_void_p, C.POINTER(FrameOpts), C.c_void_p, C.c_ssize_t, C.c_int, C.c_int]
        L.rtb_scene_render_device.argtypes = [C.c_void_p, C.POINTER(FrameOpts), C.c_void_p, C.c_ssize_t, C.c_int,
                                              C.c_int, C.c_void_p]
        L.rtb_scene_sync.argtypes = [C.c_void_p]
        L.rtb_render.argtypes = [C.POINTER(SceneStruct), C.c_void_p, C.c_int, C.c_ssize_t, C.c_int, C.c_int]
        L.rtb_scene_hit_ids.argtypes = [C.c_void_p, C.POINTER(FrameOpts), C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.rtb_scene_trace_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                           C.c_void_p]
        L.rtb_scene_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.rtb_scene_dump_tree.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_size_t), C.c_void_p,
                                          C.POINTER(C.c_size_t)]
        L.rtb_scene_root.argtypes = [C.c_void_p]
        L.rtb_set_margin_k.argtypes = [C.c_float]
        L.rtb_set_margin_k.restype = None
        L.rtb_set_schedule.argtypes = [C.c_int]
        L.rtb_set_schedule.restype = None
        L.rtb_scene_timer_start.argtypes = [C.c_void_p]
        L.rtb_scene_timer_stop.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.rtb_radix_sort_pairs_u32.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.rtb_radix_sort_pairs_u64.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.rtb_inclusive_scan_u32.argtypes = [C.c_void_p, C.c_size_t]
        L.rtb_inclusive_scan_u64.argtypes = [C.c_void_p, C.c_size_t]
        L.rtb_dev_alloc.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]
        L.rtb_dev_free.argtypes = [C.c_void_p]
        L.rtb_dev_memset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
        L.rtb_dev_download.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.rtb_dev_download_async.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.rtb_dev_download_wait.argtypes = [C.c_void_p]
        L.rtb_ipc_export.argtypes = [C.c_void_p, C.c_void_p]
        L.rtb_ipc_open.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.rtb_ipc_close.argtypes = [C.c_void_p]
        L.rtb_parse_svati.argtypes = [C.c_char_p, C.POINTER(SceneStruct), C.c_char_p, C.c_size_t]
        L.rtb_free_scene.argtypes = [C.POINTER(SceneStruct)]
        L.rtb_free_scene.restype = None
        L.rtb_write_png.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_ssize_t, C.c_char_p]
        L.rtb_synth_terrain.argtypes = [C.POINTER(SceneStruct), C.c_int, C.c_uint64, C.c_int, C.c_int]
        L.rtb_write_svati.argtypes = [C.POINTER(SceneStruct), C.c_char_p, C.c_int, C.c_int]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RtbError(rc, lib().rtb_last_error().decode("utf-8", "replace"))


def device_count():
    return lib().rtb_device_count()


class HostScene:
    """A `struct scene` with host pointers (the argument of render(), gpu/headers/raytracer.h:6)."""

    def __init__(self):
        self.s = SceneStruct()
        self._owned_by_lib = False
        self._keep = []

    # ---- constructors ----
    @staticmethod
    def parse(path):
        hs = HostScene()
        err = C.create_string_buffer(512)
        rc = lib().rtb_parse_svati(os.fsencode(path), C.byref(hs.s), err, 512)
        if rc != 0:
            raise ValueError(err.value.decode("utf-8", "replace").strip())
        hs._owned_by_lib = True
        return hs

    @staticmethod
    def terrain(grid, width, height, seed=0xB200):
        hs = HostScene()
        if lib().rtb_synth_terrain(C.byref(hs.s), grid, seed, width, height) != 0:
            raise ValueError("rtb_synth_terrain(grid=%d) failed" % grid)
        hs._owned_by_lib = True
        return hs

    @staticmethod
    def from_arrays(vertex, normal, tri_counts, materials, lights, camera):
        """vertex/normal: (T,3,3) float32 in memory order; tri_counts: per object;
        materials: (n_obj, 13) = ka3 kd3 ks3 ns ni nr d; lights: (n_light, 7) = type r g b vx vy vz;
        camera: (12,) = width height pos3 u3 v3 fov (width/height NOT multiplied by aliasing)."""
        hs = HostScene()
        v = np.ascontiguousarray(vertex, dtype=np.float32).reshape(-1, 3)
        n = np.ascontiguousarray(normal, dtype=np.float32).reshape(-1, 3)
        tri_counts = [int(c) for c in tri_counts]
        assert v.shape == n.shape and v.shape[0] == 3 * sum(tri_counts)
        materials = np.asarray(materials, dtype=np.float32).reshape(len(tri_counts), 13)
        lights = np.asarray(lights, dtype=np.float32).reshape(-1, 7)
        objs = (Object * max(1, len(tri_counts)))()
        vp = v.ctypes.data_as(C.POINTER(Vec3))
        np_ = n.ctypes.data_as(C.POINTER(Vec3))
        first = 0
        for i, cnt in enumerate(tri_counts):
            o = objs[i]
            o.triangles.vertex = C.cast(C.addressof(vp.contents) + 36 * first, C.POINTER(Vec3)) if v.size else vp
            o.triangles.normal = C.cast(C.addressof(np_.contents) + 36 * first, C.POINTER(Vec3)) if n.size else np_
            o.triangle_count = cnt
            m = materials[i]
            o.ka = Vec3(*m[0:3]); o.kd = Vec3(*m[3:6]); o.ks = Vec3(*m[6:9])
            o.ns, o.ni, o.nr, o.d = float(m[9]), float(m[10]), float(m[11]), float(m[12])
            first += cnt
        ls = (Light * max(1, len(lights)))()
        for i, l in enumerate(lights):
            ls[i].type = int(l[0]); ls[i].r, ls[i].g, ls[i].b = float(l[1]), float(l[2]), float(l[3])
            ls[i].v = Vec3(float(l[4]), float(l[5]), float(l[6]))
        s = hs.s
        s.vertex, s.normal = vp, np_
        s.objects, s.object_count = objs, len(tri_counts)
        s.lights, s.light_count = ls, len(lights)
        cam = np.asarray(camera, dtype=np.float64)
        s.camera.width, s.camera.height = int(cam[0]), int(cam[1])
        s.camera.position = Vec3(*[float(x) for x in cam[2:5]])
        s.camera.u = Vec3(*[float(x) for x in cam[5:8]])
        s.camera.v = Vec3(*[float(x) for x in cam[8:11]])
        s.camera.fov = float(cam[11])
        s.triangle_count = first
        hs._keep = [v, n, objs, ls]
        return hs

    # ---- views ----
    @property
    def width(self):
        return self.s.camera.width

    @property
    def height(self):
        return self.s.camera.height

    @property
    def triangle_count(self):
        return int(self.s.triangle_count)

    def arrays(self):
        """The inverse of from_arrays (copies)."""
        T = self.triangle_count
        v = np.zeros((T, 3, 3), np.float32)
        n = np.zeros((T, 3, 3), np.float32)
        counts, mats = [], []
        first = 0
        for i in range(self.s.object_count):
            o = self.s.objects[i]
            cnt = int(o.triangle_count)
            if cnt:
                v[first:first + cnt] = np.ctypeslib.as_array(C.cast(o.triangles.vertex, C.POINTER(C.c_float)),
                                                             shape=(cnt, 3, 3))
                n[first:first + cnt] = np.ctypeslib.as_array(C.cast(o.triangles.normal, C.POINTER(C.c_float)),
                                                             shape=(cnt, 3, 3))
            first += cnt
            counts.append(cnt)
            mats.append([o.ka.x, o.ka.y, o.ka.z, o.kd.x, o.kd.y, o.kd.z, o.ks.x, o.ks.y, o.ks.z, o.ns, o.ni, o.nr, o.d])
        lights = []
        for i in range(self.s.light_count):
            l = self.s.lights[i]
            lights.append([l.type, l.r, l.g, l.b, l.v.x, l.v.y, l.v.z])
        c = self.s.camera
        cam = [c.width, c.height, c.position.x, c.position.y, c.position.z, c.u.x, c.u.y, c.u.z, c.v.x, c.v.y, c.v.z,
               c.fov]
        return dict(vertex=v, normal=n, tri_counts=np.array(counts, np.int64),
                    materials=np.array(mats, np.float32).reshape(len(counts), 13),
                    lights=np.array(lights, np.float32).reshape(len(lights), 7), camera=np.array(cam, np.float64))

    def with_resolution(self, width, height):
        """Same scene, other camera resolution (what rewriting the `camera W H` tokens does)."""
        self.s.camera.width, self.s.camera.height = int(width), int(height)
        return self

    def scaled(self, aliasing):
        """Copy of the struct with camera.width/height multiplied by aliasing (gpu/rt.cpp:78-79)."""
        s2 = SceneStruct()
        C.memmove(C.byref(s2), C.byref(self.s), C.sizeof(SceneStruct))
        s2.camera.width *= aliasing
        s2.camera.height *= aliasing
        return s2

    def write_svati(self, path):
        if lib().rtb_write_svati(C.byref(self.s), os.fsencode(path), self.width, self.height) != 0:
            raise OSError("cannot write %s" % path)

    def close(self):
        if self._owned_by_lib:
            lib().rtb_free_scene(C.byref(self.s))
            self._owned_by_lib = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_multi(hs, n_devices, aliasing=3, out=None):
    """rtb_render_multi: render() on n_devices GPUs of this process, frame composed on device 0."""
    w, h = hs.width, hs.height
    if out is None:
        out = np.empty((h, w, 4), np.uint8)
    s = hs.scaled(aliasing)
    fn = lib().rtb_render_multi
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_ssize_t, C.c_int, C.c_int, C.c_int]
    _check(fn(C.byref(s), out.ctypes.data, aliasing, w * 4, h, w, n_devices))
    return out


def render_oneshot(hs, aliasing=3, out=None):
    """rtb_render == render() of the reference: upload + build + frame + download."""
    w, h = hs.width, hs.height
    if out is None:
        out = np.empty((h, w, 4), np.uint8)
    s = hs.scaled(aliasing)
    _check(lib().rtb_render(C.byref(s), out.ctypes.data, aliasing, w * 4, h, w))
    return out


class PinnedBuffer:
    def __init__(self, nbytes):
        self.ptr = lib().rtb_host_alloc(nbytes)
        if not self.ptr:
            raise MemoryError("rtb_host_alloc(%d) failed" % nbytes)
        self.nbytes = nbytes

    def array(self, shape, dtype=np.uint8):
        buf = (C.c_uint8 * self.nbytes).from_address(self.ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    def free(self):
        if self.ptr:
            lib().rtb_host_free(self.ptr)
            self.ptr = None


class DeviceScene:
    """rtb_scene_t: the resident scene + hierarchy (replaces to_cuda() + create_octree())."""

    def __init__(self, hs, aliasing=3, device=0, flags=BUILD_DEFAULT):
        self.h = C.c_void_p()
        self.aliasing = aliasing
        self.frame_kernel = 0
        self.primary = 0
        self.out_w, self.out_h = hs.width, hs.height
        s = hs.scaled(aliasing)
        _check(lib().rtb_scene_create(C.byref(s), device, flags, C.byref(self.h)))

    def _opts(self, tracer, rank, nranks, count):
        o = FrameOpts()
        o.aliasing, o.tracer, o.rank, o.nranks, o.count_rays = self.aliasing, tracer, rank, nranks, int(count)
        o.frame_kernel = self.frame_kernel    # 0: per-sample kernel (product path), 1: wavefront kernel
        o.primary = self.primary              # 0: auto, 1: tree walk, 2: projection (raster.cuh)
        return o

    def set_camera(self, cam_struct):
        _check(lib().rtb_scene_set_camera(self.h, C.byref(cam_struct)))

    def render(self, tracer=TRACER_TREE, rank=0, nranks=1, count=False, out=None):
        if out is None:
            out = np.zeros((self.out_h, self.out_w, 4), np.uint8)
        o = self._opts(tracer, rank, nranks, count)
        _check(lib().rtb_scene_render(self.h, C.byref(o), out.ctypes.data, out.strides[0], self.out_h, self.out_w))
        return out

    def render_device(self, d_ptr, pitch, tracer=TRACER_TREE, rank=0, nranks=1, count=False, stream=None):
        """stream: a cudaStream_t as an integer (None: the handle's own stream)"""
        o = self._opts(tracer, rank, nranks, count)
        _check(lib().rtb_scene_render_device(self.h, C.byref(o), d_ptr, pitch, self.out_h, self.out_w,
                                             C.c_void_p(stream) if stream else None))

    def sync(self):
        _check(lib().rtb_scene_sync(self.h))

    def timer_start(self):
        _check(lib().rtb_scene_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_double()
        _check(lib().rtb_scene_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def hit_ids(self, tracer=TRACER_TREE, want_obj=True):
        W, H = self.out_w * self.aliasing, self.out_h * self.aliasing
        tri = np.empty((H, W), np.int32)
        obj = np.empty((H, W), np.int32) if want_obj else None
        o = self._opts(tracer, 0, 1, False)
        _check(lib().rtb_scene_hit_ids(self.h, C.byref(o), tri.ctypes.data, obj.ctypes.data if want_obj else None,
                                       self.out_h, self.out_w))
        return tri, obj

    def trace_rays(self, rays, tracer=TRACER_TREE, any_hit=False):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        tri = np.empty(n, np.int32)
        dist = np.empty(n, np.float32)
        _check(lib().rtb_scene_trace_rays(self.h, tracer, int(any_hit), rays.ctypes.data, n, tri.ctypes.data,
                                          dist.ctypes.data))
        return tri, dist

    def stats(self):
        st = Stats()
        _check(lib().rtb_scene_stats(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in Stats._fields_}

    def raster_state(self):
        """rtb_scene_raster_state: items, gave-up flag, samples the items cover, triangles with two-part regions, light grids"""
        out = (C.c_uint * 8)()
        fn = lib().rtb_scene_raster_state
        fn.argtypes = [C.c_void_p, C.c_void_p]
        _check(fn(self.h, out))
        return dict(items=out[0], gave_up=out[1], area=out[2] | (out[3] << 32), two_part_regions=out[4], light_grids=out[5],
                    grid_entries_k=out[6], grid_cells_k=out[7])

    def exchange_create(self, nranks):
        """rtb_scene_exchange_create: this rank's exchange block (device pointer, bytes)"""
        ptr, n = C.c_void_p(), C.c_size_t()
        fn = lib().rtb_scene_exchange_create
        fn.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        _check(fn(self.h, nranks, C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def exchange_attach(self, rank, nranks, blocks):
        arr = (C.c_void_p * nranks)(*[C.c_void_p(b) for b in blocks])
        fn = lib().rtb_scene_exchange_attach
        fn.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        _check(fn(self.h, rank, nranks, arr))

    def profile(self, on=True):
        fn = lib().rtb_scene_profile
        fn.argtypes = [C.c_void_p, C.c_int]
        _check(fn(self.h, int(on)))

    def kernel_ms(self):
        """(setup phase, k_rast_items, frame kernel, whole frame, then the parts of the setup phase: clears +
        k_rast_setup, k_rast_big, k_xchg_sync, 0) of the last profiled frame, ms"""
        out = (C.c_double * 8)()
        fn = lib().rtb_scene_kernel_ms
        fn.argtypes = [C.c_void_p, C.c_void_p]
        _check(fn(self.h, out))
        return [float(x) for x in out]

    def dump_tree(self):
        nn, nt = C.c_size_t(), C.c_size_t()
        _check(lib().rtb_scene_dump_tree(self.h, None, C.byref(nn), None, C.byref(nt)))
        st = self.stats()
        floats = int(st["tree_bytes"] // max(1, st["nodes"])) // 4      # 32 per 4-wide node (16: binary A/B build)
        nodes = np.empty((nn.value, floats), np.float32)
        perm = np.empty(nt.value, np.uint32)
        _check(lib().rtb_scene_dump_tree(self.h, nodes.ctypes.data, C.byref(nn), perm.ctypes.data, C.byref(nt)))
        return nodes, perm, lib().rtb_scene_root(self.h)

    def close(self):
        if self.h:
            lib().rtb_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def radix_sort_pairs(keys, vals=None):
    keys = np.ascontiguousarray(keys)
    out_k = keys.copy()
    out_v = None if vals is None else np.ascontiguousarray(vals, np.uint32).copy()
    fn = {np.dtype(np.uint32): lib().rtb_radix_sort_pairs_u32, np.dtype(np.uint64): lib().rtb_radix_sort_pairs_u64}[
        keys.dtype]
    _check(fn(out_k.ctypes.data, None if out_v is None else out_v.ctypes.data, out_k.size))
    return out_k, out_v


def inclusive_scan(data):
    out = np.ascontiguousarray(data).copy()
    fn = {np.dtype(np.uint32): lib().rtb_inclusive_scan_u32, np.dtype(np.uint64): lib().rtb_inclusive_scan_u64}[
        out.dtype]
    _check(fn(out.ctypes.data, out.size))
    return out
