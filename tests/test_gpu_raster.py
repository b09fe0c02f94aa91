"""Primary visibility by projection (csrc/raster.cuh) and shadow rays through the light grids
(csrc/lightgrid.cuh) against EXHAUSTIVE search on the GPU (RTB_TRACER_BRUTE: the reference's
collide() loop, gpu/hit.cu:84-114, pinned bit for bit to the oracle in tests/test_gpu_parity.py).
The culls of both passes are bounds (csrc/region.cuh), so ids and frame bytes must be those of the
exhaustive tracer exactly; the tree walk culls with padded boxes (a heuristic)
and is allowed a few samples on needle soups.  Frames where the projection does not pay fall back to
the tree by themselves; rtb_scene_hit_ids(RASTER) then reports -2 everywhere, which the tests below
tell apart from a pass."""
import numpy as np
import pytest

import pyrtb
from util import golden_names, load_golden, random_scene

pytestmark = pytest.mark.gpu
AUTO, TREE, RASTER = 0, 1, 2


@pytest.fixture(scope="module")
def gpu():
    if pyrtb.device_count() < 1:
        pytest.skip("no CUDA device")


def _ids(ds, primary):
    ds.primary = primary
    tri, obj = ds.hit_ids(pyrtb.TRACER_TREE)
    ds.primary = AUTO
    return tri, obj


def _frame(ds, primary, **kw):
    ds.primary = primary
    img = ds.render(**kw)
    ds.primary = AUTO
    return img


@pytest.mark.parametrize("aliasing", [1, 3])
def test_terrain_ids_and_frames(gpu, aliasing):
    hs = pyrtb.HostScene.terrain(200, 640, 360)          # 80 000 triangles
    ds = pyrtb.DeviceScene(hs, aliasing)
    try:
        want, wobj = ds.hit_ids(pyrtb.TRACER_BRUTE)
        got, gobj = _ids(ds, RASTER)
        assert (got != -2).all(), "the projection must be active on the terrain: %r" % (ds.raster_state(),)
        assert np.array_equal(got, want) and np.array_equal(gobj, wobj)
        assert (want >= 0).mean() > 0.3
        tree, _ = _ids(ds, TREE)
        assert int((tree != want).sum()) <= 2
        fb = ds.render(tracer=pyrtb.TRACER_BRUTE)
        assert np.array_equal(_frame(ds, RASTER), fb)
        assert np.array_equal(_frame(ds, AUTO), fb)
        assert ds.stats()["kernel_launches"] == 4, "setup + big + items + frame"
        ft = _frame(ds, TREE)
        assert ds.stats()["kernel_launches"] == 1
        assert int((ft != fb).any(axis=-1).sum()) <= 2
        # the counting pass (device ray counters) always walks the tree
        assert np.array_equal(ds.render(count=True), ft)
    finally:
        ds.close()


def test_terrain_camera_moves_and_odd_sizes(gpu):
    hs = pyrtb.HostScene.terrain(200, 347, 199)           # not a multiple of 32 samples nor of the warp tile
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        for k in range(3):
            import ctypes as C
            c2 = pyrtb.Camera()
            C.memmove(C.byref(c2), C.byref(hs.scaled(3).camera), C.sizeof(pyrtb.Camera))
            c2.position.x += 40.0 * k
            c2.position.y -= 30.0 * k
            ds.set_camera(c2)
            want, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
            got, _ = _ids(ds, RASTER)
            assert (got != -2).all(), ds.raster_state()
            assert np.array_equal(got, want), "camera %d" % k
            assert np.array_equal(_frame(ds, RASTER), ds.render(tracer=pyrtb.TRACER_BRUTE))
    finally:
        ds.close()


@pytest.mark.parametrize("name", golden_names())
def test_golden_scenes_through_the_projection(gpu, name):
    """every shipped scene goes through the projection (no fall-back) and matches the fixtures"""
    hs, z = load_golden(name)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        got, _ = _ids(ds, RASTER)
        assert (got != -2).all(), ds.raster_state()
        assert np.array_equal(got, z["tri"].astype(np.int32))
        fr = _frame(ds, RASTER)
        assert ds.stats()["kernel_launches"] == 4
        assert np.array_equal(fr, _frame(ds, AUTO)), "the projection is what AUTO runs"
        assert np.array_equal(fr, _frame(ds, TREE))
    finally:
        ds.close()


def test_multi_rank_tiles_compose_with_the_projection(gpu):
    hs = pyrtb.HostScene.terrain(250, 640, 400)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        full = _frame(ds, RASTER)
        for nranks in (2, 3, 8):
            acc = np.zeros_like(full)
            for r in range(nranks):
                part = _frame(ds, RASTER, rank=r, nranks=nranks)
                assert not ds.raster_state()["gave_up"]
                mask = part[..., 3] != 0
                acc[mask] = part[mask]
            assert np.array_equal(acc, full), nranks
    finally:
        ds.close()


def _floor_scene(n_floor, n_small=400):
    """a floor of n_floor strips under and behind the camera (their boxes cross the plane through the
    camera position) and n_small triangles in front of it"""
    tris, norms = [], []
    # floor strips along x, each crossing z = camera z
    for k in range(n_floor):
        x0, x1 = -8 + 16 * k / n_floor, -8 + 16 * (k + 1) / n_floor
        tris += [[[x0, -1, -30], [x1, -1, -30], [x1, -1, 30]], [[x0, -1, -30], [x1, -1, 30], [x0, -1, 30]]]
        norms += [[[0, 1, 0]] * 3] * 2
    rng = np.random.default_rng(5)
    for _ in range(n_small):                           # small triangles in front of the camera
        c = np.array([rng.uniform(-6, 6), rng.uniform(-0.9, 3), rng.uniform(2, 20)])
        t = c + rng.standard_normal((3, 3)) * 0.3
        tris.append(t.tolist()); norms.append([[0, 0, -1]] * 3)
    v = np.array(tris, np.float32); n = np.array(norms, np.float32)
    mats = [[0.2, 0.2, 0.2, 0.7, 0.7, 0.7, 0.3, 0.3, 0.3, 20, 1, 0, 1]]
    lights = [[pyrtb.AMBIENT, .3, .3, .3, 0, 0, 0], [pyrtb.DIRECTIONAL, 1, 1, 1, 0.3, -1, 0.5]]
    cam = [160, 120, 0, 1, -4, 1, 0, 0, 0, -1, 0, 70]
    return pyrtb.HostScene.from_arrays(v, n, [len(tris)], mats, lights, cam)


def test_triangles_across_the_camera_plane(gpu):
    """a floor under and behind the camera: its triangles cross the plane through the camera position
    parallel to the image plane, where a projected vertex is meaningless.  The edge functions of
    region.cuh do not project vertices: no special list, no fall-back, exhaustive-search ids."""
    for n_floor in (3, 20):
        hs = _floor_scene(n_floor)
        ds = pyrtb.DeviceScene(hs, 3)
        try:
            want, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
            got, _ = _ids(ds, RASTER)
            assert (got != -2).all(), ds.raster_state()
            assert np.array_equal(got, want)
            assert (want[want >= 0] < 2 * n_floor).any(), "the floor is visible"
            assert np.array_equal(_frame(ds, RASTER), ds.render(tracer=pyrtb.TRACER_BRUTE))
        finally:
            ds.close()


@pytest.mark.parametrize("T", [3000, 40000])
def test_needle_soups(gpu, T):
    """30 % needles 0.003 units wide, mirrors: the projection has the exhaustive-search ids exactly;
    the tree walk (bounce rays, and everything with primary = TREE) may lose <= 3 samples"""
    hs = random_scene(900 + T, T, 200, 120)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        got, _ = _ids(ds, RASTER)
        brute, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
        if (got == -2).all():                             # a soup that covers the frame > 20 times keeps the tree walk
            assert ds.raster_state()["area"] > 20 * 600 * 360
            return
        assert np.array_equal(got, brute)
        want, _ = _ids(ds, TREE)
        assert int((want != brute).sum()) <= 3
        fb = ds.render(tracer=pyrtb.TRACER_BRUTE)
        assert int((_frame(ds, RASTER) != fb).any(axis=-1).sum()) <= 3      # bounce and shadow rays walk the tree
    finally:
        ds.close()


def _stack_scene(n):
    """n big triangles one behind the other, each covering the whole frame: the rectangles cover the
    frame n times"""
    tris = []
    for k in range(n):
        z = 5.0 + 0.01 * k
        tris.append([[-60, -40, z], [60, -40, z], [0, 80, z]])
    v = np.array(tris, np.float32)
    nr = np.zeros_like(v); nr[..., 2] = -1
    mats = [[0.2, 0.2, 0.2, 0.7, 0.7, 0.7, 0.3, 0.3, 0.3, 20, 1, 0, 1]]
    lights = [[pyrtb.AMBIENT, .3, .3, .3, 0, 0, 0], [pyrtb.DIRECTIONAL, 1, 1, 1, 0.3, -1, 0.5]]
    cam = [160, 120, 0, 1, -4, 1, 0, 0, 0, -1, 0, 70]
    return pyrtb.HostScene.from_arrays(v, nr, [n], mats, lights, cam)


def test_auto_does_not_retry_a_camera_whose_frame_fell_back(gpu):
    import ctypes as C
    hs = _stack_scene(40)                                 # rectangles cover the frame 40 times (> 20): the pass gives up
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        want = _frame(ds, TREE)
        assert np.array_equal(want, ds.render(tracer=pyrtb.TRACER_BRUTE))
        a = ds.render()
        st = ds.raster_state()
        assert st["area"] > 20 * 480 * 360 and ds.stats()["kernel_launches"] == 4
        b = ds.render()
        assert ds.stats()["kernel_launches"] == 1, "same camera: straight to the tree"
        assert np.array_equal(a, want) and np.array_equal(b, want)
        ids, _ = _ids(ds, RASTER)
        assert (ids == -2).all()
        ids, _ = _ids(ds, AUTO)
        assert np.array_equal(ids, ds.hit_ids(pyrtb.TRACER_BRUTE)[0]), "AUTO reports what the frame uses: the tree walk"
        cam = pyrtb.Camera()
        C.memmove(C.byref(cam), C.byref(hs.scaled(3).camera), C.sizeof(pyrtb.Camera))
        ds.set_camera(cam)                                # a new camera generation is a new chance
        ds.render()
        assert ds.stats()["kernel_launches"] == 4
        assert np.array_equal(_frame(ds, RASTER), want), "on request the pass runs (and falls back) every time"
        assert ds.stats()["kernel_launches"] == 4
    finally:
        ds.close()


NO_LIGHT_GRIDS = 16          # include/rtb200.h RTB_BUILD_NO_LIGHT_GRIDS


@pytest.mark.parametrize("scene", ["terrain", "terrain-low-light", "soup", "floor"])
def test_shadow_ray_grids_give_the_tree_walk_frames(gpu, scene):
    """per-light projection grids (csrc/lightgrid.cuh) against shadow rays that walk the tree"""
    if scene == "terrain":
        hs = pyrtb.HostScene.terrain(220, 640, 360)
    elif scene == "terrain-low-light":
        # the point light down among the hills and the directional light grazing: long shadows,
        # boxes across the light's plane, rays whose point falls off the grid
        hs = pyrtb.HostScene.terrain(220, 640, 360)
        a = hs.arrays()
        lights = a["lights"].copy()
        for l in lights:
            if int(l[0]) == pyrtb.POINT:
                l[4:7] = [30.0, 12.0, 420.0]
            elif int(l[0]) == pyrtb.DIRECTIONAL:
                l[4:7] = [1.0, -0.12, 0.3]
        hs = pyrtb.HostScene.from_arrays(a["vertex"], a["normal"], a["tri_counts"], a["materials"], lights, a["camera"])
    elif scene == "soup":
        hs = random_scene(77, 40000, 320, 200)           # needles, mirrors (bounce hits cast shadow rays too)
    else:
        hs = _floor_scene(6, 40000)
    with_grids = pyrtb.DeviceScene(hs, 3)
    without = pyrtb.DeviceScene(hs, 3, flags=NO_LIGHT_GRIDS)
    try:
        assert without.raster_state()["light_grids"] == 0
        n_grids = with_grids.raster_state()["light_grids"]
        if scene == "terrain":
            assert n_grids == 2
        elif scene == "terrain-low-light":
            assert n_grids >= 1      # the light among the hills may keep the tree walk: boxes cross its plane
        for primary in (TREE, AUTO):
            a = _frame(with_grids, primary)
            b = _frame(without, primary)
            diff = int((a != b).any(axis=-1).sum())
            # needle soups: the grid culls with each triangle's own box, the tree with leaf boxes, so the
            # ill-conditioned acceptances of tests/test_gpu_random_scenes.py may fall differently
            assert diff <= (3 if scene == "soup" else 0), (scene, primary, diff, n_grids)
    finally:
        with_grids.close()
        without.close()


ONE_FRAME = 8                # include/rtb200.h RTB_BUILD_ONE_FRAME


def test_one_frame_builds_make_light_grids_only_when_the_frame_repays_them(gpu):
    """rtb_render() builds for one frame: light grids from 4 samples per triangle on, and the same bytes"""
    hs = pyrtb.HostScene.terrain(200, 640, 360)          # 80 000 triangles, 2.07 M samples at 3x3
    ref = pyrtb.DeviceScene(hs, 3, flags=NO_LIGHT_GRIDS)
    one = pyrtb.DeviceScene(hs, 3, flags=ONE_FRAME)
    try:
        assert one.raster_state()["light_grids"] == 2
        want = _frame(ref, TREE)
        assert np.array_equal(one.render(), want)
        assert np.array_equal(pyrtb.render_oneshot(hs, 3), want)
    finally:
        ref.close(); one.close()
    small = pyrtb.HostScene.terrain(200, 96, 54)         # 46 656 samples: a tree walk per shadow ray is cheaper
    ds = pyrtb.DeviceScene(small, 3, flags=ONE_FRAME)
    ref = pyrtb.DeviceScene(small, 3, flags=NO_LIGHT_GRIDS)
    try:
        assert ds.raster_state()["light_grids"] == 0
        assert np.array_equal(ds.render(), _frame(ref, TREE))
    finally:
        ds.close(); ref.close()


def test_grid_extent_from_a_sample_of_the_triangles(gpu):
    """from 2^20 triangles on the grids' extent comes from every 16th rectangle: rays whose point
    falls off the grid walk the tree, triangles beyond it are clipped -- same frame"""
    hs = pyrtb.HostScene.terrain(750, 480, 270)          # 1 125 000 triangles
    a = pyrtb.DeviceScene(hs, 3)
    b = pyrtb.DeviceScene(hs, 3, flags=NO_LIGHT_GRIDS)
    try:
        assert a.raster_state()["light_grids"] == 2
        for primary in (TREE, AUTO):
            assert np.array_equal(_frame(a, primary), _frame(b, primary))
    finally:
        a.close(); b.close()


def test_frames_wider_than_the_item_corners_walk_the_tree(gpu):
    """items hold 16-bit sample corners: a 70 000-sample-wide frame keeps the tree walk, on request too"""
    hs = pyrtb.HostScene.terrain(200, 70000, 2)
    ds = pyrtb.DeviceScene(hs, 1)
    try:
        got, _ = _ids(ds, RASTER)
        want, _ = _ids(ds, TREE)
        assert np.array_equal(got, want), "the request is ignored, not half honoured"
        assert np.array_equal(_frame(ds, RASTER), _frame(ds, TREE))
        assert ds.stats()["kernel_launches"] == 1
    finally:
        ds.close()


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_sharded_projection_pass_with_item_exchange(gpu, nranks):
    """rtb_scene_exchange_*: every rank projects 1/N of the triangles and stores the items into the owners'
    inboxes (peer stores between GPUs; here N handles on ONE device, each on its own stream, so the
    device-side epoch barrier really waits for the other ranks' kernels).  Runs in a process of its own
    with CUDA_DEVICE_MAX_CONNECTIONS=32: with the default 8 hardware queues the streams of 8 "ranks" on one
    device alias, and a spinning barrier kernel can sit in front of the kernel it waits for (real ranks
    have a device each).  tests/native/exchange_probe.py has the body."""
    import os
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS="32")
    r = subprocess.run([sys.executable, os.path.join(here, "native", "exchange_probe.py"), str(nranks)], env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "exchange ok" in r.stdout
