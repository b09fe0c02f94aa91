"""GPU parity tests (run on the B200 box): the CUDA path through the C ABI against
 (1) the committed golden fixtures of the reference's 20 scenes (tests/golden, produced from
     the reference's own sources, see make_golden.py) and
 (2) the CPU oracle run on the box for synthetic scenes and arbitrary rays.
Bars: primary hit ids bit-exact on 100 % of samples; RGB within +-1/255 on >= 99.9 % of
pixels (powf is the only non-reproducible function)."""
import numpy as np
import pytest

import oracle_py
import pyrtb
from util import golden_names, load_golden, primary_rays, rgb_report

pytestmark = pytest.mark.gpu

RGB_TOL_FRACTION = 1e-3   # at most 0.1 % of pixels may differ by more than 1/255


@pytest.mark.parametrize("name", golden_names())
def test_golden_scene(gpu, name):
    hs, z = load_golden(name)
    ds = pyrtb.DeviceScene(hs, aliasing=3)
    try:
        tri, obj = ds.hit_ids(pyrtb.TRACER_TREE)
        assert np.array_equal(tri, z["tri"].astype(np.int32)), "%d samples with a wrong triangle id" % int(
            (tri != z["tri"]).sum())
        assert np.array_equal(obj, z["obj"].astype(np.int32))
        img = ds.render(pyrtb.TRACER_TREE, count=True)
        bad, same, mx = rgb_report(img, z["rgba"])
        print("%s: identical %.5f, >1 off %.6f, max diff %d" % (name, same, bad, mx))
        assert bad <= RGB_TOL_FRACTION
        assert (img[..., 3] == 255).all()
        st = ds.stats()
        assert [st["rays_primary"], st["rays_shadow"], st["rays_bounce"]] == [int(c) for c in z["counts"]]
        # the timed (non-counting) kernel gives the same bytes
        assert np.array_equal(ds.render(pyrtb.TRACER_TREE), img)
        # the staged wavefront kernel (opt-in) produces the same bytes and the same ray counts
        ds.frame_kernel = 1
        assert np.array_equal(ds.render(pyrtb.TRACER_TREE, count=True), img)
        st2 = ds.stats()
        assert [st2["rays_primary"], st2["rays_shadow"], st2["rays_bounce"]] == [int(c) for c in z["counts"]]
    finally:
        ds.close()


@pytest.mark.parametrize("name", [n for n in golden_names() if n in (
    "triangle", "cube", "dir-light-shadows", "point-light", "island_smooth", "lighthouse")])
def test_golden_scene_brute_force_tracer(gpu, name):
    hs, z = load_golden(name)
    ds = pyrtb.DeviceScene(hs, aliasing=3, flags=pyrtb.BUILD_NO_TREE)
    try:
        tri, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
        assert np.array_equal(tri, z["tri"].astype(np.int32))
        bad, same, mx = rgb_report(ds.render(pyrtb.TRACER_BRUTE), z["rgba"])
        assert bad <= RGB_TOL_FRACTION
    finally:
        ds.close()


def test_oneshot_render_is_the_resident_render(gpu):
    hs, z = load_golden("cube")
    a = pyrtb.render_oneshot(hs, 3)
    ds = pyrtb.DeviceScene(hs, 3)
    b = ds.render()
    ds.close()
    assert np.array_equal(a, b)
    assert rgb_report(a, z["rgba"])[0] <= RGB_TOL_FRACTION


@pytest.mark.parametrize("aliasing", [1, 2, 3])
def test_terrain_small_against_oracle(gpu, aliasing):
    hs = pyrtb.HostScene.terrain(48, 160, 90)
    o = oracle_py.orc_render(hs, aliasing)
    ds = pyrtb.DeviceScene(hs, aliasing)
    try:
        tri, obj = ds.hit_ids()
        assert np.array_equal(tri, o["tri"])
        assert np.array_equal(obj, o["obj"])
        img = ds.render(count=True)
        assert rgb_report(img, o["rgba"])[0] <= RGB_TOL_FRACTION
        st = ds.stats()
        assert [st["rays_primary"], st["rays_shadow"], st["rays_bounce"]] == [int(c) for c in o["counts"]]
        assert (o["tri"] >= 0).mean() > 0.3, "camera should see terrain"
    finally:
        ds.close()


def test_odd_resolution_and_tile_edges(gpu):
    hs = pyrtb.HostScene.terrain(16, 37, 23)      # not a multiple of the 8x4 warp tile nor of 32
    o = oracle_py.orc_render(hs, 3)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        assert np.array_equal(ds.hit_ids()[0], o["tri"])
        assert rgb_report(ds.render(), o["rgba"])[0] <= RGB_TOL_FRACTION
    finally:
        ds.close()


def test_tree_matches_brute_force_on_large_terrain_samples(gpu):
    """Size-independent property: culling never changes the winner.  ~0.5 M triangles,
    primary + shadow + random rays, hierarchy vs exhaustive search on the GPU."""
    hs = pyrtb.HostScene.terrain(500, 1920, 1080)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        rng = np.random.default_rng(7)
        n = 60000
        xy = np.stack([rng.integers(0, 1920 * 3, n), rng.integers(0, 1080 * 3, n)], axis=1)
        rays = primary_rays(hs, 3, xy)
        t_tree, d_tree = ds.trace_rays(rays, pyrtb.TRACER_TREE)
        t_brute, d_brute = ds.trace_rays(rays, pyrtb.TRACER_BRUTE)
        assert np.array_equal(t_tree, t_brute)
        assert np.array_equal(d_tree.view(np.uint32), d_brute.view(np.uint32))
        assert (t_tree >= 0).mean() > 0.3
        # shadow-like rays from the hit points towards both lights + random directions
        hitp = rays[:, :3] + rays[:, 3:] * d_tree[:, None]
        sel = t_tree >= 0
        org = hitp[sel][:20000]
        dirs = np.concatenate([np.tile(np.array([[-0.5, 1.0, -0.5]], np.float32), (len(org) // 2, 1)),
                               (np.array([[0, 300, 500]], np.float32) - org[len(org) // 2:])]).astype(np.float32)
        srays = np.concatenate([org, dirs], axis=1)
        a_tree, _ = ds.trace_rays(srays, pyrtb.TRACER_TREE, any_hit=True)
        a_brute, _ = ds.trace_rays(srays, pyrtb.TRACER_BRUTE, any_hit=True)
        assert np.array_equal(a_tree, a_brute)
        rdir = rng.normal(size=(len(org), 3)).astype(np.float32)
        rr = np.concatenate([org, rdir], axis=1)
        n_tree, nd_tree = ds.trace_rays(rr, pyrtb.TRACER_TREE)
        n_brute, nd_brute = ds.trace_rays(rr, pyrtb.TRACER_BRUTE)
        assert np.array_equal(n_tree, n_brute)
        assert np.array_equal(nd_tree.view(np.uint32), nd_brute.view(np.uint32))
    finally:
        ds.close()


def test_trace_rays_against_oracle(gpu):
    hs, _ = load_golden("dir-light-shadows")
    rng = np.random.default_rng(3)
    org = rng.uniform(-6, 6, size=(4000, 3)).astype(np.float32)
    d = rng.normal(size=(4000, 3)).astype(np.float32)
    rays = np.concatenate([org, d], axis=1)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        for tracer in (pyrtb.TRACER_TREE, pyrtb.TRACER_BRUTE):
            tri, dist = ds.trace_rays(rays, tracer)
            otri, odist, _ = oracle_py.orc_trace_rays(hs, rays)
            assert np.array_equal(tri, otri)
            assert np.array_equal(dist.view(np.uint32), odist.view(np.uint32))
            occ, _ = ds.trace_rays(rays, tracer, any_hit=True)
            assert np.array_equal(occ >= 0, otri >= 0)
    finally:
        ds.close()


def test_multi_rank_tiles_compose_the_frame(gpu):
    hs, z = load_golden("cube")
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        full = ds.render()
        for nranks in (2, 4, 8):
            acc = np.zeros_like(full)
            cover = np.zeros(full.shape[:2], np.int32)
            for r in range(nranks):
                part = ds.render(rank=r, nranks=nranks)
                mask = part[..., 3] != 0
                cover += mask
                acc[mask] = part[mask]
            assert (cover == 1).all(), "every pixel belongs to exactly one rank"
            assert np.array_equal(acc, full)
    finally:
        ds.close()


def test_build_is_deterministic_and_structurally_sound(gpu):
    hs = pyrtb.HostScene.terrain(120, 64, 64)
    a = pyrtb.DeviceScene(hs, 1)
    b = pyrtb.DeviceScene(hs, 1)
    try:
        na, pa, ra = a.dump_tree()
        nb, pb, rb = b.dump_tree()
        assert ra == rb and np.array_equal(pa, pb) and np.array_equal(na.view(np.uint32), nb.view(np.uint32))
        T = hs.triangle_count
        assert sorted(pa.tolist()) == list(range(T)), "leaf order is a permutation of the triangles"
        # 64-byte binary nodes (product) or 128-byte 4-wide nodes (RTB_BVH4 A/B build): rtb_internal.h
        wide = na.shape[1] == 32
        fan = 4 if wide else 2
        refs = (na[:, 24:28] if wide else na[:, 12:14]).copy().view(np.int32)
        rng_lo = na[:, 28 if wide else 14].copy().view(np.int32)
        rng_cnt = na[:, 29 if wide else 15].copy().view(np.int32)
        zoff = 16 if wide else 8
        seen = np.zeros(T, np.int32)
        verts = hs.arrays()["vertex"][pa]        # leaf order
        stack = [ra]
        visited = 0
        tight = planes = 0
        while stack:
            k = stack.pop()
            visited += 1
            lo, cnt = int(rng_lo[k]), int(rng_cnt[k])
            covered = 0
            for slot in range(fan):
                ref = int(refs[k, slot])
                # (min.x, max.x, min.y, max.y) per child first, then the (min.z, max.z) pairs
                box = (na[k, 4 * slot], na[k, 4 * slot + 1], na[k, 4 * slot + 2], na[k, 4 * slot + 3],
                       na[k, zoff + 2 * slot], na[k, zoff + 2 * slot + 1])
                if ref >= 0:
                    stack.append(ref)
                    f, c = int(rng_lo[ref]), int(rng_cnt[ref])
                else:
                    packed = ~ref
                    f, c = packed >> 3, packed & 7
                    if c == 0:                  # unused slot: inverted box
                        assert box[0] > box[1]
                        continue
                    assert 1 <= c <= 3
                    seen[f:f + c] += 1
                assert lo <= f and f + c <= lo + cnt
                covered += c
                pts = verts[f:f + c].reshape(-1, 3)
                # boxes contain their triangles; they are tight except for the needle padding
                # (build.cu), which never exceeds the triangle's longest edge
                slack = 1.5     # terrain edges are < 1.5 units at this grid
                for ax in range(3):
                    assert box[2 * ax] <= pts[:, ax].min() <= box[2 * ax] + slack
                    assert box[2 * ax + 1] - slack <= pts[:, ax].max() <= box[2 * ax + 1]
                    tight += int(box[2 * ax] == pts[:, ax].min()) + int(box[2 * ax + 1] == pts[:, ax].max())
                    planes += 2
            assert covered == cnt, "the children partition the node's slot range"
        assert (seen == 1).all(), "every triangle is in exactly one leaf"
        assert visited == len(na)
        assert tight > 0.5 * planes, "most boxes are exact min/max of their vertices"
        st = a.stats()
        assert st["nodes"] == len(na) and st["max_leaf_tris"] <= 3
    finally:
        a.close()
        b.close()


def test_rt_cli_end_to_end(gpu, tmp_path):
    """rt file.svati output.png (gpu/rt.cpp:54-97): parse -> render -> PNG, and the .ppm variant."""
    import os
    import subprocess
    from PIL import Image
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rt = os.path.join(root, "raytracing-gpu_b200", "rt")
    for name in ("cube", "dir-light-shadows"):
        hs, z = load_golden(name)
        svati, png = str(tmp_path / (name + ".svati")), str(tmp_path / (name + ".png"))
        hs.write_svati(svati)
        r = subprocess.run([rt, svati, png], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        assert "Using structure of arrays (SOA) layout" in r.stdout and "lancement." in r.stdout and "done.." in r.stdout
        img = np.asarray(Image.open(png).convert("RGBA"))
        assert rgb_report(img, z["rgba"])[0] <= RGB_TOL_FRACTION and img.shape == z["rgba"].shape
    ppm = str(tmp_path / "cube.ppm")
    assert subprocess.run([rt, str(tmp_path / "cube.svati"), ppm], capture_output=True, timeout=300).returncode == 0
    tok = open(ppm).read().split()
    assert tok[:4] == ["P3", "512", "512", "255"]
    body = np.array(tok[4:], dtype=np.int32).reshape(512, 512, 3)
    assert np.array_equal(body, load_golden("cube")[1]["rgba"][..., :3])


def test_set_camera_and_limit(gpu):
    import ctypes as C
    hs, z = load_golden("cube")
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        a = ds.render()
        cam = pyrtb.Camera()
        C.memmove(C.byref(cam), C.byref(hs.scaled(3).camera), C.sizeof(pyrtb.Camera))
        cam.position.x += 0.5
        ds.set_camera(cam)
        b = ds.render()
        assert not np.array_equal(a, b)
        hs2 = pyrtb.HostScene.from_arrays(z["vertex"], z["normal"], z["tri_counts"], z["materials"], z["lights"], z["camera"])
        hs2.s.camera.position.x += 0.5
        want = oracle_py.orc_render(hs2, 3, want_ids=False)["rgba"]
        assert rgb_report(b, want)[0] <= RGB_TOL_FRACTION
        cam.position.x = 1e7      # far outside what the hierarchy was padded for
        with pytest.raises(pyrtb.RtbError) as e:
            ds.set_camera(cam)
        assert e.value.code == pyrtb.RTB_ERR_LIMIT
    finally:
        ds.close()


def test_longest_first_schedule_never_changes_the_frame(gpu):
    """api.cu render_async: frame 2 of a handle records item costs, later frames take the items
    longest first; a camera change re-records.  Bytes must equal the unscheduled frame."""
    hs = pyrtb.HostScene.terrain(96, 200, 120)
    lib = pyrtb.lib()
    try:
        lib.rtb_set_schedule(0)
        ds = pyrtb.DeviceScene(hs, 3)
        want = ds.render().copy()
        want_r = [ds.render(rank=r, nranks=3).copy() for r in range(3)]
        ds.close()
        lib.rtb_set_schedule(1)
        ds = pyrtb.DeviceScene(hs, 3)
        for _ in range(4):
            assert np.array_equal(ds.render(), want)
        for r in range(3):
            for _ in range(3):
                assert np.array_equal(ds.render(rank=r, nranks=3), want_r[r])
        ds.set_camera(hs.scaled(3).camera)      # same camera, new generation: re-record path
        for _ in range(3):
            assert np.array_equal(ds.render(), want)
        ds.close()
    finally:
        lib.rtb_set_schedule(1)


def test_render_multi_equals_single_device(gpu):
    """rtb_render_multi (one host thread per device, tiles stored into device 0's frame over
    peer access) must give the bytes of rtb_render.  Needs >= 2 devices."""
    n = pyrtb.device_count()
    if n < 2:
        pytest.skip("one device only")
    for hs in (pyrtb.HostScene.terrain(160, 640, 360), load_golden("car-on-road")[0]):
        want = pyrtb.render_oneshot(hs, 3)
        for k in sorted({2, min(n, 4), n}):
            got = pyrtb.render_multi(hs, k, 3)
            assert np.array_equal(got, want), "%d devices" % k
