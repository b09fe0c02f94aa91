"""The hierarchy builders (GPU box): sort-key layout chosen by the surface-area cost of sample
trees, clusters rebuilt with binned surface-area splits, the plain radix tree as the A/B
baseline.  Every builder must give the brute-force tracer's answer (same ids, same distance
bits) and the same frame bytes; the builders only differ in how much they cull.
Reference semantics: collide (gpu/hit.cu:84-114), create_octree (gpu/partitioning/octree.cu:362-411)."""
import numpy as np
import pytest

import pyrtb
from util import load_golden, primary_rays, random_scene

pytestmark = pytest.mark.gpu

CUBIC_KEYS, RADIX_ONLY, ONE_FRAME = 2, 4, 8          # include/rtb200.h RTB_BUILD_*


@pytest.fixture(scope="module")
def gpu():
    if pyrtb.device_count() < 1:
        pytest.skip("no CUDA device")


def _check_tree(ds, hs):
    """structure: every triangle in exactly one leaf, child ranges inside the parent's, boxes hold the vertices"""
    nodes, perm, root = ds.dump_tree()
    T = hs.triangle_count
    assert sorted(perm.tolist()) == list(range(T))
    refs = nodes[:, 12:14].copy().view(np.int32)
    lo = nodes[:, 14].copy().view(np.int32)
    cnt = nodes[:, 15].copy().view(np.int32)
    verts = hs.arrays()["vertex"][perm]
    seen = np.zeros(T, np.int32)
    stack = [(root, 0)]
    depth = 0
    while stack:
        k, d = stack.pop()
        depth = max(depth, d + 1)
        covered = 0
        for slot in range(2):
            ref = int(refs[k, slot])
            box = (nodes[k, 4 * slot], nodes[k, 4 * slot + 1], nodes[k, 4 * slot + 2], nodes[k, 4 * slot + 3],
                   nodes[k, 8 + 2 * slot], nodes[k, 9 + 2 * slot])
            if ref >= 0:
                f, c = int(lo[ref]), int(cnt[ref])
                stack.append((ref, d + 1))
            else:
                f, c = (~ref) >> 3, (~ref) & 7
                if c == 0:
                    continue
                assert 1 <= c <= 3
                seen[f:f + c] += 1
            assert int(lo[k]) <= f and f + c <= int(lo[k]) + int(cnt[k])
            covered += c
            v = verts[f:f + c].reshape(-1, 3)
            assert (v[:, 0] >= box[0]).all() and (v[:, 0] <= box[1]).all()
            assert (v[:, 1] >= box[2]).all() and (v[:, 1] <= box[3]).all()
            assert (v[:, 2] >= box[4]).all() and (v[:, 2] <= box[5]).all()
        assert covered == int(cnt[k])
    assert (seen == 1).all()
    assert depth <= ds.stats()["max_depth"], "the reported depth bounds the traversal stack"
    return depth


def _rays(hs, n, seed):
    rng = np.random.default_rng(seed)
    W, H = hs.width * 3, hs.height * 3
    xy = np.stack([rng.integers(0, W, n), rng.integers(0, H, n)], axis=1)
    return primary_rays(hs, 3, xy)


def test_terrain_picks_the_height_axis_last_and_every_builder_agrees(gpu):
    hs = pyrtb.HostScene.terrain(300, 256, 144)          # 180 000 triangles: sample trees are the full trees
    frames, stats = {}, {}
    rays = _rays(hs, 20000, 1)
    want = None
    for name, flags in (("default", 0), ("cubic", CUBIC_KEYS), ("radix", RADIX_ONLY), ("cubic+radix", CUBIC_KEYS | RADIX_ONLY),
                        ("one-frame", ONE_FRAME)):
        ds = pyrtb.DeviceScene(hs, 3, flags=flags)
        try:
            stats[name] = ds.stats()
            _check_tree(ds, hs)
            frames[name] = ds.render()
            ds.render(count=True)
            stats[name]["visits"] = ds.stats()["node_visits"]
            tri, dist = ds.trace_rays(rays, pyrtb.TRACER_TREE)
            if want is None:
                want = ds.trace_rays(rays, pyrtb.TRACER_BRUTE)
            assert np.array_equal(tri, want[0]) and np.array_equal(dist.view(np.uint32), want[1].view(np.uint32)), name
        finally:
            ds.close()
    assert stats["default"]["key_mode"] == 2 and stats["default"]["builder"] == 1
    assert 0 < (stats["default"]["mode_cost"] >> 16) & 0xffff < 950, "the y-last sample tree is at least 5 % cheaper"
    assert stats["cubic"]["key_mode"] == 0 and stats["radix"]["builder"] == 0 and stats["radix"]["key_mode"] == 2
    assert stats["one-frame"]["builder"] == 1, "below 2^18 triangles a one-frame build still rebuilds the clusters"
    for name in frames:
        assert np.array_equal(frames[name], frames["default"]), name
    # what the better trees are for
    assert stats["default"]["visits"] < 0.92 * stats["cubic+radix"]["visits"]


def test_axis_choice_follows_the_surface_not_the_axis_name(gpu):
    """the same height field standing on its side: x is the axis that must come last"""
    hs = pyrtb.HostScene.terrain(150, 64, 64)
    a = hs.arrays()
    rot = lambda p: np.stack([p[..., 1], -p[..., 0], p[..., 2]], axis=-1)      # y -> x
    cam = list(a["camera"])
    cam[2:5] = rot(np.array(cam[2:5]))
    cam[5:8] = rot(np.array(cam[5:8]))
    cam[8:11] = rot(np.array(cam[8:11]))
    hs2 = pyrtb.HostScene.from_arrays(rot(a["vertex"]), rot(a["normal"]), a["tri_counts"], a["materials"], a["lights"], cam)
    ds = pyrtb.DeviceScene(hs2, 1)
    try:
        assert ds.stats()["key_mode"] == 1
        _check_tree(ds, hs2)
    finally:
        ds.close()


@pytest.mark.parametrize("name", ["car-on-road", "spheres", "island_smooth", "susans_smooth"])
def test_cluster_rebuild_keeps_golden_hits(gpu, name):
    """binned surface-area clusters against the plain radix tree on meshes with needles and duplicates"""
    hs, z = load_golden(name)
    got = {}
    for flags in (0, RADIX_ONLY):
        ds = pyrtb.DeviceScene(hs, 3, flags=flags)
        try:
            _check_tree(ds, hs)
            tri, _ = ds.hit_ids(pyrtb.TRACER_TREE)
            assert np.array_equal(tri, z["tri"].astype(np.int32))
            got[flags] = ds.render()
        finally:
            ds.close()
    assert np.array_equal(got[0], got[RADIX_ONLY])


def test_identical_and_degenerate_triangles_fall_back_to_halving(gpu):
    """all centroids equal: no cut has two non-empty sides, the cluster is halved by position"""
    T = 700
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    v = np.repeat(tri[None], T, axis=0)
    v[600:] = 0.25                                        # 100 point-triangles at one spot
    n = np.zeros_like(v); n[..., 2] = -1
    mats = [[0.2, 0.2, 0.2, 0.8, 0.6, 0.4, 0.3, 0.3, 0.3, 20, 1, 0, 1]]
    lights = [[pyrtb.AMBIENT, .3, .3, .3, 0, 0, 0], [pyrtb.DIRECTIONAL, 1, 1, 1, 0.2, -0.3, 1]]
    hs = pyrtb.HostScene.from_arrays(v, n, [T], mats, lights, [64, 48, 0.3, 0.3, -3, 1, 0, 0, 0, -1, 0, 60])
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        depth = _check_tree(ds, hs)
        assert depth <= 12
        tri_t, _ = ds.hit_ids(pyrtb.TRACER_TREE)
        tri_b, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
        assert np.array_equal(tri_t, tri_b) and (tri_t >= 0).any()
        assert set(np.unique(tri_t[tri_t >= 0]).tolist()) == {0}, "ties go to the lowest triangle id"
    finally:
        ds.close()


@pytest.mark.parametrize("T", [4, 5, 33, 1024, 1025, 3000])
def test_soups_around_the_cluster_limit(gpu, T):
    hs = random_scene(100 + T, T, 160, 100)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        _check_tree(ds, hs)
        rays = _rays(hs, 6000, T)
        tri, dist = ds.trace_rays(rays, pyrtb.TRACER_TREE)
        bt, bd = ds.trace_rays(rays, pyrtb.TRACER_BRUTE)
        # needle soups: the documented ill-conditioned exceptions (tests/test_gpu_random_scenes.py), nothing else
        assert int((tri != bt).sum()) <= 1
        same = tri == bt
        assert np.array_equal(dist[same].view(np.uint32), bd[same].view(np.uint32))
        a = pyrtb.DeviceScene(hs, 3)
        try:
            na, pa, ra = a.dump_tree()
            nb, pb, rb = ds.dump_tree()
            assert ra == rb and np.array_equal(pa, pb) and np.array_equal(na.view(np.uint32), nb.view(np.uint32)), "deterministic"
        finally:
            a.close()
    finally:
        ds.close()
