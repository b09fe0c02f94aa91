"""Edge cases of the frame path against the oracle (GPU box): empty and light-less scenes,
objects without triangles, every kind of degenerate triangle, mirrors facing each other
(the 11-iteration bounce cap), aliasing factors, padded row stride."""
import ctypes as C

import numpy as np
import pytest

import oracle_py
import pyrtb
from util import rgb_report

pytestmark = pytest.mark.gpu
CAM = [96, 64, 0, 0, -4, 1, 0, 0, 0, -1, 0, 90]       # width height pos u v fov


@pytest.fixture(scope="module")
def gpu():
    if pyrtb.device_count() < 1:
        pytest.skip("no CUDA device")


def _mat(kd=(0.8, 0.6, 0.4), nr=0.0, ks=(0.3, 0.3, 0.3), ns=20.0):
    return [0.2, 0.2, 0.2, *kd, *ks, ns, 1.0, nr, 1.0]


def _quad(z, half=2.0, flip=False):
    a, b, c, d = [-half, -half, z], [half, -half, z], [half, half, z], [-half, half, z]
    t = [[a, b, c], [a, c, d]]
    return [tri[::-1] for tri in t] if flip else t


def _scene(tris_per_object, materials, lights, camera=CAM, normals=None):
    """normals: one vertex normal per object (default: facing the camera)"""
    v = np.array([t for ob in tris_per_object for t in ob], np.float32).reshape(-1, 3, 3)
    normals = normals or [(0, 0, -1)] * len(tris_per_object)
    n = np.array([[nm] * 3 for ob, nm in zip(tris_per_object, normals) for _ in ob], np.float32).reshape(-1, 3, 3)
    return pyrtb.HostScene.from_arrays(v, n, [len(ob) for ob in tris_per_object], materials, lights, camera)


def _wall(x, z0=0.0, z1=60.0, half=8.0):
    a, b, c, d = [x, -half, z0], [x, -half, z1], [x, half, z1], [x, half, z0]
    return [[a, b, c], [a, c, d]]


def _check(hs, aliasing=3):
    want = oracle_py.orc_render(hs, aliasing)
    ds = pyrtb.DeviceScene(hs, aliasing)
    try:
        got = ds.render()
        tri, obj = ds.hit_ids(pyrtb.TRACER_TREE)
        brute, _ = ds.hit_ids(pyrtb.TRACER_BRUTE)
    finally:
        ds.close()
    assert np.array_equal(tri, want["tri"]) and np.array_equal(obj, want["obj"]) and np.array_equal(brute, want["tri"])
    off, same, worst = rgb_report(got, want["rgba"])
    assert off == 0.0 and worst <= 1, (off, same, worst)
    assert np.array_equal(pyrtb.render_oneshot(hs, aliasing), got)
    return got, want


LIGHTS = [[pyrtb.AMBIENT, 0.3, 0.3, 0.3, 0, 0, 0], [pyrtb.DIRECTIONAL, 1, 1, 1, 0.3, -0.5, 1],
          [pyrtb.POINT, 1, 0.9, 0.8, 1, 2, -3]]


def test_empty_scene_is_black(gpu):
    hs = _scene([[]], [_mat()], LIGHTS)
    got, _ = _check(hs)
    assert got[..., :3].max() == 0 and (got[..., 3] == 255).all()


def test_scene_without_lights(gpu):
    got, _ = _check(_scene([_quad(2.0)], [_mat()], np.zeros((0, 7), np.float32)))
    assert got[..., :3].max() == 0


def test_objects_without_triangles_between_others(gpu):
    hs = _scene([[], _quad(3.0), [], _quad(2.0, half=0.7), []], [_mat(), _mat((1, 0, 0)), _mat(), _mat((0, 1, 0)), _mat()],
                LIGHTS)
    _check(hs)


def test_degenerate_and_duplicate_triangles(gpu):
    p = [0.5, 0.5, 1.5]
    tris = _quad(2.5) + [[p, p, p],                                   # a point
                         [[-1, 0, 1.5], [0, 0, 1.5], [1, 0, 1.5]],    # collinear
                         [[-1, -1, 2], [1, -1, 2], [-1, -1, 2]],      # two equal vertices
                         ] + _quad(2.5) + _quad(2.5, half=1.0)        # exact duplicates, coplanar overlap
    _check(_scene([tris], [_mat()], LIGHTS))


def test_facing_mirrors_hit_the_bounce_cap(gpu):
    # two Nr 1.0 walls facing each other along the view direction: oblique rays zig-zag between
    # them until the 11-iteration cap of gpu/raytracer.cu:116-122 ends the chain
    hs = _scene([_wall(-1.5, z0=-3.5, half=50.0), _wall(1.5, z0=-3.5, half=50.0), _quad(59.0, half=70.0)],
                [_mat(nr=1.0), _mat((0.2, 0.4, 0.9), nr=1.0), _mat((1, 1, 0))], LIGHTS,
                normals=[(1, 0, 0), (-1, 0, 0), (0, 0, -1)])
    want = oracle_py.orc_render(hs, 3)
    samples = 96 * 64 * 9
    assert int(want["counts"][2]) > 3 * samples, "bounce chains really happen"
    assert int(want["counts"][2]) <= 10 * samples, "and the cap holds"
    _check(hs)


@pytest.mark.parametrize("aliasing", [1, 2, 4, 5])
def test_aliasing_factors(gpu, aliasing):
    hs = _scene([_quad(3.0), _quad(2.0, half=0.6)], [_mat(), _mat((0.1, 0.9, 0.2), nr=0.3)], LIGHTS, camera=[70, 45, 0.3, 0.2, -4, 1, 0, 0, 0, -1, 0, 75])
    _check(hs, aliasing)


def test_padded_row_stride(gpu):
    hs = _scene([_quad(3.0)], [_mat()], LIGHTS)
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        tight = ds.render()
        padded = np.full((64, 96 + 13, 4), 7, np.uint8)
        ds.render(out=padded[:, :96])            # a view: stride = (96 + 13) * 4 bytes
        assert np.array_equal(padded[:, :96], tight) and (padded[:, 96:] == 7).all()
    finally:
        ds.close()


def test_unit_box_known_answer_of_the_reference(gpu):
    """gpu/test.cu:362-375: the ray (0,0,0) -> (0,0,1) along an edge of the unit box [0,1]^3 "hits" the box
    (hit_aabb, gpu/partitioning/aabb.cu:203-243): origin ON the boundary, two direction components exactly 0
    (1/d = inf, 0 * inf = NaN in a naive slab test).  Replayed on the product's slab test through the tree:
    the unit cube's 12 faces as the scene, axis-parallel rays on its faces and edges -- a box test that
    culled any of them would lose a hit the exhaustive tracer finds (ids and distance bits must agree)."""
    c = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [1, 1, 1], [0, 1, 1]], np.float32)
    quads = [(0, 1, 2, 3), (4, 5, 6, 7), (0, 1, 5, 4), (3, 2, 6, 7), (0, 3, 7, 4), (1, 2, 6, 5)]
    tris = []
    for a, b, cc, d in quads:
        tris += [[c[a], c[b], c[cc]], [c[a], c[cc], c[d]]]
    v = np.array(tris, np.float32)
    n = np.ones_like(v)
    mats = [[0.2, 0.2, 0.2, 0.7, 0.7, 0.7, 0.3, 0.3, 0.3, 20, 1, 0, 1]]
    lights = [[pyrtb.AMBIENT, .3, .3, .3, 0, 0, 0]]
    cam = [64, 48, 0.5, 0.5, -4, 1, 0, 0, 0, -1, 0, 60]
    hs = pyrtb.HostScene.from_arrays(v, n, [len(tris)], mats, lights, cam)
    ds = pyrtb.DeviceScene(hs, 1)
    try:
        assert ds.stats()["nodes"] > 1, "the scene must have a hierarchy to cull with"
        rays = [[0, 0, 0, 0, 0, 1]]                                     # the reference's own vector
        for o in ([0, 0, -1], [1, 1, -1], [0, 1, -2], [0.5, 0, -1], [0, 0.5, -1], [1, 0.25, -3], [0.5, 0.5, -1]):
            rays.append([*o, 0, 0, 1])                                  # along z, on edges / faces / through the middle
        for o in ([-1, 0, 0], [-1, 1, 1], [-2, 0.5, 0], [-1, 0.5, 0.5]):
            rays.append([*o, 1, 0, 0])                                  # along x
        for o in ([0, 2, 0], [1, 3, 1], [0.5, 2, 0.5]):
            rays.append([*o, 0, -1, 0])                                 # along -y
        rays = np.array(rays, np.float32)
        tb, db = ds.trace_rays(rays, pyrtb.TRACER_BRUTE)
        tt, dt = ds.trace_rays(rays, pyrtb.TRACER_TREE)
        assert np.array_equal(tt, tb) and np.array_equal(dt.view(np.uint32), db.view(np.uint32))
        assert (tb >= 0).sum() >= 5, tb
        at, _ = ds.trace_rays(rays, pyrtb.TRACER_TREE, any_hit=True)
        ab, _ = ds.trace_rays(rays, pyrtb.TRACER_BRUTE, any_hit=True)
        assert np.array_equal(at >= 0, ab >= 0)
    finally:
        ds.close()


def test_frames_on_a_callers_stream_are_ordered_with_the_handles_own(gpu):
    """rtb_scene_render_device with a cudaStream_t of the caller: rtb_scene_sync() covers the frame, the frame
    time and the projection pass's state words are the frame's, and frames that alternate between the
    caller's stream and the handle's reuse the same scratch without stepping on each other"""
    import torch
    torch.cuda.set_device(0)
    hs = pyrtb.HostScene.terrain(132, 320, 180)            # >= 2^15 triangles: projection pass + light grids
    ds = pyrtb.DeviceScene(hs, 3)
    try:
        want = ds.render()
        side = torch.cuda.Stream()
        frame = torch.zeros((180, 320, 4), dtype=torch.uint8, device="cuda:0")
        torch.cuda.synchronize()
        for k in range(6):
            frame.zero_()
            torch.cuda.synchronize()
            ds.render_device(frame.data_ptr(), 320 * 4, stream=side.cuda_stream if k % 2 == 0 else None)
            ds.sync()                                       # the handle's sync alone, whichever stream ran the frame
            st = ds.raster_state()
            assert not st["gave_up"] and st["items"] > 0 and ds.stats()["frame_ms"] > 0
            assert np.array_equal(frame.cpu().numpy(), want), k
    finally:
        ds.close()
