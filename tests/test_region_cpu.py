"""The cull of the projection passes (csrc/region.cuh) against the rounded test it must cover (CPU).

raster.cuh and lightgrid.cuh only run the reference's ray/triangle test (gpu/hit.cu:8-48) on the
samples / rays inside a per-triangle region; the parity contract (same hit as the exhaustive
collide(), gpu/hit.cu:84-114) holds iff that region contains every ray the ROUNDED test accepts.
region.cuh is plain C++: this file compiles it for the host (tests/native/region_host.cpp), replays
primary_ray / the shadow-ray set-up and the test of tri_candidate in numpy float32, operation by
operation, and requires every accepted ray to lie inside the rectangle and the edge record --
including triangles seen edge-on, triangles across the camera plane or behind the camera, needles
and degenerate triangles, i.e. the cases the first version of the cull (boxes + padding) got wrong
(profiles/r1_pad_study_v2.txt)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from test_accept_region_bound import _cross, _dot, _rounded_test

f32 = np.float32
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EPS24 = 2.0 ** -24


@pytest.fixture(scope="module")
def rg():
    src = os.path.join(HERE, "native", "region_host.cpp")
    out_dir = os.path.join(HERE, "native", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libregion_host.so")
    hdr = os.path.join(ROOT, "raytracing-gpu_b200", "csrc", "region.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-I" + os.path.dirname(hdr),
                        "-o", so, src], check=True)
    L = C.CDLL(so)
    vp = C.c_void_p
    L.rg_primary.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, vp, vp]
    L.rg_primary_lines.argtypes = [vp, vp, C.c_int, vp, vp]
    L.rg_dir.argtypes = [vp, vp, vp, C.c_int, vp, vp]
    L.rg_point.argtypes = [vp, vp, vp, C.c_int, vp, vp]
    L.rg_primary_rows.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, vp, C.c_int]
    L.rg_primary_rows.restype = C.c_int
    L.rg_span.argtypes = [vp, vp, vp, vp, C.c_int, vp]
    L.rg_band.argtypes = [vp, vp, vp, vp, C.c_int, vp, vp]
    return L


def _unit(v):
    return v / np.linalg.norm(v)


def make_camera(rng, W, H, fov, pos_scale=300.0):
    """float32 camera as api.cu camera_frame builds it, and the double constants of CamFrame"""
    w = _unit(rng.standard_normal(3))
    u = _unit(np.cross(w, rng.standard_normal(3)))
    v = np.cross(w, u)
    u32, v32 = u.astype(f32), v.astype(f32)
    w32 = np.array([u32[1] * v32[2] - u32[2] * v32[1], u32[2] * v32[0] - u32[0] * v32[2], u32[0] * v32[1] - u32[1] * v32[0]], f32)
    pos = (rng.uniform(-1, 1, 3) * pos_scale).astype(f32)
    L = f32(W / (2 * np.tan(fov * np.pi / 360)))
    Cc = (pos + L * w32).astype(f32)
    return dict(u=u32, v=v32, w=w32, pos=pos, C=Cc, L=L, W=W, H=H)


def cam_consts(cam):
    """the host side of the cull (api.cu cam_frame_consts), float64"""
    u, v, pos, Cc = (cam[k].astype(np.float64) for k in ("u", "v", "pos", "C"))
    W, H = cam["W"], cam["H"]
    D0 = pos - Cc
    X = max(W // 2, W - 1 - W // 2)
    Y = max(H // 2, H - 1 - H // 2)
    M = np.abs(Cc) + np.abs(u) * X + np.abs(v) * Y
    Mo = float(np.linalg.norm(M))
    dmax = 0.0
    for x in (-(W // 2), W - 1 - W // 2):
        for y in (-(H // 2), H - 1 - H // 2):
            dmax = max(dmax, float(np.linalg.norm(D0 - u * x - v * y)))
    Dmax = dmax * (1 + 1e-6) + 8 * EPS24 * Mo
    return np.concatenate([pos, D0, u, v, [Dmax, Mo]]).astype(np.float64)


def primary_rays(cam, px, py):
    """primary_ray of csrc/trace.cu (gpu/raytracer.cu:97-105), float32 operation by operation"""
    fx = (px - cam["W"] // 2).astype(f32)[:, None]
    fy = (py - cam["H"] // 2).astype(f32)[:, None]
    ui = cam["u"][None, :] * fx
    vj = cam["v"][None, :] * fy
    o = (cam["C"][None, :] + ui) + vj
    dd = cam["pos"][None, :] - o
    ln = np.sqrt((dd[:, 0] * dd[:, 0] + dd[:, 1] * dd[:, 1]) + dd[:, 2] * dd[:, 2])
    return o, (dd / ln[:, None]).astype(f32)


def record_inside(rec, dx, dy):
    """region::record_inside in float32 (rec: (n, 12), dx, dy: (n,))"""
    dx = dx.astype(f32); dy = dy.astype(f32)
    f0 = (rec[:, 0] * dx + rec[:, 1] * dy) + rec[:, 2]
    f1 = (rec[:, 3] * dx + rec[:, 4] * dy) + rec[:, 5]
    f2 = (rec[:, 6] * dx + rec[:, 7] * dy) + rec[:, 8]
    flag = rec[:, 11]
    in1 = (f0 <= 0) & (f1 <= 0) & (f2 <= 0)
    in2 = (flag < 0) & (f0 >= flag) & (f1 >= flag) & (f2 >= flag)
    return in1 | in2 | (flag == 0)


def cull(rg, cam, tri):
    n = tri.shape[0]
    tri = np.ascontiguousarray(tri, f32)
    cc = cam_consts(cam)
    out = np.zeros((n, 5), np.int32)
    rec = np.zeros((n, 12), f32)
    rg.rg_primary(cc.ctypes.data, cam["W"], cam["H"], tri.ctypes.data, n, out.ctypes.data, rec.ctypes.data)
    return out, rec


def lines(rg, cam, tri):
    n = tri.shape[0]
    tri = np.ascontiguousarray(tri, f32)
    cc = cam_consts(cam)
    ln = np.zeros((n, 12), np.float64)
    ok = np.zeros(n, np.int32)
    rg.rg_primary_lines(cc.ctypes.data, tri.ctypes.data, n, ln.ctypes.data, ok.ctypes.data)
    return ln, ok


def check_camera(rg, cam, tri, want_accepts=0):
    """every sample of the frame against every triangle: accepted => inside rectangle and record.
    Returns (accepted pairs, pairs inside the cull, worst use of the allowance)."""
    W, H = cam["W"], cam["H"]
    n = tri.shape[0]
    out, rec = cull(rg, cam, tri)
    ln, ok = lines(rg, cam, tri)
    py, px = np.mgrid[0:H, 0:W]
    px = px.ravel(); py = py.ravel()
    o, d = primary_rays(cam, px, py)
    S = px.size
    accepted = 0
    inside_total = 0
    worst = 0.0
    x = (px - W // 2).astype(np.float64); y = (py - H // 2).astype(np.float64)
    for t in range(n):
        v0 = np.broadcast_to(tri[t, 0:3], (S, 3)).astype(f32)
        e1 = np.broadcast_to(tri[t, 3:6], (S, 3)).astype(f32)
        e2 = np.broadcast_to(tri[t, 6:9], (S, 3)).astype(f32)
        acc = _rounded_test(o, d, v0, e1, e2)
        kind, x0, y0, x1, y1 = (int(v) for v in out[t])
        if kind == 0:
            assert not acc.any(), "triangle %d culled but %d samples accept it" % (t, int(acc.sum()))
            continue
        in_rect = (px >= x0) & (px <= x1) & (py >= y0) & (py <= y1)
        r = np.broadcast_to(rec[t], (S, 12))
        assert rec[t, 9] == x0 and rec[t, 10] == y0
        ins = in_rect & record_inside(r, px - x0, py - y0)
        bad = acc & ~ins
        assert not bad.any(), "triangle %d: %d accepted samples outside the cull, e.g. %r (kind %d rect %r)" % (
            t, int(bad.sum()), (int(px[bad][0]), int(py[bad][0])), kind, (x0, y0, x1, y1))
        accepted += int(acc.sum())
        inside_total += int(ins.sum())
        if acc.any():
            # how much of the allowance the accepted rays use: S+ needs F_k >= -E_k, S- F_k <= E_k
            F = ln[t, 6:9][None, :] + x[acc, None] * ln[t, 0:3][None, :] + y[acc, None] * ln[t, 3:6][None, :]
            E = ln[t, 9:12][None, :]
            use = np.minimum((-F / E).max(axis=1), (F / E).max(axis=1))       # the better of the two sides
            worst = max(worst, float(use.max()))
    assert accepted >= want_accepts, accepted
    return accepted, inside_total, worst


def tris_on_screen(rng, cam, n, size_samples=(0.5, 20.0), depth=(30.0, 3000.0)):
    W, H = cam["W"], cam["H"]
    sx = rng.integers(0, W, n); sy = rng.integers(0, H, n)
    o, d = primary_rays(cam, sx, sy)
    dep = np.exp(rng.uniform(np.log(depth[0]), np.log(depth[1]), n))
    centre = cam["pos"].astype(np.float64)[None, :] + d.astype(np.float64) * dep[:, None]
    size = rng.uniform(*size_samples, n) * dep / float(cam["L"])
    e1 = rng.standard_normal((n, 3)); e1 *= (size / np.linalg.norm(e1, axis=1))[:, None]
    e2 = rng.standard_normal((n, 3)); e2 *= (size * rng.uniform(0.5, 1.5, n) / np.linalg.norm(e2, axis=1))[:, None]
    v0 = centre - (e1 + e2) / 3
    return np.concatenate([v0, e1, e2], axis=1).astype(f32)


def tris_grazing(rng, cam, n, depth=(30.0, 3000.0)):
    """triangles whose plane passes within a hair of the camera position: seen edge-on"""
    t = tris_on_screen(rng, cam, n, (2.0, 30.0), depth).astype(np.float64)
    pos = cam["pos"].astype(np.float64)
    v0, e1, e2 = t[:, 0:3], t[:, 3:6], t[:, 6:9]
    centre = v0 + (e1 + e2) / 3
    view = centre - pos
    dist = np.linalg.norm(view, axis=1)
    view /= dist[:, None]
    # e1 stays; e2 is rebuilt in the plane spanned by e1 and (view tilted by theta towards the old normal)
    nrm = _cross(e1, e2); nrm /= np.linalg.norm(nrm, axis=1)[:, None]
    theta = np.exp(rng.uniform(np.log(1e-9), np.log(3e-2), n)) * rng.choice([-1, 1], n)
    inplane = view * np.cos(theta)[:, None] + nrm * np.sin(theta)[:, None]
    e2n = inplane * np.linalg.norm(e2, axis=1)[:, None] * rng.uniform(0.5, 3.0, n)[:, None] + 0.3 * e1
    v0n = centre - (e1 + e2n) / 3
    return np.concatenate([v0n, e1, e2n], axis=1).astype(f32)


def tris_around_camera(rng, cam, n):
    """big triangles anywhere around the camera: across the camera plane, behind it, between the
    image plane and the camera"""
    pos = cam["pos"].astype(np.float64)
    L = float(cam["L"])
    scale = np.exp(rng.uniform(np.log(1.0), np.log(2 * L), n))
    c = pos[None, :] + rng.standard_normal((n, 3)) * scale[:, None]
    e1 = rng.standard_normal((n, 3)) * scale[:, None] * rng.uniform(0.05, 2.0, n)[:, None]
    e2 = rng.standard_normal((n, 3)) * scale[:, None] * rng.uniform(0.05, 2.0, n)[:, None]
    return np.concatenate([c - (e1 + e2) / 3, e1, e2], axis=1).astype(f32)


def tris_needles(rng, cam, n):
    t = tris_on_screen(rng, cam, n, (5.0, 60.0)).astype(np.float64)
    e1, e2 = t[:, 3:6], t[:, 6:9]
    s = rng.random(n)[:, None]
    width = np.exp(rng.uniform(np.log(1e-7), np.log(1e-2), n))[:, None]
    e2n = e1 * s + e2 * width                          # third vertex almost on the first edge
    e2n[: n // 4] = e1[: n // 4] * s[: n // 4]         # a quarter exactly collinear (up to float rounding)
    e2n[: n // 8] = e1[: n // 8] * 0.5                 # some exactly degenerate
    t[:, 6:9] = e2n
    return t.astype(f32)


CAMERAS = [
    # W, H, fov: narrow fields of view put the image plane thousands of units behind the camera, as
    # at 3840x2160x3 (L = 8 226), with few enough samples to test every one of them
    (96, 64, 60.0), (128, 80, 0.9), (160, 100, 0.5), (113, 67, 5.0), (64, 64, 120.0),
]


@pytest.mark.parametrize("W,H,fov", CAMERAS)
def test_primary_cull_contains_every_accepted_sample(rg, W, H, fov):
    rng = np.random.default_rng(hash((W, H)) & 0xffff)
    worst = 0.0
    acc_total = in_total = 0
    for rep in range(2):
        cam = make_camera(rng, W, H, fov)
        for gen, n, want in ((tris_on_screen, 60, 500), (tris_grazing, 60, 1), (tris_around_camera, 40, 0), (tris_needles, 40, 0)):
            tri = gen(rng, cam, n)
            a, i, w = check_camera(rg, cam, tri, want)
            worst = max(worst, w)
            if gen is tris_on_screen:
                acc_total += a; in_total += i
    # the allowance has room: SAFETY, and the first-order worst case adds what random roundings
    # mostly cancel (the worst accepted ray of these runs uses ~0.12 of it)
    assert worst <= 0.8, worst
    assert acc_total > 1000


def test_the_cull_is_not_vacuous_on_the_benchmark_geometry(rg):
    """the C5 frame: image plane 8 000 units behind the camera, triangles 300-1500 units in front of it,
    a few samples across -- the region holds what is accepted plus a rim of a fraction of a sample"""
    rng = np.random.default_rng(3)
    cam = make_camera(rng, 128, 80, 0.9)
    tri = tris_on_screen(rng, cam, 150, size_samples=(2.0, 8.0), depth=(300.0, 1500.0))
    acc, ins, worst = check_camera(rg, cam, tri, 500)
    assert ins <= 2.5 * acc, (acc, ins)        # triangles of ~6 samples: the rim is as large as the triangle
    out, _ = cull(rg, cam, tri)
    area = ((out[:, 3] - out[:, 1] + 1) * (out[:, 4] - out[:, 2] + 1))[out[:, 0] > 0].sum()
    assert area <= 8.0 * acc, (acc, int(area))


def test_far_camera_positions(rg):
    """a camera far from the origin: the rounding of the origin scales with |C|, not with L"""
    rng = np.random.default_rng(99)
    cam = make_camera(rng, 96, 64, 40.0, pos_scale=20000.0)
    for gen in (tris_on_screen, tris_grazing):
        check_camera(rg, cam, gen(rng, cam, 50, depth=(20.0, 500.0)))


def test_tile_rows_cover_the_region(rg):
    """primary_row_tiles: every accepted sample's tile is listed for its tile row"""
    rng = np.random.default_rng(5)
    cam = make_camera(rng, 160, 100, 0.7)
    W, H = cam["W"], cam["H"]
    cc = cam_consts(cam)
    py, px = np.mgrid[0:H, 0:W]
    px = px.ravel(); py = py.ravel()
    o, d = primary_rays(cam, px, py)
    S = px.size
    tile = 8
    hits = 0
    for gen in (tris_on_screen, tris_grazing, tris_around_camera):
        tri = gen(rng, cam, 30)
        for t in range(tri.shape[0]):
            one = np.ascontiguousarray(tri[t], f32)
            rows = np.zeros((64, 3), np.int32)
            n = rg.rg_primary_rows(cc.ctypes.data, W, H, one.ctypes.data, tile, rows.ctypes.data, 64)
            allowed = np.zeros((H // tile + 1, W // tile + 1), bool)
            for r in rows[:n]:
                allowed[r[0], r[1]:r[2] + 1] = True
            acc = _rounded_test(o, d, np.broadcast_to(one[0:3], (S, 3)).astype(f32), np.broadcast_to(one[3:6], (S, 3)).astype(f32),
                                np.broadcast_to(one[6:9], (S, 3)).astype(f32))
            assert allowed[py[acc] // tile, px[acc] // tile].all(), t
            hits += int(acc.sum())
    assert hits > 1000


def test_row_spans_hold_every_sample_the_record_passes(rg):
    """record_span (k_rast_items walks big items row by row) is a superset of record_inside"""
    rng = np.random.default_rng(17)
    checked = passing = 0
    for W, H, fov in ((160, 100, 0.7), (96, 64, 60.0), (128, 80, 3.0)):
        cam = make_camera(rng, W, H, fov)
        for gen in (tris_on_screen, tris_grazing, tris_around_camera, tris_needles):
            tri = gen(rng, cam, 60)
            out, rec = cull(rg, cam, tri)
            for t in np.where(out[:, 0] > 0)[0]:
                kind, x0, y0, x1, y1 = (int(v) for v in out[t])
                # an item: a random sub-rectangle of at most 32 x 32 samples
                ix0 = int(rng.integers(x0, x1 + 1)); iy0 = int(rng.integers(y0, y1 + 1))
                w = min(32, x1 - ix0 + 1); h = min(32, y1 - iy0 + 1)
                rows = np.arange(h)
                r = np.broadcast_to(rec[t], (h, 12)).copy()
                dx0 = np.full(h, ix0 - x0, f32); dy = (iy0 - y0 + rows).astype(f32)
                ab = np.zeros((h, 2), np.int32)
                ws = np.full(h, w, np.int32)
                rg.rg_span(r.ctypes.data, dx0.ctypes.data, ws.ctypes.data, dy.ctypes.data, h, ab.ctypes.data)
                for j in range(h):
                    xs = np.arange(w)
                    ins = record_inside(np.broadcast_to(rec[t], (w, 12)), (xs + (ix0 - x0)).astype(f32), np.full(w, dy[j], f32))
                    inspan = (xs >= ab[j, 0]) & (xs <= ab[j, 1])
                    assert not (ins & ~inspan).any(), (t, j, ab[j], np.where(ins)[0])
                    checked += w; passing += int(ins.sum())
    assert checked > 100000 and passing > 10000, (checked, passing)


def test_bands_hold_every_point_the_record_passes(rg):
    """record_band (the light grids list a region's cells with it) is a superset of record_inside"""
    rng = np.random.default_rng(23)
    inside_total = 0
    for W, H, fov in ((160, 100, 0.7), (96, 64, 60.0)):
        cam = make_camera(rng, W, H, fov)
        for gen in (tris_on_screen, tris_grazing, tris_around_camera, tris_needles):
            tri = gen(rng, cam, 50)
            out, rec = cull(rg, cam, tri)
            for t in np.where(out[:, 0] > 0)[0]:
                kind, x0, y0, x1, y1 = (int(v) for v in out[t])
                nb = 12
                edges = np.sort(rng.uniform(-3.0, (y1 - y0) + 3.0, nb + 1)).astype(f32)
                dy0, dy1 = edges[:-1].copy(), edges[1:].copy()
                xext = np.full(nb, float(x1 - x0) + 8.0, f32)
                r = np.broadcast_to(rec[t], (nb, 12)).copy()
                ok = np.zeros(nb, np.int32); lohi = np.zeros((nb, 2), f32)
                rg.rg_band(r.ctypes.data, dy0.ctypes.data, dy1.ctypes.data, xext.ctypes.data, nb, ok.ctypes.data, lohi.ctypes.data)
                # dense points of every band
                for j in range(nb):
                    xs = rng.uniform(-4.0, (x1 - x0) + 4.0, 400).astype(f32)
                    ys = rng.uniform(dy0[j], dy1[j], 400).astype(f32)
                    ys = np.clip(ys, dy0[j], dy1[j])
                    ins = record_inside(np.broadcast_to(rec[t], (400, 12)), xs, ys)
                    if not ok[j]:
                        assert not ins.any(), (t, j)
                    else:
                        assert ((xs[ins] >= lohi[j, 0]) & (xs[ins] <= lohi[j, 1])).all(), (t, j, lohi[j])
                    inside_total += int(ins.sum())
    assert inside_total > 20000, inside_total


# ---- shadow rays ----
def _frame_from_axis(az32):
    """the float frame build_light_grids makes around a light's axis"""
    az = az32.astype(np.float64)
    k = int(np.argmin(np.abs(az)))
    h = np.zeros(3); h[k] = 1
    ax = _unit(np.cross(h, az)).astype(f32)
    ay = np.cross(az32.astype(np.float64), ax.astype(np.float64))
    return ax, _unit(ay).astype(f32)


def _shadow_test(P, d, tri):
    S = P.shape[0]
    return _rounded_test(P, d, np.broadcast_to(tri[0:3], (S, 3)).astype(f32), np.broadcast_to(tri[3:6], (S, 3)).astype(f32),
                         np.broadcast_to(tri[6:9], (S, 3)).astype(f32))


def _scene_tris(rng, n, extent, light_dir=None, org=None):
    c = rng.uniform(-extent, extent, (n, 3))
    size = np.exp(rng.uniform(np.log(extent * 1e-3), np.log(extent * 0.3), n))
    e1 = rng.standard_normal((n, 3)) * size[:, None]
    e2 = rng.standard_normal((n, 3)) * size[:, None]
    # a third of them edge-on to the light within 1e-9 .. 1e-2 rad
    k = n // 3
    for i in range(k):
        ld = light_dir if light_dir is not None else _unit(c[i] - org)
        nrm = _unit(np.cross(e1[i], np.cross(ld, e1[i])))              # normal of the plane through e1 and ld
        theta = np.exp(rng.uniform(np.log(1e-9), np.log(1e-2)))
        e2[i] = (ld * np.cos(theta) + nrm * np.sin(theta)) * np.linalg.norm(e2[i]) + 0.3 * e1[i]
    return np.concatenate([c - (e1 + e2) / 3, e1, e2], axis=1).astype(f32)


def test_directional_light_cull(rg):
    rng = np.random.default_rng(11)
    accepted = 0
    for rep in range(3):
        extent = [1.0, 50.0, 800.0][rep]
        lv = (rng.standard_normal(3) * rng.uniform(0.2, 3)).astype(f32)
        d32 = (f32(-1) * lv).astype(f32)
        az32 = (lv.astype(np.float64) / np.linalg.norm(lv.astype(np.float64))).astype(f32)
        ax, ay = _frame_from_axis(az32)
        tri = _scene_tris(rng, 90, extent, light_dir=_unit(lv.astype(np.float64)))
        Pmax = np.sqrt(3) * extent * 1.5
        smax = 2 * Pmax
        axd, ayd = ax.astype(np.float64), ay.astype(np.float64)
        fr = np.concatenate([d32.astype(np.float64), axd, ayd, [axd @ axd, axd @ ayd, ayd @ ayd, smax, Pmax]])
        dom = np.array([-1e9, 1e9, -1e9, 1e9])
        out = np.zeros((tri.shape[0], 5)); rec = np.zeros((tri.shape[0], 12), f32)
        rg.rg_dir(fr.ctypes.data, dom.ctypes.data, tri.ctypes.data, tri.shape[0], out.ctypes.data, rec.ctypes.data)
        for t in range(tri.shape[0]):
            tv = tri[t].astype(np.float64)
            # origins: around the triangle shifted along the light, dense near its outline
            S = 40000
            uu = rng.uniform(-0.6, 1.6, S); vv = rng.uniform(-0.6, 1.6, S)
            edge = rng.integers(0, 4, S)
            uu = np.where(edge == 1, rng.normal(0, 1e-4, S), uu); vv = np.where(edge == 2, rng.normal(0, 1e-4, S), vv)
            vv = np.where(edge == 3, 1 - uu + rng.normal(0, 1e-4, S), vv)
            X = tv[0:3] + uu[:, None] * tv[3:6] + vv[:, None] * tv[6:9]
            back = rng.uniform(0.01, 1.0, S) * extent
            P = (X + lv.astype(np.float64)[None, :] / np.linalg.norm(lv) * back[:, None]).astype(f32)
            keep = np.abs(P.astype(np.float64)).max(axis=1) <= extent * 1.5
            P = P[keep]
            d = np.broadcast_to(d32, P.shape).astype(f32)
            acc = _shadow_test(P, d, tri[t])
            x = ((P[:, 0] * ax[0] + P[:, 1] * ax[1]) + P[:, 2] * ax[2]).astype(f32)     # lg_project, float32
            y = ((P[:, 0] * ay[0] + P[:, 1] * ay[1]) + P[:, 2] * ay[2]).astype(f32)
            if out[t, 0] == 0:
                assert not acc.any(), t
                continue
            inb = (x >= out[t, 1]) & (x <= out[t, 2]) & (y >= out[t, 3]) & (y <= out[t, 4])
            r = np.broadcast_to(rec[t], (P.shape[0], 12))
            ins = inb & record_inside(r, x - rec[t, 9], y - rec[t, 10])
            assert not (acc & ~ins).any(), (rep, t, int((acc & ~ins).sum()))
            accepted += int(acc.sum())
    assert accepted > 100000, accepted


def test_point_light_cull(rg):
    rng = np.random.default_rng(12)
    accepted = 0
    for rep in range(3):
        extent = [1.0, 50.0, 800.0][rep]
        org = (_unit(rng.standard_normal(3)) * extent * rng.uniform(0.2, 2.5)).astype(f32)
        az32 = _unit(-org.astype(np.float64)).astype(f32)
        ax, ay = _frame_from_axis(az32)
        M = np.stack([ax, ay, az32]).astype(np.float64)
        Minv = np.linalg.inv(M)                       # columns: dual basis
        tri = _scene_tris(rng, 90, extent, org=org.astype(np.float64))
        Pmax = np.sqrt(3) * extent * 1.5
        lim = 4.0
        ratio = np.sqrt(1 + 2 * (lim + 0.01) ** 2) * 1.001
        fr = np.concatenate([org.astype(np.float64), Minv[:, 0], Minv[:, 1], Minv[:, 2], [2 * Pmax, np.linalg.norm(org) + Pmax, ratio]])
        dom = np.array([-lim, lim, -lim, lim], np.float64)
        out = np.zeros((tri.shape[0], 5)); rec = np.zeros((tri.shape[0], 12), f32)
        rg.rg_point(fr.ctypes.data, dom.ctypes.data, tri.ctypes.data, tri.shape[0], out.ctypes.data, rec.ctypes.data)
        for t in range(tri.shape[0]):
            tv = tri[t].astype(np.float64)
            S = 40000
            uu = rng.uniform(-0.6, 1.6, S); vv = rng.uniform(-0.6, 1.6, S)
            edge = rng.integers(0, 4, S)
            uu = np.where(edge == 1, rng.normal(0, 1e-4, S), uu); vv = np.where(edge == 2, rng.normal(0, 1e-4, S), vv)
            vv = np.where(edge == 3, 1 - uu + rng.normal(0, 1e-4, S), vv)
            X = tv[0:3] + uu[:, None] * tv[3:6] + vv[:, None] * tv[6:9]
            # origins on the line from the light through X, on either side of X and of the light
            lam = rng.uniform(-1.5, 3.0, S)
            P = (org.astype(np.float64)[None, :] + (X - org.astype(np.float64)[None, :]) * lam[:, None]).astype(f32)
            keep = np.abs(P.astype(np.float64)).max(axis=1) <= extent * 1.5
            P = P[keep]
            d = (org[None, :] - P).astype(f32)                                           # gpu/light.cu:103
            acc = _shadow_test(P, d, tri[t])
            r3 = (P - org[None, :]).astype(f32)                                          # lg_project, float32
            den = ((r3[:, 0] * az32[0] + r3[:, 1] * az32[1]) + r3[:, 2] * az32[2]).astype(f32)
            with np.errstate(divide="ignore", invalid="ignore"):
                x = (((r3[:, 0] * ax[0] + r3[:, 1] * ax[1]) + r3[:, 2] * ax[2]).astype(f32) / den).astype(f32)
                y = (((r3[:, 0] * ay[0] + r3[:, 1] * ay[1]) + r3[:, 2] * ay[2]).astype(f32) / den).astype(f32)
            # rays the grid answers: inside the limit, not grazing the light's plane (lg_any)
            l1 = np.abs(r3).sum(axis=1)
            ongrid = (np.abs(x) <= lim) & (np.abs(y) <= lim) & (np.abs(den) > 1e-6 * l1)
            acc &= ongrid
            if out[t, 0] == 0:
                assert not acc.any(), t
                continue
            inb = (x >= out[t, 1]) & (x <= out[t, 2]) & (y >= out[t, 3]) & (y <= out[t, 4])
            r = np.broadcast_to(rec[t], (P.shape[0], 12))
            ins = inb & record_inside(r, x - rec[t, 9], y - rec[t, 10])
            assert not (acc & ~ins).any(), (rep, t, int((acc & ~ins).sum()))
            accepted += int(acc.sum())
    assert accepted > 100000, accepted
