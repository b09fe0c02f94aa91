"""The screen-space cull of the projection pass, replayed end to end on the CPU (numpy).

csrc/raster.cuh k_rast_setup turns a triangle into a rectangle of samples and three edge lines;
k_rast_items only runs the exact per-ray test on samples inside both.  This test rebuilds that
rectangle and those lines in float64 with the constants of the source (RAST_HULL_C, RAST_RAY_C,
the slack terms), generates the primary rays of the samples around them exactly as the kernel does
(float32, operation by operation: gpu/raytracer.cu:91-110), runs the rounded Moeller-Trumbore test
of the reference on them and requires every ACCEPTED sample to be inside the rectangle and the
edge lines -- for random cameras, depths and incidences.  It validates the method and its constants
(the ray's own rounding included), not the CUDA code: that is tests/test_gpu_raster.py."""
import numpy as np

from test_accept_region_bound import _cross, _dot, _rounded_test

f32 = np.float32
EPS24 = 2.0 ** -24
HULL_C, RAY_C, MIN_COS = 14.0, 8.0, 1.0 / 256      # csrc/pad.cuh


def _unit(v):
    return v / np.linalg.norm(v)


def _camera(rng, W, H, fov):
    w = _unit(rng.standard_normal(3))
    u = _unit(np.cross(w, rng.standard_normal(3)))
    v = np.cross(w, u)
    w = np.cross(u, v)                                  # as camera_frame: w = u x v
    pos = rng.uniform(-300, 300, 3)
    L = W / (2 * np.tan(fov * np.pi / 360))
    return dict(u=u.astype(f32), v=v.astype(f32), w=w.astype(f32), pos=pos.astype(f32), L=f32(L), W=W, H=H,
                C=(pos.astype(f32) + w.astype(f32) * f32(L)).astype(f32))


def _rays(cam, px, py):
    """primary_ray of csrc/trace.cu in float32"""
    fx = (px - cam["W"] // 2).astype(f32)[:, None]
    fy = (py - cam["H"] // 2).astype(f32)[:, None]
    ui = cam["u"][None, :] * fx
    vj = cam["v"][None, :] * fy
    o = (cam["C"][None, :] + ui) + vj
    dd = cam["pos"][None, :] - o
    ln = np.sqrt(dd[:, 0] * dd[:, 0] + dd[:, 1] * dd[:, 1] + dd[:, 2] * dd[:, 2])
    return o, (dd / ln[:, None]).astype(f32)


def _project(cam, X):
    u, v, w, pos = (cam[k].astype(np.float64) for k in ("u", "v", "w", "pos"))
    L = float(cam["L"])
    r = pos[None, :] - X
    den = r @ w
    t = L * (w @ w) / den
    q = r * t[:, None] - w[None, :] * L
    uv = u @ v
    det = 1 - uv * uv
    qu, qv = q @ u, q @ v
    return (qu - uv * qv) / det, (qv - uv * qu) / det, den


def _triangles(rng, cam, n):
    """well-shaped triangles of 3..15 samples across, on screen, 30..3000 units in front of the camera"""
    W, H = cam["W"], cam["H"]
    sx = rng.uniform(0.1 * W, 0.9 * W, n); sy = rng.uniform(0.1 * H, 0.9 * H, n)
    o, d = _rays(cam, sx.astype(np.int64), sy.astype(np.int64))
    depth = np.exp(rng.uniform(np.log(30), np.log(3000), n))
    centre = cam["pos"].astype(np.float64)[None, :] + d.astype(np.float64) * depth[:, None]
    size = rng.uniform(3, 15, n) * depth / float(cam["L"])
    a = rng.standard_normal((n, 3)); a /= np.linalg.norm(a, axis=1)[:, None]
    b = _cross(a, rng.standard_normal((n, 3))); b /= np.linalg.norm(b, axis=1)[:, None]
    ang = rng.uniform(np.radians(50), np.radians(100), n)         # corner at vertex 0; the others stay above 35 degrees
    r2 = rng.uniform(0.8, 1.25, n)
    e1 = a * size[:, None]
    e2 = (a * np.cos(ang)[:, None] + b * np.sin(ang)[:, None]) * (size * r2)[:, None]
    v0 = centre - (e1 + e2) / 3
    return v0.astype(f32), e1.astype(f32), e2.astype(f32)


def _cull(cam, v0, e1, e2):
    """rectangle and edge lines of k_rast_setup for triangles without needle padding (float64)"""
    pos = cam["pos"].astype(np.float64); L = float(cam["L"]); W, H = cam["W"], cam["H"]
    cx, cy = float(W // 2), float(H // 2)
    v0, e1, e2 = (x.astype(np.float64) for x in (v0, e1, e2))
    v1, v2 = v0 + e1, v0 + e2
    n = _cross(e1, e2)
    ln, l1, l2 = (np.linalg.norm(x, axis=1) for x in (n, e1, e2))
    far = np.maximum(np.linalg.norm(pos - v0, axis=1), np.maximum(np.linalg.norm(pos - v1, axis=1), np.linalg.norm(pos - v2, axis=1))) + 6 * (l1 + l2)
    cos = np.abs(_dot(n, pos[None, :] - v0)) / (ln * far)
    smax = np.sqrt(L * L + cx * cx + cy * cy) + far
    du = HULL_C * EPS24 * smax * l2 / (ln * cos)
    dv = HULL_C * EPS24 * smax * l1 / (ln * cos)
    ok = (cos >= MIN_COS) & (du < 2) & (dv < 2)
    dw = du + dv + 2.4e-7
    P = [v0 - du[:, None] * e1 - dv[:, None] * e2, v0 + (1 + dw + dv)[:, None] * e1 - dv[:, None] * e2,
         v0 - du[:, None] * e1 + (1 + dw + du)[:, None] * e2]
    ax, ay, den = zip(*(_project(cam, p) for p in P))
    ax, ay, den = np.stack(ax), np.stack(ay), np.stack(den)
    ok &= (np.sign(den) == np.sign(den[0])).all(axis=0)
    dnear = np.abs(den).min(axis=0)
    big = np.maximum(np.abs(ax).max(axis=0), np.abs(ay).max(axis=0))
    delta = RAY_C * EPS24 * smax
    ok &= dnear > 8 * delta
    sr = 1.3 * delta * np.sqrt(L * L + big * big) / dnear
    sb = 0.02 + 1e-6 * big
    se = sr + sb
    rect = (np.ceil(ax.min(axis=0) + cx - se), np.floor(ax.max(axis=0) + cx + se),
            np.ceil(ay.min(axis=0) + cy - se), np.floor(ay.max(axis=0) + cy + se))
    edges = []
    for k in range(3):
        kb, kc = (k + 1) % 3, (k + 2) % 3
        nx, ny = ay[kb] - ay[k], ax[k] - ax[kb]
        ll = np.sqrt(nx * nx + ny * ny)
        nx, ny = nx / ll, ny / ll
        side = nx * (ax[kc] - ax[k]) + ny * (ay[kc] - ay[k])
        ok &= (ll > 1e-3) & (np.abs(side) > 1e-3)
        flip = side > 0
        nx, ny = np.where(flip, -nx, nx), np.where(flip, -ny, ny)
        off = sr + sb * (np.abs(nx) + np.abs(ny))
        # inside  <=>  nx (px - cx - ax_k) + ny (py - cy - ay_k) - off <= 0
        edges.append((nx, ny, ax[k] + cx, ay[k] + cy, off))
    return ok, rect, edges


def test_every_accepted_sample_is_inside_rectangle_and_edge_lines():
    rng = np.random.default_rng(424242)
    accepted = outside_exact = 0
    for W, H, fov in ((11520, 6480, 70.0), (2880, 1620, 55.0), (640, 480, 100.0)):
        for _ in range(3):
            cam = _camera(rng, W, H, fov)
            v0, e1, e2 = _triangles(rng, cam, 3000)
            ok, (x0, x1, y0, y1), edges = _cull(cam, v0, e1, e2)
            assert ok.mean() > 0.8
            idx = np.nonzero(ok)[0]
            # every sample of the rectangle grown by 2 on each side (the cull may only be wrong outwards)
            for i in idx:
                xs = np.arange(int(x0[i]) - 2, int(x1[i]) + 3); ys = np.arange(int(y0[i]) - 2, int(y1[i]) + 3)
                px, py = (g.ravel() for g in np.meshgrid(xs, ys))
                keep = (px >= 0) & (px < W) & (py >= 0) & (py < H)
                px, py = px[keep], py[keep]
                if not len(px):
                    continue
                o, d = _rays(cam, px, py)
                m = len(px)
                acc = _rounded_test(o, d, np.repeat(v0[i:i + 1], m, 0), np.repeat(e1[i:i + 1], m, 0), np.repeat(e2[i:i + 1], m, 0))
                if not acc.any():
                    continue
                pxa, pya = px[acc].astype(np.float64), py[acc].astype(np.float64)
                assert ((pxa >= x0[i]) & (pxa <= x1[i]) & (pya >= y0[i]) & (pya <= y1[i])).all(), (W, i, "rectangle")
                for nx, ny, ex, ey, off in edges:
                    assert (nx[i] * (pxa - ex[i]) + ny[i] * (pya - ey[i]) - off[i] <= 0).all(), (W, i, "edge")
                accepted += int(acc.sum())
    assert accepted > 200_000, accepted
