"""The light-space cull of the shadow-ray grids, replayed on the CPU (numpy).

csrc/lightgrid.cuh lg_region / k_lg_records give every (light, triangle) three edge lines in the
light's projection; lg_any skips a triangle when the ray's point is outside one of them.  This test
rebuilds those lines in float64 with the constants of the source (RAST_HULL_C, the grid's slack of
trace.cu build_light_grids), makes shadow rays the way apply_light does (origin = a float32 point
of the scene, direction -v for a directional light resp. v - P for a point light, float32, NOT
normalised: gpu/light.cu:68-120), runs the rounded Moeller-Trumbore test on them and requires the
projected origin of every ACCEPTED ray to be inside the three lines.  Method and constants, not
the CUDA code: that is tests/test_gpu_raster.py::test_shadow_ray_grids_give_the_tree_walk_frames."""
import numpy as np

from test_accept_region_bound import _cross, _dot, _rounded_test

f32 = np.float32
EPS24 = 2.0 ** -24
HULL_C, MIN_COS = 14.0, 1.0 / 256
ABS_MAX = 500.0
DIAG = 1.001 * np.sqrt(3.0) * 2 * ABS_MAX


def _unit32(v):
    v = np.asarray(v, f32)
    r = np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
    return (v / r).astype(f32)


def _frame32(az):
    """build_light_grids: any orthonormal frame around az, float32"""
    a = np.abs(az)
    h = np.array([1, 0, 0], f32) if a[0] <= a[1] and a[0] <= a[2] else (np.array([0, 1, 0], f32) if a[1] <= a[2] else np.array([0, 0, 1], f32))
    ax = _unit32(np.cross(h, az).astype(f32))
    ay = np.cross(az, ax).astype(f32)
    return ax, ay


def _triangles(rng, n):
    centre = rng.uniform(-300, 300, (n, 3))
    size = np.exp(rng.uniform(np.log(0.3), np.log(5.0), n))
    a = rng.standard_normal((n, 3)); a /= np.linalg.norm(a, axis=1)[:, None]
    b = _cross(a, rng.standard_normal((n, 3))); b /= np.linalg.norm(b, axis=1)[:, None]
    ang = rng.uniform(np.radians(50), np.radians(100), n)
    r2 = rng.uniform(0.8, 1.25, n)
    e1 = a * size[:, None]
    e2 = (a * np.cos(ang)[:, None] + b * np.sin(ang)[:, None]) * (size * r2)[:, None]
    v0 = centre - (e1 + e2) / 3
    return v0.astype(f32), e1.astype(f32), e2.astype(f32)


def _edge_targets(rng, v0, e1, e2, band):
    """points near a random edge of each triangle, band = offset in barycentric units"""
    n = len(v0)
    t = rng.random(n); which = rng.integers(0, 3, n)
    uu = np.where(which == 0, 0.0, t) + rng.uniform(-1, 1, n) * band
    vv = np.where(which == 0, t, np.where(which == 1, 0.0, 1.0 - t)) + rng.uniform(-1, 1, n) * band
    return v0.astype(np.float64) + uu[:, None] * e1.astype(np.float64) + vv[:, None] * e2.astype(np.float64)


def _region_lines(kind, org, ax, ay, az, slack, v0, e1, e2):
    """lg_region + k_lg_records in float64; returns ok and the three (nx, ny, c) with inside <=> n.(x,y) + c <= 0"""
    v0, e1, e2 = (x.astype(np.float64) for x in (v0, e1, e2))
    ax, ay, az, org = (x.astype(np.float64) for x in (ax, ay, az, org))
    nrm = _cross(e1, e2)
    ln, l1, l2 = (np.linalg.norm(x, axis=1) for x in (nrm, e1, e2))
    if kind == 1:
        cos = np.abs(nrm @ az) / ln
        far = np.zeros(len(v0))
    else:
        p0 = org[None, :] - v0
        far = np.maximum(np.linalg.norm(p0, axis=1), np.maximum(np.linalg.norm(org - (v0 + e1), axis=1), np.linalg.norm(org - (v0 + e2), axis=1))) + 6 * (l1 + l2)
        cos = np.abs(_dot(nrm, p0)) / (ln * far)
    du = HULL_C * EPS24 * DIAG * l2 / (ln * cos)
    dv = HULL_C * EPS24 * DIAG * l1 / (ln * cos)
    ok = (cos >= MIN_COS) & (du < 2) & (dv < 2)
    dw = du + dv + 2.4e-7
    P = [v0 - du[:, None] * e1 - dv[:, None] * e2, v0 + (1 + dw + dv)[:, None] * e1 - dv[:, None] * e2,
         v0 - du[:, None] * e1 + (1 + dw + du)[:, None] * e2]

    def proj(X):
        if kind == 1:
            return X @ ax, X @ ay, np.ones(len(X))
        r = X - org[None, :]
        den = r @ az
        return (r @ ax) / den, (r @ ay) / den, den
    px, py, den = (np.stack(z) for z in zip(*(proj(p) for p in P)))
    if kind == 2:
        ok &= (np.sign(den) == np.sign(den[0])).all(axis=0) & (np.abs(den).min(axis=0) > 1e-3 * far)
    big = np.maximum(np.abs(px).max(axis=0), np.abs(py).max(axis=0))
    longest = np.zeros(len(v0))
    for k in range(3):
        kb = (k + 1) % 3
        longest = np.maximum(longest, np.maximum(np.abs(px[kb] - px[k]), np.abs(py[kb] - py[k])))
    lines = []
    for k in range(3):
        kb, kc = (k + 1) % 3, (k + 2) % 3
        nx, ny = py[kb] - py[k], px[k] - px[kb]
        ll = np.sqrt(nx * nx + ny * ny)
        nx, ny = nx / ll, ny / ll
        side = nx * (px[kc] - px[k]) + ny * (py[kc] - py[k])
        ok &= (ll > 0) & (np.abs(side) > 1e-3 * longest)
        flip = side > 0
        nx, ny = np.where(flip, -nx, nx), np.where(flip, -ny, ny)
        off = (slack + 1e-6 * big) * (np.abs(nx) + np.abs(ny))
        lines.append((nx, ny, -(nx * px[k] + ny * py[k]) - off))
    return ok, lines


def _dot32(A, b):
    return A[:, 0] * b[0] + A[:, 1] * b[1] + A[:, 2] * b[2]


def _run(kind, rng, n):
    v0, e1, e2 = _triangles(rng, n)
    if kind == 1:
        lv = (rng.standard_normal(3) * 1.3).astype(f32)
        az = _unit32(lv)
        org = np.zeros(3, f32)
        slack = 4e-6 * (2.0 * ABS_MAX)
    else:
        lv = np.array([rng.uniform(-400, 400), rng.uniform(250, 480), rng.uniform(-400, 400)], f32)      # the light's position
        az = _unit32(-lv)                                                                                 # towards the scene's centre
        org = lv
        slack = 2e-5
    ax, ay = _frame32(az)
    ok, lines = _region_lines(kind, org, ax, ay, az, slack, v0, e1, e2)
    # rays aimed a few bound-widths around the edges; origins up to 600 units before the triangle, inside the scene's box
    nrm = _cross(e1.astype(np.float64), e2.astype(np.float64))
    size = np.linalg.norm(e1.astype(np.float64), axis=1)
    band = 6 * HULL_C * EPS24 * DIAG / (size * 0.3)
    target = _edge_targets(rng, v0, e1, e2, band)
    if kind == 1:
        dirn = -lv.astype(np.float64)[None, :].repeat(n, 0)
    else:
        dirn = lv.astype(np.float64)[None, :] - target
    dirn_u = dirn / np.linalg.norm(dirn, axis=1)[:, None]
    back = np.exp(rng.uniform(np.log(0.5), np.log(600.0), n))
    P = (target - dirn_u * back[:, None])
    inside_box = (np.abs(P) <= ABS_MAX).all(axis=1)
    P32 = P.astype(f32)
    d32 = np.repeat((-lv)[None, :], n, 0).astype(f32) if kind == 1 else (lv[None, :] - P32).astype(f32)
    acc = _rounded_test(P32, d32, v0, e1, e2) & ok & inside_box
    # the ray's point as lg_any computes it (float32)
    if kind == 1:
        x, y = _dot32(P32, ax), _dot32(P32, ay)
    else:
        r = P32 - org[None, :]
        den = _dot32(r, az)
        x, y = _dot32(r, ax) / den, _dot32(r, ay) / den
    x, y = x.astype(np.float64), y.astype(np.float64)
    worst = -np.inf
    for nx, ny, c in lines:
        e = nx * x + ny * y + c
        assert (e[acc] <= 0).all(), (kind, float(e[acc].max()))
        worst = max(worst, float(e[acc].max()))
    return int(acc.sum()), int(ok.sum()), worst


def test_directional_light_regions_hold_every_accepted_shadow_ray():
    rng = np.random.default_rng(99)
    total = 0
    for _ in range(6):
        acc, ok, worst = _run(1, rng, 300_000)
        assert ok > 250_000
        total += acc
    assert total > 300_000, total


def test_point_light_regions_hold_every_accepted_shadow_ray():
    rng = np.random.default_rng(100)
    total = 0
    for _ in range(6):
        acc, ok, worst = _run(2, rng, 300_000)
        total += acc
    assert total > 200_000, total
