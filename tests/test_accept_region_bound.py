"""The rounding model behind the accept regions of the projection passes (CPU, numpy).

csrc/region.cuh bounds where the ROUNDED Moeller-Trumbore test of the reference (gpu/hit.cu:8-60,
un-fused float32) can accept a ray from the error of the computed numerators:
|fl(s . h) - s . h| < 7 eps |s| |d| |e2| (and the same with e1 for v), eps = 2^-24, s = origin -
vertex0, h = d x e2, times SAFETY = 1.25 for the second order.  In barycentric terms an accepted
ray therefore has  u >= -du, v >= -dv, u + v <= 1 + du + dv + 4 eps  with du = C eps |s||e2||d| / |a|,
dv = C eps |s||e1||d| / |a|, C = 7 x 1.25, a = e1 . h.  This test replays the test in numpy float32,
operation by operation, on rays aimed at the edges of triangles seen from far away (the geometry of
the C5 frame: |s| of thousands of triangle sizes) down to grazing incidence, and checks the exact
(float64) u, v of every ACCEPTED ray against that bound.  The regions themselves (edge functions,
clipping, records, spans) are tested on the shipped code in tests/test_region_cpu.py."""
import numpy as np

f32 = np.float32
EPS24 = 2.0 ** -24
C = 7.0 * 1.25                # region.cuh: 7 eps per numerator, SAFETY
MIN_COS = 1.0 / 256           # flatter than this the sample below would need |s| beyond any frame


def _cross(a, b):
    return np.stack([a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1], a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2],
                     a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]], axis=1)


def _dot(a, b):
    return a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1] + a[:, 2] * b[:, 2]


def _rounded_test(o, d, v0, e1, e2):
    """tri_candidate of csrc/trace.cu up to the u, v tests, float32 throughout (numpy never fuses)"""
    assert o.dtype == d.dtype == v0.dtype == e1.dtype == e2.dtype == f32
    h = _cross(d, e2)
    a = _dot(e1, h)
    ok = ~((a > f32(-1e-7)) & (a < f32(1e-7)))
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        f = f32(1) / a
        s = o - v0
        u = f * _dot(s, h)
        ok &= ~((u < 0) | (u > 1))
        q = _cross(s, e1)
        v = f * _dot(d, q)
        ok &= ~((v < 0) | (u + v > 1))
    return ok


def _exact_uv(o, d, v0, e1, e2):
    o, d, v0, e1, e2 = (x.astype(np.float64) for x in (o, d, v0, e1, e2))
    h = _cross(d, e2)
    a = _dot(e1, h)
    s = o - v0
    u = _dot(s, h) / a
    v = _dot(d, _cross(s, e1)) / a
    return u, v, a, s


def _scenes(rng, n, dist, size):
    """triangles of about `size` around the origin region, ray origins `dist` away, targets within a
    few bound-widths of a random edge point (inside and outside)"""
    v0 = rng.uniform(-500, 500, (n, 3))
    e1 = rng.standard_normal((n, 3)) * size
    e2 = rng.standard_normal((n, 3)) * size
    # keep the corner at vertex 0 and the one opposite wider than ~29 degrees: no needle padding
    c = np.abs(_dot(e1, e2)) / (np.linalg.norm(e1, axis=1) * np.linalg.norm(e2, axis=1))
    e2 = np.where((c > 0.8)[:, None], _cross(e1, rng.standard_normal((n, 3))) * (size / np.maximum(np.linalg.norm(e1, axis=1), 1e-9))[:, None], e2)
    dirs = rng.standard_normal((n, 3))
    dirs /= np.linalg.norm(dirs, axis=1)[:, None]
    o = v0 + dirs * dist
    # a point on one of the three edges
    t = rng.random(n)
    which = rng.integers(0, 3, n)
    uu = np.where(which == 0, 0.0, np.where(which == 1, t, t))
    vv = np.where(which == 0, t, np.where(which == 1, 0.0, 1.0 - t))
    nrm = _cross(e1, e2)
    cosv = np.abs(_dot(nrm, dirs)) / np.linalg.norm(nrm, axis=1)
    # offsets in barycentric units, a few times the bound at this incidence
    scale = 4 * C * EPS24 * dist / (size * np.maximum(cosv, MIN_COS))
    uu = uu + rng.uniform(-1, 1, n) * scale
    vv = vv + rng.uniform(-1, 1, n) * scale
    target = v0 + uu[:, None] * e1 + vv[:, None] * e2
    o32, v032, e132, e232 = (x.astype(f32) for x in (o, v0, e1, e2))
    dd = (target.astype(f32) - o32)
    ln = np.sqrt(dd[:, 0] * dd[:, 0] + dd[:, 1] * dd[:, 1] + dd[:, 2] * dd[:, 2])
    d32 = (dd / ln[:, None]).astype(f32)                 # normalised like primary_ray
    return o32, d32, v032, e132, e232


def _check(o, d, v0, e1, e2):
    ok = _rounded_test(o, d, v0, e1, e2)
    u, v, a, s = _exact_uv(o, d, v0, e1, e2)
    l1 = np.linalg.norm(e1.astype(np.float64), axis=1)
    l2 = np.linalg.norm(e2.astype(np.float64), axis=1)
    ld = np.linalg.norm(d.astype(np.float64), axis=1)
    ls = np.linalg.norm(s, axis=1)
    nrm = np.linalg.norm(_cross(e1.astype(np.float64), e2.astype(np.float64)), axis=1)
    cos = np.abs(a) / (nrm * ld)
    keep = ok & (cos >= MIN_COS)
    du = C * EPS24 * ls * l2 * ld / np.abs(a)
    dv = C * EPS24 * ls * l1 * ld / np.abs(a)
    keep &= (du < 2) & (dv < 2)                          # (beyond that the allowance is wider than the triangle)
    # how much of the allowance the worst accepted ray uses (1 = at the bound)
    used_u = np.where(keep, np.maximum(-u, 0) / du, 0)
    used_v = np.where(keep, np.maximum(-v, 0) / dv, 0)
    used_w = np.where(keep, np.maximum(u + v - 1, 0) / (du + dv + 4 * EPS24), 0)
    return int(keep.sum()), float(used_u.max()), float(used_v.max()), float(used_w.max())


def test_accepted_rays_cross_the_plane_inside_the_grown_triangle():
    rng = np.random.default_rng(20240611)
    worst = 0.0
    accepted = 0
    # the C5 frame (origins 8-9 thousand units away, triangles of 0.6), a near camera, shadow rays
    # (origin on a neighbouring surface, unnormalised direction is covered by scale invariance)
    for dist, size in ((9000.0, 0.6), (9000.0, 5.0), (700.0, 0.6), (30.0, 0.6), (1500.0, 40.0)):
        n, uu, vv, ww = _check(*_scenes(rng, 400_000, dist, size))
        accepted += n
        worst = max(worst, uu, vv, ww)
        assert n > 50_000, "the sample must contain accepted rays near the edges"
        assert max(uu, vv, ww) <= 1.0, (dist, size, uu, vv, ww)
    # a bound, so it has room: random rays use about 0.18 of it (1.5 eps |s||e||d| / |a|) on these two million
    assert worst <= 0.5, worst
    assert accepted > 500_000


def test_the_bound_is_not_vacuous():
    """without any allowance a visible share of the accepted rays lies outside the exact triangle"""
    rng = np.random.default_rng(7)
    o, d, v0, e1, e2 = _scenes(rng, 200_000, 9000.0, 0.6)
    ok = _rounded_test(o, d, v0, e1, e2)
    u, v, a, s = _exact_uv(o, d, v0, e1, e2)
    outside = ok & ((u < 0) | (v < 0) | (u + v > 1))
    assert outside.sum() > 100, int(outside.sum())
