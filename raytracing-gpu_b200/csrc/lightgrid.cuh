// Shadow rays by projection (included by trace.cu).
//
// All shadow rays of a directional light are parallel, all shadow rays of a point light lie on
// lines through the light's position (gpu/light.cu:68-120: direction -v, resp. v - P), and a
// shadow ray only asks "any candidate on the ray" (gpu/light.cu:29-36).  Seen along the light,
// a ray is a POINT and a triangle's culling box is a rectangle, so the candidates of a ray are the
// triangles whose rectangle covers that point: a uniform grid in the light's projection, built
// once per scene (lights are part of the scene), each cell listing the triangles whose box
// (vertex bounds + needle padding + ray-time margin, the boxes the tree culls with) touches it --
// for a triangle without needle padding the box's rectangle is cut down to the bounds of the
// region the rounded test can reach (lg_region), and an edge record per (light, triangle) holds
// that region's three edge lines.  A ray looks up one cell, checks the entries' rectangles, then
// the survivors' edge records, and runs the exact per-ray test on what is left: ~10 byte compares,
// ~5 records and ~2 triangle tests instead of ~30-40 node visits on the 10 M-triangle terrain.  Triangles that cover too many
// cells, and (point lights) boxes across the plane through the light, are tested by every ray;
// too many of those, or a list volume out of proportion, and the light keeps the tree walk.  Rays
// whose point falls outside the grid walk the tree as well.
#pragma once

namespace {

constexpr unsigned LG_CELLS_MAX = 1024;      // per triangle
constexpr float LG_PERSP_LIMIT = 4.0f;       // point lights: the grid spans |x|, |y| <= 4 on the plane at distance 1

// x, y: position in the light's projection; den: depth along az for a point light (sign = side of the light)
__device__ __forceinline__ void lg_project(const LightGrid &g, const V3 X, float &x, float &y, float &den) {
  if (g.kind == 1) {
    x = vdot(X, mk(g.ax)); y = vdot(X, mk(g.ay)); den = 1.0f;
  } else {
    const V3 r = vsub(X, mk(g.org));
    den = vdot(r, mk(g.az));
    x = vdot(r, mk(g.ax)) / den; y = vdot(r, mk(g.ay)) / den;
  }
}

// The region of triangle slot s's plane where the rounded test can accept a shadow ray of this
// light, as a triangle in the light's projection (corners ax, ay).  Derivation: raster.cuh,
// k_rast_setup; the direction of a shadow ray is not normalised (gpu/light.cu:68-120) but u and v do
// not depend on its length; |origin - vertex| <= smax, the scene's diagonal (origins are hit
// points); the incidence is that of the light's direction (1) or of the line from the light to the
// farthest corner (2).  false: no such region (needle-padded triangle, edge-on, near the light's
// plane) -- the triangle is culled by its box alone.
__device__ __forceinline__ bool lg_region(const LightGrid &g, float pad_unit, float smax, const float4 a4, const float4 b4,
                                          const float4 c4, float (&ax)[3], float (&ay)[3]) {
  const V3 v[3] = {mk(a4), vadd(mk(a4), mk(b4)), vadd(mk(a4), mk(c4))};
  V3 lo = mk(fminf(v[0].x, fminf(v[1].x, v[2].x)), fminf(v[0].y, fminf(v[1].y, v[2].y)), fminf(v[0].z, fminf(v[1].z, v[2].z)));
  V3 hi = mk(fmaxf(v[0].x, fmaxf(v[1].x, v[2].x)), fmaxf(v[0].y, fmaxf(v[1].y, v[2].y)), fmaxf(v[0].z, fmaxf(v[1].z, v[2].z)));
  const V3 lo0 = lo, hi0 = hi;
  sliver_pad(v, pad_unit, lo, hi);
  if (!(lo.x == lo0.x && lo.y == lo0.y && lo.z == lo0.z && hi.x == hi0.x && hi.y == hi0.y && hi.z == hi0.z)) return false;
  const V3 n = vcross(mk(b4), mk(c4));
  const float ln = vlen(n), l1 = vlen(mk(b4)), l2 = vlen(mk(c4));
  float cosine, far = 0.0f;
  if (g.kind == 1) cosine = fabsf(vdot(n, mk(g.az))) / ln;
  else {
    const V3 p0 = vsub(mk(g.org), v[0]);
    far = fmaxf(vlen(p0), fmaxf(vlen(vsub(mk(g.org), v[1])), vlen(vsub(mk(g.org), v[2])))) + 6.0f * (l1 + l2);
    cosine = fabsf(vdot(n, p0)) / (ln * far);
  }
  const float du = RAST_HULL_C * 5.9604645e-8f * smax * l2 / (ln * cosine);
  const float dv = RAST_HULL_C * 5.9604645e-8f * smax * l1 / (ln * cosine);
  if (!(ln > 0.0f && cosine >= RAST_HULL_MIN_COS && du < 2.0f && dv < 2.0f)) return false;       // NaN -> false
  const float dw = du + dv + 2.4e-7f;
  const V3 P[3] = {vsub(vsub(v[0], vscale(mk(b4), du)), vscale(mk(c4), dv)),
                   vsub(vadd(v[0], vscale(mk(b4), 1.0f + dw + dv)), vscale(mk(c4), dv)),
                   vadd(vsub(v[0], vscale(mk(b4), du)), vscale(mk(c4), 1.0f + dw + du))};
  float den[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) lg_project(g, P[k], ax[k], ay[k], den[k]);
  if (g.kind == 2) {
    const float near = fminf(fabsf(den[0]), fminf(fabsf(den[1]), fabsf(den[2])));
    if (!(den[0] * den[1] > 0.0f && den[0] * den[2] > 0.0f && near > 1e-3f * far)) return false;
  }
  const float big = fmaxf(fmaxf(fabsf(ax[0]), fabsf(ay[0])), fmaxf(fmaxf(fabsf(ax[1]), fabsf(ay[1])), fmaxf(fabsf(ax[2]), fabsf(ay[2]))));
  return big < 1e30f;
}

// The projected rectangle of triangle slot s's culling box, intersected with the bounds of its
// region (lg_region) when it has one.  false: the box has no rectangle
// (it crosses the plane through a point light) and must be tested by every ray.
__device__ __forceinline__ bool lg_rect(const DevScene &sc, const LightGrid &g, float eps, float smax, uint32_t s, float &x0,
                                        float &x1, float &y0, float &y1) {
  const float INFF = __builtin_huge_valf();
  const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
  const V3 v[3] = {mk(a4), vadd(mk(a4), mk(b4)), vadd(mk(a4), mk(c4))};
  V3 lo = mk(fminf(v[0].x, fminf(v[1].x, v[2].x)), fminf(v[0].y, fminf(v[1].y, v[2].y)), fminf(v[0].z, fminf(v[1].z, v[2].z)));
  V3 hi = mk(fmaxf(v[0].x, fmaxf(v[1].x, v[2].x)), fmaxf(v[0].y, fmaxf(v[1].y, v[2].y)), fmaxf(v[0].z, fmaxf(v[1].z, v[2].z)));
  sliver_pad(v, sc.pad_unit, lo, hi);
  lo = mk(lo.x - eps, lo.y - eps, lo.z - eps);
  hi = mk(hi.x + eps, hi.y + eps, hi.z + eps);
  float dmin = INFF, dmax = -INFF;
  x0 = y0 = INFF; x1 = y1 = -INFF;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    float x, y, den;
    lg_project(g, mk((k & 1) ? hi.x : lo.x, (k & 2) ? hi.y : lo.y, (k & 4) ? hi.z : lo.z), x, y, den);
    dmin = fminf(dmin, den); dmax = fmaxf(dmax, den);
    x0 = fminf(x0, x); x1 = fmaxf(x1, x); y0 = fminf(y0, y); y1 = fmaxf(y1, y);
  }
  if (g.kind == 2) {
    const float near = fminf(fabsf(dmin), fabsf(dmax));
    if (!(dmin * dmax > 0.0f) || !(near > 1e-3f * (dmax - dmin))) return false;
  }
  if (!(x0 > -1e30f && x1 < 1e30f && y0 > -1e30f && y1 < 1e30f)) return false;
  // the rounding of the projection itself, ray side and box side (g.slack: build_light_grids)
  x0 -= g.slack; x1 += g.slack; y0 -= g.slack; y1 += g.slack;
  float ax[3], ay[3];
  if (smax > 0.0f && lg_region(g, sc.pad_unit, smax, a4, b4, c4, ax, ay)) {
    const float big = fmaxf(fmaxf(fabsf(ax[0]), fabsf(ay[0])), fmaxf(fmaxf(fabsf(ax[1]), fabsf(ay[1])), fmaxf(fabsf(ax[2]), fabsf(ay[2]))));
    const float sl = g.slack + 1e-6f * big;
    x0 = fmaxf(x0, fminf(ax[0], fminf(ax[1], ax[2])) - sl); x1 = fminf(x1, fmaxf(ax[0], fmaxf(ax[1], ax[2])) + sl);
    y0 = fmaxf(y0, fminf(ay[0], fminf(ay[1], ay[2])) - sl); y1 = fminf(y1, fmaxf(ay[0], fmaxf(ay[1], ay[2])) + sl);
  }
  return true;
}

// cells [i0, i1] x [j0, j1] of the rectangle, clipped to the grid; false: nothing inside
__device__ __forceinline__ bool lg_cells(const LightGrid &g, float x0, float x1, float y0, float y1, int &i0, int &i1,
                                         int &j0, int &j1) {
  const float fx0 = floorf((x0 - g.x0) * g.inv_cell), fx1 = floorf((x1 - g.x0) * g.inv_cell);
  const float fy0 = floorf((y0 - g.y0) * g.inv_cell), fy1 = floorf((y1 - g.y0) * g.inv_cell);
  if (fx1 < 0.0f || fy1 < 0.0f || fx0 > (float)(g.nx - 1) || fy0 > (float)(g.ny - 1)) return false;
  i0 = (int)fmaxf(fx0, 0.0f); i1 = (int)fminf(fx1, (float)(g.nx - 1));
  j0 = (int)fmaxf(fy0, 0.0f); j1 = (int)fminf(fy1, (float)(g.ny - 1));
  return true;
}

// bounds of the rectangles of every stride-th triangle: out[0..3] = min x, max x, min y, max y
// (ordered uint).  The grid does not have to cover every rectangle: k_lg_scatter clips to the grid
// and a ray whose point is off the grid walks the tree, so big scenes look at a sample.
__global__ void __launch_bounds__(256) k_lg_bounds(const DevScene sc, const LightGrid g, float eps, float smax, uint32_t stride,
                                                   uint32_t *out) {
  const uint32_t s = (blockIdx.x * 256 + threadIdx.x) * stride;
  const float INFF = __builtin_huge_valf();
  float x0 = INFF, x1 = -INFF, y0 = INFF, y1 = -INFF;
  if (s < sc.n_tris) {
    float a0, a1, b0, b1;
    if (lg_rect(sc, g, eps, smax, s, a0, a1, b0, b1)) { x0 = a0; x1 = a1; y0 = b0; y1 = b1; }
  }
  const uint32_t m0 = __reduce_min_sync(0xffffffffu, f2ord(x0)), m1 = __reduce_max_sync(0xffffffffu, f2ord(x1));
  const uint32_t m2 = __reduce_min_sync(0xffffffffu, f2ord(y0)), m3 = __reduce_max_sync(0xffffffffu, f2ord(y1));
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out + 0, m0); atomicMax(out + 1, m1); atomicMin(out + 2, m2); atomicMax(out + 3, m3);
  }
}

// FILL = false: counts[cell] += 1 per covered cell, state[0] = number of large triangles (listed in
// state[2..]), state[1] != 0: give up.  FILL = true: entries[cell_start[cell] + cursor[cell]++] = slot.
// An entry is (slot, the rectangle clipped to the cell in 1/127 of the cell, rounded outwards), the
// rectangle as four 7-bit bytes x0, 127 - x1, y0, 127 - y1: a ray turns its point into the bytes
// floor(qx), 127 - ceil(qx), floor(qy), 127 - ceil(qy) with bit 7 set in each, and "inside" is
// "every byte of the entry <= the ray's byte", i.e. no byte of (ray - entry) loses its bit 7: one
// subtraction, one AND, one compare on 4 sequential bytes, before the ray touches the triangle's
// edge record or the triangle itself (scattered 16-byte loads and the exact test).
constexpr float LG_Q = 127.0f;
__device__ __forceinline__ uint32_t lg_pack(const LightGrid &g, int i, int j, float x0, float x1, float y0, float y1) {
  const float cx = g.x0 + (float)i / g.inv_cell, cy = g.y0 + (float)j / g.inv_cell;
  auto q = [&](float v, float c, bool up) {
    float t = (v - c) * g.inv_cell * LG_Q;
    t = up ? ceilf(t + 0.02f) : floorf(t - 0.02f);
    return (uint32_t)fminf(fmaxf(t, 0.0f), LG_Q);
  };
  return q(x0, cx, false) | ((127u - q(x1, cx, true)) << 8) | (q(y0, cy, false) << 16) | ((127u - q(y1, cy, true)) << 24);
}

template <bool FILL>
__global__ void __launch_bounds__(256) k_lg_scatter(const DevScene sc, const LightGrid g, float eps, float smax, uint32_t *counts,
                                                    uint32_t *state, const uint32_t *cell_start, uint2 *entries) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  if (s >= sc.n_tris) return;
  float x0, x1, y0, y1;
  int i0, i1, j0, j1;
  bool large = !lg_rect(sc, g, eps, smax, s, x0, x1, y0, y1);
  if (!large) {
    if (!lg_cells(g, x0, x1, y0, y1, i0, i1, j0, j1)) return;       // entirely outside the grid
    large = (unsigned)(i1 - i0 + 1) * (unsigned)(j1 - j0 + 1) > LG_CELLS_MAX;
  }
  if (large) {
    if (!FILL) {
      const unsigned k = atomicAdd(state, 1u);
      if (k < (unsigned)LG_LARGE_MAX) state[2 + k] = s; else state[1] = 1u;
    }
    return;
  }
  for (int j = j0; j <= j1; ++j)
    for (int i = i0; i <= i1; ++i) {
      const size_t c = (size_t)j * g.nx + i;
      const uint32_t k = atomicAdd(counts + c, 1u);
      if (FILL) entries[cell_start[c] + k] = make_uint2(s, lg_pack(g, i, j, x0, x1, y0, y1));
    }
}

// Edge record of a triangle in one light's projection: rec[3 slot + 0..2] = (n0x, n0y, c0, n1x)
// (n1y, c1, n2x, n2y) (c2, -, -, -); a ray whose point (x, y) has n . (x, y) + c > 0 for one of the
// three edges of the triangle's region (lg_region) cannot be accepted by the rounded test and
// skips the triangle without touching it.  Triangles without a region, and regions seen edge-on,
// get the record that accepts every point.
__global__ void __launch_bounds__(256) k_lg_records(const DevScene sc, const LightGrid g, float smax, float4 *rec) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  if (s >= sc.n_tris) return;
  float4 r0 = make_float4(0, 0, -1, 0), r1 = make_float4(0, -1, 0, 0), r2 = make_float4(-1, 0, 0, 0);
  const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
  float ax[3], ay[3];
  if (lg_region(g, sc.pad_unit, smax, a4, b4, c4, ax, ay)) {
    const float big = fmaxf(fmaxf(fabsf(ax[0]), fabsf(ay[0])), fmaxf(fmaxf(fabsf(ax[1]), fabsf(ay[1])), fmaxf(fabsf(ax[2]), fabsf(ay[2]))));
    bool ok = true;
    float en[3][3], longest = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int kb = (k + 1) % 3;
      longest = fmaxf(longest, fmaxf(fabsf(ax[kb] - ax[k]), fabsf(ay[kb] - ay[k])));
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int kb = (k + 1) % 3, kc = (k + 2) % 3;
      float nx = ay[kb] - ay[k], ny = ax[k] - ax[kb];
      const float len = sqrtf(nx * nx + ny * ny);
      nx /= len; ny /= len;
      const float side = nx * (ax[kc] - ax[k]) + ny * (ay[kc] - ay[k]);
      ok = ok && len > 0.0f && fabsf(side) > 1e-3f * longest;       // edge-on in this projection: no record
      if (side > 0.0f) { nx = -nx; ny = -ny; }
      // g.slack: rounding of the projection, ray side and triangle side, and of the ray's direction
      // (lg_rect); the last term: rounding of n . (x, y) + c at coordinates of this size
      const float off = (g.slack + 1e-6f * big) * (fabsf(nx) + fabsf(ny));
      en[k][0] = nx; en[k][1] = ny; en[k][2] = -(nx * ax[k] + ny * ay[k]) - off;
    }
    if (ok) {
      r0 = make_float4(en[0][0], en[0][1], en[0][2], en[1][0]);
      r1 = make_float4(en[1][1], en[1][2], en[2][0], en[2][1]);
      r2 = make_float4(en[2][2], 0, 0, 0);
    }
  }
  rec[3 * (size_t)s] = r0; rec[3 * (size_t)s + 1] = r1; rec[3 * (size_t)s + 2] = r2;
}

// 1: some candidate on the ray, 0: none, -1: this ray has to walk the tree (its point is off the grid)
__device__ __forceinline__ int lg_any(const DevScene &sc, const LightGrid &g, const V3 o, const V3 d) {
  float x, y, den;
  lg_project(g, o, x, y, den);
  const float fi = floorf((x - g.x0) * g.inv_cell), fj = floorf((y - g.y0) * g.inv_cell);
  if (!(fi >= 0.0f && fj >= 0.0f && fi <= (float)(g.nx - 1) && fj <= (float)(g.ny - 1))) return -1;
  if (g.kind == 2 && !(fabsf(den) > 1e-6f * (fabsf(o.x - g.org.x) + fabsf(o.y - g.org.y) + fabsf(o.z - g.org.z)))) return -1;
  Best b;
  b.dist = INF; b.id = 0x7fffffff; b.slot = -1;
  for (uint32_t k = 0; k < g.n_large; ++k)
    if (consider<true>(sc, o, d, (int)g.large[k], b)) return 1;
  const size_t c = (size_t)(int)fj * g.nx + (int)fi;
  const uint32_t e0 = __ldg(g.cell_start + c), e1 = __ldg(g.cell_start + c + 1);
  // the ray's point inside its cell as the four bytes the entries compare with (lg_pack); 0.01:
  // rounding of the point's own position inside the cell
  const float qx = ((x - g.x0) * g.inv_cell - fi) * LG_Q, qy = ((y - g.y0) * g.inv_cell - fj) * LG_Q;
  auto byte = [](float v) { return (uint32_t)fminf(fmaxf(v, 0.0f), LG_Q); };
  const uint32_t P = (byte(floorf(qx + 0.01f)) | ((127u - byte(ceilf(qx - 0.01f))) << 8) | (byte(floorf(qy + 0.01f)) << 16) |
                      ((127u - byte(ceilf(qy - 0.01f))) << 24)) | 0x80808080u;
  // Three phases per chunk of 32 entries, because the lanes of a warp walk different lists in
  // lockstep: first only the cheap rectangle checks (a bit per entry that survives), ...
  for (uint32_t base = e0; base < e1; base += 32) {
    const uint32_t cnt = min(32u, e1 - base);
    uint32_t mask = 0;
#pragma unroll 4
    for (uint32_t k = 0; k < cnt; ++k) {
      const uint32_t r = __ldg(&g.entries[base + k].y);
      if (((P - r) & 0x80808080u) == 0x80808080u) mask |= 1u << k;
    }
    // ... then the edge records of the survivors (two 16-byte loads and three dot products each),
    // again a loop of its own: what is left for the exact test is a fraction of the rectangle hits
    if (g.rec) {
      uint32_t m = mask;
      while (m) {
        const uint32_t k = (uint32_t)__ffs((int)m) - 1u;
        m &= m - 1u;
        const uint32_t slot = __ldg(&g.entries[base + k].x);
        const float4 r0 = __ldg(g.rec + 3 * (size_t)slot), r1 = __ldg(g.rec + 3 * (size_t)slot + 1);
        const float c2 = __ldg(&g.rec[3 * (size_t)slot + 2].x);
        if (r0.x * x + r0.y * y + r0.z > 0.0f || r0.w * x + r1.x * y + r1.y > 0.0f || r1.z * x + r1.w * y + c2 > 0.0f)
          mask &= ~(1u << k);
      }
    }
    while (mask) {
      const uint32_t k = (uint32_t)__ffs((int)mask) - 1u;
      mask &= mask - 1u;
      if (consider<true>(sc, o, d, (int)__ldg(&g.entries[base + k].x), b)) return 1;
    }
  }
  return 0;
}

}  // namespace
