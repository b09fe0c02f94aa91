// Primary visibility by projection (included by trace.cu).
//
// Every primary ray passes through the camera position (gpu/raytracer.cu:97-105: origin on the
// image plane, direction towards `pos`), and sample (px, py) is a regular grid on that plane.
// So "which rays can the rounded Moeller-Trumbore test accept for this triangle" is a set of
// samples on the screen: inside the projection through `pos` of the triangle's culling box, the
// very box the hierarchy culls with (vertex bounds + needle padding of pad.cuh + the ray-time
// margin of trace_tree), and, for a triangle without needle padding, inside the projection of
// the region of its plane that the rounding of u and v can reach (k_rast_setup).  One pass over
// the triangles emits (triangle, rectangle) items and a per-triangle edge record, one warp per
// item tests the rectangle's sample centres against the record and queues the survivors, full
// warps run the exact per-ray test of tri_candidate on them, and atomicMin on
// (new_dist bits << 32 | triangle id) keeps the reference's winner: lexicographic minimum of
// (new_dist, id), gpu/hit.cu:66-76.  k_frame then starts every sample from that record instead
// of walking the tree for its first ray; bounce rays still walk the tree (shadow rays: lightgrid.cuh).
//
// Never fewer candidates than the reference's exhaustive search accepts: a ray the tree would
// test against a triangle pierces that triangle's inflated box, and a line through `pos` that
// pierces a convex box projects inside the hull of the box's projected corners, as long as the
// box lies entirely on one side of the plane through `pos` parallel to the image plane; the
// region argument is above k_rast_setup and in DESIGN.md section 4.  Boxes that straddle that
// plane (their projection is unbounded) go to a short list every sample tests directly; more than
// RAST_STRADDLE_MAX of them, a triangle that spans more than RAST_TILES_MAX tiles, more items
// than the list holds or more rectangle area than RAST_AREA_FACTOR x the frame make the frame
// fall back to the tree for its primary rays (flag in device memory, no host round trip).
#pragma once

namespace {

constexpr int RAST_TILE = 32;                 // samples
#ifndef RTB_RAST_HULL
#define RTB_RAST_HULL 1
#endif
constexpr bool RAST_HULL = RTB_RAST_HULL != 0;
constexpr unsigned RAST_STRADDLE_MAX = 16;
constexpr unsigned RAST_TILES_MAX = 1u << 17;  // per triangle (a 4K x 3 frame has 73 000 tiles)
// rast[] words: 0 item count, 1 fallback flag, 2-3 rectangle area (u64), 4 straddler count, 8.. straddler slots
constexpr int RAST_WORDS = 8 + RAST_STRADDLE_MAX;

__device__ __forceinline__ bool rast_active(const FrameArgs &fa) {
  if (!fa.prim) return false;
  const unsigned *r = fa.rast;
  if (r[1]) return false;
  const unsigned long long area = *reinterpret_cast<const unsigned long long *>(r + 2);
  const unsigned long long samples = (unsigned long long)fa.cam.W * (unsigned long long)fa.cam.H / (unsigned)fa.nranks;
  return area <= (unsigned long long)fa.rast_area_factor * samples + 65536ull;
}

// sample tile (tx, ty) belongs to this rank?  Rank tiles are RANK_TILE pixels = RANK_TILE * a samples wide.
__device__ __forceinline__ bool rast_owned(const FrameArgs &fa, int tx, int ty) {
  if (fa.nranks == 1) return true;
  const int per = RANK_TILE * fa.aliasing / RAST_TILE;      // sample tiles per rank tile edge (a: 32a / 32)
  const int Tx = tx / per, Ty = ty / per;
  return (Tx + 3 * Ty) % fa.nranks == fa.rank;
}

// k_rast_setup: one thread per triangle slot.  Rectangle, per-triangle edge record, items.
//
// Rectangle = (projection of the culling box) INTERSECTED with, for a triangle without needle
// padding seen at a fair angle, (projection of the region of its plane where the rounded test can
// accept).  That region: the test accepts when the COMPUTED (u, v) lies in the unit simplex.  With
// eps = 2^-24, |d| = 1, s = origin - vertex0, h = d x e2, a = e1 . h (un-fused, as the reference):
//   |dh| <= 2 sqrt2 eps |e2|,  |d(s . h)| <= |s| |dh| + 3 eps |s| |h| + eps |s| |h| <= 6.9 eps |s| |e2|,
//   |da| <= 5.9 eps |e1| |e2|,  so  |du| <= 6.9 eps |s| |e2| / |a| + |u| 5.9 eps |e1| |e2| / |a| + 2 eps |u|
// and the same for v with e1; |a| = |e1 x e2| cos(incidence).  RAST_HULL_C = 14 doubles the leading
// constant (the |u| terms are smaller by |e| / |s|).  The true crossing point of an accepted ray
// therefore lies in  u >= -du, v >= -dv, u + v <= 1 + du + dv + 4 eps : a triangle in the plane
// whose corners are projected.  The cosine is (distance of pos from the plane) / |pos - point|,
// smallest at the farthest corner; |s| is bounded by the frame's farthest origin.
//
// Edge record: the three edge lines of that projected triangle, pushed out by the projected size
// of the ray's own rounding; k_rast_items tests the sample centre against them
// before it spends a ray on it.  rec[3 slot + 0..2] = (n0x, n0y, c0, n1x) (n1y, c1, n2x, n2y)
// (c2, X0, Y0, on): inside <=> n . (px - X0, py - Y0) + c <= 0 for the three edges.
__global__ void __launch_bounds__(256) k_rast_setup(const DevScene sc, const FrameArgs fa) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31;
  unsigned *rast = fa.rast;
  const float INF = __builtin_huge_valf();
  int x0 = 0, y0 = 0, x1 = -1, y1 = -1;        // empty rectangle
  bool straddle = false;
  float4 r0 = make_float4(0, 0, -1, 0), r1 = make_float4(0, -1, 0, 0), r2 = make_float4(-1, 0, 0, 0);
  if (s < sc.n_tris) {
    const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
    const V3 v[3] = {mk(a4), vadd(mk(a4), mk(b4)), vadd(mk(a4), mk(c4))};
    V3 lo = mk(fminf(v[0].x, fminf(v[1].x, v[2].x)), fminf(v[0].y, fminf(v[1].y, v[2].y)), fminf(v[0].z, fminf(v[1].z, v[2].z)));
    V3 hi = mk(fmaxf(v[0].x, fmaxf(v[1].x, v[2].x)), fmaxf(v[0].y, fmaxf(v[1].y, v[2].y)), fmaxf(v[0].z, fmaxf(v[1].z, v[2].z)));
    const V3 lo0 = lo, hi0 = hi;
    sliver_pad(v, sc.pad_unit, lo, hi);
    bool plain = RAST_HULL && lo.x == lo0.x && lo.y == lo0.y && lo.z == lo0.z && hi.x == hi0.x && hi.y == hi0.y && hi.z == hi0.z;
    // ray-time margin of trace_tree at the largest origin of this frame, plus the rounding of
    // vertex0 + edge (the build boxes come from the stored vertices themselves)
    const float e = fa.rast_eps;
    lo = mk(lo.x - e, lo.y - e, lo.z - e);
    hi = mk(hi.x + e, hi.y + e, hi.z + e);
    const V3 u = mk(fa.cam.u), vv = mk(fa.cam.v), w = mk(fa.cam.w), pos = mk(fa.cam.pos);
    const float ww = vdot(w, w), uv = vdot(u, vv);
    const float det = 1.0f - uv * uv;                      // Gram determinant of the unit vectors u, v
    const float Lw = fa.cam.L, inv_det = 1.0f / det;
    auto project = [&](const V3 X, float &x, float &y, float &den) {
      const V3 r = vsub(pos, X);
      den = vdot(r, w);
      // fast division and a reciprocal: a few more spacings of x, inside the slack below (4e-7 |x|)
      const float t = __fdividef(Lw * ww, den);
      const V3 q = vsub(vscale(r, t), vscale(w, Lw));     // offset from the plane's centre C
      const float qu = vdot(q, u), qv = vdot(q, vv);
      x = (qu - uv * qv) * inv_det;
      y = (qv - uv * qu) * inv_det;
    };
    float dmin = INF, dmax = -INF, xmin = INF, xmax = -INF, ymin = INF, ymax = -INF;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float x, y, den;
      project(mk((k & 1) ? hi.x : lo.x, (k & 2) ? hi.y : lo.y, (k & 4) ? hi.z : lo.z), x, y, den);
      dmin = fminf(dmin, den); dmax = fmaxf(dmax, den);
      xmin = fminf(xmin, x); xmax = fmaxf(xmax, x); ymin = fminf(ymin, y); ymax = fmaxf(ymax, y);
    }
    // one side of the plane through pos, and not grazing it (relative to the box's own depth range)
    const float near = fminf(fabsf(dmin), fabsf(dmax));
    straddle = !(dmin * dmax > 0.0f) || !(near > 1e-3f * (dmax - dmin)) || !(det > 1e-6f) ||
               !(xmin > -1e9f && xmax < 1e9f && ymin > -1e9f && ymax < 1e9f);
    if (!straddle) {
      // px - W/2 = x (integer division as in primary_ray: W / 2, H / 2).  Slack for the rounding of
      // the projection itself: a few float spacings of a coordinate that is at most ~1e5 samples
      const float cx = (float)(fa.cam.W / 2), cy = (float)(fa.cam.H / 2);
      const float sx = 0.02f + 1e-6f * fmaxf(fabsf(xmin), fabsf(xmax)), sy = 0.02f + 1e-6f * fmaxf(fabsf(ymin), fabsf(ymax));
      float fx0 = ceilf(xmin + cx - sx), fx1 = floorf(xmax + cx + sx);
      float fy0 = ceilf(ymin + cy - sy), fy1 = floorf(ymax + cy + sy);
      // nothing of the box on this rank's part of the screen: no region, no record, no items
      if (plain) {
        const int bx0 = (int)fmaxf(fx0, 0.0f), bx1 = (int)fminf(fx1, (float)(fa.cam.W - 1));
        const int by0 = (int)fmaxf(fy0, 0.0f), by1 = (int)fminf(fy1, (float)(fa.cam.H - 1));
        plain = bx1 >= bx0 && by1 >= by0;
        if (plain && fa.nranks > 1) {
          const int t0 = bx0 / RAST_TILE, t1 = bx1 / RAST_TILE, u0 = by0 / RAST_TILE, u1 = by1 / RAST_TILE;
          if ((t1 - t0 + 1) * (u1 - u0 + 1) <= 16) {
            bool mine = false;
            for (int ty = u0; ty <= u1; ++ty)
              for (int tx = t0; tx <= t1; ++tx) mine = mine || rast_owned(fa, tx, ty);
            plain = mine;
          }
        }
      }
      // the accepted region in the triangle's plane (see above)
      float du = 0.0f, dv = 0.0f, smax = 0.0f;
      if (plain) {
        const V3 n = vcross(mk(b4), mk(c4)), p0 = vsub(pos, v[0]);
        const float ln = vlen(n), l1 = vlen(mk(b4)), l2 = vlen(mk(c4));
        const float far0 = fmaxf(vlen(p0), fmaxf(vlen(vsub(pos, v[1])), vlen(vsub(pos, v[2]))));
        const float far = far0 + 6.0f * (l1 + l2);                        // the grown corners are farther (du, dv < 2 below)
        const float cosine = fabsf(vdot(n, p0)) / (ln * far);
        // |origin - vertex| <= |origin - pos| + |pos - vertex|, origins lie within the frame's corners
        smax = sqrtf(Lw * Lw + cx * cx + cy * cy) + far;
        du = RAST_HULL_C * 5.9604645e-8f * smax * l2 / (ln * cosine);
        dv = RAST_HULL_C * 5.9604645e-8f * smax * l1 / (ln * cosine);
        plain = ln > 0.0f && cosine >= RAST_HULL_MIN_COS && du < 2.0f && dv < 2.0f;       // NaN -> box only
      }
      if (plain) {
        const float dw = du + dv + 2.4e-7f;
        const V3 P[3] = {vsub(vsub(v[0], vscale(mk(b4), du)), vscale(mk(c4), dv)),
                         vsub(vadd(v[0], vscale(mk(b4), 1.0f + dw + dv)), vscale(mk(c4), dv)),
                         vadd(vsub(v[0], vscale(mk(b4), du)), vscale(mk(c4), 1.0f + dw + du))};
        float ax[3], ay[3];
        float hx0 = INF, hx1 = -INF, hy0 = INF, hy1 = -INF, dnear = INF;
        bool ok = true;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          float den;
          project(P[k], ax[k], ay[k], den);
          ok = ok && den * dmin > 0.0f;
          dnear = fminf(dnear, fabsf(den));
          hx0 = fminf(hx0, ax[k]); hx1 = fmaxf(hx1, ax[k]); hy0 = fminf(hy0, ay[k]); hy1 = fmaxf(hy1, ay[k]);
        }
        const float big = fmaxf(fmaxf(fabsf(hx0), fabsf(hx1)), fmaxf(fabsf(hy0), fabsf(hy1)));
        // The ray itself is rounded.  Its origin: three roundings at the size of |o|, about 1.5
        // spacings of |o| per coordinate -- thousandths of a sample (the image plane's unit is the
        // sample), and the line pivots about pos, so behind pos that shrinks by depth / L.  Its
        // direction normalize(pos - o): the subtraction and the three divisions round each coordinate
        // by eps relative (the length is a common factor), an angle of about 2 eps, i.e. 2 eps |s| across
        // the ray at the triangle.  RAST_RAY_C = 8 eps |s| bounds the sum with a factor of three; a
        // displacement of that size at depth >= dnear moves the projection by at most
        // |delta| sqrt(L^2 + x^2) / depth (1.3: second order, dnear > 8 delta).
        const float delta = RAST_RAY_C * 5.9604645e-8f * smax;
        ok = ok && dnear > 0.25f * near && dnear > 8.0f * delta && big < 1e6f;
        if (ok) {
          const float sr = 1.3f * delta * sqrtf(Lw * Lw + big * big) / dnear;      // a distance: any direction
          const float sb = 0.02f + 1e-6f * big;                                    // per coordinate
          const float se = sr + sb;
          fx0 = fmaxf(fx0, ceilf(hx0 + cx - se)); fx1 = fminf(fx1, floorf(hx1 + cx + se));
          fy0 = fmaxf(fy0, ceilf(hy0 + cy - se)); fy1 = fminf(fy1, floorf(hy1 + cy + se));
          const float X0 = fmaxf(fx0, 0.0f), Y0 = fmaxf(fy0, 0.0f);
          float en[3][3];
          bool edges = true;
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            const int kb = (k + 1) % 3, kc = (k + 2) % 3;
            float nx = ay[kb] - ay[k], ny = ax[k] - ax[kb];
            const float len = sqrtf(nx * nx + ny * ny);
            nx /= len; ny /= len;
            const float side = nx * (ax[kc] - ax[k]) + ny * (ay[kc] - ay[k]);
            edges = edges && len > 1e-3f && fabsf(side) > 1e-3f;      // edge-on on screen: rectangle only
            if (side > 0.0f) { nx = -nx; ny = -ny; }
            const float rx = (X0 - cx) - ax[k], ry = (Y0 - cy) - ay[k];
            const float off = sr + sb * (fabsf(nx) + fabsf(ny)) + 1e-6f * (fabsf(rx) + fabsf(ry));
            en[k][0] = nx; en[k][1] = ny; en[k][2] = nx * rx + ny * ry - off;
          }
          if (edges) {
            r0 = make_float4(en[0][0], en[0][1], en[0][2], en[1][0]);
            r1 = make_float4(en[1][1], en[1][2], en[2][0], en[2][1]);
            r2 = make_float4(en[2][2], X0, Y0, 1.0f);
          }
        }
      }
      x0 = (int)fmaxf(fx0, 0.0f); x1 = (int)fminf(fx1, (float)(fa.cam.W - 1));
      y0 = (int)fmaxf(fy0, 0.0f); y1 = (int)fminf(fy1, (float)(fa.cam.H - 1));
    }
  }
  if (straddle) {
    const unsigned k = atomicAdd(&rast[4], 1u);
    if (k < RAST_STRADDLE_MAX) rast[8 + k] = s; else rast[1] = 1u;
  }
  const bool any = x1 >= x0 && y1 >= y0;
  const int tx0 = x0 / RAST_TILE, tx1 = x1 / RAST_TILE, ty0 = y0 / RAST_TILE, ty1 = y1 / RAST_TILE;
  unsigned mine = 0, area = 0;
  // A rectangle of at most 32 x 32 samples inside one rank tile stays one item even across the
  // 32-sample tile lines (tiles only exist to cut big triangles into warp-sized work and to give
  // every rank its own part): a third fewer items on C5.
  const int rt = fa.nranks == 1 ? (1 << 30) : RANK_TILE * fa.aliasing;      // rank tile edge in samples
  const bool single = any && x1 - x0 < RAST_TILE && y1 - y0 < RAST_TILE && x0 / rt == x1 / rt && y0 / rt == y1 / rt;
  if (single) {
    if (!*reinterpret_cast<volatile unsigned *>(rast + 1) && rast_owned(fa, tx0, ty0)) {
      mine = 1;
      area = (unsigned)((x1 - x0 + 1) * (y1 - y0 + 1));
    }
  } else if (any) {
    const unsigned tiles = (unsigned)(tx1 - tx0 + 1) * (unsigned)(ty1 - ty0 + 1);
    if (*reinterpret_cast<volatile unsigned *>(rast + 1)) { /* the frame has given up: emit nothing more */ }
    else if (tiles > RAST_TILES_MAX) { rast[1] = 1u; }
    else {
      for (int ty = ty0; ty <= ty1; ++ty)
        for (int tx = tx0; tx <= tx1; ++tx)
          if (rast_owned(fa, tx, ty)) {
            ++mine;
            const int cx0 = max(x0, tx * RAST_TILE), cx1 = min(x1, tx * RAST_TILE + RAST_TILE - 1);
            const int cy0 = max(y0, ty * RAST_TILE), cy1 = min(y1, ty * RAST_TILE + RAST_TILE - 1);
            area += (unsigned)((cx1 - cx0 + 1) * (cy1 - cy0 + 1));
          }
    }
  }
  // one atomic per warp for the item slots and for the area
  unsigned incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  const unsigned total = __shfl_sync(0xffffffffu, incl, 31);
  const unsigned wa = __reduce_add_sync(0xffffffffu, min(area, 1u << 26));
  unsigned base = 0;
  if (lane == 31) {
    if (total) base = atomicAdd(&rast[0], total);
    if (wa) atomicAdd(reinterpret_cast<unsigned long long *>(rast + 2), (unsigned long long)wa);
  }
  base = __shfl_sync(0xffffffffu, base, 31);
  if (!mine) return;
  unsigned k = base + incl - mine;
  if (base + total > fa.rast_cap) { rast[1] = 1u; return; }
  fa.rast_rec[3 * (size_t)s] = r0; fa.rast_rec[3 * (size_t)s + 1] = r1; fa.rast_rec[3 * (size_t)s + 2] = r2;
  if (single) {
    fa.rast_items[k] = make_uint4(s, (unsigned)x0 | ((unsigned)y0 << 16), (unsigned)x1 | ((unsigned)y1 << 16), 0u);
    return;
  }
  for (int ty = ty0; ty <= ty1; ++ty)
    for (int tx = tx0; tx <= tx1; ++tx)
      if (rast_owned(fa, tx, ty)) {
        const int cx0 = max(x0, tx * RAST_TILE), cx1 = min(x1, tx * RAST_TILE + RAST_TILE - 1);
        const int cy0 = max(y0, ty * RAST_TILE), cy1 = min(y1, ty * RAST_TILE + RAST_TILE - 1);
        fa.rast_items[k++] = make_uint4(s, (unsigned)cx0 | ((unsigned)cy0 << 16), (unsigned)cx1 | ((unsigned)cy1 << 16), 0u);
      }
}

// k_rast_items: one warp per item.  Cheap stage: the rectangle's samples against the edge record,
// 32 at a time; the survivors go to the warp's queue in shared memory.  Whenever the queue holds
// 32, every lane takes one (triangle, sample) pair and runs the exact test on it -- full warps for
// the expensive stage whatever the rectangles look like.  Queue entries carry their triangle, so
// the queue simply runs on across items.
constexpr int RAST_WARPS = 8;
// k / w == (k * c_rast_rw[w]) >> 16 for k < 1024, w <= 32: 65536 / w + 1 (the division costs 20 instructions per item)
__constant__ unsigned c_rast_rw[33] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2049};

__device__ __forceinline__ void rast_test(const DevScene &sc, const FrameArgs &fa, const uint2 q) {
  const int px = (int)(q.y & 0xffffu), py = (int)(q.y >> 16);
  const float4 a4 = __ldg(sc.v0 + q.x), b4 = __ldg(sc.e1 + q.x), c4 = __ldg(sc.e2 + q.x);
  V3 o, d;
  primary_ray(fa.cam, px, py, o, d);
  float dist;
  if (tri_candidate(o, d, a4, b4, c4, dist)) {
    const unsigned long long key = ((unsigned long long)__float_as_uint(dist) << 32) | __float_as_uint(a4.w);
    unsigned long long *dst = fa.prim + (size_t)py * fa.cam.W + px;
    if (key < *dst) atomicMin(dst, key);
  }
}

__global__ void __launch_bounds__(32 * RAST_WARPS) k_rast_items(const DevScene sc, const FrameArgs fa) {
  __shared__ uint2 queue[RAST_WARPS][64];
  if (!rast_active(fa)) return;
  const unsigned n = fa.rast[0];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const unsigned warps = gridDim.x * RAST_WARPS;
  uint2 *Q = queue[wid];
  int qn = 0;                                   // warp-uniform
  for (unsigned i = blockIdx.x * RAST_WARPS + wid; i < n; i += warps) {
    const uint4 it = __ldg(fa.rast_items + i);
    const int x0 = (int)(it.y & 0xffffu), y0 = (int)(it.y >> 16), x1 = (int)(it.z & 0xffffu), y1 = (int)(it.z >> 16);
    const int w = x1 - x0 + 1, cnt = w * (y1 - y0 + 1);
    const unsigned rw = c_rast_rw[w];
    const float4 r0 = __ldg(fa.rast_rec + 3 * (size_t)it.x), r1 = __ldg(fa.rast_rec + 3 * (size_t)it.x + 1),
                 r2 = __ldg(fa.rast_rec + 3 * (size_t)it.x + 2);
    const bool edges = r2.w != 0.0f;
    const float ox = (float)x0 - r2.y, oy = (float)y0 - r2.z;     // item corner relative to the record's origin
    for (int k0 = 0; k0 < cnt; k0 += 32) {
      const int k = k0 + lane;
      const int ry = (int)(((unsigned)k * rw) >> 16), rx = k - ry * w;
      bool in = k < cnt;
      if (edges) {
        const float dx = ox + (float)rx, dy = oy + (float)ry;
        in = in && r0.x * dx + r0.y * dy + r0.z <= 0.0f && r0.w * dx + r1.x * dy + r1.y <= 0.0f &&
             r1.z * dx + r1.w * dy + r2.x <= 0.0f;
      }
      const unsigned mask = __ballot_sync(0xffffffffu, in);
      if (in) Q[qn + __popc(mask & lt)] = make_uint2(it.x, (unsigned)(x0 + rx) | ((unsigned)(y0 + ry) << 16));
      qn += __popc(mask);
      __syncwarp();
      if (qn >= 32) {
        qn -= 32;
        const uint2 q = Q[qn + lane];
        __syncwarp();
        rast_test(sc, fa, q);
      }
    }
  }
  if (lane < qn) rast_test(sc, fa, Q[lane]);
}

// The primary hit of sample (px, py) from the record (+ the straddlers, tested like any leaf).
__device__ __forceinline__ bool prim_lookup(const DevScene &sc, const FrameArgs &fa, int px, int py, const V3 o,
                                            const V3 d, Best &best) {
  const unsigned long long key = fa.prim[(size_t)py * fa.cam.W + px];
  best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
  if (key != ~0ull) {
    best.dist = __uint_as_float((unsigned)(key >> 32));
    best.id = (int)(unsigned)key;
    best.slot = (int)__ldg(sc.inv_order + best.id);
  }
  const unsigned ns = fa.rast[4];
  for (unsigned k = 0; k < ns; ++k) consider<false>(sc, o, d, (int)fa.rast[8 + k], best);
  return best.slot >= 0;
}

// rtb_scene_hit_ids through the same records (tests)
__global__ void __launch_bounds__(256) k_prim_ids(const DevScene sc, const FrameArgs fa, int32_t *tri, int32_t *obj) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)fa.cam.W * fa.cam.H) return;
  const int px = (int)(i % fa.cam.W), py = (int)(i / fa.cam.W);
  if (!rast_active(fa)) { tri[i] = -2; if (obj) obj[i] = -2; return; }      // this frame would walk the tree
  V3 o, d;
  primary_ray(fa.cam, px, py, o, d);
  Best b;
  const bool hit = prim_lookup(sc, fa, px, py, o, d, b);
  tri[i] = hit ? b.id : -1;
  if (obj) obj[i] = hit ? __float_as_int(__ldg(sc.e1 + b.slot).w) : -1;
}

}  // namespace
