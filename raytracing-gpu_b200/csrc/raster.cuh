// Primary visibility by projection (included by trace.cu).
//
// Every primary ray passes through the camera position (gpu/raytracer.cu:97-105: origin on the
// image plane, direction towards `pos`), and sample (px, py) is a regular grid on that plane.
// So "which rays can the rounded Moeller-Trumbore test accept for this triangle" is a screen
// rectangle: the projection through `pos` of the triangle's culling box, the very box the
// hierarchy culls with (vertex bounds + needle padding of pad.cuh + the ray-time margin of
// trace_tree).  One pass over the triangles emits (triangle, 32x32-sample tile) items, one warp
// per item runs the exact per-ray test of tri_candidate on the rectangle's samples, and
// atomicMin on (new_dist bits << 32 | triangle id) keeps the reference's winner: lexicographic
// minimum of (new_dist, id), gpu/hit.cu:66-76.  k_frame then starts every sample from that record
// instead of walking the tree for its first ray; shadow and bounce rays still walk the tree.
//
// Same candidates as the tree: a ray the tree would test against a triangle pierces that
// triangle's inflated box, and a line through `pos` that pierces a convex box projects inside
// the hull of the box's projected corners, as long as the box lies entirely on one side of the
// plane through `pos` parallel to the image plane.  Boxes that straddle that plane (their
// projection is unbounded) go to a short list every sample tests directly; more than
// RAST_STRADDLE_MAX of them, a triangle that spans more than RAST_TILES_MAX tiles, more items
// than the list holds or more rectangle area than RAST_AREA_FACTOR x the frame make the frame
// fall back to the tree for its primary rays (flag in device memory, no host round trip).
#pragma once

namespace {

constexpr int RAST_TILE = 32;                 // samples
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

__global__ void __launch_bounds__(256) k_rast_setup(const DevScene sc, const FrameArgs fa) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  const int lane = threadIdx.x & 31;
  unsigned *rast = fa.rast;
  const float INF = __builtin_huge_valf();
  int x0 = 0, y0 = 0, x1 = -1, y1 = -1;        // empty rectangle
  bool straddle = false;
  if (s < sc.n_tris) {
    const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
    const V3 v[3] = {mk(a4), vadd(mk(a4), mk(b4)), vadd(mk(a4), mk(c4))};
    V3 lo = mk(fminf(v[0].x, fminf(v[1].x, v[2].x)), fminf(v[0].y, fminf(v[1].y, v[2].y)), fminf(v[0].z, fminf(v[1].z, v[2].z)));
    V3 hi = mk(fmaxf(v[0].x, fmaxf(v[1].x, v[2].x)), fmaxf(v[0].y, fmaxf(v[1].y, v[2].y)), fmaxf(v[0].z, fmaxf(v[1].z, v[2].z)));
    sliver_pad(v, sc.pad_unit, lo, hi);
    // ray-time margin of trace_tree at the largest origin of this frame, plus the rounding of
    // vertex0 + edge (the build boxes come from the stored vertices themselves)
    const float e = fa.rast_eps;
    lo = mk(lo.x - e, lo.y - e, lo.z - e);
    hi = mk(hi.x + e, hi.y + e, hi.z + e);
    const V3 u = mk(fa.cam.u), vv = mk(fa.cam.v), w = mk(fa.cam.w), pos = mk(fa.cam.pos);
    const float ww = vdot(w, w), uv = vdot(u, vv);
    const float det = 1.0f - uv * uv;                      // Gram determinant of the unit vectors u, v
    float dmin = INF, dmax = -INF, xmin = INF, xmax = -INF, ymin = INF, ymax = -INF;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const V3 X = mk((k & 1) ? hi.x : lo.x, (k & 2) ? hi.y : lo.y, (k & 4) ? hi.z : lo.z);
      const V3 r = vsub(pos, X);
      const float den = vdot(r, w);
      dmin = fminf(dmin, den); dmax = fmaxf(dmax, den);
      const float t = fa.cam.L * ww / den;
      const V3 q = vsub(vscale(r, t), vscale(w, fa.cam.L));     // offset from the plane's centre C
      const float qu = vdot(q, u), qv = vdot(q, vv);
      const float x = (qu - uv * qv) / det, y = (qv - uv * qu) / det;
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
      const float sx = 0.02f + 4e-7f * fmaxf(fabsf(xmin), fabsf(xmax)), sy = 0.02f + 4e-7f * fmaxf(fabsf(ymin), fabsf(ymax));
      const float fx0 = ceilf(xmin + cx - sx), fx1 = floorf(xmax + cx + sx);
      const float fy0 = ceilf(ymin + cy - sy), fy1 = floorf(ymax + cy + sy);
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
  if (any) {
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
  for (int ty = ty0; ty <= ty1; ++ty)
    for (int tx = tx0; tx <= tx1; ++tx)
      if (rast_owned(fa, tx, ty)) {
        const int cx0 = max(x0, tx * RAST_TILE), cx1 = min(x1, tx * RAST_TILE + RAST_TILE - 1);
        const int cy0 = max(y0, ty * RAST_TILE), cy1 = min(y1, ty * RAST_TILE + RAST_TILE - 1);
        fa.rast_items[k++] = make_uint4(s, (unsigned)cx0 | ((unsigned)cy0 << 16), (unsigned)cx1 | ((unsigned)cy1 << 16), 0u);
      }
}

__global__ void __launch_bounds__(256) k_rast_items(const DevScene sc, const FrameArgs fa) {
  if (!rast_active(fa)) return;
  const unsigned n = fa.rast[0];
  const int lane = threadIdx.x & 31;
  const unsigned warps = gridDim.x * (256 / 32);
  for (unsigned i = blockIdx.x * (256 / 32) + (threadIdx.x >> 5); i < n; i += warps) {
    const uint4 it = __ldg(fa.rast_items + i);
    const int x0 = (int)(it.y & 0xffffu), y0 = (int)(it.y >> 16), x1 = (int)(it.z & 0xffffu), y1 = (int)(it.z >> 16);
    const int w = x1 - x0 + 1, cnt = w * (y1 - y0 + 1);
    const float4 a4 = __ldg(sc.v0 + it.x), b4 = __ldg(sc.e1 + it.x), c4 = __ldg(sc.e2 + it.x);
    const unsigned id = __float_as_uint(a4.w);
    for (int k = lane; k < cnt; k += 32) {
      const int py = y0 + k / w, px = x0 + k % w;
      V3 o, d;
      primary_ray(fa.cam, px, py, o, d);
      float dist;
      if (tri_candidate(o, d, a4, b4, c4, dist)) {
        const unsigned long long key = ((unsigned long long)__float_as_uint(dist) << 32) | id;
        unsigned long long *dst = fa.prim + (size_t)py * fa.cam.W + px;
        if (key < *dst) atomicMin(dst, key);
      }
    }
  }
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
