// Shadow rays by projection (included by trace.cu).
//
// All shadow rays of a directional light are parallel, all shadow rays of a point light lie on
// lines through the light's position (gpu/light.cu:68-120: direction -v, resp. v - P), and a
// shadow ray only asks "any candidate on the ray" (gpu/light.cu:29-36).  Seen along the light a
// ray is a POINT (x, y), and the set of rays the ROUNDED ray/triangle test of the reference can
// accept for one triangle is a region of that plane bounded by three offset lines: region.cuh
// (dir_lines / point_lines) derives it from the operation sequence of gpu/hit.cu:20-37 -- a bound,
// not a heuristic; tests/test_region_cpu.py replays it in float32 on the host, triangles seen
// edge-on from the light included.  Per light and triangle k_lg_records stores that region as a
// float record (region::Record: three lines, rounded outwards) plus its bounding box; a uniform
// grid in the light's projection, built once per scene (lights are part of the scene), lists for
// every cell the triangles whose region can touch it (k_lg_scatter, from the same float records).
// A ray looks up one cell, checks the entries' rectangles (four 7-bit bytes inside the cell), then
// the survivors' records, and runs the exact per-ray test on what is left: ~10 byte compares, ~5
// records and ~2 triangle tests instead of ~35 node visits on the 10 M-triangle terrain.
//
// Triangles whose region covers too many cells are tested by every ray (a short list); too many of
// those, or a list volume out of proportion, and the light keeps the tree walk.  Rays whose point
// falls outside the grid, or whose origin lies outside the scene's box (the bound on |origin -
// vertex| the regions were built for), walk the tree as well.
#pragma once

namespace {

constexpr unsigned LG_CELLS_MAX = 1u << 15;  // per triangle (a region that is a line across a 4 400-cell grid has ~6 000)
constexpr float LG_PERSP_LIMIT = 4.0f;       // point lights: the grid spans |x|, |y| <= 4 on the plane at distance 1
constexpr float LG_Q = 127.0f;               // rectangle bytes: 1 / 127 of a cell
constexpr float LG_DILATE = 2.0e-3f;         // cells: a ray's cell index is computed in float (error < 1e-3 cells at 30 000 cells)
constexpr float LG_BYTE_SLACK = 0.75f;       // bytes: rounding of the ray's position inside its cell (< 0.2 at 30 000 cells)

// build-time view of a light: the frames region.cuh works with (doubles) and the domain the regions are clipped to
struct LgBuild {
  region::DirFrame df;
  region::PointFrame pf;
  double dom[4];               // x0, x1, y0, y1 in the light's projection
};

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

__device__ __forceinline__ float round_down(double v) {
  float f = (float)v;
  if ((double)f > v) f = nextafterf(f, -__builtin_huge_valf());
  return f;
}
__device__ __forceinline__ float round_up(double v) {
  float f = (float)v;
  if ((double)f < v) f = nextafterf(f, __builtin_huge_valf());
  return f;
}

// rec[4 slot + 0..2]: the region's record (region::Record, origin = the box's centre), rec[4 slot + 3]:
// its bounding box (x0, x1, y0, y1), x0 > x1: no ray of this light can be accepted for the triangle.
__global__ void __launch_bounds__(128) k_lg_records(const DevScene sc, const LightGrid g, const LgBuild lb, float4 *rec) {
  const uint32_t s = blockIdx.x * 128 + threadIdx.x;
  if (s >= sc.n_tris) return;
  const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
  const float v0[3] = {a4.x, a4.y, a4.z}, e1[3] = {b4.x, b4.y, b4.z}, e2[3] = {c4.x, c4.y, c4.z};
  region::Lines L;
  region::Region R;
  R.kind = 0;
  const bool any = g.kind == 1 ? region::dir_lines(lb.df, v0, e1, e2, L) : region::point_lines(lb.pf, v0, e1, e2, L);
  if (any) region::build(L, lb.dom[0], lb.dom[1], lb.dom[2], lb.dom[3], R);
  float4 r0 = make_float4(0, 0, 0, 0), r1 = r0, r2 = r0, box = make_float4(1.0f, 0.0f, 1.0f, 0.0f);
  if (R.kind) {
    box = make_float4(round_down(R.x0), round_up(R.x1), round_down(R.y0), round_up(R.y1));
    const float X0 = 0.5f * box.x + 0.5f * box.y, Y0 = 0.5f * box.z + 0.5f * box.w;
    region::Record rc;
    region::make_record(R, (double)X0, (double)Y0, ((double)box.y - (double)box.x) + ((double)box.w - (double)box.z), rc);
    r0 = make_float4(rc.v[0], rc.v[1], rc.v[2], rc.v[3]);
    r1 = make_float4(rc.v[4], rc.v[5], rc.v[6], rc.v[7]);
    r2 = make_float4(rc.v[8], X0, Y0, rc.v[11]);
  }
  rec[4 * (size_t)s] = r0; rec[4 * (size_t)s + 1] = r1; rec[4 * (size_t)s + 2] = r2; rec[4 * (size_t)s + 3] = box;
}

// bounds of the boxes of every stride-th triangle: out[0..3] = min x, max x, min y, max y (ordered
// uint).  The grid does not have to cover every region: k_lg_scatter clips to the grid and a ray
// whose point is off the grid walks the tree, so big scenes look at a sample.
__global__ void __launch_bounds__(256) k_lg_bounds(const float4 *rec, uint32_t n, uint32_t stride, float huge, uint32_t *out) {
  const uint32_t s = (blockIdx.x * 256 + threadIdx.x) * stride;
  const float INFF = __builtin_huge_valf();
  float x0 = INFF, x1 = -INFF, y0 = INFF, y1 = -INFF;
  if (s < n) {
    const float4 b = __ldg(rec + 4 * (size_t)s + 3);
    // regions that are lines across the whole domain say nothing about where the geometry is
    if (b.x <= b.y && b.y - b.x < huge && b.w - b.z < huge) { x0 = b.x; x1 = b.y; y0 = b.z; y1 = b.w; }
  }
  const uint32_t m0 = __reduce_min_sync(0xffffffffu, f2ord(x0)), m1 = __reduce_max_sync(0xffffffffu, f2ord(x1));
  const uint32_t m2 = __reduce_min_sync(0xffffffffu, f2ord(y0)), m3 = __reduce_max_sync(0xffffffffu, f2ord(y1));
  if ((threadIdx.x & 31) == 0) {
    atomicMin(out + 0, m0); atomicMax(out + 1, m1); atomicMin(out + 2, m2); atomicMax(out + 3, m3);
  }
}

// An entry is (slot, the region's rectangle inside the cell in 1/127 of the cell, rounded outwards),
// the rectangle as four 7-bit bytes x0, 127 - x1, y0, 127 - y1: a ray turns its point into the bytes
// floor(qx), 127 - ceil(qx), floor(qy), 127 - ceil(qy) with bit 7 set in each, and "inside" is
// "every byte of the entry <= the ray's byte", i.e. no byte of (ray - entry) loses its bit 7: one
// subtraction, one AND, one compare on 4 sequential bytes, before the ray touches the triangle's
// record or the triangle itself (scattered 16-byte loads and the exact test).
__device__ __forceinline__ uint32_t lg_pack(const LightGrid &g, int i, int j, float x0, float x1, float y0, float y1) {
  const float cx = g.x0 + (float)i * g.cell, cy = g.y0 + (float)j * g.cell;
  auto q = [&](float v, float c, bool up) {
    float t = (v - c) * g.inv_cell * LG_Q;
    t = up ? ceilf(t + LG_BYTE_SLACK) : floorf(t - LG_BYTE_SLACK);
    return (uint32_t)fminf(fmaxf(t, 0.0f), LG_Q);
  };
  return q(x0, cx, false) | ((127u - q(x1, cx, true)) << 8) | (q(y0, cy, false) << 16) | ((127u - q(y1, cy, true)) << 24);
}

// The cells of triangle slot s, row by row: f(i, j, x0, x1, y0, y1) for every cell the region can touch
// (region::record_band on the float record the rays are tested against, dilated by the rounding of a
// ray's own cell index).  Returns the number of cells, or ~0u when it exceeds LG_CELLS_MAX.
// (jfirst, jstep): the rows j0 + jfirst, j0 + jfirst + jstep, ... only -- k_lg_big splits a triangle's rows over a CTA.
template <typename F>
__device__ __forceinline__ unsigned lg_cells(const LightGrid &g, const float4 *rec, uint32_t s, bool count_only, F f,
                                             int jfirst = 0, int jstep = 1, unsigned limit = LG_CELLS_MAX, int *rows = nullptr) {
  const float4 box = __ldg(rec + 4 * (size_t)s + 3);
  if (!(box.x <= box.y)) return 0;
  const float4 r0 = __ldg(rec + 4 * (size_t)s), r1 = __ldg(rec + 4 * (size_t)s + 1), r2 = __ldg(rec + 4 * (size_t)s + 2);
  const float X0 = r2.y, Y0 = r2.z;
  const float dil = LG_DILATE * g.cell;
  const float fj0 = floorf((box.z - dil - g.y0) * g.inv_cell), fj1 = floorf((box.w + dil - g.y0) * g.inv_cell);
  if (fj1 < 0.0f || fj0 > (float)(g.ny - 1)) return 0;
  const int j0 = (int)fmaxf(fj0, 0.0f), j1 = (int)fminf(fj1, (float)(g.ny - 1));
  const float xext = fmaxf(fabsf(box.x - X0), fabsf(box.y - X0)) + 2.0f * g.cell;
  unsigned total = 0;
  if (rows) *rows = j1 - j0 + 1;
  for (int j = j0 + jfirst; j <= j1; j += jstep) {
    const float ya = fmaxf(g.y0 + (float)j * g.cell - dil, box.z - dil), yb = fminf(g.y0 + (float)(j + 1) * g.cell + dil, box.w + dil);
    float lo, hi;
    if (!region::record_band(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w, ya - Y0, yb - Y0, xext, lo, hi)) continue;
    const float xa = fmaxf(X0 + lo, box.x) - dil, xb = fminf(X0 + hi, box.y) + dil;
    if (!(xa <= xb)) continue;
    const float fi0 = floorf((xa - g.x0) * g.inv_cell), fi1 = floorf((xb - g.x0) * g.inv_cell);
    if (fi1 < 0.0f || fi0 > (float)(g.nx - 1)) continue;
    const int i0 = (int)fmaxf(fi0, 0.0f), i1 = (int)fminf(fi1, (float)(g.nx - 1));
    total += (unsigned)(i1 - i0 + 1);
    if (total > limit) return ~0u;
    if (!count_only)
      for (int i = i0; i <= i1; ++i) f(i, j, xa, xb, ya, yb);
  }
  return total;
}

// FILL = false: counts[cell] += 1 per covered cell, state[0] = number of large triangles (listed in
// state[2..]), state[1] != 0: give up.  FILL = true: entries[cell_start[cell] + cursor[cell]++] = (slot, rectangle).
constexpr int LG_BIG_ROWS = 64;              // a triangle over more grid rows is listed for k_lg_big
constexpr unsigned LG_BIG_CAP = 1u << 18;    // slots on that list (state[32..])
constexpr int LG_STATE_WORDS = 32;

template <bool FILL>
__global__ void __launch_bounds__(256) k_lg_scatter(const DevScene sc, const LightGrid g, const float4 *rec, uint32_t *counts,
                                                    uint32_t *state, const uint32_t *cell_start, uint2 *entries) {
  const uint32_t s = blockIdx.x * 256 + threadIdx.x;
  if (s >= sc.n_tris) return;
  // triangles over many rows (regions that are lines across the grid): one CTA each, rows in parallel (k_lg_big)
  {
    const float4 box = __ldg(rec + 4 * (size_t)s + 3);
    if (box.x <= box.y && (box.w - box.z) * g.inv_cell > (float)LG_BIG_ROWS) {
      if (FILL) return;                               // k_lg_big<true> fills them
      const unsigned k = atomicAdd(state + 24, 1u);
      if (k < LG_BIG_CAP) state[LG_STATE_WORDS + k] = s; else state[1] = 1u;      // list full: the light keeps the tree walk
      return;
    }
  }
  const unsigned cells = lg_cells(g, rec, s, true, [](int, int, float, float, float, float) {});
  if (cells == 0) return;
  if (cells == ~0u) {
    if (!FILL) {
      const unsigned k = atomicAdd(state, 1u);
      if (k < (unsigned)LG_LARGE_MAX) state[2 + k] = s; else state[1] = 1u;
    }
    return;
  }
  lg_cells(g, rec, s, false, [&](int i, int j, float xa, float xb, float ya, float yb) {
    const size_t c = (size_t)j * g.nx + i;
    const uint32_t k = atomicAdd(counts + c, 1u);
    if (FILL) entries[cell_start[c] + k] = make_uint2(s, lg_pack(g, i, j, xa, xb, ya, yb));
  });
}

// The listed triangles: one CTA per triangle, thread t takes the grid rows j0 + t, j0 + t + 128, ...
// Pass 1 (all threads): cells per row -> the CTA's total decides "large" (more than LG_CELLS_MAX cells:
// tested by every ray instead); pass 2: count or fill, exactly the cells k_lg_scatter would produce.
template <bool FILL>
__global__ void __launch_bounds__(128) k_lg_big(const DevScene sc, const LightGrid g, const float4 *rec, uint32_t *counts,
                                                uint32_t *state, const uint32_t *cell_start, uint2 *entries) {
  __shared__ unsigned s_total;
  const unsigned n_big = min(state[24], LG_BIG_CAP);
  for (unsigned bi = blockIdx.x; bi < n_big; bi += gridDim.x) {
    const uint32_t s = state[LG_STATE_WORDS + bi];
    if (threadIdx.x == 0) s_total = 0;
    __syncthreads();
    const unsigned mine = lg_cells(g, rec, s, true, [](int, int, float, float, float, float) {}, (int)threadIdx.x, 128, 0xfffffff0u);
    if (mine) atomicAdd(&s_total, mine);
    __syncthreads();
    const unsigned total = s_total;
    __syncthreads();
    if (total == 0) continue;
    if (total > LG_CELLS_MAX) {
      if (!FILL && threadIdx.x == 0) {
        const unsigned k = atomicAdd(state, 1u);
        if (k < (unsigned)LG_LARGE_MAX) state[2 + k] = s; else state[1] = 1u;
      }
      continue;
    }
    lg_cells(g, rec, s, false, [&](int i, int j, float xa, float xb, float ya, float yb) {
      const size_t c = (size_t)j * g.nx + i;
      const uint32_t k = atomicAdd(counts + c, 1u);
      if (FILL) entries[cell_start[c] + k] = make_uint2(s, lg_pack(g, i, j, xa, xb, ya, yb));
    }, (int)threadIdx.x, 128, 0xfffffff0u);
  }
}

// statistics of a grid (RTB_BUILD_TRACE): out[0] longest cell list, out[1] cells with more than 64 entries,
// out[2] sum of squares of the list lengths (a random entry's list length is that / the entry count)
__global__ void __launch_bounds__(256) k_lg_stats(const uint32_t *cell_start, size_t cells, unsigned long long *out) {
  unsigned mx = 0, big = 0;
  unsigned long long sq = 0;
  for (size_t c = (size_t)blockIdx.x * 256 + threadIdx.x; c < cells; c += (size_t)gridDim.x * 256) {
    const unsigned n = cell_start[c + 1] - cell_start[c];
    mx = max(mx, n); big += n > 64u; sq += (unsigned long long)n * n;
  }
  mx = __reduce_max_sync(0xffffffffu, mx);
  big = __reduce_add_sync(0xffffffffu, big);
#pragma unroll
  for (int d = 16; d; d >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, d);
  if ((threadIdx.x & 31) == 0) { atomicMax(out, (unsigned long long)mx); atomicAdd(out + 1, (unsigned long long)big); atomicAdd(out + 2, sq); }
}

// ---- the lookup ----
// 1: some candidate on the ray, 0: none, < 0: this ray has to walk the tree (-2: its origin is outside the box the
// regions were bounded for, -3: its point is off the grid, -4: a point light's projection is not usable for it).
//
// Three loops per chunk of 32 entries, because the lanes of a warp walk different lists in lockstep and a
// single loop made the warp pay the most expensive branch in almost every iteration: (1) the entries'
// rectangle words against the point, (2) the records of the survivors, (3) the exact test on what is left,
// leaving at the first accepted candidate.  A warp-cooperative variant (survivors of all 32 rays through
// shared-memory queues so that stages 2 and 3 run 32 at a time) was built and measured: lanes per
// instruction 27 instead of 23, but 11.9 G instead of 8.3 G warp-instructions and 18.6 instead of 11.9 ms
// on C5 -- a record test is ~35 instructions, the queue traffic around it costs more than the idle lanes
// (profiles/r2_ncu_kframe_coop.txt).
__device__ __forceinline__ int lg_any(const DevScene &sc, const LightGrid &g, const V3 o, const V3 d, int &hit) {
  if (!(o.x >= g.bmin.x && o.x <= g.bmax.x && o.y >= g.bmin.y && o.y <= g.bmax.y && o.z >= g.bmin.z && o.z <= g.bmax.z)) return -2;
  float x, y, den;
  lg_project(g, o, x, y, den);
  const float fx = (x - g.x0) * g.inv_cell, fy = (y - g.y0) * g.inv_cell;
  const float fi = floorf(fx), fj = floorf(fy);
  if (!(fi >= 0.0f && fj >= 0.0f && fi <= (float)(g.nx - 1) && fj <= (float)(g.ny - 1))) return -3;      // NaN too
  if (g.kind == 2 && !(fabsf(den) > 1e-6f * (fabsf(o.x - g.org.x) + fabsf(o.y - g.org.y) + fabsf(o.z - g.org.z)) &&
                       fabsf(x) <= LG_PERSP_LIMIT && fabsf(y) <= LG_PERSP_LIMIT)) return -4;
  Best b;
  b.dist = INF; b.id = 0x7fffffff; b.slot = -1;
  for (uint32_t k = 0; k < g.n_large; ++k)
    if (consider<true>(sc, o, d, (int)g.large[k], b)) { hit = (int)g.large[k]; return 1; }
  const size_t c = (size_t)(int)fj * g.nx + (int)fi;
  const uint32_t e0 = __ldg(g.cell_start + c), e1 = __ldg(g.cell_start + c + 1);
  // the ray's point inside its cell as the four bytes the entries compare with (lg_pack)
  const float qx = (fx - fi) * LG_Q, qy = (fy - fj) * LG_Q;
  auto byte = [](float v) { return (uint32_t)fminf(fmaxf(v, 0.0f), LG_Q); };
  const uint32_t P = (byte(floorf(qx)) | ((127u - byte(ceilf(qx))) << 8) | (byte(floorf(qy)) << 16) | ((127u - byte(ceilf(qy))) << 24)) | 0x80808080u;
#ifdef RTB_LG_WHILE_WHILE      /* A/B only: measured 16.6 ms against 13.3 ms for k_frame on C5 */
  // "while-while": every lane runs through its list (rectangle words of 32 entries at a time, records of the
  // survivors) until it holds ONE candidate, then the lanes that hold one run the exact test together
  uint32_t base = e0, cur = e0, m = 0;
  for (;;) {
    int cand = -1;
    for (;;) {
      if (!m) {
        if (base >= e1) break;
        const uint32_t cnt = min(32u, e1 - base);
#pragma unroll 4
        for (uint32_t k = 0; k < cnt; ++k) {
          const uint32_t r = __ldg(&g.entries[base + k].y);
          if (((P - r) & 0x80808080u) == 0x80808080u) m |= 1u << k;
        }
        cur = base; base += 32;
        continue;
      }
      const uint32_t k = (uint32_t)__ffs((int)m) - 1u;
      m &= m - 1u;
      const uint32_t s = __ldg(&g.entries[cur + k].x);
      const float4 r0 = __ldg(g.rec + 4 * (size_t)s), r1 = __ldg(g.rec + 4 * (size_t)s + 1), r2 = __ldg(g.rec + 4 * (size_t)s + 2);
      if (region::record_inside(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w, x - r2.y, y - r2.z)) { cand = (int)s; break; }
    }
    if (cand < 0) break;
    if (consider<true>(sc, o, d, cand, b)) { hit = cand; return 1; }
  }
#else
  for (uint32_t base = e0; base < e1; base += 32) {
    const uint32_t cnt = min(32u, e1 - base);
    uint32_t mask = 0;
#pragma unroll 4
    for (uint32_t k = 0; k < cnt; ++k) {
      const uint32_t r = __ldg(&g.entries[base + k].y);
      if (((P - r) & 0x80808080u) == 0x80808080u) mask |= 1u << k;
    }
    uint32_t m = mask;
    while (m) {
      const uint32_t k = (uint32_t)__ffs((int)m) - 1u;
      m &= m - 1u;
      const size_t s = (size_t)__ldg(&g.entries[base + k].x);
      const float4 r0 = __ldg(g.rec + 4 * s), r1 = __ldg(g.rec + 4 * s + 1), r2 = __ldg(g.rec + 4 * s + 2);
      if (!region::record_inside(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w, x - r2.y, y - r2.z)) mask &= ~(1u << k);
    }
    while (mask) {
      const uint32_t k = (uint32_t)__ffs((int)mask) - 1u;
      mask &= mask - 1u;
      const int cand = (int)__ldg(&g.entries[base + k].x);
      if (consider<true>(sc, o, d, cand, b)) { hit = cand; return 1; }
    }
  }
#endif
  return 0;
}

}  // namespace
