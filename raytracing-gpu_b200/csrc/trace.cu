// Frame kernels: persistent warps pull 8x4-pixel tiles, trace every sample of the tile
// (primary ray, shadow rays, reflection chain) and resolve the SSAA box filter in shared
// memory, so the high-resolution framebuffer and the downscale pass of the reference
// (gpu/raytracer.cu:87-129 raytrace, :49-85 downscale, :177-253 render) never exist.
//
// Semantics follow SURVEY Appendix A exactly:
//   primary rays      gpu/raytracer.cu:91-110
//   bounce loop       gpu/raytracer.cu:107-122, trace :32-47, ray_bounce gpu/ray.cu:16-25
//   ray/triangle      gpu/hit.cu:8-48     nearest hit  gpu/hit.cu:50-78, :84-114
//   shadow rays       gpu/light.cu:29-36, gpu/hit.cu:241-253 (any candidate on the ray)
//   shading           gpu/light.cu:12-27, :50-127      colours  gpu/colors.cu:3-47
// Nearest hit = lexicographic minimum of (new_dist, global triangle id), which equals
// the reference's "first strict minimum in iteration order" and does not depend on the
// traversal order.
//
// What culls, and how far each cull is proven:
//   * first ray of a sample (raster.cuh) and shadow rays of scenes with light grids (lightgrid.cuh):
//     accept regions of region.cuh, a bound on where the rounded test can accept -- no candidate of the
//     reference's exhaustive search is dropped (0 of 74.6 M ids differ at 3840x2160x3 on all 20 scenes
//     and on the 10 M-triangle terrain's samples, tests/test_gpu_baseline_configs.py);
//   * every ray that walks the tree (bounce rays, shadow rays without a grid or off their grid, the
//     counting frame): boxes padded by pad.cuh's needle terms and the ray-time margin.  That padding is
//     a heuristic, not a bound: exact on the 20 fixtures at native resolution, but it can drop a
//     candidate that the rounded arithmetic accepts far outside a triangle seen edge-on (measured:
//     <= 3 samples per random needle soup, tests/test_gpu_random_scenes.py).
// Deviation kept on purpose: the reference treats a hit record whose interpolated normal is exactly
// (0,0,0) as a miss (gpu/hit.cu:100, gpu/raytracer.cu:36); this tracer does not (unit vertex normals
// produce it only where they cancel exactly, e.g. n0 = -n1 at u = 0.5); DESIGN.md section 4.2.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "rtb_internal.h"
#include "rtb_math.cuh"
#include "pad.cuh"
#include "scene_store.h"

namespace rtb {

namespace {

constexpr float INF = __builtin_huge_valf();

struct Counters {
  unsigned primary, shadow, bounce, nodes, tris;
  // shadow cache (-DRTB_SHADOW_CACHE, A/B only): the slot of the triangle that occluded this lane's last
  // shadow ray towards an even / odd light, tested first for the next sample.  Slower on C5.
  int occ0 = -1, occ1 = -1;
  unsigned sample = 0, item = 0;      // the sample and work item being shaded (stray shadow rays are listed with them)
};

struct Best {
  float dist;   // new_dist of the winner, INF when none
  int id;       // global triangle id (host order)
  int slot;     // position in the leaf-ordered streams
};

// gpu/hit.cu:20-41 + :66-67 acceptance of one triangle; dist = new_dist of hit.cu:66.
__device__ __forceinline__ bool tri_candidate(const V3 o, const V3 d, const float4 a4, const float4 b4,
                                              const float4 c4, float &dist) {
  const float EPS = 0.0000001f;
  const V3 v0 = mk(a4), e1 = mk(b4), e2 = mk(c4);
  const V3 h = vcross(d, e2);
  const float a = vdot(e1, h);
  if (a > -EPS && a < EPS) return false;
  const float f = 1 / a;
  const V3 s = vsub(o, v0);
  const float u = f * vdot(s, h);
  if (u < 0.0f || u > 1.0f) return false;
  const V3 q = vcross(s, e1);
  const float v = f * vdot(d, q);
  if (v < 0.0f || u + v > 1.0f) return false;
  const float t = f * vdot(e2, q);
  if (!(t > EPS)) return false;
  const V3 out = vadd(o, vscale(vnorm(d), t * vlen(d)));
  dist = vlen(vsub(out, o));
  return dist > 0.01f;
}

// Hit record of the winner: point (hit.cu:39-41) and interpolated normal (hit.cu:42-44).
__device__ __forceinline__ void hit_record(const DevScene &sc, const V3 o, const V3 d, int slot, V3 &P, V3 &N) {
  const V3 v0 = mk(__ldg(sc.v0 + slot)), e1 = mk(__ldg(sc.e1 + slot)), e2 = mk(__ldg(sc.e2 + slot));
  const V3 h = vcross(d, e2);
  const float a = vdot(e1, h);
  const float f = 1 / a;
  const V3 s = vsub(o, v0);
  const float u = f * vdot(s, h);
  const V3 q = vcross(s, e1);
  const float v = f * vdot(d, q);
  const float t = f * vdot(e2, q);
  P = vadd(o, vscale(vnorm(d), t * vlen(d)));
  const V3 n0 = mk(__ldg(sc.n0 + slot)), n1 = mk(__ldg(sc.n1 + slot)), n2 = mk(__ldg(sc.n2 + slot));
  N = vadd(vadd(vscale(n0, 1 - u - v), vscale(n1, u)), vscale(n2, v));
}

template <bool ANY>
__device__ __forceinline__ bool consider(const DevScene &sc, const V3 o, const V3 d, int slot, Best &best) {
  const float4 a4 = __ldg(sc.v0 + slot), b4 = __ldg(sc.e1 + slot), c4 = __ldg(sc.e2 + slot);
  float dist;
  if (!tri_candidate(o, d, a4, b4, c4, dist)) return false;
  if (ANY) return true;
  const int id = __float_as_int(a4.w);
  if (dist < best.dist || (dist == best.dist && id < best.id)) {
    best.dist = dist;
    best.id = id;
    best.slot = slot;
    return true;
  }
  return false;
}

// Every triangle for every ray: the semantic definition (gpu/hit.cu:84-114, PARTITIONING_NONE).
template <bool ANY, bool COUNT>
__device__ bool trace_brute(const DevScene &sc, const V3 o, const V3 d, Best &best, Counters &c) {
  best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
  for (uint32_t s = 0; s < sc.n_tris; ++s) {
    if (COUNT) ++c.tris;
    if (consider<ANY>(sc, o, d, (int)s, best) && ANY) return true;
  }
  return best.slot >= 0;
}

// Conservative slab test.  The box is inflated by eps on every side; eps is a multiple of
// the float spacing at the largest coordinate the triangle test can involve, so a ray
// that the (rounded) Moeller-Trumbore arithmetic accepts is never culled.

__device__ __forceinline__ float safe_rcp(float x) {
  const float tiny = 1e-30f;
  return 1.0f / (fabsf(x) > tiny ? x : copysignf(tiny, x));
}

struct Slabs { float ix, iy, iz, ax, ay, az, bx, by, bz; };

__device__ __forceinline__ Slabs make_slabs(const V3 o, const V3 d, float eps) {
  Slabs s;
  s.ix = safe_rcp(d.x); s.iy = safe_rcp(d.y); s.iz = safe_rcp(d.z);
  s.ax = (o.x + eps) * s.ix; s.ay = (o.y + eps) * s.iy; s.az = (o.z + eps) * s.iz;   // pairs with box min
  s.bx = (o.x - eps) * s.ix; s.by = (o.y - eps) * s.iy; s.bz = (o.z - eps) * s.iz;   // pairs with box max
  return s;
}

// USE_TMAX = false for shadow rays: nothing ever shortens them (gpu/light.cu:29-36 has no far bound).
template <bool USE_TMAX = true>
__device__ __forceinline__ bool slab_hit(const Slabs &s, float lox, float hix, float loy, float hiy, float loz,
                                         float hiz, float tmax, float &tn) {
  const float x0 = __fmaf_rn(lox, s.ix, -s.ax), x1 = __fmaf_rn(hix, s.ix, -s.bx);
  const float y0 = __fmaf_rn(loy, s.iy, -s.ay), y1 = __fmaf_rn(hiy, s.iy, -s.by);
  const float z0 = __fmaf_rn(loz, s.iz, -s.az), z1 = __fmaf_rn(hiz, s.iz, -s.bz);
  const float nx = fminf(x0, x1), fx = fmaxf(x0, x1);
  const float ny = fminf(y0, y1), fy = fmaxf(y0, y1);
  const float nz = fminf(z0, z1), fz = fmaxf(z0, z1);
  tn = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f));
  const float tf = USE_TMAX ? fminf(fminf(fx, fy), fminf(fz, tmax)) : fminf(fminf(fx, fy), fz);
  return tn <= tf;
}

// The lanes of a warp leave a traversal at different times and must meet again before the next one
// starts (the shadow rays of a sample, the next light, the next bounce).  The hardware re-converges
// at the join block after the traversal -- if there is one: with the result visible to the
// optimiser, the compiler threads every exit of the traversal straight into the branch that
// consumes it ("hit" / "miss" code), the join disappears and the following traversal runs with the
// lanes scattered (measured: 21.8 instead of 25.7 lanes per instruction and 1.8 x the frame time on
// susans_smooth).  An empty asm makes the value opaque, so the join block stays.
__device__ __forceinline__ int rejoin(int v) {
  asm volatile("" : "+r"(v));
  return v;
}
__device__ __forceinline__ bool rejoin(bool v) { return rejoin((int)v) != 0; }

constexpr int TRAVERSAL_DONE = (int)0x80000000;   // stack sentinel: not a valid leaf reference

// Speculative "while-while" traversal (Aila-Laine): every lane descends inner nodes; the first
// leaf it reaches is set aside and the descent goes on until every lane of the warp holds one,
// then the warp tests triangles together.  Keeping the two phases apart keeps the lanes of a
// warp in the same code at the same time; with a single if-inner-else-leaf loop the triangle
// code ran at 7.5 of 32 lanes (profiles/r1_ncu_frame_v1.txt).  History on C5: if-else 53.9 ms,
// plain while-while 47.4, speculative 44.5, one pop per step 40.5.
// start: the node (or leaf reference) the walk begins at; sc.root for a whole-tree query.
template <bool ANY, bool COUNT>
__device__ bool trace_tree(const DevScene &sc, const V3 o, const V3 d, const Slabs &sl, float eps, int start,
                           Best &best, Counters &c) {
  best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
  // distance -> ray parameter, rounded up
  const float inv_len = 1.0001f / sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
  float tmax = INF;
  int stack[STACK_MAX + 1];
  stack[0] = TRAVERSAL_DONE;
  int sp = 1;
  int cur = start;
  int pending = 0;                       // leaf references are negative, 0 = none held
  constexpr int NONE = 0x7fffffff;       // neither a node index nor a leaf
  for (;;) {
    while (cur >= 0) {
      if (COUNT) ++c.nodes;
      const float4 *np = sc.nodes + 4ll * cur;
      const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
      float tn0, tn1;
      const bool h0 = slab_hit<!ANY>(sl, q0.x, q0.y, q0.z, q0.w, q2.x, q2.y, tmax, tn0);
      const bool h1 = slab_hit<!ANY>(sl, q1.x, q1.y, q1.z, q1.w, q2.z, q2.w, tmax, tn1);
      const int r0 = __float_as_int(q3.x), r1 = __float_as_int(q3.y);
      int next = h0 ? r0 : (h1 ? r1 : NONE);
      if (h0 && h1) {
        // shadow rays only ask "anything?": storage order, no near-first compare
        const bool far0 = ANY ? false : tn1 < tn0;
        next = far0 ? r1 : r0;
        stack[sp++] = far0 ? r0 : r1;
      }
      if (next < 0 && pending == 0) {      // first leaf: keep it, go on
        pending = next;
        next = NONE;
      }
      cur = next != NONE ? next : stack[--sp];                  // the only pop of a step
      if (!__any_sync(__activemask(), pending == 0)) break;     // every lane here holds a leaf
    }
    while (pending != 0 || (cur < 0 && cur != TRAVERSAL_DONE)) {
      int leaf;
      if (pending != 0) { leaf = pending; pending = 0; }
      else { leaf = cur; cur = stack[--sp]; }
      const uint32_t packed = (uint32_t)~leaf;
      const int first = (int)(packed >> 3), cnt = (int)(packed & 7u);
      for (int k = 0; k < cnt; ++k) {
        if (COUNT) ++c.tris;
        if (consider<ANY>(sc, o, d, first + k, best)) {
          if (ANY) return true;
          tmax = (best.dist * 1.0001f + eps) * inv_len;
        }
      }
    }
    if (cur == TRAVERSAL_DONE) break;
  }
  return best.slot >= 0;
}

// start_slot >= 0 (ANY only): the ray starts on the triangle in that slot
template <int TRACER, bool ANY, bool COUNT>
__device__ __forceinline__ bool trace(const DevScene &sc, const V3 o, const V3 d, Best &best, Counters &c,
                                      int start_slot = -1) {
  if (TRACER == RTB_TRACER_BRUTE) return trace_brute<ANY, COUNT>(sc, o, d, best, c);
  const float amax = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)) + sc.abs_max;
  const float eps = sc.margin_k * 1.1920929e-7f * amax;
  const Slabs sl = make_slabs(o, d, eps);
  return trace_tree<ANY, COUNT>(sc, o, d, sl, eps, sc.root, best, c);
}

}  // namespace
#include "lightgrid.cuh"
namespace {

// gpu/light.cu:12-27
__device__ __forceinline__ uint32_t apply_specular(uint32_t color, const V3 inc_o, const V3 inc_d, const V3 P,
                                                   const V3 N, uint32_t ks_col, float ns) {
  V3 V = vsub(inc_o, P);
  V3 R = vsub(inc_d, vscale(N, 2 * vdot(N, inc_d)));
  R = vnorm(R);
  V = vnorm(V);
  const float dt = vdot(R, V);
  const float Ls = powf(dt > 0.0f ? dt : 0.0f, ns);
  return color_add(color, color_mul(ks_col, Ls));
}

// A shadow ray through the hierarchy (or the exhaustive loop): what every shadow ray of a scene without
// light grids does, and the rays a grid cannot answer.  One out-of-line copy: the frame kernel's hot
// path (grid lookups) stays small enough for the instruction cache.
struct WalkArgs { const float4 *v0, *e1, *e2, *nodes; int root; uint32_t n_tris; float abs_max, margin_k; };

template <int TRACER>
__device__ __noinline__ bool shadow_walk(const WalkArgs w, const V3 P, const V3 dir) {
  DevScene sc;                    // only the fields a traversal reads (arguments by value: see lightgrid.cuh lg_exact32)
  sc.v0 = w.v0; sc.e1 = w.e1; sc.e2 = w.e2; sc.nodes = w.nodes; sc.root = w.root; sc.n_tris = w.n_tris;
  sc.abs_max = w.abs_max; sc.margin_k = w.margin_k;
  Best b;
  Counters c = {0, 0, 0, 0, 0};
  return trace<TRACER, true, false>(sc, P, dir, b, c);
}

// gpu/light.cu:50-127
// Deferred tree (FrameArgs::need_tree): a shadow ray that is off its light's grid (code < 0 from lg_any, -1: the light has
// no grid).  First pass: list it and answer "not occluded" for now; second pass (its tile rendered again): the answer
// k_stray_any found.  No list, or a full one: the whole frame again.  Out of line: one ray in millions comes here.
__device__ __noinline__ int stray_shadow_ray(const FrameArgs &fa, int code, const V3 P, const V3 dir, unsigned light, unsigned sample,
                                             unsigned item) {
  unsigned *nt = fa.need_tree;
  atomicAdd(nt + 9 - code, 1u);      // rast[17]: no grid, [18] outside the box, [19] off the grid, [20] projection
  if (fa.stray_mode == 2) {
    const unsigned cnt = min(nt[14], fa.stray_cap);
    int answer = 0;
    for (unsigned k = 0; k < cnt; ++k)
      if (fa.stray[k].sample == sample && fa.stray[k].light == light) answer = (int)fa.stray[k].occluded;
    return answer;
  }
  if (fa.stray_mode == 1) {
    const unsigned k = atomicAdd(nt + 14, 1u);
    if (k < fa.stray_cap) {
      StrayQuery q;
      q.ox = P.x; q.oy = P.y; q.oz = P.z; q.dx = dir.x; q.dy = dir.y; q.dz = dir.z;
      q.sample = sample; q.light = light; q.item = item; q.occluded = 0u; q.pad[0] = q.pad[1] = 0u;
      fa.stray[k] = q;
      return 0;
    }
  }
  *nt = 1u;
  return 0;
}

template <int TRACER, bool COUNT, bool DEFER = false>
__device__ uint32_t apply_light(const DevScene &sc, int obj, int slot, const V3 P, const V3 Nrm, Counters &c, const FrameArgs *fa = nullptr) {
  const float4 m0 = __ldg(sc.materials + 2 * obj);
  const uint32_t ka = __float_as_uint(m0.x), kd = __float_as_uint(m0.y), ks = __float_as_uint(m0.z);
  const float ns = m0.w;
  uint32_t color = init_color(0, 0, 0);
  for (uint32_t i = 0; i < sc.n_lights; ++i) {
    const float4 l0 = __ldg(sc.lights + 2 * i), l1 = __ldg(sc.lights + 2 * i + 1);
    const int type = __float_as_int(l0.x);
    const uint32_t lcol = __float_as_uint(l0.y);
    const V3 lv = mk(l1);
    if (type == AMBIENT) {
      color = color_add(color, color_mults(lcol, ka));
    } else if (type == DIRECTIONAL || type == POINT) {
      // one shadow-ray call site for both light kinds
      const V3 L = vscale(lv, -1);
      const V3 dir = type == DIRECTIONAL ? L : vsub(lv, P);
      if (COUNT) ++c.shadow;
      // shadow-ray grid of this light (lightgrid.cuh) when it has one; -1: this ray walks the tree
      int occluded = -1;
      if (TRACER == RTB_TRACER_TREE && !COUNT && sc.lgrids && sc.lgrids[i].kind) {
#ifdef RTB_SHADOW_CACHE      /* A/B only: 14.08 against 13.31 ms for k_frame on C5 */
        int &cached = (i & 1u) ? c.occ1 : c.occ0;
        Best nb;
        if (cached >= 0 && consider<true>(sc, P, dir, cached, nb)) occluded = 1;
        else {
          int hit = -1;
          occluded = lg_any(sc, sc.lgrids[i], P, dir, hit);
          if (occluded > 0) cached = hit;
        }
#else
        int hit;
        occluded = lg_any(sc, sc.lgrids[i], P, dir, hit);
#endif
      }
      if (DEFER && occluded < 0) occluded = stray_shadow_ray(*fa, occluded, P, dir, i, c.sample, c.item);
      if (occluded < 0) {
        if (COUNT) { Best b; occluded = trace<TRACER, true, true>(sc, P, dir, b, c) ? 1 : 0; }      // the counting frame: inline, counters live
        else {
          const WalkArgs w = {sc.v0, sc.e1, sc.e2, sc.nodes, sc.root, sc.n_tris, sc.abs_max, sc.margin_k};
          occluded = shadow_walk<TRACER>(w, P, dir) ? 1 : 0;
        }
      }
      occluded = rejoin(occluded);
      if (!occluded) {
        uint32_t t = color_mults(lcol, kd);
        if (type == DIRECTIONAL) {
          t = color_mul(t, vdot(L, Nrm));
          t = apply_specular(t, vadd(P, vscale(lv, -10)), lv, P, Nrm, ks, ns);
        } else {
          V3 N = Nrm;
          if (vdot(L, N) < 0) N = vscale(N, -1);
          const float dist = vlen(vsub(lv, P));
          t = color_mul(t, vdot(L, N) * 1 / dist);
          t = apply_specular(t, vadd(P, vscale(dir, -10)), dir, P, Nrm, ks, ns);
        }
        color = color_add(color, t);
      }
    }
  }
  return color;
}

__device__ __forceinline__ void primary_ray(const DevCamera &cam, int px, int py, V3 &o, V3 &d) {
  const V3 ui = vscale(mk(cam.u), (float)(px - cam.W / 2));
  const V3 vj = vscale(mk(cam.v), (float)(py - cam.H / 2));
  o = vadd(vadd(mk(cam.C), ui), vj);
  d = vnorm(vsub(mk(cam.pos), o));
}

#ifndef RTB_RANK_TILE
#define RTB_RANK_TILE 32      /* rank tile edge in output pixels (multiple of 8) */
#endif
constexpr int RANK_TILE = RTB_RANK_TILE;
constexpr int RT_WX = RANK_TILE / 8, RT_WY = RANK_TILE / 4, RT_ITEMS = RT_WX * RT_WY;   // warp tiles per rank tile

static_assert(RANK_TILE % 32 == 0 || RANK_TILE == 16 || RANK_TILE == 8, "rank tiles");

}  // namespace
#include "raster.cuh"
namespace {

// raytrace kernel body (gpu/raytracer.cu:87-129) for one sample
// use_prim: the first ray's winner comes from the projection pass (raster.cuh prim_lookup)
template <int TRACER, bool COUNT, bool DEFER = false>
__device__ uint32_t shade_sample(const DevScene &sc, const FrameArgs &fa, bool use_prim, int px, int py, Counters &c) {
  const DevCamera &cam = fa.cam;
  V3 o, d;
  primary_ray(cam, px, py, o, d);
  float nr_ = 1, r = 1;
  int budget = 10;
  uint32_t color = init_color(0, 0, 0);
  bool first = true;
  do {
    if (COUNT) { if (first) ++c.primary; else ++c.bounce; }
    first = false;
    Best b;
    uint32_t tc;
    bool hit;
    if (use_prim) { hit = prim_lookup(sc, fa, px, py, o, d, b); use_prim = false; }
    else if (DEFER) { *fa.need_tree = 1u; atomicAdd(fa.need_tree + 9, 1u); hit = false; }      // rast[16]: nearest-hit rays without a tree
    else hit = trace<TRACER, false, COUNT>(sc, o, d, b, c);
    hit = rejoin(hit);
    if (hit) {
      V3 P, N;
      hit_record(sc, o, d, b.slot, P, N);
      const int obj = __float_as_int(__ldg(sc.e1 + b.slot).w);
      tc = apply_light<TRACER, COUNT, DEFER>(sc, obj, b.slot, P, N, c, DEFER ? &fa : nullptr);
      r = __ldg(sc.materials + 2 * obj + 1).x;
      if (r > 0) {
        d = vsub(d, vscale(N, 2 * vdot(N, d)));
        o = P;
      }
    } else {
      r = 0;
      tc = init_color(0, 0, 0);
    }
    tc = color_mul(tc, nr_);
    color = color_add(color, tc);
    nr_ *= r;
  } while (nr_ > 0.01f && budget-- > 0);
  return color;
}

constexpr int FRAME_THREADS = 128;
constexpr int FRAME_WARPS = FRAME_THREADS / 32;

// ---- work items of a frame ----
// A work item is one warp tile (8x4 output pixels) of a rank tile this rank owns.  Rank tile
// (Tx,Ty) belongs to rank (Tx + 3*Ty) mod N, so in tile row Ty this rank owns the columns
// Tx0(Ty) + N*j; rows are padded to J = ceil(nTx / N) slots (a slot past the row end is an
// empty item).  Only owned tiles are enumerated: one atomic per item, whatever N is.
__device__ __forceinline__ unsigned frame_items(const FrameArgs &fa) {
  const int nTx = (fa.out_w + RANK_TILE - 1) / RANK_TILE, nTy = (fa.out_h + RANK_TILE - 1) / RANK_TILE;
  const int J = (nTx + fa.nranks - 1) / fa.nranks;
  return (unsigned)J * (unsigned)nTy * (unsigned)RT_ITEMS;
}
__device__ __forceinline__ bool frame_item(const FrameArgs &fa, unsigned item, int &lx0, int &ly0) {
  const int nTx = (fa.out_w + RANK_TILE - 1) / RANK_TILE;
  const int N = fa.nranks;
  const int J = (nTx + N - 1) / N;
  const int T = (int)(item / RT_ITEMS), sub = (int)(item % RT_ITEMS);
  const int Ty = T / J, j = T - Ty * J;
  int tx0 = (fa.rank - 3 * Ty) % N;
  if (tx0 < 0) tx0 += N;
  const int Tx = tx0 + N * j;
  lx0 = Tx * RANK_TILE + (sub % RT_WX) * 8;
  ly0 = Ty * RANK_TILE + (sub / RT_WX) * 4;
  return lx0 < fa.out_w && ly0 < fa.out_h;
}

// Warp tiles: 8x4 output pixels.  32 warp tiles (4 across, 8 down) make one 32x32-pixel
// rank tile; rank tile (Tx,Ty) belongs to rank (Tx + 3*Ty) mod nranks.
#ifndef RTB_FRAME_MINBLOCKS
#define RTB_FRAME_MINBLOCKS 8
#endif
// FINE: the leading *fa.fine_hot entries of the item order (the tiles whose cost in the previous frame was
// above a quarter of a warp's fair share of the frame, launch_item_order) are cut into their a x a rounds of
// 32 samples, one work item each: the sums go to a global accumulator with atomics and the warp that finishes
// a tile's last round resolves it and leaves accumulator and counter zero for the next frame.  With 1/N of
// the tiles per rank the slowest tile sets the frame's tail (a tile over a cell of a light grid that hundreds
// of line regions cross takes milliseconds); cut into rounds taken by different warps it no longer does.
// Every other tile stays whole: its rounds share one SM's L1 (all tiles as rounds cost 19 % at any N).
template <int TRACER, bool COUNT, bool FINE, bool DEFER = false>
__global__ void __launch_bounds__(FRAME_THREADS, RTB_FRAME_MINBLOCKS) k_frame(const DevScene sc, const FrameArgs fa) {
  // Per output pixel of the warp tile: the channel sums of its a x a samples (x: r | g << 16,
  // y: b).  The reference sums the u8 channels as floats (gpu/raytracer.cu:62-70); sums of at
  // most 256 values below 256 are exact in either arithmetic, so integers give the same bits.
  // 1 KB per CTA whatever the aliasing, which leaves the unified L1 its largest share.
  __shared__ uint2 s_acc[FRAME_WARPS][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int a = fa.aliasing, aa = a * a;
  uint2 *acc = s_acc[warp];
  const unsigned n_items = fa.only_items ? fa.only_items : frame_items(fa);      // only_items: the tiles of item_order, nothing else
  const unsigned hot = FINE && fa.item_order && !fa.only_items ? min(__ldg(fa.fine_hot), n_items) : 0u;
  const unsigned n_fine = hot * (unsigned)aa, total = n_items + hot * (unsigned)(aa - 1);
  const bool use_prim = TRACER == RTB_TRACER_TREE && !COUNT && rast_active(fa);
  Counters c = {0, 0, 0, 0, 0};
  for (;;) {
    unsigned t = 0;
    if (lane == 0) t = atomicAdd(fa.tile_counter, 1u);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= total) break;
    // items are taken in the order of the previous frame's cost, longest first (api.cu); the rounds of a
    // hot tile are consecutive, so different warps take them at about the same time
    const bool fine = FINE && t < n_fine;
    const unsigned ti = fine ? t / (unsigned)aa : t - (n_fine - hot);
    const int q0 = fine ? (int)(t - ti * (unsigned)aa) : 0, q1 = fine ? q0 + 1 : aa;
    const unsigned item = fa.item_order ? __ldg(fa.item_order + ti) : ti;
    int lx0, ly0;
    if (!frame_item(fa, item, lx0, ly0)) continue;
    long long t_start = 0;
    if (fa.item_cost && lane == 0) t_start = clock64();
    if (!fine) {
      acc[lane] = make_uint2(0u, 0u);
      __syncwarp();
    }
    for (int q = q0; q < q1; ++q) {
      const int sx = 8 * (q % a) + (lane & 7), sy = 4 * (q / a) + (lane >> 3);
      const int px = a * lx0 + sx, py = a * ly0 + sy;
      if (px < fa.cam.W && py < fa.cam.H) {
        if (DEFER) { c.sample = (unsigned)px | ((unsigned)py << 16); c.item = item; }
        const uint32_t col = shade_sample<TRACER, COUNT, DEFER>(sc, fa, use_prim, px, py, c);
        uint2 *dst = fine ? fa.fine_acc + (size_t)(ly0 + sy / a) * fa.out_w + (lx0 + sx / a) : acc + (sy / a) * 8 + sx / a;
        atomicAdd(&dst->x, ch_r(col) | (ch_g(col) << 16));
        atomicAdd(&dst->y, ch_b(col));
      }
    }
    bool resolve = true;
    if (fine) {
      // the last round of the tile to finish resolves it: sums first (atomics at L2), then the count
      __threadfence();
      __syncwarp();
      unsigned before = 0;
      if (lane == 0) before = atomicAdd(fa.fine_done + item, 1u);
      before = __shfl_sync(0xffffffffu, before, 0);
      resolve = before == (unsigned)aa - 1u;
      if (resolve) __threadfence();
    } else __syncwarp();
    if (fa.item_cost && lane == 0) {
      const unsigned dt = (unsigned)((clock64() - t_start) >> 8);
      if (fine) atomicAdd(fa.item_cost + item, dt); else fa.item_cost[item] = dt;
    }
    if (!resolve) continue;
    // downscale kernel body (gpu/raytracer.cu:49-85) for this lane's output pixel
    const int lx = lx0 + (lane & 7), ly = ly0 + (lane >> 3);
    if (lx < fa.out_w && ly < fa.out_h) {
      uint2 sum = acc[lane];
      if (fine) {
        uint2 *g = fa.fine_acc + (size_t)ly * fa.out_w + lx;
        sum = __ldcg(g);
        __stcg(g, make_uint2(0u, 0u));
      }
      float r = (float)(sum.x & 0xffffu), g = (float)(sum.x >> 16), b = (float)sum.y;
      const float ali2 = 255.0f * (float)a * (float)a;
      r /= ali2;
      g /= ali2;
      b /= ali2;
      const uint32_t e = init_color(r, g, b);
      unsigned char *dst = fa.out + (long long)(fa.out_h - ly - 1) * fa.pitch + 4ll * (fa.out_w - lx - 1);
      *reinterpret_cast<uint32_t *>(dst) = e;
    }
    if (fine && lane == 0) fa.fine_done[item] = 0u;
    __syncwarp();
  }
  if (COUNT) {
    unsigned v[5] = {c.primary, c.shadow, c.bounce, c.nodes, c.tris};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      unsigned long long s = v[k];
#pragma unroll
      for (int dlt = 16; dlt; dlt >>= 1) s += __shfl_xor_sync(0xffffffffu, s, dlt);
      if (lane == 0 && s) atomicAdd(fa.counters + k, s);
    }
  }
}

template <int TRACER>
__global__ void __launch_bounds__(128) k_hit_ids(const DevScene sc, const DevCamera cam, int32_t *tri, int32_t *obj) {
  // 8x4 sample blocks per warp, like the frame kernel
  const int lane = threadIdx.x & 31;
  const long long w = (long long)blockIdx.x * 4 + (threadIdx.x >> 5);
  const int bw = (cam.W + 7) / 8;
  const int bx = (int)(w % bw), by = (int)(w / bw);
  const int px = bx * 8 + (lane & 7), py = by * 4 + (lane >> 3);
  if (px >= cam.W || py >= cam.H) return;
  V3 o, d;
  primary_ray(cam, px, py, o, d);
  Best b;
  Counters c = {0, 0, 0, 0, 0};
  const bool hit = trace<TRACER, false, false>(sc, o, d, b, c);
  const size_t k = (size_t)py * cam.W + px;
  tri[k] = hit ? b.id : -1;
  if (obj) obj[k] = hit ? __float_as_int(__ldg(sc.e1 + b.slot).w) : -1;
}

template <int TRACER, bool ANY>
__global__ void __launch_bounds__(128) k_trace_rays(const DevScene sc, const float *rays, size_t n, int32_t *tri,
                                                    float *dist) {
  const size_t i = (size_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  const V3 o = mk(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
  const V3 d = mk(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
  Best b;
  Counters c = {0, 0, 0, 0, 0};
  const bool hit = trace<TRACER, ANY, false>(sc, o, d, b, c);
  if (tri) tri[i] = hit ? (ANY ? 0 : b.id) : -1;
  if (dist) dist[i] = (hit && !ANY) ? b.dist : 0.0f;
}

#include "wave.cuh"   // staged wavefront kernel: binary node layout only (A/B builds)

float g_margin_k = 24.0f;

}  // namespace

void set_margin_k(float k) { g_margin_k = k; }
float get_margin_k() { return g_margin_k; }

// ---- shadow-ray grids (lightgrid.cuh), built once per resident scene ----
constexpr uint32_t LG_MIN_TRIS = 1u << 15;
constexpr int LG_MAX_GRIDS = 8;

// device scratch of one light's grid build, freed on every way out (kept blocks excepted)
struct LgScratch {
  cudaStream_t st;
  void *p[16] = {nullptr};
  int n = 0;
  template <typename T> int alloc(T **out, size_t bytes) {
    *out = nullptr;
    { const int rc = block_alloc((void **)out, bytes, st); if (rc) return rc; }
    p[n++] = *out;
    return RTB_OK;
  }
  void keep(void *q) { for (int k = 0; k < n; ++k) if (p[k] == q) p[k] = nullptr; }
  ~LgScratch() { for (int k = 0; k < n; ++k) if (p[k]) block_free(p[k], st); }
};

int build_light_grids(SceneStore &S, const struct light *lights, size_t n_lights, const float bmin[3],
                      const float bmax[3], cudaStream_t st) {
  if (!S.has_tree || S.n_tris < LG_MIN_TRIS || S.n_tris >= (1u << 27) || !n_lights) return RTB_OK;
  std::vector<LightGrid> grids(n_lights);
  memset(grids.data(), 0, grids.size() * sizeof(LightGrid));
  const DevScene sc = S.dev();
  const uint32_t n = S.n_tris;
  const unsigned blocks = (n + 255) / 256;
  // Shadow-ray origins are hit points: inside the scene's box up to the rounding of the hit point
  // (lg_any_warp sends everything else to the tree).  The regions bound |origin - vertex| and |origin| with it.
  double lo[3], hi[3], diag2 = 0, pmax2 = 0;
  for (int k = 0; k < 3; ++k) {
    const double pad = 1e-5 * (fabs((double)bmin[k]) + fabs((double)bmax[k]) + ((double)bmax[k] - (double)bmin[k])) + 1e-30;
    lo[k] = (double)bmin[k] - pad; hi[k] = (double)bmax[k] + pad;
    diag2 += (hi[k] - lo[k]) * (hi[k] - lo[k]);
    const double m = fmax(fabs(lo[k]), fabs(hi[k]));
    pmax2 += m * m;
  }
  const double diag = sqrt(diag2) * (1.0 + 1e-6), pmax = sqrt(pmax2) * (1.0 + 1e-6);
  int built = 0, mem = 0;
  auto unit = [](float x, float y, float z, float3 &o) {
    const float r = sqrtf(x * x + y * y + z * z);
    if (!(r > 0) || !(r < 1e30f)) return false;
    o = make_float3(x / r, y / r, z / r);
    return true;
  };
  for (size_t i = 0; i < n_lights && built < LG_MAX_GRIDS; ++i) {
    const struct light &l = lights[i];
    LightGrid g;
    memset(&g, 0, sizeof g);
    LgBuild lb;
    memset(&lb, 0, sizeof lb);
    const float cx = 0.5f * (bmin[0] + bmax[0]), cy = 0.5f * (bmin[1] + bmax[1]), cz = 0.5f * (bmin[2] + bmax[2]);
    if (l.type == DIRECTIONAL) {
      if (!unit(l.v.x, l.v.y, l.v.z, g.az)) continue;
      g.kind = 1;
    } else if (l.type == POINT) {
      g.org = make_float3(l.v.x, l.v.y, l.v.z);
      const float ext = fmaxf(bmax[0] - bmin[0], fmaxf(bmax[1] - bmin[1], bmax[2] - bmin[2]));
      if (!unit(cx - l.v.x, cy - l.v.y, cz - l.v.z, g.az)) continue;
      const float dist = sqrtf((cx - l.v.x) * (cx - l.v.x) + (cy - l.v.y) * (cy - l.v.y) + (cz - l.v.z) * (cz - l.v.z));
      if (!(dist > 1e-3f * ext)) continue;            // a light at the centre of the scene: half of it is off the grid
      g.kind = 2;
    } else continue;
    // any orthonormal frame around az (float: the frame the rays are projected with)
    {
      const float ax = fabsf(g.az.x), ay = fabsf(g.az.y), az = fabsf(g.az.z);
      float3 h = ax <= ay && ax <= az ? make_float3(1, 0, 0) : (ay <= az ? make_float3(0, 1, 0) : make_float3(0, 0, 1));
      float3 a = make_float3(h.y * g.az.z - h.z * g.az.y, h.z * g.az.x - h.x * g.az.z, h.x * g.az.y - h.y * g.az.x);
      if (!unit(a.x, a.y, a.z, g.ax)) continue;
      g.ay = make_float3(g.az.y * g.ax.z - g.az.z * g.ax.y, g.az.z * g.ax.x - g.az.x * g.ax.z, g.az.x * g.ax.y - g.az.y * g.ax.x);
    }
    g.bmin = make_float3((float)lo[0], (float)lo[1], (float)lo[2]);
    g.bmax = make_float3((float)hi[0], (float)hi[1], (float)hi[2]);
    // the same frame in double for region.cuh, and the domain the regions are clipped to
    const double AX[3] = {g.ax.x, g.ax.y, g.ax.z}, AY[3] = {g.ay.x, g.ay.y, g.ay.z}, AZ[3] = {g.az.x, g.az.y, g.az.z};
    if (g.kind == 1) {
      region::DirFrame &f = lb.df;
      // gpu/light.cu:70-74: the shadow ray's direction is scale(l.v, -1), exactly
      f.d[0] = -(double)l.v.x; f.d[1] = -(double)l.v.y; f.d[2] = -(double)l.v.z;
      for (int k = 0; k < 3; ++k) { f.ax[k] = AX[k]; f.ay[k] = AY[k]; }
      f.g11 = region::dot3(AX, AX); f.g12 = region::dot3(AX, AY); f.g22 = region::dot3(AY, AY);
      f.smax = diag; f.Pmax = pmax;
      lb.dom[0] = lb.dom[2] = 1e300; lb.dom[1] = lb.dom[3] = -1e300;
      for (int c = 0; c < 8; ++c) {
        const double X[3] = {(c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1], (c & 4) ? hi[2] : lo[2]};
        const double x = region::dot3(X, AX), y = region::dot3(X, AY);
        lb.dom[0] = fmin(lb.dom[0], x); lb.dom[1] = fmax(lb.dom[1], x); lb.dom[2] = fmin(lb.dom[2], y); lb.dom[3] = fmax(lb.dom[3], y);
      }
      const double m = 1e-3 * ((lb.dom[1] - lb.dom[0]) + (lb.dom[3] - lb.dom[2])) + 1e-30;
      lb.dom[0] -= m; lb.dom[1] += m; lb.dom[2] -= m; lb.dom[3] += m;
    } else {
      region::PointFrame &f = lb.pf;
      f.org[0] = l.v.x; f.org[1] = l.v.y; f.org[2] = l.v.z;
      // dual basis of (ax, ay, az): rows of the inverse of the matrix whose rows are the axes
      const double M[3][3] = {{AX[0], AX[1], AX[2]}, {AY[0], AY[1], AY[2]}, {AZ[0], AZ[1], AZ[2]}};
      double c0[3], c1[3], c2[3];
      region::cross3(M[1], M[2], c0); region::cross3(M[2], M[0], c1); region::cross3(M[0], M[1], c2);
      const double det = region::dot3(M[0], c0);
      if (!(fabs(det) > 0.5)) continue;
      for (int k = 0; k < 3; ++k) { f.hx[k] = c0[k] / det; f.hy[k] = c1[k] / det; f.hz[k] = c2[k] / det; }
      double far2 = 0;
      for (int k = 0; k < 3; ++k) { const double m = fmax(fabs(f.org[k] - lo[k]), fabs(f.org[k] - hi[k])); far2 += m * m; }
      f.smax = diag; f.dmax = sqrt(far2) * (1.0 + 1e-6);
      const double lim = (double)LG_PERSP_LIMIT * (1.0 + 1e-3);
      f.ratio = sqrt(1.0 + 2.0 * lim * lim) * 1.001;
      lb.dom[0] = lb.dom[2] = -lim; lb.dom[1] = lb.dom[3] = lim;
    }
    LgScratch tmp;
    tmp.st = st;
    float4 *rec = nullptr;
    uint32_t *d_b = nullptr;
    { int rc = tmp.alloc(&rec, (size_t)n * 4 * sizeof(float4)); if (rc) return rc; }
    { int rc = tmp.alloc(&d_b, ((size_t)LG_STATE_WORDS + LG_BIG_CAP) * 4); if (rc) return rc; }
    k_lg_records<<<(n + 127) / 128, 128, 0, st>>>(sc, g, lb, rec);
    // extent of the grid: the boxes of every 16th triangle from 2^20 on (slots are in spatial order: a regular
    // sample of the surface); regions that span more than a quarter of the domain are lines, not geometry
    const uint32_t init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
    uint32_t hb[4];
    RTB_CUDA(cudaMemcpyAsync(d_b, init, sizeof init, cudaMemcpyHostToDevice, st));
    const uint32_t bstride = n >= (1u << 20) ? 16u : 1u;
    const float huge = (float)(0.25 * fmax(lb.dom[1] - lb.dom[0], lb.dom[3] - lb.dom[2]));
    k_lg_bounds<<<(n / bstride + 256) / 256, 256, 0, st>>>(rec, n, bstride, huge, d_b);
    RTB_CUDA(cudaMemcpyAsync(hb, d_b, sizeof hb, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    float x0 = ord2f(hb[0]), x1 = ord2f(hb[1]), y0 = ord2f(hb[2]), y1 = ord2f(hb[3]);
    if (bstride > 1 && x1 > x0 && y1 > y0) {      // room for what the sample missed
      const float mx = 0.01f * (x1 - x0), my = 0.01f * (y1 - y0);
      x0 -= mx; x1 += mx; y0 -= my; y1 += my;
    }
    x0 = fmaxf(x0, (float)lb.dom[0]); y0 = fmaxf(y0, (float)lb.dom[2]);
    x1 = fminf(x1, (float)lb.dom[1]); y1 = fminf(y1, (float)lb.dom[3]);
    if (!(x1 > x0) || !(y1 > y0) || !(x1 - x0 < 1e30f) || !(y1 - y0 < 1e30f)) continue;
    // about one cell per triangle: a triangle of average size touches ~6 cells, and a ray scans ~6 entries
    // (C5: 0.5 -> 17.5 ms, 1 -> 17.05, 2 -> 16.8 at twice the memory)
    static const double cells_per_tri = [] { const char *e = getenv("RTB_LG_CELLS"); return e && atof(e) > 0 ? atof(e) : 1.0; }();
    const double target = fmin(fmax(cells_per_tri * (double)n, 1024.0), 64.0e6);
    double cell = sqrt((double)(x1 - x0) * (double)(y1 - y0) / target);
    cell = fmax(cell, fmax((double)(x1 - x0), (double)(y1 - y0)) / 30000.0);
    g.x0 = x0; g.y0 = y0; g.cell = (float)cell; g.inv_cell = 1.0f / g.cell;
    g.nx = (int)ceil((x1 - x0) / cell) + 1; g.ny = (int)ceil((y1 - y0) / cell) + 1;
    const size_t cells = (size_t)g.nx * g.ny;
    uint32_t *counts = nullptr, *cursor = nullptr, *scratch = nullptr;
    uint2 *entries = nullptr;
    const size_t ssw = scan_scratch_words(cells + 1);
    { int rc = tmp.alloc(&counts, (cells + 1) * 4); if (rc) return rc; }
    { int rc = tmp.alloc(&scratch, ssw * 4); if (rc) return rc; }
    RTB_CUDA(cudaMemsetAsync(counts, 0, (cells + 1) * 4, st));
    RTB_CUDA(cudaMemsetAsync(d_b, 0, LG_STATE_WORDS * 4, st));
    k_lg_scatter<false><<<blocks, 256, 0, st>>>(sc, g, rec, counts, d_b, nullptr, nullptr);
    k_lg_big<false><<<S.sm_count * 8, 128, 0, st>>>(sc, g, rec, counts, d_b, nullptr, nullptr);
    int rc = dev_exclusive_scan_u32(counts, cells + 1, scratch, ssw, false, st, nullptr);
    if (rc) return rc;
    uint32_t total = 0, state[2 + LG_LARGE_MAX];
    RTB_CUDA(cudaMemcpyAsync(&total, counts + cells, 4, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaMemcpyAsync(state, d_b, sizeof state, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    const bool ok = !state[1] && (size_t)total <= 24ull * n + (1u << 20) && total > 0;
    if (!ok) continue;
    { int rc2 = tmp.alloc(&entries, (size_t)total * sizeof(uint2)); if (rc2) return rc2; }
    { int rc2 = tmp.alloc(&cursor, cells * 4); if (rc2) return rc2; }
    RTB_CUDA(cudaMemsetAsync(cursor, 0, cells * 4, st));
    k_lg_scatter<true><<<blocks, 256, 0, st>>>(sc, g, rec, cursor, d_b, counts, entries);
    k_lg_big<true><<<S.sm_count * 8, 128, 0, st>>>(sc, g, rec, cursor, d_b, counts, entries);
    RTB_CUDA(cudaGetLastError());
    g.cell_start = counts; g.entries = entries; g.rec = rec;
    tmp.keep(counts); tmp.keep(entries); tmp.keep(rec);
    static const bool lg_trace = [] { const char *e = getenv("RTB_BUILD_TRACE"); return e && atoi(e) != 0; }();
    if (lg_trace) {
      unsigned long long *d_st = nullptr, h_st[3] = {0, 0, 0};
      unsigned h_big = 0;
      if (tmp.alloc(&d_st, 32) == RTB_OK) {
        cudaMemsetAsync(d_st, 0, 32, st);
        k_lg_stats<<<S.sm_count * 4, 256, 0, st>>>(counts, cells, d_st);
        cudaMemcpyAsync(h_st, d_st, 24, cudaMemcpyDeviceToHost, st);
        cudaMemcpyAsync(&h_big, d_b + 24, 4, cudaMemcpyDeviceToHost, st);
        cudaStreamSynchronize(st);
        fprintf(stderr, "rtb light grid %zu (%s): %d x %d cells, %u entries (%.2f per cell), %u triangles over more than %d rows, "
                "%u tested by every ray, longest list %llu, %llu cells with > 64 entries, a random entry's list has %.1f entries\n",
                i, g.kind == 1 ? "directional" : "point", g.nx, g.ny, total, (double)total / (double)cells, h_big, LG_BIG_ROWS,
                state[0], h_st[0], h_st[1], (double)h_st[2] / (double)(total ? total : 1));
      }
    }
    g.n_large = state[0] < (uint32_t)LG_LARGE_MAX ? state[0] : (uint32_t)LG_LARGE_MAX;
    for (uint32_t k = 0; k < g.n_large; ++k) g.large[k] = state[2 + k];
    S.lgrid_mem_bytes[mem] = (size_t)sc.n_tris * 64; S.lgrid_mem[mem++] = rec;
    S.lgrid_mem_bytes[mem] = ((size_t)cells + 1) * 4; S.lgrid_mem[mem++] = counts;
    S.lgrid_mem_bytes[mem] = (size_t)total * 4; S.lgrid_mem[mem++] = entries;
    S.lgrid_entries += total; S.lgrid_cells += cells;
    grids[i] = g;
    ++built;
  }
  if (!built) return RTB_OK;
  RTB_CUDA(cudaMallocAsync((void **)&S.lgrids, n_lights * sizeof(LightGrid), st));
  RTB_CUDA(cudaMemcpyAsync(S.lgrids, grids.data(), n_lights * sizeof(LightGrid), cudaMemcpyHostToDevice, st));
  RTB_CUDA(cudaStreamSynchronize(st));      // grids lives on this stack frame
  S.lgrid_margin_k = get_margin_k();
  S.n_lgrids = built;
  return RTB_OK;
}

template <int TRACER, bool COUNT, bool FINE = false, bool DEFER = false>
static int launch_frame_t(const DevScene &sc, const FrameArgs &fa, int sm_count, cudaStream_t st) {
  const size_t smem = 0;
  auto kern = k_frame<TRACER, COUNT, FINE, DEFER>;
#ifdef RTB_CARVEOUT   /* A/B only: shared-memory share of the unified L1, percent */
  RTB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, RTB_CARVEOUT));
#endif
  int per_sm = 0;
  RTB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FRAME_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  kern<<<sm_count * per_sm, FRAME_THREADS, smem, st>>>(sc, fa);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

constexpr int WAVE_MAX_BLOCKS_PER_SM = 16;

size_t wave_scratch_floats(int aliasing, int sm_count) {
  return (size_t)sm_count * WAVE_MAX_BLOCKS_PER_SM * WAVE_WARPS * 7 * 32 * aliasing * aliasing;
}

template <bool COUNT>
static int launch_wave_t(const DevScene &sc, const FrameArgs &fa, int sm_count, cudaStream_t st) {
  const size_t smem = (size_t)WAVE_WARPS * 16 * 32 * fa.aliasing * fa.aliasing;
  auto kern = k_frame_wave<COUNT>;
  if (smem > 48 * 1024) RTB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  RTB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WAVE_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  if (per_sm > WAVE_MAX_BLOCKS_PER_SM) per_sm = WAVE_MAX_BLOCKS_PER_SM;
  kern<<<sm_count * per_sm, WAVE_THREADS, smem, st>>>(sc, fa);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}


static int launch_frame_kernel(const DevScene &sc, const FrameArgs &fa, int tracer, bool count, int sm_count, cudaStream_t st);

// ItemEstimate (rtb_internal.h): one warp per work item, one lane per pixel of the 8x4 tile, the pixel's centre sample
__global__ void __launch_bounds__(256) k_item_estimate(const DevScene sc, const FrameArgs fa, unsigned *cost, unsigned n_items) {
  const int lane = threadIdx.x & 31;
  const unsigned warps = gridDim.x * 8;
  for (unsigned item = blockIdx.x * 8 + (threadIdx.x >> 5); item < n_items; item += warps) {
    int lx0, ly0;
    unsigned est = 0;
    if (frame_item(fa, item, lx0, ly0)) {
      const int a = fa.aliasing;
      const int px = a * (lx0 + (lane & 7)) + a / 2, py = a * (ly0 + (lane >> 3)) + a / 2;
      if (px < fa.cam.W && py < fa.cam.H) {
        est = 2;                                   // a sample that hits nothing
        const unsigned long long key = fa.prim[(size_t)py * fa.cam.W + px];
        if (key != ~0ull) {
          est = 8;
          V3 o, d;
          primary_ray(fa.cam, px, py, o, d);
          const V3 P = vadd(o, vscale(d, __uint_as_float((unsigned)(key >> 32))));
          for (uint32_t i = 0; i < sc.n_lights && sc.lgrids; ++i) {
            const LightGrid &g = sc.lgrids[i];
            if (!g.kind) continue;
            float x, y, den;
            lg_project(g, P, x, y, den);
            const float fi = floorf((x - g.x0) * g.inv_cell), fj = floorf((y - g.y0) * g.inv_cell);
            if (fi >= 0.0f && fj >= 0.0f && fi <= (float)(g.nx - 1) && fj <= (float)(g.ny - 1)) {
              const size_t c = (size_t)(int)fj * g.nx + (int)fi;
              est += 4u + (__ldg(g.cell_start + c + 1) - __ldg(g.cell_start + c));
            }
          }
        }
      }
    }
    est = __reduce_add_sync(0xffffffffu, est);
    if (lane == 0) cost[item] = est;
  }
}

int launch_item_order(const unsigned int *d_cost, unsigned n, uint64_t *d_keys, uint32_t *d_vals,
                      uint64_t *d_keys_alt, uint32_t *d_vals_alt, uint32_t *d_scratch, size_t scratch_words,
                      cudaStream_t st, int *result_in_alt, unsigned *d_hot, float per_share);

int launch_frame(const DevScene &sc, const FrameArgs &fa, int tracer, bool count, int sm_count, cudaStream_t st,
                 unsigned long long *launches, cudaEvent_t *marks, const ItemEstimate *est) {
  if (fa.aliasing < 1 || fa.aliasing > 16) { set_error("aliasing must be in 1..16"); return RTB_ERR_ARG; }
  if (marks) RTB_CUDA(cudaEventRecord(marks[0], st));
  RTB_CUDA(cudaMemsetAsync(fa.tile_counter, 0, sizeof(unsigned int), st));
  if (count) RTB_CUDA(cudaMemsetAsync(fa.counters, 0, 8 * sizeof(unsigned long long), st));
  if (launches) *launches = 1;
  if (fa.prim && tracer == RTB_TRACER_TREE && !count) {
    // projection pass: records of the nearest primary candidate per sample (raster.cuh)
    // rank tiles are whole sample tiles when 32 a is a multiple of 32 (always): a rank clears what it owns
    if (fa.nranks > 1) k_prim_clear<<<sm_count * 8, 256, 0, st>>>(fa);
    else RTB_CUDA(cudaMemsetAsync(fa.prim, 0xff, (size_t)fa.cam.W * fa.cam.H * 8, st));
    RTB_CUDA(cudaMemsetAsync(fa.rast, 0, RAST_STATE_WORDS * 4, st));
    if (sc.n_tris) {
      const unsigned blocks = (sc.n_tris + 255) / 256;
      k_rast_setup<<<fa.x.sharded ? (blocks + fa.nranks - 1) / fa.nranks : blocks, 256, 0, st>>>(sc, fa);
      if (marks) RTB_CUDA(cudaEventRecord(marks[1], st));
      k_rast_big<<<sm_count * 8, RAST_BIG_THREADS, 0, st>>>(sc, fa);
      if (marks) RTB_CUDA(cudaEventRecord(marks[2], st));
    } else if (marks) { RTB_CUDA(cudaEventRecord(marks[1], st)); RTB_CUDA(cudaEventRecord(marks[2], st)); }
    // sharded frames: counts and epoch to every rank, wait for theirs (the items went over NVLink already)
    if (fa.x.sharded) k_xchg_sync<<<1, 32, 0, st>>>(fa);
    if (marks) RTB_CUDA(cudaEventRecord(marks[3], st));
    if (sc.n_tris) k_rast_items<<<sm_count * 8, 256, 0, st>>>(sc, fa);
    RTB_CUDA(cudaGetLastError());
    if (launches) *launches = fa.x.sharded ? 5 : 4;
  } else if (marks)
    for (int k = 1; k <= 3; ++k) RTB_CUDA(cudaEventRecord(marks[k], st));
  if (marks) RTB_CUDA(cudaEventRecord(marks[4], st));
  if (est && fa.prim && tracer == RTB_TRACER_TREE && !count && !fa.item_order && est->n_items) {
    k_item_estimate<<<sm_count * 8, 256, 0, st>>>(sc, fa, est->cost, est->n_items);
    int in_alt = 0;
    const int rc2 = launch_item_order(est->cost, est->n_items, est->sk, est->sv, est->sk2, est->sv2, est->ss, est->ssw, st, &in_alt, est->hot, 0.0f);
    if (rc2) return rc2;
    FrameArgs fb = fa;
    fb.item_order = in_alt ? est->sv2 : est->sv;
    const int rc3 = launch_frame_kernel(sc, fb, tracer, count, sm_count, st);
    if (!rc3 && marks) RTB_CUDA(cudaEventRecord(marks[5], st));
    return rc3;
  }
  const int rc = launch_frame_kernel(sc, fa, tracer, count, sm_count, st);
  if (!rc && marks) RTB_CUDA(cudaEventRecord(marks[5], st));
  return rc;
}

static int launch_frame_kernel(const DevScene &sc, const FrameArgs &fa, int tracer, bool count, int sm_count, cudaStream_t st) {
  // RTB_KERNEL_WAVEFRONT was requested and is applicable (api.cu sets bounce_scratch only then);
  // everything else runs the persistent-warp per-sample kernel, the faster of the two
  if (tracer == RTB_TRACER_TREE && fa.bounce_scratch && fa.n_shadow_lights <= 32 && fa.aliasing <= WAVE_MAX_ALIASING)
    return count ? launch_wave_t<true>(sc, fa, sm_count, st) : launch_wave_t<false>(sc, fa, sm_count, st);
  if (tracer == RTB_TRACER_BRUTE)
    return count ? launch_frame_t<RTB_TRACER_BRUTE, true>(sc, fa, sm_count, st)
                 : launch_frame_t<RTB_TRACER_BRUTE, false>(sc, fa, sm_count, st);
  if (count) return launch_frame_t<RTB_TRACER_TREE, true>(sc, fa, sm_count, st);
  if (fa.need_tree) return launch_frame_t<RTB_TRACER_TREE, false, false, true>(sc, fa, sm_count, st);      // deferred tree (one-shot)
  if (fa.fine_acc) return launch_frame_t<RTB_TRACER_TREE, false, true>(sc, fa, sm_count, st);
  return launch_frame_t<RTB_TRACER_TREE, false>(sc, fa, sm_count, st);
}

// ---- deferred tree: the few shadow rays that are off their light's grid, against every triangle ----
// (the definition of has_direct_hit, gpu/light.cu:29-36 + gpu/hit.cu:241-253, with no cull at all; a thread per triangle
// slot, the rays from shared memory; 26 rays x 10 M triangles on C5: 0.2 ms)
__global__ void __launch_bounds__(256) k_stray_any(const DevScene sc, StrayQuery *q, const unsigned *count) {
  __shared__ float s_ray[STRAY_CAP][6];
  __shared__ unsigned s_hit[STRAY_CAP];
  const unsigned n = min(*count, STRAY_CAP);
  for (unsigned k = threadIdx.x; k < n; k += 256) {
    s_ray[k][0] = q[k].ox; s_ray[k][1] = q[k].oy; s_ray[k][2] = q[k].oz;
    s_ray[k][3] = q[k].dx; s_ray[k][4] = q[k].dy; s_ray[k][5] = q[k].dz;
    s_hit[k] = 0u;
  }
  __syncthreads();
  for (uint32_t s = blockIdx.x * 256 + threadIdx.x; s < sc.n_tris; s += gridDim.x * 256) {
    const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
    for (unsigned k = 0; k < n; ++k) {
      float dist;
      if (tri_candidate(mk(s_ray[k][0], s_ray[k][1], s_ray[k][2]), mk(s_ray[k][3], s_ray[k][4], s_ray[k][5]), a4, b4, c4, dist)) s_hit[k] = 1u;
    }
  }
  __syncthreads();
  for (unsigned k = threadIdx.x; k < n; k += 256)
    if (s_hit[k]) q[k].occluded = 1u;
}

int launch_stray_any(const DevScene &sc, StrayQuery *q, const unsigned *count, int sm_count, cudaStream_t st) {
  k_stray_any<<<sm_count * 4, 256, 0, st>>>(sc, q, count);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int launch_frame_only(const DevScene &sc, const FrameArgs &fa, int sm_count, cudaStream_t st) {
  RTB_CUDA(cudaMemsetAsync(fa.tile_counter, 0, sizeof(unsigned int), st));
  return launch_frame_kernel(sc, fa, RTB_TRACER_TREE, false, sm_count, st);
}

// ---- longest-first item order from the costs the previous frame left (api.cu) ----
// Keys are coarse cost classes (half octaves), not the costs themselves: the radix sort is
// stable, so items of one class keep their spatial order and neighbouring warps keep working
// on neighbouring tiles (a full sort by cost quadrupled the DRAM traffic of the frame).
#ifndef RTB_COST_CLASSES_PER_OCTAVE
#define RTB_COST_CLASSES_PER_OCTAVE 2
#endif
__device__ __forceinline__ int cost_class(unsigned cost) {      // 0 .. 24 * classes per octave
  return (int)((float)RTB_COST_CLASSES_PER_OCTAVE * log2f((float)min(cost, 0xffffffu) + 1.0f));
}

// hot[0..1]: sum of the costs (u64); hot[2]: items in a class above that of the threshold (k_item_hot)
__global__ void __launch_bounds__(256) k_item_keys(const unsigned int *cost, unsigned n, uint64_t *keys,
                                                   uint32_t *vals, unsigned *hot) {
  __shared__ unsigned long long s_sum;
  if (threadIdx.x == 0) s_sum = 0ull;
  __syncthreads();
  const unsigned i = blockIdx.x * 256 + threadIdx.x;
  unsigned long long mine = 0;
  if (i < n) {
    const unsigned c = cost[i];
    mine = c;
    keys[i] = (uint64_t)(255 - min(cost_class(c), 255));   // ascending sort = heaviest class first; empty items (cost 0) last
    vals[i] = i;
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, d);
  if ((threadIdx.x & 31) == 0 && mine) atomicAdd(&s_sum, mine);
  __syncthreads();
  if (threadIdx.x == 0 && s_sum) atomicAdd(reinterpret_cast<unsigned long long *>(hot), s_sum);
}

// hot tiles: cost class above the class of (sum of costs / resident warps) / hot_div -- a tile that keeps one
// warp for more than 1 / hot_div of the time the frame takes when every warp gets the same share.  They are
// a prefix of the sorted order (whole classes).
__global__ void __launch_bounds__(256) k_item_hot(const unsigned int *cost, unsigned n, unsigned *hot, float per_share) {
  const unsigned i = blockIdx.x * 256 + threadIdx.x;
  const float sum = (float)*reinterpret_cast<const unsigned long long *>(hot);
  const int cls_t = cost_class((unsigned)fminf(sum * per_share, 16777215.0f));
  const bool is_hot = i < n && cost[i] > 0u && cost_class(cost[i]) > cls_t;
  const unsigned m = __ballot_sync(0xffffffffu, is_hot);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(hot + 2, (unsigned)__popc(m));
}

unsigned frame_item_count(int out_w, int out_h, int nranks) {
  const int nTx = (out_w + RANK_TILE - 1) / RANK_TILE, nTy = (out_h + RANK_TILE - 1) / RANK_TILE;
  return (unsigned)((nTx + nranks - 1) / nranks) * (unsigned)nTy * (unsigned)RT_ITEMS;
}

int launch_item_order(const unsigned int *d_cost, unsigned n, uint64_t *d_keys, uint32_t *d_vals,
                      uint64_t *d_keys_alt, uint32_t *d_vals_alt, uint32_t *d_scratch, size_t scratch_words,
                      cudaStream_t st, int *result_in_alt, unsigned *d_hot, float per_share) {
  RTB_CUDA(cudaMemsetAsync(d_hot, 0, 4 * sizeof(unsigned), st));
  k_item_keys<<<(n + 255) / 256, 256, 0, st>>>(d_cost, n, d_keys, d_vals, d_hot);
  if (per_share > 0.0f) k_item_hot<<<(n + 255) / 256, 256, 0, st>>>(d_cost, n, d_hot, per_share);
  RTB_CUDA(cudaGetLastError());
  return dev_radix_sort_pairs_u64(d_keys, d_vals, d_keys_alt, d_vals_alt, n, 8, d_scratch, scratch_words, st,
                                  result_in_alt, nullptr);
}

int launch_hit_ids(const DevScene &sc, const DevCamera &cam, int tracer, int32_t *d_tri, int32_t *d_obj,
                   cudaStream_t st) {
  const long long bw = (cam.W + 7) / 8, bh = (cam.H + 3) / 4;
  const long long warps = bw * bh;
  const unsigned blocks = (unsigned)((warps + 3) / 4);
  if (!blocks) return RTB_OK;
  if (tracer == RTB_TRACER_BRUTE) k_hit_ids<RTB_TRACER_BRUTE><<<blocks, 128, 0, st>>>(sc, cam, d_tri, d_obj);
  else k_hit_ids<RTB_TRACER_TREE><<<blocks, 128, 0, st>>>(sc, cam, d_tri, d_obj);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int launch_prim_ids(const DevScene &sc, const FrameArgs &fa, int32_t *d_tri, int32_t *d_obj, int sm_count, cudaStream_t st) {
  static_assert(RAST_WORDS <= RAST_STATE_WORDS, "raster state");
  RTB_CUDA(cudaMemsetAsync(fa.prim, 0xff, (size_t)fa.cam.W * fa.cam.H * 8, st));
  RTB_CUDA(cudaMemsetAsync(fa.rast, 0, RAST_STATE_WORDS * 4, st));
  if (sc.n_tris) {
    k_rast_setup<<<(sc.n_tris + 255) / 256, 256, 0, st>>>(sc, fa);
    k_rast_big<<<sm_count * 8, RAST_BIG_THREADS, 0, st>>>(sc, fa);
    k_rast_items<<<sm_count * 8, 256, 0, st>>>(sc, fa);
  }
  const size_t n = (size_t)fa.cam.W * fa.cam.H;
  if (n) k_prim_ids<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(sc, fa, d_tri, d_obj);
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int launch_trace_rays(const DevScene &sc, int tracer, int any_hit, const float *d_rays, size_t n, int32_t *d_tri,
                      float *d_dist, cudaStream_t st) {
  if (!n) return RTB_OK;
  const unsigned blocks = (unsigned)((n + 127) / 128);
  if (tracer == RTB_TRACER_BRUTE) {
    if (any_hit) k_trace_rays<RTB_TRACER_BRUTE, true><<<blocks, 128, 0, st>>>(sc, d_rays, n, d_tri, d_dist);
    else k_trace_rays<RTB_TRACER_BRUTE, false><<<blocks, 128, 0, st>>>(sc, d_rays, n, d_tri, d_dist);
  } else {
    if (any_hit) k_trace_rays<RTB_TRACER_TREE, true><<<blocks, 128, 0, st>>>(sc, d_rays, n, d_tri, d_dist);
    else k_trace_rays<RTB_TRACER_TREE, false><<<blocks, 128, 0, st>>>(sc, d_rays, n, d_tri, d_dist);
  }
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

}  // namespace rtb
