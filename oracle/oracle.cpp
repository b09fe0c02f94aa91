/*
 * oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's GPU-semantics ray tracer, used only by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs as the
 * checker.  Nothing under raytracing-gpu_b200/ may link, import or call this file.
 *
 * Pinning: the reference ships no golden images.  This restatement is pinned
 * against the reference's OWN device sources compiled for the host
 * (oracle/Makefile -> oracle/_ref/libref_gpusem.so, driven by oracle/ref_driver.cpp):
 * tests/test_oracle_pin.py demands bit-identical RGBA on the shipped scenes, and
 * tests/golden/ holds hit-ID + RGBA fixtures generated from that build
 * (tests/golden/make_golden.py).
 *
 * Semantics followed (all binary32, evaluation order as written, no FMA contraction —
 * build with -ffp-contract=off):
 *   camera / primary rays   gpu/raytracer.cu:178-182 (host), :87-110 (kernel)
 *   bounce loop             gpu/raytracer.cu:107-122, trace :32-47, ray_bounce gpu/ray.cu:16-25
 *   ray/triangle            gpu/hit.cu:8-48 (ray_intersect)
 *   nearest hit             gpu/hit.cu:50-78 (triangle_collide), :84-114 (collide, PARTITIONING_NONE)
 *   shadow rays             gpu/light.cu:29-36 (has_direct_hit), gpu/hit.cu:241-253 (collide_dist)
 *   shading                 gpu/light.cu:12-27 (apply_specular), :50-127 (apply_light)
 *   colours                 gpu/colors.cu:3-47
 *   SSAA resolve            gpu/raytracer.cu:49-85 (downscale)
 *   vector ops              gpu/vector3.cu:6-48, gpu/vector3-extern.cu:5-23
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#include "../include/rtb200.h"

namespace {

struct V3 { float x, y, z; };
struct Ray { V3 o, d; };
struct Col { uint8_t r, g, b, a; };

inline V3 ld(const vector3 &v) { return V3{v.x, v.y, v.z}; }
inline V3 vsub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }           /* vector3.cu:6 */
inline V3 vadd(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }           /* vector3.cu:11 */
inline V3 vcross(V3 a, V3 b) {                                                       /* vector3.cu:22 */
  return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline V3 vscale(V3 a, float r) { return V3{r * a.x, r * a.y, r * a.z}; }            /* vector3.cu:31 */
inline float vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }          /* vector3-extern.cu:5 */
inline float vlen(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }         /* vector3-extern.cu:10 */
inline V3 vnorm(V3 a) { float r = vlen(a); return V3{a.x / r, a.y / r, a.z / r}; }   /* vector3.cu:40 */
inline bool vzero(V3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }                 /* vector3-extern.cu:20 */

/* colors.cu:3-21 */
inline uint8_t quant(float x) {
  x = x * 255;
  if (x > 255) x = 255;
  if (x < 0) x = 0;
  /* CUDA's float->u8 conversion gives 0 for NaN; x86 cvttss2si gives INT_MIN -> low byte 0 */
  if (x != x) return 0;
  return (uint8_t)x;
}
inline Col init_color(float r, float g, float b) { return Col{quant(r), quant(g), quant(b), 255}; }
inline Col color_add(Col a, Col b) {                                                  /* colors.cu:23-35 */
  int r = a.r + b.r, g = a.g + b.g, bl = a.b + b.b;
  return Col{(uint8_t)(r > 255 ? 255 : r), (uint8_t)(g > 255 ? 255 : g), (uint8_t)(bl > 255 ? 255 : bl), 255};
}
inline Col color_mul(Col a, float k) {                                                /* colors.cu:37-40 */
  return init_color(float(a.r) / 255 * k, float(a.g) / 255 * k, float(a.b) / 255 * k);
}
inline Col color_mults(Col a, Col b) {                                                /* colors.cu:42-47 */
  return init_color((float(a.r) / 255) * (float(b.r) / 255), (float(a.g) / 255) * (float(b.g) / 255),
                    (float(a.b) / 255) * (float(b.b) / 255));
}

struct Hit {
  bool hit = false;
  V3 p{0, 0, 0};      /* hit point  (ray.origin of the reference's hit record) */
  V3 n{0, 0, 0};      /* interpolated normal (ray.direction of the record) */
  int obj = -1;
  int64_t tri = -1;   /* global triangle index, host memory order */
  float dist = 0;
};

struct Counters { uint64_t primary = 0, shadow = 0, bounce = 0; };

/* hit.cu:20-41 without the normals (they do not influence acceptance) */
inline bool mt_test(const Ray &ray, const vector3 *vtx, V3 *out, float *uo, float *vo) {
  const float EPS = 0.0000001f;
  V3 v0 = ld(vtx[0]), v1 = ld(vtx[1]), v2 = ld(vtx[2]);
  V3 e1 = vsub(v1, v0), e2 = vsub(v2, v0);
  V3 h = vcross(ray.d, e2);
  float a = vdot(e1, h);
  if (a > -EPS && a < EPS) return false;
  float f = 1 / a;
  V3 s = vsub(ray.o, v0);
  float u = f * vdot(s, h);
  if (u < 0.0f || u > 1.0f) return false;
  V3 q = vcross(s, e1);
  float v = f * vdot(ray.d, q);
  if (v < 0.0f || u + v > 1.0f) return false;
  float t = f * vdot(e2, q);
  if (!(t > EPS)) return false;
  V3 t2 = vscale(vnorm(ray.d), t * vlen(ray.d));
  *out = vadd(ray.o, t2);
  *uo = u;
  *vo = v;
  return true;
}

/* hit.cu:15-17,42-44 */
inline V3 interp_normal(const vector3 *nrm, float u, float v) {
  V3 n0 = vnorm(ld(nrm[0])), n1 = vnorm(ld(nrm[1])), n2 = vnorm(ld(nrm[2]));
  return vadd(vadd(vscale(n0, 1 - u - v), vscale(n1, u)), vscale(n2, v));
}

/* triangle_collide, hit.cu:50-78.  any_hit: return at the first candidate (legal for
 * shadow rays, SURVEY A.5: occlusion == "some candidate exists"). */
Hit object_nearest(const struct object &o, const Ray &ray, bool any_hit) {
  Hit best;
  float distance = 0;
  float bu = 0, bv = 0;
  uint32_t bi = 0;
  for (uint32_t i = 0; i < o.triangle_count; ++i) {
    V3 out;
    float u, v;
    if (!mt_test(ray, o.triangles.vertex + 3 * (size_t)i, &out, &u, &v)) continue;
    float nd = vlen(vsub(out, ray.o));
    if (nd > 0.01f && (nd < distance || distance == 0)) {
      distance = nd;
      best.hit = true;
      best.p = out;
      best.dist = nd;
      bu = u; bv = v; bi = i;
      if (any_hit) break;
    }
  }
  if (best.hit) {
    best.n = interp_normal(o.triangles.normal + 3 * (size_t)bi, bu, bv);
    best.tri = bi;
  }
  return best;
}

/* collide, hit.cu:84-114 (PARTITIONING_NONE: the semantic definition of nearest hit) */
Hit scene_nearest(const struct scene *sc, const Ray &ray, bool any_hit) {
  Hit best;
  float distance = 0;
  int64_t first_tri = 0;
  for (size_t i = 0; i < sc->object_count; ++i) {
    const struct object &o = sc->objects[i];
    Hit h = object_nearest(o, ray, any_hit);
    /* a record whose normal is exactly zero counts as "no hit in this object" (hit.cu:100) */
    if (h.hit && !vzero(h.n)) {
      float nd = vlen(vsub(h.p, ray.o));
      if (nd > 0.01f && (nd < distance || distance == 0)) {
        distance = nd;
        best = h;
        best.obj = (int)i;
        best.tri = first_tri + h.tri;
        best.dist = nd;
        if (any_hit) return best;
      }
    }
    first_tri += o.triangle_count;
  }
  return best;
}

/* has_direct_hit, light.cu:29-36: occluded iff collide_dist != 0 */
inline bool occluded(const struct scene *sc, const Ray &r, Counters &c) {
  c.shadow++;
  return scene_nearest(sc, r, true).hit;
}

/* light.cu:12-27 */
inline void apply_specular(Col *color, const Ray &incident, V3 P, V3 N, const struct object &o) {
  Col k = init_color(o.ks.x, o.ks.y, o.ks.z);
  V3 V = vsub(incident.o, P);
  V3 R = vsub(incident.d, vscale(N, 2 * vdot(N, incident.d)));
  R = vnorm(R);
  V = vnorm(V);
  float d = vdot(R, V);
  float Ls = powf(d > 0.0f ? d : 0.0f, o.ns);
  k = color_mul(k, Ls);
  *color = color_add(*color, k);
}

/* light.cu:50-127 */
Col apply_light(const struct scene *sc, const struct object &o, V3 P, V3 Nrm, Counters &c) {
  Col color = init_color(0, 0, 0);
  for (size_t i = 0; i < sc->light_count; ++i) {
    const struct light &l = sc->lights[i];
    V3 lv = ld(l.v);
    if (l.type == AMBIENT) {
      Col t = color_mults(init_color(l.r, l.g, l.b), init_color(o.ka.x, o.ka.y, o.ka.z));
      color = color_add(color, t);
    } else if (l.type == DIRECTIONAL) {
      Ray lr{P, vscale(lv, -1)};
      if (!occluded(sc, lr, c)) {
        V3 L = vscale(lv, -1);
        Col t = color_mults(init_color(l.r, l.g, l.b), init_color(o.kd.x, o.kd.y, o.kd.z));
        t = color_mul(t, vdot(L, Nrm));
        Ray inc;
        inc.d = lv;
        inc.o = vadd(P, vscale(lv, -10));
        apply_specular(&t, inc, P, Nrm, o);
        color = color_add(color, t);
      }
    } else if (l.type == POINT) {
      V3 L = vscale(lv, -1);
      V3 N = Nrm;
      if (vdot(L, N) < 0) N = vscale(N, -1);
      Ray lr{P, vsub(lv, P)};
      float dist = vlen(vsub(lv, P));
      if (!occluded(sc, lr, c)) {
        Col t = color_mults(init_color(l.r, l.g, l.b), init_color(o.kd.x, o.kd.y, o.kd.z));
        t = color_mul(t, vdot(L, N) * 1 / dist);
        Ray inc;
        inc.d = vsub(lv, P);
        inc.o = vadd(P, vscale(inc.d, -10));
        apply_specular(&t, inc, P, Nrm /* unflipped: `point` is passed, light.cu:115 */, o);
        color = color_add(color, t);
      }
    }
  }
  return color;
}

struct CameraFrame { V3 u, v, C, pos; int W, H; };

/* host part of render(), raytracer.cu:178-182 */
CameraFrame camera_frame(const struct camera &cam) {
  CameraFrame f;
  f.u = vnorm(ld(cam.u));
  f.v = vnorm(ld(cam.v));
  V3 w = vcross(f.u, f.v);
  float L = (float)(cam.width / (2 * tan(cam.fov * M_PI / 360)));
  f.pos = ld(cam.position);
  f.C = vadd(f.pos, vscale(w, L));
  f.W = cam.width;
  f.H = cam.height;
  return f;
}

/* raytrace kernel body, raytracer.cu:87-129, for sample (px,py) */
Col shade_sample(const struct scene *sc, const CameraFrame &f, int px, int py, int32_t *tri_id,
                 int32_t *obj_id, Counters &c) {
  V3 ui = vscale(f.u, (float)(px - f.W / 2));
  V3 vj = vscale(f.v, (float)(py - f.H / 2));
  V3 point = vadd(vadd(f.C, ui), vj);
  Ray ray{point, vnorm(vsub(f.pos, point))};
  float nr_ = 1, r = 1;
  int budget = 10;
  Col color = init_color(0, 0, 0);
  bool first = true;
  do {
    /* trace(), raytracer.cu:32-47 */
    if (first) c.primary++; else c.bounce++;
    Hit h = scene_nearest(sc, ray, false);
    if (first) {
      if (tri_id) *tri_id = h.hit ? (int32_t)h.tri : -1;
      if (obj_id) *obj_id = h.hit ? h.obj : -1;
      first = false;
    }
    Col tc;
    if (h.hit) {
      const struct object &o = sc->objects[h.obj];
      tc = apply_light(sc, o, h.p, h.n, c);
      if (o.nr > 0) {
        Ray refl;
        refl.o = h.p;
        refl.d = vsub(ray.d, vscale(h.n, 2 * vdot(h.n, ray.d)));   /* ray.cu:16-25 */
        ray = refl;
      }
      r = o.nr;
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

template <class F>
void parallel_rows(int rows, int nthreads, F fn) {
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  if (nthreads <= 0) nthreads = 1;
  if (nthreads > rows) nthreads = rows > 0 ? rows : 1;
  std::atomic<int> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t)
    pool.emplace_back([&, t]() {
      for (;;) {
        int r = next.fetch_add(1);
        if (r >= rows) break;
        fn(r, t);
      }
    });
  for (auto &th : pool) th.join();
}

}  // namespace

extern "C" {

/* Full frame with GPU semantics.  scene->camera.width/height are the SAMPLE grid
 * (already multiplied by aliasing, like render()).  rgba: out_h*out_w*4 tight.
 * tri_ids/obj_ids: optional, W'*H' each, indexed [py*W'+px].  counts[3]: primary,
 * shadow, bounce.  hires: optional W'*H' RGBA8 of the per-sample colours indexed like
 * the ids (not flipped). */
int orc_render(const struct scene *sc, int aliasing, int out_w, int out_h, uint8_t *rgba,
               int32_t *tri_ids, int32_t *obj_ids, uint8_t *hires, uint64_t *counts, int nthreads) {
  CameraFrame f = camera_frame(sc->camera);
  if (f.W != out_w * aliasing || f.H != out_h * aliasing) return -1;
  std::vector<Col> hi((size_t)f.W * f.H);
  int nt = nthreads > 0 ? nthreads : (int)std::thread::hardware_concurrency();
  if (nt <= 0) nt = 1;
  std::vector<Counters> cnt((size_t)nt);
  parallel_rows(f.H, nt, [&](int py, int t) {
    for (int px = 0; px < f.W; ++px) {
      size_t k = (size_t)py * f.W + px;
      hi[k] = shade_sample(sc, f, px, py, tri_ids ? tri_ids + k : nullptr, obj_ids ? obj_ids + k : nullptr, cnt[t]);
    }
  });
  if (hires) memcpy(hires, hi.data(), hi.size() * 4);
  /* downscale kernel, raytracer.cu:49-85 */
  if (rgba) {
    for (int ly = 0; ly < out_h; ++ly)
      for (int lx = 0; lx < out_w; ++lx) {
        float r = 0, g = 0, b = 0;
        for (int py = aliasing * ly; py < aliasing * ly + aliasing; ++py)
          for (int px = aliasing * lx; px < aliasing * lx + aliasing; ++px) {
            Col c = hi[(size_t)py * f.W + px];
            r += (float)c.r;
            g += (float)c.g;
            b += (float)c.b;
          }
        float ali2 = 255.0f * (float)aliasing * (float)aliasing;
        r /= ali2;
        g /= ali2;
        b /= ali2;
        Col e = init_color(r, g, b);
        uint8_t *dst = rgba + ((size_t)(out_h - ly - 1) * out_w + (out_w - lx - 1)) * 4;
        dst[0] = e.r; dst[1] = e.g; dst[2] = e.b; dst[3] = e.a;
      }
  }
  if (counts) {
    counts[0] = counts[1] = counts[2] = 0;
    for (auto &c : cnt) { counts[0] += c.primary; counts[1] += c.shadow; counts[2] += c.bounce; }
  }
  return 0;
}

/* Sparse samples (for scenes where a full brute-force frame is out of reach):
 * xy = n pairs (px,py) on the sample grid; color_out = n RGBA8 sample colours. */
int orc_render_samples(const struct scene *sc, const int32_t *xy, size_t n, int32_t *tri_ids,
                       int32_t *obj_ids, uint8_t *color_out, uint64_t *counts, int nthreads) {
  CameraFrame f = camera_frame(sc->camera);
  int nt = nthreads > 0 ? nthreads : (int)std::thread::hardware_concurrency();
  if (nt <= 0) nt = 1;
  std::vector<Counters> cnt((size_t)nt);
  parallel_rows((int)n, nt, [&](int i, int t) {
    Col c = shade_sample(sc, f, xy[2 * i], xy[2 * i + 1], tri_ids ? tri_ids + i : nullptr,
                         obj_ids ? obj_ids + i : nullptr, cnt[t]);
    if (color_out) memcpy(color_out + 4 * (size_t)i, &c, 4);
  });
  if (counts) {
    counts[0] = counts[1] = counts[2] = 0;
    for (auto &c : cnt) { counts[0] += c.primary; counts[1] += c.shadow; counts[2] += c.bounce; }
  }
  return 0;
}

/* collide() / collide_dist() on arbitrary rays: rays = n x 6 floats (origin, direction). */
int orc_trace_rays(const struct scene *sc, int any_hit, const float *rays, size_t n, int32_t *tri_out,
                   float *dist_out, float *hit_out /* n x 6: point, normal; zeros on miss */, int nthreads) {
  parallel_rows((int)n, nthreads, [&](int i, int) {
    Ray r{V3{rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]}, V3{rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]}};
    Hit h = scene_nearest(sc, r, any_hit != 0);
    if (tri_out) tri_out[i] = h.hit ? (any_hit ? 0 : (int32_t)h.tri) : -1;
    if (dist_out) dist_out[i] = h.hit ? h.dist : 0.0f;
    if (hit_out) {
      float *o = hit_out + 6 * (size_t)i;
      o[0] = h.p.x; o[1] = h.p.y; o[2] = h.p.z; o[3] = h.n.x; o[4] = h.n.y; o[5] = h.n.z;
    }
  });
  return 0;
}

/* Primary-ray hit records (point, normal) + ids for the whole sample grid; compared
 * bitwise with the reference's collide() in tests/test_oracle_pin.py. */
int orc_primary_hits(const struct scene *sc, float *hit_out, int32_t *tri_ids, int nthreads) {
  CameraFrame f = camera_frame(sc->camera);
  parallel_rows(f.H, nthreads, [&](int py, int) {
    for (int px = 0; px < f.W; ++px) {
      V3 ui = vscale(f.u, (float)(px - f.W / 2));
      V3 vj = vscale(f.v, (float)(py - f.H / 2));
      V3 point = vadd(vadd(f.C, ui), vj);
      Ray ray{point, vnorm(vsub(f.pos, point))};
      Hit h = scene_nearest(sc, ray, false);
      size_t k = (size_t)py * f.W + px;
      if (tri_ids) tri_ids[k] = h.hit ? (int32_t)h.tri : -1;
      if (hit_out) {
        float *o = hit_out + 6 * k;
        o[0] = h.p.x; o[1] = h.p.y; o[2] = h.p.z; o[3] = h.n.x; o[4] = h.n.y; o[5] = h.n.z;
      }
    }
  });
  return 0;
}

/* Colour helpers exposed for unit tests of the u8 pipeline (SURVEY App. D probes). */
uint32_t orc_init_color(float r, float g, float b) { Col c = init_color(r, g, b); uint32_t w; memcpy(&w, &c, 4); return w; }
uint32_t orc_color_mul(uint32_t a, float k) { Col c; memcpy(&c, &a, 4); c = color_mul(c, k); uint32_t w; memcpy(&w, &c, 4); return w; }
uint32_t orc_color_mults(uint32_t a, uint32_t b) { Col x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4); x = color_mults(x, y); uint32_t w; memcpy(&w, &x, 4); return w; }
uint32_t orc_color_add(uint32_t a, uint32_t b) { Col x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4); x = color_add(x, y); uint32_t w; memcpy(&w, &x, 4); return w; }

}  /* extern "C" */
