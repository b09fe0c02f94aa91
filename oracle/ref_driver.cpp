/*
 * ref_driver.cpp — TEST INFRASTRUCTURE.  Host driver around the REFERENCE's own device
 * sources (gpu/{vector3,vector3-extern,colors,ray,light,hit,scene}.cu), which
 * oracle/Makefile compiles as plain host C++ (-DLAYOUT_SOA -DPARTITIONING_NONE,
 * -ffp-contract=off) straight from /root/reference into oracle/_ref/libref_gpusem.so.
 * All arithmetic executed here is the reference's: collide, apply_light, ray_bounce,
 * init_color, color_mul, color_add.  Only the two __global__ kernel bodies
 * (gpu/raytracer.cu:87-129 raytrace incl. static trace :32-47, :49-85 downscale) and the
 * host camera set-up (:178-182) are restated as loops, because raytracer.cu itself
 * needs spdlog and a device.
 * Used to pin oracle/oracle.cpp and to generate tests/golden/.
 */
#include <cmath>
#include <cstring>
#include <atomic>
#include <thread>
#include <vector>
#include "scene.h"
#include "hit.h"
#include "light.h"
#include "ray.h"
#include "colors.h"
#include "vector3.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

namespace {

/* trace(), gpu/raytracer.cu:32-47 */
struct color ref_trace(struct scene *scene, struct ray ray, struct ray *reflection, float *loc_nr) {
  struct object obj;
  struct ray new_ray = collide(scene, ray, &obj);
  if (!vector3_is_zero(new_ray.direction)) {
    struct color object = apply_light(scene, &obj, new_ray);
    if (obj.nr > 0) *reflection = ray_bounce(&ray, &new_ray);
    *loc_nr = obj.nr;
    return object;
  }
  *loc_nr = 0;
  return init_color(0, 0, 0);
}

struct Frame { vector3 u, v, C; };

Frame frame_of(const struct scene *s) {
  Frame f;
  f.u = vector3_normalize(s->camera.u);   /* host_v_norm == vector3_normalize op-for-op */
  f.v = vector3_normalize(s->camera.v);
  vector3 w = vector3_cross(f.u, f.v);
  float L = s->camera.width / (2 * tan(s->camera.fov * M_PI / 360));
  f.C = vector3_add(s->camera.position, vector3_scale(w, L));
  return f;
}

/* raytrace kernel body for one sample */
struct color sample_color(struct scene *scene, const Frame &f, int px, int py) {
  int width = scene->camera.width, height = scene->camera.height;
  vector3 ui = vector3_scale(f.u, px - width / 2);
  vector3 vj = vector3_scale(f.v, py - height / 2);
  vector3 point = vector3_add(vector3_add(f.C, ui), vj);
  vector3 direction = vector3_normalize(vector3_sub(scene->camera.position, point));
  struct ray ray;
  float nr_ = 1;
  float r_rn = nr_;
  ray.origin = point;
  ray.direction = direction;
  int MAX_BOUNCE = 10;
  struct color color = init_color(0, 0, 0);
  struct color tmp_color;
  do {
    tmp_color = ref_trace(scene, ray, &ray, &r_rn);
    tmp_color = color_mul(&tmp_color, nr_);
    color = color_add(&color, &tmp_color);
    nr_ *= r_rn;
  } while (nr_ > 0.01f && MAX_BOUNCE-- > 0);
  return color;
}

template <class F>
void rows(int n, int nthreads, F fn) {
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  if (nthreads <= 0) nthreads = 1;
  std::atomic<int> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t)
    pool.emplace_back([&]() {
      for (;;) {
        int r = next.fetch_add(1);
        if (r >= n) break;
        fn(r);
      }
    });
  for (auto &th : pool) th.join();
}

}  // namespace

extern "C" {

/* `scene_abi` is a scene in the SOA+OCTREE ABI layout (120 bytes); with PARTITIONING_NONE
 * the reference's struct scene is its 104-byte prefix, so the fields are copied over. */
int ref_render(const void *scene_abi, int aliasing, int out_w, int out_h, unsigned char *rgba,
               unsigned char *hires, int nthreads) {
  struct scene sc;
  memcpy(&sc, scene_abi, sizeof sc);
  int W = sc.camera.width, H = sc.camera.height;
  if (W != out_w * aliasing || H != out_h * aliasing) return -1;
  Frame f = frame_of(&sc);
  std::vector<struct color> hi((size_t)W * H);
  rows(H, nthreads, [&](int py) {
    for (int px = 0; px < W; ++px) hi[(size_t)py * W + px] = sample_color(&sc, f, px, py);
  });
  if (hires) memcpy(hires, hi.data(), hi.size() * 4);
  if (rgba)
    for (int py = 0; py < out_h; ++py)        /* downscale kernel body */
      for (int px = 0; px < out_w; ++px) {
        float r = 0, g = 0, b = 0;
        for (int h_py = aliasing * py; h_py < aliasing * py + aliasing; ++h_py)
          for (int h_px = aliasing * px; h_px < aliasing * px + aliasing; ++h_px) {
            struct color tmp = hi[(size_t)h_py * W + h_px];
            r += (float)tmp.r;
            g += (float)tmp.g;
            b += (float)tmp.b;
          }
        float ali2 = 255.0f * (float)aliasing * (float)aliasing;
        r /= ali2;
        g /= ali2;
        b /= ali2;
        struct color end_color = init_color(r, g, b);
        memcpy(rgba + ((size_t)(out_h - py - 1) * out_w + (out_w - px - 1)) * 4, &end_color, 4);
      }
  return 0;
}

/* reference collide() on arbitrary rays; hit_out = n x 6 floats (hit point, normal),
 * all zero on a miss. */
int ref_collide(const void *scene_abi, const float *rays, size_t n, float *hit_out, int nthreads) {
  struct scene sc;
  memcpy(&sc, scene_abi, sizeof sc);
  rows((int)n, nthreads, [&](int i) {
    struct ray r;
    r.origin = make_float3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    r.direction = make_float3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
    struct object obj;
    struct ray h = collide(&sc, r, &obj);
    float *o = hit_out + 6 * (size_t)i;
    o[0] = h.origin.x; o[1] = h.origin.y; o[2] = h.origin.z;
    o[3] = h.direction.x; o[4] = h.direction.y; o[5] = h.direction.z;
  });
  return 0;
}

/* primary-ray hit records of the whole sample grid (for hit-ID pinning through the
 * bit pattern of point+normal) */
int ref_primary_hits(const void *scene_abi, float *hit_out, int nthreads) {
  struct scene sc;
  memcpy(&sc, scene_abi, sizeof sc);
  int W = sc.camera.width, H = sc.camera.height;
  Frame f = frame_of(&sc);
  rows(H, nthreads, [&](int py) {
    for (int px = 0; px < W; ++px) {
      vector3 ui = vector3_scale(f.u, px - W / 2);
      vector3 vj = vector3_scale(f.v, py - H / 2);
      vector3 point = vector3_add(vector3_add(f.C, ui), vj);
      struct ray r;
      r.origin = point;
      r.direction = vector3_normalize(vector3_sub(sc.camera.position, point));
      struct object obj;
      struct ray h = collide(&sc, r, &obj);
      float *o = hit_out + 6 * ((size_t)py * W + px);
      o[0] = h.origin.x; o[1] = h.origin.y; o[2] = h.origin.z;
      o[3] = h.direction.x; o[4] = h.direction.y; o[5] = h.direction.z;
    }
  });
  return 0;
}

}  /* extern "C" */
