/*
 * refcpu_driver.c — TEST INFRASTRUCTURE.  Driver around the REFERENCE's CPU renderer
 * sources (cpu/{hit,light,colors,ray,vector3,vector3-extern}.c), compiled unmodified from
 * /root/reference by oracle/Makefile into oracle/_ref/libref_cpu.so.
 *   refcpu_collide       — cpu/hit.c:72 collide() on caller-supplied rays: the bitwise
 *                          geometric cross-check of the oracle ("Oracle-C").
 *   refcpu_trace_samples — cpu/raytracer.c:19-34 trace() (restated: it is static) over
 *                          caller-chosen samples on N pthreads: the timed CPU baseline
 *                          for scenes where a full cpu/rt frame is out of reach (C5).
 * Input scenes arrive in the GPU ABI layout (SoA vertex/normal arrays) and are
 * converted to the CPU copy's AoS `struct triangle`.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "scene.h"
#include "ray.h"
#include "hit.h"
#include "light.h"
#include "colors.h"
#include "vector3.h"

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

/* GPU-ABI mirror (include/rtb200.h), kept local so the reference headers own the names */
struct abi_vec3 { float x, y, z; };
struct abi_object {
  struct abi_vec3 *vertex, *normal;
  uint32_t triangle_count;
  struct abi_vec3 ka, kd, ks;
  float ns, ni, nr, d;
};
struct abi_light { int type; float r, g, b; struct abi_vec3 v; };
struct abi_camera { int width, height; struct abi_vec3 position, u, v; float fov; };
struct abi_scene {
  struct abi_vec3 *vertex, *normal;
  struct abi_object *objects;
  size_t object_count;
  struct abi_light *lights;
  size_t light_count;
  struct abi_camera camera;
  size_t triangle_count;
  void *aabbs, *octree;
};

static struct vector3 cv(struct abi_vec3 v) { struct vector3 r; r.x = v.x; r.y = v.y; r.z = v.z; return r; }

void *refcpu_scene_create(const struct abi_scene *in) {
  struct scene *s = calloc(1, sizeof *s);
  s->object_count = in->object_count;
  s->light_count = in->light_count;
  s->objects = calloc(in->object_count ? in->object_count : 1, sizeof(struct object));
  s->lights = calloc(in->light_count ? in->light_count : 1, sizeof(struct light));
  for (size_t i = 0; i < in->object_count; i++) {
    const struct abi_object *o = &in->objects[i];
    struct object *d = &s->objects[i];
    d->triangle_count = o->triangle_count;
    d->triangles = malloc(sizeof(struct triangle) * (o->triangle_count ? o->triangle_count : 1));
    for (uint32_t t = 0; t < o->triangle_count; t++)
      for (int j = 0; j < 3; j++) {
        d->triangles[t].vertex[j] = cv(o->vertex[3 * (size_t)t + j]);
        d->triangles[t].normal[j] = cv(o->normal[3 * (size_t)t + j]);
      }
    d->ka = cv(o->ka); d->kd = cv(o->kd); d->ks = cv(o->ks);
    d->ns = o->ns; d->ni = o->ni; d->nr = o->nr; d->d = o->d;
  }
  for (size_t i = 0; i < in->light_count; i++) {
    s->lights[i].type = (enum light_type)in->lights[i].type;
    s->lights[i].r = in->lights[i].r; s->lights[i].g = in->lights[i].g; s->lights[i].b = in->lights[i].b;
    s->lights[i].v = cv(in->lights[i].v);
  }
  s->camera.width = in->camera.width; s->camera.height = in->camera.height;
  s->camera.position = cv(in->camera.position); s->camera.u = cv(in->camera.u); s->camera.v = cv(in->camera.v);
  s->camera.fov = in->camera.fov;
  return s;
}

void refcpu_scene_destroy(void *h) {
  struct scene *s = h;
  if (!s) return;
  for (size_t i = 0; i < s->object_count; i++) free(s->objects[i].triangles);
  free(s->objects); free(s->lights); free(s);
}

/* hit_out: n x 6 floats (hit point, normal) exactly as collide() returns them */
int refcpu_collide(void *h, const float *rays, size_t n, float *hit_out) {
  struct scene *s = h;
  for (size_t i = 0; i < n; i++) {
    struct ray r;
    r.origin.x = rays[6 * i]; r.origin.y = rays[6 * i + 1]; r.origin.z = rays[6 * i + 2];
    r.direction.x = rays[6 * i + 3]; r.direction.y = rays[6 * i + 4]; r.direction.z = rays[6 * i + 5];
    struct object obj;
    struct ray hit = collide(*s, r, &obj);
    float *o = hit_out + 6 * i;
    o[0] = hit.origin.x; o[1] = hit.origin.y; o[2] = hit.origin.z;
    o[3] = hit.direction.x; o[4] = hit.direction.y; o[5] = hit.direction.z;
  }
  return 0;
}

/* cpu/raytracer.c:19-34 */
/* rays[0] += 1 per collide() issued here (primary / reflection), rays[1] += 1 per shadow
 * ray apply_light casts (one per DIRECTIONAL / POINT light, cpu/light.c:50-110) */
static struct color cpu_trace(struct scene scene, struct ray ray, float coef, uint64_t *rays) {
  if (coef < 0.01)
    return init_color(0, 0, 0);
  struct object obj;
  rays[0]++;
  struct ray new_ray = collide(scene, ray, &obj);
  if (!vector3_is_zero(new_ray.direction)) {
    for (size_t i = 0; i < scene.light_count; i++)
      if (scene.lights[i].type == DIRECTIONAL || scene.lights[i].type == POINT) rays[1]++;
    struct color object = apply_light(scene, obj, new_ray);
    struct ray reflection_ray = ray_bounce(ray, new_ray);
    struct color reflection = cpu_trace(scene, reflection_ray, obj.nr * coef, rays);
    return color_add(reflection, color_mul(object, coef));
  }
  return init_color(0, 0, 0);
}

struct job { struct scene *s; const float *xy; size_t begin, end; float *rgb; struct vector3 u, v, C; uint64_t rays[2]; };

static void *worker(void *p) {
  struct job *j = p;
  for (size_t i = j->begin; i < j->end; i++) {
    /* cpu/raytracer.c:57-64 for the sub-sample position (k,l) = xy[i] */
    struct vector3 ui = vector3_scale(j->u, j->xy[2 * i]);
    struct vector3 vj = vector3_scale(j->v, j->xy[2 * i + 1]);
    struct vector3 point = vector3_add(vector3_add(j->C, ui), vj);
    struct ray ray;
    ray.origin = point;
    ray.direction = vector3_normalize(vector3_sub(j->s->camera.position, point));
    struct color c = cpu_trace(*j->s, ray, 1, j->rays);
    j->rgb[3 * i] = c.r; j->rgb[3 * i + 1] = c.g; j->rgb[3 * i + 2] = c.b;
  }
  return NULL;
}

/* xy: n (k,l) image-plane coordinates in the CPU renderer's convention
 * (cpu/raytracer.c:50-58: pixel offsets from the image centre, step 0.5). */
int refcpu_trace_samples(void *h, const float *xy, size_t n, float *rgb_out, int nthreads, uint64_t *rays_out) {
  struct scene *s = h;
  if (nthreads < 1) nthreads = 1;
  if (nthreads > 256) nthreads = 256;
  struct vector3 u = vector3_normalize(s->camera.u);            /* cpu/raytracer.c:82-86 */
  struct vector3 v = vector3_normalize(s->camera.v);
  struct vector3 w = vector3_cross(u, v);
  float L = s->camera.width / (2 * tan(s->camera.fov * M_PI / 360));
  struct vector3 C = vector3_add(s->camera.position, vector3_scale(w, L));
  pthread_t tid[256];
  struct job jobs[256];
  for (int t = 0; t < nthreads; t++) {
    jobs[t].s = s; jobs[t].xy = xy; jobs[t].rgb = rgb_out; jobs[t].u = u; jobs[t].v = v; jobs[t].C = C;
    jobs[t].begin = n * (size_t)t / (size_t)nthreads;
    jobs[t].end = n * (size_t)(t + 1) / (size_t)nthreads;
    jobs[t].rays[0] = jobs[t].rays[1] = 0;
    pthread_create(&tid[t], NULL, worker, &jobs[t]);
  }
  for (int t = 0; t < nthreads; t++) pthread_join(tid[t], NULL);
  if (rays_out) {
    rays_out[0] = rays_out[1] = 0;
    for (int t = 0; t < nthreads; t++) { rays_out[0] += jobs[t].rays[0]; rays_out[1] += jobs[t].rays[1]; }
  }
  return 0;
}
