// Host scene container behind the reference's scene API
// (gpu/headers/scene.h:40-61, semantics of gpu/scene.cu:39-47,146-218, SOA layout).
//
// Storage: one scene-global vertex[] and normal[] array, 3 consecutive vector3 per
// triangle, objects appended contiguously in insertion order.  Unlike the reference
// (one realloc per object) the arrays grow geometrically: capacity is a pure function
// of the triangle count, so no extra field is needed in the ABI struct.
#include <cstdlib>
#include <cstring>
#include "rt_host.h"

// Vertex / normal arrays are the bulk of what render() uploads.  When the library that
// hosts this file also provides page-locked allocation (librtb200.so: api.cu), the arrays
// live in pinned memory and the upload runs at PCIe speed instead of being staged by the
// driver; the host-only build (librt_host.so) has no such hook and uses malloc.
extern "C" void *rtb_scene_mem_alloc(size_t bytes) __attribute__((weak));
extern "C" void rtb_scene_mem_free(void *p) __attribute__((weak));

namespace {

// Every bulk array carries its capacity (in vector3) in a 64-byte header in front of the
// data, so the ABI struct needs no extra field and a caller that knows the final size can
// reserve once (rtb_scene_reserve) instead of growing by doubling.
constexpr size_t BULK_HEADER = 64;

// Page-locking costs about a second per GB: worth it for a scene that is uploaded many times
// (bench, fly-through), not for the loader of a one-shot `rt` run, which asks for plain memory.
thread_local bool t_pageable = false;

void *raw_alloc(size_t bytes) {
  return rtb_scene_mem_alloc && !t_pageable ? rtb_scene_mem_alloc(bytes) : std::malloc(bytes);
}
void raw_free(void *p) {
  if (rtb_scene_mem_free) rtb_scene_mem_free(p);
  else std::free(p);
}
size_t bulk_capacity(const vector3 *p) {
  return p ? *reinterpret_cast<const size_t *>(reinterpret_cast<const char *>(p) - BULK_HEADER) : 0;
}
void bulk_free(vector3 *p) {
  if (p) raw_free(reinterpret_cast<char *>(p) - BULK_HEADER);
}
vector3 *bulk_grow(vector3 *old, size_t old_count, size_t new_count) {
  char *base = static_cast<char *>(raw_alloc(BULK_HEADER + new_count * sizeof(vector3)));
  if (!base) return nullptr;
  *reinterpret_cast<size_t *>(base) = new_count;
  vector3 *p = reinterpret_cast<vector3 *>(base + BULK_HEADER);
  if (old) std::memcpy(p, old, old_count * sizeof(vector3));
  bulk_free(old);
  return p;
}

size_t capacity_for(size_t triangles) {
  size_t cap = 1024;
  while (cap < triangles) cap <<= 1;
  return cap;
}

// Re-anchor every object's vertex/normal pointers on the (possibly moved) global arrays.
void reanchor(struct scene *s) {
  size_t first = 0;
  for (size_t i = 0; i < s->object_count; ++i) {
    s->objects[i].triangles.vertex = s->objects_data.vertex + 3 * first;
    s->objects[i].triangles.normal = s->objects_data.normal + 3 * first;
    first += s->objects[i].triangle_count;
  }
}

}  // namespace

struct scene empty_scene() {
  struct scene s;
  std::memset(&s, 0, sizeof s);
  return s;
}

struct object *add_object_to_scene(struct scene *s, uint32_t nb_triangles) {
  const size_t old_total = s->triangle_count;
  const size_t new_total = old_total + nb_triangles;
  if (s->objects_data.vertex == nullptr || bulk_capacity(s->objects_data.vertex) < 3 * new_total) {
    const size_t count = capacity_for(new_total) * 3;
    s->objects_data.vertex = bulk_grow(s->objects_data.vertex, 3 * old_total, count);
    s->objects_data.normal = bulk_grow(s->objects_data.normal, 3 * old_total, count);
  }
  // objects[] grows by powers of two as well
  size_t ocap = 16;
  while (ocap < s->object_count + 1) ocap <<= 1;
  size_t ocap_old = 16;
  while (ocap_old < s->object_count) ocap_old <<= 1;
  if (s->objects == nullptr || ocap != ocap_old)
    s->objects = static_cast<struct object *>(std::realloc(s->objects, ocap * sizeof(struct object)));

  struct object *o = &s->objects[s->object_count];
  std::memset(o, 0, sizeof *o);
  o->triangle_count = nb_triangles;
  o->ni = 1.0f;  // defaults of gpu/scene.cu:199-209: ka=kd=ks=0, ns=0, ni=1, nr=0, d=1
  o->d = 1.0f;
  s->object_count += 1;
  s->triangle_count = new_total;
  reanchor(s);
  return o;
}

vector3 *get_vertex(const struct triangles_layout t, uint32_t i) { return t.vertex + 3 * (size_t)i; }
vector3 *get_normal(const struct triangles_layout t, uint32_t i) { return t.normal + 3 * (size_t)i; }

void free_scene(struct scene *s) {
  if (!s) return;
  bulk_free(s->objects_data.vertex);
  bulk_free(s->objects_data.normal);
  std::free(s->objects);
  std::free(s->lights);
  std::memset(s, 0, sizeof *s);
}

extern "C" void rtb_free_scene(struct scene *s) { free_scene(s); }

// Extension (loaders): 1 = the bulk arrays this thread allocates from now on are plain memory, 0 = page-locked when possible.
extern "C" void rtb_scene_mem_pageable(int on) { t_pageable = on != 0; }

// Extension: make room for `triangles` triangles in total now (exact size, one allocation),
// so that the add_object_to_scene calls that follow never move the arrays.
extern "C" int rtb_scene_reserve(struct scene *s, size_t triangles) {
  if (!s) return -1;
  if (bulk_capacity(s->objects_data.vertex) >= 3 * triangles && s->objects_data.vertex) return 0;
  const size_t have = 3 * s->triangle_count, count = 3 * (triangles ? triangles : 1);
  vector3 *v = bulk_grow(s->objects_data.vertex, have, count);
  vector3 *n = bulk_grow(s->objects_data.normal, have, count);
  if (!v || !n) return -1;
  s->objects_data.vertex = v;
  s->objects_data.normal = n;
  reanchor(s);
  return 0;
}
