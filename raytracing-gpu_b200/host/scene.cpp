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

namespace {

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
  const bool grow = s->objects_data.vertex == nullptr ||
                    capacity_for(new_total) != capacity_for(old_total);
  if (grow) {
    const size_t bytes = capacity_for(new_total) * 3 * sizeof(vector3);
    s->objects_data.vertex = static_cast<vector3 *>(std::realloc(s->objects_data.vertex, bytes));
    s->objects_data.normal = static_cast<vector3 *>(std::realloc(s->objects_data.normal, bytes));
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
  std::free(s->objects_data.vertex);
  std::free(s->objects_data.normal);
  std::free(s->objects);
  std::free(s->lights);
  std::memset(s, 0, sizeof *s);
}

extern "C" void rtb_free_scene(struct scene *s) { free_scene(s); }
