// Test infrastructure: drives the REFERENCE's own scene loader (gpu/parser.cpp:64-112 parser(),
// gpu/parse_obj.cpp:15-62 parse_object(), gpu/scene.cu:146-218 scene container), compiled as host
// C++ from /root/reference by oracle/Makefile (_ref/libref_parser.so), so that the repo's loader
// (raytracing-gpu_b200/host/parser.cpp) and the golden fixtures can be checked against what the
// reference itself reads from a .svati file.  Arrays are copied out through plain pointers: the
// layout of the reference's `struct scene` (which depends on its build macros) never crosses.
#include <stdint.h>
#include <string.h>
#include "parser.h"
#include "scene.h"

// The reference's host container re-points its objects with cudaMemcpy(..., cudaMemcpyDefault)
// (gpu/scene.cu:57-113) even when both sides are host memory; that call needs a CUDA device, which
// the build container does not have.  Between two host pointers cudaMemcpyDefault IS memcpy, and this
// library only ever holds host scenes, so the two CUDA entry points scene.cu refers to are provided
// here (to_cuda(), the only caller of cudaMalloc, is never called) and the CUDA runtime is not linked.
extern "C" cudaError_t cudaMemcpy(void *dst, const void *src, size_t count, enum cudaMemcpyKind kind) {
  (void)kind;
  memcpy(dst, src, count);
  return cudaSuccess;
}
extern "C" cudaError_t cudaMalloc(void **p, size_t) { *p = nullptr; return cudaErrorMemoryAllocation; }

extern "C" {

void *refparser_parse(const char *path) {
  struct scene *s = new struct scene;
  *s = parser(path);              // errx(1, ...) on failure, like the reference
  return s;
}

void refparser_counts(void *h, uint64_t out[3]) {
  const struct scene *s = (const struct scene *)h;
  out[0] = s->object_count; out[1] = s->triangle_count; out[2] = s->light_count;
}

// vertex, normal: [9 * triangles] in the scene's memory order; tri_counts: [objects];
// materials: [13 * objects] = ka kd ks ns ni nr d; lights: [7 * lights] = type r g b v; camera: [12]
void refparser_copy(void *h, float *vertex, float *normal, uint32_t *tri_counts, float *materials, float *lights,
                    double *camera) {
  const struct scene *s = (const struct scene *)h;
  size_t first = 0;
  for (size_t o = 0; o < s->object_count; ++o) {
    const struct object &ob = s->objects[o];
    tri_counts[o] = ob.triangle_count;
    for (uint32_t t = 0; t < ob.triangle_count; ++t) {
      const vector3 *v = get_vertex(ob.triangles, t), *n = get_normal(ob.triangles, t);
      memcpy(vertex + 9 * (first + t), v, 36);
      memcpy(normal + 9 * (first + t), n, 36);
    }
    first += ob.triangle_count;
    float *m = materials + 13 * o;
    m[0] = ob.ka.x; m[1] = ob.ka.y; m[2] = ob.ka.z; m[3] = ob.kd.x; m[4] = ob.kd.y; m[5] = ob.kd.z;
    m[6] = ob.ks.x; m[7] = ob.ks.y; m[8] = ob.ks.z; m[9] = ob.ns; m[10] = ob.ni; m[11] = ob.nr; m[12] = ob.d;
  }
  for (size_t i = 0; i < s->light_count; ++i) {
    const struct light &l = s->lights[i];
    float *m = lights + 7 * i;
    m[0] = (float)l.type; m[1] = l.r; m[2] = l.g; m[3] = l.b; m[4] = l.v.x; m[5] = l.v.y; m[6] = l.v.z;
  }
  const struct camera &c = s->camera;
  const double cam[12] = {(double)c.width, (double)c.height, c.position.x, c.position.y, c.position.z, c.u.x, c.u.y, c.u.z,
                          c.v.x, c.v.y, c.v.z, c.fov};
  memcpy(camera, cam, sizeof cam);
}

}
