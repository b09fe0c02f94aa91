/*
 * refgpu_driver.cpp — TEST INFRASTRUCTURE.  C entry point around the REFERENCE's own GPU
 * renderer: render() of gpu/raytracer.cu:177-253 with everything below it (to_cuda, the
 * object-octree build, raytrace<<<>>>, downscale<<<>>>), compiled UNMODIFIED for sm_100 from
 * /root/reference/gpu by oracle/Makefile into oracle/_ref/libref_gpu.so (default nvcc flags
 * like gpu/CMakeLists.txt:51-54: FMA contraction on, -rdc).  Used on the GPU box to time
 * "the existing kernel on the same box" and to count how far its frames are from the
 * un-fused semantic definition (tools/ref_gpu_compare.py).  Never loaded by the product.
 */
#include <cstddef>
#include "scene.h"
#include "raytracer.h"

extern "C" int refgpu_render(const struct scene *s, char *buffer, int aliasing, long stride, int pre_h, int pre_w) {
  render(*s, buffer, aliasing, (std::ptrdiff_t)stride, pre_h, pre_w);
  return 0;
}
