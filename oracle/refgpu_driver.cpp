/*
 * refgpu_driver.cpp — TEST INFRASTRUCTURE.  C entry point around the REFERENCE's own GPU
 * renderer: render() of gpu/raytracer.cu:177-253 with everything below it (to_cuda, the
 * object-octree build, raytrace<<<>>>, downscale<<<>>>), compiled UNMODIFIED for sm_100 from
 * /root/reference/gpu by oracle/Makefile into oracle/_ref/libref_gpu.so (default nvcc flags
 * like gpu/CMakeLists.txt:51-54: FMA contraction on, -rdc).  Used on the GPU box to time
 * "the existing kernel on the same box" and to count how far its frames are from the
 * un-fused semantic definition (tests/reports/ref_gpu_compare.py, tests/reports/microbench.py).  Never loaded by the product.
 */
#include <cstddef>
#include "scene.h"
#include "raytracer.h"

extern "C" int refgpu_render(const struct scene *s, char *buffer, int aliasing, long stride, int pre_h, int pre_w) {
  render(*s, buffer, aliasing, (std::ptrdiff_t)stride, pre_h, pre_w);
  return 0;
}

/*
 * Micro-benchmark entries: the reference's own build primitives, timed the way
 * gpu/bench.cu does (BM_parallel_radix_sort :224-257, BM_prefix_sum_parallel, to_cuda for the
 * per-scene FULL_BENCHMARK list :435-442), host wall clock around the call + device sync,
 * because the reference functions allocate and launch from the host.
 */
#include <chrono>
#include <cstdint>
#include <cuda_runtime.h>
#include "partitioning/sort.h"
#include "partitioning/prefix_sum.h"

static double now_ms() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// keys: host, n random u32.  best-of-iters ms of parallel_radix_sort alone (the H2D copy of
// bench.cu:244 is outside the timer).  sorted_out (host, n) receives the result.
extern "C" int refgpu_bench_sort_u32(const uint32_t *keys, size_t n, int iters, double *ms_out, uint32_t *sorted_out) {
  uint32_t *d_keys = nullptr;
  size_t *d_values = nullptr;
  uint16_t *d_second = nullptr;
  if (cudaMalloc(&d_keys, sizeof(uint32_t) * n) != cudaSuccess) return -1;
  if (cudaMalloc(&d_values, sizeof(size_t) * n) != cudaSuccess) return -1;
  // sort.tuh:153 sizes the second double buffer with sizeof(O); give the first the same room
  if (cudaMalloc(&d_second, sizeof(size_t) * n) != cudaSuccess) return -1;
  cudaMemset(d_values, 0, sizeof(size_t) * n);
  cudaMemset(d_second, 0, sizeof(size_t) * n);
  double best = 1e300;
  for (int it = 0; it < iters; ++it) {
    cudaMemcpy(d_keys, keys, sizeof(uint32_t) * n, cudaMemcpyDefault);
    cudaDeviceSynchronize();
    const double t0 = now_ms();
    parallel_radix_sort(d_keys, d_values, d_second, n);
    cudaDeviceSynchronize();
    const double t1 = now_ms();
    if (t1 - t0 < best) best = t1 - t0;
  }
  *ms_out = best;
  cudaMemcpy(sorted_out, d_keys, sizeof(uint32_t) * n, cudaMemcpyDefault);
  const cudaError_t e = cudaGetLastError();
  cudaFree(d_keys); cudaFree(d_values); cudaFree(d_second);
  return e == cudaSuccess ? 0 : -2;
}

// values: host, n u64.  best-of-iters ms of shared_prefix_sum alone; scanned_out receives the inclusive scan.
extern "C" int refgpu_bench_scan_u64(const uint64_t *values, size_t n, int iters, double *ms_out, uint64_t *scanned_out) {
  static_assert(sizeof(size_t) == sizeof(uint64_t), "size_t is 64-bit");
  size_t *d = nullptr;
  if (cudaMalloc(&d, sizeof(size_t) * n) != cudaSuccess) return -1;
  double best = 1e300;
  for (int it = 0; it < iters; ++it) {
    cudaMemcpy(d, values, sizeof(size_t) * n, cudaMemcpyDefault);
    cudaDeviceSynchronize();
    const double t0 = now_ms();
    shared_prefix_sum(d, n);
    cudaDeviceSynchronize();
    const double t1 = now_ms();
    if (t1 - t0 < best) best = t1 - t0;
  }
  *ms_out = best;
  cudaMemcpy(scanned_out, d, sizeof(size_t) * n, cudaMemcpyDefault);
  const cudaError_t e = cudaGetLastError();
  cudaFree(d);
  return e == cudaSuccess ? 0 : -2;
}

// Upload + object-octree build of the reference (to_cuda, gpu/scene.cu:224-352, which calls
// create_octree): best-of-iters ms.  The device scene it returns is leaked, as in the reference.
extern "C" int refgpu_bench_to_cuda(const struct scene *s, int iters, double *ms_out) {
  double best = 1e300;
  for (int it = 0; it < iters; ++it) {
    cudaDeviceSynchronize();
    const double t0 = now_ms();
    struct scene *d = to_cuda(s);
    cudaDeviceSynchronize();
    const double t1 = now_ms();
    (void)d;
    if (t1 - t0 < best) best = t1 - t0;
  }
  *ms_out = best;
  return cudaGetLastError() == cudaSuccess ? 0 : -2;
}
