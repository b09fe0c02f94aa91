// Device arithmetic with the reference's exact operation order.
//
// This file (and every .cu that includes it) is compiled with -fmad=false: no
// multiply-add contraction anywhere, IEEE division and square root (nvcc defaults
// -prec-div=true -prec-sqrt=true; never -use_fast_math), so geometry and shading
// produce the same binary32 results as the reference's sources built without FMA
// (SURVEY Appendix A).  Fused multiply-adds are used only where written explicitly
// (__fmaf_rn in the conservative box test, which does not influence results).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

namespace rtb {

struct V3 { float x, y, z; };

__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 mk(float4 v) { return mk(v.x, v.y, v.z); }
__device__ __forceinline__ V3 mk(float3 v) { return mk(v.x, v.y, v.z); }
// gpu/vector3.cu:6-48
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 vadd(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 vcross(V3 a, V3 b) {
  return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ V3 vscale(V3 a, float r) { return mk(r * a.x, r * a.y, r * a.z); }
// gpu/vector3-extern.cu:5-23
__device__ __forceinline__ float vdot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float vlen(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }
__device__ __forceinline__ V3 vnorm(V3 a) { float r = vlen(a); return mk(a.x / r, a.y / r, a.z / r); }

// ---- colours: every value is an 8-bit channel (gpu/colors.cu:3-47).  Packed r | g<<8 | b<<16 | a<<24.
__device__ __forceinline__ uint32_t quant(float x) {
  x = x * 255;
  if (x > 255) x = 255;
  if (x < 0) x = 0;
  return (uint32_t)__float2int_rz(x);   // truncation; NaN -> 0 like the float->u8 conversion
}
__device__ __forceinline__ uint32_t pack_col(uint32_t r, uint32_t g, uint32_t b) {
  return r | (g << 8) | (b << 16) | 0xff000000u;
}
__device__ __forceinline__ uint32_t init_color(float r, float g, float b) { return pack_col(quant(r), quant(g), quant(b)); }
__device__ __forceinline__ uint32_t ch_r(uint32_t c) { return c & 0xffu; }
__device__ __forceinline__ uint32_t ch_g(uint32_t c) { return (c >> 8) & 0xffu; }
__device__ __forceinline__ uint32_t ch_b(uint32_t c) { return (c >> 16) & 0xffu; }
__device__ __forceinline__ uint32_t color_add(uint32_t a, uint32_t b) {
  uint32_t r = ch_r(a) + ch_r(b), g = ch_g(a) + ch_g(b), bl = ch_b(a) + ch_b(b);
  return pack_col(r > 255u ? 255u : r, g > 255u ? 255u : g, bl > 255u ? 255u : bl);
}
__device__ __forceinline__ uint32_t color_mul(uint32_t a, float k) {
  return init_color((float)ch_r(a) / 255 * k, (float)ch_g(a) / 255 * k, (float)ch_b(a) / 255 * k);
}
__device__ __forceinline__ uint32_t color_mults(uint32_t a, uint32_t b) {
  return init_color(((float)ch_r(a) / 255) * ((float)ch_r(b) / 255), ((float)ch_g(a) / 255) * ((float)ch_g(b) / 255),
                    ((float)ch_b(a) / 255) * ((float)ch_b(b) / 255));
}

// monotone float <-> uint mapping for atomicMin / atomicMax on floats
__device__ __forceinline__ uint32_t f2ord(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

}  // namespace rtb
