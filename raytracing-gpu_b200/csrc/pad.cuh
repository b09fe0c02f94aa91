// Needle padding of triangle boxes (see the derivation above k_emit in build.cu and DESIGN.md
// section 4).  Shared by the hierarchy build (build.cu) and the primary-visibility rasteriser
// (raster.cuh), which must cull with the same boxes.
#pragma once
#include "rtb_math.cuh"

namespace rtb {
namespace {

constexpr float SLIVER_KAPPA0 = 4.0f;
#ifndef RTB_SLIVER_K
#define RTB_SLIVER_K 3.0f
#endif
constexpr float SLIVER_K = RTB_SLIVER_K;
#ifndef RTB_SLIVER_MODE
#define RTB_SLIVER_MODE 5   /* 0: none, 1: isotropic by the sharpest corner, 2: corner spikes, 3: spikes of the
                               sharpest corner at every corner, 5 (product): third-edge slop + corner spikes */
#endif

// 1 / sin(theta / 2) for the corner between unit-length-normalised edges (cosine given)
__device__ __forceinline__ float corner_kappa(float cosine) {
  const float c = fminf(cosine, 1.0f);
  return fminf(1.0f / sqrtf(fmaxf(0.5f * (1.0f - c), 1e-12f)), 1.0e4f);
}

__device__ __forceinline__ void grow(V3 &lo, V3 &hi, const V3 p) {
  lo = mk(fminf(lo.x, p.x), fminf(lo.y, p.y), fminf(lo.z, p.z));
  hi = mk(fmaxf(hi.x, p.x), fmaxf(hi.y, p.y), fmaxf(hi.z, p.z));
}

// Grows [lo, hi] (the triangle's box).  pad_unit = SLIVER_K * 2^-23 * scale.
// Corner spikes: at a corner v with unit edge directions a, b and interior angle theta, edge
// lines pushed out by delta_a, delta_b <= delta meet at v - (delta_b a + delta_a b) / sin(theta),
// inside the parallelogram v - m (s a + t b), s, t in [0, 1], m = delta / sin(theta)
// <= pad_unit * kappa / 2.  That spike is added to the box with m = pad_unit * kappa.
__device__ __forceinline__ void sliver_pad(const V3 (&v)[3], float pad_unit, V3 &lo, V3 &hi) {
#if RTB_SLIVER_MODE == 0
  return;
#else
  const V3 ab = vsub(v[1], v[0]), bc = vsub(v[2], v[1]), ca = vsub(v[0], v[2]);
  const float lab = vlen(ab), lbc = vlen(bc), lca = vlen(ca);
  if (!(lab > 0) || !(lbc > 0) || !(lca > 0)) return;   // two equal vertices: a == 0 exactly, never accepted
  const float longest = fmaxf(fmaxf(lab, lbc), lca);
  // unit edge directions leaving each corner, and the corner cosines
  const V3 d_ab = vscale(ab, 1.0f / lab), d_bc = vscale(bc, 1.0f / lbc), d_ca = vscale(ca, 1.0f / lca);
  const V3 out0[2] = {d_ab, vscale(d_ca, -1.0f)};       // from a: towards b, towards c
  const V3 out1[2] = {d_bc, vscale(d_ab, -1.0f)};       // from b: towards c, towards a
  const V3 out2[2] = {d_ca, vscale(d_bc, -1.0f)};       // from c: towards a, towards b
  const float kap[3] = {corner_kappa(vdot(out0[0], out0[1])), corner_kappa(vdot(out1[0], out1[1])),
                        corner_kappa(vdot(out2[0], out2[1]))};
#if RTB_SLIVER_MODE == 1
  const float kappa = fmaxf(fmaxf(kap[0], kap[1]), kap[2]);
  if (kappa > SLIVER_KAPPA0) {
    // the u,v <= 1 tests keep accepted rays within about one edge length of the triangle
    const float pad = fminf(pad_unit * kappa, longest);
    lo = mk(lo.x - pad, lo.y - pad, lo.z - pad);
    hi = mk(hi.x + pad, hi.y + pad, hi.z + pad);
  }
#else
  const V3 *outs[3] = {out0, out1, out2};
#if RTB_SLIVER_MODE == 5
  // u + v <= 1 is the edge opposite vertex 0; u and v carry an absolute error of
  // delta / (height over their edge) each, so that edge line is pushed out by
  // delta * (|e1| + |e2|) / |e3| in ray space: large when the SHORT edge of a needle is the
  // one opposite vertex 0.  A distance in ray space is a distance in 3D between ray and
  // triangle, whatever the incidence: isotropic pad.
  {
    const float ratio = (lab + lca) / lbc;
    if (ratio > SLIVER_KAPPA0) {
      const float pad = fminf(pad_unit * ratio, longest);
      lo = mk(lo.x - pad, lo.y - pad, lo.z - pad);
      hi = mk(hi.x + pad, hi.y + pad, hi.z + pad);
    }
  }
#endif
#if RTB_SLIVER_MODE == 3   /* every corner gets the spike of the sharpest one */
  const float kmax = fmaxf(fmaxf(kap[0], kap[1]), kap[2]);
#endif
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#if RTB_SLIVER_MODE == 3
    if (!(kmax > SLIVER_KAPPA0)) continue;
    const float m = fminf(pad_unit * kmax, longest);
#else
    if (!(kap[c] > SLIVER_KAPPA0)) continue;
    const float m = fminf(pad_unit * kap[c], longest);
#endif
    const V3 sa = vscale(outs[c][0], m), sb = vscale(outs[c][1], m);
    grow(lo, hi, vsub(v[c], sa));
    grow(lo, hi, vsub(v[c], sb));
    grow(lo, hi, vsub(vsub(v[c], sa), sb));
  }
#endif
#endif
}


__device__ __forceinline__ void sliver_pad(const float *raw_v, uint32_t i, float pad_unit, V3 &lo, V3 &hi) {
  const float *p = raw_v + 9ull * i;
  const V3 v[3] = {mk(p[0], p[1], p[2]), mk(p[3], p[4], p[5]), mk(p[6], p[7], p[8])};
  sliver_pad(v, pad_unit, lo, hi);
}

}  // namespace
}  // namespace rtb
