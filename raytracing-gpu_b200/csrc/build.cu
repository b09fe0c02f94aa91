// Device-side hierarchy build over TRIANGLES.
//
// Replaces the reference's object-level pipeline (gpu/partitioning/octree.cu:362-411:
// compute_bounding_box -> find_scene_scale -> position_object -> parallel_radix_sort ->
// nodes_difference_array -> shared_prefix_sum -> create_octree, with blocking copies,
// per-step cudaMalloc and three cross-block races) by:
//   k_bounds      triangle boxes -> scene box (warp shuffles + ordered-int atomics, race-free)
//   k_morton      63-bit Morton code of the box centre (21 bits per axis: the octree cell
//                 path of the triangle, 21 levels x 3 bits, top level most significant)
//   radix sort    sort.cu (key + index payload)
//   k_permute     triangles into leaf order as float4 streams (v0|id, e1|obj, e2; n0,n1,n2)
//   k_emit        ONE bottom-up pass over the sorted codes: every leaf climbs towards the
//                 root, the second thread to reach a node merges the two child boxes and
//                 goes on (Apetrei-style agglomeration; the tree is the Morton radix tree,
//                 i.e. the octree with each level split into its three binary cuts)
//   k_compact     subtrees of <= LEAF_MAX triangles become leaves; surviving nodes are
//                 written to the flat 64-byte node array with remapped child references
// The result is deterministic: topology depends only on the sorted keys, boxes are exact
// min/max of vertex coordinates.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "rtb_internal.h"
#include "rtb_math.cuh"
#include "pad.cuh"
#include "scene_store.h"

namespace rtb {

namespace {

constexpr int BT = 256;
constexpr uint32_t MODE_CHOICE_MIN = 8192;      // scenes below this keep the cubic key layout
constexpr uint32_t MODE_SAMPLE_MAX = 1u << 18;   // triangles in a sample tree
constexpr uint32_t ONE_FRAME_SAH_MAX = 1u << 18; // RTB_BUILD_ONE_FRAME: larger scenes keep the radix tree

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int d = 16; d; d >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, d));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int d = 16; d; d >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, d));
  return v;
}

__device__ __forceinline__ void tri_box(const float *raw_v, uint32_t i, V3 &lo, V3 &hi) {
  const float *p = raw_v + 9ull * i;
  float x0 = p[0], y0 = p[1], z0 = p[2], x1 = p[3], y1 = p[4], z1 = p[5], x2 = p[6], y2 = p[7], z2 = p[8];
  lo = mk(fminf(x0, fminf(x1, x2)), fminf(y0, fminf(y1, y2)), fminf(z0, fminf(z1, z2)));
  hi = mk(fmaxf(x0, fmaxf(x1, x2)), fmaxf(y0, fmaxf(y1, y2)), fmaxf(z0, fmaxf(z1, z2)));
}

// bounds[0..2] = min (ordered-uint), bounds[3..5] = max.  Initialised by the host to
// 0xffffffff / 0 before the launch, so there is no in-kernel init to race with.
__global__ void __launch_bounds__(BT) k_bounds(const float *raw_v, uint32_t n, uint32_t *bounds) {
  __shared__ float red[6][BT / 32];
  const float INF = __int_as_float(0x7f800000);
  V3 lo = mk(INF, INF, INF), hi = mk(-INF, -INF, -INF);
  for (uint32_t i = blockIdx.x * BT + threadIdx.x; i < n; i += gridDim.x * BT) {
    V3 a, b;
    tri_box(raw_v, i, a, b);
    lo = mk(fminf(lo.x, a.x), fminf(lo.y, a.y), fminf(lo.z, a.z));
    hi = mk(fmaxf(hi.x, b.x), fmaxf(hi.y, b.y), fmaxf(hi.z, b.z));
  }
  lo = mk(warp_min(lo.x), warp_min(lo.y), warp_min(lo.z));
  hi = mk(warp_max(hi.x), warp_max(hi.y), warp_max(hi.z));
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    red[0][warp] = lo.x; red[1][warp] = lo.y; red[2][warp] = lo.z;
    red[3][warp] = hi.x; red[4][warp] = hi.y; red[5][warp] = hi.z;
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    float v = red[threadIdx.x][0];
    for (int w = 1; w < BT / 32; ++w) v = threadIdx.x < 3 ? fminf(v, red[threadIdx.x][w]) : fmaxf(v, red[threadIdx.x][w]);
    if (threadIdx.x < 3) atomicMin(&bounds[threadIdx.x], f2ord(v));
    else atomicMax(&bounds[threadIdx.x], f2ord(v));
  }
}

__device__ __forceinline__ uint64_t spread21(uint32_t v) {
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

// 21 bits -> every second bit (two-way interleave)
__device__ __forceinline__ uint64_t spread21_2(uint32_t v) {
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 16) & 0x0000ffff0000ffffull;
  x = (x | x << 8) & 0x00ff00ff00ff00ffull;
  x = (x | x << 4) & 0x0f0f0f0f0f0f0f0full;
  x = (x | x << 2) & 0x3333333333333333ull;
  x = (x | x << 1) & 0x5555555555555555ull;
  return x;
}

// Key layouts (all 63 bits, cubic cells).  Mode 0: x, y, z bits interleaved, the octree cell
// path.  Mode 1 / 2 / 3: axis x / y / z comes LAST: the other two axes interleaved (42 bits,
// a quadtree path), then the 21 bits of that axis.  A surface that is mostly perpendicular to
// one axis (a height field) gains nothing from cuts along it: such a cut separates the
// triangles of a cell by height into two sets that both span the cell, a wasted level for
// every ray (C5: 42.8 -> 37.1 node visits and 5.2 -> 4.6 triangle tests per ray,
// tools/tree_quality.cpp).  build_scene picks the mode by the surface-area cost of sample
// trees (k_emit sah_acc, k_choose_mode).
__device__ __forceinline__ uint64_t morton_key(int mode, uint32_t qx, uint32_t qy, uint32_t qz) {
  if (mode == 1) return ((spread21_2(qy) | (spread21_2(qz) << 1)) << 21) | qx;
  if (mode == 2) return ((spread21_2(qx) | (spread21_2(qz) << 1)) << 21) | qy;
  if (mode == 3) return ((spread21_2(qx) | (spread21_2(qy) << 1)) << 21) | qz;
  return spread21(qx) | (spread21(qy) << 1) | (spread21(qz) << 2);
}

// Entry i is triangle i * stride (stride > 1: the sample trees of the mode choice).
__global__ void __launch_bounds__(BT) k_morton(const float *raw_v, uint32_t n, uint32_t stride, const uint32_t *bounds,
                                               int mode, const int *d_mode, uint64_t *keys, uint32_t *vals) {
  const uint32_t i = blockIdx.x * BT + threadIdx.x;
  if (i >= n) return;
  if (d_mode) mode = *d_mode;
  const uint32_t tri = i * stride;
  const V3 smin = mk(ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]));
  const V3 smax = mk(ord2f(bounds[3]), ord2f(bounds[4]), ord2f(bounds[5]));
  V3 lo, hi;
  tri_box(raw_v, tri, lo, hi);
  const float S = 2097152.0f;   // 2^21 cells per axis
  // cubic cells: one scale for the three axes (the scene box's longest side), so a code
  // prefix is an octree cell and flat scenes do not waste their top splits on a thin axis
  const float ext = fmaxf(fmaxf(smax.x - smin.x, smax.y - smin.y), smax.z - smin.z);
  const float cx = 0.5f * lo.x + 0.5f * hi.x, cy = 0.5f * lo.y + 0.5f * hi.y, cz = 0.5f * lo.z + 0.5f * hi.z;
  const float fx = ext > 0 ? (cx - smin.x) / ext * S : 0.0f;
  const float fy = ext > 0 ? (cy - smin.y) / ext * S : 0.0f;
  const float fz = ext > 0 ? (cz - smin.z) / ext * S : 0.0f;
  uint32_t qx = (uint32_t)fminf(fmaxf(fx, 0.0f), S - 1.0f);
  uint32_t qy = (uint32_t)fminf(fmaxf(fy, 0.0f), S - 1.0f);
  uint32_t qz = (uint32_t)fminf(fmaxf(fz, 0.0f), S - 1.0f);
  keys[i] = morton_key(mode, qx, qy, qz);
  vals[i] = tri;
}

// inv[order[s]] = s
__global__ void __launch_bounds__(BT) k_invert(const uint32_t *order, uint32_t n, uint32_t *inv) {
  const uint32_t s = blockIdx.x * BT + threadIdx.x;
  if (s < n) inv[order[s]] = s;
}

// Identity order (RTB_BUILD_NO_TREE): slot == global id.
__global__ void __launch_bounds__(BT) k_iota(uint32_t *vals, uint32_t n) {
  const uint32_t i = blockIdx.x * BT + threadIdx.x;
  if (i < n) vals[i] = i;
}

__device__ __forceinline__ uint32_t object_of(const uint32_t *obj_first, uint32_t n_obj, uint32_t tri) {
  // largest o with obj_first[o] <= tri  (obj_first has n_obj + 1 entries, ascending)
  uint32_t lo = 0, hi = n_obj;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (obj_first[mid] <= tri) lo = mid; else hi = mid;
  }
  return lo;
}

// POS: positions (v0|id, e1|obj, e2), NRM: unit normals.  Two launches for big scenes, where
// the normals are still on their way over PCIe while the hierarchy is built (build_scene).
template <bool POS, bool NRM>
__global__ void __launch_bounds__(BT) k_permute(const float *raw_v, const float *raw_n, const uint32_t *order,
                                                uint32_t n, const uint32_t *obj_first, uint32_t n_obj,
                                                float4 *v0, float4 *e1, float4 *e2, float4 *n0, float4 *n1,
                                                float4 *n2) {
  const uint32_t s = blockIdx.x * BT + threadIdx.x;
  if (s >= n) return;
  const uint32_t i = order[s];
  if (POS) {
    const float *p = raw_v + 9ull * i;
    const V3 a = mk(p[0], p[1], p[2]), b = mk(p[3], p[4], p[5]), c = mk(p[6], p[7], p[8]);
    const V3 ab = vsub(b, a), ac = vsub(c, a);   // edge1/edge2 of gpu/hit.cu:20-21, same subtraction
    v0[s] = make_float4(a.x, a.y, a.z, __uint_as_float(i));
    e1[s] = make_float4(ab.x, ab.y, ab.z, __uint_as_float(object_of(obj_first, n_obj, i)));
    e2[s] = make_float4(ac.x, ac.y, ac.z, 0.0f);
  }
  if (NRM) {
    const float *q = raw_n + 9ull * i;
    const V3 na = vnorm(mk(q[0], q[1], q[2])), nb = vnorm(mk(q[3], q[4], q[5])), nc = vnorm(mk(q[6], q[7], q[8]));
    n0[s] = make_float4(na.x, na.y, na.z, 0.0f);
    n1[s] = make_float4(nb.x, nb.y, nb.z, 0.0f);
    n2[s] = make_float4(nc.x, nc.y, nc.z, 0.0f);
  }
}

// "How different are sorted neighbours j and j+1": position of the highest differing key
// bit (+64), or of the highest differing index bit when the keys are equal.
__device__ __forceinline__ int delta(const uint64_t *keys, uint32_t j) {
  const uint64_t x = keys[j] ^ keys[j + 1];
  if (x) return 128 - __clzll((long long)x);
  return 32 - __clz((int)(j ^ (j + 1)));
}

__device__ __forceinline__ int leaf_ref(uint32_t first, uint32_t count) { return ~(int)((first << 3) | count); }

struct RawNode {           // same 64-byte layout as the final node
  float4 q0, q1, q2;
  int ref0, ref1, lo, cnt;
};

__device__ __forceinline__ void store_child(RawNode *nd, int side, V3 lo, V3 hi, int ref) {
  float *f = reinterpret_cast<float *>(nd);
  // q0/q1 hold x and y of child 0/1; q2 holds z of both
  float *qxy = f + 4 * side;
  qxy[0] = lo.x; qxy[1] = hi.x; qxy[2] = lo.y; qxy[3] = hi.y;
  f[8 + 2 * side] = lo.z;
  f[9 + 2 * side] = hi.z;
  (&nd->ref0)[side] = ref;
}
__device__ __forceinline__ void load_child(const volatile RawNode *nd, int side, V3 &lo, V3 &hi) {
  const volatile float *f = reinterpret_cast<const volatile float *>(nd);
  const volatile float *qxy = f + 4 * side;
  lo.x = qxy[0]; hi.x = qxy[1]; lo.y = qxy[2]; hi.y = qxy[3];
  lo.z = f[8 + 2 * side];
  hi.z = f[9 + 2 * side];
}

// One thread per sorted triangle.  visit[j] (initialised to -1) receives the range bound
// of the first child to arrive at inner node j; the second child continues upwards.
// Needle padding.  In ray space (the plane perpendicular to the ray) the rounded
// Moeller-Trumbore test is "inside three half-planes", each edge line displaced by the rounding
// noise: u = s.(d x e2) / a is |d| |e2'| dist(ray, line of e2) / a, and the numerator is
// computed with an absolute error of c 2^-24 |s| |d x e2| (c about 3..10), so the lines of e1
// and e2 move by delta = c 2^-24 |s|, |s| = |origin - vertex0|, whatever the triangle.  The
// third edge is tested as u + v <= 1: u and v carry delta / (their height), which moves that
// line by delta (|e1| + |e2|) / |e3|: 47 delta for the 2.4-degree needle of
// tests/golden island_smooth, whose short edge is the one opposite vertex 0 (the sample that
// found this: accepted 0.033 from the triangle, 1 wrong id in 4.67 M before the padding).
// Distances in ray space are 3D distances between ray and triangle, so the cure is an isotropic
// pad of the leaf box by pad_unit (|e1| + |e2|) / |e3| when that ratio exceeds SLIVER_KAPPA0
// (the ray-time margin of trace.cu, 48 2^-24 (|o| + |scene|), covers the rest).  Displaced
// lines meet delta / sin(theta) beyond a corner of interior angle theta: corners sharper than
// 1 / sin(theta/2) > SLIVER_KAPPA0 get that spike added to the box (direction-aware, so a
// needle's box stays thin).  pad_unit = SLIVER_K 2^-23 (|scene| + origin_radius), i.e.
// c = 2 SLIVER_K = 6 at the largest |s| the scene handle accepts; pads are capped at the
// longest edge and inherited by inner boxes through the union.  What no finite pad covers:
// triangles seen edge-on (a -> 0), accepted by the reference arithmetic anywhere in their
// plane.
#ifndef RTB_EMIT_WARP_PHASE
#define RTB_EMIT_WARP_PHASE 1
#endif
#ifndef RTB_EMIT_RELEASE_ATOM
#define RTB_EMIT_RELEASE_ATOM 0
#endif
#ifndef RTB_EMIT_ACQUIRE_FENCE
#define RTB_EMIT_ACQUIRE_FENCE 1
#endif
// A subtree of the radix tree that k_sah_clusters rebuilds: slots [start, start + count) and
// the child reference of the radix node above it.
struct Cluster { uint32_t start, count; int parent, side; };

__device__ __forceinline__ void push_cluster(Cluster *clusters, uint32_t *n_clusters, uint32_t cl_max, uint32_t start,
                                             uint32_t count, int parent, int side) {
  if (count <= (uint32_t)LEAF_MAX || count > cl_max) return;
  const uint32_t k = atomicAdd(n_clusters, 1u);
  clusters[k].start = start; clusters[k].count = count; clusters[k].parent = parent; clusters[k].side = side;
}

// Surface-area cost of one merge (sample trees of the mode choice): the merged node is visited
// in proportion to its box area; a child that stays a leaf costs 1.5 visits per triangle (a
// triangle test against a node visit in k_frame).  Fixed point, so the sum does not depend on
// the order of the atomics and every build of a scene takes the same decision.
constexpr int SAH_SLOTS = 64;
__device__ __forceinline__ float box_area(V3 lo, V3 hi) {
  const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
  return dx * dy + dy * dz + dz * dx;
}
__device__ __forceinline__ void sah_add(unsigned long long *acc, int node, float scale, V3 alo, V3 ahi, uint32_t acnt,
                                        V3 blo, V3 bhi, uint32_t bcnt) {
  if (acnt + bcnt <= (uint32_t)LEAF_MAX) return;      // stays inside a leaf
  const V3 mlo = mk(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z));
  const V3 mhi = mk(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z));
  float c = box_area(mlo, mhi);
  if (acnt <= (uint32_t)LEAF_MAX) c += 1.5f * (float)acnt * box_area(alo, ahi);
  if (bcnt <= (uint32_t)LEAF_MAX) c += 1.5f * (float)bcnt * box_area(blo, bhi);
  atomicAdd(acc + (node & (SAH_SLOTS - 1)), (unsigned long long)(c * scale));
}

// Which axis could come last: the one the surface is most perpendicular to.  acc[a] = sum over
// the sample of |(e1 x e2)_a| = 2 x area x |normal_a|, in fixed point (order-independent).
__global__ void __launch_bounds__(BT) k_normal_stat(const float *raw_v, uint32_t n, uint32_t stride,
                                                    const uint32_t *bounds, unsigned long long *acc) {
  const uint32_t i = blockIdx.x * BT + threadIdx.x;
  unsigned long long v[3] = {0, 0, 0};
  if (i < n) {
    const float *p = raw_v + 9ull * (i * stride);
    const V3 a = mk(p[0], p[1], p[2]), e1 = vsub(mk(p[3], p[4], p[5]), a), e2 = vsub(mk(p[6], p[7], p[8]), a);
    const V3 c = vcross(e1, e2);
    const float ext = fmaxf(fmaxf(ord2f(bounds[3]) - ord2f(bounds[0]), ord2f(bounds[4]) - ord2f(bounds[1])),
                            ord2f(bounds[5]) - ord2f(bounds[2]));
    const float sc = ext > 0 ? 1099511627776.0f / ext / ext : 0.0f;     // 2^40 / extent^2
    const float q[3] = {fabsf(c.x) * sc, fabsf(c.y) * sc, fabsf(c.z) * sc};
#pragma unroll
    for (int k = 0; k < 3; ++k) v[k] = q[k] < 1.0e13f ? (unsigned long long)q[k] : 0ull;    // NaN / overflow: ignored
  }
#pragma unroll
  for (int k = 0; k < 3; ++k) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], d);
    if ((threadIdx.x & 31) == 0 && v[k]) atomicAdd(acc + k, v[k]);
  }
}

// cand[0] = key mode of the second candidate (1 + dominant normal axis; ties: lowest axis)
__global__ void k_pick_axis(const unsigned long long *acc, int *cand) {
  if (threadIdx.x || blockIdx.x) return;
  int a = 0;
  if (acc[1] > acc[a]) a = 1;
  if (acc[2] > acc[a]) a = 2;
  cand[0] = 1 + a;
}

// cost[c * SAH_SLOTS + k]: slot sums of candidate c.  Mode 0 (cubic) stays unless another
// candidate's sample tree (one axis last, cand[0]) is at least 5 % cheaper.
// out[0] = mode, out[1 + mode] = cost >> 8 of the trees that were built.
__global__ void k_choose_mode(const unsigned long long *cost, const int *cand, int *out) {
  if (threadIdx.x || blockIdx.x) return;
  unsigned long long s0 = 0, s1 = 0;
  for (int k = 0; k < SAH_SLOTS; ++k) { s0 += cost[k]; s1 += cost[SAH_SLOTS + k]; }
  const int m = cand[0];
  for (int c = 1; c < 5; ++c) out[c] = 0;
  out[1] = (int)min(s0 >> 8, 0x7fffffffull);
  out[1 + m] = (int)min(s1 >> 8, 0x7fffffffull);
  out[0] = s1 < s0 - s0 / 20 ? m : 0;
}

__global__ void __launch_bounds__(BT) k_emit(const float *raw_v, const uint32_t *order, const uint64_t *keys,
                                             uint32_t n, RawNode *raw, int *visit, int *child_height,
                                             int *root_info /* [0]=raw root, [1]=height */,
                                             const uint32_t *bounds, float origin_radius, int *raw_parent,
                                             unsigned long long *sah_acc /* nullptr or [SAH_SLOTS] */,
                                             Cluster *clusters, uint32_t *n_clusters, uint32_t cl_max) {
  // cl_max > 0: subtrees of at most cl_max triangles are rebuilt by k_sah_clusters; they are
  // listed here (at the merge that first exceeds cl_max) and count as height 0
  const uint32_t flat_max = cl_max > (uint32_t)LEAF_MAX ? cl_max : (uint32_t)LEAF_MAX;
  const uint32_t i = blockIdx.x * BT + threadIdx.x;
  [[maybe_unused]] const unsigned lane = threadIdx.x & 31u;
  const unsigned wmask = __ballot_sync(0xffffffffu, i < n);     // lanes of this warp that hold a triangle
  if (i >= n) return;
  V3 lo, hi;
  tri_box(raw_v, order[i], lo, hi);
  {
    float amax = 0;
    for (int k = 0; k < 6; ++k) amax = fmaxf(amax, fabsf(ord2f(bounds[k])));
    sliver_pad(raw_v, order[i], SLIVER_K * 1.1920929e-7f * (amax + origin_radius), lo, hi);
  }
  // surface-area cost of the tree in units of 2^-24 of the scene box's area (sah_acc only)
  float sah_scale = 0.0f;
  if (sah_acc) {
    const float ex = ord2f(bounds[3]) - ord2f(bounds[0]), ey = ord2f(bounds[4]) - ord2f(bounds[1]),
                ez = ord2f(bounds[5]) - ord2f(bounds[2]);
    const float ra = ex * ey + ey * ez + ez * ex;
    sah_scale = ra > 0 ? 16777216.0f / ra : 0.0f;
  }
  uint32_t l = i, r = i;
  int cur = -1;       // raw inner index of the node we stand on (-1: still the single triangle)
  int height = 0;     // inner nodes on the longest path below (collapsed subtrees count 0)
#if RTB_EMIT_WARP_PHASE
  // ---- warp phase: subtrees whose two children are both held by lanes of this warp are
  // merged through shuffles, no flag, no fence (31 of 32 merges of the whole build; the
  // fences of the global phase below were 60 % of this kernel's stall cycles).  Two lanes
  // are siblings exactly when they computed the same parent.
  {
    bool holding = true;
    for (;;) {
      int parent = -1, side = 0;
      if (holding) {
        if (r - l + 1 == n) {            // the whole scene fits in this warp
          root_info[0] = cur; root_info[1] = height;
          holding = false;
        } else if (l == 0 || (r != n - 1 && delta(keys, r) < delta(keys, l - 1))) { parent = (int)r; side = 0; }
        else { parent = (int)l - 1; side = 1; }
      }
      const unsigned hm = __ballot_sync(wmask, holding);
      if (!hm) return;
      unsigned grp = 0;
      if (holding) grp = __match_any_sync(hm, parent);
      const bool pair = holding && __popc(grp) == 2;
      const unsigned pm = __ballot_sync(wmask, pair);
      if (!pm) break;                    // every sibling still missing lives in another warp
      if (pair) {
        const int mate = __ffs((int)(grp & ~(1u << lane))) - 1;
        const uint32_t cnt = r - l + 1;
        const int my_ref = cnt <= (uint32_t)LEAF_MAX ? leaf_ref(l, cnt) : cur;
        const int my_h = cnt <= flat_max ? 0 : height;
        const float mlx = __shfl_sync(pm, lo.x, mate), mly = __shfl_sync(pm, lo.y, mate), mlz = __shfl_sync(pm, lo.z, mate);
        const float mhx = __shfl_sync(pm, hi.x, mate), mhy = __shfl_sync(pm, hi.y, mate), mhz = __shfl_sync(pm, hi.z, mate);
        const int m_ref = __shfl_sync(pm, my_ref, mate), m_h = __shfl_sync(pm, my_h, mate);
        const uint32_t m_r = __shfl_sync(pm, r, mate);
        if (side == 0) {                 // the left child writes the node and carries on
          r = m_r;
          float4 *dst = reinterpret_cast<float4 *>(&raw[parent]);
          dst[0] = make_float4(lo.x, hi.x, lo.y, hi.y);
          dst[1] = make_float4(mlx, mhx, mly, mhy);
          dst[2] = make_float4(lo.z, hi.z, mlz, mhz);
          dst[3] = make_float4(__int_as_float(my_ref), __int_as_float(m_ref), __int_as_float((int)l),
                               __int_as_float((int)(r - l + 1)));
          if (sah_acc)
            sah_add(sah_acc, parent, sah_scale, lo, hi, cnt, mk(mlx, mly, mlz), mk(mhx, mhy, mhz), r - l + 1 - cnt);
          if (cl_max && r - l + 1 > cl_max) {
            push_cluster(clusters, n_clusters, cl_max, l, cnt, parent, 0);
            push_cluster(clusters, n_clusters, cl_max, l + cnt, r - l + 1 - cnt, parent, 1);
          }
          lo = mk(fminf(lo.x, mlx), fminf(lo.y, mly), fminf(lo.z, mlz));
          hi = mk(fmaxf(hi.x, mhx), fmaxf(hi.y, mhy), fmaxf(hi.z, mhz));
          height = (my_h > m_h ? my_h : m_h) + 1;
          cur = parent;
        } else {
          holding = false;
        }
      }
    }
    if (!holding) return;
  }
#endif
  for (;;) {
    const uint32_t cnt = r - l + 1;
    if (cnt == n) {
      root_info[0] = cur; root_info[1] = height;
      return;
    }
    const int my_ref = cnt <= (uint32_t)LEAF_MAX ? leaf_ref(l, cnt) : cur;
    const int my_h = cnt <= flat_max ? 0 : height;
    int parent, side;
    if (l == 0 || (r != n - 1 && delta(keys, r) < delta(keys, l - 1))) { parent = (int)r; side = 0; }
    else { parent = (int)l - 1; side = 1; }
    store_child(&raw[parent], side, lo, hi, my_ref);
    child_height[2 * parent + side] = my_h;
#if RTB_EMIT_RELEASE_ATOM
    int other;
    asm volatile("atom.release.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(other) : "l"(visit + parent), "r"(side == 0 ? (int)l : (int)r) : "memory");
    if (other < 0) return;
#else
    __threadfence();                             // our child record is visible before the flag is
    const int other = atomicExch(&visit[parent], side == 0 ? (int)l : (int)r);
    if (other < 0) return;                       // first to arrive: the sibling will carry on
#if RTB_EMIT_ACQUIRE_FENCE
    __threadfence();
#endif
#endif
    // The sibling's record is read with volatile loads (L2, never a stale L1 line).  Tried and
    // not faster on 10 M triangles: one acq_rel exchange (libcu++), a release exchange, no
    // acquire fence; what pays is needing the flag protocol for 1 merge in 32 only (warp phase).
    V3 slo, shi;
    load_child(&raw[parent], side ^ 1, slo, shi);
    const int oh = ((volatile int *)child_height)[2 * parent + (side ^ 1)];
    if (side == 0) r = (uint32_t)other; else l = (uint32_t)other;
    if (sah_acc) sah_add(sah_acc, parent, sah_scale, lo, hi, cnt, slo, shi, r - l + 1 - cnt);
    if (cl_max && r - l + 1 > cl_max) {
      const uint32_t ocnt = r - l + 1 - cnt;
      // this child is [l, l + cnt) when it is the left one, else the sibling's ocnt triangles come first
      push_cluster(clusters, n_clusters, cl_max, side == 0 ? l : l + ocnt, cnt, parent, side);
      push_cluster(clusters, n_clusters, cl_max, side == 0 ? l + cnt : l, ocnt, parent, side ^ 1);
    }
    lo = mk(fminf(lo.x, slo.x), fminf(lo.y, slo.y), fminf(lo.z, slo.z));
    hi = mk(fmaxf(hi.x, shi.x), fmaxf(hi.y, shi.y), fmaxf(hi.z, shi.z));
    raw[parent].lo = (int)l;
    raw[parent].cnt = (int)(r - l + 1);
    height = (my_h > oh ? my_h : oh) + 1;
    cur = parent;
  }
}


// ---- binned surface-area splits inside clusters ----
// The radix tree cuts a cell where the Morton curve says, whatever its contents.  Below
// SAH_CLUSTER_MAX triangles the subtree is rebuilt top-down with the binned surface-area
// heuristic (8 bins on each axis of the centroid box, cost = area x count of both sides),
// which is what the traversal cost actually follows.  One CTA per cluster, everything in
// shared memory: padded triangle boxes, the slot permutation, per-warp bins and the segment
// lists of two successive levels.  A level's segments are taken by the CTA's warps; a warp
// bins its segment, picks the cheapest of the 21 cuts, partitions the slots (stable) and writes
// the node at raw[first slot of the right part - 1], the index the radix tree uses for the same
// cut position, so the node array stays collision-free and k_live / k_compact run unchanged.
// Deterministic: node indices and slot order depend on the geometry only.  CPU study of the
// tree quality: tools/tree_quality.cpp (small scenes: 25-40 % fewer node visits per ray).
#ifndef RTB_SAH_CLUSTER
#define RTB_SAH_CLUSTER 1024
#endif
constexpr uint32_t SAH_CLUSTER_MAX = RTB_SAH_CLUSTER;
// Scenes up to this size are ONE cluster: no sort, no radix tree.  5120 (the <5120, 512> instance,
// 191 KB of shared memory, one CTA) was measured and is not the default: the build takes 0.5-1.0 ms
// on one SM and the frames at 3840x2160 are a mixed bag against 1024-triangle clusters under a
// radix top (car-on-road 18.1 -> 17.3 ms, secret 17.1 -> 15.7, susans 5.00 -> 5.13, spheres 22.3 -> 26.1).
#ifndef RTB_SAH_WHOLE
#define RTB_SAH_WHOLE RTB_SAH_CLUSTER
#endif
constexpr uint32_t SAH_WHOLE_MAX = RTB_SAH_WHOLE;
constexpr int SAH_BINS = 8;              // 16 bins: same tree quality within 1 % (tools/tree_quality.cpp), twice the sweep
static_assert(SAH_WHOLE_MAX >= SAH_CLUSTER_MAX && SAH_WHOLE_MAX < 65536, "cluster-local indices are 16 bits wide");
constexpr int SAH_MEDIAN_LEVEL = 24;      // deeper segments are halved by position: depth <= 24 + log2(cluster)

struct SahSeg { unsigned short start, count; int parent_side; };   // parent_side = 2 * raw index + side, < 0: scene root

// warp-wide float min / max through the order-preserving integer mapping: one REDUX instruction
__device__ __forceinline__ float redux_min(float v) { return ord2f(__reduce_min_sync(0xffffffffu, f2ord(v))); }
__device__ __forceinline__ float redux_max(float v) { return ord2f(__reduce_max_sync(0xffffffffu, f2ord(v))); }

__device__ __forceinline__ int sah_bin(float c, float lo, float scale) {
  const int b = (int)((c - lo) * scale);
  return b < 0 ? 0 : (b > SAH_BINS - 1 ? SAH_BINS - 1 : b);
}

// CLMAX: largest cluster, T: threads per CTA.  <SAH_CLUSTER_MAX, 256> (42 KB of shared memory,
// 5 CTAs per SM) serves the clusters of a radix tree and scenes small enough to be one cluster;
// <SAH_WHOLE_MAX, 512> exists only when RTB_SAH_WHOLE is raised (A/B builds).
template <int CLMAX>
constexpr size_t sah_seg_cap() { return CLMAX / (LEAF_MAX + 1) + 1; }
template <int CLMAX, int T>
constexpr size_t sah_smem_bytes() {
  return (size_t)CLMAX * (6 * 4 + 4 + 2 * 2) + (size_t)(T / 32) * 3 * SAH_BINS * 7 * 4 + 2 * sah_seg_cap<CLMAX>() * sizeof(SahSeg) + 16;
}

#define BX(k, i) s_box[(k) * CLMAX + (i)]
template <int CLMAX, int T>
__global__ void __launch_bounds__(T) k_sah_clusters(const float *raw_v, uint32_t *order, const Cluster *clusters,
                                                    const uint32_t *n_clusters, RawNode *raw, int *root_info,
                                                    const uint32_t *bounds, float origin_radius) {
  constexpr int SAH_T = T, SAH_W = T / 32;
  constexpr int SEG_CAP = CLMAX / (LEAF_MAX + 1) + 1;      // = sah_seg_cap<CLMAX>()
  extern __shared__ __align__(16) unsigned char sah_smem[];
  float *s_box = reinterpret_cast<float *>(sah_smem);                          // [6][CLMAX] lo.xyz, hi.xyz by local id
  uint32_t *s_tri = reinterpret_cast<uint32_t *>(s_box + 6 * CLMAX);           // local id -> global triangle id
  uint32_t *s_bins = s_tri + CLMAX;                                            // [warps][3][SAH_BINS][7] min xyz, max xyz (ordered), count
  SahSeg *s_seg = reinterpret_cast<SahSeg *>(s_bins + SAH_W * 3 * SAH_BINS * 7);   // [2][SEG_CAP]
  unsigned *s_nseg = reinterpret_cast<unsigned *>(s_seg + 2 * SEG_CAP);        // [2]
  unsigned short *s_idx0 = reinterpret_cast<unsigned short *>(s_nseg + 2);     // slot -> local id
  unsigned short *s_idx1 = s_idx0 + CLMAX;                                     // partition scratch
  if (blockIdx.x >= *n_clusters) return;
  const Cluster cl = clusters[blockIdx.x];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float INF = __int_as_float(0x7f800000);
  {
    float amax = 0;
    for (int k = 0; k < 6; ++k) amax = fmaxf(amax, fabsf(ord2f(bounds[k])));
    const float pad_unit = SLIVER_K * 1.1920929e-7f * (amax + origin_radius);
    for (uint32_t i = threadIdx.x; i < cl.count; i += SAH_T) {
      const uint32_t tri = order[cl.start + i];
      V3 lo, hi;
      tri_box(raw_v, tri, lo, hi);
      sliver_pad(raw_v, tri, pad_unit, lo, hi);
      BX(0, i) = lo.x; BX(1, i) = lo.y; BX(2, i) = lo.z;
      BX(3, i) = hi.x; BX(4, i) = hi.y; BX(5, i) = hi.z;
      s_tri[i] = tri;
      s_idx0[i] = (unsigned short)i;
      if (i + 1 < cl.count) raw[cl.start + i].cnt = 0;      // the radix subtree goes away
    }
  }
  if (threadIdx.x == 0) {
    s_seg[0].start = 0; s_seg[0].count = (unsigned short)cl.count;
    s_seg[0].parent_side = cl.parent < 0 ? -1 : 2 * cl.parent + cl.side;
    s_nseg[0] = 1; s_nseg[1] = 0;
  }
  int cur = 0, level = 0;
  uint32_t *bins = s_bins + warp * 3 * SAH_BINS * 7;
  for (;;) {
    __syncthreads();                        // segments pushed by the previous level are complete
    const unsigned n_seg = s_nseg[cur];
    __syncthreads();
    if (!n_seg) break;
    if (threadIdx.x == 0) s_nseg[cur ^ 1] = 0;
    __syncthreads();
    for (unsigned sg = warp; sg < n_seg; sg += SAH_W) {
      const SahSeg seg = s_seg[cur * SEG_CAP + sg];
      const uint32_t st = seg.start, cnt = seg.count;
      // centroid box (warp reductions on order-preserving integers: one REDUX each)
      float cl0 = INF, cl1 = INF, cl2 = INF, ch0 = -INF, ch1 = -INF, ch2 = -INF;
      for (uint32_t i = lane; i < cnt; i += 32) {
        const int id = s_idx0[st + i];
        const float cx = 0.5f * BX(0, id) + 0.5f * BX(3, id), cy = 0.5f * BX(1, id) + 0.5f * BX(4, id),
                    cz = 0.5f * BX(2, id) + 0.5f * BX(5, id);
        cl0 = fminf(cl0, cx); cl1 = fminf(cl1, cy); cl2 = fminf(cl2, cz);
        ch0 = fmaxf(ch0, cx); ch1 = fmaxf(ch1, cy); ch2 = fmaxf(ch2, cz);
      }
      cl0 = redux_min(cl0); cl1 = redux_min(cl1); cl2 = redux_min(cl2);
      ch0 = redux_max(ch0); ch1 = redux_max(ch1); ch2 = redux_max(ch2);
      const float clo[3] = {cl0, cl1, cl2};
      float scale[3];
      scale[0] = ch0 > cl0 ? (float)SAH_BINS / (ch0 - cl0) : 0.0f;
      scale[1] = ch1 > cl1 ? (float)SAH_BINS / (ch1 - cl1) : 0.0f;
      scale[2] = ch2 > cl2 ? (float)SAH_BINS / (ch2 - cl2) : 0.0f;
      int cut = -1;                          // axis * SAH_BINS + last bin of the left part; -1: halve by position
      if (level < SAH_MEDIAN_LEVEL && (scale[0] > 0 || scale[1] > 0 || scale[2] > 0)) {
        for (int k = lane; k < 3 * SAH_BINS * 7; k += 32) bins[k] = (k % 7) < 3 ? 0xffffffffu : 0u;
        __syncwarp();
        for (uint32_t i = lane; i < cnt; i += 32) {
          const int id = s_idx0[st + i];
          const float lo[3] = {BX(0, id), BX(1, id), BX(2, id)}, hi[3] = {BX(3, id), BX(4, id), BX(5, id)};
          const uint32_t ol[3] = {f2ord(lo[0]), f2ord(lo[1]), f2ord(lo[2])}, oh[3] = {f2ord(hi[0]), f2ord(hi[1]), f2ord(hi[2])};
#pragma unroll
          for (int a = 0; a < 3; ++a) {
            if (!(scale[a] > 0)) continue;
            uint32_t *b = bins + (a * SAH_BINS + sah_bin(0.5f * lo[a] + 0.5f * hi[a], clo[a], scale[a])) * 7;
            atomicMin(b + 0, ol[0]); atomicMin(b + 1, ol[1]); atomicMin(b + 2, ol[2]);
            atomicMax(b + 3, oh[0]); atomicMax(b + 4, oh[1]); atomicMax(b + 5, oh[2]);
            atomicAdd(b + 6, 1u);
          }
        }
        __syncwarp();
        // sweep: lane = axis * SAH_BINS + bin holds its bin; prefix (left part) and suffix (right
        // part) unions by shuffles inside the groups of SAH_BINS lanes, all in ordered integers
        const int bb = lane & (SAH_BINS - 1);
        uint32_t pl[7], sf[7];
        {
          const uint32_t *q = bins + (lane < 3 * SAH_BINS ? lane : 0) * 7;
          const uint32_t m = lane < 3 * SAH_BINS ? q[6] : 0u;
#pragma unroll
          for (int k = 0; k < 3; ++k) {
            pl[k] = m ? q[k] : 0xff800000u;          // f2ord(+inf)
            pl[3 + k] = m ? q[3 + k] : 0x007fffffu;  // f2ord(-inf)
          }
          pl[6] = m;
#pragma unroll
          for (int k = 0; k < 7; ++k) sf[k] = pl[k];
        }
#pragma unroll
        for (int d = 1; d < SAH_BINS; d <<= 1) {
#pragma unroll
          for (int k = 0; k < 7; ++k) {
            const uint32_t up = __shfl_up_sync(0xffffffffu, pl[k], d, SAH_BINS);
            const uint32_t dn = __shfl_down_sync(0xffffffffu, sf[k], d, SAH_BINS);
            if (bb >= d) pl[k] = k < 3 ? min(pl[k], up) : (k < 6 ? max(pl[k], up) : pl[k] + up);
            if (bb + d < SAH_BINS) sf[k] = k < 3 ? min(sf[k], dn) : (k < 6 ? max(sf[k], dn) : sf[k] + dn);
          }
        }
        uint32_t rt[7];                      // right part of the cut after bin bb: suffix of the next lane
#pragma unroll
        for (int k = 0; k < 7; ++k) rt[k] = __shfl_down_sync(0xffffffffu, sf[k], 1, SAH_BINS);
        uint32_t key = 0xffffffffu;
        if (lane < 3 * SAH_BINS && bb < SAH_BINS - 1 && pl[6] && rt[6]) {
          const float cost =
              box_area(mk(ord2f(pl[0]), ord2f(pl[1]), ord2f(pl[2])), mk(ord2f(pl[3]), ord2f(pl[4]), ord2f(pl[5]))) * (float)pl[6] +
              box_area(mk(ord2f(rt[0]), ord2f(rt[1]), ord2f(rt[2])), mk(ord2f(rt[3]), ord2f(rt[4]), ord2f(rt[5]))) * (float)rt[6];
          // costs are >= 0, so their bit patterns order like integers; the low 5 bits carry the lane
          if (cost >= 0.0f && cost < INF) key = (__float_as_uint(cost) & ~31u) | (uint32_t)lane;
        }
        key = __reduce_min_sync(0xffffffffu, key);
        if (key != 0xffffffffu) cut = (int)(key & 31u);
      }
      // stable partition of the segment's slots: left part first
      uint32_t nleft = 0;
      V3 llo, lhi, rlo, rhi;
      for (int attempt = 0; attempt < 2; ++attempt) {
        const int axis = cut < 0 ? 0 : cut / SAH_BINS, kbin = cut < 0 ? 0 : cut % SAH_BINS;
        llo = mk(INF, INF, INF); lhi = mk(-INF, -INF, -INF); rlo = llo; rhi = lhi;
        uint32_t run = 0;
        for (int side = 0; side < 2; ++side) {
          for (uint32_t base = 0; base < cnt; base += 32) {
            const uint32_t i = base + lane;
            bool take = false;
            int id = 0;
            if (i < cnt) {
              id = s_idx0[st + i];
              bool left;
              if (cut < 0) left = i < cnt / 2;
              else left = sah_bin(0.5f * BX(axis, id) + 0.5f * BX(3 + axis, id), clo[axis], scale[axis]) <= kbin;
              take = left == (side == 0);
            }
            const unsigned m = __ballot_sync(0xffffffffu, take);
            if (take) {
              s_idx1[st + run + __popc(m & ((1u << lane) - 1u))] = (unsigned short)id;
              const V3 blo = mk(BX(0, id), BX(1, id), BX(2, id)), bhi = mk(BX(3, id), BX(4, id), BX(5, id));
              if (side == 0) { grow(llo, lhi, blo); grow(llo, lhi, bhi); }
              else { grow(rlo, rhi, blo); grow(rlo, rhi, bhi); }
            }
            run += __popc(m);
          }
          if (side == 0) nleft = run;
        }
        if (nleft > 0 && nleft < cnt) break;
        cut = -1;                            // cannot happen for a cut with both sides non-empty; be safe
      }
      __syncwarp();
      for (uint32_t i = lane; i < cnt; i += 32) s_idx0[st + i] = s_idx1[st + i];
      llo = mk(redux_min(llo.x), redux_min(llo.y), redux_min(llo.z)); lhi = mk(redux_max(lhi.x), redux_max(lhi.y), redux_max(lhi.z));
      rlo = mk(redux_min(rlo.x), redux_min(rlo.y), redux_min(rlo.z)); rhi = mk(redux_max(rhi.x), redux_max(rhi.y), redux_max(rhi.z));
      if (lane == 0) {
        const uint32_t first = cl.start + st, nright = cnt - nleft;
        const int p = (int)(first + nleft - 1);
        const int ref0 = nleft <= (uint32_t)LEAF_MAX ? leaf_ref(first, nleft) : 0;
        const int ref1 = nright <= (uint32_t)LEAF_MAX ? leaf_ref(first + nleft, nright) : 0;
        float4 *dst = reinterpret_cast<float4 *>(&raw[p]);
        dst[0] = make_float4(llo.x, lhi.x, llo.y, lhi.y);
        dst[1] = make_float4(rlo.x, rhi.x, rlo.y, rhi.y);
        dst[2] = make_float4(llo.z, lhi.z, rlo.z, rhi.z);
        dst[3] = make_float4(__int_as_float(ref0), __int_as_float(ref1), __int_as_float((int)first), __int_as_float((int)cnt));
        if (seg.parent_side >= 0) (&raw[seg.parent_side >> 1].ref0)[seg.parent_side & 1] = p;
        else root_info[0] = p;
        if (nleft > (uint32_t)LEAF_MAX) {
          SahSeg &c = s_seg[(cur ^ 1) * SEG_CAP + atomicAdd(&s_nseg[cur ^ 1], 1u)];
          c.start = (unsigned short)st; c.count = (unsigned short)nleft; c.parent_side = 2 * p;
        }
        if (nright > (uint32_t)LEAF_MAX) {
          SahSeg &c = s_seg[(cur ^ 1) * SEG_CAP + atomicAdd(&s_nseg[cur ^ 1], 1u)];
          c.start = (unsigned short)(st + nleft); c.count = (unsigned short)nright; c.parent_side = 2 * p + 1;
        }
      }
      __syncwarp();
    }
    cur ^= 1;
    ++level;
  }
  for (uint32_t i = threadIdx.x; i < cl.count; i += SAH_T) order[cl.start + i] = s_tri[s_idx0[i]];
  if (threadIdx.x == 0) atomicMax(root_info + 2, level);
}
#undef BX

// live[j] = 1 when raw node j survives (its range is larger than a leaf)
__global__ void __launch_bounds__(BT) k_live(const RawNode *raw, uint32_t n_inner, uint32_t *live) {
  const uint32_t j = blockIdx.x * BT + threadIdx.x;
  if (j < n_inner) live[j] = raw[j].cnt > LEAF_MAX ? 1u : 0u;
}


// new_index = exclusive scan of live.  stats: [0] leaves, [1] max leaf size
// node_up / slot_up: the upward links of rtb_internal.h (the root's own entry is set by the host)
__global__ void __launch_bounds__(BT) k_compact(const RawNode *raw, uint32_t n_inner, const uint32_t *new_index,
                                                float4 *nodes, unsigned int *stats, uint32_t *node_up, uint32_t *slot_up) {
  const uint32_t j = blockIdx.x * BT + threadIdx.x;
  if (j >= n_inner) return;
  RawNode nd = raw[j];
  if (nd.cnt <= LEAF_MAX) return;
  unsigned leaves = 0, maxleaf = 0;
  const uint32_t me = new_index[j];
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    int &ref = side ? nd.ref1 : nd.ref0;
    if (ref >= 0) {
      ref = (int)new_index[ref];
      if (node_up) node_up[ref] = 2u * me + (uint32_t)side;
    } else {
      const uint32_t packed = (uint32_t)~ref, first = packed >> 3, c = packed & 7u;
      ++leaves; maxleaf = max(maxleaf, c);
      if (slot_up) for (uint32_t k = 0; k < c; ++k) slot_up[first + k] = 2u * me + (uint32_t)side;
    }
  }
  float4 *dst = nodes + 4ull * me;
  dst[0] = nd.q0;
  dst[1] = nd.q1;
  dst[2] = nd.q2;
  dst[3] = make_float4(__int_as_float(nd.ref0), __int_as_float(nd.ref1), __int_as_float(nd.lo), __int_as_float(nd.cnt));
  if (leaves) { atomicAdd(&stats[0], leaves); atomicMax(&stats[1], maxleaf); }
}


// Scenes of at most LEAF_MAX triangles (and brute-force-only scenes): one node whose first
// child is a leaf over everything with an unbounded box (nothing to cull, so nothing to pad),
// second child empty.
__global__ void k_single_node(uint32_t n, float4 *nodes) {
  const float INF = __int_as_float(0x7f800000);
  float lx = INF, ly = INF, lz = INF, hx = -INF, hy = -INF, hz = -INF;
  if (n) { lx = ly = lz = -INF; hx = hy = hz = INF; }
  nodes[0] = make_float4(lx, hx, ly, hy);
  nodes[1] = make_float4(INF, -INF, INF, -INF);
  nodes[2] = make_float4(lz, hz, INF, -INF);
  nodes[3] = make_float4(__int_as_float(leaf_ref(0, n)), __int_as_float(leaf_ref(0, 0)), __int_as_float(0),
                         __int_as_float((int)n));
}

double ms_between(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  cudaEventElapsedTime(&ms, a, b);
  return ms;
}

uint32_t pack_host_color(float r, float g, float b) {
  // init_color, gpu/colors.cu:3-21 (host build uses -ffp-contract=off)
  auto q = [](float x) -> uint32_t {
    x = x * 255;
    if (x > 255) x = 255;
    if (x < 0) x = 0;
    if (x != x) return 0u;
    return (uint32_t)(uint8_t)x;
  };
  return q(r) | (q(g) << 8) | (q(b) << 16) | 0xff000000u;
}
float bits_to_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

}  // namespace

SceneStore::~SceneStore() { release(); }

// Scene arrays come from the device's stream-ordered memory pool (cudaMallocAsync; api.cu
// raises its release threshold), so a one-shot render() does not pay ~20 cudaMalloc/cudaFree
// round trips per call the way the reference does (gpu/scene.cu:236-308, octree.cu:367-411).
// Stream, events and frame buffers go back to a per-device cache (recycle_frame_ctx).
void SceneStore::release() {
  if (device >= 0) cudaSetDevice(device);
  // (d_prim, d_rast_items, d_rast go back to the per-device cache with the stream: recycle_frame_ctx)
  void *ptrs[] = {v0, e1, e2, n0, n1, n2, nodes, materials, lights, order, up, inv_order, d_sched, lgrids, d_fine_acc, d_fine_done, d_stray};
  for (void *p : ptrs) if (p) block_free(p, stream);
  for (void *&p : lgrid_mem) { if (p) block_free(p, stream); p = nullptr; }
  lgrids = nullptr; n_lgrids = 0; lgrid_entries = lgrid_cells = 0;
  v0 = e1 = e2 = n0 = n1 = n2 = nodes = materials = lights = nullptr;
  order = nullptr; up = nullptr; inv_order = nullptr;
  d_fine_acc = nullptr; d_fine_done = nullptr; fine_pixels = fine_items = 0;
  d_stray = nullptr; stray_pass = 0; stray_items = 0; tree_deferred = false;
  if (h_rast) { cudaFreeHost(h_rast); h_rast = nullptr; }
  if (xchg_block) { cudaStreamSynchronize(stream); cudaFree(xchg_block); xchg_block = nullptr; xchg_nranks = 0; xchg_rank = -1; }
  rast_pending = false; rast_useless = false;
  d_sched = nullptr; d_item_order = nullptr; sched_items = 0;
  if (stream) cudaStreamSynchronize(stream);
  for (cudaEvent_t &e : kev) { if (e) cudaEventDestroy(e); e = nullptr; }
  if (xev) { cudaEventDestroy(xev); xev = nullptr; }
  recycle_frame_ctx(*this);
}

#define CK(call) do { int rc__ = (call); if (rc__) return rc__; } while (0)

static thread_local cudaStream_t tl_alloc_stream = nullptr;   // set by build_scene

template <typename T>
static int dev_alloc(T **p, size_t count) {
  *p = nullptr;
  return block_alloc((void **)p, (count ? count : 1) * sizeof(T), tl_alloc_stream);
}

struct Temp {   // build-time scratch, freed (stream-ordered) when build_scene returns
  std::vector<void *> ptrs;
  ~Temp() { for (void *p : ptrs) block_free(p, tl_alloc_stream); }
  template <typename T> int alloc(T **p, size_t count) {
    int rc = dev_alloc(p, count);
    if (!rc) ptrs.push_back(*p);
    return rc;
  }
};

// How far from the world origin a primary-ray origin of this camera can be: the image plane
// lies L = W'/(2 tan(fov/2)) behind the camera (gpu/raytracer.cu:181-182).
float camera_origin_radius(const struct camera &cm) {
  const double L = cm.fov > 0 && cm.fov < 360 ? cm.width / (2 * tan(cm.fov * M_PI / 360)) : 0;
  const double pm = fmax(fmax(fabs(cm.position.x), fabs(cm.position.y)), fabs(cm.position.z));
  return (float)(pm + fabs(L) + 0.5 * (abs(cm.width) + abs(cm.height)));
}

int build_scene(const struct scene *hs, unsigned flags, SceneStore *S, UploadGroup *grp, int member) {
  if (!hs) { set_error("scene is NULL"); return RTB_ERR_ARG; }
  const size_t n_obj = hs->object_count, n_light = hs->light_count;
  if (n_obj && !hs->objects) { set_error("object_count > 0 but objects is NULL"); return RTB_ERR_ARG; }
  if (n_light && !hs->lights) { set_error("light_count > 0 but lights is NULL"); return RTB_ERR_ARG; }
  std::vector<uint32_t> first(n_obj + 1, 0);
  size_t total = 0;
  for (size_t o = 0; o < n_obj; ++o) {
    first[o] = (uint32_t)total;
    total += hs->objects[o].triangle_count;
    if (hs->objects[o].triangle_count && (!hs->objects[o].triangles.vertex || !hs->objects[o].triangles.normal)) {
      set_error("object %zu has triangles but NULL vertex/normal arrays", o);
      return RTB_ERR_ARG;
    }
    if (total >= (1ull << 28)) { set_error("more than 2^28-1 triangles"); return RTB_ERR_LIMIT; }
  }
  first[n_obj] = (uint32_t)total;
  if (hs->triangle_count != total) {
    set_error("scene.triangle_count (%zu) != sum of object triangle counts (%zu)", hs->triangle_count, total);
    return RTB_ERR_ARG;
  }
  const uint32_t n = (uint32_t)total;
  cudaStream_t st = S->stream;
  tl_alloc_stream = st;
  S->origin_radius = 2.0f * camera_origin_radius(hs->camera);   // x2 head-room for camera moves
  S->one_frame = (flags & RTB_BUILD_ONE_FRAME) != 0;
  Temp tmp;
  cudaEvent_t e_up0 = S->ev0, e_up1 = S->ev1;
  cudaEvent_t e_b1 = nullptr;
  RTB_CUDA(cudaEventCreate(&e_b1));
  struct EvGuard { cudaEvent_t e; ~EvGuard() { cudaEventDestroy(e); } } evg{e_b1};

  // ---- upload (to_cuda, gpu/scene.cu:224-352): raw float3 arrays, object table, lights ----
  // Big scenes: only the vertex arrays go up on the build stream; the normals (needed by the
  // last kernel of the build alone) follow on a second stream and cross PCIe while the
  // hierarchy is being built.
  const bool overlap = n >= (1u << 20);
  struct Side {
    cudaStream_t st = nullptr;
    cudaEvent_t verts = nullptr, normals = nullptr;
    ~Side() {
      if (st) { cudaStreamSynchronize(st); cudaStreamDestroy(st); }
      if (verts) cudaEventDestroy(verts);
      if (normals) cudaEventDestroy(normals);
    }
  } side;
  if (overlap) {
    RTB_CUDA(cudaStreamCreateWithFlags(&side.st, cudaStreamNonBlocking));
    RTB_CUDA(cudaEventCreateWithFlags(&side.verts, cudaEventDisableTiming));
    RTB_CUDA(cudaEventCreateWithFlags(&side.normals, cudaEventDisableTiming));
  }
  struct Span { const vector3 *src; size_t first, tri; };
  std::vector<Span> normal_spans;
  RTB_CUDA(cudaEventRecord(e_up0, st));
  float *raw_v = nullptr, *raw_n = nullptr;
  uint32_t *d_first = nullptr;
  if (grp && member == 0) { raw_v = grp->src_v; raw_n = grp->src_n; }      // shared upload (UploadGroup, scene_store.h)
  else {
    CK(tmp.alloc(&raw_v, 9ull * n));
    CK(tmp.alloc(&raw_n, 9ull * n));
  }
  CK(tmp.alloc(&d_first, n_obj + 1));
  // objects whose arrays are adjacent in host memory (the SoA scene) go up in one copy
  struct Run { const vector3 *pv, *pn; size_t first, tri; };
  std::vector<Run> runs;
  for (size_t o = 0; o < n_obj;) {
    size_t e = o;
    const vector3 *pv = hs->objects[o].triangles.vertex, *pn = hs->objects[o].triangles.normal;
    size_t tri = hs->objects[o].triangle_count;
    while (e + 1 < n_obj && hs->objects[e + 1].triangles.vertex == pv + 3 * tri &&
           hs->objects[e + 1].triangles.normal == pn + 3 * tri) {
      ++e;
      tri += hs->objects[e].triangle_count;
    }
    if (tri) runs.push_back({pv, pn, first[o], tri});
    o = e + 1;
  }
  const bool fetch = grp && member != 0;      // the arrays come from member 0's device, not from the host
  if (!fetch)
    for (const Run &r : runs) {
      RTB_CUDA(cudaMemcpyAsync(raw_v + 9ull * r.first, r.pv, 36ull * r.tri, cudaMemcpyHostToDevice, st));
      if (overlap) normal_spans.push_back({r.pn, r.first, r.tri});
      else RTB_CUDA(cudaMemcpyAsync(raw_n + 9ull * r.first, r.pn, 36ull * r.tri, cudaMemcpyHostToDevice, st));
    }
  else {
    if (!grp->barrier()) { set_error("device %d of the call failed during the upload", grp->device[0]); return RTB_ERR_CUDA; }
    RTB_CUDA(cudaStreamWaitEvent(st, grp->ready_v, 0));
    RTB_CUDA(cudaMemcpyPeerAsync(raw_v, grp->device[member], grp->src_v, grp->device[0], 36ull * n, st));
    if (!overlap) {
      RTB_CUDA(cudaStreamWaitEvent(st, grp->ready_n, 0));
      RTB_CUDA(cudaMemcpyPeerAsync(raw_n, grp->device[member], grp->src_n, grp->device[0], 36ull * n, st));
    }
  }
  if (overlap) {
    RTB_CUDA(cudaEventRecord(side.verts, st));               // also orders raw_n's allocation before its use
    RTB_CUDA(cudaStreamWaitEvent(side.st, side.verts, 0));   // vertices first: they gate the build
    for (const Span &sp : normal_spans)
      RTB_CUDA(cudaMemcpyAsync(raw_n + 9ull * sp.first, sp.src, 36ull * sp.tri, cudaMemcpyHostToDevice, side.st));
    if (fetch) {
      RTB_CUDA(cudaStreamWaitEvent(side.st, grp->ready_n, 0));
      RTB_CUDA(cudaMemcpyPeerAsync(raw_n, grp->device[member], grp->src_n, grp->device[0], 36ull * n, side.st));
    }
    RTB_CUDA(cudaEventRecord(side.normals, side.st));
  }
  if (grp && member == 0) {
    // the peers may start: vertices are behind ready_v on the build stream, normals behind ready_n
    RTB_CUDA(cudaEventRecord(grp->ready_v, st));
    RTB_CUDA(cudaEventRecord(grp->ready_n, overlap ? side.st : st));
    if (!grp->barrier()) { set_error("a device of the call failed during the upload"); return RTB_ERR_CUDA; }
  }
  RTB_CUDA(cudaMemcpyAsync(d_first, first.data(), (n_obj + 1) * 4, cudaMemcpyHostToDevice, st));
  std::vector<float> mat(8 * (n_obj ? n_obj : 1), 0.0f);
  for (size_t o = 0; o < n_obj; ++o) {
    const struct object &ob = hs->objects[o];
    float *m = &mat[8 * o];
    m[0] = bits_to_float(pack_host_color(ob.ka.x, ob.ka.y, ob.ka.z));
    m[1] = bits_to_float(pack_host_color(ob.kd.x, ob.kd.y, ob.kd.z));
    m[2] = bits_to_float(pack_host_color(ob.ks.x, ob.ks.y, ob.ks.z));
    m[3] = ob.ns;
    m[4] = ob.nr;
  }
  std::vector<float> lig(8 * (n_light ? n_light : 1), 0.0f);
  for (size_t i = 0; i < n_light; ++i) {
    const struct light &l = hs->lights[i];
    float *m = &lig[8 * i];
    m[0] = bits_to_float((uint32_t)l.type);
    m[1] = bits_to_float(pack_host_color(l.r, l.g, l.b));
    m[4] = l.v.x; m[5] = l.v.y; m[6] = l.v.z;
  }
  CK(dev_alloc(&S->materials, 2 * (n_obj ? n_obj : 1)));
  CK(dev_alloc(&S->lights, 2 * (n_light ? n_light : 1)));
  RTB_CUDA(cudaMemcpyAsync(S->materials, mat.data(), mat.size() * 4, cudaMemcpyHostToDevice, st));
  RTB_CUDA(cudaMemcpyAsync(S->lights, lig.data(), lig.size() * 4, cudaMemcpyHostToDevice, st));
  RTB_CUDA(cudaEventRecord(e_up1, st));

  // ---- bounds ----
  uint32_t *d_bounds = nullptr;
  CK(tmp.alloc(&d_bounds, 8));
  const uint32_t init_bounds[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
  RTB_CUDA(cudaMemcpyAsync(d_bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, st));
  unsigned long long launches = 0;
  if (n) {
    unsigned g = (n + BT - 1) / BT;
    if (g > 148u * 8u) g = 148u * 8u;
    k_bounds<<<g, BT, 0, st>>>(raw_v, n, d_bounds);
    ++launches;
  }

  CK(dev_alloc(&S->v0, n)); CK(dev_alloc(&S->e1, n)); CK(dev_alloc(&S->e2, n));
  CK(dev_alloc(&S->n0, n)); CK(dev_alloc(&S->n1, n)); CK(dev_alloc(&S->n2, n));
  CK(dev_alloc(&S->order, n));

  const bool tree = !(flags & RTB_BUILD_NO_TREE) && n > (uint32_t)LEAF_MAX;
  const unsigned gN = (n + BT - 1) / BT;
  uint64_t *keys = nullptr;
  // build scratch of the hierarchy (also used by the sample trees of the key-mode choice)
  RawNode *raw = nullptr;
  int *visit = nullptr, *child_h = nullptr, *root_info = nullptr, *raw_parent = nullptr;
  int *d_mode = nullptr;                      // [0] chosen key mode, [1..4] sample-tree costs
  S->key_mode = 0;
  for (int k = 0; k < 4; ++k) S->mode_cost[k] = 0;
  const uint32_t cl_max =
      (flags & RTB_BUILD_RADIX_ONLY) || ((flags & RTB_BUILD_ONE_FRAME) && n > ONE_FRAME_SAH_MAX) ? 0u : SAH_CLUSTER_MAX;
  const bool sah_only = tree && cl_max && n <= SAH_WHOLE_MAX;    // the whole scene is one cluster: no sort, no radix tree
  // Deferred tree (rtb_render on big scenes whose frame will take its first rays from the projection pass and its shadow
  // rays from the light grids): everything the frame needs except the nodes -- keys, sort, slot order, permuted
  // triangles, grids.  k_emit, k_live, the scan, k_compact, their 1.3 GB of scratch and two host round trips are left out.
  const bool one_frame_grids = (unsigned long long)hs->camera.width * (unsigned long long)hs->camera.height >= 4ull * n &&
                               !(flags & RTB_BUILD_NO_LIGHT_GRIDS);
  const bool defer = tree && !sah_only && cl_max == 0 && (flags & RTB_BUILD_DEFER_TREE) && (flags & RTB_BUILD_ONE_FRAME) &&
                     n >= (1u << 15) && one_frame_grids;
  Cluster *d_clusters = nullptr;
  uint32_t *d_n_clusters = nullptr;
  const uint32_t clusters_cap = cl_max ? 2u * (n / cl_max) + 2u : 1u;   // sibling clusters hold > cl_max triangles together
  S->builder = 0;
  if (tree) {
    const uint32_t n_inner = defer ? 1u : n - 1;   // deferred: no tree, no sample trees
    CK(tmp.alloc(&raw, n_inner)); CK(tmp.alloc(&visit, n_inner)); CK(tmp.alloc(&child_h, 2ull * n_inner));
    CK(tmp.alloc(&root_info, 4)); CK(tmp.alloc(&raw_parent, n_inner));
    CK(tmp.alloc(&d_mode, 8));
    CK(tmp.alloc(&d_clusters, clusters_cap)); CK(tmp.alloc(&d_n_clusters, 1));
    RTB_CUDA(cudaMemsetAsync(d_mode, 0, 8 * sizeof(int), st));
    RTB_CUDA(cudaMemsetAsync(d_n_clusters, 0, 4, st));
    const int init_root[4] = {-1, 0, 0, 0};      // [0] raw root, [1] radix height, [2] deepest cluster
    RTB_CUDA(cudaMemcpyAsync(root_info, init_root, sizeof init_root, cudaMemcpyHostToDevice, st));
  }
  if (tree && !sah_only) {
    uint64_t *keys_a = nullptr, *keys_b = nullptr;
    uint32_t *vals_b = nullptr, *scratch = nullptr;
    const size_t sw = radix_sort_scratch_words(n);
    CK(tmp.alloc(&keys_a, n)); CK(tmp.alloc(&keys_b, n)); CK(tmp.alloc(&vals_b, n)); CK(tmp.alloc(&scratch, sw));
    // ---- key mode: surface-area cost of two candidate trees over a strided sample: cubic keys,
    //      and the axis the surface is most perpendicular to last ----
    // (without a tree the layout only orders the slots, and the frame runs alike on either order: 19.89 / 19.98 ms on C5)
    const bool choose = n >= MODE_CHOICE_MIN && !(flags & RTB_BUILD_CUBIC_KEYS) && !defer;
    if (choose) {
      const uint32_t stride = (n + MODE_SAMPLE_MAX - 1) / MODE_SAMPLE_MAX;
      const uint32_t m = (n + stride - 1) / stride;
      unsigned long long *d_cost = nullptr;     // [2][SAH_SLOTS] tree costs, then [3] normal statistic
      int *d_cand = nullptr;
      CK(tmp.alloc(&d_cost, 2 * SAH_SLOTS + 3));
      CK(tmp.alloc(&d_cand, 1));
      RTB_CUDA(cudaMemsetAsync(d_cost, 0, (2 * SAH_SLOTS + 3) * sizeof(unsigned long long), st));
      const unsigned gM = (m + BT - 1) / BT;
      k_normal_stat<<<gM, BT, 0, st>>>(raw_v, m, stride, d_bounds, d_cost + 2 * SAH_SLOTS);
      k_pick_axis<<<1, 32, 0, st>>>(d_cost + 2 * SAH_SLOTS, d_cand);
      launches += 2;
      for (int c = 0; c < 2; ++c) {
        k_morton<<<gM, BT, 0, st>>>(raw_v, m, stride, d_bounds, 0, c ? d_cand : nullptr, keys_a, S->order);
        int in_alt = 0;
        CK(dev_radix_sort_pairs_u64(keys_a, S->order, keys_b, vals_b, m, MORTON_BITS, scratch, sw, st, &in_alt, &launches));
        RTB_CUDA(cudaMemsetAsync(visit, 0xff, 4ull * (m - 1), st));
        k_emit<<<gM, BT, 0, st>>>(raw_v, in_alt ? vals_b : S->order, in_alt ? keys_b : keys_a, m, raw, visit, child_h,
                                  root_info, d_bounds, S->origin_radius, raw_parent, d_cost + c * SAH_SLOTS, nullptr,
                                  nullptr, 0u);
        launches += 2;
      }
      k_choose_mode<<<1, 32, 0, st>>>(d_cost, d_cand, d_mode);
      ++launches;
    }
    k_morton<<<gN, BT, 0, st>>>(raw_v, n, 1, d_bounds, 0, d_mode, keys_a, S->order);
    ++launches;
    int in_alt = 0;
    // without a tree the order only has to keep neighbours together: the top 32 of the 63 key bits, four passes instead of eight
    CK(dev_radix_sort_pairs_u64(keys_a, S->order, keys_b, vals_b, n, MORTON_BITS, scratch, sw, st, &in_alt, &launches, defer ? MORTON_BITS - 32 : 0));
    keys = keys_a;
    if (in_alt) {
      keys = keys_b;
      RTB_CUDA(cudaMemcpyAsync(S->order, vals_b, 4ull * n, cudaMemcpyDeviceToDevice, st));
    }
  } else if (n) {
    k_iota<<<gN, BT, 0, st>>>(S->order, n);
    ++launches;
  }

  S->n_nodes = 1; S->n_leaves = n ? 1 : 0; S->max_leaf = n; S->height = 1; S->root = 0;
  uint32_t *live = nullptr, *scan_scratch = nullptr;
  unsigned int *d_stats = nullptr;
  size_t ssw = 0;
  if (tree && !defer) {
    // ---- topology: radix tree over the sorted keys, clusters rebuilt with surface-area splits ----
    const uint32_t n_inner = n - 1;
    ssw = scan_scratch_words(n_inner + 1);
    CK(tmp.alloc(&live, n_inner + 1)); CK(tmp.alloc(&scan_scratch, ssw));
    CK(tmp.alloc(&d_stats, 2));
    RTB_CUDA(cudaMemsetAsync(visit, 0xff, 4ull * n_inner, st));
    RTB_CUDA(cudaMemsetAsync(raw, 0, sizeof(RawNode) * (size_t)n_inner, st));
    RTB_CUDA(cudaMemsetAsync(d_stats, 0, 8, st));
    RTB_CUDA(cudaMemsetAsync(live, 0, 4ull * (n_inner + 1), st));
    if (sah_only) {
      const Cluster whole = {0u, n, -1, 0};
      const uint32_t one = 1;
      RTB_CUDA(cudaMemcpyAsync(d_clusters, &whole, sizeof whole, cudaMemcpyHostToDevice, st));
      RTB_CUDA(cudaMemcpyAsync(d_n_clusters, &one, 4, cudaMemcpyHostToDevice, st));
    } else {
      k_emit<<<gN, BT, 0, st>>>(raw_v, S->order, keys, n, raw, visit, child_h, root_info, d_bounds, S->origin_radius,
                                raw_parent, nullptr, d_clusters, d_n_clusters, cl_max);
      ++launches;
    }
    if (cl_max) {
      if (sah_only && n > SAH_CLUSTER_MAX) {
        constexpr size_t smem = sah_smem_bytes<(int)SAH_WHOLE_MAX, 512>();
        static_assert(smem <= 227 * 1024, "shared memory of the whole-scene instance");
        auto kern = k_sah_clusters<(int)SAH_WHOLE_MAX, 512>;
        RTB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<1, 512, smem, st>>>(raw_v, S->order, d_clusters, d_n_clusters, raw, root_info, d_bounds, S->origin_radius);
      } else {
        constexpr size_t smem = sah_smem_bytes<(int)SAH_CLUSTER_MAX, 256>();
        static_assert(smem <= 48 * 1024, "shared memory of the cluster instance");
        k_sah_clusters<(int)SAH_CLUSTER_MAX, 256><<<sah_only ? 1u : clusters_cap, 256, smem, st>>>(
            raw_v, S->order, d_clusters, d_n_clusters, raw, root_info, d_bounds, S->origin_radius);
      }
      RTB_CUDA(cudaGetLastError());
      ++launches;
      S->builder = 1;
    }
  }
  // the slot order is final now: triangles into leaf order
  if (n && tree) {
    CK(dev_alloc(&S->inv_order, n));
    k_invert<<<gN, BT, 0, st>>>(S->order, n, S->inv_order);
    ++launches;
  }
  if (n) {
    if (overlap)
      k_permute<true, false><<<gN, BT, 0, st>>>(raw_v, raw_n, S->order, n, d_first, (uint32_t)n_obj, S->v0, S->e1,
                                                S->e2, S->n0, S->n1, S->n2);
    else
      k_permute<true, true><<<gN, BT, 0, st>>>(raw_v, raw_n, S->order, n, d_first, (uint32_t)n_obj, S->v0, S->e1,
                                               S->e2, S->n0, S->n1, S->n2);
    ++launches;
  }
  if (tree && defer) {
    // no nodes: one dummy node keeps the pointers valid, the frame never reads it (FrameArgs::need_tree)
    int h_mode[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    RTB_CUDA(cudaMemcpyAsync(h_mode, d_mode, sizeof h_mode, cudaMemcpyDeviceToHost, st));
    CK(dev_alloc(&S->nodes, NODE_QUADS));
    k_single_node<<<1, 1, 0, st>>>(0u, S->nodes);
    ++launches;
    RTB_CUDA(cudaStreamSynchronize(st));
    S->key_mode = h_mode[0];
    for (int k = 0; k < 4; ++k) S->mode_cost[k] = h_mode[1 + k];
    S->n_nodes = 1; S->root = 0; S->height = 1; S->n_leaves = 0; S->max_leaf = 0;
  } else if (tree) {
    const uint32_t n_inner = n - 1;
    k_live<<<(n_inner + BT - 1) / BT, BT, 0, st>>>(raw, n_inner, live);
    ++launches;
    CK(dev_exclusive_scan_u32(live, n_inner + 1, scan_scratch, ssw, false, st, &launches));
    uint32_t n_live = 0;
    int h_root[4] = {0, 0, 0, 0};
    RTB_CUDA(cudaMemcpyAsync(&n_live, live + n_inner, 4, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaMemcpyAsync(h_root, root_info, sizeof h_root, cudaMemcpyDeviceToHost, st));
    int h_mode[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    RTB_CUDA(cudaMemcpyAsync(h_mode, d_mode, sizeof h_mode, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    S->key_mode = h_mode[0];
    for (int k = 0; k < 4; ++k) S->mode_cost[k] = h_mode[1 + k];
    const int height = h_root[1] + h_root[2];     // radix levels above the clusters + the deepest cluster
    if (n_live == 0 || h_root[0] < 0) { set_error("hierarchy build produced no root"); return RTB_ERR_CUDA; }
    if (height > HEIGHT_MAX) { set_error("hierarchy depth %d exceeds %d", height, HEIGHT_MAX); return RTB_ERR_LIMIT; }
    CK(dev_alloc(&S->nodes, (size_t)NODE_QUADS * n_live));
    k_compact<<<(n_inner + BT - 1) / BT, BT, 0, st>>>(raw, n_inner, live, S->nodes, d_stats, S->up,
                                                      S->up ? S->up + n_live : nullptr);
    ++launches;
    uint32_t new_root = 0;
    unsigned int h_stats[2] = {0, 0};
    RTB_CUDA(cudaMemcpyAsync(&new_root, live + h_root[0], 4, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaMemcpyAsync(h_stats, d_stats, 8, cudaMemcpyDeviceToHost, st));
    RTB_CUDA(cudaStreamSynchronize(st));
    if (S->up) RTB_CUDA(cudaMemsetAsync(S->up + new_root, 0xff, 4, st));      // the root has no parent
    S->n_nodes = n_live; S->root = (int)new_root; S->height = height;
    S->n_leaves = h_stats[0]; S->max_leaf = h_stats[1];
  } else {
    CK(dev_alloc(&S->nodes, NODE_QUADS));
    k_single_node<<<1, 1, 0, st>>>(n, S->nodes);
    ++launches;
  }
  // RTB_BUILD_TRACE=1: phase times of the build on stderr (tree | light grids | wait for normals + permute)
  static const bool trace_build = [] { const char *e = getenv("RTB_BUILD_TRACE"); return e && atoi(e) != 0; }();
  cudaEvent_t e_t[3] = {nullptr, nullptr, nullptr};
  if (trace_build) { for (auto &e : e_t) cudaEventCreate(&e); cudaEventRecord(e_t[0], st); }
  uint32_t h_bounds[8];
  RTB_CUDA(cudaMemcpyAsync(h_bounds, d_bounds, sizeof h_bounds, cudaMemcpyDeviceToHost, st));
  RTB_CUDA(cudaStreamSynchronize(st));
  RTB_CUDA(cudaGetLastError());
  float amax = 0;
  if (n) for (int k = 0; k < 6; ++k) { float f = ord2f(h_bounds[k]); f = f < 0 ? -f : f; if (f > amax) amax = f; }
  S->abs_max = amax;
  S->pad_unit = SLIVER_K * 1.1920929e-7f * (amax + S->origin_radius);     // as k_emit / k_sah_clusters compute it
  S->n_tris = n; S->n_objects = (uint32_t)n_obj; S->n_lights = (uint32_t)n_light;
  S->has_tree = tree;
  S->tree_deferred = defer;
  S->camera = hs->camera;
  // shadow-ray grids: resident scenes, and single frames with enough samples per triangle to repay
  // them (C5: 1.9 ms of build per light per 10 M triangles against 80 ps saved per shadow ray, i.e.
  // from about 3 samples per triangle; 7.5 there: one-shot render() 43.7 -> 38 ms).  They need the
  // positions only, so they are built here, while the normals of a big scene are still crossing PCIe.
  const bool one_frame_pays = (unsigned long long)hs->camera.width * (unsigned long long)hs->camera.height >= 4ull * n;
  if (tree && !(flags & RTB_BUILD_NO_LIGHT_GRIDS) && (!(flags & RTB_BUILD_ONE_FRAME) || one_frame_pays)) {
    float bmin[3], bmax[3];
    for (int k = 0; k < 3; ++k) { bmin[k] = ord2f(h_bounds[k]); bmax[k] = ord2f(h_bounds[3 + k]); }
    const int rc = build_light_grids(*S, hs->lights, n_light, bmin, bmax, st);
    if (rc) return rc;
  }
  if (trace_build) cudaEventRecord(e_t[1], st);
  if (overlap) {
    RTB_CUDA(cudaStreamWaitEvent(st, side.normals, 0));
    if (trace_build) cudaEventRecord(e_t[2], st);
    k_permute<false, true><<<gN, BT, 0, st>>>(raw_v, raw_n, S->order, n, d_first, (uint32_t)n_obj, S->v0, S->e1, S->e2,
                                              S->n0, S->n1, S->n2);
    ++launches;
  }
  RTB_CUDA(cudaEventRecord(e_b1, st));
  RTB_CUDA(cudaStreamSynchronize(st));
  RTB_CUDA(cudaGetLastError());
  S->upload_ms = ms_between(e_up0, e_up1);
  S->build_ms = ms_between(e_up1, e_b1);
  if (trace_build) {
    fprintf(stderr, "rtb build trace: upload %.2f | tree %.2f | light grids %.2f | normals arrived +%.2f | permute normals %.2f ms\n",
            S->upload_ms, ms_between(e_up1, e_t[0]), ms_between(e_t[0], e_t[1]), overlap ? ms_between(e_t[1], e_t[2]) : 0.0f,
            overlap ? ms_between(e_t[2], e_b1) : ms_between(e_t[1], e_b1));
    for (auto &e : e_t) cudaEventDestroy(e);
  }
  S->build_launches = launches;
  S->n_shadow_lights = 0;
  for (size_t i = 0; i < n_light; ++i)
    if (hs->lights[i].type == DIRECTIONAL || hs->lights[i].type == POINT) {
      if (S->n_shadow_lights < 32) S->shadow_light[S->n_shadow_lights] = (unsigned char)i;
      ++S->n_shadow_lights;
    }
  return RTB_OK;
}

}  // namespace rtb
