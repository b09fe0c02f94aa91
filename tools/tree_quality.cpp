// CPU study of hierarchy quality on the benchmark terrain: node visits / triangle tests per
// ray for (A) the Morton radix tree the device builds (build.cu), (B) a binned-SAH top-down
// tree, (C) PLOC (parallel locally-ordered clustering) over the same Morton order.
// Counts only; the arithmetic here is plain float and does not have to match the tracer's.
//   g++ -O2 -fopenmp -std=c++17 -I include tools/tree_quality.cpp -L raytracing-gpu_b200 -lrt_host \
//       -Wl,-rpath,$PWD/raytracing-gpu_b200 -o /tmp/tree_quality && /tmp/tree_quality 2237
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>
#include "rt_host.h"

struct V3 { float x, y, z; };
static inline V3 mk(float x, float y, float z) { return {x, y, z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline float len(V3 a) { return std::sqrt(dot(a, a)); }
static inline V3 vmin(V3 a, V3 b) { return {std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)}; }
static inline V3 vmax(V3 a, V3 b) { return {std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)}; }

struct Box { V3 lo, hi; };
static inline Box empty_box() { return {{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}}; }
static inline Box merge(const Box &a, const Box &b) { return {vmin(a.lo, b.lo), vmax(a.hi, b.hi)}; }
static inline float area(const Box &b) {
  const V3 d = b.hi - b.lo;
  return d.x * d.y + d.y * d.z + d.z * d.x;
}

struct Node { Box b[2]; int ref[2]; };   // ref >= 0 inner, < 0 leaf ~((first << 3) | count)
struct Tree {
  std::vector<Node> nodes;
  std::vector<uint32_t> order;   // slot -> triangle
  int root = 0;
};
static inline int leaf_ref(uint32_t first, uint32_t cnt) { return ~(int)((first << 3) | cnt); }

static int LEAF_MAX = 3;
static float g_yscale = 1.0f;
static float g_margin_k = 24.0f;
static uint32_t g_cluster = 0;   // > 0: radix ranges of at most this many triangles are built with binned SAH
static int g_axes = 3;           // 3: bin all axes, 1: longest centroid axis only
static int sah_build(struct Tree &t, uint32_t l, uint32_t r, struct Box &out, int depth, int &maxdepth);
static uint32_t g_small = 0;     // segments of at most this many triangles are halved by position (no binning)
static int g_pad = 1;            // needle padding of build.cu (sliver_pad, mode 5)
static int g_mode = 0;          // 0 cubic xyz interleave, 1/2/3: axis x/y/z last (2D Morton of the other two first)
static std::vector<V3> g_v;       // 3 per triangle
static std::vector<Box> g_tb;     // triangle boxes
static uint32_t g_n;


// build.cu sliver_pad (mode 5): third-edge slop + corner spikes
static float corner_kappa(float cosine) {
  const float c = std::min(cosine, 1.0f);
  return std::min(1.0f / std::sqrt(std::max(0.5f * (1.0f - c), 1e-12f)), 1.0e4f);
}
static void grow(Box &b, V3 p) { b.lo = vmin(b.lo, p); b.hi = vmax(b.hi, p); }
static void sliver_pad(uint32_t i, float pad_unit, Box &b) {
  const V3 v[3] = {g_v[3 * i], g_v[3 * i + 1], g_v[3 * i + 2]};
  const V3 ab = v[1] - v[0], bc = v[2] - v[1], ca = v[0] - v[2];
  const float lab = len(ab), lbc = len(bc), lca = len(ca);
  if (!(lab > 0) || !(lbc > 0) || !(lca > 0)) return;
  const float longest = std::max(std::max(lab, lbc), lca);
  const V3 d_ab = ab * (1.0f / lab), d_bc = bc * (1.0f / lbc), d_ca = ca * (1.0f / lca);
  const V3 outs[3][2] = {{d_ab, d_ca * -1.0f}, {d_bc, d_ab * -1.0f}, {d_ca, d_bc * -1.0f}};
  const float ratio = (lab + lca) / lbc;
  if (ratio > 4.0f) {
    const float pad = std::min(pad_unit * ratio, longest);
    b.lo = b.lo - mk(pad, pad, pad); b.hi = b.hi + mk(pad, pad, pad);
  }
  for (int c = 0; c < 3; ++c) {
    const float kap = corner_kappa(dot(outs[c][0], outs[c][1]));
    if (!(kap > 4.0f)) continue;
    const float m = std::min(pad_unit * kap, longest);
    const V3 sa = outs[c][0] * m, sb = outs[c][1] * m;
    grow(b, v[c] - sa); grow(b, v[c] - sb); grow(b, v[c] - sa - sb);
  }
}

// ---------- (A) Morton radix tree, as build.cu ----------
static uint64_t spread21(uint32_t v) {
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 32) & 0x1f00000000ffffull;
  x = (x | x << 16) & 0x1f0000ff0000ffull;
  x = (x | x << 8) & 0x100f00f00f00f00full;
  x = (x | x << 4) & 0x10c30c30c30c30c3ull;
  x = (x | x << 2) & 0x1249249249249249ull;
  return x;
}

static uint64_t spread21_2(uint32_t v) {   // 21 bits -> every second bit
  uint64_t x = v & 0x1fffffu;
  x = (x | x << 16) & 0x0000ffff0000ffffull;
  x = (x | x << 8) & 0x00ff00ff00ff00ffull;
  x = (x | x << 4) & 0x0f0f0f0f0f0f0f0full;
  x = (x | x << 2) & 0x3333333333333333ull;
  x = (x | x << 1) & 0x5555555555555555ull;
  return x;
}
static uint64_t make_key(uint32_t qx, uint32_t qy, uint32_t qz) {
  if (g_mode == 0) return spread21(qx) | (spread21((uint32_t)(qy * g_yscale)) << 1) | (spread21(qz) << 2);
  const uint32_t q[3] = {qx, qy, qz};
  const int a = g_mode - 1, p = (a + 1) % 3, r = (a + 2) % 3;
  const int lo = p < r ? p : r, hi = p < r ? r : p;
  return ((spread21_2(q[lo]) | (spread21_2(q[hi]) << 1)) << 21) | q[a];
}

static void morton_sort(std::vector<uint64_t> &keys, std::vector<uint32_t> &order, uint32_t sstride = 1) {
  Box sb = empty_box();
  for (uint32_t i = 0; i < g_n; ++i) sb = merge(sb, g_tb[i]);
  const V3 e = sb.hi - sb.lo;
  const float ext = std::max(e.x, std::max(e.y, e.z));
  const float S = 2097152.0f;
  std::vector<uint64_t> k(g_n);
  for (uint32_t i = 0; i < g_n; ++i) {
    const Box &b = g_tb[i];
    const float cx = 0.5f * b.lo.x + 0.5f * b.hi.x, cy = 0.5f * b.lo.y + 0.5f * b.hi.y, cz = 0.5f * b.lo.z + 0.5f * b.hi.z;
    auto q = [&](float c, float lo) { return (uint32_t)std::min(std::max((c - lo) / ext * S, 0.0f), S - 1.0f); };
    k[i] = make_key(q(cx, sb.lo.x), q(cy, sb.lo.y), q(cz, sb.lo.z));
  }
  order.clear();
  for (uint32_t i = 0; i < g_n; i += sstride) order.push_back(i);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return k[a] < k[b]; });
  keys.resize(order.size());
  for (uint32_t s = 0; s < order.size(); ++s) keys[s] = k[order[s]];
}

static int delta(const std::vector<uint64_t> &keys, uint32_t j) {
  const uint64_t x = keys[j] ^ keys[j + 1];
  if (x) return 128 - __builtin_clzll(x);
  return 32 - __builtin_clz(j ^ (j + 1));
}

// recursive top-down restatement of the radix tree (same topology as the bottom-up emit)
static int radix_build(Tree &t, const std::vector<uint64_t> &keys, uint32_t l, uint32_t r, Box &out) {
  // returns ref for range [l, r]
  if (r - l + 1 <= (uint32_t)LEAF_MAX) {
    out = empty_box();
    for (uint32_t s = l; s <= r; ++s) out = merge(out, g_tb[t.order[s]]);
    return leaf_ref(l, r - l + 1);
  }
  if (g_cluster && r - l + 1 <= g_cluster) { int md = 0; return sah_build(t, l, r, out, 0, md); }
  // split where delta is maximal within [l, r-1]
  int best = -1; uint32_t bj = l;
  // the radix tree splits at the position of the highest differing bit between keys[l], keys[r]
  const uint64_t x = keys[l] ^ keys[r];
  if (x) {
    const int bit = 63 - __builtin_clzll(x);
    // first index whose key has that bit set
    uint32_t lo = l, hi = r;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) / 2;
      if ((keys[mid] >> bit) & 1) hi = mid; else lo = mid;
    }
    bj = lo;
  } else {
    for (uint32_t j = l; j < r; ++j) { const int d = delta(keys, j); if (d > best) { best = d; bj = j; } }
  }
  const int me = (int)t.nodes.size();
  t.nodes.push_back(Node());
  Box b0, b1;
  const int r0 = radix_build(t, keys, l, bj, b0);
  const int r1 = radix_build(t, keys, bj + 1, r, b1);
  t.nodes[me].b[0] = b0; t.nodes[me].b[1] = b1; t.nodes[me].ref[0] = r0; t.nodes[me].ref[1] = r1;
  out = merge(b0, b1);
  return me;
}

// ---------- (B) binned SAH ----------
static int sah_build(Tree &t, uint32_t l, uint32_t r, Box &out, int depth, int &maxdepth) {
  maxdepth = std::max(maxdepth, depth);
  const uint32_t cnt = r - l + 1;
  Box nb = empty_box(), cb = empty_box();
  for (uint32_t s = l; s <= r; ++s) {
    const Box &b = g_tb[t.order[s]];
    nb = merge(nb, b);
    const V3 c = (b.lo + b.hi) * 0.5f;
    cb.lo = vmin(cb.lo, c); cb.hi = vmax(cb.hi, c);
  }
  out = nb;
  if (cnt <= (uint32_t)LEAF_MAX) return leaf_ref(l, cnt);
  const int NB = getenv("NB") ? atoi(getenv("NB")) : 16;
  float best_cost = INFINITY; int best_axis = -1, best_bin = 0;
  const bool skip_bins = cnt <= g_small;
  int longest = 0;
  { const V3 e = cb.hi - cb.lo; longest = e.x >= e.y ? (e.x >= e.z ? 0 : 2) : (e.y >= e.z ? 1 : 2); }
  for (int ax = 0; ax < 3; ++ax) {
    if (skip_bins) break;
    if (g_axes == 1 && ax != longest) continue;
    const float lo = (&cb.lo.x)[ax], hi = (&cb.hi.x)[ax];
    if (!(hi > lo)) continue;
    Box bb[64]; uint32_t bc[64];
    for (int k = 0; k < NB; ++k) { bb[k] = empty_box(); bc[k] = 0; }
    const float sc = NB / (hi - lo);
    for (uint32_t s = l; s <= r; ++s) {
      const Box &b = g_tb[t.order[s]];
      const float c = 0.5f * ((&b.lo.x)[ax] + (&b.hi.x)[ax]);
      int k = (int)((c - lo) * sc); if (k >= NB) k = NB - 1; if (k < 0) k = 0;
      bb[k] = merge(bb[k], b); ++bc[k];
    }
    float ra[64]; uint32_t rc[64];
    Box acc = empty_box(); uint32_t n = 0;
    for (int k = NB - 1; k > 0; --k) { acc = merge(acc, bb[k]); n += bc[k]; ra[k] = n ? area(acc) : 0; rc[k] = n; }
    acc = empty_box(); n = 0;
    for (int k = 0; k < NB - 1; ++k) {
      acc = merge(acc, bb[k]); n += bc[k];
      if (!n || !rc[k + 1]) continue;
      const float cost = area(acc) * n + ra[k + 1] * rc[k + 1];
      if (cost < best_cost) { best_cost = cost; best_axis = ax; best_bin = k; }
    }
  }
  uint32_t mid;
  if (best_axis < 0) {
    mid = l + cnt / 2 - 1;   // all centroids equal: median
  } else {
    const float lo = (&cb.lo.x)[best_axis], hi = (&cb.hi.x)[best_axis];
    const float sc = NB / (hi - lo);
    auto it = std::partition(t.order.begin() + l, t.order.begin() + r + 1, [&](uint32_t i) {
      const Box &b = g_tb[i];
      const float c = 0.5f * ((&b.lo.x)[best_axis] + (&b.hi.x)[best_axis]);
      int k = (int)((c - lo) * sc); if (k >= NB) k = NB - 1; if (k < 0) k = 0;
      return k <= best_bin;
    });
    mid = (uint32_t)(it - t.order.begin()) - 1;
    if (mid < l || mid >= r) mid = l + cnt / 2 - 1;
  }
  const int me = (int)t.nodes.size();
  t.nodes.push_back(Node());
  Box b0, b1;
  const int r0 = sah_build(t, l, mid, b0, depth + 1, maxdepth);
  const int r1 = sah_build(t, mid + 1, r, b1, depth + 1, maxdepth);
  t.nodes[me].b[0] = b0; t.nodes[me].b[1] = b1; t.nodes[me].ref[0] = r0; t.nodes[me].ref[1] = r1;
  return me;
}

// ---------- (C) PLOC over the Morton order ----------
// clusters: box + subtree id.  Subtree: binary tree over triangles (temp), then laid out
// with leaves collapsed (<= LEAF_MAX) and triangles in DFS order.
struct TNode { int c0, c1; Box b; uint32_t cnt; };   // c < 0: triangle ~c
static int ploc_layout(Tree &t, const std::vector<TNode> &tn, int id, std::vector<uint32_t> &new_order, Box &out, int depth, int &maxdepth) {
  maxdepth = std::max(maxdepth, depth);
  if (id < 0) { out = g_tb[~id]; const uint32_t f = new_order.size(); new_order.push_back(~id); return leaf_ref(f, 1); }
  const TNode &n = tn[id];
  out = n.b;
  if (n.cnt <= (uint32_t)LEAF_MAX) {
    const uint32_t f = new_order.size();
    // gather triangles
    std::vector<int> st{id};
    while (!st.empty()) { int x = st.back(); st.pop_back(); if (x < 0) new_order.push_back(~x); else { st.push_back(tn[x].c1); st.push_back(tn[x].c0); } }
    return leaf_ref(f, n.cnt);
  }
  const int me = (int)t.nodes.size();
  t.nodes.push_back(Node());
  Box b0, b1;
  const int r0 = ploc_layout(t, tn, n.c0, new_order, b0, depth + 1, maxdepth);
  const int r1 = ploc_layout(t, tn, n.c1, new_order, b1, depth + 1, maxdepth);
  t.nodes[me].b[0] = b0; t.nodes[me].b[1] = b1; t.nodes[me].ref[0] = r0; t.nodes[me].ref[1] = r1;
  return me;
}

static void ploc_build(Tree &t, const std::vector<uint32_t> &morton_order, int radius, int &iters, int &maxdepth) {
  struct Cl { Box b; int id; uint32_t cnt; };
  std::vector<Cl> cur(g_n), nxt;
  for (uint32_t s = 0; s < g_n; ++s) cur[s] = {g_tb[morton_order[s]], ~(int)morton_order[s], 1};
  std::vector<TNode> tn;
  tn.reserve(g_n);
  std::vector<int> nn;
  iters = 0;
  while (cur.size() > 1) {
    const int m = (int)cur.size();
    nn.assign(m, -1);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < m; ++i) {
      float best = INFINITY; int bj = -1;
      const int a = std::max(0, i - radius), b = std::min(m - 1, i + radius);
      for (int j = a; j <= b; ++j) {
        if (j == i) continue;
        const float ar = area(merge(cur[i].b, cur[j].b));
        if (ar < best) { best = ar; bj = j; }
      }
      nn[i] = bj;
    }
    nxt.clear();
    for (int i = 0; i < m; ++i) {
      const int j = nn[i];
      if (nn[j] == i) {
        if (i < j) {
          tn.push_back({cur[i].id, cur[j].id, merge(cur[i].b, cur[j].b), cur[i].cnt + cur[j].cnt});
          nxt.push_back({tn.back().b, (int)tn.size() - 1, tn.back().cnt});
        }
      } else nxt.push_back(cur[i]);
    }
    cur.swap(nxt);
    ++iters;
  }
  std::vector<uint32_t> new_order;
  new_order.reserve(g_n);
  Box rb;
  maxdepth = 0;
  t.root = ploc_layout(t, tn, cur[0].id, new_order, rb, 0, maxdepth);
  t.order.swap(new_order);
}

// ---------- traversal counters ----------
struct Cnt { double nodes = 0, tris = 0, rays = 0, hits = 0; };

static inline bool tri_hit(V3 o, V3 d, uint32_t tri, float &dist) {
  const float EPS = 1e-7f;
  const V3 v0 = g_v[3 * tri], e1 = g_v[3 * tri + 1] - v0, e2 = g_v[3 * tri + 2] - v0;
  const V3 h = cross(d, e2);
  const float a = dot(e1, h);
  if (a > -EPS && a < EPS) return false;
  const float f = 1 / a;
  const V3 s = o - v0;
  const float u = f * dot(s, h);
  if (u < 0 || u > 1) return false;
  const V3 q = cross(s, e1);
  const float v = f * dot(d, q);
  if (v < 0 || u + v > 1) return false;
  const float t = f * dot(e2, q);
  if (!(t > EPS)) return false;
  dist = t * len(d);
  return dist > 0.01f;
}

static inline bool slab(const Box &b, V3 o, V3 inv, float eps, float tmax, float &tn) {
  const float x0 = (b.lo.x - eps - o.x) * inv.x, x1 = (b.hi.x + eps - o.x) * inv.x;
  const float y0 = (b.lo.y - eps - o.y) * inv.y, y1 = (b.hi.y + eps - o.y) * inv.y;
  const float z0 = (b.lo.z - eps - o.z) * inv.z, z1 = (b.hi.z + eps - o.z) * inv.z;
  tn = std::max(std::max(std::min(x0, x1), std::min(y0, y1)), std::max(std::min(z0, z1), 0.0f));
  const float tf = std::min(std::min(std::max(x0, x1), std::max(y0, y1)), std::min(std::max(z0, z1), tmax));
  return tn <= tf;
}

static float g_absmax;

template <bool ANY>
static bool traverse(const Tree &t, V3 o, V3 d, Cnt &c, float &best, uint32_t &best_tri) {
  auto rcp = [](float x) { return 1.0f / (std::fabs(x) > 1e-30f ? x : std::copysign(1e-30f, x)); };
  const V3 inv = {rcp(d.x), rcp(d.y), rcp(d.z)};
  const float amax = std::max(std::max(std::fabs(o.x), std::fabs(o.y)), std::fabs(o.z)) + g_absmax;
  const float eps = g_margin_k * 1.1920929e-7f * amax;
  const float inv_len = 1.0001f / len(d);
  float tmax = INFINITY;
  best = INFINITY; best_tri = ~0u;
  int stack[128]; int sp = 0;
  int cur = t.root;
  c.rays += 1;
  for (;;) {
    if (cur >= 0) {
      c.nodes += 1;
      const Node &n = t.nodes[cur];
      float t0, t1;
      const bool h0 = slab(n.b[0], o, inv, eps, ANY ? INFINITY : tmax, t0);
      const bool h1 = slab(n.b[1], o, inv, eps, ANY ? INFINITY : tmax, t1);
      if (h0 && h1) {
        const bool far0 = ANY ? false : t1 < t0;
        stack[sp++] = n.ref[far0 ? 0 : 1];
        cur = n.ref[far0 ? 1 : 0];
        continue;
      }
      if (h0) { cur = n.ref[0]; continue; }
      if (h1) { cur = n.ref[1]; continue; }
    } else {
      const uint32_t p = (uint32_t)~cur;
      const uint32_t first = p >> 3, cnt = p & 7;
      for (uint32_t k = 0; k < cnt; ++k) {
        c.tris += 1;
        float dist;
        const uint32_t tri = t.order[first + k];
        if (tri_hit(o, d, tri, dist)) {
          if (ANY) { c.hits += 1; return true; }
          if (dist < best) { best = dist; best_tri = tri; tmax = (best * 1.0001f + eps) * inv_len; }
        }
      }
    }
    if (!sp) break;
    cur = stack[--sp];
  }
  if (best_tri != ~0u) c.hits += 1;
  return best_tri != ~0u;
}

static double sah_cost(const Tree &t) {
  // sum of child box areas / root area (inner visits) ; leaves weighted by count
  Box rb = merge(t.nodes[t.root].b[0], t.nodes[t.root].b[1]);
  const double ra = area(rb);
  double ci = 0, cl = 0;
  for (const Node &n : t.nodes)
    for (int s = 0; s < 2; ++s) {
      const double a = area(n.b[s]) / ra;
      if (n.ref[s] >= 0) ci += a; else cl += a * ((uint32_t)~n.ref[s] & 7);
    }
  return 1.0 + ci + 1.5 * cl;   // node visits + 1.5 x triangle tests
}

static void evaluate(const char *name, const Tree &t, const struct scene &sc, int stride) {
  // camera (gpu/raytracer.cu:178-182) at 3x3 SSAA
  const int W = sc.camera.width * 3, H = sc.camera.height * 3;
  auto nrm = [](V3 a) { const float l = len(a); return mk(a.x / l, a.y / l, a.z / l); };
  const V3 u = nrm(mk(sc.camera.u.x, sc.camera.u.y, sc.camera.u.z)), v = nrm(mk(sc.camera.v.x, sc.camera.v.y, sc.camera.v.z));
  const V3 w = cross(u, v);
  const float L = (float)(W / (2 * std::tan(sc.camera.fov * M_PI / 360)));
  const V3 pos = mk(sc.camera.position.x, sc.camera.position.y, sc.camera.position.z);
  const V3 C = pos + w * L;
  Cnt prim, sdir, spt;
#pragma omp parallel
  {
    Cnt p, sd, sp_;
#pragma omp for schedule(dynamic, 4)
    for (int py = stride / 2; py < H; py += stride)
      for (int px = stride / 2; px < W; px += stride) {
        const V3 o = C + u * (float)(px - W / 2) + v * (float)(py - H / 2);
        const V3 d = nrm(pos - o);
        float best; uint32_t tri;
        if (!traverse<false>(t, o, d, p, best, tri)) continue;
        const V3 P = o + d * best;
        for (size_t i = 0; i < sc.light_count; ++i) {
          const struct light &l = sc.lights[i];
          const V3 lv = mk(l.v.x, l.v.y, l.v.z);
          float b2; uint32_t t2;
          if (l.type == DIRECTIONAL) traverse<true>(t, P, lv * -1.0f, sd, b2, t2);
          else if (l.type == POINT) traverse<true>(t, P, lv - P, sp_, b2, t2);
        }
      }
#pragma omp critical
    {
      prim.nodes += p.nodes; prim.tris += p.tris; prim.rays += p.rays; prim.hits += p.hits;
      sdir.nodes += sd.nodes; sdir.tris += sd.tris; sdir.rays += sd.rays; sdir.hits += sd.hits;
      spt.nodes += sp_.nodes; spt.tris += sp_.tris; spt.rays += sp_.rays; spt.hits += sp_.hits;
    }
  }
  const double rays = prim.rays + sdir.rays + spt.rays;
  printf("%-22s nodes %8zu sah %7.2f | primary %6.2f n %5.2f t (hit %.3f) | dir %6.2f n %5.2f t (occl %.3f) | point %6.2f n %5.2f t (occl %.3f) | all %6.2f n %5.2f t\n",
         name, t.nodes.size(), sah_cost(t), prim.nodes / prim.rays, prim.tris / prim.rays, prim.hits / prim.rays,
         sdir.nodes / sdir.rays, sdir.tris / sdir.rays, sdir.hits / sdir.rays, spt.nodes / spt.rays, spt.tris / spt.rays,
         spt.hits / spt.rays, (prim.nodes + sdir.nodes + spt.nodes) / rays, (prim.tris + sdir.tris + spt.tris) / rays);
  fflush(stdout);
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char **argv) {
  const char *what = argc > 1 ? argv[1] : "700";
  const int stride = argc > 2 ? atoi(argv[2]) : 24;
  const char *modes = argc > 3 ? argv[3] : "0123";
  const uint32_t sstride = argc > 4 ? (uint32_t)atoi(argv[4]) : 1;
  const bool full = !(argc > 5 && atoi(argv[5]) == 0);
  if (argc > 6) g_cluster = (uint32_t)atoi(argv[6]);
  if (argc > 7) g_axes = atoi(argv[7]);
  if (argc > 8) g_pad = atoi(argv[8]);
  if (getenv("SMALL")) g_small = (uint32_t)atoi(getenv("SMALL"));
  struct scene sc;
  if (strstr(what, ".svati")) {
    char err[256];
    if (rtb_parse_svati(what, &sc, err, sizeof err)) { fprintf(stderr, "%s\n", err); return 1; }
  } else if (rtb_synth_terrain(&sc, atoi(what), 0xB200, 3840, 2160)) return 1;
  if (argc > 10) { sc.camera.width = atoi(argv[9]); sc.camera.height = atoi(argv[10]); }
  g_n = (uint32_t)sc.triangle_count;
  g_v.resize(3ull * g_n);
  memcpy(g_v.data(), sc.objects_data.vertex, 36ull * g_n);
  g_tb.resize(g_n);
  g_absmax = 0;
  double nx = 0, ny = 0, nz = 0;
  for (uint32_t i = 0; i < g_n; ++i) {
    Box b = empty_box();
    for (int k = 0; k < 3; ++k) { b.lo = vmin(b.lo, g_v[3 * i + k]); b.hi = vmax(b.hi, g_v[3 * i + k]); }
    g_tb[i] = b;
    const V3 c = cross(g_v[3 * i + 1] - g_v[3 * i], g_v[3 * i + 2] - g_v[3 * i]);
    nx += std::fabs(c.x); ny += std::fabs(c.y); nz += std::fabs(c.z);
    g_absmax = std::max(g_absmax, std::max(std::max(std::fabs(b.lo.x), std::fabs(b.hi.x)), std::max(std::max(std::fabs(b.lo.y), std::fabs(b.hi.y)), std::max(std::fabs(b.lo.z), std::fabs(b.hi.z)))));
  }
  if (g_pad) {
    // pad_unit = SLIVER_K 2^-23 (|scene| + 2 * camera_origin_radius) at 3x3 SSAA (build.cu)
    const double W3 = sc.camera.width * 3.0, H3 = sc.camera.height * 3.0;
    const double L = W3 / (2 * std::tan(sc.camera.fov * M_PI / 360));
    const double pm = std::max((double)std::max(std::fabs(sc.camera.position.x), std::fabs(sc.camera.position.y)), (double)std::fabs(sc.camera.position.z));
    const float R = 2.0f * (float)(pm + std::fabs(L) + 0.5 * (W3 + H3));
    const float pad_unit = (getenv("PADSCALE") ? (float)atof(getenv("PADSCALE")) : 1.0f) * 3.0f * 1.1920929e-7f * (g_absmax + R);
    uint32_t padded = 0; double grow_sum = 0;
    for (uint32_t i = 0; i < g_n; ++i) {
      Box b = g_tb[i];
      sliver_pad(i, pad_unit, b);
      const V3 e0 = g_tb[i].hi - g_tb[i].lo, e1 = b.hi - b.lo;
      if (e1.x > e0.x || e1.y > e0.y || e1.z > e0.z) { ++padded; grow_sum += std::max(std::max(e1.x - e0.x, e1.y - e0.y), e1.z - e0.z); }
      g_tb[i] = b;
    }
    printf("needle padding: pad unit %.5f, %u triangles padded, mean growth %.4f (scene |max| %.1f)\n", pad_unit, padded, padded ? grow_sum / padded : 0.0, g_absmax);
  }
  const double nn = nx + ny + nz;
  printf("%s: triangles %u, area-weighted |normal| x %.3f y %.3f z %.3f, ray stride %d, sample stride %u\n", what, g_n, nx / nn, ny / nn, nz / nn, stride, sstride);
  for (const char *m = modes; *m; ++m) {
    if (*m < '0' || *m > '3') continue;
    g_mode = *m - '0';
    std::vector<uint64_t> keys; std::vector<uint32_t> morder;
    char nm[64];
    if (sstride > 1) {
      const uint32_t keep = g_n;
      morton_sort(keys, morder, sstride);
      Tree t; t.order = morder; Box rb;
      g_n = (uint32_t)morder.size();
      t.root = radix_build(t, keys, 0, g_n - 1, rb);
      g_n = keep;
      printf("mode %d sample (%zu tris): sah %.3f\n", g_mode, morder.size(), sah_cost(t));
    }
    if (!full) continue;
    morton_sort(keys, morder);
    Tree t; t.order = morder; t.nodes.reserve(g_n);
    Box rb;
    t.root = radix_build(t, keys, 0, g_n - 1, rb);
    snprintf(nm, sizeof nm, "radix tree mode %d", g_mode);
    evaluate(nm, t, sc, stride);
  }
  if (strchr(modes, 's')) {
    g_mode = 0; g_cluster = 0;
    std::vector<uint64_t> keys; std::vector<uint32_t> morder;
    morton_sort(keys, morder);
    Tree t; t.order = morder; t.nodes.reserve(g_n);
    Box rb; int md = 0;
    t.root = sah_build(t, 0, g_n - 1, rb, 0, md);
    evaluate("binned SAH", t, sc, stride);
  }
  return 0;
}
