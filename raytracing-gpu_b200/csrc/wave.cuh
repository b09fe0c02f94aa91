// Wavefront frame kernel.  Included by trace.cu inside its anonymous namespace (it uses
// the traversal / shading primitives defined there).
//
// A persistent warp owns one 8x4-pixel tile = S = 32*a*a samples at a time and runs the
// reference's per-sample loop (gpu/raytracer.cu:107-122) as a sequence of warp-wide stages
// over small queues in shared memory:
//   trace stage   nearest hit of every live sample's current ray        -> slot[s]
//   shadow stage  one any-hit ray per (hit sample, directional/point light) -> occ[s] bits
//   shade stage   hit record, apply_light with the occlusion bits, colour accumulation,
//                 reflection ray -> next depth's queue (ballot compaction)
// Inside the two tracing stages the lanes are persistent threads: a lane whose ray has
// finished pulls the next ray of the stage's queue (ballot + popc, no atomics) whenever
// the number of lanes still traversing drops below a threshold, so the warp stays full
// until the queue runs dry instead of idling behind its slowest ray (profiles/r1_v2:
// 16-22 of 32 lanes were active in the node loop of the per-sample kernel).
// Per-sample arithmetic, its order, and therefore every result bit are those of
// shade_sample(); only the scheduling differs.
#pragma once

constexpr int WAVE_THREADS = 128;
constexpr int WAVE_WARPS = WAVE_THREADS / 32;
#ifndef RTB_WAVE_REFILL
#define RTB_WAVE_REFILL 0   /* measured on C5: 0 (static batches) 49.5 ms, 12: 55.4 ms, 24: 64.2 ms */
#endif
constexpr int WAVE_REFILL_BELOW = RTB_WAVE_REFILL;   // pull new rays when fewer lanes than this are traversing

// sample index inside the tile <-> sample coordinates (8x4 blocks stay contiguous)
__device__ __forceinline__ int wave_sx(int s, int a) { return 8 * ((s >> 5) % a) + (s & 7); }
__device__ __forceinline__ int wave_sy(int s, int a) { return 4 * ((s >> 5) / a) + ((s >> 3) & 3); }
__device__ __forceinline__ int wave_index(int sx, int sy, int a) {
  return (((sy >> 2) * a + (sx >> 3)) << 5) + ((sy & 3) << 3) + (sx & 7);
}

struct WaveTile {
  int *slot;            // [S] winning slot of the sample's current ray, -1 = miss
  uint32_t *occ;        // [S] bit j set: shadow light j is occluded
  uint32_t *col;        // [S] accumulated colour
  uint16_t *queue;      // [S] live samples of the current depth
  uint16_t *hitq;       // [S] samples whose current ray hit something
  float *scratch;       // global, 7 x S: origin, direction, nr_ of reflected rays (depth >= 1)
  int S, a, px0, py0;   // px0,py0: sample coordinates of the tile's corner
};

__device__ __forceinline__ void wave_ray(const DevCamera &cam, const WaveTile &t, int s, int depth, V3 &o, V3 &d) {
  if (depth == 0) {
    primary_ray(cam, t.px0 + wave_sx(s, t.a), t.py0 + wave_sy(s, t.a), o, d);
  } else {
    const float *p = t.scratch + s;
    o = mk(p[0], p[t.S], p[2 * t.S]);
    d = mk(p[3 * t.S], p[4 * t.S], p[5 * t.S]);
  }
}

// Traces `n` rays of a stage with persistent lanes.  fetch(k, o, d) produces ray k,
// done(k, hit, best) consumes its result.  Warp-synchronous: all 32 lanes call it.
template <bool ANY, bool COUNT, class Fetch, class Done>
__device__ __forceinline__ void wave_trace(const DevScene &sc, int n, Counters &c, Fetch fetch, Done done) {
  const int lane = threadIdx.x & 31;
  const uint32_t lt = (1u << lane) - 1u;
  int next = 0;                 // warp-uniform: first ray nobody has taken yet
  bool active = false;
  int my_k = 0;
  V3 o = mk(0, 0, 0), d = mk(0, 0, 1);
  Slabs sl = make_slabs(o, d, 0.0f);
  float eps = 0, inv_len = 1, tmax = INF;
  Best best;
  best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
  int stack[STACK_MAX + 1];
  int sp = 1, cur = TRAVERSAL_DONE;
  for (;;) {
    // ---- refill idle lanes ----
    const uint32_t idle = __ballot_sync(0xffffffffu, !active);
    if (idle && next < n) {
      const int k = next + __popc(idle & lt);
      if (!active && k < n) {
        fetch(k, o, d);
        my_k = k;
        const float amax = fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)) + sc.abs_max;
        eps = sc.margin_k * 1.1920929e-7f * amax;
        sl = make_slabs(o, d, eps);
        inv_len = 1.0001f / sqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
        tmax = INF;
        best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
        stack[0] = TRAVERSAL_DONE;
        sp = 1;
        cur = sc.root;
        active = true;
      }
      next += __popc(idle);
    }
    if (!__any_sync(0xffffffffu, active)) break;
    // ---- traverse until the ray ends or too few lanes are left traversing ----
    const bool more = next < n;
    while (active) {
      while (cur >= 0) {
        if (COUNT) ++c.nodes;
        const float4 *np = sc.nodes + 4ll * cur;
        const float4 q0 = __ldg(np), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
        float tn0, tn1;
        const bool h0 = slab_hit(sl, q0.x, q0.y, q0.z, q0.w, q2.x, q2.y, tmax, tn0);
        const bool h1 = slab_hit(sl, q1.x, q1.y, q1.z, q1.w, q2.z, q2.w, tmax, tn1);
        const int r0 = __float_as_int(q3.x), r1 = __float_as_int(q3.y);
        if (h0 && h1) {
          const bool far0 = tn1 < tn0;
          cur = far0 ? r1 : r0;
          stack[sp++] = far0 ? r0 : r1;
        } else if (h0) cur = r0;
        else if (h1) cur = r1;
        else cur = stack[--sp];
      }
      bool finished = cur == TRAVERSAL_DONE;
      if (!finished) {
        const uint32_t packed = (uint32_t)~cur;
        const int first = (int)(packed >> 3), cnt = (int)(packed & 7u);
        for (int k = 0; k < cnt; ++k) {
          if (COUNT) ++c.tris;
          if (consider<ANY>(sc, o, d, first + k, best)) {
            if (ANY) { finished = true; best.slot = first + k; break; }
            tmax = (best.dist * 1.0001f + eps) * inv_len;
          }
        }
        cur = stack[--sp];
      }
      if (finished) {
        done(my_k, best.slot >= 0, best);
        active = false;
      } else if (more && __popc(__activemask()) < WAVE_REFILL_BELOW) {
        break;   // keep the traversal state, let the idle lanes fetch
      }
    }
  }
}

// apply_light (gpu/light.cu:50-127) with the shadow-ray outcomes taken from `occ`
__device__ __forceinline__ uint32_t wave_apply_light(const DevScene &sc, int obj, const V3 P, const V3 Nrm, uint32_t occ) {
  const float4 m0 = __ldg(sc.materials + 2 * obj);
  const uint32_t ka = __float_as_uint(m0.x), kd = __float_as_uint(m0.y), ks = __float_as_uint(m0.z);
  const float ns = m0.w;
  uint32_t color = init_color(0, 0, 0);
  int j = 0;   // running index among the shadow-casting lights
  for (uint32_t i = 0; i < sc.n_lights; ++i) {
    const float4 l0 = __ldg(sc.lights + 2 * i), l1 = __ldg(sc.lights + 2 * i + 1);
    const int type = __float_as_int(l0.x);
    const uint32_t lcol = __float_as_uint(l0.y);
    const V3 lv = mk(l1);
    if (type == AMBIENT) {
      color = color_add(color, color_mults(lcol, ka));
    } else if (type == DIRECTIONAL) {
      const bool blocked = (occ >> j) & 1u;
      ++j;
      if (!blocked) {
        const V3 L = vscale(lv, -1);
        uint32_t t = color_mults(lcol, kd);
        t = color_mul(t, vdot(L, Nrm));
        t = apply_specular(t, vadd(P, vscale(lv, -10)), lv, P, Nrm, ks, ns);
        color = color_add(color, t);
      }
    } else if (type == POINT) {
      const bool blocked = (occ >> j) & 1u;
      ++j;
      if (!blocked) {
        const V3 L = vscale(lv, -1);
        V3 N = Nrm;
        if (vdot(L, N) < 0) N = vscale(N, -1);
        const V3 dir = vsub(lv, P);
        const float dist = vlen(vsub(lv, P));
        uint32_t t = color_mults(lcol, kd);
        t = color_mul(t, vdot(L, N) * 1 / dist);
        t = apply_specular(t, vadd(P, vscale(dir, -10)), dir, P, Nrm, ks, ns);
        color = color_add(color, t);
      }
    }
  }
  return color;
}

// shadow ray of light `li` from P (gpu/light.cu:68-75 directional, :96-104 point)
__device__ __forceinline__ V3 wave_shadow_dir(const DevScene &sc, int li, const V3 P) {
  const float4 l0 = __ldg(sc.lights + 2 * li), l1 = __ldg(sc.lights + 2 * li + 1);
  const V3 lv = mk(l1);
  return __float_as_int(l0.x) == DIRECTIONAL ? vscale(lv, -1) : vsub(lv, P);
}

template <bool COUNT>
__global__ void __launch_bounds__(WAVE_THREADS, RTB_FRAME_MINBLOCKS) k_frame_wave(const DevScene sc, const FrameArgs fa) {
  extern __shared__ uint32_t s_wave[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt = (1u << lane) - 1u;
  const int a = fa.aliasing, S = 32 * a * a;
  WaveTile t;
  {
    uint32_t *base = s_wave + (size_t)warp * 4 * S;       // 16 bytes per sample
    t.slot = reinterpret_cast<int *>(base);
    t.occ = base + S;
    t.col = base + 2 * S;
    t.queue = reinterpret_cast<uint16_t *>(base + 3 * S);
    t.hitq = t.queue + S;
    t.scratch = fa.bounce_scratch + ((size_t)blockIdx.x * WAVE_WARPS + warp) * 7 * S;
    t.S = S;
    t.a = a;
  }
  const int nl = fa.n_shadow_lights;
  const unsigned total = frame_items(fa);
  Counters c = {0, 0, 0, 0, 0};
  for (;;) {
    unsigned tile = 0;
    if (lane == 0) tile = atomicAdd(fa.tile_counter, 1u);
    tile = __shfl_sync(0xffffffffu, tile, 0);
    if (tile >= total) break;
    int lx0, ly0;
    if (!frame_item(fa, tile, lx0, ly0)) continue;
    t.px0 = a * lx0;
    t.py0 = a * ly0;

    // depth 0: every sample inside the image is live, colour (0,0,0), nr_ = 1
    int n_live = 0;
    for (int s = lane; s < S; s += 32) {
      t.col[s] = init_color(0, 0, 0);
      const bool in = t.px0 + wave_sx(s, a) < fa.cam.W && t.py0 + wave_sy(s, a) < fa.cam.H;
      const uint32_t m = __ballot_sync(0xffffffffu, in);
      if (in) t.queue[n_live + __popc(m & lt)] = (uint16_t)s;
      n_live += __popc(m);
    }
    __syncwarp();

    for (int depth = 0; depth <= 10 && n_live > 0; ++depth) {
      // ---- trace stage ----
      wave_trace<false, COUNT>(
          sc, n_live, c,
          [&](int k, V3 &o, V3 &d) {
            if (COUNT) { if (depth == 0) ++c.primary; else ++c.bounce; }
            wave_ray(fa.cam, t, t.queue[k], depth, o, d);
          },
          [&](int k, bool hit, const Best &b) { t.slot[t.queue[k]] = hit ? b.slot : -1; });
      __syncwarp();
      // ---- hit queue ----
      int n_hit = 0;
      for (int i = lane; i < ((n_live + 31) & ~31); i += 32) {
        const int s = i < n_live ? t.queue[i] : 0;
        const bool hit = i < n_live && t.slot[s] >= 0;
        const uint32_t m = __ballot_sync(0xffffffffu, hit);
        if (hit) { t.hitq[n_hit + __popc(m & lt)] = (uint16_t)s; t.occ[s] = 0u; }
        n_hit += __popc(m);
      }
      __syncwarp();
      // ---- shadow stage: ray k = (light k / n_hit, hit sample k % n_hit) ----
      if (nl > 0 && n_hit > 0) {
        wave_trace<true, COUNT>(
            sc, nl * n_hit, c,
            [&](int k, V3 &o, V3 &d) {
              if (COUNT) ++c.shadow;
              const int s = t.hitq[k % n_hit];
              V3 ro, rd, N;
              wave_ray(fa.cam, t, s, depth, ro, rd);
              hit_record(sc, ro, rd, t.slot[s], o, N);
              d = wave_shadow_dir(sc, fa.shadow_light[k / n_hit], o);
            },
            [&](int k, bool hit, const Best &) {
              if (hit) atomicOr(&t.occ[t.hitq[k % n_hit]], 1u << (k / n_hit));
            });
        __syncwarp();
      }
      // ---- shade stage (trace(), gpu/raytracer.cu:32-47, and the loop tail :118-122) ----
      int n_next = 0;
      for (int i = lane; i < ((n_hit + 31) & ~31); i += 32) {
        bool live = false;
        int s = 0;
        if (i < n_hit) {
          s = t.hitq[i];
          V3 o, d, P, N;
          wave_ray(fa.cam, t, s, depth, o, d);
          const int slot = t.slot[s];
          hit_record(sc, o, d, slot, P, N);
          const int obj = __float_as_int(__ldg(sc.e1 + slot).w);
          uint32_t tc = wave_apply_light(sc, obj, P, N, t.occ[s]);
          const float r = __ldg(sc.materials + 2 * obj + 1).x;
          float nr_ = depth == 0 ? 1.0f : t.scratch[6 * S + s];
          tc = color_mul(tc, nr_);
          t.col[s] = color_add(t.col[s], tc);
          nr_ *= r;
          live = nr_ > 0.01f && depth < 10;
          if (live) {
            // obj.nr > 0 here (nr_ > 0.01), so the ray was replaced by its reflection (ray.cu:16-25)
            const V3 rd = vsub(d, vscale(N, 2 * vdot(N, d)));
            float *p = t.scratch + s;
            p[0] = P.x; p[S] = P.y; p[2 * S] = P.z;
            p[3 * S] = rd.x; p[4 * S] = rd.y; p[5 * S] = rd.z;
            p[6 * S] = nr_;
          }
        }
        __syncwarp();   // all reads of hitq/queue entries of this round precede the compacted writes
        const uint32_t m = __ballot_sync(0xffffffffu, live);
        if (live) t.queue[n_next + __popc(m & lt)] = (uint16_t)s;
        n_next += __popc(m);
      }
      __syncwarp();
      n_live = n_next;
    }

    // ---- SSAA resolve (downscale kernel body, gpu/raytracer.cu:49-85) ----
    const int lx = lx0 + (lane & 7), ly = ly0 + (lane >> 3);
    if (lx < fa.out_w && ly < fa.out_h) {
      float r = 0, g = 0, b = 0;
      for (int j = 0; j < a; ++j)
        for (int i = 0; i < a; ++i) {
          const uint32_t col = t.col[wave_index(a * (lane & 7) + i, a * (lane >> 3) + j, a)];
          r += (float)ch_r(col);
          g += (float)ch_g(col);
          b += (float)ch_b(col);
        }
      const float ali2 = 255.0f * (float)a * (float)a;
      r /= ali2;
      g /= ali2;
      b /= ali2;
      const uint32_t e = init_color(r, g, b);
      unsigned char *dst = fa.out + (long long)(fa.out_h - ly - 1) * fa.pitch + 4ll * (fa.out_w - lx - 1);
      *reinterpret_cast<uint32_t *>(dst) = e;
    }
    __syncwarp();
  }
  if (COUNT) {
    unsigned v[5] = {c.primary, c.shadow, c.bounce, c.nodes, c.tris};
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      unsigned long long s = v[k];
#pragma unroll
      for (int dlt = 16; dlt; dlt >>= 1) s += __shfl_xor_sync(0xffffffffu, s, dlt);
      if (lane == 0 && s) atomicAdd(fa.counters + k, s);
    }
  }
}
