// C ABI of librtb200.so (include/rtb200.h).  Host orchestration only; kernels live in
// build.cu / sort.cu / trace.cu.  There is NO CPU fallback: without a usable sm_100
// device every compute entry point returns RTB_ERR_NOGPU.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <mutex>
#include <string>
#include <thread>
#include <algorithm>
#include <unordered_map>
#include <vector>
#include "rtb_internal.h"
#include "scene_store.h"

struct rtb_scene_handle {
  rtb::SceneStore store;
};

namespace rtb {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
  const char *base = strrchr(file, '/');
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), base ? base + 1 : file, line, what);
  cudaGetLastError();   // clear the sticky flag of non-fatal errors
  return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? RTB_ERR_NOGPU : RTB_ERR_CUDA;
}

static int pick_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    set_error("no CUDA device available (%s); librtb200 has no CPU fallback",
              e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return RTB_ERR_NOGPU;
  }
  if (device < 0 || device >= n) { set_error("device %d out of range (0..%d)", device, n - 1); return RTB_ERR_ARG; }
  int major = 0, minor = 0;   // (cudaGetDeviceProperties costs ~2 ms per call; attributes are cheap)
  RTB_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  RTB_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) {
    set_error("device %d is sm_%d%d; this library is built for sm_100a only", device, major, minor);
    return RTB_ERR_NOGPU;
  }
  RTB_CUDA(cudaSetDevice(device));
  return RTB_OK;
}

// host part of render(), gpu/raytracer.cu:178-182
static DevCamera camera_frame(const struct camera &cam) {
  auto norm = [](vector3 a) {
    float r = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
    vector3 o; o.x = a.x / r; o.y = a.y / r; o.z = a.z / r;
    return o;
  };
  const vector3 u = norm(cam.u), v = norm(cam.v);
  vector3 w;
  w.x = u.y * v.z - u.z * v.y;
  w.y = u.z * v.x - u.x * v.z;
  w.z = u.x * v.y - u.y * v.x;
  const float L = (float)(cam.width / (2 * tan(cam.fov * M_PI / 360)));
  DevCamera d;
  d.u = make_float3(u.x, u.y, u.z);
  d.v = make_float3(v.x, v.y, v.z);
  d.pos = make_float3(cam.position.x, cam.position.y, cam.position.z);
  d.C = make_float3(cam.position.x + L * w.x, cam.position.y + L * w.y, cam.position.z + L * w.z);
  d.W = cam.width;
  d.H = cam.height;
  d.w = make_float3(w.x, w.y, w.z);
  d.L = L;
  return d;
}

static int check_frame(const SceneStore &S, const rtb_frame_opts *o, int out_h, int out_w, rtb_frame_opts *eff) {
  rtb_frame_opts d;
  memset(&d, 0, sizeof d);
  d.aliasing = 3;
  d.nranks = 1;
  if (o) d = *o;
  if (d.nranks < 1) d.nranks = 1;
  if (d.aliasing < 1 || d.aliasing > 16) { set_error("aliasing %d out of range 1..16", d.aliasing); return RTB_ERR_ARG; }
  if (d.rank < 0 || d.rank >= d.nranks) { set_error("rank %d not in [0,%d)", d.rank, d.nranks); return RTB_ERR_ARG; }
  if (out_w <= 0 || out_h <= 0) { set_error("empty output image"); return RTB_ERR_ARG; }
  if (S.camera.width != out_w * d.aliasing || S.camera.height != out_h * d.aliasing) {
    set_error("camera is %dx%d but output %dx%d with aliasing %d needs %dx%d (camera width/height must be "
              "pre-multiplied by aliasing, gpu/rt.cpp:78-79)",
              S.camera.width, S.camera.height, out_w, out_h, d.aliasing, out_w * d.aliasing, out_h * d.aliasing);
    return RTB_ERR_ARG;
  }
  if (d.tracer != RTB_TRACER_TREE && d.tracer != RTB_TRACER_BRUTE) { set_error("unknown tracer %d", d.tracer); return RTB_ERR_ARG; }
  if (d.tracer == RTB_TRACER_TREE && !S.has_tree && S.n_tris > (uint32_t)LEAF_MAX) {
    set_error("scene was created with RTB_BUILD_NO_TREE: only RTB_TRACER_BRUTE is available");
    return RTB_ERR_ARG;
  }
  *eff = d;
  return RTB_OK;
}

static bool g_schedule = true;   // rtb_set_schedule (tests, A/B)

// The projection pass is the default for the first ray of every sample whatever the scene's size:
// its cull is a bound (region.cuh), the hierarchy's padded boxes are a heuristic.
constexpr uint32_t RASTER_AUTO_MIN_TRIS = 1u;

// Buffers of the projection pass, owned by the scene handle (stream-ordered pool: a one-shot
// render gets them in microseconds), and the box inflation of this frame.
static int ensure_raster_resources(SceneStore &S, FrameArgs &fa, cudaStream_t st) {
  const size_t samples = (size_t)fa.cam.W * fa.cam.H;
  if (S.d_prim_count < samples) {
    if (S.d_prim) block_free(S.d_prim, st);
    S.d_prim = nullptr; S.d_prim_count = 0;
    { const int rc = block_alloc((void **)&S.d_prim, samples * 8, st); if (rc) return rc; }
    S.d_prim_count = samples;
  }
  const size_t cap = 2 * (size_t)S.n_tris + (4u << 20);
  if (S.d_rast_cap < cap) {
    if (S.d_rast_items) block_free(S.d_rast_items, st);
    S.d_rast_items = nullptr; S.d_rast_cap = 0;
    { const int rc = block_alloc((void **)&S.d_rast_items, cap * sizeof(RastItem), st); if (rc) return rc; }
    S.d_rast_cap = cap;
  }
  // state words, the big list, the big list's regions
  constexpr size_t big_off = ((RAST_STATE_WORDS + (1u << 16)) * 4 + 255) & ~(size_t)255;
  if (!S.d_rast) RTB_CUDA(cudaMallocAsync((void **)&S.d_rast, big_off + (size_t)(1u << 16) * sizeof(region::PrimaryCull), st));
  // the camera in double precision and the rounding scales of this frame's rays (region::CamFrame)
  {
    region::CamFrame &cf = fa.camf;
    const double u[3] = {fa.cam.u.x, fa.cam.u.y, fa.cam.u.z}, v[3] = {fa.cam.v.x, fa.cam.v.y, fa.cam.v.z};
    const double pos[3] = {fa.cam.pos.x, fa.cam.pos.y, fa.cam.pos.z}, Cc[3] = {fa.cam.C.x, fa.cam.C.y, fa.cam.C.z};
    const int W = fa.cam.W, H = fa.cam.H;
    const double X = (double)(W / 2 > W - 1 - W / 2 ? W / 2 : W - 1 - W / 2), Y = (double)(H / 2 > H - 1 - H / 2 ? H / 2 : H - 1 - H / 2);
    double mo = 0;
    for (int k = 0; k < 3; ++k) {
      cf.pos[k] = pos[k]; cf.u[k] = u[k]; cf.v[k] = v[k];
      cf.D0[k] = pos[k] - Cc[k];
      const double m = fabs(Cc[k]) + fabs(u[k]) * X + fabs(v[k]) * Y;
      mo += m * m;
    }
    cf.Mo = sqrt(mo);
    double dmax = 0;
    for (int c = 0; c < 4; ++c) {
      const double x = (c & 1) ? (double)(W - 1 - W / 2) : -(double)(W / 2), y = (c & 2) ? (double)(H - 1 - H / 2) : -(double)(H / 2);
      double d2 = 0;
      for (int k = 0; k < 3; ++k) { const double d = cf.D0[k] - u[k] * x - v[k] * y; d2 += d * d; }
      dmax = fmax(dmax, sqrt(d2));
    }
    cf.Dmax = dmax * (1.0 + 1e-6) + 8.0 * region::EPS24 * cf.Mo;
  }
  // a tree walk costs about 3 000 instructions per sample, a rectangle sample about 100
  static const unsigned area_factor = [] { const char *e = getenv("RTB_RASTER_AREA"); return e && atoi(e) > 0 ? (unsigned)atoi(e) : 20u; }();
  fa.rast_area_factor = area_factor;
  S.rast_area_factor = area_factor;
  fa.prim = S.d_prim;
  memset(&fa.x, 0, sizeof fa.x);
  fa.x.cap = (unsigned)(S.d_rast_cap > 0xfffffff0ull ? 0xfffffff0ull : S.d_rast_cap);
  fa.x.seg[0] = S.d_rast_items;
  fa.x.in[0] = S.d_rast_items;
  fa.rast = S.d_rast;
  fa.rast_big = S.d_rast + RAST_STATE_WORDS;
  fa.rast_big_cull = reinterpret_cast<region::PrimaryCull *>(reinterpret_cast<unsigned char *>(S.d_rast) + big_off);
  return RTB_OK;
}

static int ensure_frame_resources(SceneStore &S) {
  if (!S.d_tile_counter) RTB_CUDA(cudaMalloc((void **)&S.d_tile_counter, sizeof(unsigned int)));
  if (!S.d_counters) RTB_CUDA(cudaMalloc((void **)&S.d_counters, 8 * sizeof(unsigned long long)));
  return RTB_OK;
}

static int render_async(SceneStore &S, const rtb_frame_opts &o, unsigned char *d_out, ptrdiff_t pitch, int out_h,
                        int out_w, cudaStream_t st) {
  static const bool tr_ = [] { const char *e = getenv("RTB_TRACE"); return e && *e == '2'; }();
  const auto tr0_ = std::chrono::steady_clock::now();
  auto TR = [&](const char *what) { if (tr_) fprintf(stderr, "  render_async %s +%.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tr0_).count()); };
  int rc = ensure_frame_resources(S);
  if (rc) return rc;
  TR("frame resources");
  // a caller's stream (rtb_scene_render_device): the frame runs behind everything the handle's own stream
  // holds (the scratch it reuses), and the handle's stream waits for the frame, so that rtb_scene_sync, the
  // frame time and the state words cover it whichever stream it ran on
  struct Rejoin {
    SceneStore &S; cudaStream_t st;
    ~Rejoin() {
      if (st == S.stream || !S.xev) return;
      if (cudaEventRecord(S.xev, st) == cudaSuccess) cudaStreamWaitEvent(S.stream, S.xev, 0);
    }
  } rejoin_streams{S, st};
  if (st != S.stream) {
    if (!S.xev) RTB_CUDA(cudaEventCreateWithFlags(&S.xev, cudaEventDisableTiming));
    RTB_CUDA(cudaEventRecord(S.xev, S.stream));
    RTB_CUDA(cudaStreamWaitEvent(st, S.xev, 0));
  }
  FrameArgs fa;
  fa.cam = camera_frame(S.camera);
  fa.aliasing = o.aliasing;
  fa.out_w = out_w;
  fa.out_h = out_h;
  fa.rank = o.rank;
  fa.nranks = o.nranks;
  fa.out = d_out;
  fa.pitch = (long long)pitch;
  fa.tile_counter = S.d_tile_counter;
  fa.counters = S.d_counters;
  fa.item_order = nullptr;
  fa.item_cost = nullptr;
  fa.bounce_scratch = nullptr;
  fa.n_shadow_lights = S.n_shadow_lights;
  memcpy(fa.shadow_light, S.shadow_light, sizeof fa.shadow_light);
  fa.prim = nullptr; fa.rast = nullptr; fa.rast_big = nullptr; fa.rast_big_cull = nullptr; fa.rast_area_factor = 0;
  memset(&fa.x, 0, sizeof fa.x);
  // Primary visibility by projection (raster.cuh): big scenes by default, any scene on request.
  // The frame falls back to the tree by itself when the projection does not pay or is not defined.
  const bool want_raster = o.primary == RTB_PRIMARY_RASTER || (o.primary == RTB_PRIMARY_AUTO && S.n_tris >= RASTER_AUTO_MIN_TRIS);
  if (want_raster && (S.inv_order || !S.has_tree) && o.tracer == RTB_TRACER_TREE && !o.count_rays &&
      o.frame_kernel == RTB_KERNEL_PER_SAMPLE && fa.cam.W <= 65535 && fa.cam.H <= 65535) {      // items hold 16-bit sample corners
    // a frame of this camera that ran the pass and still walked the tree (regions covering too much, the item
    // list full) says the next one will too: AUTO does not try again
    const unsigned long long key[4] = {S.camera_gen, (unsigned long long)fa.cam.W << 32 | (unsigned)fa.cam.H,
                                       (unsigned long long)o.rank << 32 | (unsigned)o.nranks, 0ull};
    const bool same = memcmp(key, S.rast_key, sizeof key) == 0;
    if (!(o.primary == RTB_PRIMARY_AUTO && same && S.rast_useless)) {
      rc = ensure_raster_resources(S, fa, st);
      TR("raster resources");
      if (rc) return rc;
      memcpy(S.rast_key, key, sizeof key);
      S.rast_useless = false;
      // Sharded projection pass: the ranks of this frame are attached to each other's exchange blocks
      // (rtb_scene_exchange_attach).  Every rank must take this branch on the same frames: the epoch counts them.
      if (o.nranks > 1 && S.xchg_block && S.xchg_nranks == o.nranks && S.xchg_rank == o.rank) {
        const int N = o.nranks;
        fa.x.sharded = 1;
        S.rast_key[3] = 1;            // finish_frame: the area the frame left is the whole frame's (k_xchg_sync)
        fa.x.cap = S.xchg_cap;
        fa.x.epoch = ++S.xchg_epoch;
        const size_t p = fa.x.epoch & 1u;
        for (int q = 0; q < N; ++q) {
          fa.x.ctl[q] = reinterpret_cast<unsigned *>(S.xchg_peer[q]);
          fa.x.seg[q] = reinterpret_cast<RastItem *>(S.xchg_peer[q] + XCHG_CTL_WORDS * 4) + (p * N + (size_t)o.rank) * S.xchg_cap;
          fa.x.in[q] = reinterpret_cast<const RastItem *>(S.xchg_block + XCHG_CTL_WORDS * 4) + (p * N + (size_t)q) * S.xchg_cap;
        }
      }
    }
  }
  if (o.frame_kernel == RTB_KERNEL_WAVEFRONT && o.tracer == RTB_TRACER_TREE && o.aliasing <= WAVE_MAX_ALIASING && S.n_shadow_lights <= 32 &&
      S.n_lights <= 255) {
    const size_t need = wave_scratch_floats(o.aliasing, S.sm_count);
    if (S.d_bounce_floats < need) {
      if (S.d_bounce) cudaFree(S.d_bounce);
      S.d_bounce = nullptr;
      S.d_bounce_floats = 0;
      RTB_CUDA(cudaMalloc((void **)&S.d_bounce, need * sizeof(float)));
      S.d_bounce_floats = need;
    }
    fa.bounce_scratch = S.d_bounce;
  }
  // Longest-first scheduling.  Warps take work items from an atomic counter; near the end of
  // the frame the slowest item in flight sets the tail (one item is 1/55 of a warp's frame at
  // 4K on one GPU, 1/7 on eight).  The second frame of a handle (and the first after every
  // camera change) records each item's cost; a 2-pass radix sort of the 16-bit costs on the
  // same stream gives the order the following frames use.  A one-shot render never gets here.
  const bool schedule = g_schedule && !fa.bounce_scratch && o.tracer == RTB_TRACER_TREE && !o.count_rays &&
                        S.frames_done >= 1 && !S.tree_deferred;
  bool record = false;
  uint64_t *sk = nullptr, *sk2 = nullptr;
  uint32_t *sv = nullptr, *sv2 = nullptr, *ss = nullptr;
  size_t ssw = 0;
  unsigned n_items = 0;
  TR("before schedule");
  if (schedule) {
    n_items = frame_item_count(out_w, out_h, o.nranks);
    const int key[5] = {o.rank, o.nranks, out_w, out_h, o.aliasing};
    const size_t cap = ((size_t)n_items + 255) & ~(size_t)255;
    if (S.sched_items < cap) {
      if (S.d_sched) cudaFreeAsync(S.d_sched, st);
      S.d_sched = nullptr; S.d_item_order = nullptr; S.sched_items = 0;
      const size_t bytes = cap * (4 + 8 + 8 + 4 + 4) + radix_sort_scratch_words(cap) * 4 + 1024;
      RTB_CUDA(cudaMallocAsync((void **)&S.d_sched, bytes, st));
      S.sched_items = cap;
    }
    if (memcmp(key, S.sched_key, sizeof key) != 0) {
      memcpy(S.sched_key, key, sizeof key);
      S.d_item_order = nullptr;
    }
    const size_t c = S.sched_items;
    sk = (uint64_t *)S.d_sched; sk2 = sk + c;
    sv = (uint32_t *)(sk2 + c); sv2 = sv + c;
    ss = sv2 + c;
    ssw = radix_sort_scratch_words(c);
    unsigned int *cost = (unsigned int *)(ss + ssw + 64);
    record = !S.d_item_order || S.order_camera_gen != S.camera_gen;
    fa.item_order = S.d_item_order;
    if (record) {
      // the current order (if any) lives in sv or sv2 and is read by this frame: the new one is
      // sorted after the frame on the same stream, so overwriting either is safe
      RTB_CUDA(cudaMemsetAsync(cost, 0, (size_t)n_items * 4, st));
      fa.item_cost = cost;
    }
  }
  // fine work items (one round of a tile each) for multi-rank frames of the per-sample kernel
  fa.fine_acc = nullptr; fa.fine_done = nullptr; fa.fine_hot = nullptr;
  static const int fine_min_ranks = [] { const char *e = getenv("RTB_FINE_ITEMS_FROM"); return e ? atoi(e) : 1; }();
  static const float hot_div = [] { const char *e = getenv("RTB_HOT_DIV"); return e ? (float)atof(e) : 4.0f; }();
  unsigned *d_hot = nullptr;
  const bool fine = schedule && o.nranks >= fine_min_ranks && fine_min_ranks > 0 && o.tracer == RTB_TRACER_TREE && !o.count_rays &&
                    !fa.bounce_scratch && o.aliasing > 1;
  if (schedule)      // 4 words behind the costs, aligned for the 64-bit sum
    d_hot = (unsigned *)(((uintptr_t)((unsigned *)(ss + ssw + 64) + S.sched_items) + 15u) & ~(uintptr_t)15u);
  TR("before fine");
  if (fine) {
    const size_t pixels = (size_t)out_w * out_h, items = frame_item_count(out_w, out_h, o.nranks);
    if (S.fine_pixels < pixels) {
      if (S.d_fine_acc) block_free(S.d_fine_acc, st);
      S.d_fine_acc = nullptr; S.fine_pixels = 0;
      { const int rc2 = block_alloc((void **)&S.d_fine_acc, pixels * sizeof(uint2), st); if (rc2) return rc2; }
      RTB_CUDA(cudaMemsetAsync(S.d_fine_acc, 0, pixels * sizeof(uint2), st));      // once: k_frame leaves it zero
      S.fine_pixels = pixels;
    }
    if (S.fine_items < items) {
      if (S.d_fine_done) cudaFreeAsync(S.d_fine_done, st);
      S.d_fine_done = nullptr; S.fine_items = 0;
      RTB_CUDA(cudaMallocAsync((void **)&S.d_fine_done, items * 4, st));
      RTB_CUDA(cudaMemsetAsync(S.d_fine_done, 0, items * 4, st));
      S.fine_items = items;
    }
    fa.fine_acc = S.d_fine_acc; fa.fine_done = S.d_fine_done; fa.fine_hot = d_hot + 2;
  }
  // Experiment: RTB_L2_PERSIST=<array>:<MB> pins part of one array in L2 for the frame's stream
  // (cudaAccessPolicyWindow, persisting hits / streaming misses).  Arrays: prim, nodes, tris, rec<g>, cells<g>, entries<g>.
  static const char *l2p = getenv("RTB_L2_PERSIST");
  if (l2p && *l2p) {
    const char *colon = strchr(l2p, ':');
    const size_t mb = colon ? (size_t)atoi(colon + 1) : 32;
    const void *ptr = nullptr; size_t bytes = 0;
    auto grid = [&](int which, const char *digit) {
      const int g = (*digit >= '0' && *digit <= '7') ? *digit - '0' : 0;
      ptr = S.lgrid_mem[3 * g + which]; bytes = S.lgrid_mem_bytes[3 * g + which];
    };
    if (!strncmp(l2p, "prim", 4)) { ptr = fa.prim; bytes = (size_t)fa.cam.W * fa.cam.H * 8; }
    else if (!strncmp(l2p, "nodes", 5)) { ptr = S.nodes; bytes = (size_t)S.n_nodes * NODE_QUADS * 16; }
    else if (!strncmp(l2p, "tris", 4)) { ptr = S.v0; bytes = (size_t)S.n_tris * 16; }
    else if (!strncmp(l2p, "rec", 3)) grid(0, l2p + 3);
    else if (!strncmp(l2p, "cells", 5)) grid(1, l2p + 5);
    else if (!strncmp(l2p, "entries", 7)) grid(2, l2p + 7);
    if (ptr && bytes) {
      cudaDeviceProp prop;
      RTB_CUDA(cudaGetDeviceProperties(&prop, S.device));
      const size_t want = mb << 20, cap = (size_t)prop.persistingL2CacheMaxSize;
      RTB_CUDA(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want < cap ? want : cap));
      cudaStreamAttrValue v;
      memset(&v, 0, sizeof v);
      v.accessPolicyWindow.base_ptr = const_cast<void *>(ptr);
      v.accessPolicyWindow.num_bytes = bytes < (size_t)prop.accessPolicyMaxWindowSize ? bytes : (size_t)prop.accessPolicyMaxWindowSize;
      v.accessPolicyWindow.hitRatio = (float)fmin(1.0, (double)(want < cap ? want : cap) / (double)v.accessPolicyWindow.num_bytes);
      v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
      v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
      RTB_CUDA(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v));
      static bool said = false;
      if (!said) {
        said = true;
        fprintf(stderr, "[rtb200] L2 window on %s: %zu of %zu bytes, hit ratio %.3f, persisting L2 %zu MB (device max %d MB, max window %d MB)\n", l2p,
                (size_t)v.accessPolicyWindow.num_bytes, bytes, v.accessPolicyWindow.hitRatio, (want < cap ? want : cap) >> 20,
                prop.persistingL2CacheMaxSize >> 20, prop.accessPolicyMaxWindowSize >> 20);
      }
    }
  }
  RTB_CUDA(cudaEventRecord(S.ev0, st));
  if (S.profile && !S.kev[0])
    for (cudaEvent_t &e : S.kev) RTB_CUDA(cudaEventCreate(&e));
  fa.need_tree = nullptr; fa.stray = nullptr; fa.stray_cap = 0; fa.stray_mode = 0; fa.only_items = 0;
  if (S.tree_deferred) {
    // only a frame of the per-sample tree tracer with the projection pass on can do without the nodes
    if (!fa.prim || o.tracer != RTB_TRACER_TREE || o.count_rays || fa.bounce_scratch) return RTB_NEED_TREE;
    fa.need_tree = fa.rast + 7;      // cleared with the other state words at the start of the frame
    if (!S.d_stray) RTB_CUDA(cudaMallocAsync((void **)&S.d_stray, (size_t)STRAY_CAP * (sizeof(StrayQuery) + 4), st));
    fa.stray = S.d_stray; fa.stray_cap = STRAY_CAP; fa.stray_mode = S.stray_pass == 2 ? 2 : 1;
    if (S.stray_pass == 2) {
      fa.item_order = reinterpret_cast<const unsigned *>(S.d_stray + STRAY_CAP);
      fa.only_items = S.stray_items;
      fa.item_cost = nullptr;
    }
  }
  TR("before launches");
  if (S.stray_pass == 2) {
    RTB_CUDA(cudaEventRecord(S.ev0, st));
    rc = launch_frame_only(S.dev(), fa, S.sm_count, st);
    if (rc) return rc;
    RTB_CUDA(cudaEventRecord(S.ev1, st));
    return RTB_OK;
  }
  // one-shot frames (deferred tree: rtb_render) have no previous frame to order their items by: an estimate instead
  ItemEstimate est_args, *est = nullptr;
  if (S.tree_deferred && g_schedule && fa.prim && !fa.item_order) {
    const unsigned n_est = frame_item_count(out_w, out_h, o.nranks);
    const size_t cap = ((size_t)n_est + 255) & ~(size_t)255;
    if (S.sched_items < cap) {
      if (S.d_sched) cudaFreeAsync(S.d_sched, st);
      S.d_sched = nullptr; S.d_item_order = nullptr; S.sched_items = 0;
      const size_t bytes = cap * (4 + 8 + 8 + 4 + 4) + radix_sort_scratch_words(cap) * 4 + 1024;
      RTB_CUDA(cudaMallocAsync((void **)&S.d_sched, bytes, st));
      S.sched_items = cap;
    }
    const size_t c = S.sched_items;
    est_args.sk = (uint64_t *)S.d_sched; est_args.sk2 = est_args.sk + c;
    est_args.sv = (uint32_t *)(est_args.sk2 + c); est_args.sv2 = est_args.sv + c;
    est_args.ss = est_args.sv2 + c;
    est_args.ssw = radix_sort_scratch_words(c);
    est_args.cost = (unsigned *)(est_args.ss + est_args.ssw + 64);
    est_args.hot = (unsigned *)(((uintptr_t)(est_args.cost + c) + 15u) & ~(uintptr_t)15u);
    est_args.n_items = n_est;
    est = &est_args;
  }
  rc = launch_frame(S.dev(), fa, o.tracer, o.count_rays != 0, S.sm_count, st, &S.frame_launches, S.profile ? S.kev : nullptr, est);
  TR("after launches");
  if (rc) return rc;
  RTB_CUDA(cudaEventRecord(S.ev1, st));
  if (record) {
    int in_alt = 0;
    // a warp's fair share of the frame: the costs are what the warps of the persistent grid spent, stalls included
    const float per_share = fine ? 1.0f / ((float)S.sm_count * 32.0f /* 8 CTAs of 4 warps per SM, trace.cu k_frame */ * hot_div) : 0.0f;
    rc = launch_item_order(fa.item_cost, n_items, sk, sv, sk2, sv2, ss, ssw, st, &in_alt, d_hot, per_share);
    if (rc) return rc;
    S.d_item_order = in_alt ? sv2 : sv;
    S.order_camera_gen = S.camera_gen;
  }
  ++S.frames_done;
  // (a handle built for one frame has no next frame to spare the attempt, and pinning memory costs milliseconds)
  if (fa.prim && !S.one_frame) {
    if (!S.h_rast) RTB_CUDA(cudaMallocHost((void **)&S.h_rast, 8 * sizeof(unsigned)));     // pinned: the copy must not block the host
    RTB_CUDA(cudaMemcpyAsync(S.h_rast, S.d_rast, 8 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    S.rast_pending = true;
  }
  if (o.count_rays)
    RTB_CUDA(cudaMemcpyAsync(S.counters, S.d_counters, sizeof S.counters, cudaMemcpyDeviceToHost, st));
  return RTB_OK;
}

static int finish_frame(SceneStore &S, cudaStream_t st) {
  RTB_CUDA(cudaStreamSynchronize(st));
  if (S.rast_pending) {
    S.rast_pending = false;
    const unsigned long long area = (unsigned long long)S.h_rast[2] | ((unsigned long long)S.h_rast[3] << 32);
    const unsigned long long ranks = S.rast_key[3] ? 1ull : ((S.rast_key[2] & 0xffffffffull) ? (S.rast_key[2] & 0xffffffffull) : 1ull);
    const unsigned long long samples = (S.rast_key[1] >> 32) * (S.rast_key[1] & 0xffffffffull) / ranks;      // as rast_active() counts
    S.rast_useless = S.h_rast[1] != 0 || area > (unsigned long long)S.rast_area_factor * samples + 65536ull;
  }
  float ms = 0;
  if (cudaEventElapsedTime(&ms, S.ev0, S.ev1) == cudaSuccess) S.frame_ms = ms;
  else cudaGetLastError();
  return RTB_OK;
}

// ---- per-device cache: what a scene handle needs besides the scene itself ----
struct FrameCtx {
  int device, sm_count;
  cudaStream_t stream;
  cudaEvent_t ev0, ev1, tev0, tev1;
  unsigned char *d_frame, *h_frame;
  size_t d_frame_bytes, h_frame_bytes;
  unsigned int *d_tile_counter;
  unsigned long long *d_counters;
  float *d_bounce;
  size_t d_bounce_floats;
  // the projection pass's scratch (samples x 8 B, items, state): GB-sized blocks cost the stream-ordered pool
  // milliseconds each to map even when it holds the memory (6 ms of a one-shot render() of C5)
  unsigned long long *d_prim;
  size_t d_prim_count;
  RastItem *d_rast_items;
  size_t d_rast_cap;
  unsigned *d_rast;
};
static std::mutex g_ctx_mu;
static std::vector<FrameCtx> g_ctx_pool;
static std::vector<int> g_pool_tuned;   // devices whose memory pool keeps freed blocks

// ---- block cache (rtb_internal.h) ----
namespace {
struct ParkedBlock { void *p; size_t bytes; cudaStream_t st; cudaEvent_t ev; };
struct BlockCache {
  std::vector<ParkedBlock> parked;                    // oldest first
  std::unordered_map<void *, size_t> live;            // big blocks handed out: their sizes
  size_t parked_bytes = 0;
};
std::mutex g_block_mu;
std::unordered_map<int, BlockCache> g_block_cache;    // per device
constexpr size_t BLOCK_CACHE_MIN = 8u << 20;
size_t block_cache_cap() {
  static const size_t cap = [] { const char *e = getenv("RTB_BLOCK_CACHE_MB"); return (size_t)(e ? atoll(e) : 24576) << 20; }();
  return cap;
}
}  // namespace

int block_alloc(void **p, size_t bytes, cudaStream_t st) {
  *p = nullptr;
  if (bytes >= BLOCK_CACHE_MIN && block_cache_cap()) {
    int dev = 0;
    RTB_CUDA(cudaGetDevice(&dev));
    ParkedBlock hit = {nullptr, 0, nullptr, nullptr};
    {
      std::lock_guard<std::mutex> lk(g_block_mu);
      BlockCache &c = g_block_cache[dev];
      for (size_t i = c.parked.size(); i-- > 0;)
        if (c.parked[i].bytes == bytes) {
          hit = c.parked[i];
          c.parked.erase(c.parked.begin() + (long)i);
          c.parked_bytes -= bytes;
          c.live[hit.p] = bytes;
          break;
        }
    }
    if (hit.p) {
      // whatever the freeing stream still had queued on the block comes first
      if (hit.st != st) RTB_CUDA(cudaStreamWaitEvent(st, hit.ev, 0));
      cudaEventDestroy(hit.ev);
      *p = hit.p;
      return RTB_OK;
    }
  }
  RTB_CUDA(cudaMallocAsync(p, bytes ? bytes : 1, st));
  if (bytes >= BLOCK_CACHE_MIN && block_cache_cap()) {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_block_mu);
    g_block_cache[dev].live[*p] = bytes;
  }
  return RTB_OK;
}

void block_free(void *p, cudaStream_t st) {
  if (!p) return;
  int dev = 0;
  cudaGetDevice(&dev);
  size_t bytes = 0;
  std::vector<ParkedBlock> evict;
  {
    std::lock_guard<std::mutex> lk(g_block_mu);
    BlockCache &c = g_block_cache[dev];
    auto it = c.live.find(p);
    if (it != c.live.end()) { bytes = it->second; c.live.erase(it); }
    if (bytes) {
      cudaEvent_t ev = nullptr;
      if (bytes <= block_cache_cap() && cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess &&
          cudaEventRecord(ev, st) == cudaSuccess) {
        c.parked.push_back({p, bytes, st, ev});
        c.parked_bytes += bytes;
        while (c.parked_bytes > block_cache_cap() && !c.parked.empty()) {
          evict.push_back(c.parked.front());
          c.parked_bytes -= c.parked.front().bytes;
          c.parked.erase(c.parked.begin());
        }
      } else {
        if (ev) cudaEventDestroy(ev);
        cudaGetLastError();
        bytes = 0;
      }
    }
  }
  if (!bytes) cudaFreeAsync(p, st);
  for (ParkedBlock &b : evict) {
    cudaStreamWaitEvent(st, b.ev, 0);
    cudaEventDestroy(b.ev);
    cudaFreeAsync(b.p, st);
  }
}

// cudaMallocAsync's default pool hands freed blocks back to the driver at every
// synchronisation unless told to keep them; keeping them makes scratch allocation ~1 us.
static void tune_pool_locked(int device) {
  for (int d : g_pool_tuned) if (d == device) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  }
  cudaGetLastError();
  g_pool_tuned.push_back(device);
}

int acquire_frame_ctx(int device, SceneStore &S) {
  {
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    tune_pool_locked(device);
    for (size_t i = 0; i < g_ctx_pool.size(); ++i)
      if (g_ctx_pool[i].device == device) {
        const FrameCtx c = g_ctx_pool[i];
        g_ctx_pool.erase(g_ctx_pool.begin() + i);
        S.sm_count = c.sm_count; S.stream = c.stream; S.ev0 = c.ev0; S.ev1 = c.ev1; S.tev0 = c.tev0; S.tev1 = c.tev1;
        S.d_frame = c.d_frame; S.h_frame = c.h_frame; S.d_frame_bytes = c.d_frame_bytes; S.h_frame_bytes = c.h_frame_bytes;
        S.d_tile_counter = c.d_tile_counter; S.d_counters = c.d_counters; S.d_bounce = c.d_bounce;
        S.d_bounce_floats = c.d_bounce_floats;
        S.d_prim = c.d_prim; S.d_prim_count = c.d_prim_count; S.d_rast_items = c.d_rast_items; S.d_rast_cap = c.d_rast_cap;
        S.d_rast = c.d_rast;
        return RTB_OK;
      }
  }
  RTB_CUDA(cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking));
  RTB_CUDA(cudaEventCreate(&S.ev0));
  RTB_CUDA(cudaEventCreate(&S.ev1));
  RTB_CUDA(cudaEventCreate(&S.tev0));
  RTB_CUDA(cudaEventCreate(&S.tev1));
  cudaDeviceGetAttribute(&S.sm_count, cudaDevAttrMultiProcessorCount, device);
  return RTB_OK;
}

void recycle_frame_ctx(SceneStore &S) {
  if (!S.stream) return;
  FrameCtx c;
  c.device = S.device; c.sm_count = S.sm_count; c.stream = S.stream; c.ev0 = S.ev0; c.ev1 = S.ev1; c.tev0 = S.tev0;
  c.tev1 = S.tev1; c.d_frame = S.d_frame; c.h_frame = S.h_frame; c.d_frame_bytes = S.d_frame_bytes;
  c.h_frame_bytes = S.h_frame_bytes; c.d_tile_counter = S.d_tile_counter; c.d_counters = S.d_counters;
  c.d_bounce = S.d_bounce; c.d_bounce_floats = S.d_bounce_floats;
  c.d_prim = S.d_prim; c.d_prim_count = S.d_prim_count; c.d_rast_items = S.d_rast_items; c.d_rast_cap = S.d_rast_cap; c.d_rast = S.d_rast;
  S.d_prim = nullptr; S.d_prim_count = 0; S.d_rast_items = nullptr; S.d_rast_cap = 0; S.d_rast = nullptr;
  S.stream = nullptr; S.ev0 = S.ev1 = S.tev0 = S.tev1 = nullptr; S.d_frame = S.h_frame = nullptr;
  S.d_frame_bytes = S.h_frame_bytes = 0; S.d_tile_counter = nullptr; S.d_counters = nullptr; S.d_bounce = nullptr;
  S.d_bounce_floats = 0;
  std::lock_guard<std::mutex> lk(g_ctx_mu);
  if (g_ctx_pool.size() < 16) { g_ctx_pool.push_back(c); return; }
  if (c.d_frame) cudaFree(c.d_frame);
  if (c.h_frame) cudaFreeHost(c.h_frame);
  if (c.d_tile_counter) cudaFree(c.d_tile_counter);
  if (c.d_counters) cudaFree(c.d_counters);
  if (c.d_bounce) cudaFree(c.d_bounce);
  if (c.d_prim) block_free(c.d_prim, c.stream);
  if (c.d_rast_items) block_free(c.d_rast_items, c.stream);
  if (c.d_rast) cudaFreeAsync(c.d_rast, c.stream);
  cudaStreamSynchronize(c.stream);
  cudaEventDestroy(c.ev0); cudaEventDestroy(c.ev1); cudaEventDestroy(c.tev0); cudaEventDestroy(c.tev1);
  cudaStreamDestroy(c.stream);
}

}  // namespace rtb

using namespace rtb;

extern "C" {

const char *rtb_last_error(void) { return g_err; }
const char *rtb_version(void) { return "rtb200 0.1 (sm_100a)"; }

int rtb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

static int scene_create_impl(const struct scene *scene, int device, unsigned flags, rtb_scene_t **out, UploadGroup *grp, int member);

int rtb_scene_create(const struct scene *scene, int device, unsigned flags, rtb_scene_t **out) {
  return scene_create_impl(scene, device, flags & 0xffffu, out, nullptr, 0);      // (the upper bits are internal)
}

// rtb_render / rtb_render_multi: may this scene's one frame be tried without the node array (build.cu `defer`)?  Only
// when no object reflects (a bounce ray walks the tree; gpu/raytracer.cu:107-122 bounces while nr > 0.01); the size and
// light-grid conditions are checked by the build itself.  RTB_DEFER_TREE=0 switches it off, =2 skips this check (tests:
// the frame then finds out on the device and is rendered again).
static unsigned defer_tree_flag(const struct scene *scene) {
  static const int mode = [] { const char *e = getenv("RTB_DEFER_TREE"); return e ? atoi(e) : 1; }();
  if (mode == 0 || !scene) return 0u;
  if (mode != 2)
    for (size_t o = 0; o < scene->object_count; ++o)
      if (scene->objects && scene->objects[o].nr > 0.0f) return 0u;
  return RTB_BUILD_DEFER_TREE;
}

// After the frame of a deferred-tree handle (build.cu `defer`): did a ray need the nodes?  Shadow rays that were merely
// off their light's grid are listed (FrameArgs::stray): they are resolved against every triangle (k_stray_any) and the
// warp tiles they belong to are rendered once more into d_out with the answers looked up.  *rebuild: anything else
// needed the tree (or the list overflowed): the caller builds the scene again, with its tree.
static int finish_deferred(rtb_scene_t *h, const rtb_frame_opts &o, unsigned char *d_out, ptrdiff_t pitch, int out_h, int out_w,
                           bool *rebuild, bool *fixed) {
  *rebuild = false;
  *fixed = false;
  SceneStore &S = h->store;
  if (!S.tree_deferred || !S.d_rast) return RTB_OK;
  unsigned w[15];
  RTB_CUDA(cudaMemcpyAsync(w, S.d_rast + 7, sizeof w, cudaMemcpyDeviceToHost, S.stream));
  RTB_CUDA(cudaStreamSynchronize(S.stream));
  static const bool trace = [] { const char *tr = getenv("RTB_TRACE"); return tr && *tr == '1'; }();
  const unsigned strays = w[14];
  if (trace && (w[0] || strays))
    fprintf(stderr, "[rtb200] rays that needed the tree: %u nearest-hit; shadow rays off their grid: %u (%u light without grid, %u origin outside "
            "the scene's box, %u off the grid, %u outside a point light's projection)\n", w[9], strays, w[10], w[11], w[12], w[13]);
  if (w[0] || strays > STRAY_CAP) { *rebuild = true; return RTB_OK; }
  if (!strays) return RTB_OK;
  int rc = launch_stray_any(S.dev(), S.d_stray, S.d_rast + 21, S.sm_count, S.stream);
  if (rc) return rc;
  std::vector<StrayQuery> q(strays);
  RTB_CUDA(cudaMemcpyAsync(q.data(), S.d_stray, strays * sizeof(StrayQuery), cudaMemcpyDeviceToHost, S.stream));
  RTB_CUDA(cudaStreamSynchronize(S.stream));
  std::vector<unsigned> items;
  for (const StrayQuery &e : q) items.push_back(e.item);
  std::sort(items.begin(), items.end());
  items.erase(std::unique(items.begin(), items.end()), items.end());
  RTB_CUDA(cudaMemcpyAsync(S.d_stray + STRAY_CAP, items.data(), items.size() * 4, cudaMemcpyHostToDevice, S.stream));
  rtb_frame_opts eff;
  rc = check_frame(S, &o, out_h, out_w, &eff);
  if (rc) return rc;
  S.stray_pass = 2;
  S.stray_items = (unsigned)items.size();
  rc = render_async(S, eff, d_out, pitch, out_h, out_w, S.stream);
  S.stray_pass = 0;
  if (rc) return rc;
  RTB_CUDA(cudaStreamSynchronize(S.stream));      // (items' host copy must outlive the upload)
  *fixed = true;
  return RTB_OK;
}

// the frame in S.d_frame to the caller's buffer (page-locked: direct; else through the pinned staging buffer)
static int download_frame(SceneStore &S, unsigned char *rgba_out, ptrdiff_t stride, int out_h, int out_w) {
  const size_t row = (size_t)out_w * 4, bytes = row * out_h;
  cudaPointerAttributes attr;
  const bool pinned = cudaPointerGetAttributes(&attr, rgba_out) == cudaSuccess && attr.type == cudaMemoryTypeHost;
  cudaGetLastError();
  if (pinned) {
    if ((size_t)stride == row) RTB_CUDA(cudaMemcpyAsync(rgba_out, S.d_frame, bytes, cudaMemcpyDeviceToHost, S.stream));
    else RTB_CUDA(cudaMemcpy2DAsync(rgba_out, (size_t)stride, S.d_frame, row, row, out_h, cudaMemcpyDeviceToHost, S.stream));
    RTB_CUDA(cudaStreamSynchronize(S.stream));
    return RTB_OK;
  }
  if (S.h_frame_bytes < bytes) {
    if (S.h_frame) cudaFreeHost(S.h_frame);
    S.h_frame = nullptr;
    S.h_frame_bytes = 0;
    RTB_CUDA(cudaMallocHost((void **)&S.h_frame, bytes));
    S.h_frame_bytes = bytes;
  }
  RTB_CUDA(cudaMemcpyAsync(S.h_frame, S.d_frame, bytes, cudaMemcpyDeviceToHost, S.stream));
  RTB_CUDA(cudaStreamSynchronize(S.stream));
  for (int y = 0; y < out_h; ++y) memcpy(rgba_out + (size_t)y * stride, S.h_frame + (size_t)y * row, row);
  return RTB_OK;
}

static int scene_create_impl(const struct scene *scene, int device, unsigned flags, rtb_scene_t **out, UploadGroup *grp, int member) {
  if (!out) { set_error("out is NULL"); return RTB_ERR_ARG; }
  *out = nullptr;
  const char *tr = getenv("RTB_TRACE");
  const bool trace = tr && *tr == '1';
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  int rc = pick_device(device);
  if (rc) return rc;
  const auto t1 = clk::now();
  rtb_scene_handle *h = new rtb_scene_handle();
  SceneStore &S = h->store;
  S.device = device;
  rc = acquire_frame_ctx(device, S);
  if (rc) { delete h; return rc; }
  const auto t2 = clk::now();
  rc = build_scene(scene, flags, &S, grp, member);
  if (rc) { delete h; return rc; }
  if (trace) {
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    fprintf(stderr, "rtb_scene_create: device %.2f ms, stream/events/pool %.2f ms, upload+build %.2f ms\n", ms(t0, t1),
            ms(t1, t2), ms(t2, clk::now()));
  }
  *out = h;
  return RTB_OK;
}

void rtb_scene_destroy(rtb_scene_t *h) { delete h; }

int rtb_scene_set_camera(rtb_scene_t *h, const struct camera *cam) {
  if (!h || !cam) { set_error("NULL argument"); return RTB_ERR_ARG; }
  if (camera_origin_radius(*cam) > h->store.origin_radius) {
    set_error("camera puts ray origins %.0f units out, the hierarchy was padded for %.0f: create the scene again "
              "with this camera", camera_origin_radius(*cam), h->store.origin_radius);
    return RTB_ERR_LIMIT;
  }
  h->store.camera = *cam;
  ++h->store.camera_gen;
  return RTB_OK;
}

int rtb_scene_render_device(rtb_scene_t *h, const rtb_frame_opts *opts, void *d_rgba, ptrdiff_t pitch, int out_h,
                            int out_w, void *stream) {
  if (!h || !d_rgba) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  rtb_frame_opts o;
  int rc = check_frame(S, opts, out_h, out_w, &o);
  if (rc) return rc;
  if (pitch < (ptrdiff_t)out_w * 4) { set_error("pitch smaller than a row"); return RTB_ERR_ARG; }
  return render_async(S, o, (unsigned char *)d_rgba, pitch, out_h, out_w, stream ? (cudaStream_t)stream : S.stream);
}

int rtb_scene_sync(rtb_scene_t *h) {
  if (!h) { set_error("NULL argument"); return RTB_ERR_ARG; }
  RTB_CUDA(cudaSetDevice(h->store.device));
  return finish_frame(h->store, h->store.stream);
}

int rtb_scene_render(rtb_scene_t *h, const rtb_frame_opts *opts, unsigned char *rgba_out, ptrdiff_t stride, int out_h,
                     int out_w) {
  if (!h || !rgba_out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  rtb_frame_opts o;
  int rc = check_frame(S, opts, out_h, out_w, &o);
  if (rc) return rc;
  if (stride < (ptrdiff_t)out_w * 4) { set_error("stride smaller than a row"); return RTB_ERR_ARG; }
  const size_t row = (size_t)out_w * 4, bytes = row * out_h;
  if (S.d_frame_bytes < bytes) {
    if (S.d_frame) cudaFree(S.d_frame);
    S.d_frame = nullptr;
    S.d_frame_bytes = 0;
    RTB_CUDA(cudaMalloc((void **)&S.d_frame, bytes));
    S.d_frame_bytes = bytes;
  }
  static const bool trace = [] { const char *tr = getenv("RTB_TRACE"); return tr && *tr == '1'; }();
  const auto t0 = std::chrono::steady_clock::now();
  if (o.nranks > 1) RTB_CUDA(cudaMemsetAsync(S.d_frame, 0, bytes, S.stream));
  rc = render_async(S, o, S.d_frame, (ptrdiff_t)row, out_h, out_w, S.stream);
  if (rc) return rc;
  const auto t1 = std::chrono::steady_clock::now();
  // D2H: straight into the caller's buffer when it is page-locked, else through our pinned staging
  cudaPointerAttributes attr;
  bool pinned = cudaPointerGetAttributes(&attr, rgba_out) == cudaSuccess &&
                (attr.type == cudaMemoryTypeHost);
  cudaGetLastError();
  if (pinned) {
    if ((size_t)stride == row) RTB_CUDA(cudaMemcpyAsync(rgba_out, S.d_frame, bytes, cudaMemcpyDeviceToHost, S.stream));
    else RTB_CUDA(cudaMemcpy2DAsync(rgba_out, (size_t)stride, S.d_frame, row, row, out_h, cudaMemcpyDeviceToHost, S.stream));
    rc = finish_frame(S, S.stream);
    if (trace) {
      auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
      fprintf(stderr, "rtb_scene_render: launches enqueued after %.2f ms, frame + download done after %.2f ms (kernels %.2f)\n", ms(t0, t1),
              ms(t0, std::chrono::steady_clock::now()), S.frame_ms);
    }
    return rc;
  }
  if (S.h_frame_bytes < bytes) {
    if (S.h_frame) cudaFreeHost(S.h_frame);
    S.h_frame = nullptr;
    S.h_frame_bytes = 0;
    RTB_CUDA(cudaMallocHost((void **)&S.h_frame, bytes));
    S.h_frame_bytes = bytes;
  }
  RTB_CUDA(cudaMemcpyAsync(S.h_frame, S.d_frame, bytes, cudaMemcpyDeviceToHost, S.stream));
  rc = finish_frame(S, S.stream);
  if (rc) return rc;
  for (int y = 0; y < out_h; ++y) memcpy(rgba_out + (size_t)y * stride, S.h_frame + (size_t)y * row, row);
  return RTB_OK;
}

int rtb_render(const struct scene *scene, unsigned char *rgba_out, int aliasing, ptrdiff_t stride, int out_h,
               int out_w) {
  // RTB_TRACE=1: wall-clock of the three phases on stderr (first-call costs of a fresh process show up here)
  const char *tr = getenv("RTB_TRACE");
  const bool trace = tr && *tr == '1';
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  rtb_scene_t *h = nullptr;
  int rc = scene_create_impl(scene, 0, RTB_BUILD_ONE_FRAME | defer_tree_flag(scene), &h, nullptr, 0);
  if (rc) return rc;
  auto t1 = clk::now();
  rtb_frame_opts o;
  memset(&o, 0, sizeof o);
  o.aliasing = aliasing;
  o.nranks = 1;
  bool again = false, fixed = false;
  if (h->store.tree_deferred) {
    // frame, then the shadow rays it could not answer without the tree (finish_deferred), then one download
    SceneStore &S = h->store;
    rtb_frame_opts eff;
    rc = check_frame(S, &o, out_h, out_w, &eff);
    if (!rc && (!rgba_out || stride < (ptrdiff_t)out_w * 4)) { set_error(rgba_out ? "stride smaller than a row" : "NULL argument"); rc = RTB_ERR_ARG; }
    const size_t bytes = (size_t)out_w * 4 * out_h;
    if (!rc && S.d_frame_bytes < bytes) {
      if (S.d_frame) cudaFree(S.d_frame);
      S.d_frame = nullptr;
      S.d_frame_bytes = 0;
      RTB_CUDA(cudaMalloc((void **)&S.d_frame, bytes));
      S.d_frame_bytes = bytes;
    }
    if (!rc) rc = render_async(S, eff, S.d_frame, (ptrdiff_t)out_w * 4, out_h, out_w, S.stream);
    again = rc == RTB_NEED_TREE;
    if (!rc) rc = finish_frame(S, S.stream);
    if (!rc) rc = finish_deferred(h, eff, S.d_frame, (ptrdiff_t)out_w * 4, out_h, out_w, &again, &fixed);
    if (!rc && !again) rc = download_frame(S, rgba_out, stride, out_h, out_w);
    if (again) rc = RTB_OK;
  } else
    rc = rtb_scene_render(h, &o, rgba_out, stride, out_h, out_w);
  if (again) {
    // some ray of the frame needed the hierarchy after all (off a light's grid, a fall-back of the projection pass):
    // the scene once more, with its tree
    if (trace) fprintf(stderr, "rtb_render: the frame needed the tree; building it\n");
    rtb_scene_destroy(h);
    h = nullptr;
    rc = scene_create_impl(scene, 0, RTB_BUILD_ONE_FRAME, &h, nullptr, 0);
    if (rc) return rc;
    t1 = clk::now();
    rc = rtb_scene_render(h, &o, rgba_out, stride, out_h, out_w);
  }
  const auto t2 = clk::now();
  rtb_stats st;
  memset(&st, 0, sizeof st);
  if (trace) rtb_scene_stats(h, &st);
  rtb_scene_destroy(h);
  if (trace) {
    auto ms = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double, std::milli>(b - a).count(); };
    fprintf(stderr, "rtb_render: create %.2f ms (device: upload %.2f + build %.2f), render %.2f ms (frame kernel %.2f), destroy %.2f ms\n",
            ms(t0, t1), st.upload_ms, st.build_ms, ms(t1, t2), st.frame_ms, ms(t2, clk::now()));
  }
  return rc;
}

// render() on several GPUs of this process (SURVEY 8e process model; `rt` reads RT_GPUS):
// one host thread per device uploads the scene and builds the hierarchy there (the build is
// deterministic, every device holds the same tree), renders the rank tiles it owns and stores
// the finished pixels straight into device 0's frame through peer access (NVLink); device 0
// downloads the frame.  No collective, no intermediate buffers.
int rtb_render_multi(const struct scene *scene, unsigned char *rgba_out, int aliasing, ptrdiff_t stride, int out_h,
                     int out_w, int n_devices) {
  if (n_devices <= 1) return rtb_render(scene, rgba_out, aliasing, stride, out_h, out_w);
  if (!scene || !rgba_out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  if (out_w <= 0 || out_h <= 0 || stride < (ptrdiff_t)out_w * 4) { set_error("bad frame geometry"); return RTB_ERR_ARG; }
  int have = 0;
  if (cudaGetDeviceCount(&have) != cudaSuccess || have < n_devices) {
    cudaGetLastError();
    set_error("%d devices requested, %d present", n_devices, have);
    return have > 0 ? RTB_ERR_ARG : RTB_ERR_NOGPU;
  }
  int rc = pick_device(0);
  if (rc) return rc;
  const size_t row = (size_t)out_w * 4, bytes = row * out_h;
  unsigned char *d_frame = nullptr;
  RTB_CUDA(cudaMalloc((void **)&d_frame, bytes));
  RTB_CUDA(cudaMemset(d_frame, 0, bytes));
  RTB_CUDA(cudaDeviceSynchronize());      // the other devices' streams know nothing of this one
  std::vector<int> rcs((size_t)n_devices, RTB_OK);
  std::vector<std::string> errs((size_t)n_devices);
  std::vector<std::thread> workers;
  const char *tr = getenv("RTB_TRACE");
  const bool trace = tr && *tr == '1';
  using clk = std::chrono::steady_clock;
  const auto t_start = clk::now();
  // shared upload (UploadGroup, scene_store.h): device 0 reads the host, the others read device 0; RTB_SHARED_UPLOAD=0: A/B
  UploadGroup grp;
  grp.n = n_devices;
  const char *su = getenv("RTB_SHARED_UPLOAD");
  bool share_upload = n_devices <= 8 && !(su && *su == '0');
  for (int b = 1; b < n_devices && share_upload; ++b) {
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, b, 0) != cudaSuccess || !can) share_upload = false;
  }
  cudaGetLastError();
  for (int dev = 0; dev < n_devices && dev < 8; ++dev) grp.device[dev] = dev;
  struct GroupGuard {
    UploadGroup &g;
    ~GroupGuard() {
      cudaSetDevice(0);
      if (g.ready_v) cudaEventDestroy(g.ready_v);
      if (g.ready_n) cudaEventDestroy(g.ready_n);
      if (g.src_v) cudaFree(g.src_v);
      if (g.src_n) cudaFree(g.src_n);
    }
  } group_guard{grp};
  if (share_upload) {
    const size_t raw = 36ull * (scene->triangle_count ? scene->triangle_count : 1);
    RTB_CUDA(cudaMalloc((void **)&grp.src_v, raw));
    RTB_CUDA(cudaMalloc((void **)&grp.src_n, raw));
    RTB_CUDA(cudaEventCreateWithFlags(&grp.ready_v, cudaEventDisableTiming));
    RTB_CUDA(cudaEventCreateWithFlags(&grp.ready_n, cudaEventDisableTiming));
  }
  for (int dev = 0; dev < n_devices; ++dev)
    workers.emplace_back([&, dev] {
      auto ms = [&](clk::time_point b) { return std::chrono::duration<double, std::milli>(b - t_start).count(); };
      const auto t0 = clk::now();
      rtb_scene_t *h = nullptr;
      // every device reads the scene from device 0 (UploadGroup) and stores its pixels into device 0's frame
      int r = pick_device(dev);
      if (!r && dev != 0) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(0, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) r = cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        cudaGetLastError();
      }
      if (!r) r = scene_create_impl(scene, dev, RTB_BUILD_ONE_FRAME | defer_tree_flag(scene), &h, share_upload ? &grp : nullptr, dev);
      if (r) grp.fail();
      const auto t1 = clk::now();
      const auto t2 = clk::now();
      if (!r) {
        rtb_frame_opts o;
        memset(&o, 0, sizeof o);
        o.aliasing = aliasing;
        o.rank = dev;
        o.nranks = n_devices;
        r = rtb_scene_render_device(h, &o, d_frame, (ptrdiff_t)row, out_h, out_w, nullptr);
        if (!r) r = rtb_scene_sync(h);
        bool again = r == RTB_NEED_TREE, fixed = false;
        if (!r) r = finish_deferred(h, o, d_frame, (ptrdiff_t)row, out_h, out_w, &again, &fixed);
        if (again) {
          // this device's tiles met a ray that needs the hierarchy: its scene once more, with the tree (from the host:
          // the shared upload was for the first build)
          rtb_scene_destroy(h);
          h = nullptr;
          r = scene_create_impl(scene, dev, RTB_BUILD_ONE_FRAME, &h, nullptr, 0);
          if (!r) r = rtb_scene_render_device(h, &o, d_frame, (ptrdiff_t)row, out_h, out_w, nullptr);
          if (!r) r = rtb_scene_sync(h);
        }
      }
      const auto t3 = clk::now();
      if (r) errs[(size_t)dev] = g_err;
      rtb_stats st;
      memset(&st, 0, sizeof st);
      if (trace && h) rtb_scene_stats(h, &st);
      if (h) rtb_scene_destroy(h);
      if (trace)
        fprintf(stderr, "rtb_render_multi dev %d: thread start +%.2f, create until +%.2f (device: upload %.2f + build %.2f), peer access +%.2f, "
                "frame until +%.2f (kernels %.2f), destroy until +%.2f ms\n", dev, ms(t0), ms(t1), st.upload_ms, st.build_ms, ms(t2), ms(t3),
                st.frame_ms, ms(clk::now()));
      rcs[(size_t)dev] = r;
    });
  for (std::thread &t : workers) t.join();
  for (int dev = 0; dev < n_devices; ++dev)
    if (rcs[(size_t)dev]) {
      set_error("device %d: %s", dev, errs[(size_t)dev].c_str());
      cudaSetDevice(0);
      cudaFree(d_frame);
      return rcs[(size_t)dev];
    }
  RTB_CUDA(cudaSetDevice(0));
  cudaError_t e = cudaMemcpy2D(rgba_out, (size_t)stride, d_frame, row, row, (size_t)out_h, cudaMemcpyDeviceToHost);
  cudaFree(d_frame);
  if (e != cudaSuccess) return cuda_fail(e, "frame download", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_scene_hit_ids(rtb_scene_t *h, const rtb_frame_opts *opts, int32_t *tri_out, int32_t *obj_out, int out_h,
                      int out_w) {
  if (!h || !tri_out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  rtb_frame_opts o;
  int rc = check_frame(S, opts, out_h, out_w, &o);
  if (rc) return rc;
  const size_t n = (size_t)S.camera.width * S.camera.height;
  int32_t *d_tri = nullptr, *d_obj = nullptr;
  RTB_CUDA(cudaMalloc((void **)&d_tri, n * 4));
  if (obj_out && cudaMalloc((void **)&d_obj, n * 4) != cudaSuccess) {
    cudaFree(d_tri);
    return cuda_fail(cudaGetLastError(), "cudaMalloc(obj ids)", __FILE__, __LINE__);
  }
  // opts->primary == RTB_PRIMARY_RASTER: the ids the projection pass records (tests); a frame that
  // falls back to the tree reports -2 everywhere so that a test cannot pass by accident
  if (o.primary != RTB_PRIMARY_TREE && o.tracer == RTB_TRACER_TREE && (S.inv_order || !S.has_tree) && S.n_tris >= RASTER_AUTO_MIN_TRIS &&
      S.camera.width <= 65535 && S.camera.height <= 65535) {
    FrameArgs fa;
    memset(&fa, 0, sizeof fa);
    fa.cam = camera_frame(S.camera);
    fa.aliasing = o.aliasing; fa.out_w = out_w; fa.out_h = out_h; fa.rank = 0; fa.nranks = 1;
    rc = ensure_raster_resources(S, fa, S.stream);
    if (!rc) rc = launch_prim_ids(S.dev(), fa, d_tri, d_obj, S.sm_count, S.stream);
    if (!rc && o.primary == RTB_PRIMARY_AUTO) {
      // AUTO reports what a frame would use: the tree walk when the projection pass gave up
      unsigned st8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      cudaError_t e2 = cudaMemcpyAsync(st8, S.d_rast, sizeof st8, cudaMemcpyDeviceToHost, S.stream);
      if (e2 == cudaSuccess) e2 = cudaStreamSynchronize(S.stream);
      if (e2 != cudaSuccess) rc = cuda_fail(e2, "raster state readback", __FILE__, __LINE__);
      const unsigned long long area = (unsigned long long)st8[2] | ((unsigned long long)st8[3] << 32);
      if (!rc && (st8[1] || area > (unsigned long long)fa.rast_area_factor * n + 65536ull))
        rc = launch_hit_ids(S.dev(), camera_frame(S.camera), o.tracer, d_tri, d_obj, S.stream);
    }
  } else
    rc = launch_hit_ids(S.dev(), camera_frame(S.camera), o.tracer, d_tri, d_obj, S.stream);
  cudaError_t e = cudaSuccess;
  if (!rc) {
    e = cudaMemcpyAsync(tri_out, d_tri, n * 4, cudaMemcpyDeviceToHost, S.stream);
    if (e == cudaSuccess && obj_out) e = cudaMemcpyAsync(obj_out, d_obj, n * 4, cudaMemcpyDeviceToHost, S.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(S.stream);
  }
  cudaFree(d_tri);
  if (d_obj) cudaFree(d_obj);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "hit-id readback", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_scene_trace_rays(rtb_scene_t *h, int tracer, int any_hit, const float *rays, size_t n, int32_t *tri_out,
                         float *dist_out) {
  if (!h || (n && !rays)) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  if (tracer != RTB_TRACER_TREE && tracer != RTB_TRACER_BRUTE) { set_error("unknown tracer %d", tracer); return RTB_ERR_ARG; }
  if (tracer == RTB_TRACER_TREE && !S.has_tree && S.n_tris > (uint32_t)LEAF_MAX) {
    set_error("scene was created with RTB_BUILD_NO_TREE");
    return RTB_ERR_ARG;
  }
  if (!n) return RTB_OK;
  float *d_rays = nullptr, *d_dist = nullptr;
  int32_t *d_tri = nullptr;
  cudaError_t e = cudaMalloc((void **)&d_rays, n * 24);
  if (e == cudaSuccess) e = cudaMalloc((void **)&d_tri, n * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&d_dist, n * 4);
  int rc = RTB_OK;
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_rays, rays, n * 24, cudaMemcpyHostToDevice, S.stream);
  if (e == cudaSuccess) rc = launch_trace_rays(S.dev(), tracer, any_hit, d_rays, n, d_tri, d_dist, S.stream);
  if (e == cudaSuccess && !rc && tri_out) e = cudaMemcpyAsync(tri_out, d_tri, n * 4, cudaMemcpyDeviceToHost, S.stream);
  if (e == cudaSuccess && !rc && dist_out) e = cudaMemcpyAsync(dist_out, d_dist, n * 4, cudaMemcpyDeviceToHost, S.stream);
  if (e == cudaSuccess && !rc) e = cudaStreamSynchronize(S.stream);
  cudaFree(d_rays); cudaFree(d_tri); cudaFree(d_dist);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "trace_rays", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_scene_stats(rtb_scene_t *h, rtb_stats *out) {
  if (!h || !out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  const SceneStore &S = h->store;
  memset(out, 0, sizeof *out);
  out->triangles = S.n_tris; out->objects = S.n_objects; out->lights = S.n_lights;
  out->nodes = S.n_nodes; out->leaves = S.n_leaves; out->max_depth = (uint64_t)S.height; out->max_leaf_tris = S.max_leaf;
  out->tree_bytes = (uint64_t)S.n_nodes * NODE_QUADS * 16;
  out->triangle_bytes = (uint64_t)S.n_tris * 96;
  out->upload_ms = S.upload_ms; out->build_ms = S.build_ms; out->frame_ms = S.frame_ms;
  out->rays_primary = S.counters[0]; out->rays_shadow = S.counters[1]; out->rays_bounce = S.counters[2];
  out->node_visits = S.counters[3]; out->tri_tests = S.counters[4];
  out->kernel_launches = S.frame_launches;
  out->build_launches = S.build_launches;
  out->key_mode = (uint64_t)S.key_mode;
  out->mode_cost = 0;
  if (S.mode_cost[0] > 0)
    for (int k = 1; k < 4; ++k) {
      unsigned long long rel = (unsigned long long)(1000.0 * S.mode_cost[k] / S.mode_cost[0] + 0.5);
      if (rel > 0xffffull) rel = 0xffffull;
      out->mode_cost |= rel << (16 * (k - 1));
    }
  out->builder = (uint64_t)S.builder;
  return RTB_OK;
}

// ---- item exchange of sharded multi-rank frames (raster.cuh) ----
int rtb_scene_exchange_create(rtb_scene_t *h, int nranks, void **d_block, size_t *bytes) {
  if (!h || !d_block) { set_error("NULL argument"); return RTB_ERR_ARG; }
  if (nranks < 2 || nranks > XCHG_MAX_RANKS) { set_error("exchange: %d ranks not in 2..%d", nranks, XCHG_MAX_RANKS); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  if (S.xchg_block) { cudaFree(S.xchg_block); S.xchg_block = nullptr; S.xchg_nranks = 0; S.xchg_rank = -1; }
  // a pair (source, destination) sees 1 / N^2 of the frame's items on average (tiles interleave over the
  // ranks, triangle blocks are dealt round robin): four times that, but never more than the whole list
  const size_t all = 2 * (size_t)S.n_tris + (4u << 20);
  size_t cap = 4 * all / ((size_t)nranks * nranks) + 65536;
  if (cap > all) cap = all;
  const size_t total = (size_t)XCHG_CTL_WORDS * 4 + 2 * (size_t)nranks * cap * sizeof(RastItem);
  RTB_CUDA(cudaMalloc((void **)&S.xchg_block, total));          // cudaMalloc, not the pool: the block is exported over CUDA IPC
  RTB_CUDA(cudaMemset(S.xchg_block, 0, (size_t)XCHG_CTL_WORDS * 4));
  RTB_CUDA(cudaDeviceSynchronize());
  S.xchg_bytes = total;
  S.xchg_cap = (unsigned)cap;
  S.xchg_epoch = 0;
  *d_block = S.xchg_block;
  if (bytes) *bytes = total;
  return RTB_OK;
}

int rtb_scene_exchange_attach(rtb_scene_t *h, int rank, int nranks, void *const *blocks) {
  if (!h || !blocks) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  if (!S.xchg_block) { set_error("exchange: rtb_scene_exchange_create first"); return RTB_ERR_ARG; }
  if (nranks < 2 || nranks > XCHG_MAX_RANKS || rank < 0 || rank >= nranks) { set_error("exchange: rank %d of %d", rank, nranks); return RTB_ERR_ARG; }
  for (int q = 0; q < nranks; ++q) {
    if (q != rank && !blocks[q]) { set_error("exchange: block of rank %d is NULL", q); return RTB_ERR_ARG; }
    S.xchg_peer[q] = q == rank ? S.xchg_block : (unsigned char *)blocks[q];
  }
  S.xchg_nranks = nranks;
  S.xchg_rank = rank;
  S.xchg_epoch = 0;
  return RTB_OK;
}

int rtb_scene_profile(rtb_scene_t *h, int on) {
  if (!h) { set_error("NULL argument"); return RTB_ERR_ARG; }
  h->store.profile = on != 0;
  return RTB_OK;
}

int rtb_scene_kernel_ms(rtb_scene_t *h, double out[8]) {
  if (!h || !out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  for (int k = 0; k < 8; ++k) out[k] = 0.0;
  if (!S.kev[0]) { set_error("rtb_scene_profile was not switched on before the frame"); return RTB_ERR_ARG; }
  RTB_CUDA(cudaSetDevice(S.device));
  RTB_CUDA(cudaEventSynchronize(S.kev[5]));
  auto span = [&](int a, int b, double &ms_out) -> int {
    float ms = 0;
    RTB_CUDA(cudaEventElapsedTime(&ms, S.kev[a], S.kev[b]));
    ms_out = ms;
    return RTB_OK;
  };
  int rc = span(0, 3, out[0]);                 // clears + k_rast_setup + k_rast_big + k_xchg_sync
  if (!rc) rc = span(3, 4, out[1]);            // k_rast_items
  if (!rc) rc = span(4, 5, out[2]);            // the frame kernel
  if (!rc) rc = span(0, 5, out[3]);
  if (!rc) rc = span(0, 1, out[4]);            // clears + k_rast_setup
  if (!rc) rc = span(1, 2, out[5]);            // k_rast_big
  if (!rc) rc = span(2, 3, out[6]);            // k_xchg_sync: the wait for the slowest rank's items
  return rc;
}

int rtb_scene_raster_state(rtb_scene_t *h, unsigned out[8]) {
  if (!h || !out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  memset(out, 0, 8 * sizeof(unsigned));
  const unsigned grids[3] = {(unsigned)S.n_lgrids, (unsigned)(S.lgrid_entries >> 10), (unsigned)(S.lgrid_cells >> 10)};
  memcpy(out + 5, grids, sizeof grids);
  if (!S.d_rast) return RTB_OK;
  RTB_CUDA(cudaSetDevice(S.device));
  unsigned st16[16];
  RTB_CUDA(cudaMemcpyAsync(st16, S.d_rast, sizeof st16, cudaMemcpyDeviceToHost, S.stream));
  RTB_CUDA(cudaStreamSynchronize(S.stream));
  memcpy(out, st16, 5 * sizeof(unsigned));
  if (getenv("RTB_BUILD_TRACE"))
    fprintf(stderr, "[rtb200] projection pass: %u triangles on the big list, %u with two-part regions, exchange timeouts %u\n", st16[5], st16[4], st16[6]);
  out[0] = 0;
  for (int q = 0; q < XCHG_MAX_RANKS; ++q) out[0] += st16[8 + q];      // items this rank wrote (for every destination)
  return RTB_OK;
}

int rtb_scene_dump_tree(rtb_scene_t *h, void *nodes_out, size_t *n_nodes, uint32_t *perm_out, size_t *n_tris) {
  if (!h) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  if (n_nodes) *n_nodes = S.n_nodes;
  if (n_tris) *n_tris = S.n_tris;
  if (nodes_out) RTB_CUDA(cudaMemcpy(nodes_out, S.nodes, (size_t)S.n_nodes * NODE_QUADS * 16, cudaMemcpyDeviceToHost));
  if (perm_out && S.n_tris) RTB_CUDA(cudaMemcpy(perm_out, S.order, (size_t)S.n_tris * 4, cudaMemcpyDeviceToHost));
  return RTB_OK;
}

/* extension: index of the root node in the dumped array */
int rtb_scene_root(rtb_scene_t *h) { return h ? h->store.root : -1; }

/* extension (tests): box inflation factor of the conservative slab test */
void rtb_set_margin_k(float k) { set_margin_k(k); }
void rtb_set_schedule(int on) { g_schedule = on != 0; }

// ---- bracket timer on the library stream (bench: K frames between two events) ----
int rtb_scene_timer_start(rtb_scene_t *h) {
  if (!h) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  RTB_CUDA(cudaEventRecord(S.tev0, S.stream));
  return RTB_OK;
}

int rtb_scene_timer_stop(rtb_scene_t *h, double *ms_out) {
  if (!h || !ms_out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  SceneStore &S = h->store;
  RTB_CUDA(cudaSetDevice(S.device));
  if (!S.tev0) { set_error("timer not started"); return RTB_ERR_ARG; }
  RTB_CUDA(cudaEventRecord(S.tev1, S.stream));
  RTB_CUDA(cudaEventSynchronize(S.tev1));
  float ms = 0;
  RTB_CUDA(cudaEventElapsedTime(&ms, S.tev0, S.tev1));
  *ms_out = ms;
  return RTB_OK;
}

// ---- bulk memory of the host scene container (host/scene.cpp): pinned when a device exists ----
static std::mutex g_pin_mu;
static std::vector<void *> g_pinned;

void *rtb_scene_mem_alloc(size_t bytes) {
  void *p = nullptr;
  if (bytes >= (1u << 20)) {      // small arrays: plain memory, and no CUDA call at all (the first one initialises the driver)
    if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) == cudaSuccess) {
      std::lock_guard<std::mutex> lk(g_pin_mu);
      g_pinned.push_back(p);
      return p;
    }
    cudaGetLastError();
  }
  return malloc(bytes);
}

// Driver and context start-up (0.2 - 1 s in a fresh process) without doing any work: loaders
// call this on a side thread so that it overlaps with reading the scene (host/parser.cpp).
void rtb_warmup(void) {
  cudaFree(0);
  cudaGetLastError();
}

void rtb_scene_mem_free(void *p) {
  if (!p) return;
  {
    std::lock_guard<std::mutex> lk(g_pin_mu);
    for (size_t i = 0; i < g_pinned.size(); ++i)
      if (g_pinned[i] == p) {
        g_pinned.erase(g_pinned.begin() + i);
        cudaFreeHost(p);
        return;
      }
  }
  free(p);
}

// ---- page-locked host memory and CUDA IPC helpers (multi-process tile gather) ----
void *rtb_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void rtb_host_free(void *p) { if (p) cudaFreeHost(p); }

int rtb_dev_alloc(int device, size_t bytes, void **out) {
  if (!out) { set_error("NULL argument"); return RTB_ERR_ARG; }
  int rc = pick_device(device);
  if (rc) return rc;
  RTB_CUDA(cudaMalloc(out, bytes ? bytes : 1));
  return RTB_OK;
}
int rtb_dev_free(void *p) { if (p) RTB_CUDA(cudaFree(p)); return RTB_OK; }
int rtb_dev_memset(void *p, int value, size_t bytes) { RTB_CUDA(cudaMemset(p, value, bytes)); return RTB_OK; }
int rtb_dev_download(void *dst_host, const void *src_dev, size_t bytes) {
  RTB_CUDA(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
  return RTB_OK;
}
// Double-buffered frames: the download of frame k on a copy stream of its own while frame k + 1 renders.
// One copy stream per device, created on first use; the caller has made sure the frame is complete
// (rtb_scene_sync, a barrier between ranks) before it asks for the copy.
static std::mutex g_copy_mu;
static cudaStream_t g_copy_stream[64] = {nullptr};
static int copy_stream_of(const void *dev_ptr, cudaStream_t *out, int *device) {
  cudaPointerAttributes attr;
  RTB_CUDA(cudaPointerGetAttributes(&attr, dev_ptr));
  const int dev = attr.device;
  if (attr.type != cudaMemoryTypeDevice || dev < 0 || dev >= 64) { set_error("not a device pointer"); return RTB_ERR_ARG; }
  RTB_CUDA(cudaSetDevice(dev));
  std::lock_guard<std::mutex> lk(g_copy_mu);
  if (!g_copy_stream[dev]) RTB_CUDA(cudaStreamCreateWithFlags(&g_copy_stream[dev], cudaStreamNonBlocking));
  *out = g_copy_stream[dev];
  if (device) *device = dev;
  return RTB_OK;
}
int rtb_dev_download_async(void *dst_host, const void *src_dev, size_t bytes) {
  if (!dst_host || !src_dev) { set_error("NULL argument"); return RTB_ERR_ARG; }
  cudaStream_t st = nullptr;
  const int rc = copy_stream_of(src_dev, &st, nullptr);
  if (rc) return rc;
  RTB_CUDA(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, st));
  return RTB_OK;
}
int rtb_dev_download_wait(const void *src_dev) {
  if (!src_dev) { set_error("NULL argument"); return RTB_ERR_ARG; }
  cudaStream_t st = nullptr;
  const int rc = copy_stream_of(src_dev, &st, nullptr);
  if (rc) return rc;
  RTB_CUDA(cudaStreamSynchronize(st));
  return RTB_OK;
}
int rtb_ipc_export(void *d_ptr, unsigned char handle_out[64]) {
  cudaIpcMemHandle_t hnd;
  RTB_CUDA(cudaIpcGetMemHandle(&hnd, d_ptr));
  static_assert(sizeof hnd == 64, "ipc handle size");
  memcpy(handle_out, &hnd, 64);
  return RTB_OK;
}
int rtb_ipc_open(const unsigned char handle[64], void **out) {
  cudaIpcMemHandle_t hnd;
  memcpy(&hnd, handle, 64);
  RTB_CUDA(cudaIpcOpenMemHandle(out, hnd, cudaIpcMemLazyEnablePeerAccess));
  return RTB_OK;
}
int rtb_ipc_close(void *p) { RTB_CUDA(cudaIpcCloseMemHandle(p)); return RTB_OK; }

// ---- build primitives on host arrays (known-answer replays of gpu/test.cu:288-340) ----
static int sort_host(uint64_t *keys64, uint32_t *keys32, uint32_t *vals, size_t n, int bits) {
  int rc = pick_device(0);
  if (rc) return rc;
  if (!n) return RTB_OK;
  std::vector<uint64_t> wide;
  if (keys32) { wide.resize(n); for (size_t i = 0; i < n; ++i) wide[i] = keys32[i]; }
  const uint64_t *src = keys32 ? wide.data() : keys64;
  std::vector<uint32_t> iota;
  if (!vals) { iota.resize(n); for (size_t i = 0; i < n; ++i) iota[i] = (uint32_t)i; }
  uint64_t *dk = nullptr, *dk2 = nullptr;
  uint32_t *dv = nullptr, *dv2 = nullptr, *ds = nullptr;
  const size_t sw = radix_sort_scratch_words(n);
  cudaError_t e = cudaMalloc((void **)&dk, n * 8);
  if (e == cudaSuccess) e = cudaMalloc((void **)&dk2, n * 8);
  if (e == cudaSuccess) e = cudaMalloc((void **)&dv, n * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&dv2, n * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&ds, sw * 4);
  if (e == cudaSuccess) e = cudaMemcpy(dk, src, n * 8, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(dv, vals ? vals : iota.data(), n * 4, cudaMemcpyHostToDevice);
  int in_alt = 0;
  if (e == cudaSuccess) rc = dev_radix_sort_pairs_u64(dk, dv, dk2, dv2, n, bits, ds, sw, 0, &in_alt, nullptr);
  if (e == cudaSuccess && !rc) {
    std::vector<uint64_t> back(keys32 ? n : 0);
    e = cudaMemcpy(keys32 ? back.data() : keys64, in_alt ? dk2 : dk, n * 8, cudaMemcpyDeviceToHost);
    if (keys32 && e == cudaSuccess) for (size_t i = 0; i < n; ++i) keys32[i] = (uint32_t)back[i];
    if (e == cudaSuccess && vals) e = cudaMemcpy(vals, in_alt ? dv2 : dv, n * 4, cudaMemcpyDeviceToHost);
  }
  cudaFree(dk); cudaFree(dk2); cudaFree(dv); cudaFree(dv2); cudaFree(ds);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "radix sort (host arrays)", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_radix_sort_pairs_u32(uint32_t *keys, uint32_t *vals, size_t n) {
  if (n && !keys) { set_error("NULL argument"); return RTB_ERR_ARG; }
  return sort_host(nullptr, keys, vals, n, 32);
}
int rtb_radix_sort_pairs_u64(uint64_t *keys, uint32_t *vals, size_t n) {
  if (n && !keys) { set_error("NULL argument"); return RTB_ERR_ARG; }
  return sort_host(keys, nullptr, vals, n, 64);
}

int rtb_inclusive_scan_u32(uint32_t *data, size_t n) {
  int rc = pick_device(0);
  if (rc) return rc;
  if (!n) return RTB_OK;
  if (!data) { set_error("NULL argument"); return RTB_ERR_ARG; }
  uint32_t *d = nullptr, *s = nullptr;
  const size_t sw = scan_scratch_words(n);
  cudaError_t e = cudaMalloc((void **)&d, n * 4);
  if (e == cudaSuccess) e = cudaMalloc((void **)&s, sw * 4);
  if (e == cudaSuccess) e = cudaMemcpy(d, data, n * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) rc = dev_exclusive_scan_u32(d, n, s, sw, true, 0, nullptr);
  if (e == cudaSuccess && !rc) e = cudaMemcpy(data, d, n * 4, cudaMemcpyDeviceToHost);
  cudaFree(d); cudaFree(s);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "scan (host arrays)", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_inclusive_scan_u64(uint64_t *data, size_t n) {
  int rc = pick_device(0);
  if (rc) return rc;
  if (!n) return RTB_OK;
  if (!data) { set_error("NULL argument"); return RTB_ERR_ARG; }
  uint64_t *d = nullptr, *s = nullptr;
  const size_t sw = scan_scratch_words(n);
  cudaError_t e = cudaMalloc((void **)&d, n * 8);
  if (e == cudaSuccess) e = cudaMalloc((void **)&s, sw * 8);
  if (e == cudaSuccess) e = cudaMemcpy(d, data, n * 8, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) rc = dev_inclusive_scan_u64(d, n, s, sw, 0);
  if (e == cudaSuccess && !rc) e = cudaMemcpy(data, d, n * 8, cudaMemcpyDeviceToHost);
  cudaFree(d); cudaFree(s);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "scan (host arrays)", __FILE__, __LINE__);
  return RTB_OK;
}

// Scratch comes from the stream-ordered pool on the caller's stream: no device-wide
// synchronisation, and after the first call of a given size no driver allocation either.
static int pool_scratch(void **p, size_t bytes, cudaStream_t st) {
  int device = 0;
  RTB_CUDA(cudaGetDevice(&device));
  { std::lock_guard<std::mutex> lk(g_ctx_mu); tune_pool_locked(device); }
  RTB_CUDA(cudaMallocAsync(p, bytes, st));
  return RTB_OK;
}

int rtb_dev_radix_sort_pairs_u64(uint64_t *d_keys, uint32_t *d_vals, size_t n, int key_bits, void *stream) {
  if (!n) return RTB_OK;
  if (!d_keys || !d_vals) { set_error("NULL argument"); return RTB_ERR_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  const size_t sw = radix_sort_scratch_words(n);
  // one block: alternate keys | alternate values | histogram scratch
  const size_t off_v = (n * 8 + 255) & ~(size_t)255, off_s = (off_v + n * 4 + 255) & ~(size_t)255;
  unsigned char *blk = nullptr;
  int rc = pool_scratch((void **)&blk, off_s + sw * 4, st);
  if (rc) return rc;
  uint64_t *dk2 = (uint64_t *)blk;
  uint32_t *dv2 = (uint32_t *)(blk + off_v), *ds = (uint32_t *)(blk + off_s);
  int in_alt = 0;
  rc = dev_radix_sort_pairs_u64(d_keys, d_vals, dk2, dv2, n, key_bits, ds, sw, st, &in_alt, nullptr);
  cudaError_t e = cudaSuccess;
  if (!rc && in_alt) {
    e = cudaMemcpyAsync(d_keys, dk2, n * 8, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_vals, dv2, n * 4, cudaMemcpyDeviceToDevice, st);
  }
  cudaFreeAsync(blk, st);
  if (rc) return rc;
  if (e != cudaSuccess) return cuda_fail(e, "device radix sort", __FILE__, __LINE__);
  return RTB_OK;
}

int rtb_dev_inclusive_scan_u32(uint32_t *d_data, size_t n, void *stream) {
  if (!n) return RTB_OK;
  if (!d_data) { set_error("NULL argument"); return RTB_ERR_ARG; }
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t *s = nullptr;
  const size_t sw = scan_scratch_words(n);
  int rc = pool_scratch((void **)&s, (sw ? sw : 1) * 4, st);
  if (rc) return rc;
  rc = dev_exclusive_scan_u32(d_data, n, s, sw, true, st, nullptr);
  cudaFreeAsync(s, st);
  return rc;
}

}  // extern "C"
