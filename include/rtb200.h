/*
 * rtb200.h — C ABI of librtb200.so, the B200 (sm_100a) ray-tracing hot path that
 * replaces the renderer library of blink97/raytracing-gpu behind its host surface.
 *
 * Every entry point cites the reference interface it stands in for
 * (paths relative to the reference checkout).  No torch / C++ types cross this
 * boundary: plain pointers, sizes and POD structs only.
 *
 * The scene structs below are layout-compatible with the reference's default
 * build (LAYOUT_SOA + PARTITIONING_OCTREE, gpu/headers/scene.h:16-170); the
 * static asserts at the end of this file pin sizes and offsets.
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- scene model (ABI types; gpu/headers/scene.h, gpu/headers/vector3.h:7) ---- */

#if defined(__VECTOR_TYPES_H__)
typedef float3 vector3;                      /* gpu/headers/vector3.h:7 */
#else
typedef struct rtb_float3 { float x, y, z; } vector3;
#endif

#ifndef SCENE_H /* the reference's own scene.h already gives these names */
#define SCENE_H

enum light_type { AMBIENT, DIRECTIONAL, POINT, SPECULAR };   /* scene.h:9-14 */

struct light {                               /* scene.h:16-23, 28 bytes */
  enum light_type type;
  float r, g, b;
  vector3 v;                                 /* direction (d_light) or position (p_light) */
};

struct camera {                              /* scene.h:25-32, 48 bytes */
  int width, height;                         /* already multiplied by `aliasing` at render() */
  vector3 position, u, v;
  float fov;                                 /* degrees */
};

struct triangles_layout {                    /* scene.h:98-107 (SOA) */
  vector3 *vertex;                           /* 3 consecutive vector3 per triangle */
  vector3 *normal;
};

struct scene_objects_additional_data {       /* scene.h:109-114 (SOA) */
  vector3 *vertex;
  vector3 *normal;
};

struct object {                              /* scene.h:119-134, 72 bytes */
  struct triangles_layout triangles;
  uint32_t triangle_count;
  vector3 ka, kd, ks;
  float ns, ni, nr, d;
};

struct AABB;
struct octree;

struct scene {                               /* scene.h:140-170, 120 bytes */
  struct scene_objects_additional_data objects_data;
  struct object *objects;
  size_t object_count;
  struct light *lights;
  size_t light_count;
  struct camera camera;
  size_t triangle_count;
  struct AABB *aabbs;                        /* unused by this library (null on host) */
  struct octree *octree;                     /* unused by this library (null on host) */
};

#endif /* SCENE_H */

/* ---- errors ---- */

#define RTB_OK            0
#define RTB_ERR_ARG      -1   /* bad argument / inconsistent scene */
#define RTB_ERR_CUDA     -2   /* CUDA runtime failure (see rtb_last_error) */
#define RTB_ERR_NOGPU    -3   /* no usable sm_100 device: there is NO CPU fallback */
#define RTB_ERR_LIMIT    -4   /* scene exceeds a format limit */

/* Thread-local description of the last failure (replaces the spdlog::error +
 * exit(1) of gpu/raytracer.cu:21-29; the C++ shim / CLI turns it back into that). */
const char *rtb_last_error(void);

/* ---- one-shot render: the drop-in for
 *      void render(const scene&, char*, int, std::ptrdiff_t, int, int)
 *      gpu/headers/raytracer.h:6, gpu/raytracer.cu:177-253.
 * `scene` holds HOST pointers; camera.width/height are pre-multiplied by `aliasing`
 * (gpu/rt.cpp:78-79).  Writes out_h rows of out_w RGBA8 pixels, `stride` bytes apart,
 * row 0 = top of the image (gpu/rt.cpp:44-47). */
int rtb_render(const struct scene *scene, unsigned char *rgba_out, int aliasing,
               ptrdiff_t stride, int out_h, int out_w);

/* The same call on `n_devices` GPUs of this process (devices 0 .. n-1): the scene is uploaded
 * and the hierarchy built on every device, each renders the 32x32-pixel tiles it owns
 * ((Tx + 3*Ty) mod n) and stores them into device 0's frame through peer access; device 0
 * downloads the frame.  Same bytes as rtb_render.  `rt` uses it when RT_GPUS > 1. */
int rtb_render_multi(const struct scene *scene, unsigned char *rgba_out, int aliasing,
                     ptrdiff_t stride_bytes, int out_h, int out_w, int n_devices);

/* ---- persistent scene: replaces to_cuda() (gpu/scene.cu:224-352) and the whole
 *      create_octree() host pipeline (gpu/partitioning/octree.cu:362-411):
 *      upload, triangle bounds, sort keys (layout chosen per scene), radix sort, node emit,
 *      clusters rebuilt with surface-area splits. ---- */

typedef struct rtb_scene_handle rtb_scene_t;

/* flags for rtb_scene_create */
#define RTB_BUILD_DEFAULT     0u
#define RTB_BUILD_NO_TREE     1u   /* skip the hierarchy (brute-force tracer only) */
#define RTB_BUILD_CUBIC_KEYS  2u   /* keep the x,y,z-interleaved sort keys: no choice between key layouts */
#define RTB_BUILD_RADIX_ONLY  4u   /* plain Morton radix tree: no surface-area rebuild of the clusters */
#define RTB_BUILD_NO_LIGHT_GRIDS 16u /* shadow rays always walk the tree (no per-light projection grids) */
#define RTB_BUILD_ONE_FRAME   8u   /* the scene will be rendered once (rtb_render passes this): scenes of more
                                      than 2^18 triangles keep the radix tree; rebuilding the clusters of
                                      10 M triangles takes 6 ms and saves 1.3 ms per 4K frame; light grids
                                      only from 4 samples per triangle on (camera.width * height >= 4 T) */

int rtb_scene_create(const struct scene *scene, int device, unsigned flags,
                     rtb_scene_t **out);
void rtb_scene_destroy(rtb_scene_t *h);

/* Replace the camera of a resident scene (fly-through).
 * width/height pre-multiplied by aliasing, as in render(). */
int rtb_scene_set_camera(rtb_scene_t *h, const struct camera *cam);

/* Tracer selection (tests and benches; the default is the hierarchy tracer). */
#define RTB_TRACER_TREE        0   /* octree traversal (product path) */
#define RTB_TRACER_BRUTE       1   /* every triangle for every ray: semantic definition,
                                      gpu/hit.cu:84-114 (PARTITIONING_NONE) */

#define RTB_KERNEL_PER_SAMPLE  0
#define RTB_KERNEL_WAVEFRONT   1

/* Primary rays all pass through the camera position, so the rays the reference's rounded ray/triangle test
 * can accept for one triangle are a region of the screen bounded by three offset lines (csrc/region.cuh: a
 * bound derived from the operation sequence of gpu/hit.cu:20-37, not a padded box).  RASTER runs the per-ray
 * test on exactly those samples (same test, same winner as the exhaustive search of gpu/hit.cu:84-114) and
 * falls back to the tree by itself when the regions cover too much or the item list overflows;
 * AUTO = RASTER for every scene, except for a camera whose last frame fell back. */
#define RTB_PRIMARY_AUTO    0
#define RTB_PRIMARY_TREE    1
#define RTB_PRIMARY_RASTER  2

typedef struct rtb_frame_opts {
  int aliasing;        /* SSAA factor, 3 in gpu/rt.cpp:67 */
  int tracer;          /* RTB_TRACER_* */
  int rank, nranks;    /* interleaved-tile split: this call renders only tiles owned by
                          `rank` of `nranks`; 0,1 = whole frame */
  int count_rays;      /* 1: maintain the device ray counters (slower; untimed passes) */
  int frame_kernel;    /* RTB_KERNEL_*: 0 = persistent-warp per-sample kernel (default, fastest measured);
                          1 = staged wavefront kernel with in-warp ray queues (tree tracer, <= 32
                          shadow-casting lights, aliasing <= 6; same bytes out) */
  int primary;         /* RTB_PRIMARY_*: how the first ray of every sample finds its triangle */
  int reserved;
} rtb_frame_opts;

/* Render into a HOST buffer (H2D nothing, D2H the frame): timed e2e path. */
int rtb_scene_render(rtb_scene_t *h, const rtb_frame_opts *opts,
                     unsigned char *rgba_out, ptrdiff_t stride, int out_h, int out_w);

/* Render into a DEVICE buffer (cudaMalloc'ed or peer-mapped memory, e.g. rank 0's
 * framebuffer opened over CUDA IPC): only pixels of tiles owned by opts->rank are
 * written.  `stream` is a cudaStream_t (0 = the library's own stream).  Asynchronous:
 * call rtb_scene_sync() or synchronise the stream.  A caller's stream is ordered behind the
 * handle's earlier work, and the handle's own stream waits for the frame: rtb_scene_sync(),
 * rtb_scene_stats() and the next frame see it finished whichever stream it ran on. */
int rtb_scene_render_device(rtb_scene_t *h, const rtb_frame_opts *opts,
                            void *d_rgba, ptrdiff_t pitch, int out_h, int out_w,
                            void *stream);
int rtb_scene_sync(rtb_scene_t *h);

/* Parity extension (not used by `rt`): per-SAMPLE primary-ray nearest-hit IDs on the
 * aliasing*out_w x aliasing*out_h sample grid, indexed [py*W' + px] with (px,py) the
 * kernel's sample coordinates of gpu/raytracer.cu:91-92.  id = global triangle index
 * in host memory order (objects in file order), -1 = miss.  obj_out may be NULL. */
int rtb_scene_hit_ids(rtb_scene_t *h, const rtb_frame_opts *opts,
                      int32_t *tri_out, int32_t *obj_out, int out_h, int out_w);

/* Closest-hit / any-hit queries for arbitrary rays (host arrays of n rays, 6 floats
 * each: origin, direction).  Mirrors collide()/collide_dist() (gpu/headers/hit.h:7-8).
 * tri_out[i] = winning triangle or -1; dist_out[i] = new_dist (0 on miss);
 * with any_hit != 0 only "occluded or not" is meaningful (tri_out = 0 / -1). */
int rtb_scene_trace_rays(rtb_scene_t *h, int tracer, int any_hit, const float *rays,
                         size_t n, int32_t *tri_out, float *dist_out);

typedef struct rtb_stats {
  uint64_t triangles, objects, lights;
  uint64_t nodes, leaves, max_depth, max_leaf_tris;
  uint64_t tree_bytes, triangle_bytes;
  double upload_ms, build_ms;          /* last rtb_scene_create */
  double frame_ms;                     /* last render (device time, CUDA events) */
  uint64_t rays_primary, rays_shadow, rays_bounce;   /* last render with count_rays */
  uint64_t node_visits, tri_tests;                    /* idem */
  uint64_t kernel_launches;                           /* kernels launched by last render */
  uint64_t build_launches;             /* kernels launched by the last rtb_scene_create */
  uint64_t key_mode;                   /* sort-key layout of the hierarchy: 0 = x,y,z interleaved,
                                          1 / 2 / 3 = axis x / y / z last (last rtb_scene_create) */
  uint64_t mode_cost;                  /* sample-tree surface-area cost of modes 1, 2, 3 relative to
                                          mode 0 in 1/1000, 16 bits each (bits 0-15: mode 1);
                                          0 = the choice was not evaluated */
  uint64_t builder;                    /* 0 = Morton radix tree, 1 = binned surface-area splits */
} rtb_stats;

int rtb_scene_stats(rtb_scene_t *h, rtb_stats *out);

/* Sharded projection pass of multi-rank frames (no reference counterpart: the reference is
 * single-GPU).  By default every rank of a frame projects ALL triangles to find the items of its own tiles.
 * With an exchange attached, rank r projects only the triangles [r T / N, (r + 1) T / N) and stores each item
 * straight into the inbox of the rank that owns its tile -- peer-mapped device memory (NVLink stores inside the
 * projection kernel, no collective), followed by a device-side epoch barrier.  Protocol: every rank calls
 * rtb_scene_exchange_create (allocates its block: control words + inbox), the blocks' addresses are made
 * visible to the other ranks (rtb_ipc_export / rtb_ipc_open between processes; the plain pointers inside one
 * process with peer access), then rtb_scene_exchange_attach(blocks[0..N-1], own entry ignored).  From then on
 * rtb_scene_render_device with this (rank, nranks) runs sharded; ALL ranks must render the same sequence of
 * such frames (the barrier counts them), each on its own stream / process.  A rank that does not show up
 * within ~3 s makes the others fall back to the tree walk for that frame (same pixels, slower). */
int rtb_scene_exchange_create(rtb_scene_t *h, int nranks, void **d_block, size_t *bytes);
int rtb_scene_exchange_attach(rtb_scene_t *h, int rank, int nranks, void *const *blocks);

/* Measurement aid (bench.py): with profiling on, every frame records CUDA events between its
 * kernels; rtb_scene_kernel_ms waits for the last frame and returns, in milliseconds, out[0] clears +
 * k_rast_setup + k_rast_big + k_xchg_sync, out[1] k_rast_items, out[2] the frame kernel, out[3] the whole
 * frame, and the parts of out[0]: out[4] clears + k_rast_setup, out[5] k_rast_big, out[6] k_xchg_sync (the
 * wait for the other ranks' items, sharded frames only); out[7] is 0. */
int rtb_scene_profile(rtb_scene_t *h, int on);
int rtb_scene_kernel_ms(rtb_scene_t *h, double out[8]);

/* State the projection pass of the last frame (or hit-id query) left: out[0] items,
 * out[1] != 0: it gave up (too many items, a triangle over too many tiles), out[2] | out[3] << 32
 * samples the items cover (the frame walks the tree when this exceeds 20 x its samples),
 * out[4] triangles whose accept region has two parts (region.cuh; all zero when no pass ran);
 * out[5] lights with a shadow-ray grid, out[6] their list entries / 1024, out[7] their cells / 1024. */
int rtb_scene_raster_state(rtb_scene_t *h, unsigned out[8]);

/* ---- build-pipeline primitives on HOST arrays (upload, run the device kernel,
 *      download) so the reference's own known-answer vectors can be replayed:
 *      parallel_radix_sort  gpu/partitioning/sort.h:21-22 (test.cu:308-340),
 *      shared_prefix_sum    gpu/partitioning/prefix_sum.h:13 (test.cu:288-305). ---- */
int rtb_radix_sort_pairs_u32(uint32_t *keys, uint32_t *vals, size_t n);
int rtb_radix_sort_pairs_u64(uint64_t *keys, uint32_t *vals, size_t n);
int rtb_inclusive_scan_u32(uint32_t *data, size_t n);
int rtb_inclusive_scan_u64(uint64_t *data, size_t n);

/* Device-resident variants used by the build itself and by micro-benchmarks
 * (gpu/bench.cu:435-451).  Pointers are device pointers; `stream` a cudaStream_t.
 * Scratch comes from the stream-ordered memory pool on `stream`; the calls are asynchronous:
 * results are ready in stream order (synchronise the stream before reading them on the host). */
int rtb_dev_radix_sort_pairs_u64(uint64_t *d_keys, uint32_t *d_vals, size_t n,
                                 int key_bits, void *stream);
int rtb_dev_inclusive_scan_u32(uint32_t *d_data, size_t n, void *stream);

/* No counterpart in the reference: added for callers that
 * cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).
 * Same result as rtb_dev_inclusive_scan_u32, but the scan's tile sums live in a file-scope device
 * array (129 KB per GPU that loads the library, until exit) instead of pool scratch: the call
 * makes no cudaMalloc / cudaFree / pool allocation.  n <= RTB_SCAN_STATIC_MAX, else RTB_ERR_LIMIT.
 * The current device must be d_data's.  All calls share the array: calls that overlap on the GPU
 * (two streams, two threads) race -- one call at a time per device. */
#define RTB_SCAN_STATIC_MAX ((size_t)1 << 26)
int rtb_dev_inclusive_scan_u32_static(uint32_t *d_data, size_t n, void *stream);

/* Dump the flattened hierarchy for structural tests: nodes (64 B each) and the
 * leaf-ordered triangle permutation.  Pass NULL buffers to query counts. */
int rtb_scene_dump_tree(rtb_scene_t *h, void *nodes_out, size_t *n_nodes,
                        uint32_t *perm_out, size_t *n_tris);

/* Index of the root node inside the array rtb_scene_dump_tree returns. */
int rtb_scene_root(rtb_scene_t *h);

/* Bracket timer on the library stream: CUDA events around everything queued between
 * start and stop (bench: K frames back to back).  stop synchronises and returns ms. */
int rtb_scene_timer_start(rtb_scene_t *h);
int rtb_scene_timer_stop(rtb_scene_t *h, double *ms_out);

/* Box-inflation factor of the conservative slab test, in float spacings at
 * |ray origin| + |scene| (DESIGN.md "conservative culling").  Tests and probes only. */
void rtb_set_margin_k(float k);

/* Longest-first scheduling of the frame's work items from the previous frame's measured
 * costs (resident scenes, from the third frame of a handle on).  On by default; 0 turns it
 * off for A/B measurements.  The frame's bytes do not depend on it. */
void rtb_set_schedule(int on);

/* Starts the CUDA driver and the device context and returns; no other effect.  A fresh
 * process pays 0.2 - 1 s for this at its first CUDA call: a loader can spend it on a side
 * thread while it reads the scene (host/parser.cpp does). */
void rtb_warmup(void);

/* Page-locked host memory (a frame buffer the D2H copy can target directly) and plain
 * device buffers / CUDA-IPC handles for the multi-process tile gather: rank 0 exports
 * its frame, the other ranks open it and pass the mapped pointer to
 * rtb_scene_render_device, so their tiles travel as NVLink peer stores. */
void *rtb_host_alloc(size_t bytes);
void rtb_host_free(void *p);
/* Allocator behind empty_scene/add_object_to_scene (gpu/scene.cu:146-218): vertex and normal
 * arrays of 1 MiB or more are page-locked when a device is present, so render()'s upload is a
 * direct DMA; falls back to malloc.  rtb_scene_mem_free accepts either kind. */
void *rtb_scene_mem_alloc(size_t bytes);
void rtb_scene_mem_free(void *p);
int rtb_dev_alloc(int device, size_t bytes, void **out);
int rtb_dev_free(void *p);
int rtb_dev_memset(void *p, int value, size_t bytes);
int rtb_dev_download(void *dst_host, const void *src_dev, size_t bytes);
/* Double-buffered frames: start the download of a finished frame on the device's copy stream (page-locked
 * dst_host, e.g. rtb_host_alloc) and go on rendering into the other buffer; rtb_dev_download_wait returns
 * when every download started from that device has arrived.  The caller orders copy after frame
 * (rtb_scene_sync, a barrier between ranks).  (gpu/raytracer.cu:244-247 is the reference's one cudaMemcpy2D.) */
int rtb_dev_download_async(void *dst_host, const void *src_dev, size_t bytes);
int rtb_dev_download_wait(const void *src_dev);
int rtb_ipc_export(void *d_ptr, unsigned char handle_out[64]);
int rtb_ipc_open(const unsigned char handle[64], void **out);
int rtb_ipc_close(void *p);

/* Library / device info. */
int rtb_device_count(void);
const char *rtb_version(void);

#ifdef __cplusplus
} /* extern "C" */
#endif

/* ---- layout pins (SURVEY Appendix D probe of the reference headers) ---- */
#if defined(__cplusplus)
static_assert(sizeof(vector3) == 12, "vector3");
static_assert(sizeof(struct light) == 28 && offsetof(struct light, v) == 16, "light");
static_assert(sizeof(struct camera) == 48 && offsetof(struct camera, fov) == 44, "camera");
static_assert(sizeof(struct object) == 72 && offsetof(struct object, triangle_count) == 16 &&
              offsetof(struct object, ka) == 20 && offsetof(struct object, nr) == 64, "object");
static_assert(sizeof(struct scene) == 120 && offsetof(struct scene, objects) == 16 &&
              offsetof(struct scene, camera) == 48 &&
              offsetof(struct scene, triangle_count) == 96, "scene");
#endif

#endif /* RTB200_H */
