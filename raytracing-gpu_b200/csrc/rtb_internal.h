// Internal declarations shared by the translation units of librtb200.so.
// Nothing here is part of the ABI (include/rtb200.h is).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include "rtb200.h"
#include "region.cuh"

namespace rtb {

// ---- error plumbing (rtb_last_error) ----
void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define RTB_CUDA(call)                                                        \
  do {                                                                        \
    cudaError_t e__ = (call);                                                 \
    if (e__ != cudaSuccess) return ::rtb::cuda_fail(e__, #call, __FILE__, __LINE__); \
  } while (0)

// ---- hierarchy format ----
// One node = 64 bytes = 4 float4 (two children with full float boxes):
//   q0 = (c0.min.x, c0.max.x, c0.min.y, c0.max.y)
//   q1 = (c1.min.x, c1.max.x, c1.min.y, c1.max.y)
//   q2 = (c0.min.z, c0.max.z, c1.min.z, c1.max.z)
//   q3 = (as_float(ref0), as_float(ref1), as_float(first slot of range), as_float(slot count))
// ref >= 0: index of an inner node; ref < 0: leaf, ~ref = (first_slot << 3) | count, count <= LEAF_MAX (3).
// An empty child is a leaf with count 0 and an inverted box (+inf, -inf).
// Variants that were measured and dropped (4-wide nodes, centre/half-extent boxes, per-octant
// traversal, shadow rays that climb from their leaf) are in tools/ab_variants/ with their numbers in DESIGN.md section 8.
#ifndef RTB_LEAF_MAX
#define RTB_LEAF_MAX 3   /* C5, speculative traversal: 2: 43.4  3: 43.3  4: 43.9  6: 45.0 ms */
#endif
constexpr int LEAF_MAX = RTB_LEAF_MAX;
constexpr int HEIGHT_MAX = 96;       // binary inner-node depth limit (63 Morton bits + 32 index tie-break bits)
constexpr int NODE_QUADS = 4;
constexpr int STACK_MAX = HEIGHT_MAX;
constexpr int MORTON_BITS = 63;      // 21 bits per axis

// A light's shadow-ray grid (lightgrid.cuh)
constexpr int LG_LARGE_MAX = 16;
struct LightGrid {
  int kind;                     // 0: none (shadow rays walk the tree), 1: directional light, 2: point light
  float3 org, ax, ay, az;       // az: the light's direction (1) / the axis from the light towards the scene (2)
  float x0, y0, inv_cell, cell; // cell (i, j) = (x0, y0) + [i, i + 1) x [j, j + 1) * cell
  int nx, ny;
  float3 bmin, bmax;            // the scene's box (a hair inflated): the regions bound |origin - vertex| by its diagonal
  const uint32_t *cell_start;   // [nx * ny + 1]
  const uint2 *entries;         // (triangle slot, its region's rectangle inside the cell), cell by cell
  const float4 *rec;            // 4 per triangle slot: region record (3) + bounding box in this projection (lightgrid.cuh)
  uint32_t n_large;             // triangles every ray tests: their regions cover too many cells
  uint32_t large[LG_LARGE_MAX];
};

struct DevScene {
  // triangles in leaf (Morton) order, three float4 streams (SoA):
  //   v0 = (vertex0.xyz, as_float(global triangle id in host order))
  //   e1 = (vertex1 - vertex0, as_float(object index))
  //   e2 = (vertex2 - vertex0, 0)
  const float4 *v0, *e1, *e2;
  // unit vertex normals (normalize() applied once at build: gpu/hit.cu:15-17), same order
  const float4 *n0, *n1, *n2;
  const float4 *nodes;      // 4 float4 per node
  // upward links (binary layout only, else nullptr): node_up[k] = 2 * parent + (which child k is),
  // 0xffffffff for the root; slot_up[s] = the same for the leaf that holds triangle slot s.
  // Rays that start ON a triangle (shadow rays) climb from its leaf instead of descending from the root.
  const uint32_t *node_up, *slot_up;
  int root;                 // index of the root node
  uint32_t n_tris;
  uint32_t n_objects, n_lights;
  const float4 *materials;  // 3 float4 per object: (ka_col, kd_col, ks_col as packed RGBA8 bits; ns) (nr,0,0,0) ...
  const float4 *lights;     // 2 float4 per light: (type, packed colour, 0, 0) (v.xyz, 0)
  const LightGrid *lgrids;  // [n_lights] shadow-ray grids, nullptr: none
  const uint32_t *inv_order; // global triangle id -> slot (raster.cuh), nullptr when the scene has no hierarchy
  float pad_unit;           // needle-padding unit the leaf boxes were built with (pad.cuh)
  float abs_max;            // largest |coordinate| of the scene bounds (conservative-margin scale)
  float margin_k;           // box inflation, in float spacings at abs_max + |origin| (trace.cu)
};

struct DevCamera {
  float3 u, v, C, pos;      // gpu/raytracer.cu:178-182 (host part of render())
  int W, H;                 // sample grid (already multiplied by aliasing)
  float3 w;                 // u x v: C = pos + w L
  float L;                  // distance of the image plane (raster.cuh)
};

// Item exchange of the projection pass between the ranks of a frame (raster.cuh).  Non-sharded (one
// rank, or every rank projecting all triangles): one local segment.  Sharded: rank r projects the
// triangles [r T / N, (r + 1) T / N) and writes each item straight into the inbox of the rank that
// owns its tile -- peer-mapped memory, NVLink stores, no collective -- then publishes its counts and
// its epoch to every peer and waits for theirs (k_xchg_sync).
constexpr int XCHG_MAX_RANKS = 8;
// control block of a rank, 32-bit words: READY[src] epoch that source has published; per parity p (epoch & 1):
// COUNT[p][src] items in segment src, GAVE[p][src] that source gave up, AREA[p][src] (2 words) its work estimate
constexpr int XCHG_READY = 0, XCHG_COUNT = 8, XCHG_GAVE = 24, XCHG_AREA = 40, XCHG_CTL_WORDS = 128;
struct RastItem { uint4 hdr; float4 r0, r1, r2; };      // (slot, x0 | y0 << 16, x1 | y1 << 16, -) + the triangle's edge record
struct XchgView {
  int sharded;                          // 0: items stay on this rank (segment 0 only)
  unsigned cap;                         // items per segment
  unsigned epoch;                       // frame number of the exchange (all ranks count alike), parity = double buffer
  RastItem *seg[XCHG_MAX_RANKS];        // seg[q]: this rank's segment in rank q's inbox (remote unless q == rank)
  unsigned *ctl[XCHG_MAX_RANKS];        // ctl[q]: rank q's control block
  const RastItem *in[XCHG_MAX_RANKS];   // this rank's inbox: the segment of source s (local memory)
};

struct StrayQuery {       // 48 bytes
  float ox, oy, oz, dx, dy, dz;
  unsigned sample;        // px | py << 16
  unsigned light;
  unsigned item;          // the warp tile (work item) the sample belongs to
  unsigned occluded;      // written by k_stray_any
  unsigned pad[2];
};
constexpr unsigned STRAY_CAP = 1024;

struct FrameArgs {
  DevCamera cam;
  int aliasing;
  int out_w, out_h;
  int rank, nranks;
  unsigned char *out;       // device (or peer-mapped) RGBA8, row 0 = top
  long long pitch;
  unsigned int *tile_counter;
  // work-item scheduling (trace.cu frame_item): order to take the items in (nullptr = 0,1,2,..)
  // and where to leave each item's cost, in 256-cycle units (nullptr = not recorded)
  const unsigned int *item_order;
  unsigned int *item_cost;        // per item, 256-cycle units (fine items: summed over the item's rounds)
  // hot tiles as fine work items (k_frame<.., FINE>): per output pixel the channel sums of its samples, per
  // item the number of finished rounds (both zero between frames: the resolving warp clears them), and how
  // many leading entries of item_order are hot; nullptr: one warp does all rounds of every item
  uint2 *fine_acc;
  unsigned int *fine_done;
  const unsigned int *fine_hot;
  // deferred tree (SceneStore::tree_deferred): a ray that would have to walk the hierarchy sets *need_tree instead
  // (its sample is then wrong: the caller builds the tree and renders again); nullptr: the scene has its tree
  unsigned int *need_tree;
  // ... except shadow rays that are merely off their light's grid (a hit point outside the scene's box: the rounded test
  // accepts on noise far beside a triangle seen edge-on; C5: 26 of 125 M).  They are listed (stray_mode 1, count in
  // need_tree[14] = rast[21]), resolved against every triangle by k_stray_any, and the warp tiles they belong to are
  // rendered once more with the answers looked up (stray_mode 2, only_items tiles through item_order).
  StrayQuery *stray;
  unsigned stray_cap;
  int stray_mode;
  unsigned only_items;
  unsigned long long *counters;   // [8]: primary, shadow, bounce, node visits, tri tests
  // primary visibility by projection (raster.cuh); prim == nullptr: every ray walks the tree
  unsigned long long *prim;       // [H * W] new_dist bits << 32 | triangle id of the nearest primary candidate, ~0: none
  XchgView x;                     // where the items go and come from
  unsigned *rast;                 // counters and flags (raster.cuh RAST_WORDS)
  unsigned *rast_big;             // [RAST_BIG_CAP] slots of triangles whose items k_rast_big emits
  region::PrimaryCull *rast_big_cull;   // [RAST_BIG_CAP] their regions as k_rast_setup computed them
  region::CamFrame camf;          // the camera in double precision + the rounding scales of this frame (region.cuh)
  unsigned rast_area_factor;      // give up when the rectangles cover more than this many times the frame
  // wavefront kernel only
  float *bounce_scratch;          // per resident warp: 7 x S floats (origin, direction, nr_ of reflected rays)
  int n_shadow_lights;            // DIRECTIONAL + POINT lights (<= 32 for the wavefront kernel)
  unsigned char shadow_light[32]; // their indices in the light array, file order
};

constexpr int WAVE_MAX_ALIASING = 6;      // shared-memory budget of the wavefront kernel
size_t wave_scratch_floats(int aliasing, int sm_count);   // size of FrameArgs::bounce_scratch

// ---- device-side build (build.cu / sort.cu) ----
int dev_radix_sort_pairs_u64(uint64_t *d_keys, uint32_t *d_vals, uint64_t *d_keys_alt, uint32_t *d_vals_alt,
                             size_t n, int key_bits, uint32_t *d_scratch, size_t scratch_words,
                             cudaStream_t st, int *result_in_alt, unsigned long long *launches, int low_bit = 0);
size_t radix_sort_scratch_words(size_t n);
int dev_exclusive_scan_u32(uint32_t *d_data, size_t n, uint32_t *d_scratch, size_t scratch_words, bool inclusive,
                           cudaStream_t st, unsigned long long *launches);
int dev_inclusive_scan_u64(uint64_t *d_data, size_t n, uint64_t *d_scratch, size_t scratch_words, cudaStream_t st);
size_t scan_scratch_words(size_t n);

// ---- device blocks (api.cu) ----
// cudaMallocAsync / cudaFreeAsync with a per-device cache of big blocks in front.  The stream-ordered pool keeps
// freed memory (api.cu raises its release threshold), but mapping a block of hundreds of MB out of it costs the
// host 1-3 ms every time; a render() of a 10 M-triangle scene allocates 30 such blocks.  Freed blocks of >= 8 MB
// are parked with an event on the freeing stream and handed to the next request of exactly that size (the next
// render() of the same scene); RTB_BLOCK_CACHE_MB bounds what is parked per device (default 24 576, 0: off).
int block_alloc(void **p, size_t bytes, cudaStream_t st);
void block_free(void *p, cudaStream_t st);

// ---- frame kernels (trace.cu) ----
// marks: nullptr, or 6 events recorded before the frame's first operation, after k_rast_setup, after k_rast_big,
// after k_xchg_sync, after k_rast_items and after the frame kernel (rtb_scene_profile)
// est (one-shot frames: there is no previous frame whose costs could order the work items): scratch for an
// ESTIMATED cost per item -- the lengths of the light-grid lists at the hit points of the item's pixel centres, known as
// soon as the projection pass has run -- and for the sort that turns it into the item order of this very frame
struct ItemEstimate {
  unsigned *cost; unsigned n_items;
  uint64_t *sk, *sk2; uint32_t *sv, *sv2, *ss; size_t ssw;
  unsigned *hot;
};
int launch_frame(const DevScene &sc, const FrameArgs &fa, int tracer, bool count, int sm_count, cudaStream_t st,
                 unsigned long long *launches, cudaEvent_t *marks = nullptr, const ItemEstimate *est = nullptr);
unsigned frame_item_count(int out_w, int out_h, int nranks);   // work items of one rank's frame
// deferred tree: the listed shadow rays against every triangle; then the frame kernel alone (no projection pass) on fa.only_items tiles
int launch_stray_any(const DevScene &sc, StrayQuery *q, const unsigned *count, int sm_count, cudaStream_t st);
int launch_frame_only(const DevScene &sc, const FrameArgs &fa, int sm_count, cudaStream_t st);
int launch_item_order(const unsigned int *d_cost, unsigned n, uint64_t *d_keys, uint32_t *d_vals,
                      uint64_t *d_keys_alt, uint32_t *d_vals_alt, uint32_t *d_scratch, size_t scratch_words,
                      cudaStream_t st, int *result_in_alt, unsigned *d_hot, float per_share);
int launch_hit_ids(const DevScene &sc, const DevCamera &cam, int tracer, int32_t *d_tri, int32_t *d_obj,
                   cudaStream_t st);
int launch_prim_ids(const DevScene &sc, const FrameArgs &fa, int32_t *d_tri, int32_t *d_obj, int sm_count, cudaStream_t st);
constexpr int RAST_STATE_WORDS = 32;      // size of FrameArgs::rast, in 32-bit words
int launch_trace_rays(const DevScene &sc, int tracer, int any_hit, const float *d_rays, size_t n, int32_t *d_tri,
                      float *d_dist, cudaStream_t st);
struct SceneStore;
int build_light_grids(SceneStore &S, const struct light *lights, size_t n_lights, const float bmin[3],
                      const float bmax[3], cudaStream_t st);   // trace.cu, called at the end of build_scene
void set_margin_k(float k);   // conservative-box scale factor (tests only)
float get_margin_k();

}  // namespace rtb
