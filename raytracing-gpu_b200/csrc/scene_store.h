// The resident scene behind rtb_scene_t: every device allocation the library owns.
#pragma once
#include <condition_variable>
#include <mutex>
#include "rtb_internal.h"

namespace rtb {

struct SceneStore {
  int device = -1;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;     // around the last frame kernel / upload
  cudaEvent_t xev = nullptr;                    // orders a caller's stream (rtb_scene_render_device) with the handle's own
  cudaEvent_t tev0 = nullptr, tev1 = nullptr;   // bracket timer (rtb_scene_timer_*)
  int sm_count = 148;
  bool profile = false;                         // rtb_scene_profile: events between the frame's kernels
  cudaEvent_t kev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};

  float4 *v0 = nullptr, *e1 = nullptr, *e2 = nullptr, *n0 = nullptr, *n1 = nullptr, *n2 = nullptr;
  float4 *nodes = nullptr, *materials = nullptr, *lights = nullptr;
  uint32_t *order = nullptr;            // slot -> global triangle id
  LightGrid *lgrids = nullptr;          // device array [n_lights] (lightgrid.cuh)
  void *lgrid_mem[24] = {nullptr};      // edge records / cell_start / entries of up to 8 grids
  size_t lgrid_mem_bytes[24] = {0};
  int n_lgrids = 0;
  unsigned long long lgrid_entries = 0, lgrid_cells = 0;
  float lgrid_margin_k = 0;             // ray-time margin the grids were built for
  uint32_t *inv_order = nullptr;        // global triangle id -> slot
  uint32_t *up = nullptr;               // [n_nodes] parent links, then [n_tris] slot -> leaf position (rtb_internal.h)
  uint32_t n_tris = 0, n_objects = 0, n_lights = 0;
  uint32_t n_nodes = 0, n_leaves = 0, max_leaf = 0;
  int root = 0, height = 0;
  int builder = 0;                      // 0: Morton radix tree, 1: binned surface-area splits (build.cu)
  int key_mode = 0;                     // sort-key layout the build chose (build.cu morton_key)
  int mode_cost[4] = {0, 0, 0, 0};      // surface-area cost of the four sample trees (0: not evaluated)
  bool has_tree = false;
  // RTB_BUILD_DEFER_TREE (rtb_render only): sort, slot order and light grids are built, the node array is not; a
  // frame whose rays turn out to need it says so in rast[7] and rtb_render builds the scene again with the tree
  bool tree_deferred = false;
  StrayQuery *d_stray = nullptr;       // deferred tree: STRAY_CAP listed shadow rays, then STRAY_CAP work-item ids (rtb_internal.h)
  int stray_pass = 0;                  // 2: the frame kernel alone on stray_items listed tiles, answers looked up
  unsigned stray_items = 0;
  float abs_max = 0;
  float pad_unit = 0;                   // needle-padding unit of the leaf boxes (pad.cuh)
  float origin_radius = 0;              // ray-origin radius the sliver padding was built for (build.cu)
  struct camera camera;                 // current camera (sample-grid resolution)

  // frame resources (grown on demand)
  unsigned char *d_frame = nullptr;
  size_t d_frame_bytes = 0;
  unsigned char *h_frame = nullptr;     // pinned staging for the D2H copy
  size_t h_frame_bytes = 0;
  unsigned int *d_tile_counter = nullptr;
  unsigned long long *d_counters = nullptr;
  // primary visibility by projection (raster.cuh), grown on demand by api.cu
  unsigned long long *d_prim = nullptr;
  size_t d_prim_count = 0;
  RastItem *d_rast_items = nullptr;     // the local item segment of frames that are not sharded
  size_t d_rast_cap = 0;
  // item exchange of sharded multi-rank frames (raster.cuh, rtb_scene_exchange_*): this rank's block =
  // control words + inbox (2 parities x nranks segments), and the peers' blocks as mapped here
  unsigned char *xchg_block = nullptr;
  size_t xchg_bytes = 0;
  unsigned xchg_cap = 0;                // items per segment
  int xchg_nranks = 0, xchg_rank = -1;  // attached: the frame layout the peers' pointers are for
  unsigned char *xchg_peer[XCHG_MAX_RANKS] = {nullptr};
  unsigned xchg_epoch = 0;
  unsigned *d_rast = nullptr;
  unsigned *h_rast = nullptr;           // pinned: state the last projection pass left (read back with the frame)
  bool one_frame = false;               // built with RTB_BUILD_ONE_FRAME
  bool rast_pending = false;            // h_rast is being filled by the stream
  unsigned long long rast_key[4] = {0, 0, 0, 0};   // camera generation, samples, ranks, factor of the frame h_rast is for
  unsigned rast_area_factor = 20;
  bool rast_useless = false;            // that frame walked the tree although it ran the pass: do not try again
  uint2 *d_fine_acc = nullptr;          // fine work items of multi-rank frames (trace.cu k_frame<.., FINE>)
  unsigned int *d_fine_done = nullptr;
  size_t fine_pixels = 0, fine_items = 0;
  float *d_bounce = nullptr;            // wavefront kernel: reflected-ray scratch
  size_t d_bounce_floats = 0;
  int n_shadow_lights = 0;              // DIRECTIONAL + POINT lights
  unsigned char shadow_light[32] = {0};

  // longest-first scheduling of the frame's work items (api.cu render_async)
  unsigned char *d_sched = nullptr;     // one block: cost | keys | keys alt | vals | vals alt | sort scratch
  size_t sched_items = 0;               // capacity, in items
  const unsigned int *d_item_order = nullptr;   // inside d_sched; nullptr = no order yet
  int sched_key[5] = {-1, -1, -1, -1, -1};      // rank, nranks, out_w, out_h, aliasing the order is for
  unsigned long long camera_gen = 0, order_camera_gen = ~0ull, frames_done = 0;

  double upload_ms = 0, build_ms = 0, frame_ms = 0;
  unsigned long long build_launches = 0, frame_launches = 0;
  unsigned long long counters[8] = {0, 0, 0, 0, 0, 0, 0, 0};

  ~SceneStore();
  void release();
  DevScene dev() const {
    DevScene d;
    d.v0 = v0; d.e1 = e1; d.e2 = e2; d.n0 = n0; d.n1 = n1; d.n2 = n2;
    d.nodes = nodes; d.node_up = up; d.slot_up = up ? up + n_nodes : nullptr; d.root = root; d.n_tris = n_tris; d.n_objects = n_objects; d.n_lights = n_lights;
    d.materials = materials; d.lights = lights; d.abs_max = abs_max; d.margin_k = get_margin_k();
    d.inv_order = inv_order; d.pad_unit = pad_unit;
    d.lgrids = (lgrids && get_margin_k() <= lgrid_margin_k) ? lgrids : nullptr;
    return d;
  }
};

// rtb_render_multi: the devices of one call share the scene upload.  Member 0 copies the vertex and normal
// arrays from the host into src_v / src_n (plain cudaMalloc memory on its device, which every peer has mapped
// for the frame anyway) and builds from them; the other members fetch them from there over NVLink, vertices
// first, normals on their side stream while the hierarchy is being built.  The host is read once whatever the
// number of devices: N devices each pulling the whole scene over PCIe queue up on the host's memory and made
// an 8-device render() three times slower than a 1-device one.
struct UploadGroup {
  int n = 0;
  int device[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  float *src_v = nullptr, *src_n = nullptr;           // 9 floats per triangle each, on device[0]
  cudaEvent_t ready_v = nullptr, ready_n = nullptr;   // recorded by member 0 behind its host copies
  std::mutex mu;
  std::condition_variable cv;
  int arrived = 0;
  unsigned phase = 0;
  bool failed = false;
  // all members arrive (true), or one of them has failed (false, for everybody, from then on)
  bool barrier() {
    std::unique_lock<std::mutex> lk(mu);
    if (failed) return false;
    const unsigned my = phase;
    if (++arrived == n) { arrived = 0; ++phase; cv.notify_all(); return true; }
    cv.wait(lk, [&] { return phase != my || failed; });
    return !failed;
  }
  void fail() {
    std::lock_guard<std::mutex> lk(mu);
    failed = true;
    cv.notify_all();
  }
};

constexpr unsigned RTB_BUILD_DEFER_TREE = 1u << 16;      // internal (rtb_render): see SceneStore::tree_deferred
constexpr int RTB_NEED_TREE = -100;                       // internal return code: this frame cannot run on a deferred tree
int build_scene(const struct scene *hs, unsigned flags, SceneStore *S, UploadGroup *grp = nullptr, int member = 0);
// per-device cache of stream, events and frame buffers (api.cu)
int acquire_frame_ctx(int device, SceneStore &S);
void recycle_frame_ctx(SceneStore &S);
float camera_origin_radius(const struct camera &cm);

}  // namespace rtb
