// Primary visibility by projection (included by trace.cu).
//
// Every primary ray passes through the camera position (gpu/raytracer.cu:97-105: origin on the
// image plane, direction towards `pos`), and sample (px, py) is a regular grid on that plane.
// So "which rays can the rounded Moeller-Trumbore test accept for this triangle" is a set of
// samples on the screen, and region.cuh bounds it rigorously: the numerators of u, v and 1 - u - v
// are linear in (px, py) (homogeneous edge functions through `pos`), the rounding of the
// reference's operation sequence moves them by at most E_k, so every accepted sample lies inside
// three offset lines (region::primary_lines has the derivation; tests/test_region_cpu.py replays
// it in float32 on the host, triangles seen edge-on, across the camera plane and degenerate ones
// included).  One pass over the triangles emits (triangle, rectangle) items and a per-triangle
// edge record, one warp per item tests the rectangle's sample centres against the record and
// queues the survivors, full warps run the exact per-ray test of tri_candidate on them, and
// atomicMin on (new_dist bits << 32 | triangle id) keeps the reference's winner: lexicographic
// minimum of (new_dist, id), gpu/hit.cu:66-76.  k_frame then starts every sample from that record
// instead of walking the tree for its first ray; bounce rays still walk the tree (shadow rays:
// lightgrid.cuh).
//
// The cull never drops a candidate of the reference's exhaustive search (gpu/hit.cu:84-114), so
// the primary hit ids are the reference's for every scene -- no boxes, no padding heuristics, no
// special list for triangles across the camera plane.  A triangle that spans more than
// RAST_TILES_MAX tiles, more items than the list holds or more rectangle area than
// RAST_AREA_FACTOR x the frame make the frame fall back to the tree for its primary rays (flag in
// device memory, no host round trip).
#pragma once

namespace {

constexpr int RAST_TILE = 32;                 // samples
constexpr unsigned RAST_TILES_MAX = 1u << 17;  // per triangle (a 4K x 3 frame has 73 000 tiles)
// rast[] words: 0 -, 1 fallback flag, 2-3 samples the items cover (u64), 4 triangles with two regions,
// 5 triangles on the big list, 6 exchange timed out, 8..15 items written for destination rank q (8: the
// only segment when the frame is not sharded)
constexpr int RAST_WORDS = 8;
constexpr unsigned RAST_BIG_TILES = 128;       // a rectangle over more tiles goes to k_rast_big (one CTA per triangle)
constexpr unsigned RAST_BIG_CAP = 1u << 16;    // slots on that list; the rest is emitted by its own thread

__device__ __forceinline__ bool rast_active(const FrameArgs &fa) {
  if (!fa.prim) return false;
  const unsigned *r = fa.rast;
  if (r[1]) return false;
  const unsigned long long area = *reinterpret_cast<const unsigned long long *>(r + 2);
  const unsigned long long samples = (unsigned long long)fa.cam.W * (unsigned long long)fa.cam.H / (unsigned)(fa.x.sharded ? 1 : fa.nranks);
  return area <= (unsigned long long)fa.rast_area_factor * samples + 65536ull;
}

// sample tile (tx, ty) belongs to this rank?  Rank tiles are RANK_TILE pixels = RANK_TILE * a samples wide.
__device__ __forceinline__ bool rast_owned(const FrameArgs &fa, int tx, int ty) {
  if (fa.nranks == 1) return true;
  const int per = RANK_TILE * fa.aliasing / RAST_TILE;      // sample tiles per rank tile edge (a: 32a / 32)
  const int Tx = tx / per, Ty = ty / per;
  return (Tx + 3 * Ty) % fa.nranks == fa.rank;
}

// destination segment of sample tile (tx, ty): the owning rank when the frame is sharded, else segment 0
// (only owned tiles are emitted then); -1: not emitted by this rank
__device__ __forceinline__ int rast_dest(const FrameArgs &fa, int tx, int ty) {
  if (fa.nranks == 1) return 0;
  const int per = RANK_TILE * fa.aliasing / RAST_TILE;
  const int owner = (tx / per + 3 * (ty / per)) % fa.nranks;
  if (fa.x.sharded) return owner;
  return owner == fa.rank ? 0 : -1;
}

__device__ __forceinline__ void rast_store(const FrameArgs &fa, int q, unsigned k, uint32_t s, int x0, int y0, int x1, int y1,
                                           const float *rv) {
  RastItem *it = fa.x.seg[q] + k;
  it->hdr = make_uint4(s, (unsigned)x0 | ((unsigned)y0 << 16), (unsigned)x1 | ((unsigned)y1 << 16), 0u);
  it->r0 = make_float4(rv[0], rv[1], rv[2], rv[3]);
  it->r1 = make_float4(rv[4], rv[5], rv[6], rv[7]);
  it->r2 = make_float4(rv[8], rv[9], rv[10], rv[11]);
}

// prim[] = ~0 ("no candidate") on the tiles this rank owns: a multi-rank frame never touches the rest
// (one warp per 32 x 32-sample tile row segment; a single rank uses cudaMemsetAsync on the whole array)
__global__ void __launch_bounds__(256) k_prim_clear(const FrameArgs fa) {
  const int tiles_x = (fa.cam.W + RAST_TILE - 1) / RAST_TILE, tiles_y = (fa.cam.H + RAST_TILE - 1) / RAST_TILE;
  const int lane = threadIdx.x & 31;
  const long long warps = (long long)gridDim.x * 8, w0 = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int per = RANK_TILE * fa.aliasing / RAST_TILE;
  for (long long t = w0; t < (long long)tiles_x * tiles_y; t += warps) {
    const int tx = (int)(t % tiles_x), ty = (int)(t / tiles_x);
    if ((tx / per + 3 * (ty / per)) % fa.nranks != fa.rank) continue;
    const int x = tx * RAST_TILE + lane;
    if (x >= fa.cam.W) continue;
    for (int r = 0; r < RAST_TILE; ++r) {
      const int y = ty * RAST_TILE + r;
      if (y < fa.cam.H) fa.prim[(size_t)y * fa.cam.W + x] = ~0ull;
    }
  }
}

// k_rast_setup: one thread per triangle slot.  Region of the screen where the rounded test can
// accept (region::primary_cull, double precision: ~150 flops per triangle, the kernel is bound by
// its gathers), the sample rectangle around it, the float edge record k_rast_items tests sample
// centres against, and the items: one if the rectangle is at most 32 x 32 samples inside one rank
// tile, else one per 32 x 32-sample tile the region can touch (region::row_range per tile row, so a
// long thin region costs its length, not its bounding box).
__global__ void __launch_bounds__(256) k_rast_setup(const DevScene sc, const FrameArgs fa) {
  // sharded frames: this rank's share of the triangle slots, blocks of 256 dealt round robin (slots are in
  // spatial order and the triangles next to the camera make most of the items: a compact share would leave
  // one rank with most of the writing)
  const uint32_t t1 = sc.n_tris;
  const uint32_t s = fa.x.sharded ? (blockIdx.x * (unsigned)fa.nranks + (unsigned)fa.rank) * 256u + threadIdx.x : blockIdx.x * 256u + threadIdx.x;
  const int lane = threadIdx.x & 31;
  unsigned *rast = fa.rast;
  region::PrimaryCull c;
  c.kind = 0;
  if (s < t1) {
    const float4 a4 = __ldg(sc.v0 + s), b4 = __ldg(sc.e1 + s), c4 = __ldg(sc.e2 + s);
    const float v0[3] = {a4.x, a4.y, a4.z}, e1[3] = {b4.x, b4.y, b4.z}, e2[3] = {c4.x, c4.y, c4.z};
    region::primary_cull(fa.camf, fa.cam.W, fa.cam.H, v0, e1, e2, c);
  }
  const bool any = c.kind != 0;
  const int x0 = c.x0, y0 = c.y0, x1 = c.x1, y1 = c.y1;
  const int ty0 = any ? y0 / RAST_TILE : 0, ty1 = any ? y1 / RAST_TILE : -1;
  const int nd = fa.x.sharded ? fa.nranks : 1;                              // destination segments
  unsigned cnt[XCHG_MAX_RANKS];                                             // items of this triangle per destination
#pragma unroll
  for (int q = 0; q < XCHG_MAX_RANKS; ++q) cnt[q] = 0;
  unsigned mine = 0, touched = 0, area = 0, big_k = 0;
  const int rt = fa.nranks == 1 ? (1 << 30) : RANK_TILE * fa.aliasing;      // rank tile edge in samples
  const bool single = any && x1 - x0 < RAST_TILE && y1 - y0 < RAST_TILE && x0 / rt == x1 / rt && y0 / rt == y1 / rt;
  const bool gave_up = *reinterpret_cast<volatile unsigned *>(rast + 1) != 0;     // the frame has given up: emit nothing more
  if (single) {
    const int q = gave_up ? -1 : rast_dest(fa, x0 / RAST_TILE, y0 / RAST_TILE);
    if (q >= 0) {
#pragma unroll
      for (int k = 0; k < XCHG_MAX_RANKS; ++k) cnt[k] = k == q ? 1u : 0u;
      mine = 1;
      area = (unsigned)((x1 - x0 + 1) * (y1 - y0 + 1));
    }
  } else if (any && !gave_up && (unsigned)(ty1 - ty0 + 1) * (unsigned)(x1 / RAST_TILE - x0 / RAST_TILE + 1) > RAST_BIG_TILES &&
             (big_k = atomicAdd(&rast[5], 1u)) < RAST_BIG_CAP) {
    fa.rast_big[big_k] = s;                     // k_rast_big emits its items, tile rows in parallel
    fa.rast_big_cull[big_k] = c;
  } else if (any && !gave_up) {
    // "area" = the samples k_rast_items will look at: it walks these items row by row over the
    // region's span (region::record_span), so a thin region costs its length, not its rectangle
    for (int ty = ty0; ty <= ty1 && mine <= RAST_TILES_MAX; ++ty) {
      int ta, tb;
      if (!region::primary_row_tiles(c, fa.cam.W, fa.cam.H, RAST_TILE, ty, ta, tb)) continue;
      touched += (unsigned)(tb - ta + 1);
      for (int tx = ta; tx <= tb; ++tx) {
        const int q = rast_dest(fa, tx, ty);
        if (q < 0) continue;
        ++mine;
#pragma unroll
        for (int k = 0; k < XCHG_MAX_RANKS; ++k) cnt[k] += k == q ? 1u : 0u;
      }
    }
    if (mine > RAST_TILES_MAX) {
      rast[1] = 1u; mine = 0;
#pragma unroll
      for (int k = 0; k < XCHG_MAX_RANKS; ++k) cnt[k] = 0;
    }
    if (mine) {
      const double rect = (double)(x1 - x0 + 1) * (double)(y1 - y0 + 1);
      const double work = fmin(c.R.work, rect) * (double)mine / (double)touched + 32.0 * (double)mine;
      area = (unsigned)fmin(work, 67108864.0);
    }
  }
  // item slots: one atomic per CTA and destination, one for the area, one for the two-part count (all of
  // rast[] is one line of one L2 slice: an atomic per warp made the slice's ALU the kernel's bottleneck)
  __shared__ unsigned s_w[8][XCHG_MAX_RANKS + 2], s_base[XCHG_MAX_RANKS], s_tot[XCHG_MAX_RANKS];
  const int warp = threadIdx.x >> 5;
  unsigned incl_q[XCHG_MAX_RANKS];
#pragma unroll
  for (int q = 0; q < XCHG_MAX_RANKS; ++q) {
    incl_q[q] = 0;
    if (q < nd) {
      unsigned incl = cnt[q];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      incl_q[q] = incl;
      if (lane == 31) s_w[warp][q] = incl;
    }
  }
  const unsigned wa = __reduce_add_sync(0xffffffffu, min(area, 1u << 26));
  const unsigned w2 = (unsigned)__popc(__ballot_sync(0xffffffffu, c.kind == 2));
  if (lane == 31) { s_w[warp][XCHG_MAX_RANKS] = wa; s_w[warp][XCHG_MAX_RANKS + 1] = w2; }
  __syncthreads();
  if ((int)threadIdx.x < nd) {
    const int q = (int)threadIdx.x;
    unsigned run = 0;
    for (int w = 0; w < 8; ++w) { const unsigned t = s_w[w][q]; s_w[w][q] = run; run += t; }
    s_tot[q] = run;
    s_base[q] = run ? atomicAdd(&rast[8 + q], run) : 0u;
  } else if (threadIdx.x == XCHG_MAX_RANKS) {
    unsigned long long sum = 0;
    for (int w = 0; w < 8; ++w) sum += s_w[w][XCHG_MAX_RANKS];
    if (sum) atomicAdd(reinterpret_cast<unsigned long long *>(rast + 2), sum);
  } else if (threadIdx.x == XCHG_MAX_RANKS + 1) {
    unsigned sum = 0;
    for (int w = 0; w < 8; ++w) sum += s_w[w][XCHG_MAX_RANKS + 1];
    if (sum) atomicAdd(&rast[4], sum);
  }
  __syncthreads();
  unsigned pos[XCHG_MAX_RANKS];
  bool overflow = false;
#pragma unroll
  for (int q = 0; q < XCHG_MAX_RANKS; ++q) {
    pos[q] = 0;
    if (q < nd) {
      pos[q] = s_base[q] + s_w[warp][q] + incl_q[q] - cnt[q];
      if (cnt[q] && s_base[q] + s_tot[q] > fa.x.cap) overflow = true;
    }
  }
  if (!mine) return;
  if (overflow) { rast[1] = 1u; return; }
  const float *rv = c.rec.v;
  if (single) {
#pragma unroll
    for (int q = 0; q < XCHG_MAX_RANKS; ++q)
      if (cnt[q]) rast_store(fa, q, pos[q], s, x0, y0, x1, y1, rv);
    return;
  }
  for (int ty = ty0; ty <= ty1; ++ty) {
    int ta, tb;
    if (!region::primary_row_tiles(c, fa.cam.W, fa.cam.H, RAST_TILE, ty, ta, tb)) continue;
    const int cy0 = max(y0, ty * RAST_TILE), cy1 = min(y1, ty * RAST_TILE + RAST_TILE - 1);
    for (int tx = ta; tx <= tb; ++tx) {
      const int q = rast_dest(fa, tx, ty);
      if (q < 0) continue;
      const int cx0 = max(x0, tx * RAST_TILE), cx1 = min(x1, tx * RAST_TILE + RAST_TILE - 1);
      unsigned k = 0;
#pragma unroll
      for (int d = 0; d < XCHG_MAX_RANKS; ++d)
        if (d == q) k = pos[d]++;
      rast_store(fa, q, k, s, cx0, cy0, cx1, cy1, rv);
    }
  }
}

// k_rast_big: the triangles k_rast_setup listed because their rectangles span many tiles (a road
// across the frame, a triangle seen edge-on whose region is a line across the screen).  One CTA
// per triangle: region::primary_row_tiles gives the tiles of each tile row the region can touch (one
// thread per row), a prefix sum over the rows numbers the tiles, and the CTA's threads take the tiles
// in that order -- a diagonal line (200 rows of 2 tiles) and a horizontal one (2 rows of 360) keep
// all lanes busy alike.  Every 32 tiles take their slots with one atomic per destination.
constexpr int RAST_BIG_THREADS = 128;
constexpr int RAST_BIG_ROWS = 512;              // tile rows per pass (a 4K x 3 frame has 203)
__global__ void __launch_bounds__(RAST_BIG_THREADS) k_rast_big(const DevScene sc, const FrameArgs fa) {
  __shared__ unsigned s_touched, s_mine, s_total, s_over, s_cnt[XCHG_MAX_RANKS], s_pos[XCHG_MAX_RANKS];
  __shared__ int s_ta[RAST_BIG_ROWS];
  __shared__ unsigned s_off[RAST_BIG_ROWS];
  unsigned *rast = fa.rast;
  const unsigned n_big = min(rast[5], RAST_BIG_CAP);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nd = fa.x.sharded ? fa.nranks : 1;
  for (unsigned bi = blockIdx.x; bi < n_big; bi += gridDim.x) {
    if (*reinterpret_cast<volatile unsigned *>(rast + 1)) return;
    const uint32_t s = fa.rast_big[bi];
    const region::PrimaryCull &c = fa.rast_big_cull[bi];                    // as k_rast_setup computed it
    if (!c.kind) continue;
    if (threadIdx.x == 0) { s_touched = 0; s_mine = 0; s_over = 0; }
    __syncthreads();
    const int ty0 = c.y0 / RAST_TILE, ty1 = c.y1 / RAST_TILE;
    unsigned touched = 0, mine = 0;
    for (int ty = ty0 + (int)threadIdx.x; ty <= ty1; ty += RAST_BIG_THREADS) {
      int ta, tb;
      if (!region::primary_row_tiles(c, fa.cam.W, fa.cam.H, RAST_TILE, ty, ta, tb)) continue;
      touched += (unsigned)(tb - ta + 1);
      if (fa.nranks == 1 || fa.x.sharded) mine += (unsigned)(tb - ta + 1);
      else for (int tx = ta; tx <= tb; ++tx) mine += rast_dest(fa, tx, ty) >= 0 ? 1u : 0u;
    }
    if (touched) atomicAdd(&s_touched, touched);
    if (mine) atomicAdd(&s_mine, mine);
    __syncthreads();
    const unsigned all_touched = s_touched, all_mine = s_mine;
    __syncthreads();
    if (!all_mine) continue;
    if (all_mine > RAST_TILES_MAX) { if (threadIdx.x == 0) rast[1] = 1u; return; }
    if (threadIdx.x == 0) {
      const double rect = (double)(c.x1 - c.x0 + 1) * (double)(c.y1 - c.y0 + 1);
      const double work = fmin(c.R.work, rect) * (double)all_mine / (double)all_touched + 32.0 * (double)all_mine;
      atomicAdd(reinterpret_cast<unsigned long long *>(rast + 2), (unsigned long long)fmin(work, 1.0e15));
    }
    for (int r0 = ty0; r0 <= ty1; r0 += RAST_BIG_ROWS) {
      const int nrows = min(RAST_BIG_ROWS, ty1 - r0 + 1);
      for (int i = (int)threadIdx.x; i < nrows; i += RAST_BIG_THREADS) {
        int ta = 0, tb = -1;
        const bool ok = region::primary_row_tiles(c, fa.cam.W, fa.cam.H, RAST_TILE, r0 + i, ta, tb);
        s_ta[i] = ta;
        s_off[i] = ok ? (unsigned)(tb - ta + 1) : 0u;
      }
      __syncthreads();
      if (warp == 0) {                          // exclusive prefix sum over the rows: 16 consecutive rows per lane
        constexpr int PER = RAST_BIG_ROWS / 32;
        unsigned v[PER], sum = 0;
#pragma unroll
        for (int k = 0; k < PER; ++k) { const int i = lane * PER + k; v[k] = i < nrows ? s_off[i] : 0u; sum += v[k]; }
        unsigned incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const unsigned o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
        unsigned base = incl - sum;
#pragma unroll
        for (int k = 0; k < PER; ++k) { const int i = lane * PER + k; if (i < nrows) s_off[i] = base; base += v[k]; }
        if (lane == 31) s_total = incl;
      }
      __syncthreads();
      const unsigned total = s_total;
      // one rank: every tile is emitted, so the pass takes its slots with a single atomic
      if (fa.nranks == 1) {
        if (threadIdx.x == 0) s_touched = atomicAdd(&rast[8], total);
        __syncthreads();
        const unsigned base = s_touched;
        if (base + total > fa.x.cap) { if (threadIdx.x == 0) rast[1] = 1u; return; }
        for (unsigned i = threadIdx.x; i < total; i += RAST_BIG_THREADS) {
          int lo = 0, hi = nrows - 1;
          while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_off[mid] <= i) lo = mid; else hi = mid - 1; }
          const int tx = s_ta[lo] + (int)(i - s_off[lo]), ty = r0 + lo;
          const int cx0 = max(c.x0, tx * RAST_TILE), cx1 = min(c.x1, tx * RAST_TILE + RAST_TILE - 1);
          const int cy0 = max(c.y0, ty * RAST_TILE), cy1 = min(c.y1, ty * RAST_TILE + RAST_TILE - 1);
          rast_store(fa, 0, base + i, s, cx0, cy0, cx1, cy1, c.rec.v);
        }
        __syncthreads();
        continue;
      }
      // several destinations (or only the owned tiles): count per destination, take the slots with one atomic
      // per destination, then hand them out through shared counters
      if (threadIdx.x < XCHG_MAX_RANKS) s_cnt[threadIdx.x] = 0u;
      __syncthreads();
      for (int pass = 0; pass < 2; ++pass) {
        for (unsigned i0 = 0; i0 < total; i0 += RAST_BIG_THREADS) {
          const unsigned i = i0 + threadIdx.x;
          int q = -1, tx = 0, ty = 0;
          if (i < total) {
            int lo = 0, hi = nrows - 1;            // the last row whose offset is <= i (rows without tiles share offsets)
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (s_off[mid] <= i) lo = mid; else hi = mid - 1; }
            tx = s_ta[lo] + (int)(i - s_off[lo]); ty = r0 + lo;
            q = rast_dest(fa, tx, ty);
          }
          for (int d = 0; d < nd; ++d) {
            const unsigned m = __ballot_sync(0xffffffffu, q == d);
            if (!m) continue;
            const int leader = __ffs(m) - 1;
            unsigned k = 0;
            if (lane == leader) k = atomicAdd(&s_cnt[d], (unsigned)__popc(m));
            if (pass == 0) continue;
            k = __shfl_sync(0xffffffffu, k, leader);
            if (q == d) {
              const int cx0 = max(c.x0, tx * RAST_TILE), cx1 = min(c.x1, tx * RAST_TILE + RAST_TILE - 1);
              const int cy0 = max(c.y0, ty * RAST_TILE), cy1 = min(c.y1, ty * RAST_TILE + RAST_TILE - 1);
              rast_store(fa, d, s_pos[d] + k + (unsigned)__popc(m & ((1u << lane) - 1u)), s, cx0, cy0, cx1, cy1, c.rec.v);
            }
          }
        }
        __syncthreads();
        if (pass == 0) {
          if ((int)threadIdx.x < nd) {
            const unsigned n = s_cnt[threadIdx.x];
            unsigned base = n ? atomicAdd(&rast[8 + threadIdx.x], n) : 0u;
            if (base + n > fa.x.cap) { rast[1] = 1u; s_over = 1u; }
            s_pos[threadIdx.x] = base;
            s_cnt[threadIdx.x] = 0u;
          }
          __syncthreads();
          if (s_over) return;
        }
      }
      __syncthreads();
    }
  }
}

// k_xchg_sync (sharded frames, one warp): publish to every rank how many items this rank wrote into its
// inbox, whether this rank gave up and its work estimate; then the epoch (release); then wait until every
// rank has published this epoch (acquire) and fold their verdicts into the local state, so that
// rast_active() decides alike on every rank.  The items themselves were written by k_rast_setup /
// k_rast_big before this kernel on the same stream.
__global__ void __launch_bounds__(32) k_xchg_sync(const FrameArgs fa) {
  unsigned *rast = fa.rast;
  const int lane = threadIdx.x, N = fa.nranks, p = (int)(fa.x.epoch & 1u);
  if (lane < N) {
    volatile unsigned *ctl = fa.x.ctl[lane];
    const unsigned n = rast[8 + lane];
    ctl[XCHG_COUNT + p * XCHG_MAX_RANKS + fa.rank] = min(n, fa.x.cap);
    ctl[XCHG_GAVE + p * XCHG_MAX_RANKS + fa.rank] = (rast[1] || n > fa.x.cap) ? 1u : 0u;
    ctl[XCHG_AREA + 2 * (p * XCHG_MAX_RANKS + fa.rank)] = rast[2];
    ctl[XCHG_AREA + 2 * (p * XCHG_MAX_RANKS + fa.rank) + 1] = rast[3];
    __threadfence_system();
    ctl[XCHG_READY + fa.rank] = fa.x.epoch;
  }
  __syncwarp();
  bool timed_out = false;
  if (lane < N) {
    volatile unsigned *mine = fa.x.ctl[fa.rank];
    const long long t0 = clock64();
    // epochs only grow; a rank that is a whole frame ahead has published a larger one
    while ((int)(mine[XCHG_READY + lane] - fa.x.epoch) < 0) {
      if (clock64() - t0 > 6000000000ll) { timed_out = true; break; }      // ~3 s: a rank died; this frame walks the tree
      __nanosleep(200);
    }
    __threadfence_system();
  }
  const unsigned late = __ballot_sync(0xffffffffu, timed_out);
  unsigned gave = 0;
  unsigned long long area = 0;
  if (lane < N) {
    volatile unsigned *mine = fa.x.ctl[fa.rank];
    gave = mine[XCHG_GAVE + p * XCHG_MAX_RANKS + lane];
    area = (unsigned long long)mine[XCHG_AREA + 2 * (p * XCHG_MAX_RANKS + lane)] |
           ((unsigned long long)mine[XCHG_AREA + 2 * (p * XCHG_MAX_RANKS + lane) + 1] << 32);
  }
  gave = __reduce_or_sync(0xffffffffu, gave);
#pragma unroll
  for (int d = 16; d; d >>= 1) area += __shfl_xor_sync(0xffffffffu, area, d);
  if (lane == 0) {
    if (gave || late) rast[1] = 1u;
    if (late) rast[6] = 1u;
    *reinterpret_cast<unsigned long long *>(rast + 2) = area;       // the whole frame's work: rast_active compares with the whole frame
  }
}

// k_rast_items: one warp per item.  Cheap stage: the rectangle's samples against the edge record,
// 32 at a time; the survivors go to the warp's queue in shared memory.  Whenever the queue holds
// 32, every lane takes one (triangle, sample) pair and runs the exact test on it -- full warps for
// the expensive stage whatever the rectangles look like.  Queue entries carry their triangle, so
// the queue simply runs on across items.
constexpr int RAST_WARPS = 8;
#ifndef RTB_RAST_SCAN_MAX
#define RTB_RAST_SCAN_MAX 96
#endif
constexpr int RAST_SCAN_MAX = RTB_RAST_SCAN_MAX;   // rectangles of more samples are walked row by row
// k / w == (k * c_rast_rw[w]) >> 16 for k < 1024, w <= 32: 65536 / w + 1 (the division costs 20 instructions per item)
__constant__ unsigned c_rast_rw[33] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2049};

__device__ __forceinline__ void rast_test(const DevScene &sc, const FrameArgs &fa, const uint2 q) {
  const int px = (int)(q.y & 0xffffu), py = (int)(q.y >> 16);
  const float4 a4 = __ldg(sc.v0 + q.x), b4 = __ldg(sc.e1 + q.x), c4 = __ldg(sc.e2 + q.x);
  V3 o, d;
  primary_ray(fa.cam, px, py, o, d);
  float dist;
  if (tri_candidate(o, d, a4, b4, c4, dist)) {
    const unsigned long long key = ((unsigned long long)__float_as_uint(dist) << 32) | __float_as_uint(a4.w);
    unsigned long long *dst = fa.prim + (size_t)py * fa.cam.W + px;
    if (key < *dst) atomicMin(dst, key);
  }
}

__global__ void __launch_bounds__(32 * RAST_WARPS) k_rast_items(const DevScene sc, const FrameArgs fa) {
  __shared__ uint2 queue[RAST_WARPS][64];
  if (!rast_active(fa)) return;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned lt = (1u << lane) - 1u;
  const unsigned warps = gridDim.x * RAST_WARPS;
  uint2 *Q = queue[wid];
  int qn = 0;                                   // warp-uniform
  // the items of this rank: one local segment, or (sharded) one segment per source rank in this rank's inbox
  // (segment by segment: choosing the segment per item cost 40 instructions of selects over kernel parameters)
  const int n_src = fa.x.sharded ? fa.nranks : 1;
  const unsigned w0 = blockIdx.x * RAST_WARPS + wid;
  unsigned first = 0;                            // global number of the segment's first item: items go round robin over the warps
  for (int q = 0; q < n_src; ++q) {
  const RastItem *src = fa.x.in[q];
  const unsigned nq = fa.x.sharded ? fa.x.ctl[fa.rank][XCHG_COUNT + (fa.x.epoch & 1u) * XCHG_MAX_RANKS + q] : min(fa.rast[8], fa.x.cap);
  const unsigned j0 = (w0 + warps - first % warps) % warps;
  first += nq;
  for (unsigned j = j0; j < nq; j += warps) {
    const uint4 it = src[j].hdr;              // plain loads: a sharded frame's items were written by peers
    const int x0 = (int)(it.y & 0xffffu), y0 = (int)(it.y >> 16), x1 = (int)(it.z & 0xffffu), y1 = (int)(it.z >> 16);
    const int w = x1 - x0 + 1, cnt = w * (y1 - y0 + 1);
    const unsigned rw = c_rast_rw[w];
    const float4 r0 = src[j].r0, r1 = src[j].r1, r2 = src[j].r2;
    const float ox = (float)x0 - r2.y, oy = (float)y0 - r2.z;     // item corner relative to the record's origin
    if (cnt <= RAST_SCAN_MAX) {
      // small rectangle: its samples 32 at a time
      for (int k0 = 0; k0 < cnt; k0 += 32) {
        const int k = k0 + lane;
        const int ry = (int)(((unsigned)k * rw) >> 16), rx = k - ry * w;
        const bool in = k < cnt && region::record_inside(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w,
                                                           ox + (float)rx, oy + (float)ry);
        const unsigned mask = __ballot_sync(0xffffffffu, in);
        if (in) Q[qn + __popc(mask & lt)] = make_uint2(it.x, (unsigned)(x0 + rx) | ((unsigned)(y0 + ry) << 16));
        qn += __popc(mask);
        __syncwarp();
        if (qn >= 32) {
          qn -= 32;
          const uint2 q = Q[qn + lane];
          __syncwarp();
          rast_test(sc, fa, q);
        }
      }
    } else {
      // big rectangle (a tile of a big or a thin region): lane r takes row r and finds the span of
      // its row that the record can pass (region::record_span, a superset; record_inside decides);
      // the spans' samples are then numbered through (prefix sums over the lanes) and taken 32 at
      // a time, so a thin region costs its length and every round runs with full lanes
      const int h = y1 - y0 + 1;                  // <= 32
      int a = 1, b = 0;
      if (lane < h) region::record_span(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w, ox, w, oy + (float)lane, a, b);
      const int len = max(b - a + 1, 0);
      int P = len;                                // inclusive prefix sum of the span lengths
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, P, d);
        if (lane >= d) P += o;
      }
      const int T = __shfl_sync(0xffffffffu, P, 31);
      const int Pex = P - len;
      for (int k0 = 0; k0 < T; k0 += 32) {
        const int k = k0 + lane;
        int row = 0;                              // number of rows whose spans end at or before sample k
#pragma unroll
        for (int step = 16; step; step >>= 1) {
          const int pr = __shfl_sync(0xffffffffu, P, row + step - 1);
          if (pr <= k) row += step;
        }
        const int ar = __shfl_sync(0xffffffffu, a, row), pe = __shfl_sync(0xffffffffu, Pex, row);
        const int rx = ar + (k - pe);
        const bool in = k < T && region::record_inside(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.w,
                                                         ox + (float)rx, oy + (float)row);
        const unsigned mask = __ballot_sync(0xffffffffu, in);
        if (in) Q[qn + __popc(mask & lt)] = make_uint2(it.x, (unsigned)(x0 + rx) | ((unsigned)(y0 + row) << 16));
        qn += __popc(mask);
        __syncwarp();
        if (qn >= 32) {
          qn -= 32;
          const uint2 q = Q[qn + lane];
          __syncwarp();
          rast_test(sc, fa, q);
        }
      }
    }
  }
  }
  if (lane < qn) rast_test(sc, fa, Q[lane]);
}

// The primary hit of sample (px, py) from the record.
__device__ __forceinline__ bool prim_lookup(const DevScene &sc, const FrameArgs &fa, int px, int py, const V3 o,
                                            const V3 d, Best &best) {
  const unsigned long long key = fa.prim[(size_t)py * fa.cam.W + px];
  best.dist = INF; best.id = 0x7fffffff; best.slot = -1;
  if (key != ~0ull) {
    best.dist = __uint_as_float((unsigned)(key >> 32));
    best.id = (int)(unsigned)key;
    best.slot = sc.inv_order ? (int)__ldg(sc.inv_order + best.id) : best.id;      // no hierarchy: slot == id
  }
  return best.slot >= 0;
}

// rtb_scene_hit_ids through the same records (tests)
__global__ void __launch_bounds__(256) k_prim_ids(const DevScene sc, const FrameArgs fa, int32_t *tri, int32_t *obj) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)fa.cam.W * fa.cam.H) return;
  const int px = (int)(i % fa.cam.W), py = (int)(i / fa.cam.W);
  if (!rast_active(fa)) { tri[i] = -2; if (obj) obj[i] = -2; return; }      // this frame would walk the tree
  V3 o, d;
  primary_ray(fa.cam, px, py, o, d);
  Best b;
  const bool hit = prim_lookup(sc, fa, px, py, o, d, b);
  tri[i] = hit ? b.id : -1;
  if (obj) obj[i] = hit ? __float_as_int(__ldg(sc.e1 + b.slot).w) : -1;
}

}  // namespace
