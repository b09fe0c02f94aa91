// Hand-written device scan and LSD radix sort (no CUB / Thrust).
//
// Stands in for the reference's gpu/partitioning/prefix_sum.cu:50-184 (Blelloch scan on
// size_t, 256 items per block, one cudaMalloc per recursion level) and
// gpu/partitioning/sort.tuh:72-220 (2-bit LSD passes, one full-array scan per digit value,
// 96-byte payloads moved every pass).  Here: 8-bit digits, key + 32-bit index payload only,
// three launches per pass (tile histograms -> scan -> stable ranked scatter staged
// through shared memory so the global writes are runs of consecutive addresses), all
// scratch supplied by the caller.
#include <mutex>
#include "rtb_internal.h"

namespace rtb {

// ------------------------------------------------------------------------------------
// scan: SCAN_BLOCK threads x SCAN_ITEMS items per CTA, warp-shuffle based
// ------------------------------------------------------------------------------------
constexpr int SCAN_BLOCK = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T warp_inclusive(T v, int lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += o;
  }
  return v;
}

// Scans each tile of SCAN_TILE items independently; tile_sums[b] = total of tile b.
template <typename T, bool INCLUSIVE>
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_tiles(T *data, size_t n, T *tile_sums) {
  __shared__ T warp_tot[SCAN_BLOCK / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS];
  T sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    v[k] = (base + k < n) ? data[base + k] : T(0);
    sum += v[k];
  }
  T incl = warp_inclusive(sum, lane);
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    T w = lane < SCAN_BLOCK / 32 ? warp_tot[lane] : T(0);
    T wi = warp_inclusive(w, lane);
    if (lane < SCAN_BLOCK / 32) warp_tot[lane] = wi - w;   // exclusive over warps
    if (lane == SCAN_BLOCK / 32 - 1 && tile_sums) tile_sums[blockIdx.x] = wi;
  }
  __syncthreads();
  T run = warp_tot[warp] + (incl - sum);   // exclusive prefix of this thread's first item
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    T x = v[k];
    if (INCLUSIVE) { run += x; if (base + k < n) data[base + k] = run; }
    else { if (base + k < n) data[base + k] = run; run += x; }
  }
}

// data[i] += tile_offsets[i / SCAN_TILE]  (tile_offsets = exclusive scan of the tile sums)
template <typename T>
__global__ void __launch_bounds__(SCAN_BLOCK) k_scan_add(T *data, size_t n, const T *tile_offsets) {
  const T off = tile_offsets[blockIdx.x];
  const size_t base = (size_t)blockIdx.x * SCAN_TILE + (size_t)threadIdx.x * SCAN_ITEMS;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k)
    if (base + k < n) data[base + k] += off;
}

size_t scan_scratch_words(size_t n) {
  size_t words = 0;
  while (n > (size_t)SCAN_TILE) {
    n = (n + SCAN_TILE - 1) / SCAN_TILE;
    words += n;
  }
  return words + 1;
}

template <typename T>
static int scan_rec(T *d, size_t n, T *scratch, bool inclusive, cudaStream_t st, unsigned long long *launches) {
  if (n == 0) return RTB_OK;
  const size_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  T *sums = tiles > 1 ? scratch : nullptr;
  if (inclusive) k_scan_tiles<T, true><<<(unsigned)tiles, SCAN_BLOCK, 0, st>>>(d, n, sums);
  else k_scan_tiles<T, false><<<(unsigned)tiles, SCAN_BLOCK, 0, st>>>(d, n, sums);
  if (launches) ++*launches;
  if (tiles > 1) {
    int rc = scan_rec<T>(sums, tiles, scratch + tiles, false, st, launches);
    if (rc) return rc;
    k_scan_add<T><<<(unsigned)tiles, SCAN_BLOCK, 0, st>>>(d, n, sums);
    if (launches) ++*launches;
  }
  RTB_CUDA(cudaGetLastError());
  return RTB_OK;
}

int dev_exclusive_scan_u32(uint32_t *d, size_t n, uint32_t *scratch, size_t scratch_words, bool inclusive,
                           cudaStream_t st, unsigned long long *launches) {
  if (scratch_words < scan_scratch_words(n)) { set_error("scan: scratch too small"); return RTB_ERR_ARG; }
  return scan_rec<uint32_t>(d, n, scratch, inclusive, st, launches);
}

int dev_inclusive_scan_u64(uint64_t *d, size_t n, uint64_t *scratch, size_t scratch_words, cudaStream_t st) {
  if (scratch_words < scan_scratch_words(n)) { set_error("scan: scratch too small"); return RTB_ERR_ARG; }
  return scan_rec<uint64_t>((uint64_t *)d, n, (uint64_t *)scratch, true, st, nullptr);
}

// ------------------------------------------------------------------------------------
// radix sort, 8 bits per pass
// ------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 4096 keys per CTA
constexpr int RS_RADIX = 256;

// counts[digit * n_tiles + tile] = number of keys of `tile` whose digit at `shift` is `digit`
__global__ void __launch_bounds__(RS_THREADS) k_rs_histogram(const uint64_t *keys, size_t n, int shift,
                                                             uint32_t *counts, uint32_t n_tiles) {
  __shared__ uint32_t hist[RS_RADIX];
  hist[threadIdx.x] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * RS_TILE;
#pragma unroll 4
  for (int k = 0; k < RS_ITEMS; ++k) {
    size_t i = base + (size_t)k * RS_THREADS + threadIdx.x;
    if (i < n) atomicAdd(&hist[(uint32_t)(keys[i] >> shift) & 0xffu], 1u);
  }
  __syncthreads();
  counts[(size_t)threadIdx.x * n_tiles + blockIdx.x] = hist[threadIdx.x];
}

// Stable scatter.  Warp w of the CTA owns the contiguous slice [w*512, (w+1)*512) of the
// tile, read warp-striped so the stream order inside a warp is (item k, lane).  Ranks
// come from __match_any_sync groups plus per-warp digit counters; after a CTA-wide scan
// of the counters each key knows its tile-local sorted position, is staged in shared
// memory at that position and written out linearly.
__global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const uint64_t *keys_in, const uint32_t *vals_in,
                                                           uint64_t *keys_out, uint32_t *vals_out, size_t n,
                                                           int shift, const uint32_t *scanned, uint32_t n_tiles) {
  __shared__ uint32_t cnt[RS_WARPS][RS_RADIX];     // per-warp digit counters -> exclusive over warps
  __shared__ uint32_t digit_start[RS_RADIX];       // tile-local exclusive scan over digits
  __shared__ uint32_t glob_base[RS_RADIX];         // global position of the tile's first key of a digit
  __shared__ uint64_t stage[RS_TILE];
  __shared__ uint32_t warp_tot[RS_WARPS];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (int i = threadIdx.x; i < RS_WARPS * RS_RADIX; i += RS_THREADS) (&cnt[0][0])[i] = 0;
  __syncthreads();

  const size_t tile_base = (size_t)blockIdx.x * RS_TILE;
  const size_t warp_base = tile_base + (size_t)warp * (32 * RS_ITEMS);
  uint64_t key[RS_ITEMS];
  uint16_t rank[RS_ITEMS];
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    size_t i = warp_base + (size_t)k * 32 + lane;
    key[k] = i < n ? keys_in[i] : ~0ull;
  }
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const size_t i = warp_base + (size_t)k * 32 + lane;
    const bool valid = i < n;
    const uint32_t d = valid ? ((uint32_t)(key[k] >> shift) & 0xffu) : 0x100u;
    const uint32_t peers = __match_any_sync(0xffffffffu, d);
    uint32_t prior = 0;
    if (valid) prior = cnt[warp][d];
    __syncwarp();
    if (valid && (peers & lt_mask) == 0) cnt[warp][d] = prior + __popc(peers);
    __syncwarp();
    rank[k] = (uint16_t)(prior + __popc(peers & lt_mask));
  }
  __syncthreads();

  // exclusive scan of the counters across warps, one digit per thread
  uint32_t total = 0;
  {
    const int d = threadIdx.x;
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) {
      uint32_t c = cnt[w][d];
      cnt[w][d] = total;
      total += c;
    }
  }
  // exclusive scan of the digit totals across the 256 threads
  uint32_t incl = warp_inclusive(total, lane);
  if (lane == 31) warp_tot[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < RS_WARPS ? warp_tot[lane] : 0u;
    uint32_t wi = warp_inclusive(w, lane);
    if (lane < RS_WARPS) warp_tot[lane] = wi - w;
  }
  __syncthreads();
  {
    const uint32_t start = warp_tot[warp] + incl - total;
    digit_start[threadIdx.x] = start;
    glob_base[threadIdx.x] = scanned[(size_t)threadIdx.x * n_tiles + blockIdx.x] - start;
  }
  __syncthreads();

  // tile-local sorted position of every key
  uint16_t pos[RS_ITEMS];
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const size_t i = warp_base + (size_t)k * 32 + lane;
    if (i < n) {
      const uint32_t d = (uint32_t)(key[k] >> shift) & 0xffu;
      pos[k] = (uint16_t)(digit_start[d] + cnt[warp][d] + rank[k]);
      stage[pos[k]] = key[k];
    }
  }
  __syncthreads();
  const uint32_t tile_n = (uint32_t)((n - tile_base) < (size_t)RS_TILE ? (n - tile_base) : (size_t)RS_TILE);
#pragma unroll 4
  for (uint32_t p = threadIdx.x; p < tile_n; p += RS_THREADS) {
    const uint64_t kk = stage[p];
    const uint32_t d = (uint32_t)(kk >> shift) & 0xffu;
    keys_out[(size_t)glob_base[d] + p] = kk;
  }
  __syncthreads();
  // payload through the same staging buffer
  uint32_t *vstage = reinterpret_cast<uint32_t *>(stage);   // positions < RS_TILE: no overlap with dstage
  uint8_t *dstage = reinterpret_cast<uint8_t *>(vstage + RS_TILE);
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    const size_t i = warp_base + (size_t)k * 32 + lane;
    if (i < n) {
      vstage[pos[k]] = vals_in[i];
      dstage[pos[k]] = (uint8_t)((uint32_t)(key[k] >> shift) & 0xffu);
    }
  }
  __syncthreads();
#pragma unroll 4
  for (uint32_t p = threadIdx.x; p < tile_n; p += RS_THREADS)
    vals_out[(size_t)glob_base[dstage[p]] + p] = vstage[p];
}

size_t radix_sort_scratch_words(size_t n) {
  const size_t tiles = (n + RS_TILE - 1) / RS_TILE;
  const size_t counts = tiles * RS_RADIX;
  return counts + scan_scratch_words(counts);
}

// Sorts (key, value) pairs by bits [low_bit, key_bits) of the key, stable (keys equal in those bits keep their order).  Ping-pongs
// between the primary and the alternate buffers; *result_in_alt tells where the result is.
int dev_radix_sort_pairs_u64(uint64_t *d_keys, uint32_t *d_vals, uint64_t *d_keys_alt, uint32_t *d_vals_alt,
                             size_t n, int key_bits, uint32_t *d_scratch, size_t scratch_words, cudaStream_t st,
                             int *result_in_alt, unsigned long long *launches, int low_bit) {
  *result_in_alt = 0;
  if (n == 0) return RTB_OK;
  if (n >= (1ull << 32)) { set_error("radix sort: more than 2^32-1 keys"); return RTB_ERR_LIMIT; }
  if (key_bits < 1) key_bits = 1;
  if (key_bits > 64) key_bits = 64;
  if (scratch_words < radix_sort_scratch_words(n)) { set_error("radix sort: scratch too small"); return RTB_ERR_ARG; }
  const uint32_t tiles = (uint32_t)((n + RS_TILE - 1) / RS_TILE);
  const size_t counts = (size_t)tiles * RS_RADIX;
  uint32_t *d_counts = d_scratch;
  uint32_t *d_scan_scratch = d_scratch + counts;
  if (low_bit < 0 || low_bit >= key_bits) low_bit = 0;
  const int passes = (key_bits - low_bit + 7) / 8;
  uint64_t *kin = d_keys, *kout = d_keys_alt;
  uint32_t *vin = d_vals, *vout = d_vals_alt;
  for (int p = 0; p < passes; ++p) {
    const int shift = low_bit + 8 * p;
    k_rs_histogram<<<tiles, RS_THREADS, 0, st>>>(kin, n, shift, d_counts, tiles);
    if (launches) ++*launches;
    int rc = dev_exclusive_scan_u32(d_counts, counts, d_scan_scratch, scratch_words - counts, false, st, launches);
    if (rc) return rc;
    k_rs_scatter<<<tiles, RS_THREADS, 0, st>>>(kin, vin, kout, vout, n, shift, d_counts, tiles);
    if (launches) ++*launches;
    uint64_t *tk = kin; kin = kout; kout = tk;
    uint32_t *tv = vin; vin = vout; vout = tv;
  }
  RTB_CUDA(cudaGetLastError());
  *result_in_alt = (passes & 1);
  return RTB_OK;
}

}  // namespace rtb

// ------------------------------------------------------------------------------------
// rtb_dev_inclusive_scan_u32_static: the scan of rtb_dev_inclusive_scan_u32 for callers
// that may not allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
// The tile sums of every level live in ONE file-scope device array, held by
// every GPU that loads the library until the process exits:
//   n <= RTB_SCAN_STATIC_MAX = 2^26 elements  ->  2^15 + 2^4 + 1 sums, 33 024 words = 129 KB.
// The array is shared by all calls and keeps the last call's contents.  Nothing in it has to
// start at zero (k_scan_tiles writes every tile sum before k_scan_add reads it), but two calls
// that overlap on the GPU (different streams, or different host threads without ordering)
// race on it: one call at a time per device.
// In host code `g_scan_static` is only a stand-in for the symbol, NOT a device address; the
// address comes from cudaGetSymbolAddress, once per device, with that device current.
// ------------------------------------------------------------------------------------
namespace rtb {
namespace {
constexpr size_t SCAN_STATIC_MAX = (size_t)1 << 26;
constexpr size_t SCAN_STATIC_WORDS = 33024;
static __device__ __align__(256) uint32_t g_scan_static[SCAN_STATIC_WORDS];
constexpr int SCAN_STATIC_DEVICES = 64;
std::mutex g_scan_static_mu;
uint32_t *g_scan_static_ptr[SCAN_STATIC_DEVICES] = {nullptr};
}  // namespace
}  // namespace rtb

extern "C" int rtb_dev_inclusive_scan_u32_static(uint32_t *d_data, size_t n, void *stream) {
  using namespace rtb;
  if (!n) return RTB_OK;
  if (!d_data) { set_error("NULL argument"); return RTB_ERR_ARG; }
  if (n > SCAN_STATIC_MAX) {
    set_error("rtb_dev_inclusive_scan_u32_static: %zu elements exceed the static scratch's limit of %zu "
              "(RTB_SCAN_STATIC_MAX); use rtb_dev_inclusive_scan_u32", n, SCAN_STATIC_MAX);
    return RTB_ERR_LIMIT;
  }
  if (scan_scratch_words(n) > SCAN_STATIC_WORDS) { set_error("static scan scratch too small"); return RTB_ERR_LIMIT; }
  int device = -1;
  RTB_CUDA(cudaGetDevice(&device));
  if (device < 0 || device >= SCAN_STATIC_DEVICES) { set_error("device %d out of range", device); return RTB_ERR_LIMIT; }
  uint32_t *scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_scan_static_mu);
    if (!g_scan_static_ptr[device]) {
      void *p = nullptr;
      RTB_CUDA(cudaGetSymbolAddress(&p, g_scan_static));       // this device's copy of the array
      g_scan_static_ptr[device] = (uint32_t *)p;
    }
    scratch = g_scan_static_ptr[device];
  }
  return dev_exclusive_scan_u32(d_data, n, scratch, SCAN_STATIC_WORDS, true, (cudaStream_t)stream, nullptr);
}
