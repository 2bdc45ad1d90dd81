// Stages 2+3 of the forward path: binning.
//
// The reference scans tiles_touched, duplicates every Gaussian into one (tile|depth) key
// per covered tile at its scan offset, sorts ALL keys with a 6-pass device-wide radix sort
// and then finds the tile boundaries (InclusiveSum, rasterizer_impl.cu:280;
// duplicateWithKeys, :70-111; cub::DeviceRadixSort::SortPairs, :306-311;
// identifyTileRanges, :116-138).  The result is fully specified -- instances ordered by
// (tile, depth bits, emission order) -- so any algorithm that produces exactly that order
// is a drop-in.  Here the tile part of the sort is done by COUNTING:
//   preprocess_kernel   counts the instances of every tile while it computes the tile
//                       rectangles (igs_preprocess.cu);
//   tile ranges         identifyTileRanges + the reference's 4-byte num_rendered readback
//                       (igs_tile_scan.cuh): the exclusive scan of the T counts.  Up to 2048
//                       tiles every duplicate CTA does that scan for itself and a service CTA
//                       writes the ranges; above, a one-CTA tile_scan_kernel runs in front.
//                       (Measured and dropped: the preprocess CTA that finishes last doing it --
//                       the fence + counter every CTA needs cost 4 us at C2 and 22 us at C5.)
//   duplicate_kernel    drops each instance of its 512 Gaussians straight into its tile's
//                       segment (slot = the tile's cursor) as one 64-bit word
//                       depth_bits<<32 | gaussian_id; its service CTA files the tiles by list
//                       length for the render kernels;
//   tile_sort_kernel    one CTA per tile sorts the segment by that word in shared memory
//                       and writes point_list (+ the reference's 64-bit keys on request).
// Sorting by (depth, id) reproduces the reference's stable order: inside one tile the
// emission order of the reference is ascending Gaussian id (a Gaussian meets a tile at
// most once), and the words are unique, so the order inside a segment before the sort does
// not matter.
//
// Bound: HBM/L2 latency.  Algorithmic bytes: duplicate 24P + 8N (+ 4N atomics),
// tile sort 8N read + 4N written (+ 8N with keys).
#include "igs_common.cuh"
#include "igs_kernels.cuh"
#include "igs_tile_scan.cuh"

namespace igs {

constexpr int ORDER_BUCKETS = 2 * DUP_THREADS;

// One warp per 32 Gaussians; the warp's instances are spread evenly over the lanes
// (a lane finds the Gaussian that owns instance k by a 5-step search over the warp's
// inclusive counts), so a screen-filling Gaussian does not serialise a thread.
// f(tile, word) is called once per instance.
template <class F>
__device__ __forceinline__ void for_each_instance(uint32_t n, uint32_t xy, uint32_t w, uint32_t depth, int gbase,
                                                  int grid_x, F&& f) {
  const int lane = threadIdx.x & 31;
  uint32_t incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  const uint32_t excl = incl - n;
  for (uint32_t kb = 0; kb < total; kb += 32) {
    const uint32_t k = kb + lane;
    // owner = number of lanes whose inclusive count is <= k
    int o = 0;
#pragma unroll
    for (int step = 16; step > 0; step >>= 1) {
      const uint32_t v = __shfl_sync(0xffffffffu, incl, o + step - 1);
      if (v <= k) o += step;
    }
    o = min(o, 31);
    const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, o);
    const uint32_t o_xy = __shfl_sync(0xffffffffu, xy, o);
    const uint32_t o_w = __shfl_sync(0xffffffffu, w, o);
    const uint32_t o_depth = __shfl_sync(0xffffffffu, depth, o);
    if (k < total) {
      const uint32_t local = k - o_excl;
      const uint32_t row = local / o_w, col = local - row * o_w;
      const uint32_t t = ((o_xy >> 16) + row) * (uint32_t)grid_x + (o_xy & 0xffffu) + col;
      f(t, ((unsigned long long)o_depth << 32) | (uint32_t)(gbase + o));
    }
  }
}

// The unit order of the render kernels: tiles with the longest lists first (their warps are the
// critical path), by one counting pass over ORDER_BUCKETS count classes.  One CTA.
__device__ void tile_order(const BinArgs& a) {
  __shared__ uint32_t s_bucket[ORDER_BUCKETS];
  __shared__ uint32_t s_warp[DUP_WARPS], s_max[DUP_WARPS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int T = a.num_tiles;
  const int per = (T + DUP_THREADS - 1) / DUP_THREADS;
  const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
#define TILE_CNT(t) __ldcg(a.tile_count + (size_t)(t) * CNT_STRIDE)
  uint32_t cmax = 0;
  for (int t = t0; t < t1; t++) cmax = max(cmax, TILE_CNT(t));
  for (int b = tid; b < ORDER_BUCKETS; b += DUP_THREADS) s_bucket[b] = 0u;
  cmax = __reduce_max_sync(0xffffffffu, cmax);
  if (lane == 0) s_max[warp] = cmax;
  __syncthreads();
#pragma unroll
  for (int w = 0; w < DUP_WARPS; w++) cmax = max(cmax, s_max[w]);
  const float qs = (float)(ORDER_BUCKETS - 1) / (float)max(cmax, 1u);
#define ORDER_BUCKET(c) ((ORDER_BUCKETS - 1) - min(ORDER_BUCKETS - 1, (int)((float)(c) * qs)))
  for (int t = t0; t < t1; t++) atomicAdd(&s_bucket[ORDER_BUCKET(TILE_CNT(t))], 1u);
  { // exclusive scan of the bucket sizes (ORDER_BUCKETS == 2 * DUP_THREADS: two per thread)
    __syncthreads();
    const uint32_t c0 = s_bucket[2 * tid], c1 = s_bucket[2 * tid + 1];
    uint32_t tot2;
    const uint32_t base = cta_exclusive_scan(c0 + c1, s_warp, &tot2);
    s_bucket[2 * tid] = base;
    s_bucket[2 * tid + 1] = base + c0;
  }
  __syncthreads();
  for (int t = t0; t < t1; t++) {
    const uint32_t slot = atomicAdd(&s_bucket[ORDER_BUCKET(TILE_CNT(t))], 1u);
    const uint2 r = __ldcg(a.ranges + t); // (written by this CTA in front of the barriers above, or by the scan kernel)
    a.tile_order[slot] = make_uint4((uint32_t)t, r.x, r.y, 0u);
  }
#undef ORDER_BUCKET
#undef TILE_CNT
}

// Who turns the tile counts into segment starts.  Up to this many tiles every duplicate CTA scans
// the counts itself (a few KB out of L2, in parallel: nothing single-CTA sits between the two kernels)
// and CTA 0, which has no Gaussians, writes ranges / order / num_rendered for the kernels behind.
// Above, that redundancy costs more than it saves (4096 tiles x 1954 CTAs re-read 256 MB: 103 us
// against 67 + 11), and a one-CTA tile_scan_kernel in front seeds absolute cursors instead.
constexpr int SCAN_IN_DUP_MAX_TILES = 2048;

__global__ void __launch_bounds__(DUP_THREADS) tile_scan_kernel(BinArgs a) {
  if (a.prof.early) pdl_launch_dependents();
  pdl_wait(); // the preprocess kernel's tile counts
  tile_ranges(a.scan, true);
}

// HIST_SMEM: the CTA first counts its instances per tile in shared memory, reserves one
// contiguous run of slots per non-empty tile with a single global atomic, and then hands
// the slots out with shared-memory atomics -- N/(instances a CTA has per tile) global
// atomics instead of N.  (The order inside a tile segment is irrelevant: tile_sort_kernel
// sorts it by unique words.)  OWN_SCAN: see SCAN_IN_DUP_MAX_TILES; the tile cursors are then
// relative to the segment starts the CTA has worked out for itself.  One CTA of the grid has no
// Gaussians: it files the tiles by list length for the render kernels (and, with OWN_SCAN, writes the
// tile ranges first).
template <bool HIST_SMEM, bool OWN_SCAN>
__global__ void __launch_bounds__(DUP_THREADS) duplicate_kernel(BinArgs a) {
  extern __shared__ uint32_t s_slot[];
  __shared__ uint32_t s_scan[DUP_WARPS];
  const int tid = threadIdx.x, lane = tid & 31;
  const int T = a.num_tiles;
  const bool service = blockIdx.x == 0;
  if (a.prof.early) pdl_launch_dependents();
  if (HIST_SMEM && !service)
    for (int t = tid; t < T; t += DUP_THREADS) s_slot[t] = 0u;
  pdl_wait(); // everything below reads what the preprocess (and tile-scan) kernel wrote
  if (service) {
    if (OWN_SCAN) tile_ranges(a.scan, false);
    tile_order(a);
    if (tid == 0) stamp_done(a.prof, ST_DUP);
    return;
  }
  if (!OWN_SCAN && __ldcg(&a.header->overflow)) return; // the frame does not fit: nothing is placed
  const int g = (int)(blockIdx.x - 1) * DUP_THREADS + tid;
  uint32_t n = 0, xy = 0, w = 1, depth = 0;
  if (g < a.P) n = __ldg(a.tiles_touched + g);
  // OWN_SCAN: this thread's run of tiles in the CTA's scan of the tile counts (requested before the
  // Gaussian's data is used, so that the two L2 round trips overlap)
  const int per = (T + DUP_THREADS - 1) / DUP_THREADS;
  const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
  uint32_t cnt[DUP_REG_TILES], sum = 0;
  if (OWN_SCAN) {
    static_assert(SCAN_IN_DUP_MAX_TILES <= DUP_REG_TILES * DUP_THREADS, "the counts of a thread's tiles stay in registers");
#pragma unroll
    for (int k = 0; k < DUP_REG_TILES; k++) {
      cnt[k] = t0 + k < t1 ? __ldcg(a.tile_count + (size_t)(t0 + k) * CNT_STRIDE) : 0u;
      sum += cnt[k];
    }
  }
  if (n) {
    const float4 r1 = __ldg(a.splat + (size_t)g * SPLAT_F4 + 1); // x, y, depth, extra0
    int x0, y0, x1, y1;
    tile_rect(r1.x, r1.y, __ldg(a.radii + g), a.grid_x, a.grid_y, x0, y0, x1, y1);
    xy = (uint32_t)x0 | ((uint32_t)y0 << 16);
    w = (uint32_t)(x1 - x0);
    depth = __float_as_uint(r1.z);
  }
  const int gbase = g - lane;
  if (HIST_SMEM) {
    __syncthreads();
    for_each_instance(n, xy, w, depth, gbase, a.grid_x,
                      [&](uint32_t t, unsigned long long) { atomicAdd(&s_slot[t], 1u); });
    if (OWN_SCAN) {
      uint32_t total;
      uint32_t run = cta_exclusive_scan(sum, s_scan, &total); // (its barriers also order the shared-memory atomics)
      if ((int64_t)total > a.capacity) return; // the frame does not fit: nothing is placed (uniform over the grid)
#pragma unroll
      for (int k = 0; k < DUP_REG_TILES; k++) {
        if (t0 + k < t1) {
          const uint32_t c = s_slot[t0 + k];
          if (c) s_slot[t0 + k] = run + atomicAdd(a.tile_cursor + (size_t)(t0 + k) * CNT_STRIDE, c);
          run += cnt[k];
        }
      }
    } else {
      __syncthreads();
      for (int t = tid; t < T; t += DUP_THREADS) {
        const uint32_t c = s_slot[t];
        if (c) s_slot[t] = atomicAdd(a.tile_cursor + (size_t)t * CNT_STRIDE, c);
      }
    }
    __syncthreads();
    for_each_instance(n, xy, w, depth, gbase, a.grid_x, [&](uint32_t t, unsigned long long word) {
      a.pairs[atomicAdd(&s_slot[t], 1u)] = word;
    });
  } else {
    for_each_instance(n, xy, w, depth, gbase, a.grid_x, [&](uint32_t t, unsigned long long word) {
      a.pairs[atomicAdd(a.tile_cursor + (size_t)t * CNT_STRIDE, 1u)] = word;
    });
  }
  if (tid == 0) stamp_done(a.prof, ST_DUP);
}

// ---- per-tile sort --------------------------------------------------------------------
constexpr int TS_THREADS = 256;
constexpr int TS_WARPS = TS_THREADS / 32;
constexpr int TS_CAP = 2048;                 // largest segment sorted in one go in shared memory
constexpr int TS_PER = TS_CAP / TS_THREADS;  // words a thread stages in registers
constexpr int TS_BUCKETS = 2048;             // most buckets of the distribution pass
constexpr int TS_L1 = 512;                   // first-level buckets of a segment beyond TS_CAP

struct TileSortShared {
  unsigned long long word[TS_CAP];
  uint32_t cnt[TS_BUCKETS];
  uint32_t l1[TS_L1];
  uint32_t red[2][TS_WARPS];
  uint32_t part[TS_WARPS];
};

// Sorts n <= TS_CAP unique words from src into out_id (+ out_key) by ONE distribution pass: the
// segment is read ONCE (into registers); depth bits are mapped monotonically onto ~n buckets
// between the segment's min and max (so a bucket holds ~1-2 words), a counting pass groups the
// words by bucket in shared memory, and every word then finds its rank among the few words of
// its own bucket.  O(n) for any spread-out depth distribution; a segment whose depths all collide
// degrades to O(n^2 / threads) but stays exact.
__device__ void sort_segment(TileSortShared& sh, const unsigned long long* __restrict__ src, const int n,
                             uint32_t* __restrict__ out_id, uint64_t* __restrict__ out_key, const uint64_t tile_hi) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long w[TS_PER];
  uint32_t lo = 0xffffffffu, hi = 0u;
#pragma unroll
  for (int k = 0; k < TS_PER; k++) {
    const int i = tid + k * TS_THREADS;
    w[k] = i < n ? __ldcg(src + i) : ~0ull;
    if (i < n) { const uint32_t d = (uint32_t)(w[k] >> 32); lo = min(lo, d); hi = max(hi, d); }
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  int nb = 32;
  while (nb < n) nb <<= 1;            // ~1 word per bucket
  nb = min(nb, TS_BUCKETS);
  __syncthreads(); // a previous segment of this CTA is done with the shared arrays
  if (lane == 0) { sh.red[0][warp] = lo; sh.red[1][warp] = hi; }
  for (int b = tid; b < nb; b += TS_THREADS) sh.cnt[b] = 0;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < TS_WARPS; q++) { lo = min(lo, sh.red[0][q]); hi = max(hi, sh.red[1][q]); }
  // bucket(d) = (d - lo) >> shift with the smallest shift that keeps (hi - lo) >> shift < nb:
  // monotone in d, between nb/2 and nb buckets in use.  (A 32.32 fixed-point multiply would fill
  // the buckets evenly, but nvcc 12.9 folds `((u64)x * m >> 32) * 4` into `x * (m >> 30)` when it
  // becomes a shared-memory index -- wrong, and it faulted; a shift cannot be mis-folded.)
  const int need = 32 - __clz(hi - lo), have = 31 - __clz(nb); // bits of the largest offset, log2(nb)
  const int shift = need > have ? need - have : 0;
#define TS_BUCKET(d) (((d) - lo) >> shift)
  // ---- count ----
#pragma unroll
  for (int k = 0; k < TS_PER; k++)
    if (tid + k * TS_THREADS < n) atomicAdd(&sh.cnt[TS_BUCKET((uint32_t)(w[k] >> 32))], 1u);
  __syncthreads();
  // ---- exclusive scan of the counts (each thread owns nb/256 consecutive buckets) ----
  {
    const int per = nb / TS_THREADS > 0 ? nb / TS_THREADS : 1; // nb is a power of two >= 32
    const int b0 = tid * per;
    uint32_t sum = 0;
    if (b0 < nb)
      for (int b = 0; b < per; b++) sum += sh.cnt[b0 + b];
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) sh.part[warp] = inc;
    __syncthreads();
    uint32_t run = inc - sum;
    for (int q = 0; q < warp; q++) run += sh.part[q];
    if (b0 < nb)
      for (int b = 0; b < per; b++) {
        const uint32_t c = sh.cnt[b0 + b];
        sh.cnt[b0 + b] = run; // start of the bucket; becomes its end after the scatter
        run += c;
      }
  }
  __syncthreads();
  // ---- scatter into bucket order ----
#pragma unroll
  for (int k = 0; k < TS_PER; k++)
    if (tid + k * TS_THREADS < n) sh.word[atomicAdd(&sh.cnt[TS_BUCKET((uint32_t)(w[k] >> 32))], 1u)] = w[k];
  __syncthreads();
  // ---- rank inside the bucket, write out ----
  for (int p = tid; p < n; p += TS_THREADS) {
    const unsigned long long wd = sh.word[p];
    const uint32_t b = TS_BUCKET((uint32_t)(wd >> 32));
    const uint32_t start = b ? sh.cnt[b - 1] : 0u, end = sh.cnt[b];
    uint32_t rank = 0;
    for (uint32_t q = start; q < end; q++) rank += sh.word[q] < wd ? 1u : 0u;
    const uint32_t dst = start + rank;
    out_id[dst] = (uint32_t)wd;
    if (out_key) out_key[dst] = tile_hi | (wd >> 32);
  }
#undef TS_BUCKET
}

// Last resort for a run of more than TS_CAP words whose depths the first level could not spread
// (thousands of equal depths): in-place bitonic network over the run in global memory (arbitrary
// n: elements beyond n are virtual +inf, every compare-exchange keeps the smaller word at the
// lower index).  Exact, O(n log^2 n).
__device__ void sort_run_bitonic(unsigned long long* run, const int n, uint32_t* __restrict__ out_id,
                                 uint64_t* __restrict__ out_key, const uint64_t tile_hi) {
  const int tid = threadIdx.x;
  for (unsigned k = 2; (k >> 1) < (unsigned)n; k <<= 1) {
    for (unsigned d = k >> 1; d > 0; d >>= 1) {
      __syncthreads();
      for (unsigned i = tid; i < (unsigned)n; i += TS_THREADS) {
        if (i & d) continue; // i is the lower index of its pair
        const unsigned j = (d == (k >> 1)) ? (i ^ (k - 1)) : (i | d); // first step of a merge mirrors
        if (j >= (unsigned)n) continue;
        const unsigned long long x = __ldcg(run + i), y = __ldcg(run + j);
        if (x > y) { __stcg(run + i, y); __stcg(run + j, x); }
      }
    }
  }
  __syncthreads();
  for (int i = tid; i < n; i += TS_THREADS) {
    const unsigned long long wd = __ldcg(run + i);
    out_id[i] = (uint32_t)wd;
    if (out_key) out_key[i] = tile_hi | (wd >> 32);
  }
}

// Segments beyond TS_CAP (a concentrated cloud: thousands of instances in one tile) take two
// levels, still O(n): the words are first distributed by depth into up to TS_L1 runs of ~1024 in
// the scratch array `tmp` (count, scan, scatter: the same monotone bucket map), and every run is
// then sorted by sort_segment.  Runs are contiguous and in depth order, so the output lands in place.
__device__ void sort_segment_big(TileSortShared& sh, const unsigned long long* __restrict__ seg,
                                 unsigned long long* __restrict__ tmp, const int n, uint32_t* __restrict__ out_id,
                                 uint64_t* __restrict__ out_key, const uint64_t tile_hi) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint32_t lo = 0xffffffffu, hi = 0u;
  for (int i = tid; i < n; i += TS_THREADS) {
    const uint32_t d = (uint32_t)(__ldcg(seg + i) >> 32);
    lo = min(lo, d); hi = max(hi, d);
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if (lane == 0) { sh.red[0][warp] = lo; sh.red[1][warp] = hi; }
  int nb = 2;
  while (nb * 1024 < n && nb < TS_L1) nb <<= 1; // runs of ~1024 words
  for (int b = tid; b < TS_L1; b += TS_THREADS) sh.l1[b] = 0;
  __syncthreads();
#pragma unroll
  for (int q = 0; q < TS_WARPS; q++) { lo = min(lo, sh.red[0][q]); hi = max(hi, sh.red[1][q]); }
  const int need = 32 - __clz(hi - lo), have = 31 - __clz(nb);
  const int shift = need > have ? need - have : 0;
  for (int i = tid; i < n; i += TS_THREADS) atomicAdd(&sh.l1[((uint32_t)(__ldcg(seg + i) >> 32) - lo) >> shift], 1u);
  __syncthreads();
  { // exclusive scan of TS_L1 = 2 * TS_THREADS run sizes
    const uint32_t c0 = sh.l1[2 * tid], c1 = sh.l1[2 * tid + 1];
    uint32_t inc = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) sh.part[warp] = inc;
    __syncthreads();
    uint32_t base = inc - (c0 + c1);
    for (int q = 0; q < warp; q++) base += sh.part[q];
    sh.l1[2 * tid] = base;          // start of the run; becomes its end after the scatter
    sh.l1[2 * tid + 1] = base + c0;
  }
  __syncthreads();
  for (int i = tid; i < n; i += TS_THREADS) {
    const unsigned long long wd = __ldcg(seg + i);
    __stcg(tmp + atomicAdd(&sh.l1[((uint32_t)(wd >> 32) - lo) >> shift], 1u), wd);
  }
  __syncthreads(); // (also makes the CTA's own global stores visible to all of its threads)
  for (int b = 0; b < nb; b++) {
    const int start = b ? (int)sh.l1[b - 1] : 0, m = (int)sh.l1[b] - start;
    if (m == 0) continue;
    if (m <= TS_CAP) sort_segment(sh, tmp + start, m, out_id + start, out_key ? out_key + start : nullptr, tile_hi);
    else sort_run_bitonic(tmp + start, m, out_id + start, out_key ? out_key + start : nullptr, tile_hi);
  }
}

__global__ void __launch_bounds__(TS_THREADS) tile_sort_kernel(BinArgs a) {
  __shared__ TileSortShared sh;
  const int tile = blockIdx.x;
  if (a.prof.early) pdl_launch_dependents();
  pdl_wait(); // the duplicate kernel's segments and tile ranges
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  // the frame is done with this tile's counter and cursor: back to zero for the next frame
  if (threadIdx.x == 0) {
    a.tile_count[(size_t)tile * CNT_STRIDE] = 0u;
    a.tile_cursor[(size_t)tile * CNT_STRIDE] = 0u;
  }
  if (n == 0 || __ldcg(&a.header->overflow)) return;
  const uint64_t tile_hi = (uint64_t)tile << 32;
  uint32_t* out_id = a.point_list + range.x;
  uint64_t* out_key = a.keys ? a.keys + range.x : nullptr;
  if (n <= TS_CAP) sort_segment(sh, a.pairs + range.x, n, out_id, out_key, tile_hi);
  else sort_segment_big(sh, a.pairs + range.x, a.pairs2 + range.x, n, out_id, out_key, tile_hi);
  if (threadIdx.x == 0) stamp_done(a.prof, ST_SORT);
}

// The reference's point_offsets (inclusive prefix sum of tiles_touched, rasterizer_impl.cu:280),
// on request only: nothing on the path reads it.  One CTA walks the array.
__global__ void __launch_bounds__(1024) point_offsets_kernel(int P, const uint32_t* __restrict__ tiles_touched,
                                                             uint32_t* __restrict__ point_offsets) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0u;
  __syncthreads();
  for (int base = 0; base < P; base += 1024) {
    const int i = base + tid;
    const uint32_t v = i < P ? tiles_touched[i] : 0u;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t run = s_carry + inc;
    for (int q = 0; q < warp; q++) run += s_warp[q];
    if (i < P) point_offsets[i] = run;
    __syncthreads();
    if (tid == 1023) s_carry = run;
    __syncthreads();
  }
}

cudaError_t launch_binning(const BinArgs& a, cudaStream_t s) {
  if (a.P <= 0 || a.num_tiles <= 0) return cudaSuccess;
  static const cudaError_t optin = cudaFuncSetAttribute(duplicate_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                        HIST_SMEM_MAX_TILES * (int)sizeof(uint32_t));
  if (optin != cudaSuccess) return optin;
  const int dup_blocks = (a.P + DUP_THREADS - 1) / DUP_THREADS + 1; // + the service CTA
  const size_t hist = (size_t)a.num_tiles * sizeof(uint32_t);
  cudaError_t e;
  if (a.num_tiles <= SCAN_IN_DUP_MAX_TILES) {
    count_launch(2);
    e = launch_after(duplicate_kernel<true, true>, dim3(dup_blocks), dim3(DUP_THREADS), hist, s, a);
  } else {
    count_launch(3);
    e = launch_after(tile_scan_kernel, dim3(1), dim3(DUP_THREADS), 0, s, a);
    if (e != cudaSuccess) return e;
    if (a.num_tiles <= HIST_SMEM_MAX_TILES)
      e = launch_after(duplicate_kernel<true, false>, dim3(dup_blocks), dim3(DUP_THREADS), hist, s, a);
    else
      e = launch_after(duplicate_kernel<false, false>, dim3(dup_blocks), dim3(DUP_THREADS), 0, s, a);
  }
  if (e != cudaSuccess) return e;
  return launch_after(tile_sort_kernel, dim3(a.num_tiles), dim3(TS_THREADS), 0, s, a);
}

void launch_point_offsets(int P, const uint32_t* tiles_touched, uint32_t* point_offsets, cudaStream_t s) {
  if (P <= 0) return;
  point_offsets_kernel<<<1, 1024, 0, s>>>(P, tiles_touched, point_offsets);
  count_launch();
}

} // namespace igs
