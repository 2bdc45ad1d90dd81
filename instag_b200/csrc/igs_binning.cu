// Stages 2+3 of the forward path: binning.
//
// The reference duplicates every Gaussian into one (tile|depth) key per covered
// tile, sorts ALL keys with a 6-pass device-wide radix sort and then finds the
// tile boundaries (duplicateWithKeys, rasterizer_impl.cu:70-111;
// cub::DeviceRadixSort::SortPairs, :306-311; identifyTileRanges, :116-138).
// The result is fully specified -- instances ordered by (tile, depth bits,
// emission order) -- so any algorithm that produces exactly that order is a
// drop-in.  Here the tile part of the sort is done by COUNTING:
//   preprocess_kernel   counts the instances of every tile while it computes the
//                       tile rectangles, and its last CTA turns the counts into
//                       the tile ranges (igs_preprocess.cu);
//   duplicate_kernel    drops every instance straight into its tile's segment
//                       (slot = atomic cursor of the tile), as one 64-bit word
//                       depth_bits<<32 | gaussian_id;
//   tile_sort_kernel    one CTA per tile sorts the segment by that word in shared
//                       memory and writes point_list (+ the reference's 64-bit keys).
// Sorting by (depth, id) reproduces the reference's stable order: inside one tile
// the emission order of the reference is ascending Gaussian id (a Gaussian meets
// a tile at most once), and the words are unique, so the order inside a segment
// before the sort does not matter.
//
// Bound: HBM/L2 latency.  Algorithmic bytes: duplicate 24P + 8N (+ 4N atomics),
// tile sort 8N read + 12N written.
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int DUP_THREADS = PRE_THREADS;

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

// HIST_SMEM: the CTA first counts its instances per tile in shared memory, reserves one
// contiguous run of slots per non-empty tile with a single global atomic, and then hands
// the slots out with shared-memory atomics -- N/(instances a CTA has per tile) global
// atomics instead of N.  (The order inside a tile segment is irrelevant: tile_sort_kernel
// sorts it by unique words.)
template <bool HIST_SMEM>
__global__ void __launch_bounds__(DUP_THREADS) duplicate_kernel(BinArgs a) {
  extern __shared__ uint32_t s_slot[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int g = blockIdx.x * DUP_THREADS + tid;
  if (HIST_SMEM)
    for (int t = tid; t < a.num_tiles; t += DUP_THREADS) s_slot[t] = 0u;
  pdl_wait(); // everything below reads what the preprocess kernel wrote
  uint32_t n = 0, xy = 0, w = 1, depth = 0;
  if (g < a.P) {
    n = __ldg(a.tiles_touched + g);
    // reference point_offsets (InclusiveSum, rasterizer_impl.cu:280): CTA-local scan + CTA prefix
    if (a.block_prefix) a.point_offsets[g] += __ldcg(a.block_prefix + blockIdx.x);
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
    __syncthreads();
    for (int t = tid; t < a.num_tiles; t += DUP_THREADS) {
      const uint32_t c = s_slot[t];
      if (c) s_slot[t] = atomicAdd(a.tile_cursor + (size_t)t * CNT_STRIDE, c);
    }
    __syncthreads();
    for_each_instance(n, xy, w, depth, gbase, a.grid_x, [&](uint32_t t, unsigned long long word) {
      const uint32_t slot = atomicAdd(&s_slot[t], 1u);
      if ((int64_t)slot < a.capacity) a.pairs[slot] = word;
    });
  } else {
    for_each_instance(n, xy, w, depth, gbase, a.grid_x, [&](uint32_t t, unsigned long long word) {
      const uint32_t slot = atomicAdd(a.tile_cursor + (size_t)t * CNT_STRIDE, 1u);
      if ((int64_t)slot < a.capacity) a.pairs[slot] = word;
    });
  }
}

__global__ void reset_cursors_kernel(BinArgs a) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.num_tiles) return;
  // start of the segment = exclusive prefix of the counts; recomputed from the ranges of the
  // non-empty tiles is not possible for empty ones, but those never receive an instance
  a.tile_cursor[(size_t)t * CNT_STRIDE] = a.ranges[t].x;
}

// ---- per-tile sort --------------------------------------------------------------------
constexpr int TS_THREADS = 256;
constexpr int TS_CAP = 4096;      // largest segment sorted in shared memory
constexpr int TS_BUCKETS = 2048;  // most buckets of the distribution pass (40 KB of shared memory in all)

// Segments are sorted by ONE distribution pass: depth bits are mapped monotonically onto
// ~n buckets between the segment's min and max (so a bucket holds ~1-2 words), a counting
// pass groups the words by bucket, and every word then finds its rank among the few words
// of its own bucket.  O(n) for any spread-out depth distribution; a segment whose depths
// all collide degrades to O(n^2 / threads) but stays exact.
__global__ void __launch_bounds__(TS_THREADS) tile_sort_kernel(BinArgs a) {
  __shared__ unsigned long long s_word[TS_CAP];
  __shared__ uint32_t s_cnt[TS_BUCKETS];
  __shared__ uint32_t s_red[2][TS_THREADS / 32];
  __shared__ uint32_t s_part[TS_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int tile = blockIdx.x;
  pdl_wait(); // the duplicate kernel's segments
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  if (n == 0 || a.header->overflow) return;
  unsigned long long* seg = a.pairs + range.x;
  uint32_t* __restrict__ out_id = a.point_list + range.x;
  uint64_t* __restrict__ out_key = a.keys + range.x;
  const uint64_t tile_hi = (uint64_t)tile << 32;

  if (n > TS_CAP) {
    // rare: in-place bitonic network over the segment in global memory (arbitrary n: elements
    // beyond n are virtual +inf, every compare-exchange keeps the smaller word at the lower index)
    for (unsigned k = 2; (k >> 1) < (unsigned)n; k <<= 1) {
      for (unsigned d = k >> 1; d > 0; d >>= 1) {
        __syncthreads();
        for (unsigned i = tid; i < (unsigned)n; i += TS_THREADS) {
          if (i & d) continue; // i is the lower index of its pair
          const unsigned j = (d == (k >> 1)) ? (i ^ (k - 1)) : (i | d); // first step of a merge mirrors
          if (j >= (unsigned)n) continue;
          const unsigned long long x = __ldcg(seg + i), y = __ldcg(seg + j);
          if (x > y) { __stcg(seg + i, y); __stcg(seg + j, x); }
        }
      }
    }
    __syncthreads();
    for (int i = tid; i < n; i += TS_THREADS) {
      const unsigned long long wd = __ldcg(seg + i);
      out_id[i] = (uint32_t)wd;
      out_key[i] = tile_hi | (wd >> 32);
    }
    return;
  }

  // ---- min / max of the depth bits ----
  uint32_t lo = 0xffffffffu, hi = 0u;
  for (int i = tid; i < n; i += TS_THREADS) {
    const uint32_t d = (uint32_t)(seg[i] >> 32);
    lo = min(lo, d); hi = max(hi, d);
  }
  lo = __reduce_min_sync(0xffffffffu, lo);
  hi = __reduce_max_sync(0xffffffffu, hi);
  if (lane == 0) { s_red[0][warp] = lo; s_red[1][warp] = hi; }
  int nb = 32;
  while (nb < n) nb <<= 1;            // ~1 word per bucket
  nb = min(nb, TS_BUCKETS);
  for (int b = tid; b < nb; b += TS_THREADS) s_cnt[b] = 0;
  __syncthreads();
#pragma unroll
  for (int w = 0; w < TS_THREADS / 32; w++) { lo = min(lo, s_red[0][w]); hi = max(hi, s_red[1][w]); }
  // bucket(d) = (d - lo) >> shift with the smallest shift that keeps (hi - lo) >> shift < nb:
  // monotone in d, between nb/2 and nb buckets in use.  (A 32.32 fixed-point multiply would fill
  // the buckets evenly, but nvcc 12.9 folds `((u64)x * m >> 32) * 4` into `x * (m >> 30)` when it
  // becomes a shared-memory index -- wrong, and it faulted; a shift cannot be mis-folded.)
  const int need = 32 - __clz(hi - lo), have = 31 - __clz(nb); // bits of the largest offset, log2(nb)
  const int shift = need > have ? need - have : 0;
#define TS_BUCKET(d) (((d) - lo) >> shift)

  // ---- count ----
  for (int i = tid; i < n; i += TS_THREADS) atomicAdd(&s_cnt[TS_BUCKET((uint32_t)(seg[i] >> 32))], 1u);
  __syncthreads();
  // ---- exclusive scan of the counts (each thread owns nb/256 consecutive buckets) ----
  {
    const int per = nb / TS_THREADS > 0 ? nb / TS_THREADS : 1; // nb is a power of two >= 32
    const int b0 = tid * per;
    uint32_t sum = 0;
    if (b0 < nb)
      for (int b = 0; b < per; b++) sum += s_cnt[b0 + b];
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    if (lane == 31) s_part[warp] = inc;
    __syncthreads();
    uint32_t run = inc - sum;
    for (int w = 0; w < warp; w++) run += s_part[w];
    if (b0 < nb)
      for (int b = 0; b < per; b++) {
        const uint32_t c = s_cnt[b0 + b];
        s_cnt[b0 + b] = run; // start of the bucket; becomes its end after the scatter
        run += c;
      }
  }
  __syncthreads();
  // ---- scatter into bucket order ----
  for (int i = tid; i < n; i += TS_THREADS) {
    const unsigned long long wd = seg[i];
    const uint32_t slot = atomicAdd(&s_cnt[TS_BUCKET((uint32_t)(wd >> 32))], 1u);
    s_word[slot] = wd;
  }
  __syncthreads();
  // ---- rank inside the bucket, write out ----
  for (int p = tid; p < n; p += TS_THREADS) {
    const unsigned long long wd = s_word[p];
    const uint32_t b = TS_BUCKET((uint32_t)(wd >> 32));
    const uint32_t start = b ? s_cnt[b - 1] : 0u, end = s_cnt[b];
    uint32_t rank = 0;
    for (uint32_t q = start; q < end; q++) rank += s_word[q] < wd ? 1u : 0u;
    const uint32_t dst = start + rank;
    out_id[dst] = (uint32_t)wd;
    out_key[dst] = tile_hi | (wd >> 32);
  }
#undef TS_BUCKET
}

void launch_binning(const BinArgs& a, cudaStream_t s) {
  if (a.P <= 0 || a.num_tiles <= 0) return;
  const int dup_blocks = (a.P + DUP_THREADS - 1) / DUP_THREADS;
  if (a.num_tiles <= HIST_SMEM_MAX_TILES)
    launch_after(duplicate_kernel<true>, dim3(dup_blocks), dim3(DUP_THREADS), (size_t)a.num_tiles * sizeof(uint32_t), s, a);
  else
    launch_after(duplicate_kernel<false>, dim3(dup_blocks), dim3(DUP_THREADS), 0, s, a);
  prof_mark(2, s);
  launch_after(tile_sort_kernel, dim3(a.num_tiles), dim3(TS_THREADS), 0, s, a);
  prof_mark(3, s);
  prof_mark(4, s);
  count_launch(2);
}

void launch_reset_cursors(const BinArgs& a, cudaStream_t s) {
  if (a.num_tiles <= 0) return;
  reset_cursors_kernel<<<(a.num_tiles + 255) / 256, 256, 0, s>>>(a);
  count_launch(1);
}

} // namespace igs
