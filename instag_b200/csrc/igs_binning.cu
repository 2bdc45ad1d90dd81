// Stages 2+3 of the forward path and the first half of stage 4:
//   duplicate_kernel   instance-parallel tile-key duplication (reference
//                      duplicateWithKeys, rasterizer_impl.cu:70-111) fused with
//                      the digit histograms of every radix pass;
//   onesweep_kernel    hand-written stable LSD radix sort of the 64-bit
//                      tile|depth keys, one launch per 8-bit digit, chained
//                      decoupled look-back (replaces cub::DeviceRadixSort::
//                      SortPairs, rasterizer_impl.cu:306-311);
//   tile_ranges_kernel identifyTileRanges (rasterizer_impl.cu:116-138) incl.
//                      the zero fill of the ranges.
// num_rendered lives on the device (CtrlHeader): no kernel here needs the host
// to know it, grids are sized by the caller's capacity.
//
// Bound: HBM.  Algorithmic bytes: duplicate 20P + 12N; sort (8+4)*2*N per pass;
// ranges 8N + 8T.
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int DUP_THREADS = 256;
constexpr int DUP_ITEMS = 8;
constexpr int DUP_TILE = DUP_THREADS * DUP_ITEMS;

// One thread per INSTANCE (coalesced key/value stores, no per-Gaussian loop
// imbalance): instance i belongs to the Gaussian g with
// point_offsets[g-1] <= i < point_offsets[g].
__global__ void __launch_bounds__(DUP_THREADS) duplicate_kernel(BinArgs a) {
  __shared__ uint32_t s_hist[MAX_PASSES][RADIX];
  const int tid = threadIdx.x;
  for (int k = tid; k < a.passes * RADIX; k += DUP_THREADS) (&s_hist[0][0])[k] = 0;
  __syncthreads();
  const uint32_t N = min((int64_t)a.header->num_rendered, a.capacity);
  const uint32_t base = blockIdx.x * DUP_TILE;
  if (base < N) {
    // Gaussian owning the first instance of this CTA's chunk: upper_bound
    int lo = 0, hi = a.P; // first g with offsets[g] > base
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (__ldg(a.point_offsets + mid) > base) hi = mid; else lo = mid + 1;
    }
    const int g_first = lo;
#pragma unroll 1
    for (int it = 0; it < DUP_ITEMS; it++) {
      const uint32_t i = base + it * DUP_THREADS + tid;
      if (i >= N) break;
      // search forward from the CTA's first Gaussian: galloping + bisection
      int l = g_first, step = 1, h = g_first;
      while (h < a.P && __ldg(a.point_offsets + h) <= i) { l = h + 1; h += step; step <<= 1; }
      if (h > a.P) h = a.P;
      while (l < h) {
        const int mid = (l + h) >> 1;
        if (__ldg(a.point_offsets + mid) > i) h = mid; else l = mid + 1;
      }
      const int g = l;
      const uint32_t start = g == 0 ? 0u : __ldg(a.point_offsets + g - 1);
      const float4 r1 = __ldg(a.splat + (size_t)g * SPLAT_F4 + 1); // x, y, depth, extra0
      int x0, y0, x1, y1;
      tile_rect(r1.x, r1.y, __ldg(a.radii + g), a.grid_x, a.grid_y, x0, y0, x1, y1);
      const uint32_t local = i - start, w = (uint32_t)(x1 - x0);
      const uint32_t ty = y0 + local / w, tx = x0 + local % w;
      const uint64_t key = ((uint64_t)(ty * a.grid_x + tx) << 32) | __float_as_uint(r1.z);
      a.keys[0][i] = key;
      a.vals[0][i] = (uint32_t)g;
      for (int p = 0; p < a.passes; p++) atomicAdd(&s_hist[p][(key >> (RADIX_BITS * p)) & (RADIX - 1)], 1u);
    }
  }
  __syncthreads();
  for (int k = tid; k < a.passes * RADIX; k += DUP_THREADS) {
    const uint32_t c = (&s_hist[0][0])[k];
    if (c) atomicAdd(a.hist + k, c);
  }
}

// lookback word: high 2 bits = flag (1 aggregate, 2 inclusive prefix), low 30 = count
constexpr uint32_t LB_AGG = 1u << 30, LB_PREFIX = 2u << 30, LB_MASK = (1u << 30) - 1;

__device__ __forceinline__ uint32_t ld_relaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// One 8-bit digit pass.  Stability: keys are read warp-striped (a warp owns a
// contiguous chunk, item-major / lane-minor = memory order) and ranked with
// match_any in that order; warps and CTAs are ordered by prefix sums.
__global__ void __launch_bounds__(SORT_THREADS) onesweep_kernel(
    const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint64_t* __restrict__ kout,
    uint32_t* __restrict__ vout, const uint32_t* __restrict__ hist, uint32_t* lookback, CtrlHeader* header,
    int pass, int64_t capacity) {
  constexpr int WARPS = SORT_THREADS / 32;
  __shared__ uint32_t s_warp_hist[WARPS][RADIX];
  __shared__ uint64_t s_keys[SORT_TILE];
  __shared__ uint32_t s_vals[SORT_TILE];
  __shared__ uint32_t s_local_start[RADIX];  // exclusive digit offsets inside this CTA
  __shared__ uint32_t s_global_base[RADIX];  // where this CTA's run of each digit starts in the output
  __shared__ uint32_t s_scan[WARPS];
  __shared__ uint32_t s_bid;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int shift = RADIX_BITS * pass;

  const uint32_t N = min((int64_t)header->num_rendered, capacity);
  const uint32_t num_tiles = (N + SORT_TILE - 1) / SORT_TILE;
  if (tid == 0) s_bid = atomicAdd(&header->ticket[pass], 1u);
  for (int k = tid; k < WARPS * RADIX; k += SORT_THREADS) (&s_warp_hist[0][0])[k] = 0;
  __syncthreads();
  const uint32_t bid = s_bid;
  if (bid >= num_tiles) return;

  // ---- load (warp-striped) and rank within the warp ----
  const uint32_t warp_base = bid * SORT_TILE + warp * (32 * SORT_ITEMS);
  uint64_t key[SORT_ITEMS];
  uint32_t rank[SORT_ITEMS];
  const unsigned lt_mask = (1u << lane) - 1;
#pragma unroll
  for (int it = 0; it < SORT_ITEMS; it++) {
    const uint32_t i = warp_base + it * 32 + lane;
    key[it] = i < N ? kin[i] : ~0ull;
  }
#pragma unroll
  for (int it = 0; it < SORT_ITEMS; it++) {
    const uint32_t i = warp_base + it * 32 + lane;
    const bool valid = i < N;
    const uint32_t d = (uint32_t)(key[it] >> shift) & (RADIX - 1);
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? d : RADIX);
    const int leader = __ffs(peers) - 1;
    uint32_t pre = 0;
    if (valid && lane == leader) {
      pre = s_warp_hist[warp][d];
      s_warp_hist[warp][d] = pre + __popc(peers);
    }
    pre = __shfl_sync(0xffffffffu, pre, leader);
    rank[it] = pre + __popc(peers & lt_mask);
    __syncwarp();
  }
  __syncthreads();

  // ---- per digit (thread d): warp offsets, CTA count, chained look-back ----
  uint32_t count = 0;
  {
    const int d = tid;
#pragma unroll
    for (int w = 0; w < WARPS; w++) {
      const uint32_t t = s_warp_hist[w][d];
      s_warp_hist[w][d] = count;
      count += t;
    }
    uint32_t* my = lookback + (size_t)bid * RADIX + d;
    uint32_t excl = 0;
    if (bid == 0) {
      st_relaxed(my, LB_PREFIX | count);
    } else {
      st_relaxed(my, LB_AGG | count);
      int j = (int)bid - 1;
      while (true) {
        const uint32_t v = ld_relaxed(lookback + (size_t)j * RADIX + d);
        if ((v >> 30) == 0) continue; // not published yet
        excl += v & LB_MASK;
        if ((v >> 30) == 2) break;
        j--;
      }
      st_relaxed(my, LB_PREFIX | (excl + count));
    }
    // global digit start = exclusive scan of the pass histogram (256 bins)
    const uint32_t h = __ldg(hist + pass * RADIX + d);
    uint32_t hs = h, cs = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n1 = __shfl_up_sync(0xffffffffu, hs, o), n2 = __shfl_up_sync(0xffffffffu, cs, o);
      if (lane >= o) { hs += n1; cs += n2; }
    }
    // two interleaved block scans share s_scan via two rounds
    if (lane == 31) s_scan[warp] = hs;
    __syncthreads();
    uint32_t hoff = 0;
    for (int w = 0; w < warp; w++) hoff += s_scan[w];
    __syncthreads();
    if (lane == 31) s_scan[warp] = cs;
    __syncthreads();
    uint32_t coff = 0;
    for (int w = 0; w < warp; w++) coff += s_scan[w];
    const uint32_t local_start = coff + cs - count;
    s_local_start[d] = local_start;
    s_global_base[d] = (hoff + hs - h) + excl - local_start; // add the position inside the CTA to get the output index
  }
  __syncthreads();

  // ---- place into shared memory in sorted order, then coalesced scatter ----
#pragma unroll
  for (int it = 0; it < SORT_ITEMS; it++) {
    const uint32_t i = warp_base + it * 32 + lane;
    if (i < N) {
      const uint32_t d = (uint32_t)(key[it] >> shift) & (RADIX - 1);
      const uint32_t pos = s_local_start[d] + s_warp_hist[warp][d] + rank[it];
      s_keys[pos] = key[it];
      s_vals[pos] = vin[i];
    }
  }
  __syncthreads();
  const uint32_t tile_count = min((uint32_t)SORT_TILE, N - bid * SORT_TILE);
  for (uint32_t p = tid; p < tile_count; p += SORT_THREADS) {
    const uint64_t k = s_keys[p];
    const uint32_t d = (uint32_t)(k >> shift) & (RADIX - 1);
    const uint32_t dst = s_global_base[d] + p;
    kout[dst] = k;
    vout[dst] = s_vals[p];
  }
}

// ranges[t] = [first, last+1) of tile t in the sorted list, (0,0) if empty.
// One thread per tile boundary candidate: thread i looks at keys[i-1], keys[i].
__global__ void tile_ranges_kernel(const uint64_t* __restrict__ keys, const CtrlHeader* header, int64_t capacity,
                                   uint2* ranges, int num_tiles) {
  const uint32_t N = min((int64_t)header->num_rendered, capacity);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  // tiles that own no instance keep (0,0): fill the gaps between consecutive
  // tile ids here instead of a separate memset (reference rasterizer_impl.cu:313)
  if (N == 0) {
    if (i < (uint32_t)num_tiles) ranges[i] = make_uint2(0, 0);
    return;
  }
  if (i > N) return;
  const int cur = i < N ? (int)(keys[i] >> 32) : num_tiles;       // virtual end marker
  const int prev = i > 0 ? (int)(keys[i - 1] >> 32) : -1;        // virtual start marker
  if (cur != prev) {
    if (i < N) ranges[cur].x = i;
    if (i > 0) ranges[prev].y = i;
    for (int t = prev + 1; t < cur && t < num_tiles; t++) ranges[t] = make_uint2(0, 0);
  }
}

void launch_binning(const BinArgs& a, cudaStream_t s) {
  if (a.P <= 0 || a.capacity <= 0) {
    if (a.num_tiles > 0) cudaMemsetAsync(a.ranges, 0, sizeof(uint2) * a.num_tiles, s);
    return;
  }
  const int dup_blocks = (int)((a.capacity + DUP_TILE - 1) / DUP_TILE);
  duplicate_kernel<<<dup_blocks, DUP_THREADS, 0, s>>>(a);
  prof_mark(2, s);
  for (int p = 0; p < a.passes; p++) {
    const int in = p & 1, out = in ^ 1;
    onesweep_kernel<<<a.sort_blocks, SORT_THREADS, 0, s>>>(a.keys[in], a.vals[in], a.keys[out], a.vals[out],
                                                           a.hist, a.lookback + (size_t)p * a.sort_blocks * RADIX,
                                                           a.header, p, a.capacity);
  }
  prof_mark(3, s);
  const uint64_t* sorted = a.keys[a.passes & 1];
  const int64_t threads = a.capacity + 1 > a.num_tiles ? a.capacity + 1 : a.num_tiles;
  tile_ranges_kernel<<<(int)((threads + 255) / 256), 256, 0, s>>>(sorted, a.header, a.capacity, a.ranges, a.num_tiles);
  prof_mark(4, s);
  count_launch(2 + a.passes);
}

} // namespace igs
