// Tile counts -> tile ranges, cursors, num_rendered: the single-CTA step between counting the
// instances per tile (preprocess_kernel) and placing them (duplicate_kernel).  Run by the service CTA
// of the duplicate kernel (images up to SCAN_IN_DUP_MAX_TILES tiles: every duplicate CTA then scans the
// counts for itself as well and nobody waits for this one) or by tile_scan_kernel in front of it.
#pragma once
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int DUP_THREADS = PRE_THREADS;
constexpr int DUP_WARPS = DUP_THREADS / 32;
constexpr int DUP_REG_TILES = 8; // tile counts a thread keeps in registers between its two passes (T <= 4096)

// exclusive prefix of one value per thread over the CTA (DUP_THREADS threads); *total = the sum.
// Two barriers; s_warp is free for reuse after the call's second barrier has been passed by all.
__device__ __forceinline__ uint32_t cta_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads(); // s_warp may still be read from an earlier call
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t run = inc - v, all = 0;
#pragma unroll
  for (int w = 0; w < DUP_WARPS; w++) {
    const uint32_t s = s_warp[w];
    if (w < warp) run += s;
    all += s;
  }
  *total = all;
  return run;
}

// Tile counts -> tile ranges (reference identifyTileRanges, rasterizer_impl.cu:116-138 + the
// memset at :313: [exclusive prefix of the counts, +count), (0,0) for an empty tile), the tile
// cursors (= segment starts, with seed_cursors), num_rendered and the overflow flag.  One CTA.
__device__ __forceinline__ void tile_ranges(const TileScanArgs& a, const bool seed_cursors) {
  __shared__ uint32_t s_warp[DUP_WARPS];
  const int tid = threadIdx.x;
  const int T = a.num_tiles;
  const int per = (T + DUP_THREADS - 1) / DUP_THREADS;
  const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
  uint32_t cnt[DUP_REG_TILES], sum = 0;
  if (per <= DUP_REG_TILES) {
#pragma unroll
    for (int k = 0; k < DUP_REG_TILES; k++) {
      cnt[k] = t0 + k < t1 ? __ldcg(a.tile_count + (size_t)(t0 + k) * CNT_STRIDE) : 0u;
      sum += cnt[k];
    }
  } else {
    for (int t = t0; t < t1; t++) sum += __ldcg(a.tile_count + (size_t)t * CNT_STRIDE);
  }
  uint32_t total;
  uint32_t run = cta_exclusive_scan(sum, s_warp, &total);
  if (tid == 0) {
    a.header->num_rendered = total;
    a.header->overflow = (int64_t)total > a.capacity ? 1u : 0u;
    a.header->render_ticket = 0u;
    a.info[0] = total;
    a.info[1] = (int64_t)total > a.capacity ? 1u : 0u; // a backward launched without the host having seen the count stands down
    // straight into the caller's pinned host slot: a 4-byte copy-engine transfer between two kernels
    // costs ~12 us of stream time on B200, a posted store costs nothing (the host polls this slot, or
    // reads it when it next needs the count)
    if (a.num_rendered_host) *(volatile int32_t*)a.num_rendered_host = (int32_t)total;
  }
  if (per <= DUP_REG_TILES) {
#pragma unroll
    for (int k = 0; k < DUP_REG_TILES; k++) {
      if (t0 + k < t1) {
        const uint32_t c = cnt[k];
        if (seed_cursors) a.tile_cursor[(size_t)(t0 + k) * CNT_STRIDE] = run;
        a.ranges[t0 + k] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
        run += c;
      }
    }
  } else {
    for (int t = t0; t < t1; t++) {
      const uint32_t c = __ldcg(a.tile_count + (size_t)t * CNT_STRIDE);
      if (seed_cursors) a.tile_cursor[(size_t)t * CNT_STRIDE] = run;
      a.ranges[t] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
      run += c;
    }
  }
}


} // namespace igs
