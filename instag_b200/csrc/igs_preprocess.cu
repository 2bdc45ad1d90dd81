// Stage 1 of the forward path: per-Gaussian preprocess (EWA projection, 3D->2D
// covariance, SH->RGB, radius, tile rectangle) fused with the inclusive prefix
// sum of tiles_touched (single-pass decoupled look-back), so that one launch
// replaces the reference's preprocessCUDA + cub::DeviceScan::InclusiveSum
// (reference forward.cu:203-316, rasterizer_impl.cu:280).  Also markVisible.
//
// Bound: HBM.  Algorithmic bytes per Gaussian: reads 44 (+12*(D+1)^2 SH),
// writes 64 (splat record) + 4 radii + 4 tiles + 4 offsets (+1 clamp mask).
#include "igs_common.cuh"
#include "igs_kernels.cuh"
#include "igs_math.cuh"

namespace igs {

__device__ __constant__ float kSH_C0 = 0.28209479177387814f;
__device__ __constant__ float kSH_C1 = 0.4886025119029199f;
__device__ __constant__ float kSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                           -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float kSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                           0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                           -0.5900435899266435f};

// SH -> RGB (reference forward.cu:20-71); returns colour, sets clamp mask bits
// The 3*(deg+1)^2 coefficients of the Gaussian are fetched up front into registers -- with 16-byte
// loads when the row is 16-byte aligned (M % 4 == 0) -- instead of 48 dependent scalar loads at a
// 12*M-byte stride per lane.
__device__ __forceinline__ float3 sh_to_rgb(int deg, int M, const float* __restrict__ sh, float x, float y, float z,
                                            uint32_t& clamp_mask) {
  float v[48];
  const int need = 3 * (deg + 1) * (deg + 1);
  if ((M & 3) == 0) {
    const float4* p = reinterpret_cast<const float4*>(sh);
#pragma unroll
    for (int i = 0; i < 12; i++)
      if (4 * i < need) {
        const float4 t = __ldg(p + i);
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
      }
  } else {
#pragma unroll
    for (int j = 0; j < 48; j++)
      if (j < need) v[j] = __ldg(sh + j);
  }
  float res[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
#define SHK(k) v[3 * (k) + c]
    float r = __fmul_rn(kSH_C0, SHK(0));
    if (deg > 0) {
      r = __fmaf_rn(-__fmul_rn(kSH_C1, y), SHK(1), r); // the degree-0 product stays rounded
      r = __fmaf_rn(__fmul_rn(kSH_C1, z), SHK(2), r);
      r = __fmaf_rn(-__fmul_rn(kSH_C1, x), SHK(3), r);
      if (deg > 1) {
        const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
        const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
        r = __fmaf_rn(__fmul_rn(kSH_C2[0], xy), SHK(4), r);
        r = __fmaf_rn(__fmul_rn(kSH_C2[1], yz), SHK(5), r);
        r = __fmaf_rn(__fmul_rn(kSH_C2[2], __fsub_rn(__fmaf_rn(2.f, zz, -xx), yy)), SHK(6), r);
        r = __fmaf_rn(__fmul_rn(kSH_C2[3], xz), SHK(7), r);
        r = __fmaf_rn(__fmul_rn(kSH_C2[4], __fsub_rn(xx, yy)), SHK(8), r);
        if (deg > 2) {
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[0], y), __fmaf_rn(3.f, xx, -yy)), SHK(9), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[1], xy), z), SHK(10), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[2], y), __fsub_rn(__fmaf_rn(4.f, zz, -xx), yy)), SHK(11), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[3], z),
                                  __fmaf_rn(-3.f, yy, __fmaf_rn(-3.f, xx, __fmul_rn(2.f, zz)))), SHK(12), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[4], x), __fsub_rn(__fmaf_rn(4.f, zz, -xx), yy)), SHK(13), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[5], z), __fsub_rn(xx, yy)), SHK(14), r);
          r = __fmaf_rn(__fmul_rn(__fmul_rn(kSH_C3[6], x), __fmaf_rn(-3.f, yy, xx)), SHK(15), r);
        }
      }
    }
#undef SHK
    r = __fadd_rn(r, 0.5f);
    if (r < 0.f) clamp_mask |= 1u << c;
    res[c] = fmaxf(r, 0.f);
  }
  return make_float3(res[0], res[1], res[2]);
}

constexpr int ORDER_BUCKETS = 2 * PRE_THREADS;

#ifdef IGS_PRE_TRACE
__device__ unsigned long long g_pre_trace[8 * 4096];
__device__ __forceinline__ unsigned long long pre_timer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define PRE_MARK(k) do { if (threadIdx.x == 0) g_pre_trace[8 * blockIdx.x + (k)] = pre_timer(); } while (0)
#else
#define PRE_MARK(k) do { } while (0)
#endif

template <bool USE_SH, bool HIST_SMEM>
__global__ void __launch_bounds__(PRE_THREADS, 1024 / PRE_THREADS) preprocess_kernel(PreArgs a) {
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  __shared__ uint32_t s_warp_sum[PRE_THREADS / 32];
  __shared__ uint32_t s_last;
  // CTA-local histogram of the instances per tile (HIST_SMEM: the launch gave us 4*T bytes):
  // one global RED per (CTA, non-empty tile) instead of one per instance
  extern __shared__ uint32_t s_hist[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int T = a.grid_x * a.grid_y;
  if (HIST_SMEM)
    for (int t = tid; t < T; t += PRE_THREADS) s_hist[t] = 0u;
  pdl_wait(); // launched as a programmatic dependent of whatever precedes it on the stream
  if (tid < 16) { s_view[tid] = __ldg(a.viewmatrix + tid); s_proj[tid] = __ldg(a.projmatrix + tid); }
  if (tid < 3) s_cam[tid] = __ldg(a.campos + tid);
  PRE_MARK(0);
  __syncthreads();
  PRE_MARK(1);
  const int bid = (int)blockIdx.x;

  const int idx = bid * PRE_THREADS + tid;
  uint32_t tiles = 0;
  int radius = 0;
  if (idx < a.P) {
    // every input of this Gaussian is requested up front (one DRAM round trip instead of three
    // dependent ones); the few that turn out to be culled cost 44 B of bandwidth
    const float px = __ldg(a.means3D + 3 * idx), py = __ldg(a.means3D + 3 * idx + 1), pz = __ldg(a.means3D + 3 * idx + 2);
    const float3 sc = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
    const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
    const float op = __ldg(a.opacities + idx);
    const float e0 = a.extra_attrs ? __ldg(a.extra_attrs + (size_t)idx * a.ED) : 0.f;
    float3 rgb = make_float3(0.f, 0.f, 0.f);
    if (!USE_SH) rgb = make_float3(__ldg(a.colors_precomp + 3 * idx), __ldg(a.colors_precomp + 3 * idx + 1),
                                   __ldg(a.colors_precomp + 3 * idx + 2));
    const float vz = xform_row(s_view, 2, px, py, pz);
    if (vz <= NEAR_Z) { // reference auxiliary.h:155 (NaN passes, as there)
      if (a.prefiltered) { printf("Point is filtered although prefiltered is set. This shouldn't happen!"); __trap(); }
    } else {
      const float hx = xform_row(s_proj, 0, px, py, pz), hy = xform_row(s_proj, 1, px, py, pz);
      const float hw = xform_row(s_proj, 3, px, py, pz);
      const float p_w = __fdiv_rn(1.0f, __fadd_rn(hw, 0.0000001f));
      const float ndc_x = __fmul_rn(hx, p_w), ndc_y = __fmul_rn(hy, p_w);

      const Rot R = quat_to_rot(q);
      float c3[6];
      if (a.cov3Ds_precomp) {
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = __ldg(a.cov3Ds_precomp + 6 * idx + k);
      } else {
        cov3d_from_scale_rot(sc, a.scale_modifier, R, c3);
      }
      const float tx = xform_row(s_view, 0, px, py, pz), ty = xform_row(s_view, 1, px, py, pz);
      const float3 cov = cov2d(tx, ty, vz, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c3, s_view);
      const float det = __fmaf_rn(cov.x, cov.z, -__fmul_rn(cov.y, cov.y));
      if (det != 0.0f) {
        const float det_inv = __fdiv_rn(1.f, det);
        const float mid = __fmul_rn(0.5f, __fadd_rn(cov.x, cov.z));
        const float disc = __fsqrt_rn(fmaxf(0.1f, __fmaf_rn(mid, mid, -det)));
        const float lam = fmaxf(__fadd_rn(mid, disc), __fsub_rn(mid, disc));
        const float my_radius = ceilf(__fmul_rn(3.f, __fsqrt_rn(lam)));
        const float sx = ndc2pix(ndc_x, a.W), sy = ndc2pix(ndc_y, a.H);
        int x0, y0, x1, y1;
        tile_rect(sx, sy, (int)my_radius, a.grid_x, a.grid_y, x0, y0, x1, y1);
        const uint32_t n_tiles = (uint32_t)((x1 - x0) * (y1 - y0));
        if (n_tiles != 0) {
          tiles = n_tiles;
          radius = (int)my_radius;
          // instances per tile: sizes the tile segments the duplicate kernel fills
          for (int ty = y0; ty < y1; ty++)
            for (int tx = x0; tx < x1; tx++) {
              if (HIST_SMEM) atomicAdd(&s_hist[ty * a.grid_x + tx], 1u);
              else atomicAdd(a.tile_count + (size_t)(ty * a.grid_x + tx) * CNT_STRIDE, 1u);
            }
          // normal: rotated axis of the smallest RAW scale, flipped to face the camera
          float nx, ny, nz;
          if (a.norm3Ds_precomp) {
            nx = __ldg(a.norm3Ds_precomp + 3 * idx); ny = __ldg(a.norm3Ds_precomp + 3 * idx + 1);
            nz = __ldg(a.norm3Ds_precomp + 3 * idx + 2);
          } else {
            int k = 0;
            if (sc.x > sc.z && sc.y > sc.z) k = 2;
            else if (sc.x > sc.y && sc.z > sc.y) k = 1;
            nx = R.m[0][k]; ny = R.m[1][k]; nz = R.m[2][k];
            const float rx = __fsub_rn(px, s_cam[0]), ry = __fsub_rn(py, s_cam[1]), rz = __fsub_rn(pz, s_cam[2]);
            if (dot3c(rx, nx, ry, ny, rz, nz) > 0.f) { nx = -nx; ny = -ny; nz = -nz; }
          }
          if (USE_SH) {
            float dx = __fsub_rn(px, s_cam[0]), dy = __fsub_rn(py, s_cam[1]), dz = __fsub_rn(pz, s_cam[2]);
            const float len = __fsqrt_rn(dot3c(dx, dx, dy, dy, dz, dz));
            dx = __fdiv_rn(dx, len); dy = __fdiv_rn(dy, len); dz = __fdiv_rn(dz, len);
            uint32_t mask = 0;
            rgb = sh_to_rgb(a.D, a.M, a.shs + (size_t)idx * a.M * 3, dx, dy, dz, mask);
            a.clamped[idx] = (uint8_t)mask;
          }
          float4* rec = a.splat + (size_t)idx * SPLAT_F4;
          rec[0] = make_float4(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv), __fmul_rn(cov.x, det_inv), op);
          rec[1] = make_float4(sx, sy, vz, e0);
          rec[2] = make_float4(rgb.x, rgb.y, rgb.z, nx);
          rec[3] = make_float4(ny, nz, 2.f * logf(255.f * op), 0.f); // culling bound, see splat_may_touch
        }
      }
    }
    a.radii[idx] = radius;
    a.tiles_touched[idx] = tiles;
  }

  PRE_MARK(2);
  // block inclusive scan of `tiles`
  uint32_t incl = tiles;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  if (lane == 31) s_warp_sum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = lane < PRE_THREADS / 32 ? s_warp_sum[lane] : 0u;
    uint32_t wi = w;
#pragma unroll
    for (int o = 1; o < PRE_THREADS / 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += n;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, wi, PRE_THREADS / 32 - 1);
    if (lane < PRE_THREADS / 32) s_warp_sum[lane] = wi - w; // exclusive warp offsets
    // no look-back chain: the CTA publishes its total, the last CTA scans the totals, and the
    // duplicate kernel adds the CTA's prefix to these CTA-local offsets
    if (lane == 0) a.block_total[bid] = total;
  }
  __syncthreads();
  PRE_MARK(3);
  if (idx < a.P) a.point_offsets[idx] = s_warp_sum[warp] + incl;
  if (HIST_SMEM) { // (the barrier above ordered every thread's shared-memory atomics)
    for (int t = tid; t < T; t += PRE_THREADS) {
      const uint32_t c = s_hist[t];
      if (c) atomicAdd(a.tile_count + (size_t)t * CNT_STRIDE, c);
    }
  }

  // ---- the CTA that finishes last turns the tile counts into tile ranges ------------------
  // (reference identifyTileRanges, rasterizer_impl.cu:116-138 + the memset at :313: the
  // range of a tile is [exclusive prefix of the counts, +count), (0,0) for an empty tile)
  __threadfence();
  __syncthreads();
  PRE_MARK(4);
  if (tid == 0) s_last = atomicAdd(&a.header->done, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  PRE_MARK(5);
  if (!s_last) return;
  __threadfence();
  // (1) tile counts: one L2 round trip into shared memory (all requests in flight together);
  //     later phases read them from there.  Beyond HIST_SMEM_MAX_TILES they stay in L2.
  //     The counters go back to zero as they are read: the next frame needs no memset of [ctrl].
  if (HIST_SMEM) {
    for (int t = tid; t < T; t += PRE_THREADS) {
      s_hist[t] = __ldcg(a.tile_count + (size_t)t * CNT_STRIDE);
      a.tile_count[(size_t)t * CNT_STRIDE] = 0u;
    }
  }
  if (tid == 0) { a.header->done = 0u; a.header->render_ticket = 0u; }
#define TILE_CNT(t) (HIST_SMEM ? s_hist[t] : __ldcg(a.tile_count + (size_t)(t) * CNT_STRIDE))
  // (2) CTA totals of tiles_touched -> exclusive prefixes (added to point_offsets by the
  //     duplicate kernel), num_rendered, overflow flag
  const int nb = (int)gridDim.x;
  const int bper = (nb + PRE_THREADS - 1) / PRE_THREADS;
  const int b0 = min(nb, tid * bper), b1 = min(nb, b0 + bper);
  uint32_t bsum = 0;
  for (int b = b0; b < b1; b++) bsum += __ldcg(a.block_total + b);
  uint32_t binc = bsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, binc, o);
    if (lane >= o) binc += n;
  }
  __syncthreads(); // s_warp_sum is reused; s_hist is complete
  if (lane == 31) s_warp_sum[warp] = binc;
  __syncthreads();
  {
    uint32_t run = binc - bsum, all = 0;
    for (int w = 0; w < PRE_THREADS / 32; w++) {
      if (w < warp) run += s_warp_sum[w];
      all += s_warp_sum[w];
    }
    if (bper == 1) { // P <= PRE_THREADS^2: this thread's one total is already in a register
      if (b0 < b1) a.block_prefix[b0] = run;
    } else {
      for (int b = b0; b < b1; b++) {
        a.block_prefix[b] = run;
        run += __ldcg(a.block_total + b);
      }
    }
    if (tid == 0) {
      a.header->num_rendered = all;
      a.header->overflow = (int64_t)all > a.capacity ? 1u : 0u;
      // straight into the caller's pinned host slot: a 4-byte copy-engine transfer between two
      // kernels costs ~12 us of stream time on B200, a posted store costs nothing
      // No system-scope fence after it: the store is posted to the host at once, while a
      // __threadfence_system() here held the whole CTA at its next barrier for the PCIe round trip
      // (3-4 us of every frame, measured with IGS_PRE_TRACE); the host polls this slot.
      if (a.num_rendered_host) *(volatile int32_t*)a.num_rendered_host = (int32_t)all;
    }
  }
  // (3) tile counts -> ranges and cursors
  const int per = (T + PRE_THREADS - 1) / PRE_THREADS;
  const int t0 = min(T, tid * per), t1 = min(T, t0 + per);
  uint32_t sum = 0, cmax = 0;
  for (int t = t0; t < t1; t++) { const uint32_t c = TILE_CNT(t); sum += c; cmax = max(cmax, c); }
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __shared__ uint32_t s_bucket[ORDER_BUCKETS];
  __shared__ uint32_t s_max[PRE_THREADS / 32];
  cmax = __reduce_max_sync(0xffffffffu, cmax);
  __syncthreads(); // s_warp_sum is reused
  if (lane == 31) s_warp_sum[warp] = inc;
  if (lane == 0) s_max[warp] = cmax;
  for (int b = tid; b < ORDER_BUCKETS; b += PRE_THREADS) s_bucket[b] = 0u;
  __syncthreads();
  uint32_t run = inc - sum;
  for (int w = 0; w < warp; w++) run += s_warp_sum[w];
#pragma unroll
  for (int w = 0; w < PRE_THREADS / 32; w++) cmax = max(cmax, s_max[w]);
  // ---- launch order of the render kernels: tiles with the longest lists first (their warps
  // are the critical path), by one counting pass over ORDER_BUCKETS count classes ----
  const float qs = (float)(ORDER_BUCKETS - 1) / (float)max(cmax, 1u);
#define ORDER_BUCKET(c) ((ORDER_BUCKETS - 1) - min(ORDER_BUCKETS - 1, (int)((float)(c) * qs)))
  for (int t = t0; t < t1; t++) {
    const uint32_t c = TILE_CNT(t);
    a.tile_cursor[(size_t)t * CNT_STRIDE] = run;
    a.ranges[t] = c ? make_uint2(run, run + c) : make_uint2(0u, 0u);
    run += c;
    atomicAdd(&s_bucket[ORDER_BUCKET(c)], 1u);
  }
  __syncthreads();
  { // exclusive scan of the bucket sizes (ORDER_BUCKETS == 2 * PRE_THREADS: two per thread)
    const uint32_t c0 = s_bucket[2 * tid], c1 = s_bucket[2 * tid + 1];
    uint32_t in2 = c0 + c1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t n = __shfl_up_sync(0xffffffffu, in2, o);
      if (lane >= o) in2 += n;
    }
    if (lane == 31) s_max[warp] = in2;
    __syncthreads();
    uint32_t base = in2 - (c0 + c1);
    for (int w = 0; w < warp; w++) base += s_max[w];
    s_bucket[2 * tid] = base;
    s_bucket[2 * tid + 1] = base + c0;
  }
  __syncthreads();
  for (int t = t0; t < t1; t++) {
    const uint32_t slot = atomicAdd(&s_bucket[ORDER_BUCKET(TILE_CNT(t))], 1u);
    a.tile_order[slot] = (uint32_t)t;
  }
  if (!HIST_SMEM) { // the counts were read from global memory up to here
    __syncthreads();
    for (int t = tid; t < T; t += PRE_THREADS) a.tile_count[(size_t)t * CNT_STRIDE] = 0u;
  }
#undef TILE_CNT
  PRE_MARK(6);
#undef ORDER_BUCKET
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ view,
                                    uint8_t* __restrict__ present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  float m[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m[k] = __ldg(view + k);
  const float vz = xform_row(m, 2, __ldg(means3D + 3 * idx), __ldg(means3D + 3 * idx + 1), __ldg(means3D + 3 * idx + 2));
  present[idx] = !(vz <= NEAR_Z);
}

void launch_preprocess(const PreArgs& a, cudaStream_t s) {
  if (a.P <= 0) return;
  const int blocks = (a.P + PRE_THREADS - 1) / PRE_THREADS;
  const int T = a.grid_x * a.grid_y;
  const size_t hist = (size_t)T * sizeof(uint32_t);
  if (T <= HIST_SMEM_MAX_TILES) {
    if (a.shs) launch_after(preprocess_kernel<true, true>, dim3(blocks), dim3(PRE_THREADS), hist, s, a);
    else launch_after(preprocess_kernel<false, true>, dim3(blocks), dim3(PRE_THREADS), hist, s, a);
  } else { // images beyond ~1700x1700: count straight into global memory
    if (a.shs) launch_after(preprocess_kernel<true, false>, dim3(blocks), dim3(PRE_THREADS), 0, s, a);
    else launch_after(preprocess_kernel<false, false>, dim3(blocks), dim3(PRE_THREADS), 0, s, a);
  }
  count_launch();
}

#ifdef IGS_PRE_TRACE
extern "C" int igs_pre_trace_read(unsigned long long* host, int n) {
  return (int)cudaMemcpyFromSymbol(host, g_pre_trace, sizeof(unsigned long long) * n);
}
#endif

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  if (P <= 0) return;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
  count_launch();
}

} // namespace igs
