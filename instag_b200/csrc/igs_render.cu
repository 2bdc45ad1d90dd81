// Stage 4 of the forward path: per-tile front-to-back alpha compositing
// (reference renderCUDA, forward.cu:321-456).
//
// One CTA per 16x16 tile, one thread per pixel.  Differences from the
// reference, none of which changes a pixel's value:
//  * each 64-byte splat record is gathered ONCE per tile into shared memory
//    (the reference re-fetches colour/normal/depth/extra from global memory for
//    every accepted pixel-Gaussian pair, forward.cu:426-431);
//  * a warp covers an 8x4 pixel block (not 16x2) and leaves the batch loop as
//    soon as its 32 pixels are done (warp-level early termination; the
//    reference only votes per CTA, forward.cu:380);
//  * ED (extra channels) is a template parameter: no runtime-indexed arrays.
//
// Bound: FP32 issue.  Per evaluated pair ~15 flop + 1 MUFU.EX2, per blended
// pair +17 flop.  DRAM floor: 64 B per instance + 40 B per pixel.
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int BATCH = TILE_PIX; // Gaussians staged per round

// pixel owned by a thread: warp w covers the 8x4 block (w&1, w>>1)
__device__ __forceinline__ void thread_pixel(int tid, int& lx, int& ly) {
  const int w = tid >> 5, l = tid & 31;
  lx = (w & 1) * 8 + (l & 7);
  ly = (w >> 1) * 4 + (l >> 3);
}

template <int ED_T> // 0, 1, or -1 = generic (runtime ED up to IGS_MAX_EXTRA_DIMS)
__global__ void __launch_bounds__(TILE_PIX) render_kernel(RenderArgs a) {
  __shared__ float4 s_r0[BATCH]; // conic.xyz, opacity
  __shared__ float4 s_r1[BATCH]; // x, y, depth, extra0
  __shared__ float4 s_r2[BATCH]; // r, g, b, nx
  __shared__ float4 s_r3[BATCH]; // ny, nz, cut, -
  __shared__ uint32_t s_id[ED_T < 0 ? BATCH : 1];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int lx, ly;
  thread_pixel(tid, lx, ly);
  const int px = blockIdx.x * TILE + lx, py = blockIdx.y * TILE + ly;
  const bool inside = px < a.W && py < a.H;
  const float pxf = (float)px, pyf = (float)py;
  // the 8x4 pixel rectangle of this warp (culling)
  const float wx0 = (float)(blockIdx.x * TILE + (warp & 1) * 8), wx1 = wx0 + 7.f;
  const float wy0 = (float)(blockIdx.y * TILE + (warp >> 1) * 4), wy1 = wy0 + 3.f;
  const uint2 range = a.ranges[blockIdx.y * a.grid_x + blockIdx.x];
  const int total = (int)(range.y - range.x);
  const int ED = ED_T < 0 ? a.ED : ED_T;

  float T = 1.0f;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dp = 0.f, N0 = 0.f, N1 = 0.f, N2 = 0.f, E0 = 0.f;
  float Ex[ED_T < 0 ? IGS_MAX_EXTRA_DIMS : 1];
  if (ED_T < 0) {
#pragma unroll
    for (int c = 0; c < IGS_MAX_EXTRA_DIMS; c++) Ex[c] = 0.f;
  }
  uint32_t last_contributor = 0;
  bool done = !inside;

  for (int base = 0; base < total; base += BATCH) {
    // whole CTA finished?  (reference forward.cu:380)
    if (__syncthreads_and(done)) break;
    const int n = min(BATCH, total - base);
    if (tid < n) {
      const uint32_t id = __ldg(a.point_list + range.x + base + tid);
      const float4* rec = a.splat + (size_t)id * SPLAT_F4;
      s_r0[tid] = __ldg(rec);
      s_r1[tid] = __ldg(rec + 1);
      s_r2[tid] = __ldg(rec + 2);
      s_r3[tid] = __ldg(rec + 3);
      if (ED_T < 0) s_id[tid] = id;
    }
    __syncthreads();
    // warp-level early termination + warp-level culling: 32 splats at a time, each lane
    // tests one against the warp's pixel rectangle; only survivors are composited
    for (int cb = 0; cb < n; cb += 32) {
      if (__all_sync(0xffffffffu, done)) break;
      bool hit = false;
      if (cb + lane < n) {
        const float4 q0 = s_r0[cb + lane];
        const float4 q1 = s_r1[cb + lane];
        hit = splat_may_touch(q0.x, q0.y, q0.z, q1.x, q1.y, s_r3[cb + lane].z, wx0, wx1, wy0, wy1);
      }
      unsigned m = __ballot_sync(0xffffffffu, hit);
      while (m) {
        const int j = cb + __ffs(m) - 1;
        m &= m - 1;
        if (done) continue;
        const float4 r1 = s_r1[j];
        const float4 r0 = s_r0[j];
        const float dx = __fsub_rn(r1.x, pxf), dy = __fsub_rn(r1.y, pyf);
        const float power = gauss_power(r0.x, r0.y, r0.z, dx, dy);
        if (power > 0.0f) continue;
        const float alpha = fminf(0.99f, __fmul_rn(r0.w, expf(power)));
        if (alpha < ALPHA_MIN) continue;
        const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha));
        if (test_T < T_MIN) { done = true; continue; }
        // X += f*alpha*T  ->  fma(f*alpha, T, X)   (reference forward.cu:425-431)
        const float4 r2 = s_r2[j];
        const float4 r3 = s_r3[j];
        C0 = __fmaf_rn(__fmul_rn(r2.x, alpha), T, C0);
        C1 = __fmaf_rn(__fmul_rn(r2.y, alpha), T, C1);
        C2 = __fmaf_rn(__fmul_rn(r2.z, alpha), T, C2);
        Dp = __fmaf_rn(__fmul_rn(r1.z, alpha), T, Dp);
        N0 = __fmaf_rn(__fmul_rn(r2.w, alpha), T, N0);
        N1 = __fmaf_rn(__fmul_rn(r3.x, alpha), T, N1);
        N2 = __fmaf_rn(__fmul_rn(r3.y, alpha), T, N2);
        if (ED_T == 1) E0 = __fmaf_rn(__fmul_rn(r1.w, alpha), T, E0);
        if (ED_T < 0) {
          const float* e = a.extra_attrs + (size_t)s_id[j] * ED;
          for (int c = 0; c < ED; c++) Ex[c] = __fmaf_rn(__fmul_rn(__ldg(e + c), alpha), T, Ex[c]);
        }
        T = test_T;
        last_contributor = (uint32_t)(base + j + 1);
      }
    }
  }

  if (inside) {
    const size_t HW = (size_t)a.H * a.W, pix = (size_t)py * a.W + px;
    a.out_alpha[pix] = __fsub_rn(1.f, T);
    a.n_contrib[pix] = last_contributor;
    a.out_color[pix] = __fmaf_rn(T, __ldg(a.bg + 0), C0);
    a.out_color[HW + pix] = __fmaf_rn(T, __ldg(a.bg + 1), C1);
    a.out_color[2 * HW + pix] = __fmaf_rn(T, __ldg(a.bg + 2), C2);
    a.out_depth[pix] = Dp;
    a.out_norm[pix] = N0;
    a.out_norm[HW + pix] = N1;
    a.out_norm[2 * HW + pix] = N2;
    if (ED_T == 1) a.out_extra[pix] = E0;
    if (ED_T < 0)
      for (int c = 0; c < ED; c++) a.out_extra[c * HW + pix] = Ex[c];
  }
}

void launch_render(const RenderArgs& a, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return;
  const dim3 grid(a.grid_x, a.grid_y);
  if (a.ED == 0) render_kernel<0><<<grid, TILE_PIX, 0, s>>>(a);
  else if (a.ED == 1) render_kernel<1><<<grid, TILE_PIX, 0, s>>>(a);
  else render_kernel<-1><<<grid, TILE_PIX, 0, s>>>(a);
  count_launch();
}

} // namespace igs
