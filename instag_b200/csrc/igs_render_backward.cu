// Backward of the compositing stage (reference renderCUDA, backward.cu:473-712).
//
// One CTA per tile, one thread per pixel, back-to-front traversal, T recovered
// by division exactly as the reference does (backward.cu:612).  Where the
// reference issues 13+ED global float atomics per accepted pixel-Gaussian pair,
// this kernel first reduces the 14 partial gradients over the 32 pixels of a
// warp with a TRANSPOSING butterfly (16 shuffles for 16 values instead of
// 16*5), after which lane 2k holds the warp total of value k and the warp
// issues one coalesced 14-lane red.global.add into the Gaussian's 64-byte
// accumulator row.  Gaussians no lane of the warp touches are skipped with one
// ballot; the tail of the list behind every pixel's last contributor is skipped
// without being loaded.
//
// Each warp first culls the staged splats against its own 8x4 pixel rectangle
// (splat_may_touch, igs_common.cuh), 32 at a time, and only walks the survivors.
//
// Bound: FP32 issue + L2 atomics.  Per blended pair ~110 flop.
#include <cstdlib>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {


// Sum 16 per-lane values over the warp; lane l returns the total of v[l>>1].
__device__ __forceinline__ float warp_transpose_reduce16(const float (&v)[16], int lane) {
  constexpr unsigned FULL = 0xffffffffu;
  float w8[8], w4[4], w2[2];
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4, b1 = lane & 2;
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const float send = b4 ? v[i] : v[i + 8], keep = b4 ? v[i + 8] : v[i];
    w8[i] = keep + __shfl_xor_sync(FULL, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float send = b3 ? w8[i] : w8[i + 4], keep = b3 ? w8[i + 4] : w8[i];
    w4[i] = keep + __shfl_xor_sync(FULL, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const float send = b2 ? w4[i] : w4[i + 2], keep = b2 ? w4[i + 2] : w4[i];
    w2[i] = keep + __shfl_xor_sync(FULL, send, 4);
  }
  const float send = b1 ? w2[0] : w2[1], keep = b1 ? w2[1] : w2[0];
  const float w1 = keep + __shfl_xor_sync(FULL, send, 2);
  return w1 + __shfl_xor_sync(FULL, w1, 1);
}

constexpr int BW = 4;        // warps per CTA: one per SM sub-partition
constexpr int BT = BW * 32;

__device__ __forceinline__ float4 lds128b(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128b(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float rcp_approx(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// Per-pixel state of the reverse walk.  The reference keeps one running "colour
// behind me" per channel (accum_rec*, last_*, backward.cu:640-675) and sums
// (c - accum)*dL_dpixel over the channels.  Both are linear, so ONE scalar recurrence on
//   u = <attributes, dL_dpixel> + dL_dalpha_pixel
// gives the same sum:  A <- last_alpha*u_last + (1-last_alpha)*A ;  dL_dalpha = (u - A)*T + bg term.
struct PixelBack {
  float T, A, u_last, last_alpha;
};

template <int ED_T> // 0, 1, or -1 = generic
__global__ void __launch_bounds__(BT) render_backward_kernel(BwdRenderArgs a) {
  __shared__ float4 s_rec[BW][4][32];
  __shared__ uint32_t s_id[BW][32];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int num_units = a.grid_x * a.grid_y * 8;
  const int ED = ED_T < 0 ? a.ED : ED_T;
  const size_t HW = (size_t)a.H * a.W;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(&s_rec[warp][0][0]);
  const uint32_t smine = sbase + lane * 16;
  const uint32_t sidb = (uint32_t)__cvta_generic_to_shared(&s_id[warp][0]);
  const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);

 // persistent: every warp pulls (tile, 8x4 block) units from one ticket, heaviest tiles first
 for (;;) {
  int unit = 0;
  if (lane == 0) unit = (int)atomicAdd(a.ticket, 1u);
  unit = __shfl_sync(0xffffffffu, unit, 0);
  if (unit >= num_units) break;
  // consecutive tickets (= the four warps of a CTA, one SM) go to four DIFFERENT tiles, so that
  // an SM's share of the frame is a sum of independent unit costs (groups of 4 tiles x 8 blocks)
  const int num_tiles = a.grid_x * a.grid_y;
  int slot = ((unit >> 5) << 2) + (unit & 3), wb = (unit >> 2) & 7;
  if (((unit >> 5) << 2) + 4 > num_tiles) { slot = unit >> 3; wb = unit & 7; } // ragged last group
  const int tile = (int)__ldg(a.tile_order + slot);
  const int tx = tile % a.grid_x, ty = tile / a.grid_x;
  const int bx = tx * TILE + (wb & 1) * 8, by = ty * TILE + (wb >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < a.W && py < a.H;
  const float pxf = (float)px, pyf = (float)py;
  const float wx0 = (float)bx, wx1 = wx0 + 7.f, wy0 = (float)by, wy1 = wy0 + 3.f;
  const size_t pix = (size_t)py * a.W + px;
  const uint2 range = __ldg(a.ranges + tile);

  const int last_contributor = inside ? (int)a.n_contrib[pix] : 0;
  // nothing behind the last contributor of any of the warp's pixels matters
  const int wmax = (int)__reduce_max_sync(0xffffffffu, (unsigned)last_contributor);
  if (wmax == 0) continue;
  const float T_final = inside ? __fsub_rn(1.f, a.out_alpha[pix]) : 0.f;
  float dpc0 = 0, dpc1 = 0, dpc2 = 0, dpd = 0, dpn0 = 0, dpn1 = 0, dpn2 = 0, dpa = 0, dpe0 = 0;
  if (inside) {
    dpc0 = a.dL_dcolor[pix]; dpc1 = a.dL_dcolor[HW + pix]; dpc2 = a.dL_dcolor[2 * HW + pix];
    dpd = a.dL_ddepth[pix];
    dpn0 = a.dL_dnorm[pix]; dpn1 = a.dL_dnorm[HW + pix]; dpn2 = a.dL_dnorm[2 * HW + pix];
    if (a.normalize) {
      // backward of n_hat = n / max(|n|, eps): (g - n_hat <n_hat, g>) / |n|, and g / eps where the
      // clamp is active (what autograd derives for F.normalize, DGR diff_gauss/__init__.py:48)
      const float len = a.norm_len[pix];
      if (len > 1e-12f) {
        const float h0 = a.out_norm[pix], h1 = a.out_norm[HW + pix], h2 = a.out_norm[2 * HW + pix];
        const float d = h0 * dpn0 + h1 * dpn1 + h2 * dpn2, inv = 1.f / len;
        dpn0 = (dpn0 - h0 * d) * inv; dpn1 = (dpn1 - h1 * d) * inv; dpn2 = (dpn2 - h2 * d) * inv;
      } else {
        dpn0 *= 1e12f; dpn1 *= 1e12f; dpn2 *= 1e12f;
      }
    }
    dpa = a.dL_dalpha[pix];
    if (ED >= 1) dpe0 = a.dL_dextra[pix];
  }
  PixelBack st{T_final, 0.f, 0.f, 0.f};
  const float bg_dot = bg0 * dpc0 + bg1 * dpc1 + bg2 * dpc2;
  const float nbgT = -T_final * bg_dot;

  // entry e of the reverse walk is list position wmax-1-e (= its `contributor` index)
  const uint32_t* __restrict__ list = a.point_list + range.x;
  uint32_t id_cur = 0, id_nxt = 0;
  float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
  if (lane < wmax) id_cur = __ldg(list + (wmax - 1 - lane));
  if (32 + lane < wmax) id_nxt = __ldg(list + (wmax - 33 - lane));
  if (lane < wmax) {
    q0 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4);
    q1 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4 + 1);
  }

  for (int cb = 0; cb < wmax; cb += 32) {
    const bool live = cb + lane < wmax;
    const bool hit = live && splat_may_touch(q0.x, q0.y, q0.z, q1.x, q1.y, 2.f * __logf(255.f * q0.w),
                                             wx0, wx1, wy0, wy1);
    unsigned m = __ballot_sync(0xffffffffu, hit);
    const uint32_t id_mine = id_cur;
    const float4 c0 = q0, c1 = q1;
    id_cur = id_nxt;
    if (cb + 32 + lane < wmax) {
      q0 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4);
      q1 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4 + 1);
    }
    if (cb + 64 + lane < wmax) id_nxt = __ldg(list + (wmax - 65 - cb - lane));
    if (m == 0) continue;
    __syncwarp();
    if (hit) {
      const float4* rec = a.splat + (size_t)id_mine * SPLAT_F4;
      const float4 c2 = __ldg(rec + 2), c3 = __ldg(rec + 3);
      sts128b(smine, c0);
      sts128b(smine + 512, c1);
      sts128b(smine + 1024, c2);
      sts128b(smine + 1536, c3);
      s_id[warp][lane] = id_mine;
    }
    __syncwarp();
    while (m) {
      const int j = __ffs(m) - 1;
      m &= m - 1;
      const int contributor = wmax - 1 - cb - j;
      const float4 r1 = lds128b(sbase + 512 + j * 16);
      const float4 r0 = lds128b(sbase + j * 16);
      const float dx = __fsub_rn(r1.x, pxf), dy = __fsub_rn(r1.y, pyf);
      const float power = gauss_power(r0.x, r0.y, r0.z, dx, dy);
      const float G = expf(power);
      const float alpha = fminf(0.99f, __fmul_rn(r0.w, G));
      const bool valid = contributor < last_contributor && !(power > 0.0f) && !(alpha < ALPHA_MIN);
      if (!__any_sync(0xffffffffu, valid)) continue;
      // per-pixel partials; lanes that do not blend this splat contribute exact zeros
      float weight = 0.f, gp = 0.f; // alpha*T ; G * dL/dalpha
      if (valid) {
        const float4 r2 = lds128b(sbase + 1024 + j * 16);
        const float4 r3 = lds128b(sbase + 1536 + j * 16);
        const float rc = rcp_approx(__fsub_rn(1.f, alpha));
        st.T = st.T * rc;                              // reference backward.cu:612 (T / (1 - alpha))
        weight = alpha * st.T;
        float u = dpa;
        u = fmaf(r2.x, dpc0, u); u = fmaf(r2.y, dpc1, u); u = fmaf(r2.z, dpc2, u);
        u = fmaf(r1.z, dpd, u);
        u = fmaf(r2.w, dpn0, u); u = fmaf(r3.x, dpn1, u); u = fmaf(r3.y, dpn2, u);
        if (ED >= 1) u = fmaf(r1.w, dpe0, u);
        if (ED_T < 0) {
          const float* e = a.extra_attrs + (size_t)s_id[warp][j] * ED;
          for (int c = 1; c < ED; c++) u = fmaf(__ldg(e + c), a.dL_dextra[c * HW + pix], u);
        }
        st.A = fmaf(st.last_alpha, st.u_last - st.A, st.A);
        st.u_last = u;
        st.last_alpha = alpha;
        // + background term, reference backward.cu:684-687
        const float dL_dalpha = fmaf(u - st.A, st.T, nbgT * rc);
        gp = G * dL_dalpha;
      }
      // moments of gp about the splat centre: the per-Gaussian kernel turns them into
      // dL/dmean2D, dL/dconic and dL/dopacity (they only differ by per-Gaussian factors)
      float v[16];
      const float gx = gp * dx, gy = gp * dy;
      v[0] = gx; v[1] = gy; v[2] = gx * dx; v[3] = gx * dy; v[4] = gy * dy; v[5] = gp;
      v[6] = weight * dpc0; v[7] = weight * dpc1; v[8] = weight * dpc2; v[9] = weight * dpd;
      v[10] = weight * dpn0; v[11] = weight * dpn1; v[12] = weight * dpn2;
      v[13] = ED >= 1 ? weight * dpe0 : 0.f;
      v[14] = 0.f; v[15] = 0.f;
      const float tot = warp_transpose_reduce16(v, lane);
      uint32_t id;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(id) : "r"(sidb + j * 4));
      if (!(lane & 1) && (lane >> 1) < 13 + (ED >= 1 ? 1 : 0))
        atomicAdd(a.accum + (size_t)id * ACC_F + (lane >> 1), tot);
      if (ED_T < 0) { // channels 1.. : plain warp reduction into the side array
        for (int c = 1; c < ED; c++) {
          float gsum = valid ? weight * a.dL_dextra[c * HW + pix] : 0.f;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
          if (lane == 0) atomicAdd(a.accum_extra + (size_t)id * (ED - 1) + (c - 1), gsum);
        }
      }
    }
  }
  __syncwarp(); // the next unit reuses this warp's shared-memory slots
 } // unit loop
}

template <class K>
static int resident_ctas_bwd(K kernel, int threads) {
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
  // 6 x 4 warps per SM measured best on B200 (sweep 2..8 in profiles/r01_v3_notes.md): fewer
  // resident warps leave a deeper ticket queue (better balance), more hide more latency
  per_sm = min(per_sm, 6);
  if (const char* e = getenv("IGS_CTAS_PER_SM")) per_sm = max(1, atoi(e)); // experiments only
  return (sms > 0 ? sms : 148) * (per_sm > 0 ? per_sm : 1);
}

void launch_render_backward(const BwdRenderArgs& a, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return;
  static const int cap0 = resident_ctas_bwd(render_backward_kernel<0>, BT),
                   cap1 = resident_ctas_bwd(render_backward_kernel<1>, BT),
                   capx = resident_ctas_bwd(render_backward_kernel<-1>, BT);
  const int units = a.grid_x * a.grid_y * 8;
  const int want = (units + BW - 1) / BW;
  if (a.ED == 0) render_backward_kernel<0><<<min(want, cap0), BT, 0, s>>>(a);
  else if (a.ED == 1) render_backward_kernel<1><<<min(want, cap1), BT, 0, s>>>(a);
  else render_backward_kernel<-1><<<min(want, capx), BT, 0, s>>>(a);
  count_launch();
}

} // namespace igs
