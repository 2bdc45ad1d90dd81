// Backward of the compositing stage (reference renderCUDA, backward.cu:473-712).
//
// One warp per 8x4 pixel block, one lane per pixel, back-to-front traversal, T
// recovered by division exactly as the reference does (backward.cu:612).  Where
// the reference issues 13+ED global float atomics per accepted pixel-Gaussian
// pair, the sum over the warp's 32 pixels is done here as a small matrix product.
// Per blended (pixel, splat) pair a lane produces just TWO scalars, the blend
// weight w = alpha*T and g = G*dL/dalpha; every per-Gaussian gradient is linear in them:
//     dL/dattr_k(s)  = sum_p w[s][p] * dL_dpixel_k[p]                (8 channels)
//     moments(s)     = sum_p g[s][p] * {1, lx, lx^2, lx*ly, ly^2, ly}[p]   (lx, ly = pixel
//                      offset inside the block; shifted to the splat centre afterwards)
// i.e. [16 splats x 32 pixels] x [32 pixels x 8], with the right-hand matrix constant for
// the whole unit.  w and g are parked in shared memory; every 16 parked splats the warp
// runs the two products on the tensor cores (mma.sync m16n8k8 TF32, split into hi/lo
// terms so that the result keeps fp32 accuracy: 3 terms against dL_dpixel, 2 against the
// small-integer moment basis, which TF32 holds exactly) and adds the fragment to the
// Gaussians' 64-byte accumulator rows with 8-byte vector REDs: 14 atomics per
// (warp, splat) as before, but ~9 instructions per splat instead of the ~80 of a
// shuffle butterfly over 16 partial products.  Gaussians no lane of the warp touches
// are skipped with one ballot; the tail of the list behind every pixel's last
// contributor is skipped without being loaded.
//
// Each warp first culls the staged splats against its own 8x4 pixel rectangle
// (splat_may_touch, igs_common.cuh), 32 at a time, and only walks the survivors.
//
// Bound: FP32 issue + L2 atomics.  Per blended pair ~110 flop.
#include <cstdlib>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {


__device__ __forceinline__ uint32_t tf32_hi(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
// D += A(16x8, row) * B(8x8, col), TF32 inputs, fp32 accumulate
__device__ __forceinline__ void mma_tf32(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void red_add_v2(float* addr, float x, float y) {
  asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(x), "f"(y) : "memory");
}

constexpr int PEND = 16;     // parked splats per matrix product (the M of m16n8k8)

#ifndef IGS_BWD_MIN_CTAS
#define IGS_BWD_MIN_CTAS 6
#endif
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
// 4 bytes global -> shared with no register in between (LDGSTS): the list ids two chunks ahead
__device__ __forceinline__ void copy4_async(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void copy16_async_b(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void copies_wait_planes_b() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ void copies_commit_b() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void copies_wait_b() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
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
__global__ void __launch_bounds__(BT, IGS_BWD_MIN_CTAS) render_backward_kernel(BwdRenderArgs a) {
  // Everything a warp owns sits in ONE block, so that one base register and immediate offsets address it
  // all (separate arrays cost a base register each, and the kernel lives at the 80-register line):
  //   rec  [4][32] float4  the staged records of the chunk's survivors, one 512-byte plane per 16 bytes
  //   wg   [2][PEND][32]   parked blend weights [0] and g = G*dL/dalpha [1] of up to PEND splats; pixel p of
  //                        slot s sits in column p ^ ((s & 1) << 4) so that both the scalar stores and the
  //                        16-byte fragment loads are bank-conflict free (the swizzle works on address
  //                        bits: the block is 128-byte aligned and wg starts at a multiple of 128)
  //   bf   [4][32] float4  right-hand fragments, one float4 per lane: [0..1] dL_dpixel TF32 hi, [2..3] lo
  //   id   [32]            Gaussian ids of the staged records
  struct __align__(128) WarpBlock {
    float4 rec[4][32];
    float wg[2][PEND][32];
    float4 bf[4][32];
    uint32_t id[32];
  };
  constexpr uint32_t WB_WG = 2048, WB_BF = 6144, WB_ID = 8192;
  static_assert(sizeof(WarpBlock) == 8320 && sizeof(WarpBlock) % 128 == 0, "offsets below");
  __shared__ WarpBlock s_w[BW];
  // ids of the chunk after next, copied in asynchronously (double buffered; the buffers of a warp are
  // told apart by address bit 7).  In a register the prefetched id was spilled to local memory at 80
  // registers, and the spill store waited for the load it had just issued: 5 % of the kernel's stall
  // samples sat on that one instruction.
  __shared__ __align__(256) uint32_t s_idn[BW][2][32];
  // 512-byte stride per warp (PEND = 16 records used): the fill level of a warp's queue is then
  // (address & 511) >> 4, with no base address to keep in (or recompute into) a register
  __shared__ __align__(512) float4 s_pend[BW][32]; // {mean.x, mean.y, gaussian id bits, -}
  // the moment basis (the same for every warp and unit)
  __shared__ float4 s_bm[2][32];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (a.prof.early) pdl_launch_dependents();
  pdl_wait(); // whatever ran before on the stream (the forward, the caller's loss kernels) is complete
  if (threadIdx.x == 0) stamp_begun(a.prof, ST_BWD_START);
  // a frame that outgrew its instance workspaces was not rendered (a caller that launches the backward
  // without having read the count -- recorded steps, run-ahead loops -- finds out afterwards)
  if (__ldg(a.info + 1)) {
    if (lane == 0) stamp_done(a.prof, ST_RBWD);
    return;
  }
  const int num_units = a.grid_x * a.grid_y * 8;
  const int ED = ED_T < 0 ? a.ED : ED_T;
  const size_t HW = (size_t)a.H * a.W;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(&s_w[warp]);
  const uint32_t smine = sbase + lane * 16;
  const uint32_t sidb = sbase + WB_ID;
  const uint32_t sidn = (uint32_t)__cvta_generic_to_shared(&s_idn[warp][0][lane]); // + 128: the other buffer
  const float bg0 = __ldg(a.bg), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);
  // mma fragment coordinates: this lane owns rows fg, fg+8 (splat slots) of A and C, and the
  // k-columns ft, ft+4; k-step 2q+h of the product covers pixels 16q + 4*ft + {2h, 2h+1}
  const int fg = lane >> 2, ft = lane & 3;
  const uint32_t swg = sbase + WB_WG;
  const uint32_t sfrag = swg + fg * 128 + ft * 16; // + (q ^ (fg & 1)) * 64: the slot's swizzle (+8 slots: +1024, g: +2048)
  // moment basis of the pixels of this lane's k-columns, row fg of the right-hand matrix:
  // {1, lx, lx^2, lx*ly, ly^2, ly, 0, 0}; small integers, exact in TF32
  if (warp == 0) {
#pragma unroll
    for (int q = 0; q < 2; q++) {
      float v[4];
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const float lx = (float)(4 * (ft & 1) + i), ly = (float)(2 * q + (ft >> 1));
        v[i] = fg == 0 ? 1.f : fg == 1 ? lx : fg == 2 ? lx * lx : fg == 3 ? lx * ly : fg == 4 ? ly * ly : fg == 5 ? ly : 0.f;
      }
      s_bm[q][lane] = make_float4(v[0], v[1], v[2], v[3]);
    }
  }
  __syncthreads();
  const uint32_t sbf = smine + WB_BF;
  const uint32_t sbm = (uint32_t)__cvta_generic_to_shared(&s_bm[0][lane]);
  const uint32_t spend = (uint32_t)__cvta_generic_to_shared(&s_pend[warp][0]);

 // persistent: every warp pulls (tile, 8x4 block) units from one ticket, heaviest tiles first
 // (the first unit of every warp is its own number, as in the forward)
 auto next_unit = [&]() {
   int u = 0;
   if (lane == 0) u = (int)gridDim.x * BW + (int)atomicAdd(a.ticket, 1u);
   return __shfl_sync(0xffffffffu, u, 0);
 };
 for (int unit = (int)blockIdx.x * BW + warp; unit < num_units; unit = next_unit()) {
  // consecutive tickets (= the four warps of a CTA, one SM) go to four DIFFERENT tiles, so that
  // an SM's share of the frame is a sum of independent unit costs (groups of 4 tiles x 8 blocks)
  const int num_tiles = a.grid_x * a.grid_y;
  int slot = ((unit >> 5) << 2) + (unit & 3), wb = (unit >> 2) & 7;
  if (((unit >> 5) << 2) + 4 > num_tiles) { slot = unit >> 3; wb = unit & 7; } // ragged last group
  const uint4 entry = __ldg(a.tile_order + slot);
  const int tile = (int)entry.x;
  const int tx = tile % a.grid_x, ty = tile / a.grid_x;
  const int bx = tx * TILE + (wb & 1) * 8, by = ty * TILE + (wb >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < a.W && py < a.H;
  const float pxf = (float)px, pyf = (float)py;
  const float wx0 = (float)bx, wx1 = wx0 + 7.f, wy0 = (float)by, wy1 = wy0 + 3.f;
  const size_t pix = (size_t)py * a.W + px;
  const uint2 range = make_uint2(entry.y, entry.z);

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
  // right-hand matrix of the attribute product: dL_dpixel channel fg at this lane's k-columns,
  // transposed through shared memory once per unit and split into TF32 hi + lo
  {
    float* sD = &s_w[warp].wg[0][0][0];
    sD[lane] = dpc0; sD[32 + lane] = dpc1; sD[64 + lane] = dpc2; sD[96 + lane] = dpd;
    sD[128 + lane] = dpn0; sD[160 + lane] = dpn1; sD[192 + lane] = dpn2; sD[224 + lane] = dpe0;
    __syncwarp();
    const float4 d0 = *reinterpret_cast<const float4*>(sD + fg * 32 + 4 * ft);
    const float4 d1 = *reinterpret_cast<const float4*>(sD + fg * 32 + 16 + 4 * ft);
    __syncwarp();
    const float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
      hi[i] = __uint_as_float(tf32_hi(d[i]));
      lo[i] = d[i] - hi[i];
    }
    sts128b(sbf, make_float4(hi[0], hi[1], hi[2], hi[3]));
    sts128b(sbf + 512, make_float4(hi[4], hi[5], hi[6], hi[7]));
    sts128b(sbf + 1024, make_float4(lo[0], lo[1], lo[2], lo[3]));
    sts128b(sbf + 1536, make_float4(lo[4], lo[5], lo[6], lo[7]));
  }
  // the two matrix products over the parked splats + atomics into the accumulator rows
  auto flush = [&](const int n) {
    __syncwarp();
    float ca[4] = {0.f, 0.f, 0.f, 0.f}, cb[4] = {0.f, 0.f, 0.f, 0.f}, cc[4] = {0.f, 0.f, 0.f, 0.f};
    float ma[4] = {0.f, 0.f, 0.f, 0.f}, mb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < 2; q++) {
      const uint32_t sq = sfrag + ((q ^ (fg & 1)) << 6);
      const float4 wA = lds128b(sq), wB = lds128b(sq + 1024);
      const float4 gA = lds128b(sq + 2048), gB = lds128b(sq + 3072);
      const float wv[8] = {wA.x, wB.x, wA.y, wB.y, wA.z, wB.z, wA.w, wB.w};
      const float gv[8] = {gA.x, gB.x, gA.y, gB.y, gA.z, gB.z, gA.w, gB.w};
      const float4 fh = lds128b(sbf + q * 512), fl = lds128b(sbf + 1024 + q * 512), fm = lds128b(sbm + q * 512);
      const uint32_t bh[4] = {__float_as_uint(fh.x), __float_as_uint(fh.y), __float_as_uint(fh.z), __float_as_uint(fh.w)};
      const uint32_t bl[4] = {__float_as_uint(fl.x), __float_as_uint(fl.y), __float_as_uint(fl.z), __float_as_uint(fl.w)};
      const uint32_t bm[4] = {__float_as_uint(fm.x), __float_as_uint(fm.y), __float_as_uint(fm.z), __float_as_uint(fm.w)};
      uint32_t wh[8], wl[8], gh[8], gl[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        wh[i] = tf32_hi(wv[i]); wl[i] = __float_as_uint(wv[i] - __uint_as_float(wh[i]));
        gh[i] = tf32_hi(gv[i]); gl[i] = __float_as_uint(gv[i] - __uint_as_float(gh[i]));
      }
      // (Measured and dropped: a third TF32 term for g -- hi + mid + lo carry all 24 bits.  The moments
      // are shifted to the splat centre afterwards, which multiplies their rounding error by
      // (distance / block size)^2: with two terms a splat ~300 px wide shows 2.4e-4 against the reference
      // in dL_dcov3D, with three 5e-5; splats of a head are a few px wide, and the term costs 5 us.)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const int b = 2 * h, o = 4 * h;
        mma_tf32(ca, wh[o], wh[o + 1], wh[o + 2], wh[o + 3], bh[b], bh[b + 1]);
        mma_tf32(cb, wl[o], wl[o + 1], wl[o + 2], wl[o + 3], bh[b], bh[b + 1]);
        mma_tf32(cc, wh[o], wh[o + 1], wh[o + 2], wh[o + 3], bl[b], bl[b + 1]);
        mma_tf32(ma, gh[o], gh[o + 1], gh[o + 2], gh[o + 3], bm[b], bm[b + 1]);
        mma_tf32(mb, gl[o], gl[o + 1], gl[o + 2], gl[o + 3], bm[b], bm[b + 1]);
      }
    }
    const float4 pA = s_pend[warp][fg], pB = s_pend[warp][fg + 8];
    float* rowA = a.accum + (size_t)__float_as_uint(pA.z) * ACC_F;
    float* rowB = a.accum + (size_t)__float_as_uint(pB.z) * ACC_F;
    const bool liveA = fg < n, liveB = fg + 8 < n;
    // attribute gradients: C[slot][channel], this lane holds channels 2ft, 2ft+1 -> accumulator 6 + channel
    if (liveA) red_add_v2(rowA + 6 + 2 * ft, ca[0] + cb[0] + cc[0], ca[1] + cb[1] + cc[1]);
    if (liveB) red_add_v2(rowB + 6 + 2 * ft, ca[2] + cb[2] + cc[2], ca[3] + cb[3] + cc[3]);
    // moments about the block origin {S0, Sx | Sxx, Sxy | Syy, Sy | 0, 0} -> about the splat
    // centre: dx = mean.x - pixel.x = ax - lx, dy = ay - ly
    const int l0 = lane & ~3;
#pragma unroll
    for (int r = 0; r < 2; r++) {
      const float m0 = ma[2 * r] + mb[2 * r], m1 = ma[2 * r + 1] + mb[2 * r + 1];
      const float S0 = __shfl_sync(0xffffffffu, m0, l0), Sx = __shfl_sync(0xffffffffu, m1, l0);
      const float Sy = __shfl_sync(0xffffffffu, m1, l0 + 2);
      const float ax = (r ? pB.x : pA.x) - wx0, ay = (r ? pB.y : pA.y) - wy0;
      float o0, o1;
      if (ft == 0) { o0 = fmaf(ax, S0, -Sx); o1 = fmaf(ay, S0, -Sy); }                 // sum g dx, sum g dy
      else if (ft == 1) {                                                             // sum g dx^2, sum g dx dy
        o0 = fmaf(ax, fmaf(ax, S0, -2.f * Sx), m0);
        o1 = fmaf(ax, fmaf(ay, S0, -Sy), fmaf(-ay, Sx, m1));
      } else { o0 = fmaf(ay, fmaf(ay, S0, -2.f * Sy), m0); o1 = S0; }                  // sum g dy^2, sum g
      if (ft < 3 && (r ? liveB : liveA)) red_add_v2((r ? rowB : rowA) + 2 * ft, o0, o1);
    }
    __syncwarp();
  };
  PixelBack st{T_final, 0.f, 0.f, 0.f};
  const float bg_dot = bg0 * dpc0 + bg1 * dpc1 + bg2 * dpc2;
  const float nbgT = -T_final * bg_dot;

  uint32_t park_at = swg + lane * 4, pend_at = spend; // this lane's word of the next free slot; its info record
  // one splat some pixel of the warp blends: per-pixel recurrences, then park w and g
  auto blend = [&](const int j, const float4 r1, const float G, const float alpha, const bool valid) {
    float weight = 0.f, gp = 0.f; // alpha*T ; G * dL/dalpha; lanes that do not blend park exact zeros
    if (valid) {
      const float4 r2 = lds128b(sbase + 1024 + j * 16);
      const float4 r3 = lds128b(sbase + 1536 + j * 16);
      const float rc = rcp_approx(__fsub_rn(1.f, alpha));
      st.T = st.T * rc;                              // reference backward.cu:612 (T / (1 - alpha))
      weight = alpha * st.T;
      float u = fmaf(r2.x, dpc0, dpa), u2 = r2.y * dpc1;
      u = fmaf(r2.z, dpc2, u); u2 = fmaf(r1.z, dpd, u2);
      u = fmaf(r2.w, dpn0, u); u2 = fmaf(r3.x, dpn1, u2);
      u = fmaf(r3.y, dpn2, u);
      if (ED >= 1) u2 = fmaf(r1.w, dpe0, u2);
      u += u2;
      if (ED_T < 0) {
        uint32_t gid;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(gid) : "r"(sidb + j * 4));
        const float* e = a.extra_attrs + (size_t)gid * ED;
        for (int c = 1; c < ED; c++) u = fmaf(__ldg(e + c), a.dL_dextra[c * HW + pix], u);
      }
      st.A = fmaf(st.last_alpha, st.u_last - st.A, st.A);
      st.u_last = u;
      st.last_alpha = alpha;
      // + background term, reference backward.cu:684-687
      const float dL_dalpha = fmaf(u - st.A, st.T, nbgT * rc);
      gp = G * dL_dalpha;
    }
    // the per-Gaussian kernel turns the moments of g about the splat centre into dL/dmean2D,
    // dL/dconic and dL/dopacity (they only differ by per-Gaussian factors)
    uint32_t id;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(id) : "r"(sidb + j * 4));
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(park_at), "f"(weight) : "memory");
    asm volatile("st.shared.f32 [%0], %1;" ::"r"(park_at + 2048), "f"(gp) : "memory");
    if (lane == 0) sts128b(pend_at, make_float4(r1.x, r1.y, __uint_as_float(id), 0.f));
    park_at = (park_at + 128) ^ 64; // next slot, its column swizzle toggles
    pend_at += 16;
    if (ED_T < 0) { // channels 1.. : plain warp reduction into the side array
      for (int c = 1; c < ED; c++) {
        float gsum = valid ? weight * a.dL_dextra[c * HW + pix] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) gsum += __shfl_xor_sync(0xffffffffu, gsum, o);
        if (lane == 0) atomicAdd(a.accum_extra + (size_t)id * (ED - 1) + (c - 1), gsum);
      }
    }
  };

  // entry e of the reverse walk is list position wmax-1-e (= its `contributor` index)
  const uint32_t* __restrict__ list = a.point_list + range.x;
  uint32_t id_cur = 0;
  uint32_t nbuf = sidn; // where the ids of the NEXT chunk land
  float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
  if (lane < wmax) id_cur = __ldg(list + (wmax - 1 - lane));
  if (32 + lane < wmax) copy4_async(nbuf, list + (wmax - 33 - lane));
  copies_commit_b();
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
    copies_wait_b(); // the next chunk's ids (issued one chunk ago; a lane reads the word it copied itself)
    // Survivors stage their record: the half the cull used goes from its registers (which the prefetch
    // below then reuses), the colour / normal half is copied global -> shared without a register.  Nobody
    // still reads the slots: the ballot above is behind every lane's last trip of the previous chunk.
    if (hit) {
      const float4* rec = a.splat + (size_t)id_mine * SPLAT_F4;
      sts128b(smine, q0);
      sts128b(smine + 512, q1);
      copy16_async_b(smine + 1024, rec + 2);
      copy16_async_b(smine + 1536, rec + 3);
      asm volatile("st.shared.u32 [%0], %1;" ::"r"(sidb + lane * 4), "r"(id_mine) : "memory");
    }
    copies_commit_b(); // group 1 of the chunk: the planes
    if (cb + 32 + lane < wmax) {
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(id_cur) : "r"(nbuf));
      q0 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4);
      q1 = __ldg(a.splat + (size_t)id_cur * SPLAT_F4 + 1);
    }
    nbuf ^= 128u;
    if (cb + 64 + lane < wmax) copy4_async(nbuf, list + (wmax - 65 - cb - lane));
    copies_commit_b(); // group 2: the ids of the chunk after next
    if (m == 0) continue;
    copies_wait_planes_b(); // all but the newest group: the planes are in, the ids may still travel
    __syncwarp();
    const int jmin = wmax - 1 - cb - last_contributor; // entry j is in front of this pixel's last contributor iff j > jmin
    // two staged splats per trip: the two alpha evaluations are independent, the recurrences run in order
    while (m) {
      const int j0 = __ffs(m) - 1;
      m &= m - 1;
      const bool two = m != 0;
      const int j1 = two ? __ffs(m) - 1 : j0;
      m &= m - 1;
      const float4 a1 = lds128b(sbase + 512 + j0 * 16), a0 = lds128b(sbase + j0 * 16);
      const float4 b1 = lds128b(sbase + 512 + j1 * 16), b0 = lds128b(sbase + j1 * 16);
      const float dxa = __fsub_rn(a1.x, pxf), dya = __fsub_rn(a1.y, pyf);
      const float dxb = __fsub_rn(b1.x, pxf), dyb = __fsub_rn(b1.y, pyf);
      const float pa = gauss_power(a0.x, a0.y, a0.z, dxa, dya), pb = gauss_power(b0.x, b0.y, b0.z, dxb, dyb);
      const float Ga = expf(pa), Gb = expf(pb);
      const float ala = fminf(0.99f, __fmul_rn(a0.w, Ga)), alb = fminf(0.99f, __fmul_rn(b0.w, Gb));
      const bool va = j0 > jmin && !(pa > 0.0f) && !(ala < ALPHA_MIN);
      const bool vb = two && j1 > jmin && !(pb > 0.0f) && !(alb < ALPHA_MIN);
      const bool anya = __any_sync(0xffffffffu, va), anyb = __any_sync(0xffffffffu, vb);
      if (anya) blend(j0, a1, Ga, ala, va);
      if (anyb) blend(j1, b1, Gb, alb, vb);
      if ((pend_at & 511u) >= (PEND - 1) * 16) { // the next trip may park two more
        flush((int)(pend_at & 511u) >> 4);
        park_at = swg + lane * 4;
        pend_at &= ~511u;
      }
    }
  }
  if (pend_at & 511u) flush((int)(pend_at & 511u) >> 4);
  copies_wait_b();
  __syncwarp(); // the next unit reuses this warp's shared-memory slots
 } // unit loop
  if (lane == 0) stamp_done(a.prof, ST_RBWD);
}

template <class K>
static int resident_ctas_bwd(K kernel, int threads) {
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
  // 6 x 4 warps per SM measured best on B200 (sweep 2..8 in profiles/r01_v3_notes.md): fewer
  // resident warps leave a deeper ticket queue (better balance), more hide more latency
  per_sm = min(per_sm, 6);
  if (const char* e = getenv("IGS_BWD_CTAS")) per_sm = max(1, atoi(e)); // experiments only
  return (sms > 0 ? sms : 148) * (per_sm > 0 ? per_sm : 1);
}

cudaError_t launch_render_backward(const BwdRenderArgs& a, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  static const int cap0 = resident_ctas_bwd(render_backward_kernel<0>, BT),
                   cap1 = resident_ctas_bwd(render_backward_kernel<1>, BT),
                   capx = resident_ctas_bwd(render_backward_kernel<-1>, BT);
  const int units = a.grid_x * a.grid_y * 8;
  const int want = (units + BW - 1) / BW;
  count_launch();
  if (a.ED == 0) return launch_after(render_backward_kernel<0>, dim3(min(want, cap0)), dim3(BT), 0, s, a);
  if (a.ED == 1) return launch_after(render_backward_kernel<1>, dim3(min(want, cap1)), dim3(BT), 0, s, a);
  return launch_after(render_backward_kernel<-1>, dim3(min(want, capx)), dim3(BT), 0, s, a);
}

} // namespace igs
