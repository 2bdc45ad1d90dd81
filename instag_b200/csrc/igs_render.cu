// Stage 4 of the forward path: front-to-back alpha compositing
// (reference renderCUDA, forward.cu:321-456).
//
// The reference runs one 256-thread CTA per 16x16 tile, stages 256 splats at a
// time behind two __syncthreads and lets every pixel walk the whole tile list.
// Here the unit of work is ONE WARP = one 8x4 pixel block of a tile, fully
// autonomous (no CTA barrier anywhere):
//  * the warp walks its tile's depth-sorted list 32 entries at a time; lane l
//    fetches entry l's {conic, opacity | x, y, depth, extra0} (32 B) and tests it
//    against the warp's pixel rectangle (splat_may_touch, igs_common.cuh) --
//    on the BASELINE frame 2/3 of the entries are dropped here;
//  * survivors fetch the other 32 B of their record and park all 64 B in the
//    warp's private shared-memory slots; the 32 pixels then composite only the
//    survivors, reading them with broadcast LDS.128, two splats per trip so the
//    two exp/alpha chains overlap;
//  * ids are prefetched two chunks ahead and the cull data one chunk ahead, so
//    the list walk never waits for an id -> record pointer chase;
//  * a warp stops as soon as its 32 pixels are saturated (the reference only
//    votes per CTA, forward.cu:380); the kernel is persistent (4-warp CTAs, as
//    many as fit) and every warp pulls units from one ticket, longest lists first.
// None of this changes a pixel: the per-pixel arithmetic below is the
// reference's, with the same FMA contraction (igs_common.cuh gauss_power).
//
// Bound: FP32 issue.  Per evaluated pair ~15 flop + 1 MUFU.EX2, per blended
// pair +17 flop.  DRAM floor: 64 B per instance + 40 B per pixel.
#include <cstdlib>
#include <type_traits>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

static __device__ unsigned long long* g_trace = nullptr;
__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ uint32_t smid() {
  uint32_t r;
  asm volatile("mov.u32 %0, %smid;" : "=r"(r));
  return r;
}

constexpr int RW = 4;           // warps per CTA: one per SM sub-partition
constexpr int RT = RW * 32;     // threads per CTA
#ifndef IGS_FWD_RG
#define IGS_FWD_RG 4
#endif
constexpr int RG = IGS_FWD_RG;  // survivors composited per trip of the inner loop
#ifndef IGS_FWD_MIN_CTAS
#define IGS_FWD_MIN_CTAS 5       // resident CTAs per SM the register budget is cut for (launch_render caps at 5)
#endif

__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, float4 v) {
  asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
// 16 bytes global -> shared without a register in between (LDGSTS): the copy is in flight while the warp works
__device__ __forceinline__ void copy16_async(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void copies_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void copies_wait() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }


// Persistent kernel: every SM holds as many 4-warp CTAs as fit, and every WARP pulls (tile,
// 8x4 block) units from one global ticket, heaviest tiles first (tile_order), until the
// frame is done -- the four sub-partitions of an SM stay evenly loaded to the end.
template <int ED_T> // 0, 1, or -1 = generic (runtime ED up to IGS_MAX_EXTRA_DIMS)
__global__ void __launch_bounds__(RT, IGS_FWD_MIN_CTAS) render_kernel(RenderArgs a) {
  // per warp, 3 KB: planes r0, r1 of the chunk's survivors (stored by the survivors after the cull,
  // from the registers the cull data was prefetched into), and, double buffered, planes r2, r3 of
  // EVERY entry of the chunk, copied in asynchronously one chunk ahead (cp.async: no register in
  // between).  A survivor's colour / normal are then in shared memory when its first blend needs them:
  // fetched after the cull, that load was the kernel's largest single stall (11 % of the samples).
  // (Measured and dropped: all four planes through a cp.async ring two chunks ahead -- 66.0 us against
  // 63.7; 7 / 8 resident CTAs per SM at 64 registers -- 72.0 / 77.7 us: the kernel is issue bound while
  // its warps are busy, more of them only stretch the heavy units that make the tail.)
  __shared__ float4 s_rec[RW][6][32];
  __shared__ uint32_t s_id[ED_T < 0 ? RW : 1][32];

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (a.prof.early) pdl_launch_dependents();
  pdl_wait(); // the sorted tile lists
  if (a.header->overflow) return;
  const int num_units = a.grid_x * a.grid_y * 8;
  const int ED = ED_T < 0 ? a.ED : ED_T;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(&s_rec[warp][0][0]);
  const uint32_t smine = sbase + lane * 16;         // my slot in plane 0; planes are 512 B apart
  const float bg0 = __ldg(a.bg + 0), bg1 = __ldg(a.bg + 1), bg2 = __ldg(a.bg + 2);

 // the first unit of every warp is its own number (no ticket: all resident warps start at once, and that
 // many atomics on one word queue up for ~2 us); the tickets hand out the units behind those
 auto next_unit = [&]() {
   int u = 0;
   if (lane == 0) u = (int)gridDim.x * RW + (int)atomicAdd(&a.header->render_ticket, 1u);
   return __shfl_sync(0xffffffffu, u, 0);
 };
 for (int unit = (int)blockIdx.x * RW + warp; unit < num_units; unit = next_unit()) {
  unsigned long long* const trace = g_trace;
  const unsigned long long t_start = trace ? gtimer() : 0ull;
  // consecutive tickets (= the four warps of a CTA, one SM) go to four DIFFERENT tiles, so that
  // an SM's share of the frame is a sum of independent unit costs (groups of 4 tiles x 8 blocks)
  const int num_tiles = a.grid_x * a.grid_y;
  int slot = ((unit >> 5) << 2) + (unit & 3), wb = (unit >> 2) & 7;
  if (((unit >> 5) << 2) + 4 > num_tiles) { slot = unit >> 3; wb = unit & 7; } // ragged last group
  const uint4 entry = __ldg(a.tile_order + slot);   // longest lists first; the tile's range comes with it
  const int tile = (int)entry.x;
  const int tx = tile % a.grid_x, ty = tile / a.grid_x;
  const int bx = tx * TILE + (wb & 1) * 8, by = ty * TILE + (wb >> 1) * 4;
  const int px = bx + (lane & 7), py = by + (lane >> 3);
  const bool inside = px < a.W && py < a.H;
  const float pxf = (float)px, pyf = (float)py;
  const float wx0 = (float)bx, wx1 = wx0 + 7.f, wy0 = (float)by, wy1 = wy0 + 3.f;
  const uint2 range = make_uint2(entry.y, entry.z);
  const int total = (int)(range.y - range.x);
  const uint32_t* __restrict__ list = a.point_list + range.x;

  float T = 1.0f;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, Dp = 0.f, N0 = 0.f, N1 = 0.f, N2 = 0.f, E0 = 0.f;
  float Ex[ED_T < 0 ? IGS_MAX_EXTRA_DIMS : 1];
  if (ED_T < 0) {
#pragma unroll
    for (int c = 0; c < IGS_MAX_EXTRA_DIMS; c++) Ex[c] = 0.f;
  }
  uint32_t last_contributor = 0;
  bool done = !inside;

  // software pipeline: ids two chunks ahead, cull data (r0, r1: registers) and colour / normal planes
  // (r2, r3: shared memory) one chunk ahead.  The two r2 / r3 buffers sit behind the r0 / r1 planes:
  // r2 plane at +1024 (+2048), r3 plane 512 further.
  uint32_t id_cur = 0, id_nxt = 0;
  float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
  if (lane < total) id_cur = __ldg(list + lane);
  if (32 + lane < total) id_nxt = __ldg(list + 32 + lane);
  if (lane < total) {
    const float4* rec = a.splat + (size_t)id_cur * SPLAT_F4;
    q0 = __ldg(rec);
    q1 = __ldg(rec + 1);
    copy16_async(smine + 1024, rec + 2);
    copy16_async(smine + 1536, rec + 3);
  }
  copies_commit();
  uint32_t par = 0; // which r2 / r3 buffer the current chunk reads

  uint32_t tr_chunks = 0, tr_surv = 0; // (trace only)
  for (int cb = 0; cb < total; cb += 32) {
    if (__all_sync(0xffffffffu, done)) break;
    const bool live = cb + lane < total;
    // cut = 2 ln(255 opacity), the record's r3.z, recomputed here so that culling needs 32 B only
    const bool hit = live && splat_may_touch(q0.x, q0.y, q0.z, q1.x, q1.y, 2.f * __logf(255.f * q0.w),
                                             wx0, wx1, wy0, wy1);
    unsigned m = __ballot_sync(0xffffffffu, hit);
    if (trace) { tr_chunks++; tr_surv += __popc(m); }
    const uint32_t id_mine = id_cur;
    // this chunk's r2 / r3 planes have landed; everybody is done reading the previous chunk's slots
    copies_wait();
    __syncwarp();
    // survivors park the half of their record the cull used, from the registers the prefetch below reuses
    if (hit) {
      sts128(smine, q0);
      sts128(smine + 512, q1);
      if (ED_T < 0) s_id[warp][lane] = id_mine;
    }
    // prefetch the next chunk's cull data and colour / normal planes, and the ids of the chunk after it
    id_cur = id_nxt;
    const uint32_t cur = sbase + 1024 + par * 1024; // r2 plane of this chunk (r3: +512)
    par ^= 1u;
    if (cb + 32 + lane < total) {
      const float4* rec = a.splat + (size_t)id_cur * SPLAT_F4;
      q0 = __ldg(rec);
      q1 = __ldg(rec + 1);
      copy16_async(smine + 1024 + par * 1024, rec + 2);
      copy16_async(smine + 1536 + par * 1024, rec + 3);
    }
    copies_commit();
    if (cb + 64 + lane < total) id_nxt = __ldg(list + cb + 64 + lane);
    if (m == 0) continue;
    __syncwarp();
    // G survivors per trip.  Their alpha chains are independent (instruction-level parallelism for a
    // warp that is alone on its scheduler at the end of the frame); only the transmittance recurrence is
    // ordered.  A trip first runs it WITHOUT the stop test -- a splat that does not blend gets alpha 0,
    // which leaves T alone (T * (1 - 0) = T exactly) and weighs nothing: T only ever shrinks, so if it is
    // still >= T_MIN after the trip no pixel stopped inside it, and the products are the reference's, in
    // its order.  Only the trips in which a pixel does stop (at most once per pixel and unit) repeat
    // the recurrence with the reference's test after every splat.
    auto trip = [&](auto width) {
      constexpr int G = decltype(width)::value;
      int j[G];
      bool have[G];
#pragma unroll
      for (int h = 0; h < G; h++) {
        have[h] = m != 0;
        j[h] = have[h] ? __ffs(m) - 1 : j[h ? h - 1 : 0];
        m &= m - 1;
      }
      float alpha[G], zd[G], ze[G];
      bool ok[G];
#pragma unroll
      for (int h = 0; h < G; h++) {
        const float4 r1 = lds128(sbase + 512 + j[h] * 16), r0 = lds128(sbase + j[h] * 16);
        const float dx = __fsub_rn(r1.x, pxf), dy = __fsub_rn(r1.y, pyf);
        const float pw = gauss_power(r0.x, r0.y, r0.z, dx, dy);
        alpha[h] = fminf(0.99f, __fmul_rn(r0.w, expf(pw)));
        ok[h] = have[h] && !done && !(pw > 0.0f) && !(alpha[h] < ALPHA_MIN);
        zd[h] = r1.z; ze[h] = r1.w;
      }
      float w[G];
      float Tn = T;
      bool any = false;
#pragma unroll
      for (int h = 0; h < G; h++) {
        const float am = ok[h] ? alpha[h] : 0.f;
        w[h] = __fmul_rn(am, Tn);
        Tn = __fmul_rn(Tn, __fsub_rn(1.f, am));
        any = any || ok[h];
      }
      if (!__any_sync(0xffffffffu, Tn < T_MIN)) {
        T = Tn;
#pragma unroll
        for (int h = 0; h < G; h++) last_contributor = ok[h] ? (uint32_t)(cb + j[h] + 1) : last_contributor;
      } else {
        // ordered part with the stop test: T, done, last contributor (bit-exact with the reference)
        any = false;
#pragma unroll
        for (int h = 0; h < G; h++) {
          const float test_T = __fmul_rn(T, __fsub_rn(1.f, alpha[h]));
          const bool live_h = ok[h] && !done;
          const bool stop = test_T < T_MIN;
          const bool blend = live_h && !stop;
          done = done || (live_h && stop);
          w[h] = blend ? __fmul_rn(alpha[h], T) : 0.f;
          T = blend ? test_T : T;
          last_contributor = blend ? (uint32_t)(cb + j[h] + 1) : last_contributor;
          any = any || blend;
        }
      }
      if (!__any_sync(0xffffffffu, any)) return;
      // X += f*(alpha*T): one rounding away from the reference's fma(f*alpha, T, X)
      // (forward.cu:425-431), far inside the 1e-5 image tolerance
#pragma unroll
      for (int h = 0; h < G; h++) {
        const float4 r2 = lds128(cur + j[h] * 16);
        const float4 r3 = lds128(cur + 512 + j[h] * 16);
        C0 = __fmaf_rn(r2.x, w[h], C0);
        C1 = __fmaf_rn(r2.y, w[h], C1);
        C2 = __fmaf_rn(r2.z, w[h], C2);
        Dp = __fmaf_rn(zd[h], w[h], Dp);
        N0 = __fmaf_rn(r2.w, w[h], N0);
        N1 = __fmaf_rn(r3.x, w[h], N1);
        N2 = __fmaf_rn(r3.y, w[h], N2);
        if (ED_T == 1) E0 = __fmaf_rn(ze[h], w[h], E0);
        if (ED_T < 0) {
          if (w[h] != 0.f) {
            const float* e = a.extra_attrs + (size_t)s_id[warp][j[h]] * ED;
            for (int c = 0; c < ED; c++) Ex[c] = __fmaf_rn(__ldg(e + c), w[h], Ex[c]);
          }
        }
      }
    };
    // four at a time; a remainder of one or two in a trip half as wide (trips of 6 / 8 with halves of
    // 3 / 4 measured: 62.7 / 68.0 us against 60.8)
    while (m) {
      if (__popc(m) > RG / 2) trip(std::integral_constant<int, RG>{});
      else trip(std::integral_constant<int, RG / 2>{});
    }
  }

  if (inside) {
    const size_t HW = (size_t)a.H * a.W, pix = (size_t)py * a.W + px;
    a.out_alpha[pix] = __fsub_rn(1.f, T);
    a.n_contrib[pix] = last_contributor;
    a.out_color[pix] = __fmaf_rn(T, bg0, C0);
    a.out_color[HW + pix] = __fmaf_rn(T, bg1, C1);
    a.out_color[2 * HW + pix] = __fmaf_rn(T, bg2, C2);
    a.out_depth[pix] = Dp;
    if (a.normalize) { // F.normalize(norm, p=2, dim=0): n / max(|n|, 1e-12)  (DGR diff_gauss/__init__.py:48)
      const float len = __fsqrt_rn(__fmaf_rn(N2, N2, __fmaf_rn(N1, N1, __fmul_rn(N0, N0))));
      const float den = fmaxf(len, 1e-12f);
      a.norm_len[pix] = len;
      N0 = __fdiv_rn(N0, den); N1 = __fdiv_rn(N1, den); N2 = __fdiv_rn(N2, den);
    }
    a.out_norm[pix] = N0;
    a.out_norm[HW + pix] = N1;
    a.out_norm[2 * HW + pix] = N2;
    if (ED_T == 1) a.out_extra[pix] = E0;
    if (ED_T < 0)
      for (int c = 0; c < ED; c++) a.out_extra[c * HW + pix] = Ex[c];
  }
  copies_wait(); // (a unit that stopped early leaves its last prefetch in flight)
  __syncwarp();  // the next unit reuses this warp's shared-memory slots
  if (trace && lane == 0) {
    trace[4 * unit] = t_start; trace[4 * unit + 1] = gtimer();
    trace[4 * unit + 2] = (unsigned long long)smid() | ((unsigned long long)total << 32);
    trace[4 * unit + 3] = (unsigned long long)tr_chunks | ((unsigned long long)tr_surv << 32);
  }
 } // unit loop
  if (lane == 0) stamp_done(a.prof, ST_RENDER);
}

// CTAs that fit on the device at once (all SMs), per kernel variant; cached per process
template <class K>
static int resident_ctas(K kernel, int threads) {
  int dev = 0, sms = 0, per_sm = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0);
  // 5 x 4 warps per SM measured best on B200 (4 / 5 / 6: 63.4 / 62.5 / 64.0 us, profiles/r02_notes.md):
  // the kernel is issue bound while its warps are busy; fewer resident warps leave a deeper ticket
  // queue and give the heavy units that make the tail a larger share of their scheduler
  per_sm = min(per_sm, 5);
  if (const char* e = getenv("IGS_FWD_CTAS")) per_sm = max(1, atoi(e)); // experiments only
  return (sms > 0 ? sms : 148) * (per_sm > 0 ? per_sm : 1);
}

cudaError_t set_render_trace(unsigned long long* p) { return cudaMemcpyToSymbol(g_trace, &p, sizeof p); }

cudaError_t launch_render(const RenderArgs& a, cudaStream_t s) {
  if (a.W <= 0 || a.H <= 0) return cudaSuccess;
  static const int cap0 = resident_ctas(render_kernel<0>, RT), cap1 = resident_ctas(render_kernel<1>, RT),
                   capx = resident_ctas(render_kernel<-1>, RT);
  const int units = a.grid_x * a.grid_y * 8;
  const int want = (units + RW - 1) / RW;
  count_launch();
  if (a.ED == 0) return launch_after(render_kernel<0>, dim3(min(want, cap0)), dim3(RT), 0, s, a);
  if (a.ED == 1) return launch_after(render_kernel<1>, dim3(min(want, cap1)), dim3(RT), 0, s, a);
  return launch_after(render_kernel<-1>, dim3(min(want, capx)), dim3(RT), 0, s, a);
}

} // namespace igs
