// Stage 1 of the forward path: per-Gaussian preprocess (EWA projection, 3D->2D
// covariance, SH->RGB, radius, tile rectangle; reference forward.cu:203-316) fused
// with the count of instances per tile that sizes the tile segments (the role of
// cub::DeviceScan::InclusiveSum, rasterizer_impl.cu:280, in the reference: here
// instances are placed by tile, not by Gaussian).  One thread per Gaussian, no
// dependence between threads except the CTA-local tile histogram.  Also markVisible
// and the on-request prefix sum of tiles_touched (the reference's point_offsets).
//
// Bound: HBM.  Algorithmic bytes per Gaussian: reads 44 (+12*(D+1)^2 SH),
// writes 64 (splat record) + 4 radii + 4 tiles (+1 clamp mask); 4 per instance of tile counting.
#include <cstdlib>
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
__device__ __forceinline__ void sh_load_row(int deg, int M, const float* __restrict__ sh, float (&v)[48]) {
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
}

__device__ __forceinline__ float3 sh_eval(int deg, const float (&v)[48], float x, float y, float z, uint32_t& clamp_mask) {
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

#ifndef IGS_PK_THREADS
#define IGS_PK_THREADS 256
#endif
// Gaussians per preprocess CTA.  256 (4 CTAs per SM): with the prefix sum gone the only CTA barrier left
// is the one in front of the histogram flush, and smaller CTAs wait less at it (C5: 512 -> 121.0 us,
// 256 -> 117.5, 128 -> 117.1; C2: 8.5 / 7.8 / 9.4)
constexpr int PK_THREADS = IGS_PK_THREADS;

template <bool USE_SH, bool HIST_SMEM>
__global__ void __launch_bounds__(PK_THREADS, 1024 / PK_THREADS) preprocess_kernel(PreArgs a) {
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  // CTA-local histogram of the instances per tile (HIST_SMEM: the launch gave us 4*T bytes):
  // one global RED per (CTA, non-empty tile) instead of one per instance
  extern __shared__ uint32_t s_hist[];
  const int tid = threadIdx.x;
  const int T = a.grid_x * a.grid_y;
  if (a.prof.early) pdl_launch_dependents(); // the duplicate kernel's CTAs may take their places (they wait for this grid to finish)
  if (HIST_SMEM)
    for (int t = tid; t < T; t += PK_THREADS) s_hist[t] = 0u;
  pdl_wait(); // launched as a programmatic dependent of whatever precedes it on the stream
  if (tid == 0) stamp_begun(a.prof, ST_FWD_START);
  if (tid < 16) { s_view[tid] = __ldg(a.viewmatrix + tid); s_proj[tid] = __ldg(a.projmatrix + tid); }
  if (tid < 3) s_cam[tid] = __ldg(a.campos + tid);
  __syncthreads();
  const int bid = (int)blockIdx.x;

  const int idx = bid * PK_THREADS + tid;
  uint32_t tiles = 0;
  int radius = 0;
  // what the record needs, kept across the (warp-cooperative) SH section
  float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0;
  float3 rgb = make_float3(0.f, 0.f, 0.f), nrm = rgb, dir = rgb;
  if (idx < a.P) {
    // every input of this Gaussian is requested up front (one DRAM round trip instead of three
    // dependent ones); the few that turn out to be culled cost 44 B of bandwidth
    const float px = __ldg(a.means3D + 3 * idx), py = __ldg(a.means3D + 3 * idx + 1), pz = __ldg(a.means3D + 3 * idx + 2);
    const float3 sc = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
    const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
    const float op = __ldg(a.opacities + idx);
    const float e0 = a.extra_attrs ? __ldg(a.extra_attrs + (size_t)idx * a.ED) : 0.f;
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
          if (a.norm3Ds_precomp) {
            nrm = make_float3(__ldg(a.norm3Ds_precomp + 3 * idx), __ldg(a.norm3Ds_precomp + 3 * idx + 1),
                              __ldg(a.norm3Ds_precomp + 3 * idx + 2));
          } else {
            int k = 0;
            if (sc.x > sc.z && sc.y > sc.z) k = 2;
            else if (sc.x > sc.y && sc.z > sc.y) k = 1;
            nrm = make_float3(R.m[0][k], R.m[1][k], R.m[2][k]);
            const float rx = __fsub_rn(px, s_cam[0]), ry = __fsub_rn(py, s_cam[1]), rz = __fsub_rn(pz, s_cam[2]);
            if (dot3c(rx, nrm.x, ry, nrm.y, rz, nrm.z) > 0.f) nrm = make_float3(-nrm.x, -nrm.y, -nrm.z);
          }
          if (USE_SH) {
            float dx = __fsub_rn(px, s_cam[0]), dy = __fsub_rn(py, s_cam[1]), dz = __fsub_rn(pz, s_cam[2]);
            const float len = __fsqrt_rn(dot3c(dx, dx, dy, dy, dz, dz));
            dir = make_float3(__fdiv_rn(dx, len), __fdiv_rn(dy, len), __fdiv_rn(dz, len));
          }
          r0 = make_float4(__fmul_rn(cov.z, det_inv), __fmul_rn(-cov.y, det_inv), __fmul_rn(cov.x, det_inv), op);
          r1 = make_float4(sx, sy, vz, e0);
        }
      }
    }
    a.radii[idx] = radius;
    a.tiles_touched[idx] = tiles;
  }
  // ---- colour from spherical harmonics (outside the divergent nest above: every visible lane of the warp
  // issues its row loads together).  Measured and dropped: the warp fetching its 32 rows as contiguous runs
  // through shared memory -- C5 113.5 us against 110.7 with the per-lane 16-byte loads, C3 face 13.4 / 12.1.
  if (USE_SH && tiles != 0) {
    float v[48];
    sh_load_row(a.D, a.M, a.shs + (size_t)idx * a.M * 3, v);
    uint32_t mask = 0;
    rgb = sh_eval(a.D, v, dir.x, dir.y, dir.z, mask);
    a.clamped[idx] = (uint8_t)mask;
  }
  if (tiles != 0) {
    float4* rec = a.splat + (size_t)idx * SPLAT_F4;
    rec[0] = r0;
    rec[1] = r1;
    rec[2] = make_float4(rgb.x, rgb.y, rgb.z, nrm.x);
    rec[3] = make_float4(nrm.y, nrm.z, 2.f * logf(255.f * r0.w), 0.f); // culling bound, see splat_may_touch
  }

  // No prefix sum here: instances are placed by tile (per-tile counts -> ranges -> cursors,
  // igs_binning.cu), so the reference's point_offsets (InclusiveSum, rasterizer_impl.cu:280) is not
  // on the path; igs_point_offsets produces it on request.  num_rendered is the sum of the tile counts.
  if (HIST_SMEM) {
    __syncthreads();
    for (int t = tid; t < T; t += PK_THREADS) {
      const uint32_t c = s_hist[t];
      if (c) atomicAdd(a.tile_count + (size_t)t * CNT_STRIDE, c);
    }
  }
  if (tid == 0) stamp_done(a.prof, ST_PRE);
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

cudaError_t launch_preprocess(const PreArgs& a, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const int blocks = (a.P + PK_THREADS - 1) / PK_THREADS;
  const int T = a.grid_x * a.grid_y;
  const size_t hist = (size_t)T * sizeof(uint32_t);
  // the histogram may need more than the 48 KB a kernel gets without asking
  static const cudaError_t optin = [] {
    const int bytes = HIST_SMEM_MAX_TILES * (int)sizeof(uint32_t);
    cudaError_t e = cudaFuncSetAttribute(preprocess_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(preprocess_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    return e;
  }();
  if (optin != cudaSuccess) return optin;
  count_launch();
  const PreArgs& b = a;
  const size_t smem = T <= HIST_SMEM_MAX_TILES ? hist : 0;
  if (T <= HIST_SMEM_MAX_TILES) {
    if (a.shs) return launch_after(preprocess_kernel<true, true>, dim3(blocks), dim3(PK_THREADS), smem, s, b);
    return launch_after(preprocess_kernel<false, true>, dim3(blocks), dim3(PK_THREADS), smem, s, b);
  }
  // images beyond ~3500x3500: count straight into global memory
  if (a.shs) return launch_after(preprocess_kernel<true, false>, dim3(blocks), dim3(PK_THREADS), smem, s, b);
  return launch_after(preprocess_kernel<false, false>, dim3(blocks), dim3(PK_THREADS), smem, s, b);
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s) {
  if (P <= 0) return;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view, present);
  count_launch();
}

} // namespace igs
