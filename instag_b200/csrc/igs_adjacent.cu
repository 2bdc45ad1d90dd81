// The steps either side of the rasterizer that SURVEY.md 8(f) ranks next (f1, f2), as
// single HBM-bound kernels behind the same C ABI:
//
//  f1  igs_compose_frame   synthesize_fuse.py:64-76: optional 13x13 max-dilation of the mouth
//                          alpha, mouth over background, face over mouth, clamp, *255, uint8 HWC --
//                          one pass over the planes instead of ~12 element-wise launches and a
//                          blocking .cpu() per frame (the frame lands in a device ring slot).
//  f2  igs_activate_*      gaussian_renderer/__init__.py:227-235: means3D = xyz + d_xyz,
//                          scales = softplus(_scaling + d_scale), rotations = normalize(_rotation
//                          + d_rot), opacity = sigmoid(_opacity); forward and backward.
//      igs_densify_stats   train_face.py:675-682 + scene/gaussian_model.py:683-685: max_radii2D,
//                          xyz_gradient_accum, denom updated in place for the visible Gaussians
//                          (no boolean-mask gathers/scatters).
//
// All of them stream every byte once; grids are sized to the data (they are far below one
// wave per SM only for toy inputs).
#include <cfloat>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

// ---------------------------------------------------------------------------------------
// f1: frame composition
// ---------------------------------------------------------------------------------------
constexpr int CX = 32, CY = 8; // pixels per CTA
constexpr int MAX_HALO = 8;    // dilation kernels up to 17x17

// alpha planes are [1,H,W]; rgb planes [3,H,W]; background [3,H,W] in 0..255 or NULL (= 0);
// out_u8 [H,W,3]; out_image [3,H,W] float (optional, un-clamped `image` of the reference).
__global__ void __launch_bounds__(CX * CY) compose_kernel(
    const float* __restrict__ face_rgb, const float* __restrict__ face_alpha,
    const float* __restrict__ mouth_rgb, const float* __restrict__ mouth_alpha,
    const float* __restrict__ background, int H, int W, int halo, uint8_t* __restrict__ out_u8,
    float* __restrict__ out_image) {
  __shared__ float s_a[CY + 2 * MAX_HALO][CX + 2 * MAX_HALO + 1];
  __shared__ float s_r[CY + 2 * MAX_HALO][CX];
  const int tx = threadIdx.x % CX, ty = threadIdx.x / CX;
  const int x0 = blockIdx.x * CX, y0 = blockIdx.y * CY;
  const int x = x0 + tx, y = y0 + ty;
  float am = 0.f;
  if (halo > 0) {
    // F.max_pool2d(alpha, k, stride 1, padding halo): out-of-image samples never win
    const int rw = CX + 2 * halo, rh = CY + 2 * halo;
    for (int i = threadIdx.x; i < rw * rh; i += CX * CY) {
      const int ry = i / rw, rx = i % rw;
      const int gx = x0 - halo + rx, gy = y0 - halo + ry;
      s_a[ry][rx] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? __ldg(mouth_alpha + (size_t)gy * W + gx) : -FLT_MAX;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < CX * rh; i += CX * CY) { // row maxima
      const int ry = i / CX, rx = i % CX;
      float m = -FLT_MAX;
      for (int k = 0; k <= 2 * halo; k++) m = fmaxf(m, s_a[ry][rx + k]);
      s_r[ry][rx] = m;
    }
    __syncthreads();
    float m = -FLT_MAX;
    for (int k = 0; k <= 2 * halo; k++) m = fmaxf(m, s_r[ty + k][tx]);
    am = m;
  }
  if (x >= W || y >= H) return;
  const size_t HW = (size_t)H * W, pix = (size_t)y * W + x;
  if (halo <= 0) am = __ldg(mouth_alpha + pix);
  const float af = __ldg(face_alpha + pix);
  const float one_am = __fsub_rn(1.f, am), one_af = __fsub_rn(1.f, af);
  uint8_t px[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    // every operation rounded separately, in the reference's order (torch has no FMA here):
    // mouth_image = mouth + bg/255 * (1 - alpha_mouth); image = face + mouth_image * (1 - alpha)
    const float bg = background ? __fdiv_rn(__ldg(background + c * HW + pix), 255.0f) : 0.f;
    const float mouth = __fadd_rn(__ldg(mouth_rgb + c * HW + pix), __fmul_rn(bg, one_am));
    const float img = __fadd_rn(__ldg(face_rgb + c * HW + pix), __fmul_rn(mouth, one_af));
    if (out_image) out_image[c * HW + pix] = img;
    const float v = __fmul_rn(fminf(fmaxf(img, 0.f), 1.f), 255.f); // (a NaN pixel becomes 0)
    px[c] = (uint8_t)(int)v;                                          // .astype(np.uint8): truncation
  }
  uint8_t* o = out_u8 + pix * 3;
  o[0] = px[0]; o[1] = px[1]; o[2] = px[2];
}

// ---------------------------------------------------------------------------------------
// f2: activations around the rasterizer call
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ float softplus_f(float x) { // torch softplus, beta 1, threshold 20
  return x > 20.f ? x : log1pf(expf(x));
}
__device__ __forceinline__ float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

struct ActArgs {
  int P;
  const float *xyz, *d_xyz, *scaling, *d_scale, *rotation, *d_rot, *opacity; // d_* may be NULL (= 0)
  float *means3D, *scales, *rotations, *opacities;
};

__global__ void __launch_bounds__(256) activate_forward_kernel(ActArgs a) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= a.P) return;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const float dx = a.d_xyz ? __ldg(a.d_xyz + 3 * i + c) : 0.f;
    a.means3D[3 * i + c] = __fadd_rn(__ldg(a.xyz + 3 * i + c), dx);
    const float ds = a.d_scale ? __ldg(a.d_scale + 3 * i + c) : 0.f;
    a.scales[3 * i + c] = softplus_f(__fadd_rn(__ldg(a.scaling + 3 * i + c), ds));
  }
  float4 q = __ldg(reinterpret_cast<const float4*>(a.rotation) + i);
  if (a.d_rot) {
    const float4 d = __ldg(reinterpret_cast<const float4*>(a.d_rot) + i);
    q.x = __fadd_rn(q.x, d.x); q.y = __fadd_rn(q.y, d.y); q.z = __fadd_rn(q.z, d.z); q.w = __fadd_rn(q.w, d.w);
  }
  // F.normalize(q, dim=1): q / max(|q|, 1e-12)
  const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
  reinterpret_cast<float4*>(a.rotations)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
  a.opacities[i] = sigmoid_f(__ldg(a.opacity + i));
}

struct ActBwdArgs {
  int P;
  const float *scaling, *d_scale, *rotation, *d_rot, *opacity;          // forward inputs
  const float *g_means3D, *g_scales, *g_rotations, *g_opacities;         // upstream (any may be NULL = 0)
  float *g_xyz, *g_scaling, *g_rotation, *g_opacity;                     // outputs; the d_* get the same gradients
};

__global__ void __launch_bounds__(256) activate_backward_kernel(ActBwdArgs a) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= a.P) return;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    a.g_xyz[3 * i + c] = a.g_means3D ? __ldg(a.g_means3D + 3 * i + c) : 0.f;
    const float ds = a.d_scale ? __ldg(a.d_scale + 3 * i + c) : 0.f;
    const float x = __fadd_rn(__ldg(a.scaling + 3 * i + c), ds);
    const float g = a.g_scales ? __ldg(a.g_scales + 3 * i + c) : 0.f;
    a.g_scaling[3 * i + c] = x > 20.f ? g : g * sigmoid_f(x); // d softplus = sigmoid
  }
  float4 q = __ldg(reinterpret_cast<const float4*>(a.rotation) + i);
  if (a.d_rot) {
    const float4 d = __ldg(reinterpret_cast<const float4*>(a.d_rot) + i);
    q.x = __fadd_rn(q.x, d.x); q.y = __fadd_rn(q.y, d.y); q.z = __fadd_rn(q.z, d.z); q.w = __fadd_rn(q.w, d.w);
  }
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
  if (a.g_rotations) g = __ldg(reinterpret_cast<const float4*>(a.g_rotations) + i);
  const float len = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  float4 o;
  if (len > 1e-12f) { // (g - q_hat <q_hat, g>) / |q|
    const float inv = 1.f / len;
    const float hx = q.x * inv, hy = q.y * inv, hz = q.z * inv, hw = q.w * inv;
    const float d = hx * g.x + hy * g.y + hz * g.z + hw * g.w;
    o = make_float4((g.x - hx * d) * inv, (g.y - hy * d) * inv, (g.z - hz * d) * inv, (g.w - hw * d) * inv);
  } else { // the clamp is active: q / eps
    o = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
  }
  reinterpret_cast<float4*>(a.g_rotation)[i] = o;
  const float s = sigmoid_f(__ldg(a.opacity + i));
  a.g_opacity[i] = (a.g_opacities ? __ldg(a.g_opacities + i) : 0.f) * s * (1.f - s);
}

__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const int32_t* __restrict__ radii,
                                                            const float* __restrict__ grad_means2D,
                                                            float* __restrict__ max_radii2D,
                                                            float* __restrict__ xyz_gradient_accum,
                                                            float* __restrict__ denom) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const int r = __ldg(radii + i);
  if (r <= 0) return; // visibility_filter = radii > 0 (gaussian_renderer/__init__.py:131)
  max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
  const float gx = __ldg(grad_means2D + 3 * i), gy = __ldg(grad_means2D + 3 * i + 1);
  xyz_gradient_accum[i] += sqrtf(gx * gx + gy * gy); // torch.norm(grad[:, :2], dim=-1)
  denom[i] += 1.f;
}

} // namespace igs

using namespace igs;

extern "C" {

int igs_compose_frame(const float* face_rgb, const float* face_alpha, const float* mouth_rgb,
                      const float* mouth_alpha, const float* background, int32_t H, int32_t W,
                      int32_t dilate_ksize, uint8_t* out_u8, float* out_image, igs_stream_t stream) {
  if (H <= 0 || W <= 0 || !face_rgb || !face_alpha || !mouth_rgb || !mouth_alpha || !out_u8) {
    set_error("compose: bad arguments"); return IGS_ERR_INVALID;
  }
  if (dilate_ksize < 0 || (dilate_ksize > 0 && (dilate_ksize % 2 == 0 || dilate_ksize / 2 > MAX_HALO))) {
    set_error("compose: dilate_ksize must be 0 or odd <= %d", 2 * MAX_HALO + 1); return IGS_ERR_INVALID;
  }
  const dim3 grid((W + CX - 1) / CX, (H + CY - 1) / CY);
  compose_kernel<<<grid, CX * CY, 0, (cudaStream_t)stream>>>(face_rgb, face_alpha, mouth_rgb, mouth_alpha, background,
                                                             H, W, dilate_ksize / 2, out_u8, out_image);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "compose");
}

int igs_activate_forward(int32_t P, const float* xyz, const float* d_xyz, const float* scaling, const float* d_scale,
                         const float* rotation, const float* d_rot, const float* opacity, float* means3D,
                         float* scales, float* rotations, float* opacities, igs_stream_t stream) {
  if (P < 0 || (P > 0 && (!xyz || !scaling || !rotation || !opacity || !means3D || !scales || !rotations || !opacities))) {
    set_error("activate: bad arguments"); return IGS_ERR_INVALID;
  }
  if (P == 0) return IGS_OK;
  ActArgs a{P, xyz, d_xyz, scaling, d_scale, rotation, d_rot, opacity, means3D, scales, rotations, opacities};
  activate_forward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "activate forward");
}

int igs_activate_backward(int32_t P, const float* scaling, const float* d_scale, const float* rotation,
                          const float* d_rot, const float* opacity, const float* g_means3D, const float* g_scales,
                          const float* g_rotations, const float* g_opacities, float* g_xyz, float* g_scaling,
                          float* g_rotation, float* g_opacity, igs_stream_t stream) {
  if (P < 0 || (P > 0 && (!scaling || !rotation || !opacity || !g_xyz || !g_scaling || !g_rotation || !g_opacity))) {
    set_error("activate backward: bad arguments"); return IGS_ERR_INVALID;
  }
  if (P == 0) return IGS_OK;
  ActBwdArgs a{P, scaling, d_scale, rotation, d_rot, opacity, g_means3D, g_scales, g_rotations, g_opacities,
               g_xyz, g_scaling, g_rotation, g_opacity};
  activate_backward_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(a);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "activate backward");
}

int igs_densify_stats(int32_t P, const int32_t* radii, const float* grad_means2D, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, igs_stream_t stream) {
  if (P < 0 || (P > 0 && (!radii || !grad_means2D || !max_radii2D || !xyz_gradient_accum || !denom))) {
    set_error("densify stats: bad arguments"); return IGS_ERR_INVALID;
  }
  if (P == 0) return IGS_OK;
  densify_stats_kernel<<<(P + 255) / 256, 256, 0, (cudaStream_t)stream>>>(P, radii, grad_means2D, max_radii2D,
                                                                         xyz_gradient_accum, denom);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "densify stats");
}

} // extern "C"
