// the photometric loss on the rasterizer's colour output,
//     loss = l1_loss(image, gt) + lambda_dssim * (1 - ssim(image, gt))
// (reference train_face.py:421-422, utils/loss_utils.py:28-73), as two kernels that produce
// dL/dimage -- the rasterizer backward's dL_dcolor input -- directly in CHW layout:
//
//   forward : per channel, separable 11-tap Gaussian window (sigma 1.5, zero padding 5) of
//             {x, y, x^2, y^2, xy} through shared memory, 4 outputs per thread along the window -> SSIM map -> per-CTA partial sums of
//             the SSIM map and of |x - y|, plus three per-pixel derivative maps
//                 A = df/dmu1 - 2 mu1 df/ds1 - mu2 df/ds12,  B = df/ds1,  C = df/ds12
//   backward: dL/dx = g * ( sign(x - y)/N - lambda/N * (conv(A) + 2 x conv(B) + y conv(C)) )
//             (the window is symmetric, so the adjoint of the convolution is the convolution).
//
// The reference runs 5 depthwise conv2d + ~15 element-wise kernels forward and their autograd
// backward (another 5 transposed convolutions) and materialises ~12 full-frame temporaries.
// HBM bound: forward reads 2 and writes 3 floats per pixel-channel, backward reads 5, writes 1.
#include "igs_common.cuh"
#include "igs_kernels.cuh"
#include <mutex>

namespace igs {

constexpr int QT = 32;            // output tile edge
constexpr int LH = 5;             // window radius (window_size 11)
constexpr int QR = QT + 2 * LH;   // 42 rows / columns with halo
constexpr int QTHREADS = 256;     // 8 warps: a warp owns 4 output rows, a lane one column
constexpr float SSIM_C1 = 0.01f * 0.01f, SSIM_C2 = 0.03f * 0.03f;

struct Win { float w[11]; };

// gaussian(11, 1.5) of utils/loss_utils.py:42-44, in fp32 like torch.Tensor([...]) / sum
static Win make_window() {
  Win k;
  float s = 0.f;
  for (int i = 0; i < 11; i++) { k.w[i] = (float)exp(-(double)((i - 5) * (i - 5)) / (2.0 * 1.5 * 1.5)); s += k.w[i]; }
  for (int i = 0; i < 11; i++) k.w[i] = k.w[i] / s;
  return k;
}

// Four neighbouring outputs of the 11-tap window from the 14 inputs they share, all in registers:
// 14 shared-memory loads per 4 outputs instead of 44.  Taps in ascending order from 0, as a thread
// that owned one output would add them.
__device__ __forceinline__ void taps4(const Win& win, const float (&in)[14], float (&out)[4]) {
#pragma unroll
  for (int j = 0; j < 4; j++) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) s = fmaf(win.w[k], in[j + k], s);
    out[j] = s;
  }
}

__global__ void __launch_bounds__(QTHREADS) l1_ssim_forward_kernel(
    const float* __restrict__ img1, const float* __restrict__ img2, int H, int W, Win win,
    float* __restrict__ dA, float* __restrict__ dB, float* __restrict__ dC, float* __restrict__ partial) {
  __shared__ float s_x[QR][QR + 1], s_y[QR][QR + 1];
  __shared__ float s_h[5][QR][QT + 1];
  __shared__ float s_red[2][QTHREADS / 32];
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * QT, y0 = blockIdx.y * QT;
  const size_t plane = (size_t)c * H * W;
  for (int i = threadIdx.x; i < QR * QR; i += QTHREADS) {
    const int ry = i / QR, rx = i - ry * QR, gx = x0 - LH + rx, gy = y0 - LH + ry;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
    s_x[ry][rx] = in ? __ldg(img1 + plane + (size_t)gy * W + gx) : 0.f;
    s_y[ry][rx] = in ? __ldg(img2 + plane + (size_t)gy * W + gx) : 0.f;
  }
  __syncthreads();
  // horizontal pass: 42 rows x 8 groups of 4 columns; a group reads its 14 x and 14 y once
  for (int it = threadIdx.x; it < QR * (QT / 4); it += QTHREADS) {
    const int ry = it >> 3, cg = (it & 7) * 4;
    float a[14], b[14], t[14], o[4];
#pragma unroll
    for (int k = 0; k < 14; k++) { a[k] = s_x[ry][cg + k]; b[k] = s_y[ry][cg + k]; }
    taps4(win, a, o);
#pragma unroll
    for (int j = 0; j < 4; j++) s_h[0][ry][cg + j] = o[j];
    taps4(win, b, o);
#pragma unroll
    for (int j = 0; j < 4; j++) s_h[1][ry][cg + j] = o[j];
#pragma unroll
    for (int k = 0; k < 14; k++) t[k] = a[k] * a[k];
    taps4(win, t, o);
#pragma unroll
    for (int j = 0; j < 4; j++) s_h[2][ry][cg + j] = o[j];
#pragma unroll
    for (int k = 0; k < 14; k++) t[k] = b[k] * b[k];
    taps4(win, t, o);
#pragma unroll
    for (int j = 0; j < 4; j++) s_h[3][ry][cg + j] = o[j];
#pragma unroll
    for (int k = 0; k < 14; k++) t[k] = a[k] * b[k];
    taps4(win, t, o);
#pragma unroll
    for (int j = 0; j < 4; j++) s_h[4][ry][cg + j] = o[j];
  }
  __syncthreads();
  // vertical pass: lane = column, 4 rows per thread from the 14 rows they share
  const int tx = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * 4;
  float v[5][4];
#pragma unroll
  for (int m = 0; m < 5; m++) {
    float in[14];
#pragma unroll
    for (int k = 0; k < 14; k++) in[k] = s_h[m][r0 + k][tx];
    taps4(win, in, v[m]);
  }
  const int x = x0 + tx;
  float ssim_v = 0.f, l1_v = 0.f;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int y = y0 + r0 + j;
    if (x < W && y < H) {
      // utils/loss_utils.py:58-70
      const float mu1 = v[0][j], mu2 = v[1][j], e11 = v[2][j], e22 = v[3][j], e12 = v[4][j];
      const float mu1_sq = mu1 * mu1, mu2_sq = mu2 * mu2, mu12 = mu1 * mu2;
      const float s1 = e11 - mu1_sq, s2 = e22 - mu2_sq, s12 = e12 - mu12;
      const float a1 = 2.f * mu12 + SSIM_C1, a2 = 2.f * s12 + SSIM_C2;
      const float b1 = mu1_sq + mu2_sq + SSIM_C1, b2 = s1 + s2 + SSIM_C2;
      const float inv = 1.f / (b1 * b2);
      const float f = a1 * a2 * inv;
      ssim_v += f;
      const float df_dmu1 = 2.f * mu2 * a2 * inv - f * 2.f * mu1 / b1;
      const float df_ds1 = -f / b2;
      const float df_ds12 = 2.f * a1 * inv;
      const size_t pix = plane + (size_t)y * W + x;
      dA[pix] = df_dmu1 - 2.f * mu1 * df_ds1 - mu2 * df_ds12;
      dB[pix] = df_ds1;
      dC[pix] = df_ds12;
      l1_v += fabsf(s_x[r0 + j + LH][tx + LH] - s_y[r0 + j + LH][tx + LH]);
    }
  }
  // per-CTA partial sums (summed by the finish kernel in a fixed order: deterministic, no float atomics)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ssim_v += __shfl_xor_sync(0xffffffffu, ssim_v, o);
    l1_v += __shfl_xor_sync(0xffffffffu, l1_v, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_red[0][warp] = ssim_v; s_red[1][warp] = l1_v; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float a = 0.f, b = 0.f;
    for (int w = 0; w < QTHREADS / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
    const size_t cta = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
    partial[2 * cta] = a;
    partial[2 * cta + 1] = b;
  }
}

__global__ void __launch_bounds__(QTHREADS) l1_ssim_backward_kernel(
    const float* __restrict__ img1, const float* __restrict__ img2, const float* __restrict__ dA,
    const float* __restrict__ dB, const float* __restrict__ dC, int H, int W, Win win, float lambda_dssim,
    const float* __restrict__ grad_out, float inv_n, float* __restrict__ grad_img1) {
  __shared__ float s_m[3][QR][QR + 1];
  __shared__ float s_h[3][QR][QT + 1];
  const int c = blockIdx.z;
  const int x0 = blockIdx.x * QT, y0 = blockIdx.y * QT;
  const size_t plane = (size_t)c * H * W;
  for (int i = threadIdx.x; i < QR * QR; i += QTHREADS) {
    const int ry = i / QR, rx = i - ry * QR, gx = x0 - LH + rx, gy = y0 - LH + ry;
    const bool in = gx >= 0 && gx < W && gy >= 0 && gy < H;
    const size_t p = plane + (size_t)gy * W + gx;
    s_m[0][ry][rx] = in ? __ldg(dA + p) : 0.f;
    s_m[1][ry][rx] = in ? __ldg(dB + p) : 0.f;
    s_m[2][ry][rx] = in ? __ldg(dC + p) : 0.f;
  }
  __syncthreads();
  for (int it = threadIdx.x; it < QR * (QT / 4); it += QTHREADS) {
    const int ry = it >> 3, cg = (it & 7) * 4;
#pragma unroll
    for (int m = 0; m < 3; m++) {
      float in[14], o[4];
#pragma unroll
      for (int k = 0; k < 14; k++) in[k] = s_m[m][ry][cg + k];
      taps4(win, in, o);
#pragma unroll
      for (int j = 0; j < 4; j++) s_h[m][ry][cg + j] = o[j];
    }
  }
  __syncthreads();
  const int tx = threadIdx.x & 31, r0 = (threadIdx.x >> 5) * 4;
  float v[3][4];
#pragma unroll
  for (int m = 0; m < 3; m++) {
    float in[14];
#pragma unroll
    for (int k = 0; k < 14; k++) in[k] = s_h[m][r0 + k][tx];
    taps4(win, in, v[m]);
  }
  const int x = x0 + tx;
  if (x >= W) return;
  const float g = __ldg(grad_out) * inv_n;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int y = y0 + r0 + j;
    if (y >= H) break;
    const size_t pix = plane + (size_t)y * W + x;
    const float xv = __ldg(img1 + pix), yv = __ldg(img2 + pix);
    const float d = xv - yv;
    const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); // torch.abs backward: sign(0) = 0
    const float dssim = v[0][j] + 2.f * xv * v[1][j] + yv * v[2][j]; // sum_q d ssim_q / d x_p
    grad_img1[pix] = g * (sgn - lambda_dssim * dssim);
  }
}

// {loss, Ll1, ssim} from the forward kernel's per-CTA partial sums, in a fixed order (deterministic)
__global__ void __launch_bounds__(256) l1_ssim_finish_kernel(const float* __restrict__ partial, int n, float inv_n,
                                                             float lambda_dssim, float* __restrict__ out3) {
  __shared__ double s_red[2][8];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 256) { a += partial[2 * i]; b += partial[2 * i + 1]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
  if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = a; s_red[1][threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = 0.0; b = 0.0;
    for (int w = 0; w < 8; w++) { a += s_red[0][w]; b += s_red[1][w]; }
    const float ssim = (float)(a * inv_n), l1 = (float)(b * inv_n);
    out3[0] = l1 + lambda_dssim * (1.0f - ssim);
    out3[1] = l1;
    out3[2] = ssim;
  }
}

} // namespace igs

using namespace igs;

extern "C" {

int igs_l1_ssim_finish(const float* partial, int32_t C, int32_t H, int32_t W, float lambda_dssim, float* out3,
                       igs_stream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0 || !partial || !out3) { set_error("l1_ssim finish: bad arguments"); return IGS_ERR_INVALID; }
  const int n = (int)igs_l1_ssim_partials(C, H, W);
  l1_ssim_finish_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(partial, n, 1.0f / ((float)C * (float)H * (float)W), lambda_dssim, out3);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "l1_ssim finish");
}

size_t igs_l1_ssim_partials(int32_t C, int32_t H, int32_t W) {
  return (size_t)(C > 0 ? C : 0) * ((H + QT - 1) / QT) * ((W + QT - 1) / QT);
}

int igs_l1_ssim_forward(const float* img1, const float* img2, int32_t C, int32_t H, int32_t W, float* dA,
                        float* dB, float* dC, float* partial, igs_stream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !dA || !dB || !dC || !partial) {
    set_error("l1_ssim forward: bad arguments"); return IGS_ERR_INVALID;
  }
  static const Win win = make_window();
  const dim3 grid((W + QT - 1) / QT, (H + QT - 1) / QT, C);
  l1_ssim_forward_kernel<<<grid, QTHREADS, 0, (cudaStream_t)stream>>>(img1, img2, H, W, win, dA, dB, dC, partial);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "l1_ssim forward");
}

int igs_l1_ssim_backward(const float* img1, const float* img2, const float* dA, const float* dB, const float* dC,
                         int32_t C, int32_t H, int32_t W, float lambda_dssim, const float* grad_out,
                         float* grad_img1, igs_stream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !dA || !dB || !dC || !grad_out || !grad_img1) {
    set_error("l1_ssim backward: bad arguments"); return IGS_ERR_INVALID;
  }
  static const Win win = make_window();
  const dim3 grid((W + QT - 1) / QT, (H + QT - 1) / QT, C);
  const float inv_n = 1.0f / ((float)C * (float)H * (float)W);
  l1_ssim_backward_kernel<<<grid, QTHREADS, 0, (cudaStream_t)stream>>>(img1, img2, dA, dB, dC, H, W, win, lambda_dssim,
                                                                      grad_out, inv_n, grad_img1);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "l1_ssim backward");
}

} // extern "C"

// ---- igs_l1_ssim_forward_static: the forward + finish pair with NO caller scratch --------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the per-CTA partial sums live in a file-scope device array instead of a caller
// workspace.  Every GPU that loads the library holds the array (512 KB) until exit.  No counterpart in
// the reference.
constexpr size_t STATIC_PARTIALS_MAX = 65536; // {ssim, l1} pairs: up to 3 x 4096 x 4096 (49152 CTAs)
static __device__ __align__(256) float g_static_partials[2 * STATIC_PARTIALS_MAX];

// In host code g_static_partials is only a stand-in for the symbol, not a device address: the
// address a kernel can use comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
static float* static_partials_on_current_device(int* status) {
  static std::mutex mu;
  static float* ptr[STATIC_MAX_DEVICES] = {};
  int dev = -1;
  *status = check_cuda(cudaGetDevice(&dev), "l1_ssim forward_static: cudaGetDevice");
  if (*status != IGS_OK) return nullptr;
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) {
    set_error("l1_ssim forward_static: device %d beyond the %d devices supported", dev, STATIC_MAX_DEVICES);
    *status = IGS_ERR_INVALID;
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    float* p = nullptr;
    *status = check_cuda(cudaGetSymbolAddress((void**)&p, g_static_partials),
                         "l1_ssim forward_static: cudaGetSymbolAddress");
    if (*status != IGS_OK) return nullptr;
    ptr[dev] = p;
  }
  return ptr[dev];
}

extern "C" int igs_l1_ssim_forward_static(const float* img1, const float* img2, int32_t C, int32_t H, int32_t W,
                                          float lambda_dssim, float* dA, float* dB, float* dC, float* out3,
                                          igs_stream_t stream) {
  if (C <= 0 || H <= 0 || W <= 0 || !img1 || !img2 || !dA || !dB || !dC || !out3) {
    set_error("l1_ssim forward_static: bad arguments"); return IGS_ERR_INVALID;
  }
  const size_t n = igs_l1_ssim_partials(C, H, W);
  if (n > STATIC_PARTIALS_MAX) {
    set_error("l1_ssim forward_static: %d x %d x %d needs %zu partial sums, the static scratch holds %zu "
              "(use igs_l1_ssim_forward with a caller workspace)", C, H, W, n, STATIC_PARTIALS_MAX);
    return IGS_ERR_INVALID;
  }
  int status = IGS_OK;
  float* partial = static_partials_on_current_device(&status);
  if (!partial) return status;
  // Nothing in the array has to start at zero: every CTA of the forward kernel overwrites its own pair
  // and the finish kernel reads exactly the n pairs of this call, so what the last call left is dead.
  static const Win win = make_window();
  const dim3 grid((W + QT - 1) / QT, (H + QT - 1) / QT, C);
  l1_ssim_forward_kernel<<<grid, QTHREADS, 0, (cudaStream_t)stream>>>(img1, img2, H, W, win, dA, dB, dC, partial);
  l1_ssim_finish_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(partial, (int)n, 1.0f / ((float)C * (float)H * (float)W),
                                                            lambda_dssim, out3);
  count_launch(2);
  return stage_check(false, (cudaStream_t)stream, "l1_ssim forward_static");
}
