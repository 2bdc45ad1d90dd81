// SURVEY.md 8(f) row f3, the rest of it: the geometry regularisers train_face.py puts on the
// rasterizer's depth / normal / alpha images, and their gradients -- the dL_ddepth, dL_dnorm and
// dL_dalpha images igs_backward consumes -- written directly in CHW:
//
//   :469  w_n * (1 - normal_gt * normal).sum(0)[head ^ mouth].mean()
//   :491  w_d * (normalize(depth)[face ^ mouth] - normalize(depth_mono)[face ^ mouth]).abs().mean()
//         normalize (utils/loss_utils.py:17-20) works on the ROWS of the [H,W] map:
//             (x - mean_row) / (std_row + 1e-2 * std(all elements)),  unbiased standard deviations
//   :540  w_a * (((1 - alpha) * head).mean() + (alpha * ~head).mean())
//
// Forward: one CTA per image row sums what the three terms need (double accumulation), a second
// pass per row evaluates the depth term with the row / global statistics and leaves the two row sums
// its gradient needs (A = sum u, B = sum u (x - mean_row), u = d|diff| / dz), one CTA adds the rows up.
// Backward: one pass, every element once.  The reference runs ~45 element-wise / reduction /
// boolean-indexing kernels forward and as many backward, with a dozen full-frame temporaries.
// HBM bound: forward reads 11 floats + 3 mask bytes per pixel (depth and its reference twice),
// backward reads the same and writes 5 floats.
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

constexpr int GL_THREADS = 256;
constexpr int GL_ROW = 12; // doubles per image row in the workspace
// row workspace: 0 sum d | 1 sum d^2 | 2 sum m | 3 sum m^2 | 4 |head^mouth| | 5 |face^mouth| | 6 sum_{head^mouth} (3 - <n_gt, n>)
//                7 sum (1-a) head + a (1-head) | 8 sum_{face^mouth} |z - z_m| | 9 A' = sum sign | 10 B' = sum sign (d - mean_row)
// glob (8 doubles): 0 G' = sum_rows B'_r / s_r^2 | 1 |head^mouth| | 2 |face^mouth|

template <int K>
__device__ __forceinline__ void cta_sum(double (&v)[K], double* s_red /* [K][GL_THREADS/32] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if (lane == 0) s_red[k * (GL_THREADS / 32) + warp] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; k++) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < GL_THREADS / 32; w++) t += s_red[k * (GL_THREADS / 32) + w];
    v[k] = t;
  }
  __syncthreads();
}

struct RowStats {
  double mu, sd, s, mu_g, sd_g;
};

// statistics normalize() needs for row r of a map whose per-row sums sit at columns c0 (sum), c0+1 (sum of squares)
__device__ __forceinline__ RowStats row_stats(const double* rows, int H, int W, int r, int c0, double* s_red) {
  double g[2] = {0.0, 0.0};
  for (int i = threadIdx.x; i < H; i += GL_THREADS) { g[0] += rows[(size_t)i * GL_ROW + c0]; g[1] += rows[(size_t)i * GL_ROW + c0 + 1]; }
  cta_sum(g, s_red);
  RowStats st;
  const double n = (double)W, N = (double)H * (double)W;
  st.mu_g = g[0] / N;
  st.sd_g = sqrt(fmax(0.0, (g[1] - N * st.mu_g * st.mu_g) / (N - 1.0)));
  const double a = rows[(size_t)r * GL_ROW + c0], b = rows[(size_t)r * GL_ROW + c0 + 1];
  st.mu = a / n;
  st.sd = sqrt(fmax(0.0, (b - n * st.mu * st.mu) / (n - 1.0)));
  st.s = st.sd + 1e-2 * st.sd_g;
  return st;
}

__global__ void __launch_bounds__(GL_THREADS) geom_rows_kernel(const float* __restrict__ depth, const float* __restrict__ norm,
                                                               const float* __restrict__ alpha, const float* __restrict__ depth_mono,
                                                               const float* __restrict__ normal_gt, const uint8_t* __restrict__ head,
                                                               const uint8_t* __restrict__ face, const uint8_t* __restrict__ mouth,
                                                               int H, int W, double* __restrict__ rows) {
  __shared__ double s_red[8 * GL_THREADS / 32];
  const int r = blockIdx.x;
  const size_t HW = (size_t)H * W, base = (size_t)r * W;
  double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int x = threadIdx.x; x < W; x += GL_THREADS) {
    const size_t p = base + x;
    const double d = depth[p], m = depth_mono[p];
    const bool hd = head[p] != 0, fc = face[p] != 0, mo = mouth[p] != 0;
    v[0] += d; v[1] += d * d; v[2] += m; v[3] += m * m;
    if (hd != mo) {
      v[4] += 1.0;
      v[6] += 3.0 - ((double)normal_gt[p] * norm[p] + (double)normal_gt[HW + p] * norm[HW + p] +
                     (double)normal_gt[2 * HW + p] * norm[2 * HW + p]);
    }
    if (fc != mo) v[5] += 1.0;
    const double a = alpha[p];
    v[7] += hd ? 1.0 - a : a;
  }
  cta_sum(v, s_red);
  if (threadIdx.x < 8) rows[(size_t)r * GL_ROW + threadIdx.x] = v[threadIdx.x];
}

__global__ void __launch_bounds__(GL_THREADS) geom_depth_kernel(const float* __restrict__ depth, const float* __restrict__ depth_mono,
                                                                const uint8_t* __restrict__ face, const uint8_t* __restrict__ mouth,
                                                                int H, int W, double* __restrict__ rows) {
  __shared__ double s_red[3 * GL_THREADS / 32];
  const int r = blockIdx.x;
  const RowStats sd = row_stats(rows, H, W, r, 0, s_red), sm = row_stats(rows, H, W, r, 2, s_red);
  const size_t base = (size_t)r * W;
  double v[3] = {0, 0, 0};
  for (int x = threadIdx.x; x < W; x += GL_THREADS) {
    const size_t p = base + x;
    if ((face[p] != 0) == (mouth[p] != 0)) continue;
    const double d = depth[p];
    const double diff = (d - sd.mu) / sd.s - ((double)depth_mono[p] - sm.mu) / sm.s;
    const double sg = diff > 0.0 ? 1.0 : (diff < 0.0 ? -1.0 : 0.0);
    v[0] += fabs(diff); v[1] += sg; v[2] += sg * (d - sd.mu);
  }
  cta_sum(v, s_red);
  if (threadIdx.x < 3) rows[(size_t)r * GL_ROW + 8 + threadIdx.x] = v[threadIdx.x];
}

__global__ void __launch_bounds__(GL_THREADS) geom_final_kernel(int H, int W, const double* __restrict__ rows, float w_normal,
                                                                float w_depth, float w_alpha, float* __restrict__ loss,
                                                                double* __restrict__ glob) {
  __shared__ double s_red[7 * GL_THREADS / 32];
  double g[2] = {0.0, 0.0};
  for (int i = threadIdx.x; i < H; i += GL_THREADS) { g[0] += rows[(size_t)i * GL_ROW]; g[1] += rows[(size_t)i * GL_ROW + 1]; }
  cta_sum(g, s_red);
  const double n = (double)W, N = (double)H * (double)W;
  const double mu_g = g[0] / N, sd_g = sqrt(fmax(0.0, (g[1] - N * mu_g * mu_g) / (N - 1.0)));
  double v[6] = {0, 0, 0, 0, 0, 0}; // |M1|, |M2|, normal sum, alpha sum, depth sum, G'
  for (int i = threadIdx.x; i < H; i += GL_THREADS) {
    const double* row = rows + (size_t)i * GL_ROW;
    const double mu = row[0] / n, sd = sqrt(fmax(0.0, (row[1] - n * mu * mu) / (n - 1.0))), s = sd + 1e-2 * sd_g;
    v[0] += row[4]; v[1] += row[5]; v[2] += row[6]; v[3] += row[7]; v[4] += row[8];
    v[5] += s > 0.0 ? row[10] / (s * s) : 0.0;
  }
  cta_sum(v, s_red);
  if (threadIdx.x == 0) {
    loss[0] = (float)(w_normal * v[2] / v[0]);
    loss[1] = (float)(w_depth * v[4] / v[1]);
    loss[2] = (float)(w_alpha * v[3] / N);
    glob[0] = v[5]; glob[1] = v[0]; glob[2] = v[1];
  }
}

__global__ void __launch_bounds__(GL_THREADS) geom_backward_kernel(const float* __restrict__ depth, const float* __restrict__ depth_mono,
                                                                   const float* __restrict__ normal_gt, const uint8_t* __restrict__ head,
                                                                   const uint8_t* __restrict__ face, const uint8_t* __restrict__ mouth,
                                                                   int H, int W, const double* __restrict__ rows,
                                                                   const double* __restrict__ glob, float w_normal, float w_depth,
                                                                   float w_alpha, const float* __restrict__ grad_out,
                                                                   float* __restrict__ dL_ddepth, float* __restrict__ dL_dnorm,
                                                                   float* __restrict__ dL_dalpha) {
  __shared__ double s_red[2 * GL_THREADS / 32];
  const int r = blockIdx.x;
  const RowStats sd = row_stats(rows, H, W, r, 0, s_red), sm = row_stats(rows, H, W, r, 2, s_red);
  const size_t HW = (size_t)H * W, base = (size_t)r * W;
  const double g = (double)__ldg(grad_out), N = (double)H * (double)W, n = (double)W;
  const double ws = glob[2] > 0.0 ? w_depth * g / glob[2] : 0.0; // dL/dz = ws * sign(diff) on face ^ mouth
  const double A = ws * rows[(size_t)r * GL_ROW + 9];
  const double B = sd.s > 0.0 ? ws * rows[(size_t)r * GL_ROW + 10] / (sd.s * sd.s) : 0.0;
  const double G = ws * glob[0];
  const double k_row = (sd.sd > 0.0 && n > 1.0) ? B / ((n - 1.0) * sd.sd) : 0.0;
  const double k_glob = (sd.sd_g > 0.0) ? 1e-2 * G / ((N - 1.0) * sd.sd_g) : 0.0;
  const double kn = glob[1] > 0.0 ? -w_normal * g / glob[1] : 0.0;
  const float ka = (float)(w_alpha * g / N);
  for (int x = threadIdx.x; x < W; x += GL_THREADS) {
    const size_t p = base + x;
    const bool hd = head[p] != 0, fc = face[p] != 0, mo = mouth[p] != 0;
    const double d = depth[p];
    double u = 0.0;
    if (fc != mo) {
      const double diff = (d - sd.mu) / sd.s - ((double)depth_mono[p] - sm.mu) / sm.s;
      u = diff > 0.0 ? ws : (diff < 0.0 ? -ws : 0.0);
    }
    dL_ddepth[p] = (float)(u / sd.s - A / (n * sd.s) - k_row * (d - sd.mu) - k_glob * (d - sd.mu_g));
    const float m1 = hd != mo ? (float)kn : 0.f;
    dL_dnorm[p] = m1 * normal_gt[p];
    dL_dnorm[HW + p] = m1 * normal_gt[HW + p];
    dL_dnorm[2 * HW + p] = m1 * normal_gt[2 * HW + p];
    dL_dalpha[p] = hd ? -ka : ka;
  }
}

} // namespace igs

using namespace igs;

extern "C" {

size_t igs_geom_loss_workspace(int32_t H) { return ((size_t)(H > 0 ? H : 0) * GL_ROW + 8) * sizeof(double); }

int igs_geom_loss_forward(const float* depth, const float* norm, const float* alpha, const float* depth_mono,
                          const float* normal_gt, const uint8_t* head_mask, const uint8_t* face_mask,
                          const uint8_t* mouth_mask, int32_t H, int32_t W, float w_normal, float w_depth, float w_alpha,
                          void* workspace, float* loss3, igs_stream_t stream) {
  if (H <= 0 || W <= 0 || !depth || !norm || !alpha || !depth_mono || !normal_gt || !head_mask || !face_mask || !mouth_mask ||
      !workspace || !loss3) {
    set_error("geom_loss forward: bad arguments"); return IGS_ERR_INVALID;
  }
  cudaStream_t s = (cudaStream_t)stream;
  double* rows = (double*)workspace;
  double* glob = rows + (size_t)H * GL_ROW;
  geom_rows_kernel<<<H, GL_THREADS, 0, s>>>(depth, norm, alpha, depth_mono, normal_gt, head_mask, face_mask, mouth_mask, H, W, rows);
  geom_depth_kernel<<<H, GL_THREADS, 0, s>>>(depth, depth_mono, face_mask, mouth_mask, H, W, rows);
  geom_final_kernel<<<1, GL_THREADS, 0, s>>>(H, W, rows, w_normal, w_depth, w_alpha, loss3, glob);
  count_launch(3);
  return stage_check(false, s, "geom_loss forward");
}

int igs_geom_loss_backward(const float* depth, const float* depth_mono, const float* normal_gt, const uint8_t* head_mask,
                           const uint8_t* face_mask, const uint8_t* mouth_mask, int32_t H, int32_t W, float w_normal,
                           float w_depth, float w_alpha, const void* workspace, const float* grad_out, float* dL_ddepth,
                           float* dL_dnorm, float* dL_dalpha, igs_stream_t stream) {
  if (H <= 0 || W <= 0 || !depth || !depth_mono || !normal_gt || !head_mask || !face_mask || !mouth_mask || !workspace ||
      !grad_out || !dL_ddepth || !dL_dnorm || !dL_dalpha) {
    set_error("geom_loss backward: bad arguments"); return IGS_ERR_INVALID;
  }
  const double* rows = (const double*)workspace;
  geom_backward_kernel<<<H, GL_THREADS, 0, (cudaStream_t)stream>>>(depth, depth_mono, normal_gt, head_mask, face_mask, mouth_mask, H,
                                                                   W, rows, rows + (size_t)H * GL_ROW, w_normal, w_depth, w_alpha,
                                                                   grad_out, dL_ddepth, dL_dnorm, dL_dalpha);
  count_launch();
  return stage_check(false, (cudaStream_t)stream, "geom_loss backward");
}

} // extern "C"
