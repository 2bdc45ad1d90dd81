/* TEST INFRASTRUCTURE -- see cpu_raster.h.  Per-tile front-to-back alpha
 * compositing of the reference (K6, DGR forward.cu:321-456), one pixel at a
 * time.  The block-level vote / batch staging of the CUDA kernel only decide
 * how much work is skipped, never the value of a pixel, so they have no
 * counterpart here. */
#include "cpu_raster.h"
#include "cpu_math.h"

#define ALPHA_MIN (1.0f / 255.0f)
#define T_MIN 0.0001f
/* relative margins inside which a threshold test may flip between expf
 * implementations (glibc here, MUFU.EX2-based on the GPU) */
#define BORDER_ALPHA 4e-6f
#define BORDER_T 1e-4f

void orc_render(const orc_scene* s, const orc_geom* g, const uint32_t* ranges,
                const uint32_t* point_list, float* out_color, float* out_depth, float* out_norm,
                float* out_alpha, float* out_extra, uint32_t* n_contrib, uint8_t* borderline,
                int64_t* counters)
{
  const int W = s->W, H = s->H, ED = s->ED;
  const int gx = (W + ORC_BLOCK_X - 1) / ORC_BLOCK_X;
  const int gy = (H + ORC_BLOCK_Y - 1) / ORC_BLOCK_Y;
  const size_t HW = (size_t)H * W;
  const float* features = s->colors_precomp ? s->colors_precomp : g->rgb; /* rasterizer_impl.cu:324 */
  const float* norms = s->norm3Ds_precomp ? s->norm3Ds_precomp : g->norm3D;
  int64_t n_eval = 0, n_acc = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_eval, n_acc)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < ORC_BLOCK_Y; ly++)
      for (int lx = 0; lx < ORC_BLOCK_X; lx++) {
        const int px = tx * ORC_BLOCK_X + lx, py = ty * ORC_BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const size_t pix = (size_t)W * py + px;
        const float pxf = (float)px, pyf = (float)py;
        float T = 1.0f, C[3] = {0, 0, 0}, Dp = 0, N[3] = {0, 0, 0}, E[64];
        for (int ch = 0; ch < ED; ch++) E[ch] = 0;
        uint32_t contributor = 0, last_contributor = 0;
        uint8_t border = 0;
        for (uint32_t k = r0; k < r1; k++) {
          const uint32_t id = point_list[k];
          contributor++;
          n_eval++;
          const float dx = g->means2D[2 * id] - pxf, dy = g->means2D[2 * id + 1] - pyf;
          const float* co = g->conic_opacity + 4 * id;
          /* forward.cu:406; contraction as in the reference's SASS:
           * fma(fma(dx, cx*dx, (cz*dy)*dy), -0.5, -((cy*dx)*dy)) */
          const float power = fmaf(fmaf(co[0] * dx, dx, (co[2] * dy) * dy), -0.5f, -((co[1] * dx) * dy));
          if (power > 0.0f) continue;
          const float a_raw = co[3] * expf(power);
          const float alpha = fminf(0.99f, a_raw);
          if (fabsf(alpha - ALPHA_MIN) <= BORDER_ALPHA * ALPHA_MIN) border = 1;
          if (alpha < ALPHA_MIN) continue;
          const float test_T = T * (1 - alpha);
          if (fabsf(test_T - T_MIN) <= BORDER_T * T_MIN) border = 1;
          if (test_T < T_MIN) break; /* done = true */
          n_acc++;
          /* forward.cu:425-431: X += f*alpha*T  ->  fma(f*alpha, T, X) */
          for (int ch = 0; ch < 3; ch++) C[ch] = fmaf(features[3 * id + ch] * alpha, T, C[ch]);
          Dp = fmaf(g->depths[id] * alpha, T, Dp);
          for (int ch = 0; ch < 3; ch++) N[ch] = fmaf(norms[3 * id + ch] * alpha, T, N[ch]);
          for (int ch = 0; ch < ED; ch++) E[ch] = fmaf(s->extra_attrs[(size_t)ED * id + ch] * alpha, T, E[ch]);
          T = test_T;
          last_contributor = contributor;
        }
        /* forward.cu:442-455 */
        out_alpha[pix] = 1 - T;
        n_contrib[pix] = last_contributor;
        for (int ch = 0; ch < 3; ch++) out_color[ch * HW + pix] = fmaf(T, s->bg[ch], C[ch]);
        out_depth[pix] = Dp;
        for (int ch = 0; ch < 3; ch++) out_norm[ch * HW + pix] = N[ch];
        for (int ch = 0; ch < ED; ch++) out_extra[ch * HW + pix] = E[ch];
        if (borderline) borderline[pix] = border;
      }
  }
  if (counters) { counters[0] = n_eval; counters[1] = n_acc; }
}
