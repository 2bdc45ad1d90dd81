/* TEST INFRASTRUCTURE -- see cpu_raster.h.  Reverse-order compositing
 * gradients of the reference (K7, DGR backward.cu:473-712).  The per-pair
 * terms are fp32 like the reference's; the per-Gaussian sums (float atomics of
 * unspecified order there) are accumulated in double and rounded once, which
 * is the order-independent value both implementations scatter around. */
#include "cpu_raster.h"
#include "cpu_math.h"
#include <stdlib.h>

void orc_render_backward(const orc_scene* s, const orc_geom* g, const uint32_t* ranges,
                         const uint32_t* point_list, const float* out_alpha,
                         const uint32_t* n_contrib, const float* dL_dpix,
                         const float* dL_dpix_depth, const float* dL_dpix_norm,
                         const float* dL_dpix_alpha, const float* dL_dpix_extra,
                         float* dL_dmean2D, float* dL_dconic, float* dL_dopacity,
                         float* dL_dcolor, float* dL_ddepth, float* dL_dnorm3D, float* dL_dextra)
{
  const int W = s->W, H = s->H, ED = s->ED, P = s->P;
  const int gx = (W + ORC_BLOCK_X - 1) / ORC_BLOCK_X;
  const int gy = (H + ORC_BLOCK_Y - 1) / ORC_BLOCK_Y;
  const size_t HW = (size_t)H * W;
  const float* colors = s->colors_precomp ? s->colors_precomp : g->rgb; /* rasterizer_impl.cu:404-405 */
  const float* norms = s->norm3Ds_precomp ? s->norm3Ds_precomp : g->norm3D;
  const int K = 13 + ED; /* m2d.xy, conic.xyw, opacity, color3, depth, norm3, extra */
  double* acc = (double*)calloc((size_t)(P ? P : 1) * K, sizeof(double));
  const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H; /* backward.cu:564-565 */

#pragma omp parallel for schedule(dynamic, 1)
  for (int tile = 0; tile < gx * gy; tile++) {
    const int tx = tile % gx, ty = tile / gx;
    const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
    for (int ly = 0; ly < ORC_BLOCK_Y; ly++)
      for (int lx = 0; lx < ORC_BLOCK_X; lx++) {
        const int px = tx * ORC_BLOCK_X + lx, py = ty * ORC_BLOCK_Y + ly;
        if (px >= W || py >= H) continue;
        const size_t pix = (size_t)W * py + px;
        const float pxf = (float)px, pyf = (float)py;
        const float T_final = 1 - out_alpha[pix]; /* backward.cu:528 */
        float T = T_final;
        uint32_t contributor = r1 - r0;
        const uint32_t last_contributor = n_contrib[pix];
        float accum_rec[3] = {0, 0, 0}, accum_red = 0, accum_ren[3] = {0, 0, 0}, accum_rea = 0;
        float accum_ree[64], last_extra[64], dpe[64];
        float dpc[3], dpn[3];
        for (int i = 0; i < 3; i++) dpc[i] = dL_dpix[i * HW + pix];
        const float dpd = dL_dpix_depth[pix];
        for (int i = 0; i < 3; i++) dpn[i] = dL_dpix_norm[i * HW + pix];
        const float dpa = dL_dpix_alpha[pix];
        for (int i = 0; i < ED; i++) { dpe[i] = dL_dpix_extra[i * HW + pix]; accum_ree[i] = 0; last_extra[i] = 0; }
        float last_alpha = 0, last_color[3] = {0, 0, 0}, last_depth = 0, last_norm[3] = {0, 0, 0};
        float bg_dot = 0;
        for (int i = 0; i < 3; i++) bg_dot += s->bg[i] * dpc[i]; /* backward.cu:684-686 */

        for (uint32_t k = r1; k-- > r0;) {
          contributor--;
          if (contributor >= last_contributor) continue; /* backward.cu:595-597 */
          const uint32_t id = point_list[k];
          const float dx = g->means2D[2 * id] - pxf, dy = g->means2D[2 * id + 1] - pyf;
          const float* co = g->conic_opacity + 4 * id;
          const float power = fmaf(fmaf(co[0] * dx, dx, (co[2] * dy) * dy), -0.5f, -((co[1] * dx) * dy));
          if (power > 0.0f) continue;
          const float G = expf(power);
          const float alpha = fminf(0.99f, co[3] * G);
          if (alpha < 1.0f / 255.0f) continue;
          T = T / (1.f - alpha); /* backward.cu:612: T recovered by division */
          const float weight = alpha * T;
          double* a = acc + (size_t)id * K;
          float dL_dalpha = 0.0f;
          for (int ch = 0; ch < 3; ch++) {
            const float c = colors[3 * id + ch];
            accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
            last_color[ch] = c;
            dL_dalpha += (c - accum_rec[ch]) * dpc[ch];
#pragma omp atomic
            a[6 + ch] += (double)(weight * dpc[ch]);
          }
          const float dep = g->depths[id];
          accum_red = last_alpha * last_depth + (1.f - last_alpha) * accum_red;
          last_depth = dep;
          dL_dalpha += (dep - accum_red) * dpd;
#pragma omp atomic
          a[9] += (double)(weight * dpd);
          for (int ch = 0; ch < 3; ch++) {
            const float n = norms[3 * id + ch];
            accum_ren[ch] = last_alpha * last_norm[ch] + (1.f - last_alpha) * accum_ren[ch];
            last_norm[ch] = n;
            dL_dalpha += (n - accum_ren[ch]) * dpn[ch];
#pragma omp atomic
            a[10 + ch] += (double)(weight * dpn[ch]);
          }
          for (int ch = 0; ch < ED; ch++) {
            const float e = s->extra_attrs[(size_t)ED * id + ch];
            accum_ree[ch] = last_alpha * last_extra[ch] + (1.f - last_alpha) * accum_ree[ch];
            last_extra[ch] = e;
            dL_dalpha += (e - accum_ree[ch]) * dpe[ch];
#pragma omp atomic
            a[13 + ch] += (double)(weight * dpe[ch]);
          }
          accum_rea = last_alpha + (1.f - last_alpha) * accum_rea; /* backward.cu:674-675 */
          dL_dalpha += (1 - accum_rea) * dpa;
          dL_dalpha *= T;
          last_alpha = alpha;
          dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot; /* backward.cu:687 */

          const float dL_dG = co[3] * dL_dalpha;
          const float gdx = G * dx, gdy = G * dy;
          const float dG_ddelx = -gdx * co[0] - gdy * co[1];
          const float dG_ddely = -gdy * co[2] - gdx * co[1];
#pragma omp atomic
          a[0] += (double)(dL_dG * dG_ddelx * ddelx_dx);
#pragma omp atomic
          a[1] += (double)(dL_dG * dG_ddely * ddely_dy);
#pragma omp atomic
          a[2] += (double)(-0.5f * gdx * dx * dL_dG);
#pragma omp atomic
          a[3] += (double)(-0.5f * gdx * dy * dL_dG);
#pragma omp atomic
          a[4] += (double)(-0.5f * gdy * dy * dL_dG);
#pragma omp atomic
          a[5] += (double)(G * dL_dalpha);
        }
      }
  }
  for (int i = 0; i < P; i++) {
    const double* a = acc + (size_t)i * K;
    dL_dmean2D[3 * i] += (float)a[0];
    dL_dmean2D[3 * i + 1] += (float)a[1]; /* .z never written, backward.cu:700-701 */
    dL_dconic[4 * i] += (float)a[2];
    dL_dconic[4 * i + 1] += (float)a[3];
    dL_dconic[4 * i + 3] += (float)a[4]; /* .z never written, backward.cu:704-706 */
    dL_dopacity[i] += (float)a[5];
    for (int ch = 0; ch < 3; ch++) dL_dcolor[3 * i + ch] += (float)a[6 + ch];
    dL_ddepth[i] += (float)a[9];
    for (int ch = 0; ch < 3; ch++) dL_dnorm3D[3 * i + ch] += (float)a[10 + ch];
    for (int ch = 0; ch < ED; ch++) dL_dextra[(size_t)ED * i + ch] += (float)a[13 + ch];
  }
  free(acc);
}
