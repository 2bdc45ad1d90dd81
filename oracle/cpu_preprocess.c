/* TEST INFRASTRUCTURE -- see cpu_raster.h.  Per-Gaussian forward stage (K1),
 * prefix sum (K2) and markVisible (K10) of the reference, restated in C. */
#include "cpu_raster.h"
#include "cpu_math.h"
#include <omp.h>

int orc_num_threads(void) { return omp_get_max_threads(); }

/* DGR forward.cu:118-152 computeCov3D.  M = S*R, Sigma = M^T*M (glm, column
 * major).  The quaternion is NOT normalised (forward.cu:127). */
static void build_R(const float* q, m3* R)
{
  float r = q[0], x = q[1], y = q[2], z = q[3];
  /* glm::mat3(a,b,c, d,e,f, g,h,i) fills COLUMNS: R->m[0] = (a,b,c) ... */
  R->m[0][0] = 1.f - 2.f * (y * y + z * z);           /* SASS: FADD(yy,zz): both products are shared, none fused */
  R->m[0][1] = 2.f * fmaf(x, y, -(r * z));
  R->m[0][2] = 2.f * fmaf(r, y, x * z);               /* SASS: x*z is the shared rounded product */
  R->m[1][0] = 2.f * fmaf(x, y, r * z);
  R->m[1][1] = fmaf(-2.f, fmaf(x, x, z * z), 1.f);
  R->m[1][2] = 2.f * fmaf(y, z, -(r * x));
  R->m[2][0] = 2.f * fmaf(-r, y, x * z);
  R->m[2][1] = 2.f * fmaf(y, z, r * x);
  R->m[2][2] = fmaf(-2.f, fmaf(x, x, y * y), 1.f);
}

static void cov3d_from_scale_rot(const float* scale, float mod, const float* q, float* cov3D)
{
  m3 S = m3_zero(), R, M, Mt, Sigma;
  S.m[0][0] = mod * scale[0];
  S.m[1][1] = mod * scale[1];
  S.m[2][2] = mod * scale[2];
  build_R(q, &R);
  M = m3_mul(&S, &R);
  Mt = m3_transpose(&M);
  Sigma = m3_mul(&Mt, &M);
  cov3D[0] = Sigma.m[0][0];
  cov3D[1] = Sigma.m[0][1];
  cov3D[2] = Sigma.m[0][2];
  cov3D[3] = Sigma.m[1][1];
  cov3D[4] = Sigma.m[1][2];
  cov3D[5] = Sigma.m[2][2];
}

/* DGR forward.cu:156-200 computeNorm3D: axis of the smallest RAW scale
 * (strict-inequality cascade, ties fall to x), rotated, flipped to the camera. */
static void norm3d_from_scale_rot(const float* scale, const float* q, const float* mean,
                                  const float* campos, float* out)
{
  m3 R, Rt;
  float axis[3] = {1.f, 0.f, 0.f}, n[3], ray[3];
  build_R(q, &R);
  if (scale[0] > scale[2] && scale[1] > scale[2]) { axis[0] = 0.f; axis[2] = 1.f; }
  else if (scale[0] > scale[1] && scale[2] > scale[1]) { axis[0] = 0.f; axis[1] = 1.f; }
  Rt = m3_transpose(&R);
  m3_mulv(&Rt, axis, n);
  ray[0] = mean[0] - campos[0];
  ray[1] = mean[1] - campos[1];
  ray[2] = mean[2] - campos[2];
  if (dot3v(ray, n) > 0.f) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
  out[0] = n[0]; out[1] = n[1]; out[2] = n[2];
}

/* DGR forward.cu:74-113 computeCov2D (EWA, +0.3 low-pass). */
void orc_cov2d(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy,
               const float* cov3D, const float* view, float* cov /*[3]*/, m3* T_out,
               float* t_out /*[3] clamped t*/, float* txtz_tytz /*[2]*/)
{
  float t[3];
  xform4x3(mean, view, t);
  const float limx = 1.3f * tan_fovx, limy = 1.3f * tan_fovy;
  const float txtz = t[0] / t[2], tytz = t[1] / t[2];
  t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
  t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
  m3 J = m3_zero(), Wm, T, Tt, V, X, C;
  J.m[0][0] = fx / t[2];
  J.m[0][2] = -(fx * t[0]) / (t[2] * t[2]);
  J.m[1][1] = fy / t[2];
  J.m[1][2] = -(fy * t[1]) / (t[2] * t[2]);
  Wm.m[0][0] = view[0]; Wm.m[0][1] = view[4]; Wm.m[0][2] = view[8];
  Wm.m[1][0] = view[1]; Wm.m[1][1] = view[5]; Wm.m[1][2] = view[9];
  Wm.m[2][0] = view[2]; Wm.m[2][1] = view[6]; Wm.m[2][2] = view[10];
  T = m3_mul(&Wm, &J);
  V.m[0][0] = cov3D[0]; V.m[0][1] = cov3D[1]; V.m[0][2] = cov3D[2];
  V.m[1][0] = cov3D[1]; V.m[1][1] = cov3D[3]; V.m[1][2] = cov3D[4];
  V.m[2][0] = cov3D[2]; V.m[2][1] = cov3D[4]; V.m[2][2] = cov3D[5];
  Tt = m3_transpose(&T);
  m3 Vt = m3_transpose(&V);
  X = m3_mul(&Tt, &Vt);
  C = m3_mul(&X, &T);
  cov[0] = C.m[0][0] + 0.3f;
  cov[1] = C.m[0][1];
  cov[2] = C.m[1][1] + 0.3f;
  if (T_out) *T_out = T;
  if (t_out) { t_out[0] = t[0]; t_out[1] = t[1]; t_out[2] = t[2]; }
  if (txtz_tytz) { txtz_tytz[0] = txtz; txtz_tytz[1] = tytz; }
}

/* DGR forward.cu:20-71 computeColorFromSH. */
static void color_from_sh(int idx, int deg, int M, const float* means, const float* campos,
                          const float* shs, uint8_t* clamped, float* rgb)
{
  const float* pos = means + 3 * idx;
  float d[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
  float len = sqrtf(dot3v(d, d));
  float x = d[0] / len, y = d[1] / len, z = d[2] / len;
  const float* sh = shs + (size_t)idx * M * 3;
  for (int c = 0; c < 3; c++) {
#define SH(k) sh[3 * (k) + c]
    float res = SH_C0 * SH(0);
    if (deg > 0) {
      /* result - C1*y*sh1 + C1*z*sh2 - C1*x*sh3 ; the first product is the
       * left operand of the first subtraction and is the one fused. */
      res = fmaf(-(SH_C1 * y), SH(1), SH_C0 * SH(0)); /* SASS: the degree-0 product is rounded, this one fused */
      res = fmaf(SH_C1 * z, SH(2), res);
      res = fmaf(-(SH_C1 * x), SH(3), res);
      if (deg > 1) {
        float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
        res = fmaf(SH_C2[0] * xy, SH(4), res);
        res = fmaf(SH_C2[1] * yz, SH(5), res);
        res = fmaf(SH_C2[2] * ((2.f * zz - xx) - yy), SH(6), res); /* xx, yy, zz are shared rounded products */
        res = fmaf(SH_C2[3] * xz, SH(7), res);
        res = fmaf(SH_C2[4] * (xx - yy), SH(8), res);
        if (deg > 2) {
          res = fmaf(SH_C3[0] * y * fmaf(3.f, xx, -yy), SH(9), res);
          res = fmaf(SH_C3[1] * xy * z, SH(10), res);
          res = fmaf(SH_C3[2] * y * (fmaf(4.f, zz, -xx) - yy), SH(11), res);
          res = fmaf(SH_C3[3] * z * fmaf(-3.f, yy, fmaf(-3.f, xx, 2.f * zz)), SH(12), res);
          res = fmaf(SH_C3[4] * x * (fmaf(4.f, zz, -xx) - yy), SH(13), res);
          res = fmaf(SH_C3[5] * z * (xx - yy), SH(14), res);
          res = fmaf(SH_C3[6] * x * fmaf(-3.f, yy, xx), SH(15), res);
        }
      }
    }
#undef SH
    res += 0.5f;
    clamped[3 * idx + c] = (res < 0.f);
    rgb[3 * idx + c] = fmaxf(res, 0.f);
  }
}

/* DGR forward.cu:203-316 preprocessCUDA. */
void orc_preprocess(const orc_scene* s, orc_geom* g)
{
  const float focal_y = s->H / (2.0f * s->tan_fovy); /* rasterizer_impl.cu:228-229 */
  const float focal_x = s->W / (2.0f * s->tan_fovx);
  const int gx = (s->W + ORC_BLOCK_X - 1) / ORC_BLOCK_X;
  const int gy = (s->H + ORC_BLOCK_Y - 1) / ORC_BLOCK_Y;
#pragma omp parallel for schedule(static)
  for (int idx = 0; idx < s->P; idx++) {
    g->radii[idx] = 0;
    g->tiles_touched[idx] = 0;
    const float* p = s->means3D + 3 * idx;
    float pv[3], ph[4];
    /* in_frustum, auxiliary.h:140-165: near cull only */
    xform4x3(p, s->viewmatrix, pv);
    if (pv[2] <= ORC_NEAR) continue;
    xform4x4(p, s->projmatrix, ph);
    float p_w = 1.0f / (ph[3] + 0.0000001f);
    float proj_x = ph[0] * p_w, proj_y = ph[1] * p_w;

    const float* cov3D;
    if (s->cov3Ds_precomp) cov3D = s->cov3Ds_precomp + 6 * idx;
    else {
      cov3d_from_scale_rot(s->scales + 3 * idx, s->scale_modifier, s->rotations + 4 * idx,
                           g->cov3D + 6 * idx);
      cov3D = g->cov3D + 6 * idx;
    }
    if (!s->norm3Ds_precomp)
      norm3d_from_scale_rot(s->scales + 3 * idx, s->rotations + 4 * idx, p, s->campos,
                            g->norm3D + 3 * idx);

    float cov[3];
    orc_cov2d(p, focal_x, focal_y, s->tan_fovx, s->tan_fovy, cov3D, s->viewmatrix, cov, 0, 0, 0);
    float det = fmaf(cov[0], cov[2], -(cov[1] * cov[1]));
    if (det == 0.0f) continue;
    float det_inv = 1.f / det;
    float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
    float mid = 0.5f * (cov[0] + cov[2]);
    float disc = sqrtf(fmaxf(0.1f, fmaf(mid, mid, -det)));
    float lambda1 = mid + disc, lambda2 = mid - disc;
    float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
    float px = ndc2pix(proj_x, s->W), py = ndc2pix(proj_y, s->H);
    int rmin[2], rmax[2];
    get_rect(px, py, f2i_rz(my_radius), gx, gy, rmin, rmax);
    if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;

    if (!s->colors_precomp)
      color_from_sh(idx, s->D, s->M, s->means3D, s->campos, s->shs, g->clamped, g->rgb);

    g->depths[idx] = pv[2];
    g->radii[idx] = f2i_rz(my_radius);
    g->means2D[2 * idx] = px;
    g->means2D[2 * idx + 1] = py;
    g->conic_opacity[4 * idx + 0] = conic[0];
    g->conic_opacity[4 * idx + 1] = conic[1];
    g->conic_opacity[4 * idx + 2] = conic[2];
    g->conic_opacity[4 * idx + 3] = s->opacities[idx];
    g->tiles_touched[idx] = (uint32_t)((rmax[1] - rmin[1]) * (rmax[0] - rmin[0]));
  }
}

/* DGR rasterizer_impl.cu:280 (cub::DeviceScan::InclusiveSum, uint32). */
int64_t orc_scan(int P, const uint32_t* tiles_touched, uint32_t* point_offsets)
{
  uint32_t acc = 0;
  for (int i = 0; i < P; i++) { acc += tiles_touched[i]; point_offsets[i] = acc; }
  return P > 0 ? (int64_t)acc : 0;
}

/* DGR rasterizer_impl.cu:54-66 checkFrustum -> in_frustum(prefiltered=false). */
void orc_mark_visible(int P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present)
{
  (void)projmatrix;
#pragma omp parallel for schedule(static)
  for (int idx = 0; idx < P; idx++) {
    float pv[3];
    xform4x3(means3D + 3 * idx, viewmatrix, pv);
    present[idx] = pv[2] > ORC_NEAR;
  }
}
