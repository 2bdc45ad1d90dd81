/* TEST INFRASTRUCTURE -- see cpu_raster.h.  Per-Gaussian backward of the
 * reference: K8 computeCov2DCUDA (DGR backward.cu:144-274) and K9
 * preprocessCUDA (backward.cu:397-470) with its device helpers (SH :20-139,
 * cov3D :278-341, normal :344-392). */
#include "cpu_raster.h"
#include "cpu_math.h"

void orc_cov2d(const float* mean, float fx, float fy, float tan_fovx, float tan_fovy,
               const float* cov3D, const float* view, float* cov, m3* T_out, float* t_out,
               float* txtz_tytz);

/* DGR backward.cu:287-291 (same literal as forward.cu:134-138) */
static void build_R(const float* q, m3* R)
{
  float r = q[0], x = q[1], y = q[2], z = q[3];
  R->m[0][0] = 1.f - 2.f * (y * y + z * z);
  R->m[0][1] = 2.f * (x * y - r * z);
  R->m[0][2] = 2.f * (x * z + r * y);
  R->m[1][0] = 2.f * (x * y + r * z);
  R->m[1][1] = 1.f - 2.f * (x * x + z * z);
  R->m[1][2] = 2.f * (y * z - r * x);
  R->m[2][0] = 2.f * (x * z - r * y);
  R->m[2][1] = 2.f * (y * z + r * x);
  R->m[2][2] = 1.f - 2.f * (x * x + y * y);
}

/* shared tail of backward.cu:333-336 and :385-388; A is indexed A[col][row] */
static void quat_grad(const float* q, const m3* A, float* dq)
{
  const float r = q[0], x = q[1], y = q[2], z = q[3];
#define a(c, rr) A->m[c][rr]
  dq[0] = 2 * z * (a(0, 1) - a(1, 0)) + 2 * y * (a(2, 0) - a(0, 2)) + 2 * x * (a(1, 2) - a(2, 1));
  dq[1] = 2 * y * (a(1, 0) + a(0, 1)) + 2 * z * (a(2, 0) + a(0, 2)) + 2 * r * (a(1, 2) - a(2, 1)) -
          4 * x * (a(2, 2) + a(1, 1));
  dq[2] = 2 * x * (a(1, 0) + a(0, 1)) + 2 * r * (a(2, 0) - a(0, 2)) + 2 * z * (a(1, 2) + a(2, 1)) -
          4 * y * (a(2, 2) + a(0, 0));
  dq[3] = 2 * r * (a(0, 1) - a(1, 0)) + 2 * x * (a(2, 0) + a(0, 2)) + 2 * y * (a(1, 2) + a(2, 1)) -
          4 * z * (a(1, 1) + a(0, 0));
#undef a
}

/* K8, DGR backward.cu:144-274 */
static void cov2d_backward(const orc_scene* s, int idx, const float* cov3D, float fx, float fy,
                           const float* dL_dconics, float* dL_dmean /*assigned*/, float* dL_dcov)
{
  const float* view = s->viewmatrix;
  const float* mean = s->means3D + 3 * idx;
  const float dcx = dL_dconics[4 * idx], dcy = dL_dconics[4 * idx + 1], dcz = dL_dconics[4 * idx + 3];
  float cov[3], t[3], tt[2];
  m3 T;
  orc_cov2d(mean, fx, fy, s->tan_fovx, s->tan_fovy, cov3D, view, cov, &T, t, tt);
  const float limx = 1.3f * s->tan_fovx, limy = 1.3f * s->tan_fovy;
  const float x_grad_mul = (tt[0] < -limx || tt[0] > limx) ? 0.f : 1.f;
  const float y_grad_mul = (tt[1] < -limy || tt[1] > limy) ? 0.f : 1.f;
  const float a = cov[0], b = cov[1], c = cov[2];
  const float denom = a * c - b * b;
  float dL_da = 0, dL_db = 0, dL_dc = 0;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
#define Tm(cc, rr) T.m[cc][rr]
  if (denom2inv != 0) {
    dL_da = denom2inv * (-c * c * dcx + 2 * b * c * dcy + (denom - a * c) * dcz);
    dL_dc = denom2inv * (-a * a * dcz + 2 * a * b * dcy + (denom - a * c) * dcx);
    dL_db = denom2inv * 2 * (b * c * dcx - (denom + 2 * b * b) * dcy + a * b * dcz);
    dL_dcov[0] = (Tm(0, 0) * Tm(0, 0) * dL_da + Tm(0, 0) * Tm(1, 0) * dL_db + Tm(1, 0) * Tm(1, 0) * dL_dc);
    dL_dcov[3] = (Tm(0, 1) * Tm(0, 1) * dL_da + Tm(0, 1) * Tm(1, 1) * dL_db + Tm(1, 1) * Tm(1, 1) * dL_dc);
    dL_dcov[5] = (Tm(0, 2) * Tm(0, 2) * dL_da + Tm(0, 2) * Tm(1, 2) * dL_db + Tm(1, 2) * Tm(1, 2) * dL_dc);
    dL_dcov[1] = 2 * Tm(0, 0) * Tm(0, 1) * dL_da + (Tm(0, 0) * Tm(1, 1) + Tm(0, 1) * Tm(1, 0)) * dL_db +
                 2 * Tm(1, 0) * Tm(1, 1) * dL_dc;
    dL_dcov[2] = 2 * Tm(0, 0) * Tm(0, 2) * dL_da + (Tm(0, 0) * Tm(1, 2) + Tm(0, 2) * Tm(1, 0)) * dL_db +
                 2 * Tm(1, 0) * Tm(1, 2) * dL_dc;
    dL_dcov[4] = 2 * Tm(0, 2) * Tm(0, 1) * dL_da + (Tm(0, 1) * Tm(1, 2) + Tm(0, 2) * Tm(1, 1)) * dL_db +
                 2 * Tm(1, 1) * Tm(1, 2) * dL_dc;
  } else {
    for (int i = 0; i < 6; i++) dL_dcov[i] = 0;
  }
  /* Vrk[c][r], symmetric */
  const float V[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
  float dL_dT0[3], dL_dT1[3];
  for (int k = 0; k < 3; k++) {
    const float r0 = Tm(0, 0) * V[k][0] + Tm(0, 1) * V[k][1] + Tm(0, 2) * V[k][2];
    const float r1 = Tm(1, 0) * V[k][0] + Tm(1, 1) * V[k][1] + Tm(1, 2) * V[k][2];
    dL_dT0[k] = 2 * r0 * dL_da + r1 * dL_db;
    dL_dT1[k] = 2 * r1 * dL_dc + r0 * dL_db;
  }
#undef Tm
  /* W[c][r]: W[0]=(v0,v4,v8) W[1]=(v1,v5,v9) W[2]=(v2,v6,v10) */
  const float dL_dJ00 = view[0] * dL_dT0[0] + view[4] * dL_dT0[1] + view[8] * dL_dT0[2];
  const float dL_dJ02 = view[2] * dL_dT0[0] + view[6] * dL_dT0[1] + view[10] * dL_dT0[2];
  const float dL_dJ11 = view[1] * dL_dT1[0] + view[5] * dL_dT1[1] + view[9] * dL_dT1[2];
  const float dL_dJ12 = view[2] * dL_dT1[0] + view[6] * dL_dT1[1] + view[10] * dL_dT1[2];
  const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
  const float dL_dtx = x_grad_mul * -fx * tz2 * dL_dJ02;
  const float dL_dty = y_grad_mul * -fy * tz2 * dL_dJ12;
  const float dL_dtz = -fx * tz2 * dL_dJ00 - fy * tz2 * dL_dJ11 + (2 * fx * t[0]) * tz3 * dL_dJ02 +
                       (2 * fy * t[1]) * tz3 * dL_dJ12;
  /* transformVec4x3Transpose, auxiliary.h:90-98 */
  dL_dmean[0] = view[0] * dL_dtx + view[1] * dL_dty + view[2] * dL_dtz;
  dL_dmean[1] = view[4] * dL_dtx + view[5] * dL_dty + view[6] * dL_dtz;
  dL_dmean[2] = view[8] * dL_dtx + view[9] * dL_dty + view[10] * dL_dtz;
}

/* DGR backward.cu:20-139 */
static void sh_backward(const orc_scene* s, int idx, const uint8_t* clamped, const float* dL_dcolor,
                        float* dL_dmean /*+=*/, float* dL_dsh)
{
  const int deg = s->D, M = s->M;
  const float* pos = s->means3D + 3 * idx;
  const float dorig[3] = {pos[0] - s->campos[0], pos[1] - s->campos[1], pos[2] - s->campos[2]};
  const float len = sqrtf(dorig[0] * dorig[0] + dorig[1] * dorig[1] + dorig[2] * dorig[2]);
  const float x = dorig[0] / len, y = dorig[1] / len, z = dorig[2] / len;
  const float* sh = s->shs + (size_t)idx * M * 3;
  float* out = dL_dsh + (size_t)idx * M * 3;
  float dRGB[3], ddir[3] = {0, 0, 0};
  for (int c = 0; c < 3; c++) dRGB[c] = dL_dcolor[3 * idx + c] * (clamped[3 * idx + c] ? 0.f : 1.f);
  float w[16];
  int n = 1;
  w[0] = SH_C0;
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
  if (deg > 0) {
    w[1] = -SH_C1 * y; w[2] = SH_C1 * z; w[3] = -SH_C1 * x; n = 4;
    if (deg > 1) {
      w[4] = SH_C2[0] * xy; w[5] = SH_C2[1] * yz; w[6] = SH_C2[2] * (2.f * zz - xx - yy);
      w[7] = SH_C2[3] * xz; w[8] = SH_C2[4] * (xx - yy); n = 9;
      if (deg > 2) {
        w[9] = SH_C3[0] * y * (3.f * xx - yy); w[10] = SH_C3[1] * xy * z;
        w[11] = SH_C3[2] * y * (4.f * zz - xx - yy); w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
        w[13] = SH_C3[4] * x * (4.f * zz - xx - yy); w[14] = SH_C3[5] * z * (xx - yy);
        w[15] = SH_C3[6] * x * (xx - 3.f * yy); n = 16;
      }
    }
  }
  for (int k = 0; k < n; k++)
    for (int c = 0; c < 3; c++) out[3 * k + c] = w[k] * dRGB[c];
  for (int c = 0; c < 3; c++) {
#define SH(k) sh[3 * (k) + c]
    float gx = 0, gy = 0, gz = 0;
    if (deg > 0) {
      gx = -SH_C1 * SH(3); gy = -SH_C1 * SH(1); gz = SH_C1 * SH(2);
      if (deg > 1) {
        gx += SH_C2[0] * y * SH(4) + SH_C2[2] * 2.f * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * 2.f * x * SH(8);
        gy += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * 2.f * -y * SH(6) + SH_C2[4] * 2.f * -y * SH(8);
        gz += SH_C2[1] * y * SH(5) + SH_C2[2] * 2.f * 2.f * z * SH(6) + SH_C2[3] * x * SH(7);
        if (deg > 2) {
          gx += (SH_C3[0] * SH(9) * 3.f * 2.f * xy + SH_C3[1] * SH(10) * yz + SH_C3[2] * SH(11) * -2.f * xy +
                 SH_C3[3] * SH(12) * -3.f * 2.f * xz + SH_C3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) +
                 SH_C3[5] * SH(14) * 2.f * xz + SH_C3[6] * SH(15) * 3.f * (xx - yy));
          gy += (SH_C3[0] * SH(9) * 3.f * (xx - yy) + SH_C3[1] * SH(10) * xz +
                 SH_C3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12) * -3.f * 2.f * yz +
                 SH_C3[4] * SH(13) * -2.f * xy + SH_C3[5] * SH(14) * -2.f * yz + SH_C3[6] * SH(15) * -3.f * 2.f * xy);
          gz += (SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * 4.f * 2.f * yz +
                 SH_C3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13) * 4.f * 2.f * xz +
                 SH_C3[5] * SH(14) * (xx - yy));
        }
      }
    }
#undef SH
    ddir[0] += gx * dRGB[c]; ddir[1] += gy * dRGB[c]; ddir[2] += gz * dRGB[c];
  }
  /* dnormvdv, auxiliary.h:108-118 */
  const float vx = dorig[0], vy = dorig[1], vz = dorig[2];
  const float sum2 = vx * vx + vy * vy + vz * vz;
  const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
  dL_dmean[0] += ((+sum2 - vx * vx) * ddir[0] - vy * vx * ddir[1] - vz * vx * ddir[2]) * inv;
  dL_dmean[1] += (-vx * vy * ddir[0] + (sum2 - vy * vy) * ddir[1] - vz * vy * ddir[2]) * inv;
  dL_dmean[2] += (-vx * vz * ddir[0] - vy * vz * ddir[1] + (sum2 - vz * vz) * ddir[2]) * inv;
}

/* DGR backward.cu:278-341: dL_dcov3D -> dL_dscale (assigned), dL_drot (assigned) */
static void cov3d_backward(const float* scale, float mod, const float* q, const float* dc,
                           float* dL_dscale, float* dL_drot)
{
  m3 R, S = m3_zero(), M, dS, dM, twoM, Rt, dMt;
  build_R(q, &R);
  const float sv[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
  S.m[0][0] = sv[0]; S.m[1][1] = sv[1]; S.m[2][2] = sv[2];
  M = m3_mul(&S, &R);
  dS.m[0][0] = dc[0];        dS.m[0][1] = 0.5f * dc[1]; dS.m[0][2] = 0.5f * dc[2];
  dS.m[1][0] = 0.5f * dc[1]; dS.m[1][1] = dc[3];        dS.m[1][2] = 0.5f * dc[4];
  dS.m[2][0] = 0.5f * dc[2]; dS.m[2][1] = 0.5f * dc[4]; dS.m[2][2] = dc[5];
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) twoM.m[c][r] = 2.0f * M.m[c][r];
  dM = m3_mul(&twoM, &dS); /* dL_dM = 2.0f * M * dL_dSigma */
  Rt = m3_transpose(&R);
  dMt = m3_transpose(&dM);
  for (int k = 0; k < 3; k++) dL_dscale[k] = dot3v(Rt.m[k], dMt.m[k]);
  for (int k = 0; k < 3; k++)
    for (int r = 0; r < 3; r++) dMt.m[k][r] *= sv[k];
  quat_grad(q, &dMt, dL_drot);
}

/* DGR backward.cu:344-392: dL_dnorm3D -> dL_drot (+=).  The sign is resolved
 * with the SAVED forward normal (:373). */
static void norm3d_backward(const float* scale, const float* q, const float* norm3D,
                            const float* dn, float* dL_drot)
{
  m3 R, dRt;
  build_R(q, &R);
  float n[3] = {1.f, 0.f, 0.f}, Rn[3], dq[4];
  if (scale[0] > scale[2] && scale[1] > scale[2]) { n[0] = 0.f; n[2] = 1.f; }
  else if (scale[0] > scale[1] && scale[2] > scale[1]) { n[0] = 0.f; n[1] = 1.f; }
  m3_mulv(&R, n, Rn);
  if (dot3v(Rn, norm3D) < 0) { n[0] = -n[0]; n[1] = -n[1]; n[2] = -n[2]; }
  /* dL_dR[c][r] = dn[c]*n[r]; dL_dRt = transpose */
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) dRt.m[c][r] = dn[r] * n[c];
  quat_grad(q, &dRt, dq);
  for (int k = 0; k < 4; k++) dL_drot[k] += dq[k];
}

void orc_preprocess_backward(const orc_scene* s, const orc_geom* g, const float* dL_dmean2D,
                             const float* dL_dconic, const float* dL_dcolor, const float* dL_ddepth,
                             const float* dL_dnorm3D, float* dL_dmean3D, float* dL_dcov3D,
                             float* dL_dsh, float* dL_dscale, float* dL_drot)
{
  const float focal_y = s->H / (2.0f * s->tan_fovy); /* rasterizer_impl.cu:395-396 */
  const float focal_x = s->W / (2.0f * s->tan_fovx);
  const float* view = s->viewmatrix;
  const float* proj = s->projmatrix;
#pragma omp parallel for schedule(static)
  for (int idx = 0; idx < s->P; idx++) {
    float* dmean = dL_dmean3D + 3 * idx;
    /* the reference zero-fills every output (rasterize_points.cu:177-193) */
    for (int k = 0; k < 3; k++) dmean[k] = 0;
    for (int k = 0; k < 6; k++) dL_dcov3D[6 * idx + k] = 0;
    for (int k = 0; k < 3; k++) dL_dscale[3 * idx + k] = 0;
    for (int k = 0; k < 4; k++) dL_drot[4 * idx + k] = 0;
    if (s->shs) for (int k = 0; k < 3 * s->M; k++) dL_dsh[(size_t)idx * s->M * 3 + k] = 0;
    if (!(g->radii[idx] > 0)) continue;
    const float* cov3D = s->cov3Ds_precomp ? s->cov3Ds_precomp + 6 * idx : g->cov3D + 6 * idx;
    cov2d_backward(s, idx, cov3D, focal_x, focal_y, dL_dconic, dmean, dL_dcov3D + 6 * idx);

    /* K9, backward.cu:426-457 */
    const float* m = s->means3D + 3 * idx;
    float mh[4];
    xform4x4(m, proj, mh);
    const float m_w = 1.0f / (mh[3] + 0.0000001f);
    const float mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
    const float g2x = dL_dmean2D[3 * idx], g2y = dL_dmean2D[3 * idx + 1];
    dmean[0] += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
    dmean[1] += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
    dmean[2] += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
    const float mul3 = view[2] * m[0] + view[6] * m[1] + view[10] * m[2] + view[14];
    dmean[0] += (view[2] - view[3] * mul3) * dL_ddepth[idx];
    dmean[1] += (view[6] - view[7] * mul3) * dL_ddepth[idx];
    dmean[2] += (view[10] - view[11] * mul3) * dL_ddepth[idx];

    if (s->shs) sh_backward(s, idx, g->clamped, dL_dcolor, dmean, dL_dsh);
    if (s->scales)
      cov3d_backward(s->scales + 3 * idx, s->scale_modifier, s->rotations + 4 * idx,
                     dL_dcov3D + 6 * idx, dL_dscale + 3 * idx, dL_drot + 4 * idx);
    if (!s->norm3Ds_precomp) {
      const float* n3 = g->norm3D + 3 * idx;
      norm3d_backward(s->scales + 3 * idx, s->rotations + 4 * idx, n3, dL_dnorm3D + 3 * idx,
                      dL_drot + 4 * idx);
    }
  }
}
