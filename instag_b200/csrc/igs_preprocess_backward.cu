// Per-Gaussian backward: ONE kernel for what the reference does in two
// (computeCov2DCUDA, backward.cu:144-274, and preprocessCUDA, backward.cu:397-470
// with its SH / cov3D / normal helpers, :20-139, :278-392), reading the packed
// 64-byte gradient accumulator row written by render_backward_kernel and
// writing EVERY element of every output (the reference pre-zeroes 12 tensors
// with torch::zeros, rasterize_points.cu:177-193; here invisible Gaussians get
// their zeros from this kernel, so no fill launches are needed).
//
// Bound: HBM.  Algorithmic bytes per Gaussian: reads 64 accum + 64 splat + 44
// inputs (+12S SH +1 clamp), writes 4*(3+3+1+3+6+3+3M+3+4+ED).
#include "igs_common.cuh"
#include "igs_kernels.cuh"
#include "igs_math.cuh"

namespace igs {

__device__ __constant__ float bSH_C0 = 0.28209479177387814f;
__device__ __constant__ float bSH_C1 = 0.4886025119029199f;
__device__ __constant__ float bSH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                           -1.0925484305920792f, 0.5462742152960396f};
__device__ __constant__ float bSH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                           0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                           -0.5900435899266435f};

// quaternion gradient from a 3x3 matrix gradient A[col][row]
// (shared tail of reference backward.cu:333-336 and :385-388)
__device__ __forceinline__ float4 quat_grad(float4 q, const float (&A)[3][3]) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  float4 d;
  d.x = 2 * z * (A[0][1] - A[1][0]) + 2 * y * (A[2][0] - A[0][2]) + 2 * x * (A[1][2] - A[2][1]);
  d.y = 2 * y * (A[1][0] + A[0][1]) + 2 * z * (A[2][0] + A[0][2]) + 2 * r * (A[1][2] - A[2][1]) - 4 * x * (A[2][2] + A[1][1]);
  d.z = 2 * x * (A[1][0] + A[0][1]) + 2 * r * (A[2][0] - A[0][2]) + 2 * z * (A[1][2] + A[2][1]) - 4 * y * (A[2][2] + A[0][0]);
  d.w = 2 * r * (A[0][1] - A[1][0]) + 2 * x * (A[2][0] + A[0][2]) + 2 * y * (A[1][2] + A[2][1]) - 4 * z * (A[1][1] + A[0][0]);
  return d;
}

// SH backward (reference backward.cu:20-139): writes dL_dsh[0..M), returns dL/ddir
__device__ __forceinline__ float3 sh_backward(int deg, int M, const float* __restrict__ sh, float* __restrict__ dsh,
                                              float x, float y, float z, float3 g) {
  float w[16];
  int n = 1;
  w[0] = bSH_C0;
  const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
  if (deg > 0) {
    w[1] = -bSH_C1 * y; w[2] = bSH_C1 * z; w[3] = -bSH_C1 * x; n = 4;
    if (deg > 1) {
      w[4] = bSH_C2[0] * xy; w[5] = bSH_C2[1] * yz; w[6] = bSH_C2[2] * (2.f * zz - xx - yy);
      w[7] = bSH_C2[3] * xz; w[8] = bSH_C2[4] * (xx - yy); n = 9;
      if (deg > 2) {
        w[9] = bSH_C3[0] * y * (3.f * xx - yy); w[10] = bSH_C3[1] * xy * z;
        w[11] = bSH_C3[2] * y * (4.f * zz - xx - yy); w[12] = bSH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
        w[13] = bSH_C3[4] * x * (4.f * zz - xx - yy); w[14] = bSH_C3[5] * z * (xx - yy);
        w[15] = bSH_C3[6] * x * (xx - 3.f * yy); n = 16;
      }
    }
  }
  const float gc[3] = {g.x, g.y, g.z};
  if ((M & 3) == 0) {
    // the row dL_dsh[idx] is 16-byte aligned: element e = 3k+c is w[k]*g[c] (0 beyond the active
    // degree), written as 3M/4 float4 stores instead of 3M scalar ones at a 12*M-byte lane stride
    float4* o = reinterpret_cast<float4*>(dsh);
    const int nv = 3 * M / 4;
#pragma unroll
    for (int i = 0; i < 12; i++)
      if (i < nv) {
        float e[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int k = (4 * i + j) / 3, c = (4 * i + j) % 3;
          e[j] = k < n ? w[k] * gc[c] : 0.f;
        }
        o[i] = make_float4(e[0], e[1], e[2], e[3]);
      }
    for (int i = 12; i < nv; i++) o[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  } else {
#pragma unroll
    for (int k = 0; k < 16; k++)
      if (k < M) {
        const float wk = k < n ? w[k] : 0.f;
        dsh[3 * k] = wk * gc[0]; dsh[3 * k + 1] = wk * gc[1]; dsh[3 * k + 2] = wk * gc[2];
      }
    for (int k = 16; k < M; k++) { dsh[3 * k] = 0.f; dsh[3 * k + 1] = 0.f; dsh[3 * k + 2] = 0.f; }
  }
  float3 dd = make_float3(0.f, 0.f, 0.f);
  if (deg > 0) {
    float v[48];
    const int need = 3 * n;
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
#pragma unroll
    for (int c = 0; c < 3; c++) {
#define SHK(k) v[3 * (k) + c]
      float gx = -bSH_C1 * SHK(3), gy = -bSH_C1 * SHK(1), gz = bSH_C1 * SHK(2);
      if (deg > 1) {
        gx += bSH_C2[0] * y * SHK(4) + bSH_C2[2] * 2.f * -x * SHK(6) + bSH_C2[3] * z * SHK(7) + bSH_C2[4] * 2.f * x * SHK(8);
        gy += bSH_C2[0] * x * SHK(4) + bSH_C2[1] * z * SHK(5) + bSH_C2[2] * 2.f * -y * SHK(6) + bSH_C2[4] * 2.f * -y * SHK(8);
        gz += bSH_C2[1] * y * SHK(5) + bSH_C2[2] * 2.f * 2.f * z * SHK(6) + bSH_C2[3] * x * SHK(7);
        if (deg > 2) {
          gx += (bSH_C3[0] * SHK(9) * 3.f * 2.f * xy + bSH_C3[1] * SHK(10) * yz + bSH_C3[2] * SHK(11) * -2.f * xy +
                 bSH_C3[3] * SHK(12) * -3.f * 2.f * xz + bSH_C3[4] * SHK(13) * (-3.f * xx + 4.f * zz - yy) +
                 bSH_C3[5] * SHK(14) * 2.f * xz + bSH_C3[6] * SHK(15) * 3.f * (xx - yy));
          gy += (bSH_C3[0] * SHK(9) * 3.f * (xx - yy) + bSH_C3[1] * SHK(10) * xz +
                 bSH_C3[2] * SHK(11) * (-3.f * yy + 4.f * zz - xx) + bSH_C3[3] * SHK(12) * -3.f * 2.f * yz +
                 bSH_C3[4] * SHK(13) * -2.f * xy + bSH_C3[5] * SHK(14) * -2.f * yz + bSH_C3[6] * SHK(15) * -3.f * 2.f * xy);
          gz += (bSH_C3[1] * SHK(10) * xy + bSH_C3[2] * SHK(11) * 4.f * 2.f * yz +
                 bSH_C3[3] * SHK(12) * 3.f * (2.f * zz - xx - yy) + bSH_C3[4] * SHK(13) * 4.f * 2.f * xz +
                 bSH_C3[5] * SHK(14) * (xx - yy));
        }
      }
#undef SHK
      dd.x += gx * gc[c]; dd.y += gy * gc[c]; dd.z += gz * gc[c];
    }
  }
  return dd;
}

// Resident CTAs per SM the register allocation aims at.  3 (80 registers, a few spills): the kernel
// waits on its loads two thirds of the time, and 24 warps hide more of that than 16 (C5: 172.5 -> 164.3 us,
// C2 10.6 -> 8.5); 4 CTAs (64 registers) spill too much (184.3).
#ifndef IGS_PBW_MIN_CTAS
#define IGS_PBW_MIN_CTAS 3
#endif
template <bool USE_SH>
__global__ void __launch_bounds__(256, IGS_PBW_MIN_CTAS) preprocess_backward_kernel(BwdPreArgs a) {
  __shared__ float s_view[16], s_proj[16], s_cam[3];
  const int tid = threadIdx.x;
  if (a.prof.early) pdl_launch_dependents();
  if (tid < 16) { s_view[tid] = __ldg(a.viewmatrix + tid); s_proj[tid] = __ldg(a.projmatrix + tid); }
  if (tid < 3) s_cam[tid] = __ldg(a.campos + tid);
  pdl_wait(); // the accumulator rows of the render-backward kernel
  __syncthreads();
  const int idx = blockIdx.x * 256 + tid;
  if (idx >= a.P) return;
  const float* view = s_view;
  const float* proj = s_proj;

  if (idx == 0) *a.ticket = 0u; // the render-backward kernel is done with its unit counter
  // the accumulator row is read once and put back to zero: the next backward needs no memset
  float4* acc = reinterpret_cast<float4*>(a.accum) + (size_t)idx * (ACC_F / 4);
  float4 g0 = acc[0], g1 = acc[1];
  const float4 g2 = acc[2], g3 = acc[3];
  {
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    acc[0] = z; acc[1] = z; acc[2] = z; acc[3] = z;
  }
  const bool visible = __ldg(a.radii + idx) > 0;
  // The compositing backward leaves raw moments of g = G*dL/dalpha about the splat centre
  // (igs_common.cuh ACC_F); with dL/dG = opacity * dL/dalpha (reference backward.cu:690-709):
  //   dL/dmean2D.x = -0.5 W o (a Sx + b Sy)   dL/dmean2D.y = -0.5 H o (c Sy + b Sx)
  //   dL/dconic.{x,y,w} = -0.5 o {Sxx, Sxy, Syy}      dL/dopacity = S
  if (visible) {
    const float4 co = __ldg(a.splat + (size_t)idx * SPLAT_F4); // conic.xyz, opacity
    const float Sx = g0.x, Sy = g0.y, Sxx = g0.z, Sxy = g0.w, Syy = g1.x;
    g0.x = -0.5f * a.W * co.w * (co.x * Sx + co.y * Sy);
    g0.y = -0.5f * a.H * co.w * (co.z * Sy + co.y * Sx);
    g0.z = -0.5f * co.w * Sxx;
    g0.w = -0.5f * co.w * Sxy;
    g1.x = -0.5f * co.w * Syy;
  } else {
    g0 = make_float4(0.f, 0.f, 0.f, 0.f);
    g1.x = 0.f;
  }
  // g0 = (m2d.x, m2d.y, conic.x, conic.y) g1 = (conic.w, opacity, col.r, col.g)
  // g2 = (col.b, depth, n.x, n.y)         g3 = (n.z, extra0, -, -)
  a.dL_dmeans2D[3 * idx] = g0.x; a.dL_dmeans2D[3 * idx + 1] = g0.y; a.dL_dmeans2D[3 * idx + 2] = 0.f;
  a.dL_dcolors[3 * idx] = g1.z; a.dL_dcolors[3 * idx + 1] = g1.w; a.dL_dcolors[3 * idx + 2] = g2.x;
  a.dL_dopacity[idx] = g1.y;
  a.dL_dnorm3D[3 * idx] = g2.z; a.dL_dnorm3D[3 * idx + 1] = g2.w; a.dL_dnorm3D[3 * idx + 2] = g3.x;
  if (a.ED >= 1) {
    a.dL_dextra_attrs[(size_t)idx * a.ED] = g3.y;
    for (int c = 1; c < a.ED; c++) {
      float* e = a.accum_extra + (size_t)idx * (a.ED - 1) + c - 1;
      a.dL_dextra_attrs[(size_t)idx * a.ED + c] = *e;
      *e = 0.f;
    }
  }

  float dmean[3] = {0.f, 0.f, 0.f}, dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float dscale[3] = {0.f, 0.f, 0.f};
  float4 drot = make_float4(0.f, 0.f, 0.f, 0.f);
  float* dsh = USE_SH ? a.dL_dsh + (size_t)idx * a.M * 3 : nullptr;

  if (visible) {
    const float mx = __ldg(a.means3D + 3 * idx), my = __ldg(a.means3D + 3 * idx + 1), mz = __ldg(a.means3D + 3 * idx + 2);
    const float3 sc = make_float3(__ldg(a.scales + 3 * idx), __ldg(a.scales + 3 * idx + 1), __ldg(a.scales + 3 * idx + 2));
    const float4 q = __ldg(reinterpret_cast<const float4*>(a.rotations) + idx);
    const Rot R = quat_to_rot(q);
    float c3[6];
    if (a.cov3Ds_precomp) {
#pragma unroll
      for (int k = 0; k < 6; k++) c3[k] = __ldg(a.cov3Ds_precomp + 6 * idx + k);
    } else {
      cov3d_from_scale_rot(sc, a.scale_modifier, R, c3);
    }
    // ---- conic -> cov2D -> cov3D, mean (reference backward.cu:144-274) ----
    const float tx0 = xform_row(view, 0, mx, my, mz), ty0 = xform_row(view, 1, mx, my, mz), tz0 = xform_row(view, 2, mx, my, mz);
    float T0[3], T1[3], t[3], ratio[2];
    const float3 cov = cov2d(tx0, ty0, tz0, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c3, view, T0, T1, t, ratio);
    const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
    const float x_grad_mul = (ratio[0] < -limx || ratio[0] > limx) ? 0.f : 1.f;
    const float y_grad_mul = (ratio[1] < -limy || ratio[1] > limy) ? 0.f : 1.f;
    const float ca = cov.x, cb = cov.y, cc = cov.z;
    const float dcx = g0.z, dcy = g0.w, dcz = g1.x; // dL_dconic {x, y, w}
    const float denom = ca * cc - cb * cb;
    float dL_da = 0, dL_db = 0, dL_dc = 0;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    if (denom2inv != 0) {
      dL_da = denom2inv * (-cc * cc * dcx + 2 * cb * cc * dcy + (denom - ca * cc) * dcz);
      dL_dc = denom2inv * (-ca * ca * dcz + 2 * ca * cb * dcy + (denom - ca * cc) * dcx);
      dL_db = denom2inv * 2 * (cb * cc * dcx - (denom + 2 * cb * cb) * dcy + ca * cb * dcz);
      dcov[0] = (T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc);
      dcov[3] = (T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc);
      dcov[5] = (T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc);
      dcov[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
      dcov[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
      dcov[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
    }
    const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
    float dT0[3], dT1[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float r0 = T0[0] * V[k][0] + T0[1] * V[k][1] + T0[2] * V[k][2];
      const float r1 = T1[0] * V[k][0] + T1[1] * V[k][1] + T1[2] * V[k][2];
      dT0[k] = 2 * r0 * dL_da + r1 * dL_db;
      dT1[k] = 2 * r1 * dL_dc + r0 * dL_db;
    }
    const float dJ00 = view[0] * dT0[0] + view[4] * dT0[1] + view[8] * dT0[2];
    const float dJ02 = view[2] * dT0[0] + view[6] * dT0[1] + view[10] * dT0[2];
    const float dJ11 = view[1] * dT1[0] + view[5] * dT1[1] + view[9] * dT1[2];
    const float dJ12 = view[2] * dT1[0] + view[6] * dT1[1] + view[10] * dT1[2];
    const float tz = 1.f / t[2], tz2 = tz * tz, tz3 = tz2 * tz;
    const float hx = a.focal_x, hy = a.focal_y;
    const float dtx = x_grad_mul * -hx * tz2 * dJ02;
    const float dty = y_grad_mul * -hy * tz2 * dJ12;
    const float dtz = -hx * tz2 * dJ00 - hy * tz2 * dJ11 + (2 * hx * t[0]) * tz3 * dJ02 + (2 * hy * t[1]) * tz3 * dJ12;
    dmean[0] = view[0] * dtx + view[1] * dty + view[2] * dtz;
    dmean[1] = view[4] * dtx + view[5] * dty + view[6] * dtz;
    dmean[2] = view[8] * dtx + view[9] * dty + view[10] * dtz;

    // ---- mean2D and depth -> mean (reference backward.cu:426-457) ----
    const float mhw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (mhw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    dmean[0] += (proj[0] * m_w - proj[3] * mul1) * g0.x + (proj[1] * m_w - proj[3] * mul2) * g0.y;
    dmean[1] += (proj[4] * m_w - proj[7] * mul1) * g0.x + (proj[5] * m_w - proj[7] * mul2) * g0.y;
    dmean[2] += (proj[8] * m_w - proj[11] * mul1) * g0.x + (proj[9] * m_w - proj[11] * mul2) * g0.y;
    const float mul3 = view[2] * mx + view[6] * my + view[10] * mz + view[14];
    const float ddep = g2.y;
    dmean[0] += (view[2] - view[3] * mul3) * ddep;
    dmean[1] += (view[6] - view[7] * mul3) * ddep;
    dmean[2] += (view[10] - view[11] * mul3) * ddep;

    // ---- SH (reference backward.cu:20-139) ----
    if (USE_SH) {
      const float ox = mx - s_cam[0], oy = my - s_cam[1], oz = mz - s_cam[2];
      const float len = sqrtf(ox * ox + oy * oy + oz * oz);
      const uint32_t cl = a.clamped[idx];
      const float3 g = make_float3(g1.z * ((cl & 1) ? 0.f : 1.f), g1.w * ((cl & 2) ? 0.f : 1.f), g2.x * ((cl & 4) ? 0.f : 1.f));
      const float3 dd = sh_backward(a.D, a.M, a.shs + (size_t)idx * a.M * 3, dsh, ox / len, oy / len, oz / len, g);
      const float sum2 = ox * ox + oy * oy + oz * oz;
      const float inv = 1.0f / sqrtf(sum2 * sum2 * sum2);
      dmean[0] += ((+sum2 - ox * ox) * dd.x - oy * ox * dd.y - oz * ox * dd.z) * inv;
      dmean[1] += (-ox * oy * dd.x + (sum2 - oy * oy) * dd.y - oz * oy * dd.z) * inv;
      dmean[2] += (-ox * oz * dd.x - oy * oz * dd.y + (sum2 - oz * oz) * dd.z) * inv;
    }

    // ---- cov3D -> scale, rotation (reference backward.cu:278-341) ----
    {
      const float sv[3] = {a.scale_modifier * sc.x, a.scale_modifier * sc.y, a.scale_modifier * sc.z};
      float A2[3][3]; // 2*M, M[c][r] = s_r R[c][r]
#pragma unroll
      for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) A2[c][r] = 2.0f * (sv[r] * R.m[c][r]);
      const float D[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                             {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                             {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
      float dM[3][3];
#pragma unroll
      for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) dM[c][r] = A2[0][r] * D[c][0] + A2[1][r] * D[c][1] + A2[2][r] * D[c][2];
      float dMt[3][3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        dscale[k] = R.m[0][k] * dM[0][k] + R.m[1][k] * dM[1][k] + R.m[2][k] * dM[2][k];
#pragma unroll
        for (int r = 0; r < 3; r++) dMt[k][r] = dM[r][k] * sv[k];
      }
      drot = quat_grad(q, dMt);
    }
    // ---- normal -> rotation (reference backward.cu:344-392) ----
    if (!a.norm3Ds_precomp) {
      int k = 0;
      if (sc.x > sc.z && sc.y > sc.z) k = 2;
      else if (sc.x > sc.y && sc.z > sc.y) k = 1;
      const float4 r2 = __ldg(a.splat + (size_t)idx * SPLAT_F4 + 2), r3 = __ldg(a.splat + (size_t)idx * SPLAT_F4 + 3);
      // sign resolved against the SAVED forward normal, with R*axis (not R^T*axis) as the reference does
      const float dotv = R.m[k][0] * r2.w + R.m[k][1] * r3.x + R.m[k][2] * r3.y;
      const float sgn = dotv < 0 ? -1.f : 1.f;
      float dRt[3][3];
#pragma unroll
      for (int c = 0; c < 3; c++) {
        dRt[c][0] = c == k ? g2.z * sgn : 0.f;
        dRt[c][1] = c == k ? g2.w * sgn : 0.f;
        dRt[c][2] = c == k ? g3.x * sgn : 0.f;
      }
      const float4 dq = quat_grad(q, dRt);
      drot.x += dq.x; drot.y += dq.y; drot.z += dq.z; drot.w += dq.w;
    }
  } else if (USE_SH) {
    for (int k = 0; k < a.M * 3; k++) dsh[k] = 0.f;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) { a.dL_dmeans3D[3 * idx + k] = dmean[k]; a.dL_dscales[3 * idx + k] = dscale[k]; }
#pragma unroll
  for (int k = 0; k < 6; k++) a.dL_dcov3D[6 * idx + k] = dcov[k];
  reinterpret_cast<float4*>(a.dL_drotations)[idx] = drot;
  if (tid == 0) stamp_done(a.prof, ST_PBWD);
}

cudaError_t launch_preprocess_backward(const BwdPreArgs& a, cudaStream_t s) {
  if (a.P <= 0) return cudaSuccess;
  const int blocks = (a.P + 255) / 256;
  count_launch();
  if (a.shs) return launch_after(preprocess_backward_kernel<true>, dim3(blocks), dim3(256), 0, s, a);
  return launch_after(preprocess_backward_kernel<false>, dim3(blocks), dim3(256), 0, s, a);
}

} // namespace igs
