// fp32 geometry shared by the forward and backward per-Gaussian kernels.
// Every fused multiply-add is spelled out (see DESIGN.md "FMA contraction"):
// radii / tiles_touched / depth keys hang off these values bit for bit.
#pragma once
#include "igs_common.cuh"

namespace igs {

// rotation matrix entries from the RAW (un-normalised) quaternion (r,x,y,z);
// R[c][r] in glm column-major terms (reference forward.cu:127-138)
struct Rot { float m[3][3]; };
__device__ __forceinline__ Rot quat_to_rot(float4 q) {
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  Rot R;
  R.m[0][0] = __fsub_rn(1.f, __fmul_rn(2.f, __fadd_rn(__fmul_rn(y, y), __fmul_rn(z, z)))); // yy, zz shared: not fused
  R.m[0][1] = 2.f * __fmaf_rn(x, y, -__fmul_rn(r, z));
  R.m[0][2] = 2.f * __fmaf_rn(r, y, __fmul_rn(x, z));  // x*z is the shared rounded product
  R.m[1][0] = 2.f * __fmaf_rn(x, y, __fmul_rn(r, z));
  R.m[1][1] = __fmaf_rn(-2.f, __fmaf_rn(x, x, __fmul_rn(z, z)), 1.f);
  R.m[1][2] = 2.f * __fmaf_rn(y, z, -__fmul_rn(r, x));
  R.m[2][0] = 2.f * __fmaf_rn(-r, y, __fmul_rn(x, z));
  R.m[2][1] = 2.f * __fmaf_rn(y, z, __fmul_rn(r, x));
  R.m[2][2] = __fmaf_rn(-2.f, __fmaf_rn(x, x, __fmul_rn(y, y)), 1.f);
  return R;
}

// Sigma = (S R)^T (S R), upper triangle (reference forward.cu:118-152)
__device__ __forceinline__ void cov3d_from_scale_rot(float3 s, float mod, const Rot& R, float* c) {
  const float sv[3] = {__fmul_rn(mod, s.x), __fmul_rn(mod, s.y), __fmul_rn(mod, s.z)};
  float M[3][3];
#pragma unroll
  for (int cc = 0; cc < 3; cc++)
#pragma unroll
    for (int r = 0; r < 3; r++) M[cc][r] = __fmul_rn(sv[r], R.m[cc][r]);
#define SIG(cc, r) dot3c(M[r][0], M[cc][0], M[r][1], M[cc][1], M[r][2], M[cc][2])
  c[0] = SIG(0, 0); c[1] = SIG(0, 1); c[2] = SIG(0, 2);
  c[3] = SIG(1, 1); c[4] = SIG(1, 2); c[5] = SIG(2, 2);
#undef SIG
}

// EWA 2D covariance + 0.3 px low-pass (reference forward.cu:74-113)
__device__ __forceinline__ float3 cov2d(float tx, float ty, float tz, float fx, float fy, float tan_fovx,
                                        float tan_fovy, const float* c3, const float* __restrict__ v,
                                        float* T0_out = nullptr, float* T1_out = nullptr, float* t_out = nullptr,
                                        float* ratio_out = nullptr) {
  const float limx = __fmul_rn(1.3f, tan_fovx), limy = __fmul_rn(1.3f, tan_fovy);
  const float txtz = __fdiv_rn(tx, tz), tytz = __fdiv_rn(ty, tz);
  tx = __fmul_rn(fminf(limx, fmaxf(-limx, txtz)), tz);
  ty = __fmul_rn(fminf(limy, fmaxf(-limy, tytz)), tz);
  const float tz2 = __fmul_rn(tz, tz);
  const float J00 = __fdiv_rn(fx, tz), J02 = __fdiv_rn(-__fmul_rn(fx, tx), tz2);
  const float J11 = __fdiv_rn(fy, tz), J12 = __fdiv_rn(-__fmul_rn(fy, ty), tz2);
  // T = W*J ; W[0]=(v0,v4,v8) W[1]=(v1,v5,v9) W[2]=(v2,v6,v10) ; T[2][*] = 0
  float T0[3], T1[3];
#pragma unroll
  for (int r = 0; r < 3; r++) {
    T0[r] = __fmaf_rn(v[2 + 4 * r], J02, __fmul_rn(v[0 + 4 * r], J00));
    T1[r] = __fmaf_rn(v[2 + 4 * r], J12, __fmul_rn(v[1 + 4 * r], J11));
  }
  const float V[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
  // X = T^T * Vrk^T ; X[c][r] = sum_k T[r][k] V[c][k]
  float X0[3], X1[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    X0[c] = dot3c(T0[0], V[c][0], T0[1], V[c][1], T0[2], V[c][2]);
    X1[c] = dot3c(T1[0], V[c][0], T1[1], V[c][1], T1[2], V[c][2]);
  }
  // cov = X * T ; cov[c][r] = sum_k X[k][r] T[c][k]
  const float c00 = dot3c(X0[0], T0[0], X0[1], T0[1], X0[2], T0[2]);
  const float c01 = dot3c(X1[0], T0[0], X1[1], T0[1], X1[2], T0[2]);
  const float c11 = dot3c(X1[0], T1[0], X1[1], T1[1], X1[2], T1[2]);
  if (T0_out) {
#pragma unroll
    for (int r = 0; r < 3; r++) { T0_out[r] = T0[r]; T1_out[r] = T1[r]; }
    t_out[0] = tx; t_out[1] = ty; t_out[2] = tz;
    ratio_out[0] = txtz; ratio_out[1] = tytz;
  }
  return make_float3(__fadd_rn(c00, 0.3f), c01, __fadd_rn(c11, 0.3f));
}

} // namespace igs
