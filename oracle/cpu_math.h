/* TEST INFRASTRUCTURE -- see cpu_raster.h.  fp32 helpers that restate the
 * glm / auxiliary.h arithmetic of the reference with nvcc's FMA contraction
 * spelled out (this translation unit is compiled with -ffp-contract=off). */
#ifndef ORACLE_CPU_MATH_H
#define ORACLE_CPU_MATH_H
#include <math.h>
#include <stdint.h>
#include <string.h>

/* DGR cuda_rasterizer/auxiliary.h:23-40 */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

/* a0*b0 + a1*b1 + a2*b2 as nvcc contracts it: the left product of the first
 * addition is fused, the right one stays a rounded multiply. */
static inline float d3(float a0, float b0, float a1, float b1, float a2, float b2)
{
  return fmaf(a2, b2, fmaf(a0, b0, a1 * b1));
}

/* glm::dot(vec3,vec3): tmp = a*b; tmp.x + tmp.y + tmp.z */
static inline float dot3v(const float* a, const float* b)
{
  return d3(a[0], b[0], a[1], b[1], a[2], b[2]);
}

/* glm::mat3, column major: m[c][r] */
typedef struct { float m[3][3]; } m3;

static inline m3 m3_zero(void) { m3 z; memset(&z, 0, sizeof z); return z; }

/* glm operator*(mat3,mat3), third_party/glm/glm/detail/type_mat3x3.inl:486 */
static inline m3 m3_mul(const m3* A, const m3* B)
{
  m3 R;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++)
      R.m[c][r] = d3(A->m[0][r], B->m[c][0], A->m[1][r], B->m[c][1], A->m[2][r], B->m[c][2]);
  return R;
}

static inline m3 m3_transpose(const m3* A)
{
  m3 R;
  for (int c = 0; c < 3; c++)
    for (int r = 0; r < 3; r++) R.m[c][r] = A->m[r][c];
  return R;
}

/* glm operator*(mat3,vec3), type_mat3x3.inl:468 */
static inline void m3_mulv(const m3* A, const float* v, float* out)
{
  for (int r = 0; r < 3; r++)
    out[r] = d3(A->m[0][r], v[0], A->m[1][r], v[1], A->m[2][r], v[2]);
}

/* DGR auxiliary.h:59-78 transformPoint4x3 / 4x4: m0*x + m4*y + m8*z + m12 */
static inline void xform4x3(const float* p, const float* m, float* o)
{
  for (int r = 0; r < 3; r++) o[r] = d3(m[r], p[0], m[4 + r], p[1], m[8 + r], p[2]) + m[12 + r];
}
static inline void xform4x4(const float* p, const float* m, float* o)
{
  for (int r = 0; r < 4; r++) o[r] = d3(m[r], p[0], m[4 + r], p[1], m[8 + r], p[2]) + m[12 + r];
}

/* CUDA float->int conversion (cvt.rzi.s32.f32): truncates, saturates, NaN->0 */
static inline int f2i_rz(float v)
{
  if (v != v) return 0;
  if (v >= 2147483648.0f) return 2147483647;
  if (v <= -2147483648.0f) return (int)(-2147483647 - 1);
  return (int)v;
}

/* DGR auxiliary.h:42-45 ndc2Pix: evaluated in DOUBLE (literals 1.0/0.5), the
 * multiply-subtract contracted to a double FMA, then rounded to float. */
static inline float ndc2pix(float v, int S)
{
  return (float)(fma((double)v + 1.0, (double)S, -1.0) * 0.5);
}

/* DGR auxiliary.h:47-57 getRect (max_radius is an int parameter). */
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }
static inline void get_rect(float px, float py, int max_radius, int gx, int gy, int* rmin, int* rmax)
{
  const float r = (float)max_radius;
  rmin[0] = clampi(f2i_rz((px - r) / (float)16), 0, gx);
  rmin[1] = clampi(f2i_rz((py - r) / (float)16), 0, gy);
  rmax[0] = clampi(f2i_rz((((px + r) + 16.f) - 1.f) / (float)16), 0, gx);
  rmax[1] = clampi(f2i_rz((((py + r) + 16.f) - 1.f) / (float)16), 0, gy);
}

static inline uint32_t f2bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

#endif
