// TEST/BASELINE INFRASTRUCTURE -- not part of the product path.
//
// C-ABI shim around the UNMODIFIED reference rasterizer
// (/root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/*),
// compiled by oracle/Makefile from the sources where they lie into
// oracle/_ref/libdgr_ref.so.  Nothing of the reference is copied here: this
// file only calls CudaRasterizer::Rasterizer::{forward,backward,markVisible}
// (cuda_rasterizer/rasterizer.h:20-102) and re-uses the reference's own
// State::fromChunk (cuda_rasterizer/rasterizer_impl.h:22-72) to report where
// the intermediates live inside the scratch buffers, so that the parity tests
// can compare tiles_touched / point_offsets / keys / point_list / ranges /
// n_contrib bit for bit.
//
// Used by: tests/ (GPU parity oracle) and bench.py --impl reference.
#include <cstdint>
#include <cstddef>
#include <functional>
#include <cuda_runtime.h>
#include "rasterizer.h"
#include "rasterizer_impl.h"

extern "C" {

// which: 0 = geometry, 1 = binning, 2 = image.  Must return device memory of
// at least `bytes` bytes that stays alive until the matching backward.
typedef char* (*dgr_ref_alloc_fn)(size_t bytes, int which);

int dgr_ref_forward(
    dgr_ref_alloc_fn alloc,
    int P, int D, int M, int ED,
    const float* background, int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* opacities, const float* scales, float scale_modifier,
    const float* rotations, const float* cov3Ds_precomp,
    const float* norm3Ds_precomp, const float* extra_attrs,
    const float* viewmatrix, const float* projmatrix, const float* cam_pos,
    float tan_fovx, float tan_fovy, int prefiltered,
    float* out_color, float* out_depth, float* out_norm, float* out_alpha,
    float* out_extra, int* radii, int debug)
{
  std::function<char*(size_t)> g = [alloc](size_t n) { return alloc(n, 0); };
  std::function<char*(size_t)> b = [alloc](size_t n) { return alloc(n, 1); };
  std::function<char*(size_t)> i = [alloc](size_t n) { return alloc(n, 2); };
  try {
    return CudaRasterizer::Rasterizer::forward(
        g, b, i, P, D, M, ED, background, width, height, means3D, shs,
        colors_precomp, opacities, scales, scale_modifier, rotations,
        cov3Ds_precomp, norm3Ds_precomp, extra_attrs, viewmatrix, projmatrix,
        cam_pos, tan_fovx, tan_fovy, prefiltered != 0, out_color, out_depth,
        out_norm, out_alpha, out_extra, radii, debug != 0);
  } catch (...) {
    return -1;
  }
}

int dgr_ref_backward(
    int P, int D, int M, int R, int ED,
    const float* background, int width, int height,
    const float* means3D, const float* shs, const float* colors_precomp,
    const float* scales, float scale_modifier, const float* rotations,
    const float* cov3Ds_precomp, const float* norm3Ds_precomp,
    const float* extra_attrs, const float* viewmatrix,
    const float* projmatrix, const float* campos,
    float tan_fovx, float tan_fovy, const int* radii,
    char* geom_buffer, char* binning_buffer, char* image_buffer,
    const float* accum_alphas,
    const float* dL_dpix, const float* dL_dpix_depth, const float* dL_dpix_norm,
    const float* dL_dpix_alpha, const float* dL_dpix_extra,
    float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
    float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D, float* dL_dnorm3D,
    float* dL_dsh, float* dL_dscale, float* dL_drot, float* dL_dextra,
    int debug)
{
  try {
    CudaRasterizer::Rasterizer::backward(
        P, D, M, R, ED, background, width, height, means3D, shs,
        colors_precomp, scales, scale_modifier, rotations, cov3Ds_precomp,
        norm3Ds_precomp, extra_attrs, viewmatrix, projmatrix, campos, tan_fovx,
        tan_fovy, radii, geom_buffer, binning_buffer, image_buffer,
        accum_alphas, dL_dpix, dL_dpix_depth, dL_dpix_norm, dL_dpix_alpha,
        dL_dpix_extra, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
        dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dnorm3D, dL_dsh, dL_dscale,
        dL_drot, dL_dextra, debug != 0);
  } catch (...) {
    return -1;
  }
  return 0;
}

void dgr_ref_mark_visible(int P, float* means3D, float* viewmatrix,
                          float* projmatrix, bool* present)
{
  CudaRasterizer::Rasterizer::markVisible(P, means3D, viewmatrix, projmatrix,
                                          present);
}

// Byte offsets (relative to `base`) of the arrays inside the three scratch
// buffers, as laid out by the reference's own fromChunk.
// geom: depths, clamped, means2D, cov3D, conic_opacity, rgb, norm3D,
//       tiles_touched, point_offsets                              (9 entries)
void dgr_ref_geom_layout(char* base, size_t P, int64_t* off)
{
  char* c = base;
  auto s = CudaRasterizer::GeometryState::fromChunk(c, P);
  off[0] = (char*)s.depths - base;
  off[1] = (char*)s.clamped - base;
  off[2] = (char*)s.means2D - base;
  off[3] = (char*)s.cov3D - base;
  off[4] = (char*)s.conic_opacity - base;
  off[5] = (char*)s.rgb - base;
  off[6] = (char*)s.norm3D - base;
  off[7] = (char*)s.tiles_touched - base;
  off[8] = (char*)s.point_offsets - base;
}

// binning: point_list, point_list_unsorted, keys, keys_unsorted   (4 entries)
void dgr_ref_binning_layout(char* base, size_t R, int64_t* off)
{
  char* c = base;
  auto s = CudaRasterizer::BinningState::fromChunk(c, R);
  off[0] = (char*)s.point_list - base;
  off[1] = (char*)s.point_list_unsorted - base;
  off[2] = (char*)s.point_list_keys - base;
  off[3] = (char*)s.point_list_keys_unsorted - base;
}

// image: n_contrib, ranges                                        (2 entries)
void dgr_ref_image_layout(char* base, size_t HW, int64_t* off)
{
  char* c = base;
  auto s = CudaRasterizer::ImageState::fromChunk(c, HW);
  off[0] = (char*)s.n_contrib - base;
  off[1] = (char*)s.ranges - base;
}

}  // extern "C"
