// Kernel argument blocks and host-side launchers of libigs_raster.
#pragma once
#include "igs_common.cuh"

namespace igs {

// what the tile scan (igs_tile_scan.cuh) reads and writes
struct TileScanArgs {
  int num_tiles;
  int64_t capacity;
  CtrlHeader* header;
  int32_t* num_rendered_host; // device view of the caller's pinned slot (NULL: the host copies it)
  uint32_t* info;             // [4] in the img workspace: {num_rendered, ...}, stays with the frame's state
  const uint32_t* tile_count; // [T*CNT_STRIDE]
  uint32_t* tile_cursor;      // [T*CNT_STRIDE] <- segment starts (when the duplicate CTAs do not scan themselves)
  uint2* ranges;              // [T] [start, end) of each tile in the sorted list, (0,0) if empty
};

struct PreArgs {
  Prof prof;
  int P, D, M, ED, W, H, grid_x, grid_y, prefiltered;
  float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
  int64_t capacity;
  const float *viewmatrix, *projmatrix, *campos;
  const float *means3D, *shs, *colors_precomp, *opacities, *scales, *rotations;
  const float *cov3Ds_precomp, *norm3Ds_precomp, *extra_attrs;
  int32_t* radii;
  float4* splat;
  uint32_t* tiles_touched;
  uint8_t* clamped;
  uint32_t* tile_count;   // [T*CNT_STRIDE] instances per tile (atomics); zero at the start of a frame
};

struct BinArgs {
  Prof prof;
  TileScanArgs scan;
  int P, grid_x, grid_y, num_tiles;
  int64_t capacity;
  const int32_t* radii;
  const float4* splat;
  const uint32_t* tiles_touched;
  CtrlHeader* header;
  uint32_t* tile_count;       // [T*CNT_STRIDE] instances per tile, from the preprocess kernel
  uint32_t* tile_cursor;      // [T*CNT_STRIDE] slots handed out in each tile's segment: relative to the segment start
                              // when the duplicate CTAs scan the counts themselves, else absolute (tile_scan_kernel)
  const uint2* ranges;        // [T] [start, end) of each tile in the sorted list, (0,0) if empty
  uint4* tile_order;          // [T] {tile id, its range} longest list first: a render unit finds its list with ONE load
  unsigned long long* pairs;  // [capacity] (depth bits << 32 | gaussian id), grouped by tile, unordered inside
  unsigned long long* pairs2; // [capacity] scratch of the two-level sort of segments beyond TS_CAP
  uint32_t* point_list;       // [capacity] sorted
  uint64_t* keys;             // [capacity] sorted (tile << 32 | depth bits), or NULL: not wanted
};

struct RenderArgs {
  Prof prof;
  int W, H, ED, grid_x, grid_y;
  CtrlHeader* header; // overflow flag: the tile lists are incomplete, render nothing; unit ticket
  const uint2* ranges;
  const uint4* tile_order;
  const uint32_t* point_list;
  const float4* splat;
  const float* extra_attrs; // [P,ED], used for channels >= 1
  const float* bg;
  float *out_color, *out_depth, *out_norm, *out_alpha, *out_extra;
  uint32_t* n_contrib;
  float* norm_len;   // [H*W] |n| before normalisation, written when normalize != 0
  int normalize;
};

struct BwdRenderArgs {
  Prof prof;
  int W, H, ED, grid_x, grid_y;
  const uint2* ranges;
  const uint4* tile_order;
  const uint32_t* point_list;
  const float4* splat;
  const float* extra_attrs;
  const float* bg;
  const float* out_alpha;
  const uint32_t* n_contrib;
  const uint32_t* info;   // the frame's {num_rendered, overflowed}: lists and n_contrib of an overflowed frame are not there
  const float *dL_dcolor, *dL_ddepth, *dL_dnorm, *dL_dalpha, *dL_dextra;
  const float* out_norm;  // unit normals of the forward, read when normalize != 0
  const float* norm_len;  // [H*W]
  int normalize;
  uint32_t* ticket;   // unit counter, zeroed with the accumulators
  float* accum;       // [P][16]
  float* accum_extra; // [P][ED-1] or NULL
};

struct BwdPreArgs {
  Prof prof;
  int P, D, M, ED, W, H;
  float tan_fovx, tan_fovy, focal_x, focal_y, scale_modifier;
  const float *viewmatrix, *projmatrix, *campos;
  const float *means3D, *shs, *scales, *rotations, *cov3Ds_precomp, *norm3Ds_precomp;
  const int32_t* radii;
  const float4* splat;
  const uint8_t* clamped;
  float* accum;        // read, then zeroed again (the next frame finds it clean)
  float* accum_extra;
  uint32_t* ticket;    // unit counter of the render-backward kernel, reset here
  float *dL_dmeans2D, *dL_dcolors, *dL_dopacity, *dL_dmeans3D, *dL_dcov3D, *dL_dnorm3D, *dL_dsh;
  float *dL_dscales, *dL_drotations, *dL_dextra_attrs;
};

// debug hook (igs_debug_trace): when non-null, the render kernels record per unit
// {globaltimer at start, at end, smid | unit << 32} -- 3 words per unit, forward units first
cudaError_t set_render_trace(unsigned long long* device_buffer);

cudaError_t launch_preprocess(const PreArgs& a, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);
cudaError_t launch_binning(const BinArgs& a, cudaStream_t s);     // tile ranges, duplicate into tile segments, per-tile depth sort
void launch_point_offsets(int P, const uint32_t* tiles_touched, uint32_t* point_offsets, cudaStream_t s);
cudaError_t launch_render(const RenderArgs& a, cudaStream_t s);
cudaError_t launch_render_backward(const BwdRenderArgs& a, cudaStream_t s);
cudaError_t launch_preprocess_backward(const BwdPreArgs& a, cudaStream_t s);

} // namespace igs
