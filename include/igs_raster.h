/* libigs_raster -- B200 (sm_100a) differentiable tile-based Gaussian-splatting
 * rasterizer.  C ABI: plain pointers and sizes, no torch types, no C++
 * exceptions across the boundary.
 *
 * This is the drop-in boundary for the reference's native rasterizer
 * ("DGR" = /root/reference/submodules/diff-gaussian-rasterization):
 *
 *   igs_forward       replaces CudaRasterizer::Rasterizer::forward
 *                     (DGR cuda_rasterizer/rasterizer.h:35-61, called from
 *                     RasterizeGaussiansCUDA, DGR rasterize_points.cu:35-133,
 *                     pybind name rasterize_gaussians, DGR ext.cpp:16)
 *   igs_backward      replaces CudaRasterizer::Rasterizer::backward
 *                     (rasterizer.h:63-100; RasterizeGaussiansBackwardCUDA,
 *                     rasterize_points.cu:135-239; ext.cpp:17)
 *   igs_mark_visible  replaces CudaRasterizer::Rasterizer::markVisible
 *                     (rasterizer.h:24-29; rasterize_points.cu:241-260;
 *                     ext.cpp:18)
 *   igs_*_bytes / igs_*_layout replace the reference's std::function<char*(size_t)>
 *                     resize callbacks and GeometryState/BinningState/ImageState
 *                     ::fromChunk (DGR rasterizer_impl.h:22-72): the caller
 *                     asks for sizes up front and owns every buffer.
 *
 * Ownership: the caller owns all memory; the library allocates no device
 * memory.  Buffers marked [keep] must stay alive and unmodified until the
 * matching igs_backward; [scratch] may be reused by the next call on the same
 * stream.  All device pointers must be 16-byte aligned (torch allocations are).
 * All launches go to the given stream.  Re-entrant per (device, stream) as long
 * as each stream uses its own scratch/ctrl buffers.
 *
 * Errors: every entry point returns IGS_OK (0) or a negative IGS_ERR_* code;
 * igs_last_error() returns a thread-local message.  With args->debug != 0 the
 * stream is synchronised and checked after every stage (the reference's
 * CHECK_CUDA, DGR cuda_rasterizer/auxiliary.h:167-174).
 */
#ifndef IGS_RASTER_H
#define IGS_RASTER_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGS_OK 0
#define IGS_ERR_INVALID (-1)   /* bad argument */
#define IGS_ERR_CUDA (-2)      /* CUDA runtime error, see igs_last_error() */
#define IGS_ERR_CAPACITY (-3)  /* num_rendered > capacity: grow [bin]/[sort] and call igs_forward_bin_render */

#define IGS_MAX_EXTRA_DIMS 34  /* DGR cuda_rasterizer/auxiliary.h:20 */
#define IGS_TILE 16            /* DGR cuda_rasterizer/config.h:16-17 */

typedef void* igs_stream_t; /* cudaStream_t */

typedef struct IgsFwdArgs {
  /* sizes: P gaussians, D active SH degree, M SH coefficients stored per
   * gaussian (may exceed (D+1)^2), ED extra-attribute channels */
  int32_t P, D, M, ED;
  int32_t W, H;
  float tan_fovx, tan_fovy, scale_modifier;
  int32_t prefiltered, debug;
  /* != 0: out_norm receives the unit normal n / max(|n|, 1e-12) -- the F.normalize(norm, dim=0) the
   * reference applies in Python right after the call (DGR diff_gauss/__init__.py:48) -- and |n| is
   * kept in the img workspace for igs_backward */
  int32_t normalize_norm;
  /* camera (device) */
  const float* bg;          /* [3] */
  const float* viewmatrix;  /* [16] row-vector convention (transposed W2C) */
  const float* projmatrix;  /* [16] */
  const float* campos;      /* [3] */
  /* gaussians (device, contiguous fp32; NULL = not provided) */
  const float* means3D;         /* [P,3] */
  const float* shs;             /* [P,M,3] or NULL */
  const float* colors_precomp;  /* [P,3] or NULL */
  const float* opacities;       /* [P] */
  const float* scales;          /* [P,3] */
  const float* rotations;       /* [P,4] */
  const float* cov3Ds_precomp;  /* [P,6] or NULL */
  const float* norm3Ds_precomp; /* [P,3] or NULL */
  const float* extra_attrs;     /* [P,ED] or NULL */
  /* outputs (device); every element is written, no pre-zeroing needed */
  float* out_color; /* [3,H,W] */
  float* out_depth; /* [1,H,W] */
  float* out_norm;  /* [3,H,W] un-normalised (normalize_norm == 0) or unit length */
  float* out_alpha; /* [1,H,W] */
  float* out_extra; /* [ED,H,W] or NULL */
  int32_t* radii;   /* [P] */
  /* workspaces (device) */
  void* geom;   /* [keep]    igs_geom_bytes(P) */
  void* img;    /* [keep]    igs_img_bytes(W,H) */
  void* bin;    /* [keep]    igs_bin_bytes(capacity) */
  void* sort;   /* [scratch] igs_sort_bytes(capacity) */
  void* ctrl;   /* [scratch] igs_ctrl_bytes(P,W,H): header, tile counts and cursors */
  int64_t capacity; /* instance capacity of bin/sort */
  /* pinned host slot the library copies num_rendered into */
  int32_t* num_rendered_pinned;
  /* != 0: [ctrl] is clean -- all zero, or last used with the SAME P, W, H by an igs_forward /
   * igs_forward_bin_render that returned IGS_OK or IGS_ERR_CAPACITY (the kernels put every counter
   * back to zero when they are done with it) -- so the library skips its memset of [ctrl] (one
   * stream operation less per frame) */
  int32_t ctrl_clean;
  /* != 0 (and the slot is device-mapped pinned memory): igs_forward queues the whole frame and
   * returns IGS_OK at once with *num_rendered = -1, without waiting for the count.  The library
   * sets the slot to -1 before it launches; the caller reads it once the frame is known to have
   * run, and compares it with `capacity` itself: a count above capacity means
   * the frame's images are incomplete and it must be rendered again with larger bin / sort
   * workspaces.  Nothing is written out of bounds in that case.  For forward-only pipelines that
   * own the point where a frame is consumed (the synthesis driver). */
  int32_t defer_count;
  /* != 0: the sorted 64-bit keys (tile << 32 | depth bits, the reference's binningState.point_list_keys,
   * DGR rasterizer_impl.h:47-57) are written into [bin] beside point_list.  Nothing on the path reads
   * them (the instance lists are built by tile, not by a global key sort); tests compare them with the
   * reference's bit for bit.  0 saves 8 bytes of HBM writes per instance. */
  int32_t write_keys;
} IgsFwdArgs;

typedef struct IgsBwdArgs {
  int32_t P, D, M, ED;
  int32_t W, H;
  float tan_fovx, tan_fovy, scale_modifier;
  int32_t debug;
  /* same value as in the forward call; != 0: dL_dnorm is the gradient w.r.t. the UNIT normal and
   * out_norm (the forward's normalised output) must be given */
  int32_t normalize_norm;
  int64_t num_rendered;
  const float* bg;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3Ds_precomp;
  const float* norm3Ds_precomp;
  const float* extra_attrs;
  const int32_t* radii;
  /* forward state */
  const void* geom;
  const void* img;
  const void* bin;
  const float* out_alpha; /* [1,H,W] as returned by forward */
  const float* out_norm;  /* [3,H,W] as returned by forward; only read when normalize_norm != 0 */
  /* upstream gradients (device, contiguous) */
  const float* dL_dcolor; /* [3,H,W] */
  const float* dL_ddepth; /* [1,H,W] */
  const float* dL_dnorm;  /* [3,H,W] */
  const float* dL_dalpha; /* [1,H,W] */
  const float* dL_dextra; /* [ED,H,W] or NULL */
  /* [scratch] igs_accum_bytes(P,ED); zeroed by the library */
  void* accum;
  /* outputs (device); every element is written, no pre-zeroing needed */
  float* dL_dmeans2D;   /* [P,3] (.z = 0) */
  float* dL_dcolors;    /* [P,3] */
  float* dL_dopacity;   /* [P,1] */
  float* dL_dmeans3D;   /* [P,3] */
  float* dL_dcov3D;     /* [P,6] */
  float* dL_dnorm3D;    /* [P,3] */
  float* dL_dsh;        /* [P,M,3] or NULL */
  float* dL_dscales;    /* [P,3] */
  float* dL_drotations; /* [P,4] */
  float* dL_dextra_attrs; /* [P,ED] or NULL */
  /* != 0: [accum] is clean -- all zero, or last used by an igs_backward that returned IGS_OK (the
   * per-Gaussian kernel zeroes every row it has read) -- so the library skips its memset of [accum] */
  int32_t accum_clean;
} IgsBwdArgs;

const char* igs_version(void);
const char* igs_last_error(void);

size_t igs_geom_bytes(int32_t P);
size_t igs_img_bytes(int32_t W, int32_t H);
/* byte offset, inside the img workspace, of the uint32 in which the device keeps the frame's
 * num_rendered with the state (a caller that did not wait for the count and lost its pinned slot
 * can still read it from there) */
size_t igs_img_info_offset(int32_t W, int32_t H);
size_t igs_bin_bytes(int64_t capacity);
size_t igs_sort_bytes(int64_t capacity);
size_t igs_ctrl_bytes(int32_t P, int32_t W, int32_t H);
size_t igs_accum_bytes(int32_t P, int32_t ED);

/* Byte offsets of the arrays inside the workspaces (for tests / inspection).
 * geom  (4): splat records float4[P][4] = {conic.xyz,opacity | x,y,depth,extra0 |
 *            r,g,b,nx | ny,nz,-,-}, tiles_touched u32[P], point_offsets u32[P]
 *            (inclusive scan; only after igs_point_offsets), clamped u8[P] (bit c = colour channel c clamped)
 * bin   (2): point_list u32[cap], keys u64[cap]      (sorted; keys only with write_keys)
 * img   (2): ranges uint2[T], n_contrib u32[H*W]  (+ private: tile order, normal lengths) */
void igs_geom_layout(int32_t P, int64_t* off4);
void igs_bin_layout(int64_t capacity, int64_t* off2);
void igs_img_layout(int32_t W, int32_t H, int64_t* off2);

/* Full forward.  Queues the per-Gaussian kernel and, sized by `capacity`, the binning, sorting
 * and render kernels (the count of instances is posted into the pinned slot by the device on the
 * way), then waits for the count only.  *num_rendered receives it.
 * Returns IGS_ERR_CAPACITY if it exceeds capacity: the outputs are then invalid (the kernels wrote
 * nothing) and the caller calls igs_forward_bin_render with bigger bin / sort buffers. */
int igs_forward(const IgsFwdArgs* args, igs_stream_t stream, int64_t* num_rendered);

/* The second half of igs_forward's contract for a call made with defer_count != 0: waits until
 * the count has been posted into args->num_rendered_pinned (same args as the igs_forward call),
 * stores it in *num_rendered and returns IGS_OK, or IGS_ERR_CAPACITY exactly as igs_forward
 * would have.  Lets the caller do its own bookkeeping between queueing the frame and needing
 * the count. */
int igs_forward_count(const IgsFwdArgs* args, igs_stream_t stream, int64_t* num_rendered);

/* After IGS_ERR_CAPACITY: same geom / img / ctrl buffers as the call that overflowed, bigger
 * bin / sort and capacity; renders the frame again and waits for its count. */
int igs_forward_bin_render(const IgsFwdArgs* args, igs_stream_t stream);

/* The reference's point_offsets -- the inclusive prefix sum of tiles_touched
 * (cub::DeviceScan::InclusiveSum, DGR rasterizer_impl.cu:280) -- written into the geom workspace of
 * a finished igs_forward, on request: the path itself places instances by tile and never reads it. */
int igs_point_offsets(int32_t P, void* geom, igs_stream_t stream);

int igs_backward(const IgsBwdArgs* args, igs_stream_t stream);

/* Measurement hooks (no reference counterpart; SURVEY.md 5 "tracing").  With profiling enabled
 * the kernels of every igs_forward / igs_backward of this host thread leave device-side time stamps
 * (%globaltimer of each kernel's last CTA; a ring of the last 256 frames), which costs one atomic per
 * CTA and puts nothing between the kernels -- their programmatic dependent launches stay as they are.
 * igs_profile_read synchronises the device and returns the MEAN stage durations in ms over the frames
 * recorded since igs_profile_enable(1) (-1 = stage did not run).  A stage lasts from the completion of
 * its predecessor to its own completion, so the stages of a half add up to the half: 0 preprocess (from
 * its first CTA's start; its last CTA also makes the tile ranges), 1 duplicate, 2 per-tile depth sort, 3 (unused), 4 render,
 * 5 render backward (from its first CTA's start), 6 per-Gaussian backward.  igs_launch_count returns how
 * many kernels this host thread has launched through the library so far. */
#define IGS_NUM_STAGES 7
int igs_profile_enable(int on);
int igs_profile_read(float* ms /* [IGS_NUM_STAGES] */, int32_t* frames);
uint64_t igs_launch_count(void);

/* Debug hook: with a non-NULL device buffer of 4*8*T uint64 words the forward render kernel records,
 * per (tile, 8x4 block) unit, {globaltimer at start, at end, smid | list length << 32, chunks of 32
 * entries walked | entries that survived the cull << 32}; NULL switches it off.  Used by
 * profiles/tools/trace_units.py to look at the load balance; not part of the reference API. */
int igs_debug_trace(void* device_buffer);

int igs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, igs_stream_t stream);

/* ---- the steps either side of the rasterizer call ------------------------
 *
 * f1  igs_compose_frame replaces the per-frame tensor expression of the synthesis loop,
 *     /root/reference/synthesize_fuse.py:64-76:
 *         alpha_mouth = dilate_fn(alpha_mouth)                      (only with dilate_ksize > 0; :19-22, k = 13)
 *         mouth_image = mouth_rgb + background / 255 * (1 - alpha_mouth)
 *         image       = face_rgb + mouth_image * (1 - face_alpha)
 *         pred        = (clamp(image, 0, 1) * 255).astype(uint8), HWC
 *     rgb planes [3,H,W], alpha planes [1,H,W], background [3,H,W] in 0..255 or NULL (black);
 *     out_u8 [H,W,3]; out_image [3,H,W] (the un-clamped `image`) or NULL.  The frame stays on
 *     the device: the caller copies ring slots to pinned memory asynchronously
 *     (instag_b200/synthesis.py) instead of the reference's blocking .cpu() per frame.
 *
 * f2  igs_activate_forward / _backward replace gaussian_renderer/__init__.py:227-235
 *         means3D   = xyz + d_xyz
 *         scales    = softplus(_scaling + d_scale)      (scene/gaussian_model.py:43)
 *         rotations = normalize(_rotation + d_rot)      (:51, eps 1e-12)
 *         opacity   = sigmoid(_opacity)                 (:48)
 *     and their autograd backward; d_* and upstream gradients may be NULL (zero).  The
 *     gradient of a d_* input equals the gradient of its parameter.
 *     igs_densify_stats replaces train_face.py:675-682 + scene/gaussian_model.py:683-685, in
 *     place, for the Gaussians with radii > 0:
 *         max_radii2D = max(max_radii2D, radii);  xyz_gradient_accum += |grad_means2D[:, :2]|;  denom += 1
 *     grad_means2D is [P,3]; max_radii2D, xyz_gradient_accum, denom are [P] / [P,1] float. */
int igs_compose_frame(const float* face_rgb, const float* face_alpha, const float* mouth_rgb,
                      const float* mouth_alpha, const float* background, int32_t H, int32_t W,
                      int32_t dilate_ksize, uint8_t* out_u8, float* out_image, igs_stream_t stream);
int igs_activate_forward(int32_t P, const float* xyz, const float* d_xyz, const float* scaling,
                         const float* d_scale, const float* rotation, const float* d_rot,
                         const float* opacity, float* means3D, float* scales, float* rotations,
                         float* opacities, igs_stream_t stream);
int igs_activate_backward(int32_t P, const float* scaling, const float* d_scale, const float* rotation,
                          const float* d_rot, const float* opacity, const float* g_means3D,
                          const float* g_scales, const float* g_rotations, const float* g_opacities,
                          float* g_xyz, float* g_scaling, float* g_rotation, float* g_opacity,
                          igs_stream_t stream);
int igs_densify_stats(int32_t P, const int32_t* radii, const float* grad_means2D, float* max_radii2D,
                      float* xyz_gradient_accum, float* denom, igs_stream_t stream);

/* ---- the exchange step of a view-sharded batch ----------------------------
 * igs_allreduce_nvls sums n_floats fp32 values (a multiple of 4) in place over the `world` GPUs of
 * one NVSwitch box, in ONE kernel over NVLink multicast memory (multimem.ld_reduce / multimem.st,
 * flag barriers through the peers' signal pads).  It replaces the ncclAllReduce of the packed
 * per-Gaussian gradient block (the reference has no multi-GPU path: north star).
 *   multicast_ptr       : the MULTICAST address of the buffer (same offset on every rank)
 *   signal_pad_ptrs_dev : device array of `world` pointers to the ranks' signal pads (zeroed;
 *                         signal_pad_bytes each, of which world + 2 words from byte 1024 on are used:
 *                         one flag per peer, two words for the rank's own CTAs)
 * as handed out by torch.distributed._symmetric_memory (multicast_ptr, signal_pad_ptrs_dev);
 * every rank must call it with the same n_floats on its own stream (instag_b200/sharding.py). */
int igs_allreduce_nvls(void* multicast_ptr, void* signal_pad_ptrs_dev, int32_t signal_pad_bytes,
                       int32_t rank, int32_t world, int64_t n_floats, igs_stream_t stream);

/* f3  igs_l1_ssim_* replace the photometric loss on the rasterizer's colour output,
 *         Ll1 = l1_loss(image, gt);  loss = Ll1 + lambda_dssim * (1.0 - ssim(image, gt))
 *     (/root/reference/train_face.py:421-422, utils/loss_utils.py:28-73: 11x11 Gaussian window,
 *     sigma 1.5, zero padding, C1 = 0.01^2, C2 = 0.03^2), and its autograd backward, which yields
 *     the dL_dcolor image igs_backward consumes.  img1 = image, img2 = gt, both [C,H,W].
 *     forward : writes the derivative maps dA, dB, dC ([C,H,W] each, kept for backward) and
 *               igs_l1_ssim_partials(C,H,W) pairs {sum of the SSIM map, sum of |img1 - img2|}, one per
 *               CTA; the caller adds them up: ssim = sum0 / (C*H*W), Ll1 = sum1 / (C*H*W).
 *     backward: grad_img1 [C,H,W] = *grad_out * d loss / d img1 (grad_out: device scalar). */
size_t igs_l1_ssim_partials(int32_t C, int32_t H, int32_t W);
int igs_l1_ssim_forward(const float* img1, const float* img2, int32_t C, int32_t H, int32_t W, float* dA,
                        float* dB, float* dC, float* partial, igs_stream_t stream);
/*     finish  : out3 <- {loss = Ll1 + lambda_dssim * (1 - ssim), Ll1, ssim} from the partial sums (device). */
int igs_l1_ssim_finish(const float* partial, int32_t C, int32_t H, int32_t W, float lambda_dssim, float* out3,
                       igs_stream_t stream);
int igs_l1_ssim_backward(const float* img1, const float* img2, const float* dA, const float* dB,
                         const float* dC, int32_t C, int32_t H, int32_t W, float lambda_dssim,
                         const float* grad_out, float* grad_img1, igs_stream_t stream);

/* igs_l1_ssim_forward_static: igs_l1_ssim_forward + igs_l1_ssim_finish with no caller scratch.
 *     NO COUNTERPART IN THE REFERENCE: added for callers that
 *     cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
 *     cudaMalloc).  The per-CTA partial sums go to a file-scope `static __device__` array inside the
 *     library (512 KB, held by every GPU that loads the library until exit) instead of a caller
 *     workspace; the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *     Limit: igs_l1_ssim_partials(C,H,W) <= 65536, i.e. up to [3,4096,4096]; larger frames are refused
 *     with IGS_ERR_INVALID (use igs_l1_ssim_forward with a workspace).  The caller passes the inputs and
 *     the outputs: dA, dB, dC ([C,H,W] each, kept for igs_l1_ssim_backward) and out3 = {loss, Ll1, ssim}.
 *     Every call on a device shares the one array and finds the last call's contents in it (none of them
 *     is read: each call overwrites the pairs it then sums); calls that OVERLAP on the GPU -- different
 *     streams, or host threads, on one device -- race on it and must be serialised by the caller. */
int igs_l1_ssim_forward_static(const float* img1, const float* img2, int32_t C, int32_t H, int32_t W,
                               float lambda_dssim, float* dA, float* dB, float* dC, float* out3,
                               igs_stream_t stream);

/* f3 (the rest)  igs_geom_loss_* replace the geometry regularisers train_face.py computes on the
 *     rasterizer's depth / normal / alpha images (/root/reference/train_face.py:469, :481-491 with
 *     utils/loss_utils.py:17-20 `normalize`, :540):
 *         l_normal = w_normal * (1 - normal_gt * normal).sum(0)[head ^ mouth].mean()
 *         l_depth  = w_depth  * (normalize(depth[0])[face ^ mouth] - normalize(depth_mono)[face ^ mouth]).abs().mean()
 *         l_alpha  = w_alpha  * (((1 - alpha) * head).mean() + (alpha * ~head).mean())
 *     (the reference's weights: 0.01, 1e-2, 1e-3; a weight of 0 switches a term off) and their autograd
 *     backward, which yields the dL_ddepth [1,H,W], dL_dnorm [3,H,W], dL_dalpha [1,H,W] images igs_backward
 *     consumes.  depth, alpha [1,H,W]; norm, normal_gt [3,H,W]; depth_mono [H,W]; masks [H,W] bytes
 *     (torch.bool).  workspace: igs_geom_loss_workspace(H) bytes, 8-byte aligned, kept for backward.
 *     forward : loss3 <- {l_normal, l_depth, l_alpha} (device).
 *     backward: gradients of *grad_out * (l_normal + l_depth + l_alpha) (grad_out: device scalar). */
size_t igs_geom_loss_workspace(int32_t H);
int igs_geom_loss_forward(const float* depth, const float* norm, const float* alpha, const float* depth_mono,
                          const float* normal_gt, const uint8_t* head_mask, const uint8_t* face_mask,
                          const uint8_t* mouth_mask, int32_t H, int32_t W, float w_normal, float w_depth, float w_alpha,
                          void* workspace, float* loss3, igs_stream_t stream);
int igs_geom_loss_backward(const float* depth, const float* depth_mono, const float* normal_gt, const uint8_t* head_mask,
                           const uint8_t* face_mask, const uint8_t* mouth_mask, int32_t H, int32_t W, float w_normal,
                           float w_depth, float w_alpha, const void* workspace, const float* grad_out, float* dL_ddepth,
                           float* dL_dnorm, float* dL_dalpha, igs_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif
