// libigs_raster internal header: workspace layouts, error plumbing, fp32
// helpers.  sm_100a only.
#pragma once
#include <cstdint>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../include/igs_raster.h"

namespace igs {

constexpr int TILE = IGS_TILE;        // 16x16 pixel tiles (reference config.h:16-17)
constexpr int TILE_PIX = TILE * TILE; // 256 threads per render CTA, one per pixel
constexpr float NEAR_Z = 0.2f;        // reference auxiliary.h:155
constexpr float ALPHA_MIN = 1.0f / 255.0f;
constexpr float T_MIN = 0.0001f;

// ---- per-Gaussian splat record: 4 x float4 = 64 B, one aligned gather ------
//  r0 = {conic.x, conic.y, conic.z, opacity}
//  r1 = {mean2D.x, mean2D.y, view depth, extra[0]}
//  r2 = {r, g, b, normal.x}
//  r3 = {normal.y, normal.z, cut, -}   cut = 2 ln(255 opacity): the splat can only
//       blend (alpha >= 1/255) where  a dx^2 + 2 b dx dy + c dy^2 <= cut
constexpr int SPLAT_F4 = 4;

// ---- gradient accumulator: 16 floats = 64 B per Gaussian -------------------
//  with g = G * dL/dalpha summed over the pixels that blend the splat and (dx,dy) = centre - pixel:
//  0,1 sum g dx, g dy | 2,3,4 sum g dx^2, g dx dy, g dy^2 | 5 sum g (= dL/dopacity) | 6..8 dL/dcolor
//  9 dL/ddepth | 10..12 dL/dnormal | 13 dL/dextra[0] | 14,15 unused
constexpr int ACC_F = 16;

__host__ __device__ inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

struct GeomLayout {
  size_t splat, tiles_touched, point_offsets, clamped, total;
  __host__ __device__ explicit GeomLayout(int P) {
    size_t o = 0, p = (size_t)(P > 0 ? P : 1);
    splat = o;          o = align_up(o + p * SPLAT_F4 * sizeof(float4));
    tiles_touched = o;  o = align_up(o + p * sizeof(uint32_t));
    point_offsets = o;  o = align_up(o + p * sizeof(uint32_t));
    clamped = o;        o = align_up(o + p * sizeof(uint8_t));
    total = o;
  }
};

struct ImgLayout {
  size_t ranges, n_contrib, tile_order, norm_len, info, total;
  __host__ __device__ ImgLayout(int W, int H) {
    size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    size_t o = 0;
    ranges = o;     o = align_up(o + (T ? T : 1) * sizeof(uint2));
    tile_order = o; o = align_up(o + (T ? T : 1) * sizeof(uint4));    // {tile, start, end, -}, longest list first
    n_contrib = o;  o = align_up(o + (size_t)(W * H > 0 ? W * H : 1) * sizeof(uint32_t));
    norm_len = o;   o = align_up(o + (size_t)(W * H > 0 ? W * H : 1) * sizeof(float)); // |n| per pixel (fused normalize)
    info = o;       o = align_up(o + 4 * sizeof(uint32_t)); // {num_rendered of this frame, -, -, -}: kept with the state
    total = o;
  }
};

struct BinLayout { // sorted result, kept for backward
  size_t point_list, keys, total;
  __host__ __device__ explicit BinLayout(int64_t cap) {
    size_t c = (size_t)(cap > 0 ? cap : 1), o = 0;
    point_list = o;  o = align_up(o + c * sizeof(uint32_t));
    keys = o;        o = align_up(o + c * sizeof(uint64_t));
    total = o;
  }
};

struct SortLayout { // unsorted (depth|id) pairs grouped by tile, scratch
  size_t pairs, pairs2, total;
  __host__ __device__ explicit SortLayout(int64_t cap) {
    size_t c = (size_t)(cap > 0 ? cap : 1), o = 0;
    pairs = o;   o = align_up(o + c * sizeof(uint64_t));
    pairs2 = o;  o = align_up(o + c * sizeof(uint64_t)); // second level of the sort of segments beyond TS_CAP
    total = o;
  }
};

constexpr int PRE_THREADS = 512;   // Gaussians per preprocess / duplicate CTA
// Per-tile counters and cursors live one per 128-byte line: the L2 atomic units then work on
// ~T different lines instead of T/32 (measured: contended REDs to neighbouring words serialise).
constexpr int CNT_STRIDE = 32;     // in uint32 words
// CTA-local tile histograms live in dynamic shared memory (4 bytes per tile, opted in beyond 48 KB):
// 49152 tiles = 192 KB covers images up to ~3500x3500; larger ones count in global memory
constexpr int HIST_SMEM_MAX_TILES = 49152;

struct CtrlHeader {
  uint32_t num_rendered;  // sum of the tile counts, written by the service CTA of the duplicate kernel
  uint32_t overflow;      // num_rendered > capacity: the kernels behind it do nothing
  uint32_t render_ticket; // next (tile, 8x4 block) unit of the persistent render kernel
  uint32_t pad[5];
};

struct CtrlLayout {
  size_t header, tile_count, tile_cursor, total;
  __host__ __device__ CtrlLayout(int P, int W, int H) {
    (void)P;
    size_t o = 0;
    const size_t T = (size_t)((W + TILE - 1) / TILE) * ((H + TILE - 1) / TILE);
    header = o;       o = align_up(o + sizeof(CtrlHeader));
    tile_count = o;   o = align_up(o + (T ? T : 1) * CNT_STRIDE * sizeof(uint32_t));
    tile_cursor = o;  o = align_up(o + (T ? T : 1) * CNT_STRIDE * sizeof(uint32_t));
    total = o;
  }
  // a frame starts from an all-zero block; the kernels put every word back to zero when they are
  // done with it (tile_sort_kernel: counts and cursors of its tile), so a steady-state frame needs no memset
  __host__ __device__ size_t zeroed() const { return total; }
};

// [P][16] accumulators | [P][ED-1] side array | unit ticket of the persistent backward kernel
inline size_t accum_ticket_offset(int P, int ED) {
  size_t p = (size_t)(P > 0 ? P : 1);
  size_t extra = ED > 1 ? (size_t)(ED - 1) : 0; // channels 1.. live in a side array
  return align_up(p * ACC_F * sizeof(float)) + align_up(p * extra * sizeof(float));
}
inline size_t accum_bytes(int P, int ED) { return accum_ticket_offset(P, ED) + 256; }

// reference rasterizer_impl.cu:35-50 getHigherMsb: number of tile-id bits sorted
inline int higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return (int)msb;
}

// error plumbing (igs_api.cu)
void set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
int stage_check(bool debug, cudaStream_t s, const char* what);

// optional per-stage timing + launch accounting (igs_api.cu; include/igs_raster.h igs_profile_*).
// The kernels of a frame overlap their launches (programmatic dependent launch), so stage times are
// taken ON THE DEVICE: with profiling on, every CTA leaves %globaltimer in the frame's stamp block
// when it is done (atomicMax), and the first kernel of each half leaves the earliest time at which
// one of its CTAs got past griddepcontrol.wait.  A stage = from its predecessor's last CTA to its
// own last CTA; the stages of a half add up to the half.  Nothing sits between the kernels.
enum : int { ST_FWD_START = 0, ST_PRE, ST_DUP, ST_SORT, ST_RENDER, ST_BWD_START, ST_RBWD, ST_PBWD, ST_WORDS = 16 };
struct Prof {
  unsigned long long* stamps; // [ST_WORDS] of this frame, zeroed when the frame starts; NULL: not profiling
  int early;                  // != 0: release the programmatic dependents at the top of the kernel
};
Prof prof_begin(bool forward, cudaStream_t s); // advances to the next frame's block when forward
void count_launch(int n = 1);

// ---- programmatic dependent launch ------------------------------------------------
// The kernels of a frame run back to back on one stream.  Each dependent kernel is launched with
// the programmatic-stream-serialization attribute and starts with pdl_wait(): its launch is
// processed while its predecessor still runs, its CTAs land on SMs as the predecessor's CTAs
// retire, and they block in pdl_wait() until the predecessor grid has completed and its memory
// is visible -- the launch latency between kernels disappears.  (In a kernel launched without
// the attribute pdl_wait() returns immediately.)
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Lets the NEXT kernel on the stream (if it was launched as a programmatic dependent) be scheduled
// as soon as every CTA of this grid has started: its CTAs run their prologue (shared-memory
// clears, constant loads) on whatever SM resources are free and then block in pdl_wait() until
// this grid has completed.  Without it they are only scheduled as this grid's CTAs exit.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// (one thread per CTA / per warp calls these)
__device__ __forceinline__ void stamp_done(const Prof& p, int word) { if (p.stamps) atomicMax(p.stamps + word, global_ns()); }
__device__ __forceinline__ void stamp_begun(const Prof& p, int word) { if (p.stamps) atomicMax(p.stamps + word, ~global_ns()); }

template <class... KArgs, class... Args>
inline cudaError_t launch_after(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s,
                                Args&&... args) {
  // (a launch error is returned AND stays in cudaGetLastError for the stage check that follows)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  // IGS_NO_PDL=1: plain stream order (griddepcontrol.wait is then a no-op); for debugging only
  static const bool pdl = !(getenv("IGS_NO_PDL") && getenv("IGS_NO_PDL")[0] == '1');
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- device helpers ---------------------------------------------------------
// a0*b0 + a1*b1 + a2*b2 with the contraction nvcc applies to that expression
// (left product of the first sum fused, see DESIGN.md "FMA contraction");
// spelled with intrinsics so the result does not depend on compiler mood.
__device__ __forceinline__ float dot3c(float a0, float b0, float a1, float b1, float a2, float b2) {
  return __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)));
}

// m[r]*x + m[4+r]*y + m[8+r]*z + m[12+r]   (reference auxiliary.h:59-78)
__device__ __forceinline__ float xform_row(const float* __restrict__ m, int r, float x, float y, float z) {
  return __fadd_rn(dot3c(m[r], x, m[4 + r], y, m[8 + r], z), m[12 + r]);
}

// reference auxiliary.h:42-45: evaluated in double, contracted to one DFMA
__device__ __forceinline__ float ndc2pix(float v, int S) {
  return (float)(__fma_rn((double)v + 1.0, (double)S, -1.0) * 0.5);
}

// reference auxiliary.h:47-57 (max_radius is an int there)
__device__ __forceinline__ void tile_rect(float px, float py, int radius, int gx, int gy,
                                          int& x0, int& y0, int& x1, int& y1) {
  const float r = (float)radius;
  x0 = min(gx, max(0, (int)(__fsub_rn(px, r) * 0.0625f)));
  y0 = min(gy, max(0, (int)(__fsub_rn(py, r) * 0.0625f)));
  x1 = min(gx, max(0, (int)(__fsub_rn(__fadd_rn(__fadd_rn(px, r), 16.f), 1.f) * 0.0625f)));
  y1 = min(gy, max(0, (int)(__fsub_rn(__fadd_rn(__fadd_rn(py, r), 16.f), 1.f) * 0.0625f)));
}

// -0.5(a dx^2 + c dy^2) - b dx dy, contraction as in reference forward.cu:406
__device__ __forceinline__ float gauss_power(float cx, float cy, float cz, float dx, float dy) {
  const float s = __fmaf_rn(__fmul_rn(cx, dx), dx, __fmul_rn(__fmul_rn(cz, dy), dy));
  return __fmaf_rn(s, -0.5f, -__fmul_rn(__fmul_rn(cy, dx), dy));
}

// ---- warp-level culling --------------------------------------------------
// Can the splat {conic (a,b,c), mean (mx,my), cut} blend into ANY pixel of the
// pixel rectangle [x0,x1] x [y0,y1]?  A pixel blends only if
//   power = -q/2 <= 0  and  opacity*exp(power) >= 1/255   <=>   q <= cut,
// q = a dx^2 + 2 b dx dy + c dy^2.  q is convex (the conic is positive definite), so
// its minimum over the rectangle is 0 if the mean is inside, else it lies on one of the
// (at most two) edges facing the mean, where it is a clamped 1-D quadratic minimum.
// CONSERVATIVE by construction: a margin covers fp32 rounding of q here and of `power`
// in the compositing loops, every comparison is written so that NaN means "keep",
// and a conic that is not positive definite is always kept.  Culling therefore never
// changes a result, it only skips work (tests/test_gpu_parity.py checks n_contrib and
// the images against the unculled reference on every scene).
__device__ __forceinline__ bool splat_may_touch(float a, float b, float c, float mx, float my, float cut,
                                                float x0, float x1, float y0, float y1) {
  const float ex0 = x0 - mx, ex1 = x1 - mx, ey0 = y0 - my, ey1 = y1 - my;
  const float dxn = fminf(fmaxf(0.f, ex0), ex1); // offset of the nearest rectangle point
  const float dyn = fminf(fmaxf(0.f, ey0), ey1);
  // facing vertical edge (dx = dxn): minimise over dy
  const float dy1 = fminf(fmaxf(__fdividef(-b * dxn, c), ey0), ey1);
  const float p1 = a * dxn * dxn, p2 = 2.f * b * dxn * dy1, p3 = c * dy1 * dy1;
  // facing horizontal edge (dy = dyn): minimise over dx
  const float dx2 = fminf(fmaxf(__fdividef(-b * dyn, a), ex0), ex1);
  const float r1 = a * dx2 * dx2, r2 = 2.f * b * dx2 * dyn, r3 = c * dyn * dyn;
  const float q1 = p1 + p2 + p3, m1 = p1 + fabsf(p2) + p3;
  const float q2 = r1 + r2 + r3, m2 = r1 + fabsf(r2) + r3;
  const bool far1 = q1 > cut + 0.01f + 1e-5f * m1;
  const bool far2 = q2 > cut + 0.01f + 1e-5f * m2;
  const bool pd = a > 0.f && c > 0.f && a * c > b * b;
  return !(far1 && far2 && pd);
}

} // namespace igs
