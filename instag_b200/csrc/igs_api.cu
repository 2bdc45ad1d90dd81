// C ABI of libigs_raster (include/igs_raster.h): argument validation, workspace
// carving, stage orchestration, error reporting.  No device memory is
// allocated here; the only host<->device traffic is the 4-byte num_rendered
// copy, which the host waits for AFTER it has queued every forward kernel.
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <initializer_list>
#include "igs_common.cuh"
#include "igs_kernels.cuh"

namespace igs {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return IGS_OK;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return IGS_ERR_CUDA;
}

// reference CHECK_CUDA (auxiliary.h:167-174): with debug, sync + check per stage
int stage_check(bool debug, cudaStream_t s, const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && debug) e = cudaStreamSynchronize(s);
  return check_cuda(e, what);
}

// ---- profiling: device-side time stamps (igs_common.cuh "Prof"), only when enabled -------
constexpr int PROF_RING = 256; // frames kept; older ones are overwritten
static __device__ unsigned long long g_stamps[PROF_RING * ST_WORDS];
struct Profiler {
  bool enabled = false;
  int slot = -1;        // frame being recorded (advanced by every forward)
  long long frames = 0; // frames started since enable
  unsigned long long* base = nullptr;
};
static thread_local Profiler g_prof;
static thread_local unsigned long long g_launches = 0;

Prof prof_begin(bool forward, cudaStream_t s) {
  // IGS_PDL_EARLY=1: griddepcontrol.launch_dependents at the top of every kernel.  Off: measured on
  // B200, dependents that take their places early slow their predecessors (C2 frame 222 -> 239 us)
  static const int early = getenv("IGS_PDL_EARLY") && getenv("IGS_PDL_EARLY")[0] == '1';
  Prof pr{nullptr, early};
  Profiler& p = g_prof;
  if (!p.enabled || !p.base) return pr;
  if (forward) {
    p.slot = (p.slot + 1) % PROF_RING;
    p.frames++;
    if (cudaMemsetAsync(p.base + (size_t)p.slot * ST_WORDS, 0, ST_WORDS * sizeof(unsigned long long), s) != cudaSuccess) {
      cudaGetLastError();
      return pr;
    }
  }
  if (p.slot >= 0) pr.stamps = p.base + (size_t)p.slot * ST_WORDS;
  return pr;
}
void count_launch(int n) { g_launches += (unsigned long long)n; }

struct EventCache { // one reusable event per host thread
  cudaEvent_t ev = nullptr;
  cudaEvent_t get() {
    if (!ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    return ev;
  }
};
static thread_local EventCache g_event;

// device-side address of a pinned host slot (cached per host thread; NULL if it is not mapped)
static int32_t* device_view_of_pinned(int32_t* host) {
  static thread_local int32_t* last_host = nullptr;
  static thread_local int32_t* last_dev = nullptr;
  if (host == last_host) return last_dev;
  void* d = nullptr;
  if (cudaHostGetDevicePointer(&d, host, 0) != cudaSuccess) { cudaGetLastError(); d = nullptr; }
  last_host = host;
  last_dev = (int32_t*)d;
  return last_dev;
}

// the kernels use 16-byte vector accesses on the caller's arrays (include/igs_raster.h: "all device
// pointers must be 16-byte aligned"); a misaligned one would fault and poison the CUDA context
static bool misaligned(std::initializer_list<const void*> ps) {
  for (const void* p : ps)
    if ((uintptr_t)p & 15u) return true;
  return false;
}

static int validate_fwd(const IgsFwdArgs* a) {
  if (!a) { set_error("null args"); return IGS_ERR_INVALID; }
  if (a->P < 0 || a->W <= 0 || a->H <= 0) { set_error("bad sizes P=%d W=%d H=%d", a->P, a->W, a->H); return IGS_ERR_INVALID; }
  if (a->ED < 0 || a->ED > IGS_MAX_EXTRA_DIMS) { set_error("ED=%d out of range [0,%d]", a->ED, IGS_MAX_EXTRA_DIMS); return IGS_ERR_INVALID; }
  if ((a->shs == nullptr) == (a->colors_precomp == nullptr) && a->P > 0) {
    set_error("provide exactly one of shs / colors_precomp"); return IGS_ERR_INVALID;
  }
  if (a->shs && (a->D < 0 || a->D > 3 || a->M < (a->D + 1) * (a->D + 1))) {
    set_error("sh degree %d needs M >= %d, got %d", a->D, (a->D + 1) * (a->D + 1), a->M); return IGS_ERR_INVALID;
  }
  if (a->P > 0 && (!a->means3D || !a->opacities || !a->scales || !a->rotations)) {
    set_error("means3D, opacities, scales, rotations are required"); return IGS_ERR_INVALID;
  }
  if (a->ED > 0 && !a->extra_attrs && a->P > 0) { set_error("ED>0 but extra_attrs is null"); return IGS_ERR_INVALID; }
  if (!a->out_color || !a->out_depth || !a->out_norm || !a->out_alpha || (a->ED > 0 && !a->out_extra)) {
    set_error("null output"); return IGS_ERR_INVALID;
  }
  if (!a->geom || !a->img || !a->bin || !a->sort || !a->ctrl || a->capacity < 0 || !a->num_rendered_pinned) {
    set_error("null workspace"); return IGS_ERR_INVALID;
  }
  if (a->capacity >= (1ll << 30)) { set_error("capacity must be < 2^30 instances"); return IGS_ERR_INVALID; }
  if (misaligned({a->means3D, a->shs, a->colors_precomp, a->opacities, a->scales, a->rotations, a->cov3Ds_precomp,
                  a->norm3Ds_precomp, a->extra_attrs, a->out_color, a->out_depth, a->out_norm, a->out_alpha, a->out_extra,
                  a->radii, a->geom, a->img, a->bin, a->sort, a->ctrl})) {
    set_error("device pointers must be 16-byte aligned"); return IGS_ERR_INVALID;
  }
  return IGS_OK;
}

static PreArgs make_pre(const IgsFwdArgs* a) {
  PreArgs p{};
  const GeomLayout gl(a->P);
  const CtrlLayout cl(a->P, a->W, a->H);
  char* g = (char*)a->geom;
  char* c = (char*)a->ctrl;
  p.P = a->P; p.D = a->D; p.M = a->M; p.ED = a->ED; p.W = a->W; p.H = a->H;
  p.grid_x = (a->W + TILE - 1) / TILE; p.grid_y = (a->H + TILE - 1) / TILE;
  p.prefiltered = a->prefiltered;
  p.tan_fovx = a->tan_fovx; p.tan_fovy = a->tan_fovy;
  p.focal_y = a->H / (2.0f * a->tan_fovy); // reference rasterizer_impl.cu:228-229
  p.focal_x = a->W / (2.0f * a->tan_fovx);
  p.scale_modifier = a->scale_modifier;
  p.capacity = a->capacity;
  p.viewmatrix = a->viewmatrix; p.projmatrix = a->projmatrix; p.campos = a->campos;
  p.means3D = a->means3D; p.shs = a->shs; p.colors_precomp = a->colors_precomp; p.opacities = a->opacities;
  p.scales = a->scales; p.rotations = a->rotations; p.cov3Ds_precomp = a->cov3Ds_precomp;
  p.norm3Ds_precomp = a->norm3Ds_precomp; p.extra_attrs = a->extra_attrs;
  p.radii = a->radii;
  p.splat = (float4*)(g + gl.splat);
  p.tiles_touched = (uint32_t*)(g + gl.tiles_touched);
  p.clamped = (uint8_t*)(g + gl.clamped);
  p.tile_count = (uint32_t*)(c + cl.tile_count);
  return p;
}

static BinArgs make_bin(const IgsFwdArgs* a) {
  const GeomLayout gl(a->P);
  const ImgLayout il(a->W, a->H);
  const BinLayout bl(a->capacity);
  const SortLayout sl(a->capacity);
  const CtrlLayout cl(a->P, a->W, a->H);
  char* g = (char*)a->geom; char* im = (char*)a->img; char* b = (char*)a->bin;
  char* so = (char*)a->sort; char* c = (char*)a->ctrl;
  BinArgs ba{};
  ba.P = a->P;
  ba.grid_x = (a->W + TILE - 1) / TILE; ba.grid_y = (a->H + TILE - 1) / TILE;
  ba.num_tiles = ba.grid_x * ba.grid_y;
  ba.capacity = a->capacity;
  ba.radii = a->radii;
  ba.splat = (const float4*)(g + gl.splat);
  ba.tiles_touched = (const uint32_t*)(g + gl.tiles_touched);
  ba.header = (CtrlHeader*)(c + cl.header);
  ba.tile_count = (uint32_t*)(c + cl.tile_count);
  ba.tile_cursor = (uint32_t*)(c + cl.tile_cursor);
  ba.ranges = (const uint2*)(im + il.ranges);
  ba.tile_order = (uint4*)(im + il.tile_order);
  ba.scan.num_tiles = ba.num_tiles;
  ba.scan.capacity = a->capacity;
  ba.scan.header = ba.header;
  ba.scan.num_rendered_host = device_view_of_pinned(a->num_rendered_pinned);
  ba.scan.info = (uint32_t*)(im + il.info);
  ba.scan.tile_count = ba.tile_count;
  ba.scan.tile_cursor = ba.tile_cursor;
  ba.scan.ranges = (uint2*)(im + il.ranges);
  ba.pairs = (unsigned long long*)(so + sl.pairs);
  ba.pairs2 = (unsigned long long*)(so + sl.pairs2);
  ba.point_list = (uint32_t*)(b + bl.point_list);
  ba.keys = a->write_keys ? (uint64_t*)(b + bl.keys) : nullptr;
  return ba;
}

static int run_bin_render(const IgsFwdArgs* a, cudaStream_t s, const Prof& prof) {
  const ImgLayout il(a->W, a->H);
  char* im = (char*)a->img;
  const bool debug = a->debug != 0;
  BinArgs ba = make_bin(a);
  ba.prof = prof;
  if (int r = check_cuda(launch_binning(ba, s), "binning launch")) return r;
  if (int r = stage_check(debug, s, "binning (duplicate/tile sort)")) return r;

  RenderArgs ra{};
  ra.prof = prof;
  ra.W = a->W; ra.H = a->H; ra.ED = a->ED; ra.grid_x = ba.grid_x; ra.grid_y = ba.grid_y;
  ra.header = ba.header;
  ra.ranges = ba.ranges;
  ra.tile_order = ba.tile_order;
  ra.point_list = ba.point_list;
  ra.splat = ba.splat;
  ra.extra_attrs = a->extra_attrs;
  ra.bg = a->bg;
  ra.out_color = a->out_color; ra.out_depth = a->out_depth; ra.out_norm = a->out_norm;
  ra.out_alpha = a->out_alpha; ra.out_extra = a->out_extra;
  ra.n_contrib = (uint32_t*)(im + il.n_contrib);
  ra.norm_len = (float*)(im + il.norm_len);
  ra.normalize = a->normalize_norm != 0;
  if (int r = check_cuda(launch_render(ra, s), "render launch")) return r;
  return stage_check(debug, s, "render");
}

// zero the header, the tile counts and the tile cursors
static int reset_ctrl(const IgsFwdArgs* a, cudaStream_t s) {
  const CtrlLayout cl(a->P, a->W, a->H);
  return check_cuda(cudaMemsetAsync(a->ctrl, 0, cl.zeroed(), s), "memset ctrl");
}

} // namespace igs

using namespace igs;

extern "C" {

const char* igs_version(void) { return "libigs_raster 0.1 (sm_100a)"; }
const char* igs_last_error(void) { return g_err; }

size_t igs_geom_bytes(int32_t P) { return GeomLayout(P).total; }
size_t igs_img_bytes(int32_t W, int32_t H) { return ImgLayout(W, H).total; }
size_t igs_img_info_offset(int32_t W, int32_t H) { return ImgLayout(W, H).info; }
size_t igs_bin_bytes(int64_t capacity) { return BinLayout(capacity).total; }
size_t igs_sort_bytes(int64_t capacity) { return SortLayout(capacity).total; }
size_t igs_ctrl_bytes(int32_t P, int32_t W, int32_t H) { return CtrlLayout(P, W, H).total; }
size_t igs_accum_bytes(int32_t P, int32_t ED) { return accum_bytes(P, ED); }

void igs_geom_layout(int32_t P, int64_t* off) {
  const GeomLayout l(P);
  off[0] = (int64_t)l.splat; off[1] = (int64_t)l.tiles_touched; off[2] = (int64_t)l.point_offsets; off[3] = (int64_t)l.clamped;
}
void igs_bin_layout(int64_t capacity, int64_t* off) {
  const BinLayout l(capacity);
  off[0] = (int64_t)l.point_list; off[1] = (int64_t)l.keys;
}
void igs_img_layout(int32_t W, int32_t H, int64_t* off) {
  const ImgLayout l(W, H);
  off[0] = (int64_t)l.ranges; off[1] = (int64_t)l.n_contrib;
}

int igs_forward(const IgsFwdArgs* a, igs_stream_t stream, int64_t* num_rendered) {
  if (int r = validate_fwd(a)) return r;
  cudaStream_t s = (cudaStream_t)stream;
  const bool debug = a->debug != 0;
  if (num_rendered) *num_rendered = 0;
  if (!a->ctrl_clean)
    if (int r = reset_ctrl(a, s)) return r;
  if (a->P == 0) { // reference rasterize_points.cu:93: outputs stay at their fill value
    const size_t HW = (size_t)a->W * a->H;
    cudaMemsetAsync(a->out_depth, 0, HW * 4, s); cudaMemsetAsync(a->out_alpha, 0, HW * 4, s);
    cudaMemsetAsync(a->out_norm, 0, 3 * HW * 4, s); cudaMemsetAsync(a->out_color, 0, 3 * HW * 4, s);
    if (a->ED > 0) cudaMemsetAsync(a->out_extra, 0, (size_t)a->ED * HW * 4, s);
    *a->num_rendered_pinned = 0;
    return stage_check(debug, s, "empty forward");
  }
  PreArgs pa = make_pre(a);
  const Prof prof = prof_begin(true, s);
  pa.prof = prof;
  int32_t* const slot_dev = device_view_of_pinned(a->num_rendered_pinned);
  volatile int32_t* const slot = (volatile int32_t*)a->num_rendered_pinned;
  if (slot_dev) *slot = -1; // the duplicate kernel's service CTA posts the count here (device-mapped pinned memory)
  if (int r = check_cuda(launch_preprocess(pa, s), "preprocess launch")) return r;
  if (int r = stage_check(debug, s, "preprocess")) return r;
  // speculative second half: sized by capacity, bounded on the device by num_rendered
  if (int r = run_bin_render(a, s, prof)) return r;
  if (a->defer_count && slot_dev) {
    // the caller reads the posted count later (igs_forward_count, or the slot itself: -1 = not
    // there yet) and holds it against the capacity; everything is queued, nothing on the device
    // depends on the host
    if (num_rendered) *num_rendered = -1;
    return IGS_OK;
  }
  if (!slot_dev) { // slot not device-mapped: a copy-engine transfer behind the frame, and a wait for it
    const CtrlLayout cl(a->P, a->W, a->H);
    const CtrlHeader* h = (const CtrlHeader*)((const char*)a->ctrl + cl.header);
    if (int r = check_cuda(cudaMemcpyAsync(a->num_rendered_pinned, &h->num_rendered, sizeof(int32_t),
                                           cudaMemcpyDeviceToHost, s), "copy num_rendered")) return r;
    cudaEvent_t ev = g_event.get();
    if (int r = check_cuda(cudaEventRecord(ev, s), "event record")) return r;
    if (int r = check_cuda(cudaEventSynchronize(ev), "wait num_rendered")) return r;
  }
  return igs_forward_count(a, stream, num_rendered);
}

int igs_forward_count(const IgsFwdArgs* a, igs_stream_t stream, int64_t* num_rendered) {
  if (!a || !a->num_rendered_pinned) { set_error("igs_forward_count: no slot"); return IGS_ERR_INVALID; }
  cudaStream_t s = (cudaStream_t)stream;
  volatile int32_t* const slot = (volatile int32_t*)a->num_rendered_pinned;
  // Wait for the posted count itself, not for an event: nothing sits between the kernels on the
  // stream, and the host wakes when the store lands (the last preprocess CTA posts it before it
  // builds the tile ranges).  cudaStreamQuery only guards against a kernel that died without posting.
  for (unsigned spin = 1; *slot == -1; spin++) {
    if ((spin & 0x3ff) == 0) {
      const cudaError_t q = cudaStreamQuery(s);
      if (q == cudaSuccess) break; // the stream is idle; the count is there, or it never will be
      if (q != cudaErrorNotReady) return check_cuda(q, "wait num_rendered");
    }
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
  }
  if (*slot == -1) { set_error("the preprocess kernel did not post num_rendered"); return IGS_ERR_CUDA; }
  const int64_t n = (int64_t)(uint32_t)*a->num_rendered_pinned;
  if (num_rendered) *num_rendered = n;
  if (n > a->capacity) {
    set_error("num_rendered %lld exceeds capacity %lld", (long long)n, (long long)a->capacity);
    return IGS_ERR_CAPACITY;
  }
  return IGS_OK;
}

int igs_forward_bin_render(const IgsFwdArgs* a, igs_stream_t stream) {
  // After IGS_ERR_CAPACITY: same geom / img / ctrl, bigger bin / sort.  The frame is queued again
  // from the top (every kernel leaves [ctrl] clean, also when the frame did not fit, so the tile
  // counts of the first attempt are gone); the per-Gaussian kernel is a tenth of a frame, and a
  // frame that outgrows 2x the largest count seen so far is rare.
  if (int r = validate_fwd(a)) return r;
  IgsFwdArgs b = *a;
  b.ctrl_clean = 1;
  b.defer_count = 0;
  return igs_forward(&b, stream, nullptr);
}

int igs_point_offsets(int32_t P, void* geom, igs_stream_t stream) {
  if (P < 0 || (P > 0 && !geom)) { set_error("igs_point_offsets: bad arguments"); return IGS_ERR_INVALID; }
  const GeomLayout gl(P);
  launch_point_offsets(P, (const uint32_t*)((char*)geom + gl.tiles_touched), (uint32_t*)((char*)geom + gl.point_offsets),
                       (cudaStream_t)stream);
  return stage_check(false, (cudaStream_t)stream, "point_offsets");
}

int igs_backward(const IgsBwdArgs* a, igs_stream_t stream) {
  if (!a) { set_error("null args"); return IGS_ERR_INVALID; }
  if (a->P < 0 || a->W <= 0 || a->H <= 0 || a->ED < 0 || a->ED > IGS_MAX_EXTRA_DIMS) { set_error("bad sizes"); return IGS_ERR_INVALID; }
  if (a->P == 0) return IGS_OK; // reference rasterize_points.cu:194
  if (!a->accum || !a->geom || !a->img || !a->bin || !a->out_alpha || !a->dL_dcolor || !a->dL_ddepth ||
      !a->dL_dnorm || !a->dL_dalpha || (a->ED > 0 && !a->dL_dextra) || (a->normalize_norm && !a->out_norm)) {
    set_error("null buffer"); return IGS_ERR_INVALID;
  }
  if (misaligned({a->means3D, a->shs, a->colors_precomp, a->scales, a->rotations, a->cov3Ds_precomp, a->norm3Ds_precomp,
                  a->extra_attrs, a->radii, a->geom, a->img, a->bin, a->out_alpha, a->out_norm, a->dL_dcolor, a->dL_ddepth,
                  a->dL_dnorm, a->dL_dalpha, a->dL_dextra, a->accum, a->dL_dmeans2D, a->dL_dcolors, a->dL_dopacity,
                  a->dL_dmeans3D, a->dL_dcov3D, a->dL_dnorm3D, a->dL_dsh, a->dL_dscales, a->dL_drotations,
                  a->dL_dextra_attrs})) {
    set_error("device pointers must be 16-byte aligned"); return IGS_ERR_INVALID;
  }
  cudaStream_t s = (cudaStream_t)stream;
  const bool debug = a->debug != 0;
  const GeomLayout gl(a->P);
  const ImgLayout il(a->W, a->H);
  const BinLayout bl(a->num_rendered);
  const char* g = (const char*)a->geom; const char* im = (const char*)a->img; const char* b = (const char*)a->bin;
  if (!a->accum_clean)
    if (int r = check_cuda(cudaMemsetAsync(a->accum, 0, accum_bytes(a->P, a->ED), s), "memset accum")) return r;

  const Prof prof = prof_begin(false, s);
  BwdRenderArgs ra{};
  ra.prof = prof;
  ra.W = a->W; ra.H = a->H; ra.ED = a->ED;
  ra.grid_x = (a->W + TILE - 1) / TILE; ra.grid_y = (a->H + TILE - 1) / TILE;
  ra.ranges = (const uint2*)(im + il.ranges);
  ra.tile_order = (const uint4*)(im + il.tile_order);
  ra.point_list = (const uint32_t*)(b + 0); // point_list is the first array of the bin workspace for any capacity
  (void)bl;
  ra.splat = (const float4*)(g + gl.splat);
  ra.extra_attrs = a->extra_attrs; ra.bg = a->bg; ra.out_alpha = a->out_alpha;
  ra.n_contrib = (const uint32_t*)(im + il.n_contrib);
  ra.info = (const uint32_t*)(im + il.info);
  ra.norm_len = (const float*)(im + il.norm_len);
  ra.normalize = a->normalize_norm != 0;
  ra.out_norm = a->out_norm;
  ra.dL_dcolor = a->dL_dcolor; ra.dL_ddepth = a->dL_ddepth; ra.dL_dnorm = a->dL_dnorm;
  ra.dL_dalpha = a->dL_dalpha; ra.dL_dextra = a->dL_dextra;
  ra.accum = (float*)a->accum;
  ra.ticket = (uint32_t*)((char*)a->accum + accum_ticket_offset(a->P, a->ED));
  ra.accum_extra = a->ED > 1 ? (float*)((char*)a->accum + align_up((size_t)a->P * ACC_F * sizeof(float))) : nullptr;
  if (a->num_rendered != 0) // (< 0: the caller does not know the count; the kernel finds empty lists by itself)
    if (int r = check_cuda(launch_render_backward(ra, s), "render backward launch")) return r;
  if (int r = stage_check(debug, s, "render backward")) return r;

  BwdPreArgs pa{};
  pa.prof = prof;
  pa.P = a->P; pa.D = a->D; pa.M = a->M; pa.ED = a->ED; pa.W = a->W; pa.H = a->H;
  pa.tan_fovx = a->tan_fovx; pa.tan_fovy = a->tan_fovy;
  pa.focal_y = a->H / (2.0f * a->tan_fovy); pa.focal_x = a->W / (2.0f * a->tan_fovx);
  pa.scale_modifier = a->scale_modifier;
  pa.viewmatrix = a->viewmatrix; pa.projmatrix = a->projmatrix; pa.campos = a->campos;
  pa.means3D = a->means3D; pa.shs = a->shs; pa.scales = a->scales; pa.rotations = a->rotations;
  pa.cov3Ds_precomp = a->cov3Ds_precomp; pa.norm3Ds_precomp = a->norm3Ds_precomp;
  pa.radii = a->radii;
  pa.splat = ra.splat;
  pa.clamped = (const uint8_t*)(g + gl.clamped);
  pa.accum = ra.accum; pa.accum_extra = ra.accum_extra; pa.ticket = ra.ticket;
  pa.dL_dmeans2D = a->dL_dmeans2D; pa.dL_dcolors = a->dL_dcolors; pa.dL_dopacity = a->dL_dopacity;
  pa.dL_dmeans3D = a->dL_dmeans3D; pa.dL_dcov3D = a->dL_dcov3D; pa.dL_dnorm3D = a->dL_dnorm3D;
  pa.dL_dsh = a->dL_dsh; pa.dL_dscales = a->dL_dscales; pa.dL_drotations = a->dL_drotations;
  pa.dL_dextra_attrs = a->dL_dextra_attrs;
  if (int r = check_cuda(launch_preprocess_backward(pa, s), "preprocess backward launch")) return r;
  return stage_check(debug, s, "preprocess backward");
}

int igs_profile_enable(int on) {
  Profiler& p = g_prof;
  if (on && !p.base) {
    void* sym = nullptr;
    if (int r = check_cuda(cudaGetSymbolAddress(&sym, g_stamps), "profile stamps")) return r;
    p.base = (unsigned long long*)sym;
  }
  p.enabled = on != 0;
  p.slot = -1;
  p.frames = 0;
  if (on)
    if (int r = check_cuda(cudaMemset(p.base, 0, sizeof(unsigned long long) * PROF_RING * ST_WORDS), "profile reset")) return r;
  return IGS_OK;
}

int igs_profile_read(float* ms, int32_t* frames) {
  if (!ms) { set_error("null ms"); return IGS_ERR_INVALID; }
  Profiler& p = g_prof;
  for (int k = 0; k < IGS_NUM_STAGES; k++) ms[k] = -1.f;
  if (frames) *frames = 0;
  if (!p.base) return IGS_OK;
  if (int r = check_cuda(cudaDeviceSynchronize(), "profile sync")) return r;
  static thread_local unsigned long long host[PROF_RING * ST_WORDS];
  if (int r = check_cuda(cudaMemcpy(host, p.base, sizeof host, cudaMemcpyDeviceToHost), "profile read")) return r;
  // stage k: from stamp `from` (its predecessor's last CTA; inverted = a start stamp) to stamp `to`
  static const int to[IGS_NUM_STAGES] = {ST_PRE, ST_DUP, ST_SORT, -1, ST_RENDER, ST_RBWD, ST_PBWD};
  static const int from[IGS_NUM_STAGES] = {ST_FWD_START, ST_PRE, ST_DUP, -1, ST_SORT, ST_BWD_START, ST_RBWD};
  int used = 0;
  for (int k = 0; k < IGS_NUM_STAGES; k++) {
    if (to[k] < 0) continue;
    double sum = 0.0;
    int n = 0;
    const int recorded = (int)(p.frames < PROF_RING ? p.frames : PROF_RING);
    for (int f = 0; f < recorded; f++) {
      const unsigned long long* st = host + (size_t)f * ST_WORDS;
      unsigned long long t0 = st[from[k]], t1 = st[to[k]];
      if (!t0 || !t1) continue;
      if (from[k] == ST_FWD_START || from[k] == ST_BWD_START) t0 = ~t0;
      if (t1 < t0) continue;
      sum += (double)(t1 - t0) * 1e-6;
      n++;
    }
    if (n) ms[k] = (float)(sum / n);
    if (n > used) used = n;
  }
  if (frames) *frames = used;
  return IGS_OK;
}

uint64_t igs_launch_count(void) { return g_launches; }

int igs_debug_trace(void* device_buffer) {
  return check_cuda(set_render_trace((unsigned long long*)device_buffer), "set trace buffer");
}

int igs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, igs_stream_t stream) {
  (void)projmatrix; // the reference's test only uses the view-space depth (auxiliary.h:155)
  if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("bad arguments"); return IGS_ERR_INVALID; }
  launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
  return stage_check(false, (cudaStream_t)stream, "mark_visible");
}

} // extern "C"
