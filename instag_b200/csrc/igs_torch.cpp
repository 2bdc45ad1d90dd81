// Compiled host side of the drop-in: the autograd node of the rasterizer as a C++
// torch::autograd::Function over the C ABI of libigs_raster (include/igs_raster.h).
//
// Role in the reference ("DGR" = submodules/diff-gaussian-rasterization): the torch binding
// rasterize_points.cu:35-239 (RasterizeGaussiansCUDA / RasterizeGaussiansBackwardCUDA) plus the
// autograd.Function of diff_gauss/__init__.py:52-178.  The Python twin of this file is
// instag_b200/rasterizer.py (_forward_native / _backward_native / _RasterizeGaussians); both call
// the same kernels.  This one exists because an end-to-end step of the C2 workload is host bound:
// allocations, argument marshalling and the backward node cost ~150 us per step from Python and
// ~30 us from here (the backward also runs on the autograd engine's thread without the GIL).
//
// No kernel is launched from this file and there is no fallback: every entry point ends in
// igs_forward / igs_backward.
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <array>
#include <atomic>
#include <deque>
#include <map>
#include <mutex>
#include <tuple>
#include <chrono>
#include <vector>
#include <cstdlib>

#include "igs_raster.h"

namespace {

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

// Count slots in the pinned block of a Scratch (int32 each, written by the device):
//   [0]                       the call that waits for its count inside igs_forward
//   [kDeferFirst, +kDeferSlots)  calls under synthesis.DeferredCounts (forward only; read at harvest)
//   [kGradFirst, +kGradSlots)    calls that will get a backward: the count is read there (a ring;
//                                a slot that was handed out again is recognised by its serial)
//   [kGraphFirst, +kGraphSlots)  calls recorded into a CUDA graph: read by captured_overflowed()
constexpr int kDeferFirst = 1, kDeferSlots = 127;
constexpr int kGradFirst = 128, kGradSlots = 128;
constexpr int kGraphFirst = 256, kGraphSlots = 64;
constexpr int kPinnedSlots = kGraphFirst + kGraphSlots;

// instance capacity to allocate for a (P, H, W): raised from every count that is read (any thread)
struct Hint {
  std::atomic<int64_t> capacity{0};
  void raise(int64_t n);
};

// Reusable buffers of one (device, stream, host thread); see rasterizer.py::_DeviceScratch.
struct Scratch {
  Tensor sort, ctrl, accum, pinned;
  bool ctrl_dirty = false, accum_dirty = false;
  std::tuple<int64_t, int64_t, int64_t> ctrl_key{-1, -1, -1};
  std::map<std::tuple<int64_t, int64_t, int64_t>, Hint> capacity_hint;
  std::map<std::tuple<int64_t, int64_t, int64_t>, Tensor> zero_imgs;
  int defer_next = kDeferFirst, defer_in_use = 0;  // ring of count slots for calls that do not wait (see Pending)
  int grad_next = kGradFirst, graph_next = kGraphFirst;
  int64_t grad_serial = 0;
  std::array<std::atomic<int64_t>, kGradSlots> grad_slot_serial{};
  // the last call whose count a backward will read: looked at (not waited for) by the next forward,
  // so that a forward that never gets a backward does not go unchecked for ever
  volatile int32_t* last_slot = nullptr;
  int64_t last_capacity = 0, last_serial = 0;
  Hint* last_hint = nullptr;

  // `ctrl` and `accum` are handed to the library zeroed and come back clean from every successful
  // call; a failed call marks them dirty and they are cleared before the next one.
  Tensor& buf(Tensor& t, bool& dirty, size_t nbytes, const at::Device& dev, bool zeroed) {
    if (!t.defined() || (size_t)t.numel() < nbytes) {
      auto opt = at::TensorOptions().dtype(at::kByte).device(dev);
      int64_t n = (int64_t)(nbytes + nbytes / 4) + 256;
      t = zeroed ? at::zeros({n}, opt) : at::empty({n}, opt);
      dirty = false;
    } else if (dirty) {
      t.zero_();
      dirty = false;
    }
    return t;
  }
  const Tensor& zero_img(int64_t C, int64_t H, int64_t W, const at::Device& dev) {
    auto key = std::make_tuple(C, H, W);
    auto it = zero_imgs.find(key);
    if (it == zero_imgs.end())
      it = zero_imgs.emplace(key, at::zeros({C, H, W}, at::TensorOptions().dtype(at::kFloat).device(dev))).first;
    return it->second;
  }
};

// Head-room over the largest count seen: a call that does not wait for its count cannot grow its
// workspaces, so they are sized generously (12 + 16 bytes per instance; IGS_COUNT_HEADROOM overrides).
double headroom() {
  static const double h = [] {
    const char* e = getenv("IGS_COUNT_HEADROOM");
    const double v = e ? atof(e) : 2.0;
    return v < 1.05 ? 1.05 : v;
  }();
  return h;
}
void Hint::raise(int64_t n) {
  const int64_t want = std::max<int64_t>((int64_t)(n * headroom()) + 4096, 1 << 14);
  int64_t cur = capacity.load(std::memory_order_relaxed);
  while (cur < want && !capacity.compare_exchange_weak(cur, want, std::memory_order_relaxed)) {}
}
// IGS_SYNC_COUNTS=1: every call waits for its count inside the forward, as the reference does
// (rasterizer_impl.cu:284), and grows its workspaces on the spot
bool sync_counts() {
  static const bool v = getenv("IGS_SYNC_COUNTS") && getenv("IGS_SYNC_COUNTS")[0] == '1';
  return v;
}

std::map<std::pair<int, cudaStream_t>, Scratch>& all_scratch() {
  // never destroyed: the tensors must not outlive the CUDA context at thread / process exit
  thread_local auto* all = new std::map<std::pair<int, cudaStream_t>, Scratch>();
  return *all;
}

Scratch& scratch(int device, cudaStream_t stream) {
  Scratch& s = all_scratch()[{device, stream}];
  if (!s.pinned.defined()) s.pinned = at::zeros({kPinnedSlots}, at::kInt).pin_memory();
  return s;
}

bool capturing(cudaStream_t stream) {
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) { cudaGetLastError(); return false; }
  return st != cudaStreamCaptureStatusNone;
}

// calls recorded into CUDA graphs: (slot, capacity, hint); their counts are there after a replay
struct Captured { volatile int32_t* slot; int64_t capacity; Hint* hint; };
std::mutex g_captured_mutex;
std::vector<Captured>* g_captured = new std::vector<Captured>();

// Waits (briefly: the count is posted early in the forward, long before a backward can start)
// for a posted count; -1 if the stream went idle or failed without one.
int64_t wait_count(volatile int32_t* slot, cudaStream_t stream) {
  for (uint64_t spin = 1; *slot == -1; ++spin) {
    if ((spin & 0x3ff) == 0) {
      const cudaError_t q = cudaStreamQuery(stream);
      if (q != cudaErrorNotReady) {
        if (q != cudaSuccess) cudaGetLastError();
        break;
      }
    }
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
  }
  const int32_t v = *slot;
  return v == -1 ? -1 : (int64_t)(uint32_t)v;
}

// ---- deferred counts (forward-only pipelines; synthesis.DeferredCounts is the Python face) ----
// A forward call normally waits for the preprocess kernel to post num_rendered, because only then
// is it known that the instance workspaces were large enough.  A pipeline that renders frame after
// frame without gradients and owns the point where a frame is consumed can take that wait off the
// host's critical path: the call queues the frame, notes (slot, capacity) here and returns; the
// pipeline asks `deferred_overflowed(first, last)` when it harvests the frame -- by then the count
// has long been posted -- and renders the frame again if it was cut short.
struct Pending {
  volatile int32_t* slot;
  int64_t capacity;
  Hint* hint;     // the scratch's capacity hint for this (P, H, W): raised from the count when it is read
  Scratch* owner; // its ring of slots gets this one back
  cudaStream_t stream;
};
thread_local bool t_defer = false;
thread_local int64_t t_serial = 0;                              // deferred calls so far on this thread
thread_local std::deque<std::pair<int64_t, Pending>>* t_pending = nullptr;  // (serial, entry), oldest first

// fp32, contiguous and 16-byte aligned (the kernels use vector accesses; a contiguous view at an odd
// storage offset -- which the reference accepts -- is copied)
inline Tensor f32c(const Tensor& t) {
  Tensor r = t.scalar_type() == at::kFloat ? t : t.to(at::kFloat);
  if (!r.is_contiguous()) return r.contiguous();
  if (r.numel() && (reinterpret_cast<uintptr_t>(r.data_ptr()) & 15u)) return r.clone();
  return r;
}
inline const float* fptr(const Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }
inline float* fptr_mut(Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }

// Host time of the two halves of the node, by phase (steady_clock marks, ~25 ns each; read and reset
// with host_trace()): where an end-to-end step spends its host time is the question of profiles/tools/host_trace.py.
struct HostTrace {
  std::atomic<int64_t> ns[8]{};
  std::atomic<int64_t> calls[2]{};
  std::atomic<int64_t> modes[4]{};  // forwards that waited / deferred / left the count to backward / were recorded
};
HostTrace g_trace;
inline int64_t now_ns() {
  return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

[[noreturn]] void fail(int st, const char* what) {
  TORCH_CHECK(false, "libigs_raster ", what, " failed (", st, "): ", igs_last_error());
}

// Packed gradient block: same layout as rasterizer.py::gradient_layout (parameter gradients first,
// every block 16-byte aligned).  Return order: means2D, colours, opacity, means3D, cov3D, norm3D,
// sh, scales, rotations, extra.
struct GradLayout {
  std::array<int64_t, 10> start, width;
  int64_t total;
};
GradLayout gradient_layout(int64_t P, int64_t M, int64_t ED) {
  GradLayout g;
  g.width = {3, 3, 1, 3, 6, 3, 3 * M, 3, 4, ED};
  static const int order[10] = {3, 1, 2, 6, 7, 8, 9, 0, 4, 5};
  int64_t o = 0;
  for (int i : order) {
    g.start[i] = o;
    o += (P * g.width[i] + 3) / 4 * 4;
  }
  g.total = o;
  return g;
}

// Optional allocator of the packed gradient block (sharding.NvlsGradientBlock puts it in NVLink
// multicast memory): a Python callable fn(n_floats, device) -> flat float32 tensor.
py::object* g_grad_alloc = new py::object();  // leaked on purpose (no Python at static destruction)
std::mutex g_grad_alloc_mutex;
std::atomic<bool> g_have_grad_alloc{false};

struct RasterizeFn : public torch::autograd::Function<RasterizeFn> {
  static variable_list forward(AutogradContext* ctx, const Tensor& means3D, const Tensor& means2D, const Tensor& sh,
                               const Tensor& colors_precomp, const Tensor& opacities, const Tensor& scales,
                               const Tensor& rotations, const Tensor& cov3Ds_precomp, const Tensor& norm3Ds_precomp,
                               const Tensor& extra_attrs, const Tensor& bg, const Tensor& viewmatrix,
                               const Tensor& projmatrix, const Tensor& campos, double tanfovx, double tanfovy,
                               double scale_modifier, int64_t H, int64_t W, int64_t sh_degree, bool prefiltered,
                               bool fuse_normalize, bool wants_grad) {
    const int64_t t_entry = now_ns();
    (void)means2D;
    TORCH_CHECK(means3D.dim() == 2 && means3D.size(1) == 3, "means3D must have dimensions (num_points, 3)");
    TORCH_CHECK(means3D.is_cuda(), "libigs_raster is CUDA-only: means3D must be a CUDA tensor (no CPU fallback)");
    TORCH_CHECK(extra_attrs.size(0) == 0 || extra_attrs.size(1) <= IGS_MAX_EXTRA_DIMS, "too many extra attributes");
    const at::Device dev = means3D.device();
    c10::cuda::CUDAGuard guard(dev);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    Scratch& sc = scratch(dev.index(), stream);
    const int64_t P = means3D.size(0);
    const int64_t M = sh.numel() ? sh.size(1) : 0;
    const int64_t ED = extra_attrs.size(0) != 0 ? extra_attrs.size(1) : 0;
    auto f32 = at::TensorOptions().dtype(at::kFloat).device(dev);
    auto u8 = at::TensorOptions().dtype(at::kByte).device(dev);
    const bool in_graph = capturing(stream);
    // The previous call of this scratch whose count a backward was to read: if that count is there and
    // was too large, say so now (its backward will too; a forward without a backward would go unnoticed).
    if (sc.last_slot && !in_graph) {
      const int32_t v = *sc.last_slot;
      if (v != -1) {
        const int64_t n = (int64_t)(uint32_t)v, cap = sc.last_capacity;
        const bool same = sc.grad_slot_serial[(sc.last_slot - sc.pinned.data_ptr<int32_t>()) - kGradFirst].load() == sc.last_serial;
        Hint* h = sc.last_hint;
        sc.last_slot = nullptr;
        if (same) {
          h->raise(n);
          TORCH_CHECK(n <= cap, "libigs_raster: the previous rasterization produced ", n, " instances but its "
                      "workspaces held ", cap, " (sized from earlier frames): its images were not rendered. The "
                      "capacity has been raised; IGS_SYNC_COUNTS=1 makes every call wait for its count instead.");
        }
      }
    }

    Tensor color = at::empty({3, H, W}, f32), depth = at::empty({1, H, W}, f32), norm = at::empty({3, H, W}, f32),
           alpha = at::empty({1, H, W}, f32);
    Tensor extra = ED > 0 ? at::empty({ED, H, W}, f32) : at::empty({0}, f32);
    Tensor radii = at::empty({P}, at::TensorOptions().dtype(at::kInt).device(dev));

    const auto key = std::make_tuple(P, H, W);
    const size_t geom_bytes = igs_geom_bytes((int32_t)P), img_bytes = igs_img_bytes((int32_t)W, (int32_t)H);
    auto hint = sc.capacity_hint.find(key);
    const bool have_hint = hint != sc.capacity_hint.end() && hint->second.capacity.load(std::memory_order_relaxed) > 0;
    int64_t capacity = have_hint ? hint->second.capacity.load(std::memory_order_relaxed) : std::max<int64_t>(8 * P, 1 << 16);
    TORCH_CHECK(!in_graph || (have_hint && P > 0),
                "libigs_raster: rasterize these sizes once before recording them into a CUDA graph");
    // forward state kept for backward: one allocation, [geom | img]
    Tensor state = at::empty({(int64_t)(geom_bytes + img_bytes)}, u8);
    Tensor binb = at::empty({(int64_t)igs_bin_bytes(capacity)}, u8);
    bool no_dirt = false;
    Tensor& sort = sc.buf(sc.sort, no_dirt, igs_sort_bytes(capacity), dev, false);
    if (sc.ctrl_key != key) {  // another (P, H, W) lays [ctrl] out differently
      sc.ctrl_dirty = true;
      sc.ctrl_key = key;
    }
    Tensor& ctrl = sc.buf(sc.ctrl, sc.ctrl_dirty, igs_ctrl_bytes((int32_t)P, (int32_t)W, (int32_t)H), dev, true);

    const int64_t t_alloc = now_ns();
    Tensor k_bg = f32c(bg), k_view = f32c(viewmatrix), k_proj = f32c(projmatrix), k_cam = f32c(campos),
           k_means = f32c(means3D), k_sh = f32c(sh), k_col = f32c(colors_precomp), k_op = f32c(opacities),
           k_sc = f32c(scales), k_rot = f32c(rotations), k_cov = f32c(cov3Ds_precomp), k_n3 = f32c(norm3Ds_precomp),
           k_ex = f32c(extra_attrs);
    IgsFwdArgs a = {};
    a.P = (int32_t)P, a.D = (int32_t)sh_degree, a.M = (int32_t)M, a.ED = (int32_t)ED, a.W = (int32_t)W, a.H = (int32_t)H;
    a.tan_fovx = (float)tanfovx, a.tan_fovy = (float)tanfovy, a.scale_modifier = (float)scale_modifier;
    a.prefiltered = prefiltered, a.debug = 0, a.normalize_norm = fuse_normalize;
    a.bg = fptr(k_bg), a.viewmatrix = fptr(k_view), a.projmatrix = fptr(k_proj), a.campos = fptr(k_cam);
    a.means3D = fptr(k_means), a.shs = fptr(k_sh), a.colors_precomp = fptr(k_col), a.opacities = fptr(k_op);
    a.scales = fptr(k_sc), a.rotations = fptr(k_rot), a.cov3Ds_precomp = fptr(k_cov), a.norm3Ds_precomp = fptr(k_n3);
    a.extra_attrs = fptr(k_ex);
    a.out_color = color.data_ptr<float>(), a.out_depth = depth.data_ptr<float>(), a.out_norm = norm.data_ptr<float>();
    a.out_alpha = alpha.data_ptr<float>(), a.out_extra = fptr_mut(extra);
    a.radii = P > 0 ? radii.data_ptr<int32_t>() : nullptr;
    uint8_t* st_ptr = state.data_ptr<uint8_t>();
    a.geom = st_ptr, a.img = st_ptr + geom_bytes, a.bin = binb.data_ptr(), a.sort = sort.data_ptr(), a.ctrl = ctrl.data_ptr();
    a.capacity = capacity;
    a.num_rendered_pinned = sc.pinned.data_ptr<int32_t>();
    a.ctrl_clean = 1;
    sc.ctrl_dirty = true;  // until the call has returned without an error

    const int64_t t_args = now_ns();
    // Who reads the count?  (1) nobody can later: this call waits for it (and grows the workspaces on
    // the spot); (2) a synthesis.DeferredCounts pipeline at harvest; (3) the backward of this call;
    // (4) the owner of the CUDA graph being recorded, after a replay (captured_overflowed()).
    enum { kWait, kDeferred, kBackward, kGraph } mode = kWait;
    if (P > 0 && have_hint && !sync_counts()) {
      if (in_graph) mode = kGraph;
      else if (wants_grad) mode = kBackward;
      else if (t_defer && sc.defer_in_use < kDeferSlots) mode = kDeferred;
    }
    g_trace.modes[mode]++;
    if (mode != kWait) {
      int32_t* const pinned = sc.pinned.data_ptr<int32_t>();
      int slot;
      if (mode == kDeferred) {
        slot = sc.defer_next;
        sc.defer_next = (slot + 1 - kDeferFirst) % kDeferSlots + kDeferFirst;
        sc.defer_in_use++;
      } else if (mode == kBackward) {
        slot = sc.grad_next;
        sc.grad_next = (slot + 1 - kGradFirst) % kGradSlots + kGradFirst;
      } else {
        TORCH_CHECK(sc.graph_next < kGraphFirst + kGraphSlots, "libigs_raster: too many rasterizations recorded into CUDA graphs");
        slot = sc.graph_next++;
      }
      a.num_rendered_pinned = pinned + slot;
      a.defer_count = 1;
      const int64_t serial = ++sc.grad_serial;
      if (mode == kBackward) sc.grad_slot_serial[slot - kGradFirst].store(serial);
      const int st = igs_forward(&a, stream, nullptr);
      if (st != IGS_OK) {
        if (mode == kDeferred) sc.defer_in_use--;
        fail(st, "forward");
      }
      sc.ctrl_dirty = false;
      if (mode == kDeferred) {
        if (!t_pending) t_pending = new std::deque<std::pair<int64_t, Pending>>();
        t_pending->push_back({t_serial++, Pending{a.num_rendered_pinned, capacity, &hint->second, &sc, stream}});
        return {color, depth, norm, alpha, radii, extra};
      }
      if (mode == kGraph) {
        std::lock_guard<std::mutex> lock(g_captured_mutex);
        g_captured->push_back(Captured{a.num_rendered_pinned, capacity, &hint->second});
      } else {
        sc.last_slot = a.num_rendered_pinned, sc.last_capacity = capacity, sc.last_serial = serial, sc.last_hint = &hint->second;
      }
      ctx->saved_data["num_rendered"] = (int64_t)-1;  // not known here
      ctx->saved_data["count_slot"] = mode == kGraph ? (int64_t)0 : (int64_t)(intptr_t)a.num_rendered_pinned;
      ctx->saved_data["count_serial_ptr"] = mode == kGraph ? (int64_t)0 : (int64_t)(intptr_t)&sc.grad_slot_serial[slot - kGradFirst];
      ctx->saved_data["count_serial"] = serial;
      ctx->saved_data["count_capacity"] = capacity;
      ctx->saved_data["count_hint"] = (int64_t)(intptr_t)&hint->second;
    } else {
      int64_t num_rendered = 0;
      int st = igs_forward(&a, stream, &num_rendered);
      if (st == IGS_ERR_CAPACITY) {
        capacity = (int64_t)(num_rendered * 1.1) + 4096;
        binb = at::empty({(int64_t)igs_bin_bytes(capacity)}, u8);
        Tensor& sort2 = sc.buf(sc.sort, no_dirt, igs_sort_bytes(capacity), dev, false);
        a.bin = binb.data_ptr(), a.sort = sort2.data_ptr(), a.capacity = capacity;
        st = igs_forward_bin_render(&a, stream);
      }
      if (st != IGS_OK) fail(st, "forward");
      sc.ctrl_dirty = false;
      sc.capacity_hint[key].raise(num_rendered);
      ctx->saved_data["num_rendered"] = num_rendered;
      ctx->saved_data["count_slot"] = (int64_t)0;
    }
    const int64_t t_launch = now_ns();
    ctx->saved_data["tanfovx"] = tanfovx;
    ctx->saved_data["tanfovy"] = tanfovy;
    ctx->saved_data["scale_modifier"] = scale_modifier;
    ctx->saved_data["sh_degree"] = sh_degree;
    ctx->saved_data["fused_norm"] = fuse_normalize;
    ctx->set_materialize_grads(false);
    ctx->mark_non_differentiable({radii});
    ctx->save_for_backward({colors_precomp, means3D, scales, rotations, cov3Ds_precomp, norm3Ds_precomp, radii,
                            extra_attrs, sh, state, binb, alpha, bg, viewmatrix, projmatrix, campos,
                            fuse_normalize ? norm : Tensor()});
    const int64_t t_exit = now_ns();
    g_trace.ns[0] += t_alloc - t_entry, g_trace.ns[1] += t_args - t_alloc, g_trace.ns[2] += t_launch - t_args;
    g_trace.ns[3] += t_exit - t_launch, g_trace.calls[0]++;
    return {color, depth, norm, alpha, radii, extra};
  }

  static variable_list backward(AutogradContext* ctx, variable_list grad_out) {
    const int64_t t_entry = now_ns();
    const auto saved = ctx->get_saved_variables();
    const Tensor &colors_precomp = saved[0], &means3D = saved[1], &scales = saved[2], &rotations = saved[3],
                 &cov3Ds_precomp = saved[4], &norm3Ds_precomp = saved[5], &radii = saved[6], &extra_attrs = saved[7],
                 &sh = saved[8], &state = saved[9], &binb = saved[10], &alpha = saved[11], &bg = saved[12],
                 &viewmatrix = saved[13], &projmatrix = saved[14], &campos = saved[15], &norm_out = saved[16];
    const bool fused_norm = ctx->saved_data["fused_norm"].toBool();
    const at::Device dev = means3D.device();
    c10::cuda::CUDAGuard guard(dev);
    cudaStream_t stream = c10::cuda::getCurrentCUDAStream(dev.index()).stream();
    Scratch& sc = scratch(dev.index(), stream);
    const int64_t P = means3D.size(0), H = alpha.size(1), W = alpha.size(2);
    const int64_t ED = extra_attrs.numel() ? extra_attrs.size(1) : 0;
    const int64_t M = sh.numel() ? sh.size(1) : 0;
    auto f32 = at::TensorOptions().dtype(at::kFloat).device(dev);

    // The forward did not wait for its instance count: it is read here (posted early in the
    // forward, so it has long landed) and held against the capacity the forward allocated.
    if (const int64_t slot_i = ctx->saved_data["count_slot"].toInt()) {
      volatile int32_t* slot = (volatile int32_t*)(intptr_t)slot_i;
      auto* serial_ptr = (std::atomic<int64_t>*)(intptr_t)ctx->saved_data["count_serial_ptr"].toInt();
      Hint* h = (Hint*)(intptr_t)ctx->saved_data["count_hint"].toInt();
      const int64_t cap = ctx->saved_data["count_capacity"].toInt();
      int64_t n = -1;
      if (serial_ptr->load() == ctx->saved_data["count_serial"].toInt()) n = wait_count(slot, stream);
      if (n < 0 || serial_ptr->load() != ctx->saved_data["count_serial"].toInt()) {
        // the slot went to a later call (128+ forwards since) or nothing was posted: the count the
        // device kept with the forward's state decides
        int32_t v = 0;
        const cudaError_t e = cudaMemcpyAsync(&v, state.data_ptr<uint8_t>() + igs_geom_bytes((int32_t)P) + igs_img_info_offset((int32_t)W, (int32_t)H),
                                              sizeof v, cudaMemcpyDeviceToHost, stream);
        TORCH_CHECK(e == cudaSuccess && cudaStreamSynchronize(stream) == cudaSuccess, "libigs_raster: cannot read num_rendered");
        n = (int64_t)(uint32_t)v;
      }
      h->raise(n);
      TORCH_CHECK(n <= cap, "libigs_raster: this rasterization produced ", n, " instances but its workspaces held ", cap,
                  " (sized from earlier frames): its images were not rendered and there is nothing to differentiate. "
                  "The capacity has been raised; IGS_SYNC_COUNTS=1 makes every call wait for its count instead.");
    }

    const int64_t t_count = now_ns();
    const GradLayout lay = gradient_layout(P, M, ED);
    Tensor flat;
    if (g_have_grad_alloc) {
      py::gil_scoped_acquire gil;
      std::lock_guard<std::mutex> lock(g_grad_alloc_mutex);
      if (g_have_grad_alloc)
        flat = (*g_grad_alloc)(lay.total, py::cast(dev)).cast<Tensor>();
    }
    if (!flat.defined()) flat = at::empty({lay.total}, f32);
    // the kernel writes all ten blocks; tensor views are only made for the gradients autograd wants
    float* const base = lay.total ? flat.data_ptr<float>() : nullptr;
    auto gptr = [&](int i) -> float* { return (base && P * lay.width[i]) ? base + lay.start[i] : nullptr; };
    Tensor& accum = sc.buf(sc.accum, sc.accum_dirty, igs_accum_bytes((int32_t)P, (int32_t)ED), dev, true);

    auto grad = [&](const Tensor& t, int64_t C) { return t.defined() ? f32c(t) : sc.zero_img(C, H, W, dev); };
    Tensor k_bg = f32c(bg), k_view = f32c(viewmatrix), k_proj = f32c(projmatrix), k_cam = f32c(campos),
           k_means = f32c(means3D), k_sh = f32c(sh), k_col = f32c(colors_precomp), k_sc = f32c(scales),
           k_rot = f32c(rotations), k_cov = f32c(cov3Ds_precomp), k_n3 = f32c(norm3Ds_precomp),
           k_ex = f32c(extra_attrs), k_alpha = f32c(alpha);
    Tensor d_color = grad(grad_out[0], 3), d_depth = grad(grad_out[1], 1), d_norm = grad(grad_out[2], 3),
           d_alpha = grad(grad_out[3], 1), d_extra = ED > 0 ? grad(grad_out[5], ED) : Tensor();
    Tensor k_norm = fused_norm ? f32c(norm_out) : Tensor();

    IgsBwdArgs a = {};
    a.P = (int32_t)P, a.D = (int32_t)ctx->saved_data["sh_degree"].toInt(), a.M = (int32_t)M, a.ED = (int32_t)ED;
    a.W = (int32_t)W, a.H = (int32_t)H;
    a.tan_fovx = (float)ctx->saved_data["tanfovx"].toDouble(), a.tan_fovy = (float)ctx->saved_data["tanfovy"].toDouble();
    a.scale_modifier = (float)ctx->saved_data["scale_modifier"].toDouble();
    a.debug = 0, a.normalize_norm = fused_norm;
    a.num_rendered = ctx->saved_data["num_rendered"].toInt();
    a.bg = fptr(k_bg), a.viewmatrix = fptr(k_view), a.projmatrix = fptr(k_proj), a.campos = fptr(k_cam);
    a.means3D = fptr(k_means), a.shs = fptr(k_sh), a.colors_precomp = fptr(k_col), a.scales = fptr(k_sc);
    a.rotations = fptr(k_rot), a.cov3Ds_precomp = fptr(k_cov), a.norm3Ds_precomp = fptr(k_n3), a.extra_attrs = fptr(k_ex);
    a.radii = radii.numel() ? radii.data_ptr<int32_t>() : nullptr;
    const uint8_t* st_ptr = state.data_ptr<uint8_t>();
    a.geom = st_ptr, a.img = st_ptr + igs_geom_bytes((int32_t)P), a.bin = binb.data_ptr();
    a.out_alpha = fptr(k_alpha), a.out_norm = fused_norm ? fptr(k_norm) : nullptr;
    a.dL_dcolor = fptr(d_color), a.dL_ddepth = fptr(d_depth), a.dL_dnorm = fptr(d_norm), a.dL_dalpha = fptr(d_alpha);
    a.dL_dextra = ED > 0 ? fptr(d_extra) : nullptr;
    a.accum = accum.data_ptr();
    a.accum_clean = 1;
    sc.accum_dirty = true;
    a.dL_dmeans2D = gptr(0), a.dL_dcolors = gptr(1), a.dL_dopacity = gptr(2);
    a.dL_dmeans3D = gptr(3), a.dL_dcov3D = gptr(4), a.dL_dnorm3D = gptr(5);
    a.dL_dsh = gptr(6), a.dL_dscales = gptr(7), a.dL_drotations = gptr(8);
    a.dL_dextra_attrs = gptr(9);
    const int64_t t_args = now_ns();
    int st = igs_backward(&a, stream);
    if (st != IGS_OK) fail(st, "backward");
    sc.accum_dirty = false;
    const int64_t t_launch = now_ns();

    // forward's argument order: means3D, means2D, sh, colours, opacities, scales, rotations, cov3D,
    // norm3D, extra, then 12 non-differentiable settings.  Gradients nobody asked for (including
    // those of the empty 'None' sentinels) are dropped, as autograd does with the reference's.
    static const int src[10] = {3, 0, 6, 1, 2, 7, 8, 4, 5, 9};
    variable_list out(23);
    for (int i = 0; i < 10; ++i) {
      if (!ctx->needs_input_grad(i)) continue;
      const int k = src[i];
      if (k == 9 && ED == 0) out[i] = at::empty(extra_attrs.sizes(), f32);  // no elements; the shape autograd expects
      else if (k == 6) out[i] = flat.as_strided({P, M, 3}, {3 * M, 3, 1}, lay.start[k]);
      else out[i] = flat.as_strided({P, lay.width[k]}, {lay.width[k], 1}, lay.start[k]);
    }
    const int64_t t_exit = now_ns();
    g_trace.ns[4] += t_count - t_entry, g_trace.ns[5] += t_args - t_count, g_trace.ns[6] += t_launch - t_args;
    g_trace.ns[7] += t_exit - t_launch, g_trace.calls[1]++;
    return out;
  }
};

std::vector<Tensor> rasterize(const Tensor& means3D, const Tensor& means2D, const Tensor& sh, const Tensor& colors_precomp,
                              const Tensor& opacities, const Tensor& scales, const Tensor& rotations,
                              const Tensor& cov3Ds_precomp, const Tensor& norm3Ds_precomp, const Tensor& extra_attrs,
                              const Tensor& bg, const Tensor& viewmatrix, const Tensor& projmatrix, const Tensor& campos,
                              double tanfovx, double tanfovy, double scale_modifier, int64_t H, int64_t W,
                              int64_t sh_degree, bool prefiltered, bool fuse_normalize) {
  // will this call get a backward?  (asked here: inside Function::forward grad mode is always off)
  const bool wants_grad = at::GradMode::is_enabled() &&
                          (means3D.requires_grad() || sh.requires_grad() || colors_precomp.requires_grad() ||
                           opacities.requires_grad() || scales.requires_grad() || rotations.requires_grad() ||
                           cov3Ds_precomp.requires_grad() || norm3Ds_precomp.requires_grad() ||
                           extra_attrs.requires_grad() || means2D.requires_grad());
  return RasterizeFn::apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                            norm3Ds_precomp, extra_attrs, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                            scale_modifier, H, W, sh_degree, prefiltered, fuse_normalize, wants_grad);
}

bool set_deferred_counts(bool on) {
  const bool was = t_defer;
  t_defer = on;
  return was;
}
int64_t deferred_mark() { return t_serial; }

// Reads the counts of the deferred calls with serial in [.., last) that have not been read yet
// (waiting for any that has not been posted: the caller asks after the frame has been consumed,
// so there is normally nothing to wait for) and tells whether one with serial >= first exceeded
// its capacity.  Raises the capacity hints from the counts, as a waiting call would have.
bool deferred_overflowed(int64_t first, int64_t last) {
  bool over = false;
  while (t_pending && !t_pending->empty() && t_pending->front().first < last) {
    const int64_t serial = t_pending->front().first;
    const Pending p = t_pending->front().second;
    t_pending->pop_front();
    const int64_t n = wait_count(p.slot, p.stream);
    if (n < 0) p.owner->defer_in_use--;
    TORCH_CHECK(n >= 0, "libigs_raster: a deferred num_rendered was never posted (the frame's kernels failed?)");
    p.owner->defer_in_use--;
    p.hint->raise(n);
    if (serial >= first && n > p.capacity) over = true;
  }
  return over;
}

// After a replay of a CUDA graph that holds rasterizations (and a synchronize): did one of them
// produce more instances than the workspaces recorded into the graph hold?  Raises the capacity
// hints from the counts, so that a graph recorded again has room.
bool captured_overflowed() {
  std::lock_guard<std::mutex> lock(g_captured_mutex);
  bool over = false;
  for (const Captured& c : *g_captured) {
    const int32_t v = *c.slot;
    if (v == -1) continue;  // not replayed yet
    const int64_t n = (int64_t)(uint32_t)v;
    c.hint->raise(n);
    over = over || n > c.capacity;
  }
  return over;
}

// Forget what this thread has learnt about instance counts: the next call of every size waits for
// its count again and sizes the workspaces from it (tests; a scene whose counts have collapsed).
void reset_capacity_hints() {
  for (auto& kv : all_scratch())
    for (auto& h : kv.second.capacity_hint) h.second.capacity.store(0, std::memory_order_relaxed);
}

void set_gradient_allocator(py::object fn) {
  std::lock_guard<std::mutex> lock(g_grad_alloc_mutex);
  g_have_grad_alloc = !fn.is_none();
  *g_grad_alloc = g_have_grad_alloc ? fn : py::object();
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "autograd node of the libigs_raster rasterizer (C++ host side over the C ABI)";
  m.def("rasterize", &rasterize, py::call_guard<py::gil_scoped_release>(),
        "forward of the rasterizer as one autograd node: (color, depth, norm, alpha, radii, extra)");
  m.def("set_gradient_allocator", &set_gradient_allocator, "fn(n_floats, device) -> flat float32 tensor, or None");
  m.def("set_deferred_counts", &set_deferred_counts,
        "this thread's forward-only calls stop waiting for num_rendered (True) or wait again (False); returns the old setting");
  m.def("deferred_mark", &deferred_mark, "serial number the next deferred call of this thread will get");
  m.def("deferred_overflowed", &deferred_overflowed, py::call_guard<py::gil_scoped_release>(),
        "reads the counts of the deferred calls before serial `last`; True if one at or after `first` exceeded its capacity");
  m.def("host_trace", []() {
    // mean host microseconds per call since the last read: forward {allocations, argument block, igs_forward,
    // autograd bookkeeping}, backward {saved tensors + count, allocations + argument block, igs_backward, views}
    std::vector<double> out(12, 0.0);
    for (int i = 0; i < 4; i++) out[8 + i] = (double)g_trace.modes[i].exchange(0);
    const int64_t nf = std::max<int64_t>(g_trace.calls[0].exchange(0), 1), nb = std::max<int64_t>(g_trace.calls[1].exchange(0), 1);
    for (int i = 0; i < 8; i++) out[i] = g_trace.ns[i].exchange(0) / 1e3 / (i < 4 ? nf : nb);
    return out;
  }, "mean host us per phase of the node's forward [0:4] and backward [4:8] since the last call");
  m.def("reset_capacity_hints", &reset_capacity_hints,
        "this thread's next call of every size waits for its count again and sizes the workspaces from it");
  m.def("captured_overflowed", &captured_overflowed,
        "after replaying a CUDA graph that holds rasterizations: True if one of them outgrew its recorded workspaces");
  m.def("library_version", []() { return std::string(igs_version()); });
}
