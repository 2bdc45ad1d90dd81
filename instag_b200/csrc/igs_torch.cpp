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

#include "igs_raster.h"

namespace {

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

// Reusable buffers of one (device, stream, host thread); see rasterizer.py::_DeviceScratch.
struct Scratch {
  Tensor sort, ctrl, accum, pinned;
  bool ctrl_dirty = false, accum_dirty = false;
  std::tuple<int64_t, int64_t, int64_t> ctrl_key{-1, -1, -1};
  std::map<std::tuple<int64_t, int64_t, int64_t>, int64_t> capacity_hint;
  std::map<std::tuple<int64_t, int64_t, int64_t>, Tensor> zero_imgs;
  int defer_next = 1, defer_in_use = 0;  // ring of count slots for calls that do not wait (see Pending)

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

// Slot 0 of `pinned` receives the count of a call that waits for it; slots 1.. are handed out in
// turn to calls that do not (deferred counts, below).
constexpr int kDeferSlots = 255;

Scratch& scratch(int device, cudaStream_t stream) {
  // never destroyed: the tensors must not outlive the CUDA context at thread / process exit
  thread_local auto* all = new std::map<std::pair<int, cudaStream_t>, Scratch>();
  Scratch& s = (*all)[{device, stream}];
  if (!s.pinned.defined()) s.pinned = at::zeros({1 + kDeferSlots}, at::kInt).pin_memory();
  return s;
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
  int64_t* hint;  // the scratch's capacity hint for this (P, H, W): raised from the count when it is read
  Scratch* owner; // its ring of slots gets this one back
};
thread_local bool t_defer = false;
thread_local int64_t t_serial = 0;                              // deferred calls so far on this thread
thread_local std::deque<std::pair<int64_t, Pending>>* t_pending = nullptr;  // (serial, entry), oldest first

inline Tensor f32c(const Tensor& t) {
  Tensor r = t.scalar_type() == at::kFloat ? t : t.to(at::kFloat);
  return r.is_contiguous() ? r : r.contiguous();
}
inline const float* fptr(const Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }
inline float* fptr_mut(Tensor& t) { return t.numel() ? t.data_ptr<float>() : nullptr; }

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
                               bool fuse_normalize) {
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

    Tensor color = at::empty({3, H, W}, f32), depth = at::empty({1, H, W}, f32), norm = at::empty({3, H, W}, f32),
           alpha = at::empty({1, H, W}, f32);
    Tensor extra = ED > 0 ? at::empty({ED, H, W}, f32) : at::empty({0}, f32);
    Tensor radii = at::empty({P}, at::TensorOptions().dtype(at::kInt).device(dev));

    const auto key = std::make_tuple(P, H, W);
    const size_t geom_bytes = igs_geom_bytes((int32_t)P), img_bytes = igs_img_bytes((int32_t)W, (int32_t)H);
    auto hint = sc.capacity_hint.find(key);
    int64_t capacity = hint != sc.capacity_hint.end() ? hint->second : std::max<int64_t>(8 * P, 1 << 16);
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

    // no graph is being recorded, the sizes have been seen before, a slot is free: do not wait for the count
    const bool defer = t_defer && P > 0 && hint != sc.capacity_hint.end() && sc.defer_in_use < kDeferSlots &&
                       !(at::GradMode::is_enabled() &&
                         (means3D.requires_grad() || sh.requires_grad() || colors_precomp.requires_grad() ||
                          opacities.requires_grad() || scales.requires_grad() || rotations.requires_grad() ||
                          cov3Ds_precomp.requires_grad() || norm3Ds_precomp.requires_grad() ||
                          extra_attrs.requires_grad() || means2D.requires_grad()));
    if (defer) {
      const int slot = sc.defer_next;
      sc.defer_next = slot % kDeferSlots + 1;
      sc.defer_in_use++;
      a.num_rendered_pinned = sc.pinned.data_ptr<int32_t>() + slot;
      a.defer_count = 1;
      const int st = igs_forward(&a, stream, nullptr);
      if (st != IGS_OK && st != IGS_ERR_CAPACITY) {  // (a count already known to be too large is found at harvest)
        sc.defer_in_use--;
        fail(st, "forward");
      }
      sc.ctrl_dirty = false;
      if (!t_pending) t_pending = new std::deque<std::pair<int64_t, Pending>>();
      t_pending->push_back({t_serial++, Pending{a.num_rendered_pinned, capacity, &hint->second, &sc}});
      return {color, depth, norm, alpha, radii, extra};
    }

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
    sc.capacity_hint[key] = std::max<int64_t>((int64_t)(num_rendered * 1.25) + 4096, 1 << 14);

    ctx->saved_data["num_rendered"] = num_rendered;
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
    return {color, depth, norm, alpha, radii, extra};
  }

  static variable_list backward(AutogradContext* ctx, variable_list grad_out) {
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
    int st = igs_backward(&a, stream);
    if (st != IGS_OK) fail(st, "backward");
    sc.accum_dirty = false;

    // forward's argument order: means3D, means2D, sh, colours, opacities, scales, rotations, cov3D,
    // norm3D, extra, then 12 non-differentiable settings.  Gradients nobody asked for (including
    // those of the empty 'None' sentinels) are dropped, as autograd does with the reference's.
    static const int src[10] = {3, 0, 6, 1, 2, 7, 8, 4, 5, 9};
    variable_list out(22);
    for (int i = 0; i < 10; ++i) {
      if (!ctx->needs_input_grad(i)) continue;
      const int k = src[i];
      if (k == 9 && ED == 0) out[i] = at::empty(extra_attrs.sizes(), f32);  // no elements; the shape autograd expects
      else if (k == 6) out[i] = flat.as_strided({P, M, 3}, {3 * M, 3, 1}, lay.start[k]);
      else out[i] = flat.as_strided({P, lay.width[k]}, {lay.width[k], 1}, lay.start[k]);
    }
    return out;
  }
};

std::vector<Tensor> rasterize(const Tensor& means3D, const Tensor& means2D, const Tensor& sh, const Tensor& colors_precomp,
                              const Tensor& opacities, const Tensor& scales, const Tensor& rotations,
                              const Tensor& cov3Ds_precomp, const Tensor& norm3Ds_precomp, const Tensor& extra_attrs,
                              const Tensor& bg, const Tensor& viewmatrix, const Tensor& projmatrix, const Tensor& campos,
                              double tanfovx, double tanfovy, double scale_modifier, int64_t H, int64_t W,
                              int64_t sh_degree, bool prefiltered, bool fuse_normalize) {
  return RasterizeFn::apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                            norm3Ds_precomp, extra_attrs, bg, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                            scale_modifier, H, W, sh_degree, prefiltered, fuse_normalize);
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
    for (uint64_t spin = 1; *p.slot == -1; ++spin) {
      TORCH_CHECK(spin < (1ull << 33), "libigs_raster: a deferred num_rendered was never posted");
#if defined(__x86_64__) || defined(__i386__)
      __builtin_ia32_pause();
#endif
    }
    const int64_t n = (int64_t)(uint32_t)*p.slot;
    p.owner->defer_in_use--;
    *p.hint = std::max<int64_t>(*p.hint, (int64_t)(n * 1.25) + 4096);
    if (serial >= first && n > p.capacity) over = true;
  }
  return over;
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
  m.def("library_version", []() { return std::string(igs_version()); });
}
