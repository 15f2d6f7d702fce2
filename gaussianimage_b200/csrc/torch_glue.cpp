// Host glue over the C ABI (include/gi2d.h): the operator layer's host side in C++, so that the eager path (what an unchanged
// caller of the reference gets) spends its host time in torch's own dispatch, not in Python wrappers, ctypes marshalling and
// a dozen Python-level torch.empty calls per step.  It replaces, like cuda.py / rasterize_sum.py, the reference's
// bindings.cu host wrappers (bindings.cu:400-469, 706-779, 781-1004) and the Python orchestration of
// gsplat/gsplat/rasterize_sum.py:92-185; the kernels, the ABI and the results are the same ones -- this file contains no
// device code and no arithmetic.  Two layers:
//   * plain "check inputs, allocate outputs, launch" calls (project_fwd, fused_forward, raster_bwd_packed, ...) used by the
//     Python autograd Functions (GI2D_CPP_AUTOGRAD=0);
//   * the two operators as autograd nodes (ProjectNode, RasterNode) together with the whole policy of the fused
//     chain -- which tensors are known to come from one of our projections, workspace sizing, memoised binning, overflow
//     bookkeeping -- so that one operator call is one Python -> C++ transition (default when this module is built).
// When the module is absent the ctypes path of cuda.py runs (GI2D_HOST_GLUE=0 forces it).
#include <torch/extension.h>
#include <torch/csrc/autograd/function.h>
#include <torch/csrc/autograd/functions/utils.h>
#include <torch/csrc/autograd/saved_variable.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>
#include <cuda_runtime_api.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "gi2d.h"

namespace {

using at::Tensor;
using OptTensor = c10::optional<Tensor>;

inline void chk(int rc, const char *what) {
  TORCH_CHECK(rc == 0, what, " failed: ", gi2d_status_string(rc), " (status ", rc, ")");
}

inline void chk_in(const Tensor &x, const char *name, at::ScalarType want) {   // bindings.h:9-14 (CHECK_INPUT)
  TORCH_CHECK(x.is_cuda(), name, " must be a CUDA tensor");
  TORCH_CHECK(x.is_contiguous(), name, " must be contiguous");
  TORCH_CHECK(x.scalar_type() == want, name, " must be ", want, ", got ", x.scalar_type());
}

inline cudaStream_t stream_of(const Tensor &x) { return c10::cuda::getCurrentCUDAStream(x.get_device()).stream(); }
inline const float *fp(const Tensor &x) { return x.data_ptr<float>(); }
inline const int *ip(const Tensor &x) { return x.data_ptr<int>(); }
inline int64_t align256(int64_t v) { return (v + 255) / 256 * 256; }
inline bool capturing(cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return false; }
  return cs != cudaStreamCaptureStatusNone;
}

// ---------------------------------------------------------------------------------------------------------------------
// Layer 1: allocation + launch sequences (no autograd, no policy)
// ---------------------------------------------------------------------------------------------------------------------

// cuda.py: project_gaussians_2d_forward / project_gaussians_2d_scale_rot_forward.  mode 0 = Cholesky (p0 = L (N,3)),
// mode 1 = scale-rotation (p0 = scales (N,2), p1 = rotation (N,1)).  total_host: address of a mapped pinned int or 0.
std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> project_fwd_impl(int64_t mode, int64_t N, const Tensor &means2d, const Tensor &p0,
                                                                    const OptTensor &p1, int64_t H, int64_t W, int64_t TX, int64_t TY,
                                                                    int *total_host) {
  chk_in(means2d, "means2d", at::kFloat);
  chk_in(p0, mode == 0 ? "L_elements" : "scales2d", at::kFloat);
  if (mode != 0) {
    TORCH_CHECK(p1.has_value(), "rotation is required");
    chk_in(*p1, "rotation", at::kFloat);
  }
  c10::cuda::CUDAGuard guard(means2d.device());
  auto f = means2d.options().requires_grad(false);
  auto i = f.dtype(at::kInt);
  Tensor xys = at::empty({N, 2}, f), depths = at::empty({N}, f), radii = at::empty({N}, i);
  Tensor conics = at::empty({N, 3}, f), nth = at::empty({N}, i);
  cudaStream_t st = stream_of(means2d);
  chk(gi2d_project_fwd_total((int)mode, (int)N, fp(means2d), fp(p0), mode != 0 ? fp(*p1) : nullptr, (int)H, (int)W, (int)TX, (int)TY,
                             xys.data_ptr<float>(), depths.data_ptr<float>(), radii.data_ptr<int>(), conics.data_ptr<float>(),
                             nth.data_ptr<int>(), nullptr, total_host, st),
      mode == 0 ? "project_gaussians_2d_forward" : "project_gaussians_2d_scale_rot_forward");
  return {xys, depths, radii, conics, nth};
}

std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> project_fwd(int64_t mode, int64_t N, const Tensor &means2d, const Tensor &p0,
                                                               const OptTensor &p1, int64_t H, int64_t W, int64_t TX, int64_t TY) {
  return project_fwd_impl(mode, N, means2d, p0, p1, H, W, TX, TY, nullptr);
}

// cuda.py: project_backward_autograd, packed branch -- grads12 is the (N,12) row buffer of raster_bwd (v_xy's storage).
std::tuple<Tensor, Tensor, OptTensor> project_bwd_packed(int64_t mode, int64_t N, const Tensor &p0, const OptTensor &p1,
                                                         int64_t H, int64_t W, const Tensor &radii, const Tensor &conics,
                                                         const Tensor &v_xy) {
  chk_in(p0, mode == 0 ? "L_elements" : "scales2d", at::kFloat);
  chk_in(radii, "radii", at::kInt);
  chk_in(conics, "conics", at::kFloat);
  TORCH_CHECK(v_xy.is_cuda() && v_xy.scalar_type() == at::kFloat && v_xy.dim() == 2 && v_xy.stride(0) == 12 &&
                  v_xy.stride(1) == 1,
              "v_xy must be columns 0-1 of an (N,12) float32 gradient buffer");
  c10::cuda::CUDAGuard guard(conics.device());
  auto f = conics.options().requires_grad(false);
  Tensor v_mean2d = at::empty({N, 2}, f);
  cudaStream_t st = stream_of(conics);
  if (mode == 0) {
    Tensor v_L = at::empty({N, 3}, f);
    chk(gi2d_project_cholesky_bwd_packed((int)N, fp(p0), (int)H, (int)W, ip(radii), fp(conics), fp(v_xy),
                                         v_mean2d.data_ptr<float>(), v_L.data_ptr<float>(), st),
        "project_gaussians_2d_backward");
    return {v_mean2d, v_L, c10::nullopt};
  }
  TORCH_CHECK(p1.has_value(), "rotation is required");
  chk_in(*p1, "rotation", at::kFloat);
  Tensor v_scale = at::empty({N, 2}, f), v_rot = at::empty({N, 1}, f);
  chk(gi2d_project_scale_rot_bwd_packed((int)N, fp(p0), fp(*p1), (int)H, (int)W, ip(radii), fp(conics), fp(v_xy),
                                        v_mean2d.data_ptr<float>(), v_scale.data_ptr<float>(), v_rot.data_ptr<float>(), st),
      "project_gaussians_2d_scale_rot_backward");
  return {v_mean2d, v_scale, v_rot};
}

// Tensors of one pass of the fused chain.
struct Chain {
  Tensor out_img, final_Ts, final_idx, tile_bins, packed, scratch, grads12;
  int64_t o_gids = 0, o_status = 0, cap = 0;
};

// rasterize_sum.py: fused_chain without a custom rasterizer -- one scratch allocation [workspace | gids_sorted | status],
// tile_bins (T,2), packed (cap*12), the image and (optionally) final_Ts / final_idx; gi2d_bin_fused_ex followed by
// gi2d_raster_sum_fwd_packed on the current stream.  host_status = address of a pinned {n_isect, overflow} slot or null;
// sticky = device {max n, overflow} pair or null; want_grads: also allocate the (N,12) gradient rows and have the chain zero them.
Chain run_chain(const Tensor &xys, const Tensor &radii, const Tensor &conics, const Tensor &nth, const Tensor &colors,
                const Tensor &opacity, const Tensor &background, int64_t H, int64_t W, int64_t cap, int *host_status, int *sticky,
                bool want_aux, bool want_grads) {
  chk_in(xys, "xys", at::kFloat);
  chk_in(radii, "radii", at::kInt);
  chk_in(conics, "conics", at::kFloat);
  chk_in(nth, "num_tiles_hit", at::kInt);
  chk_in(colors, "colors", at::kFloat);
  chk_in(opacity, "opacity", at::kFloat);
  chk_in(background, "background", at::kFloat);
  const int64_t N = xys.size(0);
  TORCH_CHECK(radii.numel() == N && nth.numel() == N && conics.numel() == 3 * N && colors.numel() == 3 * N && opacity.numel() == N &&
                  background.numel() == 3,
              "rasterize_gaussians_sum: inconsistent tensor sizes for ", N, " Gaussians");
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  c10::cuda::CUDAGuard guard(xys.device());
  cudaStream_t st = stream_of(xys);
  const size_t ws_bytes = gi2d_bin_fused_workspace_bytes((int)N, TX, TY, (int)cap);
  Chain c;
  c.cap = cap;
  c.o_gids = align256((int64_t)ws_bytes);
  c.o_status = c.o_gids + align256(4 * cap);
  auto f = xys.options().requires_grad(false);
  auto i = f.dtype(at::kInt);
  c.scratch = at::empty({c.o_status + 256}, f.dtype(at::kByte));
  c.tile_bins = at::empty({(int64_t)TX * TY, 2}, i);
  c.packed = at::empty({cap * 12}, f);
  if (want_grads) c.grads12 = at::empty({N, 12}, f);
  uint8_t *base = c.scratch.data_ptr<uint8_t>();
  int *d_status = reinterpret_cast<int *>(base + c.o_status);
  gi2d_bin_args a;
  a.N = (int)N; a.TX = TX; a.TY = TY; a.ty_begin = 0; a.ty_end = TY; a.capacity = (int)cap;
  a.xys = fp(xys); a.radii = ip(radii); a.num_tiles_hit = ip(nth); a.conics = fp(conics); a.colors = fp(colors); a.opacities = fp(opacity);
  a.ws = base; a.ws_bytes = ws_bytes;
  a.gids_sorted = reinterpret_cast<int *>(base + c.o_gids); a.tile_bins = c.tile_bins.data_ptr<int>(); a.packed = c.packed.data_ptr<float>();
  a.status_out = d_status; a.status_host = host_status; a.status_sticky = sticky;
  a.zero_grads12 = want_grads ? c.grads12.data_ptr<float>() : nullptr;
  chk(gi2d_bin_fused_ex(&a, st), "bin_fused");
  c.out_img = at::empty({H, W, 3}, f);
  if (want_aux) {
    c.final_Ts = at::empty({H, W}, f);
    c.final_idx = at::empty({H, W}, i);
  }
  chk(gi2d_raster_sum_fwd_packed(TX, TY, (int)W, (int)H, c.tile_bins.data_ptr<int>(), c.packed.data_ptr<float>(), d_status,
                                 fp(background), c.out_img.data_ptr<float>(), want_aux ? c.final_Ts.data_ptr<float>() : nullptr,
                                 want_aux ? c.final_idx.data_ptr<int>() : nullptr, st),
      "rasterize_sum_forward");
  return c;
}

std::tuple<Tensor, OptTensor, OptTensor, Tensor, Tensor, Tensor, int64_t, int64_t> fused_forward(
    const Tensor &xys, const Tensor &radii, const Tensor &conics, const Tensor &nth, const Tensor &colors, const Tensor &opacity,
    const Tensor &background, int64_t H, int64_t W, int64_t cap, int64_t host_status, bool want_aux, int64_t sticky) {
  Chain c = run_chain(xys, radii, conics, nth, colors, opacity, background, H, W, cap, reinterpret_cast<int *>(host_status),
                      reinterpret_cast<int *>(sticky), want_aux, false);
  OptTensor Ts, idx;
  if (want_aux) { Ts = c.final_Ts; idx = c.final_idx; }
  return {c.out_img, Ts, idx, c.tile_bins, c.packed, c.scratch, c.o_gids, c.o_status};
}

// rasterize_sum.py: _RasterizeGaussiansSum.backward, fused branch -> the (N,12) gradient rows.
Tensor raster_bwd_packed(int64_t N, int64_t H, int64_t W, const Tensor &tile_bins, const Tensor &packed, const Tensor &v_out) {
  chk_in(v_out, "v_out_img", at::kFloat);
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  c10::cuda::CUDAGuard guard(packed.device());
  Tensor grads = at::empty({N, 12}, packed.options().requires_grad(false));
  chk(gi2d_raster_sum_bwd_packed((int)N, TX, TY, (int)W, (int)H, ip(tile_bins), fp(packed), fp(v_out), grads.data_ptr<float>(),
                                 stream_of(packed)),
      "rasterize_sum_backward");
  return grads;
}

// offsets of gids_sorted / status inside the scratch tensor of fused_forward (same arithmetic, for the Python side)
std::tuple<int64_t, int64_t> fused_offsets(int64_t N, int64_t H, int64_t W, int64_t cap) {
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  const int64_t o_gids = align256((int64_t)gi2d_bin_fused_workspace_bytes((int)N, TX, TY, (int)cap));
  return {o_gids, o_gids + align256(4 * cap)};
}

// ---------------------------------------------------------------------------------------------------------------------
// Layer 2: operator state and policy
// ---------------------------------------------------------------------------------------------------------------------
using Impl = c10::TensorImpl;
using WeakImpl = c10::weak_intrusive_ptr<Impl>;
using Key = std::tuple<int, int64_t, int64_t, int64_t>;          // (device, N, TX, TY)

inline uint32_t version_of(const Tensor &t) { return t.is_inference() ? 0u : t._version(); }

// What one of our projections produced: identity (TensorImpl address, kept un-recyclable by a weak reference) and version
// counters of xys / depths / radii / num_tiles_hit at the moment of production, plus where the kernel publishes the
// intersection total.  rasterize_gaussians_sum takes the fused chain ONLY for arguments found here, unmodified: that is what
// guarantees depths == 0 (quirk Q2: the key is the tile id alone), what the binning memo is keyed on, and -- in strict mode
// -- what sizes the workspace exactly, so that there is nothing to overflow.
struct Projection {
  c10::optional<WeakImpl> w[4];
  const Impl *p[4] = {nullptr, nullptr, nullptr, nullptr};
  uint32_t v[4] = {0, 0, 0, 0};
  uint64_t id = 0;
  int slot = -1;
  uint64_t slot_gen = 0;
  int64_t n_isect = -1;          // once read
  int dev = -1;
};

struct TotalRing {               // mapped pinned host words the projection kernels publish sum(num_tiles_hit) to
  static constexpr int SLOTS = 256;
  Tensor t;
  volatile int *p = nullptr;
  uint64_t gen[SLOTS] = {0};
  int next = 0;
  void ensure() {
    if (p) return;
    t = at::empty({SLOTS}, at::TensorOptions().dtype(at::kInt)).pin_memory();
    p = t.data_ptr<int>();
  }
};

struct Memo {                    // the last binning result per (device, stream): see rasterize_sum.py::_Geometry
  uint64_t proj_id = 0;
  int dev = -1;
  cudaStream_t stream = nullptr;
  int64_t N = 0, H = 0, W = 0;
  Tensor scratch, tile_bins;
  int64_t o_gids = 0, o_status = 0, cap = 0;
  bool from_capture = false;
};

struct LastStatus { Tensor scratch; int64_t o_status = 0, cap = 0; Key key; bool valid = false; };

struct State {
  std::mutex mu;
  bool sync_free = false, memoise = true;
  std::map<Key, int64_t> capacity;
  std::map<Key, Tensor> sticky;
  std::vector<Projection> projections = std::vector<Projection>(32);
  size_t proj_next = 0;
  uint64_t proj_id = 0;
  TotalRing ring;
  std::vector<Memo> memo;
  std::map<int, LastStatus> last;
  int64_t geometry_hits = 0, fast_calls = 0, total_waits = 0, total_fallbacks = 0;
};
State &S() {
  static State *s = new State();      // leaked on purpose: tensors must not be destroyed after the CUDA context at exit
  return *s;
}

inline int64_t default_capacity(int64_t N) { return std::max<int64_t>(4 * N, 1 << 14); }

// ---------------------------------------------------------------------------------------------------------------------
// The two operators as torch::autograd::Function
// ---------------------------------------------------------------------------------------------------------------------
using torch::autograd::variable_list;

inline bool packed_pair(const Tensor &v_xy, const Tensor &v_conic) {   // cuda.py: _is_packed_pair
  return v_xy.defined() && v_conic.defined() && v_xy.is_cuda() && v_xy.scalar_type() == at::kFloat &&
         v_conic.scalar_type() == at::kFloat && v_xy.dim() == 2 && v_conic.dim() == 2 && v_xy.size(1) == 2 &&
         v_conic.size(1) == 3 && v_xy.stride(0) == 12 && v_xy.stride(1) == 1 && v_conic.stride(0) == 12 &&
         v_conic.stride(1) == 1 &&
         static_cast<const char *>(v_conic.data_ptr()) == static_cast<const char *>(v_xy.data_ptr()) + 8 &&
         reinterpret_cast<uintptr_t>(v_xy.data_ptr()) % 16 == 0;
}

// project_gaussians_2d.py / project_gaussians_2d_scale_rot.py: _ProjectGaussians2d[ScaleRot].  A hand-written autograd Node
// (what torch's generated VariableType code does) rather than torch::autograd::Function: Function::apply spends ~10 us per
// call on its generic bookkeeping (AutogradContext, output wrapping, dirty / non-differentiable sets), a third of this
// operator's host time.  Inputs with edges: means2d, p0 [, p1]; outputs with history: xys (0), depths (1), conics (2).
struct ProjectNode : public torch::autograd::Node {
  torch::autograd::SavedVariable p0, p1, conics;      // conics is an OUTPUT of this node (saved without its grad_fn: no cycle)
  Tensor radii;
  int64_t mode = 0, N = 0, H = 0, W = 0;
  std::string name() const override { return mode == 0 ? "ProjectGaussians2dBackward" : "ProjectGaussians2dScaleRotBackward"; }
  void release_variables() override {
    std::lock_guard<std::mutex> lock(mutex_);
    p0.reset_data(); p1.reset_data(); conics.reset_data();
    radii.reset();
  }
  variable_list apply(variable_list &&g) override {
    std::lock_guard<std::mutex> lock(mutex_);
    TORCH_CHECK(radii.defined(), "project_gaussians_2d: backward through the graph a second time (saved tensors were freed); "
                "specify retain_graph=True");
    const Tensor p0v = p0.unpack(), conv = conics.unpack(shared_from_this());
    Tensor p1v;
    if (mode != 0) p1v = p1.unpack();
    Tensor v_xy = g[0], v_conic = g[2];
    Tensor v_mean2d, v_a, v_b;
    if (packed_pair(v_xy, v_conic)) {
      auto r = project_bwd_packed(mode, N, p0v, mode != 0 ? OptTensor(p1v) : OptTensor(), H, W, radii, conv, v_xy);
      v_mean2d = std::get<0>(r);
      v_a = std::get<1>(r);
      if (std::get<2>(r).has_value()) v_b = *std::get<2>(r);
    } else {                                       // cuda.py: project_gaussians_2d[_scale_rot]_backward without v_cov2d
      c10::cuda::CUDAGuard guard(conv.device());
      auto f = conv.options().requires_grad(false);
      v_xy = v_xy.defined() ? v_xy.contiguous() : at::zeros({N, 2}, f);
      v_conic = v_conic.defined() ? v_conic.contiguous() : at::zeros({N, 3}, f);
      chk_in(v_xy, "v_xy", at::kFloat);
      chk_in(v_conic, "v_conic", at::kFloat);
      v_mean2d = at::empty({N, 2}, f);
      cudaStream_t st = stream_of(conv);
      if (mode == 0) {
        v_a = at::empty({N, 3}, f);
        chk(gi2d_project_cholesky_bwd((int)N, fp(p0v), (int)H, (int)W, ip(radii), fp(conv), fp(v_xy), fp(v_conic), nullptr,
                                      v_mean2d.data_ptr<float>(), v_a.data_ptr<float>(), st),
            "project_gaussians_2d_backward");
      } else {
        v_a = at::empty({N, 2}, f);
        v_b = at::empty({N, 1}, f);
        chk(gi2d_project_scale_rot_bwd((int)N, fp(p0v), fp(p1v), (int)H, (int)W, ip(radii), fp(conv), fp(v_xy), fp(v_conic),
                                       nullptr, v_mean2d.data_ptr<float>(), v_a.data_ptr<float>(), v_b.data_ptr<float>(), st),
            "project_gaussians_2d_scale_rot_backward");
      }
    }
    if (mode != 0) return {v_mean2d, v_a, v_b};
    return {v_mean2d, v_a};
  }
};

variable_list project_autograd(int64_t mode, const Tensor &means2d, const Tensor &p0, const OptTensor &p1, int64_t H, int64_t W,
                               int64_t TX, int64_t TY) {
  State &s = S();
  int slot = -1;
  uint64_t gen = 0;
  int64_t host = 0;
  if (means2d.is_cuda() && !capturing(stream_of(means2d))) {     // the total is only ever read outside a capture
    std::lock_guard<std::mutex> g(s.mu);
    s.ring.ensure();
    slot = s.ring.next;
    s.ring.next = (slot + 1) % TotalRing::SLOTS;
    gen = ++s.ring.gen[slot];
    s.ring.p[slot] = -1;
    host = reinterpret_cast<int64_t>(const_cast<int *>(s.ring.p + slot));
  }
  const int64_t N = means2d.size(-2);
  std::shared_ptr<ProjectNode> node;
  if (mode != 0 ? torch::autograd::compute_requires_grad(means2d, p0, *p1) : torch::autograd::compute_requires_grad(means2d, p0)) {
    node = std::shared_ptr<ProjectNode>(new ProjectNode(), torch::autograd::deleteNode);
    node->set_next_edges(mode != 0 ? torch::autograd::collect_next_edges(means2d, p0, *p1)
                                   : torch::autograd::collect_next_edges(means2d, p0));
  }
  variable_list out;
  {
    at::AutoDispatchBelowADInplaceOrView below_autograd;
    auto r = project_fwd_impl(mode, N, means2d, p0, p1, H, W, TX, TY, reinterpret_cast<int *>(host));
    out = {std::get<0>(r), std::get<1>(r), std::get<2>(r), std::get<3>(r), std::get<4>(r)};   // xys, depths, radii, conics, num_tiles_hit
  }
  if (node) {
    node->mode = mode; node->N = N; node->H = H; node->W = W;
    node->p0 = torch::autograd::SavedVariable(p0, false);
    if (mode != 0) node->p1 = torch::autograd::SavedVariable(*p1, false);
    node->radii = out[2];
    torch::autograd::set_history(out[0], node);
    torch::autograd::set_history(out[1], node);
    torch::autograd::set_history(out[3], node);
    node->conics = torch::autograd::SavedVariable(out[3], true);
  }
  {
    std::lock_guard<std::mutex> g(s.mu);
    Projection &e = s.projections[s.proj_next];
    s.proj_next = (s.proj_next + 1) % s.projections.size();
    const Tensor *ts[4] = {&out[0], &out[1], &out[2], &out[4]};     // xys, depths, radii, num_tiles_hit
    for (int k = 0; k < 4; ++k) {
      e.w[k].emplace(ts[k]->getIntrusivePtr());
      e.p[k] = ts[k]->unsafeGetTensorImpl();
      e.v[k] = version_of(*ts[k]);
    }
    e.id = ++s.proj_id;
    e.slot = slot;
    e.slot_gen = gen;
    e.n_isect = -1;
    e.dev = means2d.get_device();
  }
  return out;
}

// Index into State::projections of the projection that produced exactly these four tensors, still unmodified; -1 otherwise.
int find_projection(State &s, const Tensor &xys, const Tensor &depths, const Tensor &radii, const Tensor &nth) {
  const Tensor *ts[4] = {&xys, &depths, &radii, &nth};
  const size_t n = s.projections.size();
  for (size_t back = 1; back <= n; ++back) {                       // newest first: the usual hit is the first probe
    const size_t i = (s.proj_next + n - back) % n;
    const Projection &e = s.projections[i];
    if (e.id == 0 || e.p[0] != xys.unsafeGetTensorImpl()) continue;
    bool ok = true;
    for (int k = 0; k < 4 && ok; ++k)
      ok = e.p[k] == ts[k]->unsafeGetTensorImpl() && e.w[k].has_value() && !e.w[k]->expired() && e.v[k] == version_of(*ts[k]);
    if (ok) return (int)i;
  }
  return -1;
}

// rasterize_sum.py: _RasterizeGaussiansSum, fused branch, as a hand-written autograd Node (see ProjectNode).  Inputs with
// edges: xys, conics, colors, opacity; output with history: out_img.
struct RasterNode : public torch::autograd::Node {
  Tensor tile_bins, packed, grads12;             // internal tensors of the forward's chain (no autograd history of their own)
  int64_t N = 0, H = 0, W = 0;
  std::vector<int64_t> opacity_shape;
  bool fresh_grads = false;                      // grads12 was zero-filled by the forward's binning chain and not used yet
  std::string name() const override { return "RasterizeGaussiansSumBackward"; }
  void release_variables() override {
    std::lock_guard<std::mutex> lock(mutex_);
    tile_bins.reset(); packed.reset(); grads12.reset();
  }
  variable_list apply(variable_list &&g) override {
    std::lock_guard<std::mutex> lock(mutex_);
    variable_list out(4);
    if (!g[0].defined()) return out;              // only out_alpha was used downstream (the reference ignores v_out_alpha, Q7)
    TORCH_CHECK(packed.defined(), "rasterize_gaussians_sum: backward through the graph a second time (saved tensors were "
                "freed); specify retain_graph=True");
    Tensor v_out = g[0].contiguous();
    Tensor grads;
    if (fresh_grads && grads12.defined()) {
      // the rows were zero-filled by the forward's binning chain: no memset in front of the kernel.  One use only -- a
      // second backward through the same node (retain_graph) takes the allocating path below.
      fresh_grads = false;
      grads = grads12;
      grads12.reset();
      chk_in(v_out, "v_out_img", at::kFloat);
      const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
      c10::cuda::CUDAGuard guard(grads.device());
      chk(gi2d_raster_sum_bwd_packed_acc((int)N, TX, TY, (int)W, (int)H, ip(tile_bins), fp(packed), fp(v_out), nullptr,
                                         grads.data_ptr<float>(), stream_of(grads)),
          "rasterize_sum_backward");
    } else {
      grads = raster_bwd_packed(N, H, W, tile_bins, packed, v_out);
    }
    out[0] = grads.slice(1, 0, 2);                                    // v_xy
    out[1] = grads.slice(1, 2, 5);                                    // v_conic
    out[2] = grads.slice(1, 5, 8);                                    // v_colors
    out[3] = grads.slice(1, 8, 9).reshape(opacity_shape);             // v_opacity
    return out;
  }
};

// One forward pass.  memo != null: re-rasterize a memoised binning with other conics / colours / opacities
// (gi2d_repack_records_ex) instead of running the chain.  Returns {out_img, final_Ts (or an undefined tensor)}.
variable_list raster_apply(const Tensor &xys, const Tensor &radii, const Tensor &conics, const Tensor &nth, const Tensor &colors,
                           const Tensor &opacity, const Tensor &background, int64_t H, int64_t W, int64_t cap, int64_t sticky,
                           bool want_aux, bool want_grads, const Memo *memo, Chain *chain_out) {
  Chain c;
  const int64_t N = xys.size(0);
  std::shared_ptr<RasterNode> node;
  if (want_grads) {
    node = std::shared_ptr<RasterNode>(new RasterNode(), torch::autograd::deleteNode);
    node->set_next_edges(torch::autograd::collect_next_edges(xys, conics, colors, opacity));
  }
  {
    at::AutoDispatchBelowADInplaceOrView below_autograd;
    if (memo) {
      const Memo &m = *memo;
      chk_in(conics, "conics", at::kFloat);
      chk_in(colors, "colors", at::kFloat);
      chk_in(opacity, "opacity", at::kFloat);
      chk_in(background, "background", at::kFloat);
      TORCH_CHECK(conics.numel() == 3 * N && colors.numel() == 3 * N && opacity.numel() == N && background.numel() == 3,
                  "rasterize_gaussians_sum: inconsistent tensor sizes for ", N, " Gaussians");
      const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
      c10::cuda::CUDAGuard guard(xys.device());
      cudaStream_t st = stream_of(xys);
      auto f = xys.options().requires_grad(false);
      c.scratch = m.scratch; c.tile_bins = m.tile_bins; c.o_gids = m.o_gids; c.o_status = m.o_status; c.cap = m.cap;
      c.packed = at::empty({m.cap * 12}, f);
      if (want_grads) c.grads12 = at::empty({N, 12}, f);
      uint8_t *base = c.scratch.data_ptr<uint8_t>();
      const int *d_status = reinterpret_cast<const int *>(base + c.o_status);
      chk(gi2d_repack_records_ex((int)m.cap, reinterpret_cast<const int *>(base + c.o_gids), d_status, fp(xys), fp(conics), fp(colors),
                                 fp(opacity), c.packed.data_ptr<float>(), (int)N, want_grads ? c.grads12.data_ptr<float>() : nullptr, st),
          "repack_records");
      c.out_img = at::empty({H, W, 3}, f);
      if (want_aux) {
        c.final_Ts = at::empty({H, W}, f);
        c.final_idx = at::empty({H, W}, f.dtype(at::kInt));
      }
      chk(gi2d_raster_sum_fwd_packed(TX, TY, (int)W, (int)H, c.tile_bins.data_ptr<int>(), c.packed.data_ptr<float>(), d_status,
                                     fp(background), c.out_img.data_ptr<float>(), want_aux ? c.final_Ts.data_ptr<float>() : nullptr,
                                     want_aux ? c.final_idx.data_ptr<int>() : nullptr, st),
          "rasterize_sum_forward");
    } else {
      c = run_chain(xys, radii, conics, nth, colors, opacity, background, H, W, cap, nullptr, reinterpret_cast<int *>(sticky), want_aux,
                    want_grads);
    }
  }
  if (chain_out) *chain_out = c;                  // bookkeeping for the caller (memo, last status)
  if (node) {
    node->N = N; node->H = H; node->W = W;
    node->opacity_shape = opacity.sizes().vec();
    node->tile_bins = c.tile_bins; node->packed = c.packed; node->grads12 = c.grads12;
    node->fresh_grads = true;
    torch::autograd::set_history(c.out_img, node);
  }
  return {c.out_img, want_aux ? c.final_Ts : Tensor()};
}

// rasterize_gaussians_sum, fused path, as ONE call: returns {} when the arguments are not the unmodified outputs of one of our
// projections (the caller then takes the reference's staged sequence), else {out_img} or {out_img, out_alpha}.
std::vector<Tensor> rasterize(const Tensor &xys, const Tensor &depths, const Tensor &radii, const Tensor &conics, const Tensor &nth,
                              const Tensor &colors, const Tensor &opacity, const Tensor &background, int64_t H, int64_t W,
                              bool return_alpha) {
  State &s = S();
  if (!xys.is_cuda() || xys.scalar_type() != at::kFloat) return {};
  const int64_t N = xys.size(0);
  const int64_t TX = (W + 15) / 16, TY = (H + 15) / 16;
  const int dev = xys.get_device();
  cudaStream_t st = stream_of(xys);
  const bool cap_mode = capturing(st);
  const Key key{dev, N, TX, TY};
  int64_t cap = 0, sticky = 0;
  const Memo *hit = nullptr;
  Memo hit_copy;
  int pi;
  uint64_t proj_id = 0;
  bool strict;
  volatile int *wait_on = nullptr;
  {
    std::lock_guard<std::mutex> g(s.mu);
    pi = find_projection(s, xys, depths, radii, nth);
    if (pi < 0) return {};
    Projection &e = s.projections[pi];
    proj_id = e.id;
    ++s.fast_calls;
    strict = !(s.sync_free || cap_mode);
    // entries recorded during a capture belong to that graph's memory pool: drop them at the first call outside of it
    for (size_t k = 0; k < s.memo.size();)
      if (s.memo[k].from_capture && !cap_mode) s.memo.erase(s.memo.begin() + k); else ++k;
    if (s.memoise)
      for (const Memo &m : s.memo)
        if (m.dev == dev && m.stream == st && m.proj_id == e.id && m.N == N && m.H == H && m.W == W) { hit_copy = m; hit = &hit_copy; break; }
    if (hit) {
      ++s.geometry_hits;
    } else if (strict) {
      if (e.n_isect >= 0) cap = e.n_isect;
      else if (e.slot >= 0 && s.ring.gen[e.slot] == e.slot_gen) wait_on = s.ring.p + e.slot;
      else cap = -1;                                 // the slot was recycled (or the projection ran inside a capture): count on demand
    } else {
      auto it = s.capacity.find(key);
      cap = it == s.capacity.end() ? default_capacity(N) : it->second;
      auto sk = s.sticky.find(key);
      if (sk != s.sticky.end()) sticky = reinterpret_cast<int64_t>(sk->second.data_ptr<int>());
    }
    if (!cap_mode && s.sticky.find(key) == s.sticky.end()) {
      // the sticky {max n, overflow} pair of this shape: allocated (and zeroed) at the first EAGER call, so that a later capture
      // records a pointer to persistent memory, never an allocation or a zero fill that would be replayed
      c10::cuda::CUDAGuard guard(xys.device());
      auto sk = s.sticky.emplace(key, at::zeros({2}, xys.options().dtype(at::kInt).requires_grad(false))).first;
      if (!strict && !hit) sticky = reinterpret_cast<int64_t>(sk->second.data_ptr<int>());
    }
  }
  // Strict mode sizes the workspace from the projection's intersection total.  If the total has not arrived yet (a render loop
  // calls us microseconds after the projection was launched) we do not wait for it here: the chain is launched with the
  // capacity this shape has learned so far, the total is awaited AFTER everything is enqueued -- by then it is there -- and
  // only if it exceeds that capacity (first calls of a shape, a sudden growth) the pass is redone with the exact size.
  auto await_total = [&]() -> int64_t {
    pybind11::gil_scoped_release nogil;
    uint64_t spins = 0;
    while (*wait_on < 0) {
      if ((++spins & ((1u << 22) - 1)) == 0) {       // seconds without an answer: surface a CUDA fault if there is one
        const cudaError_t err = cudaStreamSynchronize(st);
        TORCH_CHECK(err == cudaSuccess, "gi2d: CUDA error while waiting for the intersection total: ", cudaGetErrorString(err));
        TORCH_CHECK(*wait_on >= 0, "gi2d: the projection never published its intersection total");
      }
    }
    return *wait_on;
  };
  auto exact_cap = [](int64_t n) { return std::max<int64_t>((n + 4095) / 4096 * 4096, 4096); };
  auto learn = [&](int64_t n, bool waited) {
    std::lock_guard<std::mutex> g(s.mu);
    Projection &e = s.projections[pi];
    if (e.id == proj_id) e.n_isect = n;
    if (waited) ++s.total_waits;
    int64_t &learned = s.capacity[key];                // what a later capture / sync-free phase of this shape starts from
    learned = std::max<int64_t>(learned, n + n / 4 + 1024);
  };
  bool optimistic = false;
  if (!hit && strict) {
    int64_t n = cap;
    if (wait_on) {
      if (*wait_on >= 0) {
        n = *wait_on;
      } else {
        int64_t learned = 0;
        {
          std::lock_guard<std::mutex> g(s.mu);
          auto it = s.capacity.find(key);
          if (it != s.capacity.end()) learned = it->second;
        }
        if (learned > 0) { optimistic = true; n = -2; cap = learned; }
        else n = await_total();
      }
    } else if (n < 0) {
      n = nth.sum().item<int64_t>();                   // rare: one synchronising read, like the reference's `.item()`
      std::lock_guard<std::mutex> g(s.mu);
      ++s.total_fallbacks;
    }
    if (!optimistic) {
      cap = exact_cap(n);
      learn(n, false);
    }
  }
  const bool want_grads = at::GradMode::is_enabled() &&
                          (xys.requires_grad() || conics.requires_grad() || colors.requires_grad() || opacity.requires_grad());
  Chain chain;
  variable_list out = raster_apply(xys, radii, conics, nth, colors, opacity, background, H, W, cap, sticky, return_alpha, want_grads, hit,
                                   &chain);
  if (optimistic) {
    const int64_t n = await_total();
    learn(n, true);
    if (n > cap) {                                     // truncated: redo with the exact size (the first pass is dropped)
      cap = exact_cap(n);
      out = raster_apply(xys, radii, conics, nth, colors, opacity, background, H, W, cap, sticky, return_alpha, want_grads, hit, &chain);
    }
  }
  {
    std::lock_guard<std::mutex> g(s.mu);
    LastStatus &ls = s.last[dev];
    ls.scratch = chain.scratch; ls.o_status = chain.o_status; ls.cap = chain.cap; ls.key = key; ls.valid = true;
    if (!hit && s.memoise) {
      Memo *slot = nullptr;
      for (Memo &m : s.memo)
        if (m.dev == dev && m.stream == st) slot = &m;
      if (!slot) { s.memo.emplace_back(); slot = &s.memo.back(); }
      slot->proj_id = proj_id; slot->dev = dev; slot->stream = st; slot->N = N; slot->H = H; slot->W = W;
      slot->scratch = chain.scratch; slot->tile_bins = chain.tile_bins; slot->o_gids = chain.o_gids; slot->o_status = chain.o_status;
      slot->cap = chain.cap; slot->from_capture = cap_mode;
    }
  }
  if (return_alpha) return {out[0], at::rsub(out[1], 1)};     // reference :181-183 (all zeros unless there were no intersections)
  return {out[0]};
}

// ---- state accessors (rasterize_sum.py routes its public switches here) ------------------------------------------------
void set_sync_free(bool f) { std::lock_guard<std::mutex> g(S().mu); S().sync_free = f; }
void set_memoise(bool f) { std::lock_guard<std::mutex> g(S().mu); S().memoise = f; S().memo.clear(); }
void clear_memo() { std::lock_guard<std::mutex> g(S().mu); S().memo.clear(); }
int64_t get_capacity(int64_t dev, int64_t N, int64_t TX, int64_t TY) {
  std::lock_guard<std::mutex> g(S().mu);
  auto it = S().capacity.find(Key{(int)dev, N, TX, TY});
  return it == S().capacity.end() ? 0 : it->second;
}
void set_capacity(int64_t dev, int64_t N, int64_t TX, int64_t TY, int64_t cap) {
  std::lock_guard<std::mutex> g(S().mu);
  S().capacity[Key{(int)dev, N, TX, TY}] = cap;
}
std::map<std::string, int64_t> stats() {
  std::lock_guard<std::mutex> g(S().mu);
  return {{"fast_calls", S().fast_calls}, {"geometry_hits", S().geometry_hits}, {"total_waits", S().total_waits},
          {"total_fallbacks", S().total_fallbacks}};
}

// Synchronising check of `dev`: the sticky status pairs of every shape used in sync-free / captured mode and the status of the
// most recent chain.  Returns {n_isect of the most recent chain, overflowed}; when something overflowed the capacity of that
// shape is raised, the sticky pair cleared and the memo dropped (its lists are truncated) -- the caller raises.
std::tuple<int64_t, bool, int64_t, int64_t> check_overflow(int64_t dev) {
  State &s = S();
  std::vector<std::pair<Key, Tensor>> pairs;
  LastStatus ls;
  {
    std::lock_guard<std::mutex> g(s.mu);
    for (auto &kv : s.sticky)
      if (std::get<0>(kv.first) == (int)dev) pairs.emplace_back(kv.first, kv.second);
    auto it = s.last.find((int)dev);
    if (it != s.last.end()) ls = it->second;
  }
  int64_t n_last = 0, n_over = 0, cap_over = 0;
  bool over = false;
  if (ls.valid) {
    Tensor st = ls.scratch.slice(0, ls.o_status, ls.o_status + 8).view(at::kInt).cpu();
    n_last = st[0].item<int>();
    if (st[1].item<int>()) { over = true; n_over = n_last; cap_over = ls.cap; }
  }
  for (auto &kv : pairs) {
    Tensor h = kv.second.cpu();
    const int64_t n = h[0].item<int>();
    if (h[1].item<int>()) {
      over = true;
      n_over = std::max(n_over, n);
      std::lock_guard<std::mutex> g(s.mu);
      int64_t &c = s.capacity[kv.first];
      if (!cap_over) cap_over = c ? c : default_capacity(std::get<1>(kv.first));
      c = std::max<int64_t>(c, n + n / 4 + 1024);
      kv.second.zero_();
    }
  }
  if (over) {
    std::lock_guard<std::mutex> g(s.mu);
    if (ls.valid) {
      int64_t &c = s.capacity[ls.key];
      c = std::max<int64_t>(c, n_over + n_over / 4 + 1024);
    }
    s.memo.clear();
    s.last.erase((int)dev);
  }
  return {n_last, over, n_over, cap_over};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "gi2d host glue: the operator layer's host side over libgi2d.so (include/gi2d.h)";
  m.def("abi_version", []() { return gi2d_abi_version(); });
  m.def("project_fwd", &project_fwd);
  m.def("project_bwd_packed", &project_bwd_packed);
  m.def("fused_forward", &fused_forward);
  m.def("raster_bwd_packed", &raster_bwd_packed);
  m.def("fused_offsets", &fused_offsets);
  m.def("project_autograd", &project_autograd);
  m.def("rasterize", &rasterize);
  m.def("set_sync_free", &set_sync_free);
  m.def("set_memoise", &set_memoise);
  m.def("clear_memo", &clear_memo);
  m.def("get_capacity", &get_capacity);
  m.def("set_capacity", &set_capacity);
  m.def("stats", &stats);
  m.def("check_overflow", &check_overflow);
}
