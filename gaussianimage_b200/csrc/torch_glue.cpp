// Host glue over the C ABI (include/gi2d.h): the "allocate the outputs, look up the stream, launch" sequences of
// gaussianimage_b200/cuda.py and rasterize_sum.py::fused_chain as ONE C++ call each, so that the eager operator path
// (what an unchanged caller of the reference gets) spends its host time in torch's autograd, not in ctypes marshalling
// and a dozen Python-level torch.empty calls per step.  It replaces, like cuda.py, the reference's bindings.cu host
// wrappers (bindings.cu:400-469, 706-779, 781-1004); the kernels, the ABI and the results are the same ones -- this file
// contains no device code and no arithmetic.  Optional: when the module is absent the ctypes path of cuda.py runs.
#include <torch/extension.h>
#include <c10/cuda/CUDAGuard.h>
#include <c10/cuda/CUDAStream.h>

#include <tuple>

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

// cuda.py: project_gaussians_2d_forward / project_gaussians_2d_scale_rot_forward.  mode 0 = Cholesky (p0 = L (N,3)),
// mode 1 = scale-rotation (p0 = scales (N,2), p1 = rotation (N,1)).
std::tuple<Tensor, Tensor, Tensor, Tensor, Tensor> project_fwd(int64_t mode, int64_t N, const Tensor &means2d, const Tensor &p0,
                                                               const OptTensor &p1, int64_t H, int64_t W, int64_t TX,
                                                               int64_t TY) {
  chk_in(means2d, "means2d", at::kFloat);
  chk_in(p0, mode == 0 ? "L_elements" : "scales2d", at::kFloat);
  if (mode != 0) {
    TORCH_CHECK(p1.has_value(), "rotation is required");
    chk_in(*p1, "rotation", at::kFloat);
  }
  c10::cuda::CUDAGuard guard(means2d.device());
  auto f = means2d.options();
  auto i = f.dtype(at::kInt);
  Tensor xys = at::empty({N, 2}, f), depths = at::empty({N}, f), radii = at::empty({N}, i);
  Tensor conics = at::empty({N, 3}, f), nth = at::empty({N}, i);
  cudaStream_t st = stream_of(means2d);
  if (mode == 0)
    chk(gi2d_project_cholesky_fwd((int)N, fp(means2d), fp(p0), (int)H, (int)W, (int)TX, (int)TY, xys.data_ptr<float>(),
                                  depths.data_ptr<float>(), radii.data_ptr<int>(), conics.data_ptr<float>(),
                                  nth.data_ptr<int>(), st),
        "project_gaussians_2d_forward");
  else
    chk(gi2d_project_scale_rot_fwd((int)N, fp(means2d), fp(p0), fp(*p1), (int)H, (int)W, (int)TX, (int)TY,
                                   xys.data_ptr<float>(), depths.data_ptr<float>(), radii.data_ptr<int>(),
                                   conics.data_ptr<float>(), nth.data_ptr<int>(), st),
        "project_gaussians_2d_scale_rot_forward");
  return {xys, depths, radii, conics, nth};
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
  auto f = conics.options();
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

// rasterize_sum.py: fused_chain without a custom rasterizer -- one scratch allocation [workspace | gids_sorted | status],
// tile_bins (T,2), packed (cap*12), the image and (optionally) final_Ts / final_idx; gi2d_bin_fused_rows_notify followed
// by gi2d_raster_sum_fwd_packed on the current stream.  host_status = address of a pinned {n_isect, overflow} slot or 0.
std::tuple<Tensor, OptTensor, OptTensor, Tensor, Tensor, Tensor, int64_t, int64_t> fused_forward(
    const Tensor &xys, const Tensor &radii, const Tensor &conics, const Tensor &nth, const Tensor &colors, const Tensor &opacity,
    const Tensor &background, int64_t H, int64_t W, int64_t cap, int64_t host_status, bool want_aux) {
  const int64_t N = xys.size(0);
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  c10::cuda::CUDAGuard guard(xys.device());
  cudaStream_t st = stream_of(xys);
  const size_t ws_bytes = gi2d_bin_fused_workspace_bytes((int)N, TX, TY, (int)cap);
  const int64_t o_gids = align256((int64_t)ws_bytes), o_status = o_gids + align256(4 * cap);
  auto f = xys.options();
  auto i = f.dtype(at::kInt);
  Tensor scratch = at::empty({o_status + 256}, f.dtype(at::kByte));
  Tensor tile_bins = at::empty({(int64_t)TX * TY, 2}, i);
  Tensor packed = at::empty({cap * 12}, f);
  uint8_t *base = scratch.data_ptr<uint8_t>();
  int *d_status = reinterpret_cast<int *>(base + o_status);
  chk(gi2d_bin_fused_rows_notify((int)N, TX, TY, 0, TY, fp(xys), ip(radii), ip(nth), fp(conics), fp(colors), fp(opacity), (int)cap,
                                 base, ws_bytes, reinterpret_cast<int *>(base + o_gids), tile_bins.data_ptr<int>(),
                                 packed.data_ptr<float>(), d_status, reinterpret_cast<int *>(host_status), st),
      "bin_fused");
  Tensor out_img = at::empty({H, W, 3}, f);
  OptTensor final_Ts, final_idx;
  if (want_aux) {
    final_Ts = at::empty({H, W}, f);
    final_idx = at::empty({H, W}, i);
  }
  chk(gi2d_raster_sum_fwd_packed(TX, TY, (int)W, (int)H, tile_bins.data_ptr<int>(), packed.data_ptr<float>(), d_status,
                                 fp(background), out_img.data_ptr<float>(), want_aux ? final_Ts->data_ptr<float>() : nullptr,
                                 want_aux ? final_idx->data_ptr<int>() : nullptr, st),
      "rasterize_sum_forward");
  return {out_img, final_Ts, final_idx, tile_bins, packed, scratch, o_gids, o_status};
}

// rasterize_sum.py: _RasterizeGaussiansSum.backward, fused branch -> the (N,12) gradient rows.
Tensor raster_bwd_packed(int64_t N, int64_t H, int64_t W, const Tensor &tile_bins, const Tensor &packed, const Tensor &v_out) {
  chk_in(v_out, "v_out_img", at::kFloat);
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  c10::cuda::CUDAGuard guard(packed.device());
  Tensor grads = at::empty({N, 12}, packed.options());
  chk(gi2d_raster_sum_bwd_packed((int)N, TX, TY, (int)W, (int)H, ip(tile_bins), fp(packed), fp(v_out), grads.data_ptr<float>(),
                                 stream_of(packed)),
      "rasterize_sum_backward");
  return grads;
}


// ---------------------------------------------------------------------------------------------------------------------
// The two operators as torch::autograd::Function (opt-in from Python, GI2D_CPP_AUTOGRAD=1): the same forward / backward
// calls as above, but the autograd nodes live in C++, so a backward pass does not have to take the GIL and re-enter
// Python twice per step.  Policy (strict-mode polling, capacity retry, memoised binning) stays in rasterize_sum.py.
// ---------------------------------------------------------------------------------------------------------------------
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

inline bool packed_pair(const Tensor &v_xy, const Tensor &v_conic) {   // cuda.py: _is_packed_pair
  return v_xy.defined() && v_conic.defined() && v_xy.is_cuda() && v_xy.scalar_type() == at::kFloat &&
         v_conic.scalar_type() == at::kFloat && v_xy.dim() == 2 && v_conic.dim() == 2 && v_xy.size(1) == 2 &&
         v_conic.size(1) == 3 && v_xy.stride(0) == 12 && v_xy.stride(1) == 1 && v_conic.stride(0) == 12 &&
         v_conic.stride(1) == 1 &&
         static_cast<const char *>(v_conic.data_ptr()) == static_cast<const char *>(v_xy.data_ptr()) + 8 &&
         reinterpret_cast<uintptr_t>(v_xy.data_ptr()) % 16 == 0;
}

// project_gaussians_2d.py / project_gaussians_2d_scale_rot.py: _ProjectGaussians2d[ScaleRot]
struct ProjectFn : public torch::autograd::Function<ProjectFn> {
  static variable_list forward(AutogradContext *ctx, const Tensor &means2d, const Tensor &p0, const Tensor &p1, int64_t mode,
                               int64_t H, int64_t W, int64_t TX, int64_t TY) {
    const int64_t N = means2d.size(-2);
    auto out = project_fwd(mode, N, means2d, p0, mode != 0 ? OptTensor(p1) : OptTensor(), H, W, TX, TY);
    Tensor xys = std::get<0>(out), depths = std::get<1>(out), radii = std::get<2>(out), conics = std::get<3>(out),
           nth = std::get<4>(out);
    ctx->saved_data["mode"] = mode;
    ctx->saved_data["N"] = N;
    ctx->saved_data["H"] = H;
    ctx->saved_data["W"] = W;
    ctx->save_for_backward({p0, p1, radii, conics});
    ctx->mark_non_differentiable({radii, nth});
    ctx->set_materialize_grads(false);
    return {xys, depths, radii, conics, nth};
  }

  static variable_list backward(AutogradContext *ctx, variable_list g) {
    const auto saved = ctx->get_saved_variables();
    const Tensor &p0 = saved[0], &p1 = saved[1], &radii = saved[2], &conics = saved[3];
    const int64_t mode = ctx->saved_data["mode"].toInt(), N = ctx->saved_data["N"].toInt();
    const int64_t H = ctx->saved_data["H"].toInt(), W = ctx->saved_data["W"].toInt();
    Tensor v_xy = g[0], v_conic = g[3];
    Tensor v_mean2d, v_a, v_b;
    if (packed_pair(v_xy, v_conic)) {
      auto r = project_bwd_packed(mode, N, p0, mode != 0 ? OptTensor(p1) : OptTensor(), H, W, radii, conics, v_xy);
      v_mean2d = std::get<0>(r);
      v_a = std::get<1>(r);
      if (std::get<2>(r).has_value()) v_b = *std::get<2>(r);
    } else {                                       // cuda.py: project_gaussians_2d[_scale_rot]_backward without v_cov2d
      c10::cuda::CUDAGuard guard(conics.device());
      auto f = conics.options();
      v_xy = v_xy.defined() ? v_xy.contiguous() : at::zeros({N, 2}, f);
      v_conic = v_conic.defined() ? v_conic.contiguous() : at::zeros({N, 3}, f);
      chk_in(v_xy, "v_xy", at::kFloat);
      chk_in(v_conic, "v_conic", at::kFloat);
      v_mean2d = at::empty({N, 2}, f);
      cudaStream_t st = stream_of(conics);
      if (mode == 0) {
        v_a = at::empty({N, 3}, f);
        chk(gi2d_project_cholesky_bwd((int)N, fp(p0), (int)H, (int)W, ip(radii), fp(conics), fp(v_xy), fp(v_conic), nullptr,
                                      v_mean2d.data_ptr<float>(), v_a.data_ptr<float>(), st),
            "project_gaussians_2d_backward");
      } else {
        v_a = at::empty({N, 2}, f);
        v_b = at::empty({N, 1}, f);
        chk(gi2d_project_scale_rot_bwd((int)N, fp(p0), fp(p1), (int)H, (int)W, ip(radii), fp(conics), fp(v_xy), fp(v_conic),
                                       nullptr, v_mean2d.data_ptr<float>(), v_a.data_ptr<float>(), v_b.data_ptr<float>(), st),
            "project_gaussians_2d_scale_rot_backward");
      }
    }
    return {v_mean2d, v_a, v_b, Tensor(), Tensor(), Tensor(), Tensor(), Tensor()};
  }
};

// rasterize_sum.py: _RasterizeGaussiansSum, fused branch.  Outputs: out_img, final_Ts, final_idx (both empty (0,) unless
// want_aux), tile_bins, packed, scratch -- everything but out_img is non-differentiable bookkeeping for the Python side.
struct RasterFusedFn : public torch::autograd::Function<RasterFusedFn> {
  static variable_list forward(AutogradContext *ctx, const Tensor &xys, const Tensor &radii, const Tensor &conics,
                               const Tensor &nth, const Tensor &colors, const Tensor &opacity, const Tensor &background, int64_t H,
                               int64_t W, int64_t cap, int64_t host_status, bool want_aux) {
    auto r = fused_forward(xys, radii, conics, nth, colors, opacity, background, H, W, cap, host_status, want_aux);
    Tensor out_img = std::get<0>(r), tile_bins = std::get<3>(r), packed = std::get<4>(r), scratch = std::get<5>(r);
    Tensor final_Ts = want_aux ? *std::get<1>(r) : at::empty({0}, out_img.options());
    Tensor final_idx = want_aux ? *std::get<2>(r) : at::empty({0}, tile_bins.options());
    ctx->saved_data["N"] = xys.size(0);
    ctx->saved_data["H"] = H;
    ctx->saved_data["W"] = W;
    ctx->saved_data["opacity_shape"] = opacity.sizes().vec();
    ctx->save_for_backward({tile_bins, packed});
    ctx->mark_non_differentiable({final_Ts, final_idx, tile_bins, packed, scratch});
    ctx->set_materialize_grads(false);
    return {out_img, final_Ts, final_idx, tile_bins, packed, scratch};
  }

  static variable_list backward(AutogradContext *ctx, variable_list g) {
    variable_list out(12);
    if (!g[0].defined()) return out;              // only out_alpha was used downstream (the reference ignores v_out_alpha, Q7)
    const auto saved = ctx->get_saved_variables();
    const int64_t N = ctx->saved_data["N"].toInt(), H = ctx->saved_data["H"].toInt(), W = ctx->saved_data["W"].toInt();
    Tensor grads = raster_bwd_packed(N, H, W, saved[0], saved[1], g[0].contiguous());
    out[0] = grads.slice(1, 0, 2);                                                     // v_xy
    out[2] = grads.slice(1, 2, 5);                                                     // v_conic
    out[4] = grads.slice(1, 5, 8);                                                     // v_colors
    out[5] = grads.slice(1, 8, 9).reshape(ctx->saved_data["opacity_shape"].toIntVector());   // v_opacity
    return out;
  }
};

variable_list project_autograd(int64_t mode, const Tensor &means2d, const Tensor &p0, const OptTensor &p1, int64_t H, int64_t W,
                               int64_t TX, int64_t TY) {
  // Function::apply wants defined tensors: in Cholesky mode the unused third slot carries p0 again (its gradient slot
  // stays undefined)
  return ProjectFn::apply(means2d, p0, p1.has_value() ? *p1 : p0, mode, H, W, TX, TY);
}

variable_list raster_fused_autograd(const Tensor &xys, const Tensor &radii, const Tensor &conics, const Tensor &nth,
                                    const Tensor &colors, const Tensor &opacity, const Tensor &background, int64_t H, int64_t W,
                                    int64_t cap, int64_t host_status, bool want_aux) {
  return RasterFusedFn::apply(xys, radii, conics, nth, colors, opacity, background, H, W, cap, host_status, want_aux);
}

// offsets of gids_sorted / status inside the scratch tensor of fused_forward (same arithmetic, for the Python side)
std::tuple<int64_t, int64_t> fused_offsets(int64_t N, int64_t H, int64_t W, int64_t cap) {
  const int TX = (int)((W + 15) / 16), TY = (int)((H + 15) / 16);
  const int64_t o_gids = align256((int64_t)gi2d_bin_fused_workspace_bytes((int)N, TX, TY, (int)cap));
  return {o_gids, o_gids + align256(4 * cap)};
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "gi2d host glue: allocation + launch sequences over libgi2d.so (include/gi2d.h)";
  m.def("abi_version", []() { return gi2d_abi_version(); });
  m.def("project_fwd", &project_fwd);
  m.def("project_bwd_packed", &project_bwd_packed);
  m.def("fused_forward", &fused_forward);
  m.def("raster_bwd_packed", &raster_bwd_packed);
  m.def("project_autograd", &project_autograd);
  m.def("raster_fused_autograd", &raster_fused_autograd);
  m.def("fused_offsets", &fused_offsets);
}
