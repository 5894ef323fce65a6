// lt_b200_torch.cpp — the thin torch extension over the C ABI of include/lt_b200.h.
//
// `torch.ops.lt_b200.*`: the three functions the reference's nets call (mvn/utils/op.py:99-163 unproject_heatmaps,
// mvn/utils/op.py:84-96 integrate_tensor_3d_with_coordinates, mvn/utils/multiview.py:162-174
// triangulate_batch_of_points) as dispatcher ops with C++ autograd, plus the raw forward / backward launches they are
// made of (each with a CUDA kernel registration here and a fake-tensor registration in torch_ops.py, so torch.compile /
// export see opaque ops with known output shapes).
//
// No kernel lives here: every op allocates its outputs with ATen, takes torch's current stream and calls one or two
// `lt_*` entry points of liblt_b200.so (hand-written sm_100a kernels, no torch inside). The symbols are resolved at
// load time from the already loaded liblt_b200.so (python loads it RTLD_GLOBAL first), so this file needs no link line.
// CUDA only: there is no CPU registration, a CPU tensor ends in the dispatcher's own "no kernel for backend CPU" error.
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/autograd.h>
#include <torch/library.h>

#include "../../include/lt_b200.h"

namespace {

using at::Tensor;
using torch::autograd::AutogradContext;
using torch::autograd::variable_list;

inline void check(int rc) {
  TORCH_CHECK(rc == 0, "lt_b200: ", lt_last_error(), " (code ", rc, ")");
}
inline Tensor f32c(const Tensor& t) { return t.detach().to(at::kFloat).contiguous(); }
inline const float* fp(const Tensor& t) { return t.defined() ? t.data_ptr<float>() : nullptr; }
inline float* fpm(Tensor& t) { return t.defined() ? t.data_ptr<float>() : nullptr; }
inline void* stream() { return at::cuda::getCurrentCUDAStream().stream(); }
inline Tensor opt(const c10::optional<Tensor>& t) { return t.has_value() ? *t : Tensor(); }

// ---------------------------------------------------------------------------------------------------------------------
// raw launches (CUDA dispatch key). Tensors must be fp32 + contiguous; the differentiable ops below see to that.
// ---------------------------------------------------------------------------------------------------------------------
Tensor feat_to_hwc(const Tensor& hm) {  // (BV,C,H,W) -> flat gather layout
  TORCH_CHECK(hm.dim() == 4, "feat_to_hwc: (BV,C,H,W) expected");
  const c10::cuda::CUDAGuard g(hm.device());
  const int BV = hm.size(0), C = hm.size(1), H = hm.size(2), W = hm.size(3);
  Tensor out = at::empty({(int64_t)lt_feat_layout_elems(BV, C, H, W)}, hm.options());
  check(lt_feat_to_hwc(fp(hm), fpm(out), BV, C, H, W, stream()));
  return out;
}

Tensor feat_to_chw(const Tensor& hwcp, int64_t BV, int64_t C, int64_t H, int64_t W) {
  const c10::cuda::CUDAGuard g(hwcp.device());
  Tensor out = at::empty({BV, C, H, W}, hwcp.options());
  check(lt_feat_to_chw(fp(hwcp), fpm(out), BV, C, H, W, stream()));
  return out;
}

// coord: (B,N0,N1,N2,3) tensor, or the (B,18) spec when spec_dims = [N0,N1,N2] is given (coordinates in-kernel)
Tensor unproject_fwd(const Tensor& feat_hwc, const Tensor& proj, const Tensor& coord, const c10::optional<Tensor>& conf,
                     int64_t B, int64_t V, int64_t C, int64_t H, int64_t W, int64_t N0, int64_t N1, int64_t N2,
                     int64_t agg, bool align_corners, bool coord_is_spec, bool transfer) {
  const c10::cuda::CUDAGuard g(feat_hwc.device());
  Tensor out = at::empty({B, C, N0, N1, N2}, feat_hwc.options());
  const Tensor cf = opt(conf);
  if (coord_is_spec)
    check(lt_unproject_fwd_spec(fp(feat_hwc), fp(proj), fp(coord), transfer, fp(cf), fpm(out), B, V, C, H, W, N0, N1, N2,
                                agg, align_corners, stream()));
  else
    check(lt_unproject_fwd(fp(feat_hwc), fp(proj), fp(coord), fp(cf), fpm(out), B, V, C, H, W, N0, N1, N2, agg,
                           align_corners, stream()));
  return out;
}

std::tuple<Tensor, Tensor> unproject_bwd(const Tensor& grad_out, const Tensor& feat_hwc, const Tensor& proj,
                                         const Tensor& coord, const c10::optional<Tensor>& conf, int64_t B, int64_t V,
                                         int64_t C, int64_t H, int64_t W, int64_t N0, int64_t N1, int64_t N2,
                                         int64_t agg, bool align_corners, bool coord_is_spec, bool transfer,
                                         bool need_conf) {
  const c10::cuda::CUDAGuard g(feat_hwc.device());
  Tensor gfeat = at::zeros_like(feat_hwc);  // the kernel scatter-adds into it
  const Tensor cf = opt(conf);
  Tensor gconf, ws;
  if (need_conf && cf.defined()) {
    gconf = at::empty({B, V, C}, feat_hwc.options());
    const int64_t n = std::max<int64_t>((int64_t)lt_unproject_bwd_ws_bytes(B, V, C, N0, N1, N2) / 4, 1);
    ws = at::empty({n}, feat_hwc.options());
  }
  if (coord_is_spec)
    check(lt_unproject_bwd_spec(fp(grad_out), fp(feat_hwc), fp(proj), fp(coord), transfer, fp(cf), fpm(gfeat),
                                fpm(gconf), fpm(ws), B, V, C, H, W, N0, N1, N2, agg, align_corners, stream()));
  else
    check(lt_unproject_bwd(fp(grad_out), fp(feat_hwc), fp(proj), fp(coord), fp(cf), fpm(gfeat), fpm(gconf), fpm(ws), B,
                           V, C, H, W, N0, N1, N2, agg, align_corners, stream()));
  return {gfeat, gconf.defined() ? gconf : at::empty({0}, feat_hwc.options())};
}

std::tuple<Tensor, Tensor> softargmax3d_fwd(const Tensor& vol, const Tensor& coord, bool softmax, double multiplier,
                                            bool coord_is_spec, bool transfer) {
  TORCH_CHECK(vol.dim() == 5, "softargmax3d_fwd: (B,J,N0,N1,N2) expected");
  const c10::cuda::CUDAGuard g(vol.device());
  const int B = vol.size(0), J = vol.size(1), N0 = vol.size(2), N1 = vol.size(3), N2 = vol.size(4);
  const long long nvox = (long long)N0 * N1 * N2;
  Tensor coords = at::zeros({B, J, 3}, vol.options());
  Tensor prob = at::empty_like(vol);
  if (nvox > 0 && B * J > 0) {
    Tensor stats = at::empty({B, J, 5}, vol.options());
    Tensor ws = at::empty({std::max<int64_t>((int64_t)lt_softargmax_ws_bytes(B, J, nvox) / 4 + 4, 4)}, vol.options());
    if (coord_is_spec)
      check(lt_softargmax_fwd_spec(fp(vol), fp(coord), transfer, fpm(coords), fpm(prob), fpm(stats), fpm(ws), B, J, N0,
                                   N1, N2, softmax, (float)multiplier, stream()));
    else
      check(lt_softargmax_fwd(fp(vol), fp(coord), fpm(coords), fpm(prob), fpm(stats), fpm(ws), B, J, nvox, softmax,
                              (float)multiplier, stream()));
  }
  return {coords, prob};
}

Tensor softargmax3d_bwd(const Tensor& prob, const Tensor& coord, const Tensor& coords,
                        const c10::optional<Tensor>& grad_coords, const c10::optional<Tensor>& grad_prob, bool softmax,
                        double multiplier, bool coord_is_spec, bool transfer) {
  const c10::cuda::CUDAGuard g(prob.device());
  const int B = prob.size(0), J = prob.size(1), N0 = prob.size(2), N1 = prob.size(3), N2 = prob.size(4);
  const long long nvox = (long long)N0 * N1 * N2;
  Tensor gvol = at::empty_like(prob);
  if (nvox == 0 || B * J == 0) return gvol;
  Tensor gc = grad_coords.has_value() ? f32c(*grad_coords) : at::zeros({B, J, 3}, prob.options());
  Tensor gp = grad_prob.has_value() ? f32c(*grad_prob) : Tensor();
  Tensor dots;
  if (gp.defined() && softmax) {
    dots = at::empty({B, J}, prob.options());
    Tensor ws = at::empty({std::max<int64_t>((int64_t)lt_softargmax_ws_bytes(B, J, nvox) / 4 + 4, 4)}, prob.options());
    check(lt_softargmax_bwd_dot(fp(prob), fp(gp), fpm(dots), fpm(ws), B, J, nvox, stream()));
  }
  if (coord_is_spec)
    check(lt_softargmax_bwd_spec(fp(prob), fp(coord), transfer, fp(coords), fp(gc), fp(gp), fp(dots), fpm(gvol), B, J,
                                 N0, N1, N2, softmax, (float)multiplier, stream()));
  else
    check(lt_softargmax_bwd(fp(prob), fp(coord), fp(coords), fp(gc), fp(gp), fp(dots), fpm(gvol), B, J, nvox, softmax,
                            (float)multiplier, stream()));
  return gvol;
}

std::tuple<Tensor, Tensor> dlt_fwd(const Tensor& proj, const Tensor& pts, const c10::optional<Tensor>& conf) {
  const c10::cuda::CUDAGuard g(pts.device());
  const int B = pts.size(0), V = pts.size(1), J = pts.size(2);
  Tensor out = at::zeros({B, J, 3}, pts.options());  // multiview.py:164: always fp32
  Tensor saved = at::empty({B, J, LT_DLT_SAVED}, pts.options().dtype(at::kDouble));
  check(lt_dlt_fwd(fp(proj), fp(pts), fp(opt(conf)), fpm(out), saved.data_ptr<double>(), B, V, J, stream()));
  return {out, saved};
}

std::tuple<Tensor, Tensor> dlt_bwd(const Tensor& grad_out, const Tensor& proj, const Tensor& pts,
                                   const c10::optional<Tensor>& conf, const Tensor& saved, bool need_conf) {
  const c10::cuda::CUDAGuard g(pts.device());
  const int B = pts.size(0), V = pts.size(1), J = pts.size(2);
  Tensor gpts = at::zeros({B, V, J, 2}, pts.options());
  Tensor gconf = (need_conf && conf.has_value()) ? at::zeros({B, V, J}, pts.options()) : Tensor();
  check(lt_dlt_bwd(fp(grad_out), fp(proj), fp(pts), fp(opt(conf)), saved.data_ptr<double>(), fpm(gpts), fpm(gconf), B, V,
                   J, stream()));
  return {gpts, gconf.defined() ? gconf : at::empty({0}, pts.options())};
}

// ---------------------------------------------------------------------------------------------------------------------
// differentiable ops (Autograd key): argument normalisation + the raw ops through the dispatcher
// ---------------------------------------------------------------------------------------------------------------------
struct UnprojectFn : public torch::autograd::Function<UnprojectFn> {
  static Tensor forward(AutogradContext* ctx, const Tensor& heatmaps, const Tensor& proj, const Tensor& coord,
                        const c10::optional<Tensor>& conf, int64_t agg, bool align_corners, c10::IntArrayRef spec_dims,
                        bool transfer) {
    at::AutoDispatchBelowADInplaceOrView guard;
    const int64_t B = heatmaps.size(0), V = heatmaps.size(1), C = heatmaps.size(2), H = heatmaps.size(3),
                  W = heatmaps.size(4);
    const bool is_spec = spec_dims.size() == 3;
    const int64_t N0 = is_spec ? spec_dims[0] : coord.size(1), N1 = is_spec ? spec_dims[1] : coord.size(2),
                  N2 = is_spec ? spec_dims[2] : coord.size(3);
    const Tensor hm = f32c(heatmaps), proj_c = f32c(proj), coord_c = f32c(coord);
    c10::optional<Tensor> conf_c;
    if (conf.has_value()) conf_c = f32c(*conf);
    const Tensor feat = feat_to_hwc(hm.view({B * V, C, H, W}));  // K0: gather layout, opaque to this side
    Tensor out = unproject_fwd(feat, proj_c, coord_c, conf_c, B, V, C, H, W, N0, N1, N2, agg, align_corners, is_spec,
                               transfer);
    ctx->save_for_backward({feat, proj_c, coord_c, conf_c.has_value() ? *conf_c : Tensor()});
    ctx->saved_data["dims"] = std::vector<int64_t>{B, V, C, H, W, N0, N1, N2, agg, align_corners, is_spec, transfer};
    ctx->saved_data["hm_dtype"] = (int64_t)heatmaps.scalar_type();
    ctx->saved_data["conf_dtype"] = (int64_t)(conf.has_value() ? conf->scalar_type() : at::kFloat);
    return out;
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    const auto saved = ctx->get_saved_variables();
    const Tensor &feat = saved[0], &proj_c = saved[1], &coord_c = saved[2], &conf_c = saved[3];
    const auto d = ctx->saved_data["dims"].toIntVector();
    // input edges exist for the tensor arguments that were passed: (heatmaps, proj, coord[, conf])
    const bool need_hm = ctx->needs_input_grad(0), need_conf = conf_c.defined() && ctx->needs_input_grad(3);
    variable_list out(8);
    if (!(need_hm || need_conf)) return out;
    at::AutoDispatchBelowADInplaceOrView guard;
    c10::optional<Tensor> cf;
    if (conf_c.defined()) cf = conf_c;
    auto r = unproject_bwd(f32c(grads[0]), feat, proj_c, coord_c, cf, d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7], d[8],
                           d[9] != 0, d[10] != 0, d[11] != 0, need_conf);
    if (need_hm)
      out[0] = feat_to_chw(std::get<0>(r), d[0] * d[1], d[2], d[3], d[4])
                   .view({d[0], d[1], d[2], d[3], d[4]})
                   .to((at::ScalarType)ctx->saved_data["hm_dtype"].toInt());
    if (need_conf) out[3] = std::get<1>(r).to((at::ScalarType)ctx->saved_data["conf_dtype"].toInt());
    return out;
  }
};

Tensor unproject_heatmaps(const Tensor& heatmaps, const Tensor& proj, const Tensor& coord, int64_t agg,
                          const c10::optional<Tensor>& conf, bool align_corners, c10::IntArrayRef spec_dims,
                          bool transfer) {
  TORCH_CHECK(heatmaps.is_cuda() && proj.is_cuda() && coord.is_cuda(),
              "lt_b200: these ops only exist as CUDA kernels (no CPU fallback)");
  TORCH_CHECK(heatmaps.dim() == 5, "heatmaps must be (B,V,C,H,W)");
  return UnprojectFn::apply(heatmaps, proj, coord, conf, agg, align_corners, spec_dims, transfer);
}

struct SoftArgmax3dFn : public torch::autograd::Function<SoftArgmax3dFn> {
  static variable_list forward(AutogradContext* ctx, const Tensor& volumes, const Tensor& coord, bool softmax,
                               double multiplier, bool coord_is_spec, bool transfer) {
    at::AutoDispatchBelowADInplaceOrView guard;
    const Tensor vol = f32c(volumes), coord_c = f32c(coord);
    auto r = softargmax3d_fwd(vol, coord_c, softmax, multiplier, coord_is_spec, transfer);
    ctx->save_for_backward({std::get<1>(r), coord_c, std::get<0>(r)});
    ctx->saved_data["softmax"] = softmax;
    ctx->saved_data["multiplier"] = multiplier;
    ctx->saved_data["spec"] = coord_is_spec;
    ctx->saved_data["transfer"] = transfer;
    ctx->saved_data["dtype"] = (int64_t)volumes.scalar_type();
    ctx->set_materialize_grads(false);
    return {std::get<0>(r).to(volumes.scalar_type()), std::get<1>(r).to(volumes.scalar_type())};
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    variable_list out(6);
    if (!ctx->needs_input_grad(0)) return out;
    const auto saved = ctx->get_saved_variables();
    at::AutoDispatchBelowADInplaceOrView guard;
    c10::optional<Tensor> gc, gp;
    if (grads[0].defined()) gc = grads[0];
    if (grads[1].defined()) gp = grads[1];
    out[0] = softargmax3d_bwd(saved[0], saved[1], saved[2], gc, gp, ctx->saved_data["softmax"].toBool(),
                              ctx->saved_data["multiplier"].toDouble(), ctx->saved_data["spec"].toBool(),
                              ctx->saved_data["transfer"].toBool())
                 .to((at::ScalarType)ctx->saved_data["dtype"].toInt());
    return out;
  }
};

std::tuple<Tensor, Tensor> integrate_tensor_3d_with_coordinates(const Tensor& volumes, const Tensor& coord, bool softmax,
                                                                double multiplier, bool coord_is_spec, bool transfer) {
  TORCH_CHECK(volumes.is_cuda() && coord.is_cuda(), "lt_b200: these ops only exist as CUDA kernels (no CPU fallback)");
  TORCH_CHECK(volumes.dim() == 5, "volumes must be (B,J,N0,N1,N2)");
  TORCH_CHECK(multiplier > 0.0, "multiplier must be > 0 (scale on the caller side otherwise)");
  auto r = SoftArgmax3dFn::apply(volumes, coord, softmax, multiplier, coord_is_spec, transfer);
  return {r[0], r[1]};
}

struct DltFn : public torch::autograd::Function<DltFn> {
  static Tensor forward(AutogradContext* ctx, const Tensor& proj, const Tensor& pts, const c10::optional<Tensor>& conf) {
    at::AutoDispatchBelowADInplaceOrView guard;
    const Tensor proj_c = f32c(proj), pts_c = f32c(pts);
    c10::optional<Tensor> conf_c;
    if (conf.has_value()) conf_c = f32c(*conf);
    auto r = dlt_fwd(proj_c, pts_c, conf_c);
    ctx->save_for_backward({proj_c, pts_c, conf_c.has_value() ? *conf_c : Tensor(), std::get<1>(r)});
    ctx->saved_data["pts_dtype"] = (int64_t)pts.scalar_type();
    ctx->saved_data["conf_dtype"] = (int64_t)(conf.has_value() ? conf->scalar_type() : at::kFloat);
    return std::get<0>(r);
  }
  static variable_list backward(AutogradContext* ctx, variable_list grads) {
    variable_list out(3);
    const auto saved = ctx->get_saved_variables();
    const bool need_pts = ctx->needs_input_grad(1), need_conf = saved[2].defined() && ctx->needs_input_grad(2);
    if (!(need_pts || need_conf)) return out;
    at::AutoDispatchBelowADInplaceOrView guard;
    c10::optional<Tensor> cf;
    if (saved[2].defined()) cf = saved[2];
    auto r = dlt_bwd(f32c(grads[0]), saved[0], saved[1], cf, saved[3], need_conf);
    if (need_pts) out[1] = std::get<0>(r).to((at::ScalarType)ctx->saved_data["pts_dtype"].toInt());
    if (need_conf) out[2] = std::get<1>(r).to((at::ScalarType)ctx->saved_data["conf_dtype"].toInt());
    return out;
  }
};

Tensor triangulate_batch_of_points(const Tensor& proj, const Tensor& pts, const c10::optional<Tensor>& conf) {
  TORCH_CHECK(proj.is_cuda() && pts.is_cuda(), "lt_b200: these ops only exist as CUDA kernels (no CPU fallback)");
  TORCH_CHECK(pts.dim() == 4 && pts.size(3) == 2, "points_batch must be (B,V,J,2)");
  return DltFn::apply(proj, pts, conf);
}

}  // namespace

TORCH_LIBRARY(lt_b200, m) {
  m.def("feat_to_hwc(Tensor heatmaps) -> Tensor");
  m.def("feat_to_chw(Tensor hwcp, int BV, int C, int H, int W) -> Tensor");
  m.def("unproject_fwd(Tensor feat_hwc, Tensor proj, Tensor coord, Tensor? conf, int B, int V, int C, int H, int W, "
        "int N0, int N1, int N2, int agg, bool align_corners, bool coord_is_spec, bool transfer) -> Tensor");
  m.def("unproject_bwd(Tensor grad_out, Tensor feat_hwc, Tensor proj, Tensor coord, Tensor? conf, int B, int V, int C, "
        "int H, int W, int N0, int N1, int N2, int agg, bool align_corners, bool coord_is_spec, bool transfer, "
        "bool need_conf) -> (Tensor, Tensor)");
  m.def("softargmax3d_fwd(Tensor volumes, Tensor coord, bool softmax, float multiplier, bool coord_is_spec, "
        "bool transfer) -> (Tensor, Tensor)");
  m.def("softargmax3d_bwd(Tensor prob, Tensor coord, Tensor coords, Tensor? grad_coords, Tensor? grad_prob, "
        "bool softmax, float multiplier, bool coord_is_spec, bool transfer) -> Tensor");
  m.def("dlt_fwd(Tensor proj, Tensor points, Tensor? conf) -> (Tensor, Tensor)");
  m.def("dlt_bwd(Tensor grad_out, Tensor proj, Tensor points, Tensor? conf, Tensor saved, bool need_conf) -> "
        "(Tensor, Tensor)");
  // the reference's functions (agg: 0 sum, 1 conf*, 2 max, 3 softmax; spec_dims = [N0,N1,N2] when `coord_volumes`
  // is the (B,18) spec of lt_b200.CoordSpec)
  m.def("unproject_heatmaps(Tensor heatmaps, Tensor proj_matricies, Tensor coord_volumes, int agg, "
        "Tensor? vol_confidences, bool align_corners, int[] spec_dims, bool transfer) -> Tensor");
  m.def("integrate_tensor_3d_with_coordinates(Tensor volumes, Tensor coord_volumes, bool softmax, float multiplier, "
        "bool coord_is_spec, bool transfer) -> (Tensor, Tensor)");
  m.def("triangulate_batch_of_points(Tensor proj_matricies_batch, Tensor points_batch, Tensor? confidences_batch) "
        "-> Tensor");
}

TORCH_LIBRARY_IMPL(lt_b200, CUDA, m) {
  m.impl("feat_to_hwc", &feat_to_hwc);
  m.impl("feat_to_chw", &feat_to_chw);
  m.impl("unproject_fwd", &unproject_fwd);
  m.impl("unproject_bwd", &unproject_bwd);
  m.impl("softargmax3d_fwd", &softargmax3d_fwd);
  m.impl("softargmax3d_bwd", &softargmax3d_bwd);
  m.impl("dlt_fwd", &dlt_fwd);
  m.impl("dlt_bwd", &dlt_bwd);
}

// the differentiable ops own their autograd (C++ Function nodes); registered for every backend so that the CUDA
// check inside produces the library's own error message
TORCH_LIBRARY_IMPL(lt_b200, CompositeImplicitAutograd, m) {
  m.impl("unproject_heatmaps", &unproject_heatmaps);
  m.impl("integrate_tensor_3d_with_coordinates", &integrate_tensor_3d_with_coordinates);
  m.impl("triangulate_batch_of_points", &triangulate_batch_of_points);
}
