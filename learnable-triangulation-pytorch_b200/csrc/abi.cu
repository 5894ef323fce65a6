// abi.cu — the extern "C" surface declared in include/lt_b200.h. Argument checking, error strings and
// dispatch only; every entry point ends in a kernel launch on the caller's stream (no CPU path exists).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/lt_b200.h"
#include "kernels.h"

namespace {

thread_local char g_err[512] = "";

int fail_invalid(const char* fn, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", fn, what);
  return LT_ERR_INVALID;
}

int finish(const char* fn, cudaError_t e) {
  if (e == cudaSuccess) return 0;
  snprintf(g_err, sizeof(g_err), "%s: CUDA error %d (%s)", fn, (int)e, cudaGetErrorString(e));
  return (int)e;
}

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

bool bad_unproj_shape(int B, int V, int C, int H, int W, int N0, int N1, int N2) {
  return B < 0 || V < 1 || C < 0 || H < 1 || W < 1 || N0 < 0 || N1 < 0 || N2 < 0 ||
         (long long)(H + 2) * (W + 2) * lt::padded_channels(C > 0 ? C : 1) > 0x0fffffffLL ||
         // tap descriptors are 32-bit float offsets from an item's first map (unproject.cuh build_descs)
         (long long)V * (H + 2) * (W + 2) * lt::padded_channels(C > 0 ? C : 1) > 0x7fffffffLL;
}

}  // namespace

#define LT_REQUIRE(cond, msg) \
  do {                        \
    if (!(cond)) return fail_invalid(__func__, msg); \
  } while (0)

// buffers the kernels only ever access as float4 (include/lt_b200.h "alignment")
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

#define LT_API __attribute__((visibility("default")))
extern "C" {

LT_API int lt_abi_version(void) { return LT_ABI_VERSION; }
LT_API const char* lt_last_error(void) { return g_err; }

LT_API int lt_padded_channels(int C) { return C > 0 ? lt::padded_channels(C) : 0; }
LT_API size_t lt_feat_layout_elems(int BV, int C, int H, int W) { return lt::feat_layout_elems(BV, C, H, W); }

LT_API int lt_feat_to_hwc(const float* src_chw, float* dst_hwc, int BV, int C, int H, int W, void* stream) {
  LT_REQUIRE(BV >= 0 && C >= 0 && H >= 0 && W >= 0, "negative extent");
  LT_REQUIRE(BV <= 65535, "B*V > 65535");
  if ((long long)BV * C * H * W == 0) return 0;
  LT_REQUIRE(src_chw && dst_hwc, "null pointer");
  LT_REQUIRE(al16(dst_hwc), "the hwcp buffer must be 16-byte aligned");
  return finish(__func__, lt::launch_feat_layout(src_chw, dst_hwc, BV, C, H, W, true, as_stream(stream)));
}

LT_API int lt_feat_to_chw(const float* src_hwc, float* dst_chw, int BV, int C, int H, int W, void* stream) {
  LT_REQUIRE(BV >= 0 && C >= 0 && H >= 0 && W >= 0, "negative extent");
  LT_REQUIRE(BV <= 65535, "B*V > 65535");
  if ((long long)BV * C * H * W == 0) return 0;
  LT_REQUIRE(src_hwc && dst_chw, "null pointer");
  LT_REQUIRE(al16(src_hwc), "the hwcp buffer must be 16-byte aligned");
  return finish(__func__, lt::launch_feat_layout(src_hwc, dst_chw, BV, C, H, W, false, as_stream(stream)));
}

LT_API int lt_unproject_fwd(const float* feat_hwc, const float* proj, const float* coord, const float* conf, float* out,
                     int B, int V, int C, int H, int W, int N0, int N1, int N2, int agg, int align_corners,
                     void* stream) {
  LT_REQUIRE(!bad_unproj_shape(B, V, C, H, W, N0, N1, N2), "bad shape");
  LT_REQUIRE(agg >= LT_AGG_SUM && agg <= LT_AGG_SOFTMAX, "unknown aggregation");
  if ((long long)B * C * N0 * N1 * N2 == 0) return 0;
  LT_REQUIRE(feat_hwc && proj && coord && out, "null pointer");
  LT_REQUIRE(al16(feat_hwc) && al16(proj), "feat_hwc and proj must be 16-byte aligned");
  LT_REQUIRE(agg != LT_AGG_CONF || conf, "conf aggregation needs vol_confidences");
  lt::UnprojShape s{B, V, C, H, W, N0, N1, N2, agg, align_corners ? 1 : 0};
  return finish(__func__, lt::launch_unproject_fwd(feat_hwc, proj, coord, conf, out, s, as_stream(stream)));
}

LT_API size_t lt_unproject_bwd_ws_bytes(int B, int V, int C, int N0, int N1, int N2) {
  if (B <= 0 || V <= 0 || C <= 0) return 0;
  return lt::unproject_bwd_ws_bytes(B, V, C, N0, N1, N2);
}

LT_API int lt_unproject_bwd(const float* grad_out, const float* feat_hwc, const float* proj, const float* coord,
                     const float* conf, float* grad_feat_hwc, float* grad_conf, void* conf_ws, int B, int V, int C,
                     int H, int W, int N0, int N1, int N2, int agg, int align_corners, void* stream) {
  LT_REQUIRE(!bad_unproj_shape(B, V, C, H, W, N0, N1, N2), "bad shape");
  LT_REQUIRE(agg >= LT_AGG_SUM && agg <= LT_AGG_SOFTMAX, "unknown aggregation");
  if ((long long)B * C == 0) return 0;
  LT_REQUIRE(feat_hwc && proj && coord && grad_feat_hwc, "null pointer");
  LT_REQUIRE(al16(feat_hwc) && al16(proj) && al16(grad_feat_hwc) && al16(conf_ws),
             "feat_hwc, proj, grad_feat_hwc and conf_ws must be 16-byte aligned");
  LT_REQUIRE((long long)N0 * N1 * N2 == 0 || grad_out, "null grad_out");
  LT_REQUIRE(agg != LT_AGG_CONF || conf, "conf aggregation needs vol_confidences");
  LT_REQUIRE(!(agg == LT_AGG_CONF && grad_conf) || conf_ws, "grad_conf needs conf_ws");
  lt::UnprojShape s{B, V, C, H, W, N0, N1, N2, agg, align_corners ? 1 : 0};
  return finish(__func__, lt::launch_unproject_bwd(grad_out, feat_hwc, proj, coord, conf, grad_feat_hwc, grad_conf,
                                                   conf_ws, s, as_stream(stream)));
}

LT_API size_t lt_softargmax_ws_bytes(int B, int J, long long nvox) {
  if (B <= 0 || J <= 0 || nvox <= 0) return 0;
  return lt::softargmax_ws_bytes(B, J, nvox);
}

LT_API int lt_softargmax_partial(const float* vol, const float* coord, float* stats, void* ws, int B, int J,
                          long long nvox, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE(B <= 65535, "B > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  LT_REQUIRE(stats && (nvox == 0 || (vol && coord && ws)), "null pointer");
  return finish(__func__, lt::launch_softargmax_partial(vol, coord, stats, ws, B, J, nvox, softmax ? 1 : 0,
                                                        multiplier, as_stream(stream)));
}

LT_API int lt_softargmax_finalize(const float* vol, const float* stats, float* coords, float* prob, int B, int J,
                           long long nvox, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE((long long)B * J <= 65535, "B*J > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  LT_REQUIRE(stats && (nvox == 0 || prob == nullptr || vol), "null pointer");
  return finish(__func__, lt::launch_softargmax_finalize(vol, stats, coords, nvox == 0 ? nullptr : prob, B, J, nvox,
                                                         softmax ? 1 : 0, multiplier, as_stream(stream)));
}

LT_API int lt_softargmax_fwd(const float* vol, const float* coord, float* coords, float* prob, float* stats, void* ws,
                      int B, int J, long long nvox, int softmax, float multiplier, void* stream) {
  // One persistent launch (softargmax.cu, k_sam_fwd_fused): the statistics of item b+1 run between the two passes of
  // item b, so the second read of the volumes is an L2 hit. (Running partial+finalize per L2-sized group of batch
  // items as separate launches was measured slower at every group size: the launches no longer fill 148 SMs.)
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE(B <= 65535 && (long long)B * J <= 65535, "B*J > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  LT_REQUIRE(stats && (nvox == 0 || (vol && coord && ws)), "null pointer");
  LT_REQUIRE(((uintptr_t)ws & 15) == 0, "ws must be 16-byte aligned");
  return finish(__func__, lt::launch_softargmax_fwd(vol, coord, coords, nvox == 0 ? nullptr : prob, stats, ws, B, J, nvox,
                                                    softmax ? 1 : 0, multiplier, as_stream(stream)));
}

LT_API int lt_softargmax_merge(const float* parts, float* merged, int R, int BJ, int softmax, void* stream) {
  LT_REQUIRE(R >= 0 && BJ >= 0, "negative extent");
  if (BJ == 0) return 0;
  LT_REQUIRE(merged && (R == 0 || parts), "null pointer");
  return finish(__func__, lt::launch_softargmax_merge(parts, merged, R, BJ, softmax ? 1 : 0, as_stream(stream)));
}

LT_API int lt_softargmax_bwd_dot(const float* prob, const float* grad_prob, float* dots, void* ws, int B, int J,
                          long long nvox, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE((long long)B * J <= 65535, "B*J > 65535");
  if (B * J == 0) return 0;
  LT_REQUIRE(dots && (nvox == 0 || (prob && grad_prob && ws)), "null pointer");
  return finish(__func__, lt::launch_softargmax_bwd_dot(prob, grad_prob, dots, ws, B, J, nvox, as_stream(stream)));
}

LT_API int lt_softargmax_bwd(const float* prob, const float* coord, const float* coords, const float* grad_coords,
                      const float* grad_prob, const float* dots, float* grad_vol, int B, int J, long long nvox,
                      int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE(B <= 65535, "B > 65535");
  if ((long long)B * J * nvox == 0) return 0;
  LT_REQUIRE(prob && coord && coords && grad_coords && grad_vol, "null pointer");
  LT_REQUIRE(!(softmax && grad_prob) || dots, "grad_prob given without dots");
  return finish(__func__, lt::launch_softargmax_bwd(prob, coord, coords, grad_coords, grad_prob, dots, grad_vol, B, J,
                                                    nvox, softmax ? 1 : 0, multiplier, as_stream(stream)));
}

LT_API int lt_softargmax2d_fwd(const float* hm, float* coords, float* prob, int BJ, int H, int W, int softmax,
                        float multiplier, void* stream) {
  LT_REQUIRE(BJ >= 0 && H >= 0 && W >= 0, "negative extent");
  LT_REQUIRE((long long)H * W <= 0x7fffffffLL, "map too large");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (BJ == 0) return 0;
  LT_REQUIRE(coords && (H * W == 0 || (hm && prob)), "null pointer");
  return finish(__func__, lt::launch_softargmax2d_fwd(hm, coords, prob, BJ, H, W, softmax ? 1 : 0, multiplier,
                                                      as_stream(stream)));
}

LT_API int lt_softargmax2d_bwd(const float* prob, const float* coords, const float* grad_coords, const float* grad_prob,
                        float* grad_hm, int BJ, int H, int W, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(BJ >= 0 && H >= 0 && W >= 0, "negative extent");
  LT_REQUIRE((long long)H * W <= 0x7fffffffLL, "map too large");
  if ((long long)BJ * H * W == 0) return 0;
  LT_REQUIRE(prob && coords && grad_coords && grad_hm, "null pointer");
  return finish(__func__, lt::launch_softargmax2d_bwd(prob, coords, grad_coords, grad_prob, grad_hm, BJ, H, W,
                                                      softmax ? 1 : 0, multiplier, as_stream(stream)));
}

LT_API int lt_coord_volumes(const float* spec, float* out, int B, int N0, int N1, int N2, int transfer, void* stream) {
  LT_REQUIRE(B >= 0 && N0 >= 0 && N1 >= 0 && N2 >= 0, "negative extent");
  LT_REQUIRE(B <= 65535, "B > 65535");
  LT_REQUIRE(!transfer || N1 == N2, "the CMU->Human3.6M axis swap needs N1 == N2 (triangulation.py:337)");
  LT_REQUIRE((long long)N0 * N1 * N2 <= 0x7fffffffLL, "more than 2^31 voxels per item");
  if ((long long)B * N0 * N1 * N2 == 0) return 0;
  LT_REQUIRE(spec && out, "null pointer");
  return finish(__func__, lt::launch_coord_volumes(spec, out, B, N0, N1, N2, transfer ? 1 : 0, as_stream(stream)));
}

// ---- the same ops with the coordinates computed in the kernel from the K6 spec (coordspec.cuh) ----
static bool bad_spec_volume(int N0, int N1, int N2, int transfer) {
  return N0 < 0 || N1 < 0 || N2 < 0 || (long long)N0 * N1 * N2 > 0x7fffffffLL || (transfer && N1 != N2);
}

LT_API int lt_unproject_fwd_spec(const float* feat_hwc, const float* proj, const float* coord_spec, int transfer,
                          const float* conf, float* out, int B, int V, int C, int H, int W, int N0, int N1, int N2,
                          int agg, int align_corners, void* stream) {
  LT_REQUIRE(!bad_unproj_shape(B, V, C, H, W, N0, N1, N2), "bad shape");
  LT_REQUIRE(!bad_spec_volume(N0, N1, N2, transfer), "more than 2^31 voxels per item, or transfer with N1 != N2");
  LT_REQUIRE(agg >= LT_AGG_SUM && agg <= LT_AGG_SOFTMAX, "unknown aggregation");
  if ((long long)B * C * N0 * N1 * N2 == 0) return 0;
  LT_REQUIRE(feat_hwc && proj && coord_spec && out, "null pointer");
  LT_REQUIRE(al16(feat_hwc) && al16(proj), "feat_hwc and proj must be 16-byte aligned");
  LT_REQUIRE(agg != LT_AGG_CONF || conf, "conf aggregation needs vol_confidences");
  lt::UnprojShape s{B, V, C, H, W, N0, N1, N2, agg, align_corners ? 1 : 0};
  const lt::CoordSpecArg cs{coord_spec, N1, N2, transfer ? 1 : 0};
  return finish(__func__, lt::launch_unproject_fwd(feat_hwc, proj, nullptr, conf, out, s, as_stream(stream), &cs));
}

LT_API int lt_unproject_bwd_spec(const float* grad_out, const float* feat_hwc, const float* proj, const float* coord_spec,
                          int transfer, const float* conf, float* grad_feat_hwc, float* grad_conf, void* conf_ws,
                          int B, int V, int C, int H, int W, int N0, int N1, int N2, int agg, int align_corners,
                          void* stream) {
  LT_REQUIRE(!bad_unproj_shape(B, V, C, H, W, N0, N1, N2), "bad shape");
  LT_REQUIRE(!bad_spec_volume(N0, N1, N2, transfer), "more than 2^31 voxels per item, or transfer with N1 != N2");
  LT_REQUIRE(agg >= LT_AGG_SUM && agg <= LT_AGG_SOFTMAX, "unknown aggregation");
  if ((long long)B * C == 0) return 0;
  LT_REQUIRE(feat_hwc && proj && coord_spec && grad_feat_hwc, "null pointer");
  LT_REQUIRE(al16(feat_hwc) && al16(proj) && al16(grad_feat_hwc) && al16(conf_ws),
             "feat_hwc, proj, grad_feat_hwc and conf_ws must be 16-byte aligned");
  LT_REQUIRE((long long)N0 * N1 * N2 == 0 || grad_out, "null grad_out");
  LT_REQUIRE(agg != LT_AGG_CONF || conf, "conf aggregation needs vol_confidences");
  LT_REQUIRE(!(agg == LT_AGG_CONF && grad_conf) || conf_ws, "grad_conf needs conf_ws");
  lt::UnprojShape s{B, V, C, H, W, N0, N1, N2, agg, align_corners ? 1 : 0};
  const lt::CoordSpecArg cs{coord_spec, N1, N2, transfer ? 1 : 0};
  return finish(__func__, lt::launch_unproject_bwd(grad_out, feat_hwc, proj, nullptr, conf, grad_feat_hwc, grad_conf,
                                                   conf_ws, s, as_stream(stream), &cs));
}

LT_API int lt_softargmax_partial_spec(const float* vol, const float* coord_spec, int transfer, float* stats, void* ws,
                               int B, int J, int N0, int N1, int N2, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0, "negative extent");
  LT_REQUIRE(!bad_spec_volume(N0, N1, N2, transfer), "more than 2^31 voxels per item, or transfer with N1 != N2");
  LT_REQUIRE(B <= 65535, "B > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  const long long nvox = (long long)N0 * N1 * N2;
  LT_REQUIRE(stats && (nvox == 0 || (vol && coord_spec && ws)), "null pointer");
  const lt::CoordSpecArg cs{coord_spec, N1, N2, transfer ? 1 : 0};
  return finish(__func__, lt::launch_softargmax_partial(vol, nullptr, stats, ws, B, J, nvox, softmax ? 1 : 0,
                                                        multiplier, as_stream(stream), &cs));
}

LT_API int lt_softargmax_fwd_spec(const float* vol, const float* coord_spec, int transfer, float* coords, float* prob,
                           float* stats, void* ws, int B, int J, int N0, int N1, int N2, int softmax,
                           float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0, "negative extent");
  LT_REQUIRE(!bad_spec_volume(N0, N1, N2, transfer), "more than 2^31 voxels per item, or transfer with N1 != N2");
  LT_REQUIRE(B <= 65535 && (long long)B * J <= 65535, "B*J > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  const long long nvox = (long long)N0 * N1 * N2;
  LT_REQUIRE(stats && (nvox == 0 || (vol && coord_spec && ws)), "null pointer");
  LT_REQUIRE(((uintptr_t)ws & 15) == 0, "ws must be 16-byte aligned");
  const lt::CoordSpecArg cs{coord_spec, N1, N2, transfer ? 1 : 0};
  return finish(__func__, lt::launch_softargmax_fwd(vol, nullptr, coords, nvox == 0 ? nullptr : prob, stats, ws, B, J,
                                                    nvox, softmax ? 1 : 0, multiplier, as_stream(stream), &cs));
}

LT_API int lt_softargmax_bwd_spec(const float* prob, const float* coord_spec, int transfer, const float* coords,
                           const float* grad_coords, const float* grad_prob, const float* dots, float* grad_vol,
                           int B, int J, int N0, int N1, int N2, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0, "negative extent");
  LT_REQUIRE(!bad_spec_volume(N0, N1, N2, transfer), "more than 2^31 voxels per item, or transfer with N1 != N2");
  LT_REQUIRE(B <= 65535, "B > 65535");
  const long long nvox = (long long)N0 * N1 * N2;
  if ((long long)B * J * nvox == 0) return 0;
  LT_REQUIRE(prob && coord_spec && coords && grad_coords && grad_vol, "null pointer");
  LT_REQUIRE(!(softmax && grad_prob) || dots, "grad_prob given without dots");
  const lt::CoordSpecArg cs{coord_spec, N1, N2, transfer ? 1 : 0};
  return finish(__func__, lt::launch_softargmax_bwd(prob, nullptr, coords, grad_coords, grad_prob, dots, grad_vol, B, J,
                                                    nvox, softmax ? 1 : 0, multiplier, as_stream(stream), &cs));
}

LT_API size_t lt_volce_ws_bytes(int B, int J, long long nvox) { return lt::volce_ws_bytes(B, J, nvox); }

LT_API int lt_volce_fwd(const float* coord, const float* prob, const float* kp_gt, const float* validity, float* loss,
                 int* idx, void* ws, int B, int J, long long nvox, void* stream) {
  LT_REQUIRE(B >= 1 && J >= 1 && nvox >= 1, "need B, J, nvox >= 1 (the reference divides by B*J)");
  LT_REQUIRE(B <= 65535, "B > 65535");
  LT_REQUIRE(nvox <= 0x7fffffffLL, "nvox does not fit the int32 index");
  LT_REQUIRE(coord && prob && kp_gt && validity && loss && idx && ws, "null pointer");
  return finish(__func__, lt::launch_volce_fwd(coord, prob, kp_gt, validity, loss, idx, ws, B, J, nvox,
                                               as_stream(stream)));
}

LT_API int lt_volce_bwd(const float* prob, const int* idx, const float* validity, const float* grad_loss,
                 float* grad_prob, int B, int J, long long nvox, void* stream) {
  LT_REQUIRE(B >= 1 && J >= 1 && nvox >= 1, "need B, J, nvox >= 1");
  LT_REQUIRE(prob && idx && validity && grad_loss && grad_prob, "null pointer");
  return finish(__func__, lt::launch_volce_bwd(prob, idx, validity, grad_loss, grad_prob, B, J, nvox,
                                               as_stream(stream)));
}

LT_API int lt_volce_bwd_sparse(const float* prob, const int* idx, const float* validity, const float* grad_loss,
                        float* vals, float* dots, int B, int J, long long nvox, void* stream) {
  LT_REQUIRE(B >= 1 && J >= 1 && nvox >= 1, "need B, J, nvox >= 1");
  LT_REQUIRE(prob && idx && validity && grad_loss && vals && dots, "null pointer");
  return finish(__func__, lt::launch_volce_bwd_sparse(prob, idx, validity, grad_loss, vals, dots, B, J, nvox,
                                                      as_stream(stream)));
}

LT_API int lt_softargmax_bwd_add_sparse(const float* prob, const int* idx, const float* vals, float* grad_vol, int B,
                                 int J, long long nvox, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 1 && J >= 1 && nvox >= 1, "need B, J, nvox >= 1");
  LT_REQUIRE(prob && idx && vals && grad_vol, "null pointer");
  return finish(__func__, lt::launch_softargmax_bwd_add_sparse(prob, idx, vals, grad_vol, B, J, nvox, softmax ? 1 : 0,
                                                               multiplier, as_stream(stream)));
}

LT_API size_t lt_pwconv_ws_bytes(int BV, int Cin, int Cout, int H, int W) {
  if (BV <= 0 || Cin <= 0 || Cout <= 0 || H <= 0 || W <= 0) return 0;
  return lt::pwconv_ws_bytes(BV, Cin, Cout, H, W);
}

static bool bad_pwconv_shape(int BV, int Cin, int Cout, int H, int W) {
  return BV < 0 || Cin < 1 || Cout < 1 || H < 1 || W < 1 || BV > 65535 || (long long)H * W > 0x3fffffffLL ||
         (long long)BV * ((long long)H * W / 16 + 1) > 0x7fffffffLL;
}

LT_API int lt_pwconv_fwd(const float* x, const float* weight, const float* bias, float* out_hwcp, void* ws, int BV,
                         int Cin, int Cout, int H, int W, void* stream) {
  LT_REQUIRE(!bad_pwconv_shape(BV, Cin, Cout, H, W), "bad shape");
  if (BV == 0) return 0;
  LT_REQUIRE(x && weight && out_hwcp && ws, "null pointer");
  LT_REQUIRE(((uintptr_t)ws & 15) == 0 && ((uintptr_t)out_hwcp & 15) == 0, "ws / out_hwcp must be 16-byte aligned");
  return finish(__func__, lt::launch_pwconv_fwd(x, weight, bias, out_hwcp, ws, BV, Cin, Cout, H, W, as_stream(stream)));
}

LT_API int lt_pwconv_bwd_data(const float* grad_hwcp, const float* weight, float* grad_x, int BV, int Cin, int Cout,
                              int H, int W, void* stream) {
  LT_REQUIRE(!bad_pwconv_shape(BV, Cin, Cout, H, W), "bad shape");
  if (BV == 0) return 0;
  LT_REQUIRE(grad_hwcp && weight && grad_x, "null pointer");
  LT_REQUIRE(((uintptr_t)grad_hwcp & 15) == 0, "grad_hwcp must be 16-byte aligned");
  return finish(__func__, lt::launch_pwconv_bwd_data(grad_hwcp, weight, grad_x, BV, Cin, Cout, H, W, as_stream(stream)));
}

LT_API int lt_pwconv_bwd_weight(const float* x, const float* grad_hwcp, float* grad_weight, float* grad_bias, void* ws,
                                int BV, int Cin, int Cout, int H, int W, void* stream) {
  LT_REQUIRE(!bad_pwconv_shape(BV, Cin, Cout, H, W), "bad shape");
  LT_REQUIRE(grad_weight, "null pointer");
  if (BV == 0) {  // an empty batch has a zero gradient
    cudaError_t e = cudaMemsetAsync(grad_weight, 0, (size_t)Cin * Cout * 4, as_stream(stream));
    if (e == cudaSuccess && grad_bias) e = cudaMemsetAsync(grad_bias, 0, (size_t)Cout * 4, as_stream(stream));
    return finish(__func__, e);
  }
  LT_REQUIRE(x && grad_hwcp && ws, "null pointer");
  LT_REQUIRE(((uintptr_t)ws & 15) == 0 && ((uintptr_t)grad_hwcp & 15) == 0, "ws / grad_hwcp must be 16-byte aligned");
  return finish(__func__, lt::launch_pwconv_bwd_weight(x, grad_hwcp, grad_weight, grad_bias, ws, BV, Cin, Cout, H, W,
                                                       as_stream(stream)));
}

LT_API int lt_dlt_fwd(const float* proj, const float* pts, const float* conf, float* out, double* saved, int B, int V,
               int J, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0, "negative extent");
  LT_REQUIRE(V >= 1 && V <= LT_DLT_MAX_VIEWS, "need 1 <= V <= LT_DLT_MAX_VIEWS");
  if (B * J == 0) return 0;
  LT_REQUIRE(proj && pts && out, "null pointer");
  return finish(__func__, lt::launch_dlt_fwd(proj, pts, conf, out, saved, B, V, J, as_stream(stream)));
}

LT_API int lt_dlt_bwd(const float* grad_out, const float* proj, const float* pts, const float* conf, const double* saved,
               float* grad_pts, float* grad_conf, int B, int V, int J, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0, "negative extent");
  LT_REQUIRE(V >= 1 && V <= LT_DLT_MAX_VIEWS, "need 1 <= V <= LT_DLT_MAX_VIEWS");
  if (B * J == 0) return 0;
  LT_REQUIRE(grad_out && proj && pts && saved && grad_pts, "null pointer");
  return finish(__func__, lt::launch_dlt_bwd(grad_out, proj, pts, conf, saved, grad_pts, grad_conf, B, V, J,
                                             as_stream(stream)));
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// lt_softargmax_fwd_noalloc — K2 forward for callers that cannot allocate (CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc). No counterpart in the reference.
//
// lt_softargmax_fwd needs two caller-owned scratch buffers (the per-warp tuples `ws` and the merged `stats`); here both
// live in ONE file-scope device array of LT_NOALLOC_SCRATCH_BYTES that every GPU loading this library holds until exit.
// `g_noalloc_scratch` below is, in host code, only a stand-in for the symbol: its device address is taken once per
// device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive. The kernels are
// the ones lt_softargmax_fwd launches (lt::launch_softargmax_fwd), so results are bit-identical to it.
//
// Nothing in the scratch must start at zero on the default (split) path: k_sam_stats writes every tuple k_sam_combine
// reads, k_sam_combine writes every stats row k_sam_final reads; the opt-in fused path (LT_B200_K2FWD=fused) clears its
// own control words with cudaMemsetAsync on the caller's stream. All calls share the array and find the previous
// call's contents in it: calls that overlap on the GPU (different streams without an ordering between them, or two host
// threads on one device) RACE on it. Calls on one stream, or ordered by events, are safe.
// No cudaMalloc / cudaFree / allocating library routine is called here or below.
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[LT_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kNoallocMaxDevices = 64;

// device address of g_noalloc_scratch on the CURRENT device (one cudaGetSymbolAddress per device and process)
cudaError_t noalloc_scratch(unsigned char** out) {
  static void* ptrs[kNoallocMaxDevices] = {};  // written once per device; a racing second writer stores the same value
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kNoallocMaxDevices) return cudaErrorInvalidDevice;
  void* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    if (p == nullptr) return cudaErrorInvalidSymbol;
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
  }
  *out = static_cast<unsigned char*>(p);
  return cudaSuccess;
}

}  // namespace

extern "C" {

LT_API int lt_softargmax_fwd_noalloc(const float* vol, const float* coord, float* coords, float* prob, int B, int J,
                                     long long nvox, int softmax, float multiplier, void* stream) {
  LT_REQUIRE(B >= 0 && J >= 0 && nvox >= 0, "negative extent");
  LT_REQUIRE(B <= 65535 && (long long)B * J <= 65535, "B*J > 65535");
  LT_REQUIRE(multiplier > 0.f, "multiplier must be > 0 (scale on the caller side otherwise)");
  if (B * J == 0) return 0;
  LT_REQUIRE(coords && (nvox == 0 || (vol && coord && prob)), "null pointer");
  // layout of the static array: [stats (B,J,5) fp32, rounded up to 256 bytes][ws of lt_softargmax_ws_bytes]
  const size_t stats_bytes = ((size_t)B * J * 5 * sizeof(float) + 255) / 256 * 256;
  const size_t need = stats_bytes + lt::softargmax_ws_bytes(B, J, nvox);
  if (need > (size_t)LT_NOALLOC_SCRATCH_BYTES) {
    snprintf(g_err, sizeof(g_err),
             "lt_softargmax_fwd_noalloc: B=%d J=%d nvox=%lld needs %zu bytes of scratch, the static array holds "
             "LT_NOALLOC_SCRATCH_BYTES = %zu; use lt_softargmax_fwd with caller-owned scratch",
             B, J, nvox, need, (size_t)LT_NOALLOC_SCRATCH_BYTES);
    return LT_ERR_INVALID;
  }
  unsigned char* base = nullptr;
  cudaError_t e = noalloc_scratch(&base);
  if (e != cudaSuccess) return finish(__func__, e);
  float* stats = reinterpret_cast<float*>(base);
  void* ws = base + stats_bytes;
  return finish(__func__, lt::launch_softargmax_fwd(vol, coord, coords, nvox == 0 ? nullptr : prob, stats, ws, B, J,
                                                    nvox, softmax ? 1 : 0, multiplier, as_stream(stream)));
}

}  // extern "C"
