// kernels.h — internal C++ launch interface between abi.cu and the kernel translation units.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace lt {

constexpr int kAggSum = 0, kAggConf = 1, kAggMax = 2, kAggSoftmax = 3;  // == enum lt_agg

// coordinates computed in the consuming kernel from the per-item 18-float spec of K6 instead of read from the tensor
// (coordspec.cuh); N1, N2 = the two inner volume extents, transfer = the CMU -> Human3.6M axis swap
struct CoordSpecArg {
  const float* spec;  // (B, 18) device pointer
  int N1, N2, transfer;
};

int padded_channels(int C);
int sm_count();  // SMs of the current device (148 on B200), queried once per device
size_t feat_layout_elems(int BV, int C, int H, int W);
cudaError_t launch_feat_layout(const float* src, float* dst, int BV, int C, int H, int W, bool to_hwcp,
                               cudaStream_t st);

struct UnprojShape {
  int B, V, C, H, W, N0, N1, N2, agg, align_corners;
};
cudaError_t launch_unproject_fwd(const float* feat_hwc, const float* proj, const float* coord,
                                 const float* conf, float* out, const UnprojShape& s, cudaStream_t st,
                                 const CoordSpecArg* spec = nullptr);
size_t unproject_bwd_ws_bytes(int B, int V, int C, int N0, int N1, int N2);
cudaError_t launch_unproject_bwd(const float* grad_out, const float* feat_hwc, const float* proj,
                                 const float* coord, const float* conf, float* grad_feat_hwc,
                                 float* grad_conf, void* conf_ws, const UnprojShape& s, cudaStream_t st,
                                 const CoordSpecArg* spec = nullptr);

// on-chip merged formulation of the backward (unproject_bwd_tile.cu); UnprojArgs lives in unproject.cuh
struct UnprojArgs;
bool unproject_bwd_tile_eligible(const UnprojShape& s, const float* grad_out);
size_t unproject_bwd_tile_ws_bytes(const UnprojShape& s);
cudaError_t launch_conf_reduce_rows(const float* part, float* grad_conf, int B, int VC, int nrows, cudaStream_t st);
cudaError_t launch_unproject_bwd_tile(UnprojArgs& a, const UnprojShape& s, float* grad_conf, cudaStream_t st,
                                      const CoordSpecArg* spec);

// persistent TMA-staged formulation of the forward (unproject_fwd_tma.cu): 32-channel features, <= 4 views, N2 % 8 == 0
bool unproject_fwd_tma_eligible(const UnprojArgs& a);
cudaError_t launch_unproject_fwd_tma(const UnprojArgs& a, const CoordSpecArg* spec, int agg, cudaStream_t st);

size_t softargmax_ws_bytes(int B, int J, long long nvox);
cudaError_t launch_softargmax_partial(const float* vol, const float* coord, float* stats, void* ws, int B,
                                      int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                      const CoordSpecArg* spec = nullptr);
cudaError_t launch_softargmax_finalize(const float* vol, const float* stats, float* coords, float* prob,
                                       int B, int J, long long nvox, int softmax, float mult,
                                       cudaStream_t st);
cudaError_t launch_softargmax_fwd(const float* vol, const float* coord, float* coords, float* prob, float* stats,
                                  void* ws, int B, int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                  const CoordSpecArg* spec = nullptr);
// one-read forward: a thread-block cluster keeps one (b, joint) volume in shared memory (softargmax_cluster.cu)
int softargmax_cluster_size(long long nvox);   // 0 = does not apply
cudaError_t launch_softargmax_fwd_cluster(const float* vol, const float* coord, float* coords, float* prob, float* stats,
                                          int B, int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                          const CoordSpecArg* spec);
cudaError_t launch_softargmax_merge(const float* parts, float* merged, int R, int BJ, int softmax,
                                    cudaStream_t st);
cudaError_t launch_softargmax_bwd_dot(const float* prob, const float* grad_prob, float* dots, void* ws,
                                      int B, int J, long long nvox, cudaStream_t st);
cudaError_t launch_softargmax_bwd(const float* prob, const float* coord, const float* coords,
                                  const float* grad_coords, const float* grad_prob, const float* dots,
                                  float* grad_vol, int B, int J, long long nvox, int softmax, float mult,
                                  cudaStream_t st, const CoordSpecArg* spec = nullptr);

cudaError_t launch_coord_volumes(const float* spec, float* out, int B, int N0, int N1, int N2, int transfer,
                                 cudaStream_t st);
size_t volce_ws_bytes(int B, int J, long long nvox);
cudaError_t launch_volce_fwd(const float* coord, const float* prob, const float* kp, const float* validity, float* loss,
                             int* idx_out, void* ws, int B, int J, long long nvox, cudaStream_t st);
cudaError_t launch_volce_bwd_sparse(const float* prob, const int* idx, const float* validity, const float* gloss,
                                    float* vals, float* dots, int B, int J, long long nvox, cudaStream_t st);
cudaError_t launch_softargmax_bwd_add_sparse(const float* prob, const int* idx, const float* vals, float* gvol, int B,
                                             int J, long long nvox, int softmax, float mult, cudaStream_t st);
cudaError_t launch_volce_bwd(const float* prob, const int* idx, const float* validity, const float* gloss, float* gvol,
                             int B, int J, long long nvox, cudaStream_t st);
cudaError_t launch_softargmax2d_fwd(const float* hm, float* coords, float* prob, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st);
cudaError_t launch_softargmax2d_bwd(const float* prob, const float* coords, const float* grad_coords,
                                    const float* grad_prob, float* grad_hm, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st);

size_t pwconv_ws_bytes(int BV, int Cin, int Cout, int H, int W);
cudaError_t launch_pwconv_fwd(const float* x, const float* weight, const float* bias, float* out_hwcp, void* ws, int BV,
                              int Cin, int Cout, int H, int W, cudaStream_t st);
cudaError_t launch_pwconv_bwd_data(const float* grad_hwcp, const float* weight, float* grad_x, int BV, int Cin,
                                   int Cout, int H, int W, cudaStream_t st);
cudaError_t launch_pwconv_bwd_weight(const float* x, const float* grad_hwcp, float* grad_weight, float* grad_bias,
                                     void* ws, int BV, int Cin, int Cout, int H, int W, cudaStream_t st);

cudaError_t launch_dlt_fwd(const float* proj, const float* pts, const float* conf, float* out,
                           double* saved, int B, int V, int J, cudaStream_t st);
cudaError_t launch_dlt_bwd(const float* grad_out, const float* proj, const float* pts, const float* conf,
                           const double* saved, float* grad_pts, float* grad_conf, int B, int V, int J,
                           cudaStream_t st);

}  // namespace lt
