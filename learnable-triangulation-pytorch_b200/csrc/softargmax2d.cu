// softargmax2d.cu — K4: 2-D soft-argmax of joint heat maps, forward and backward.
//
// Replaces mvn/utils/op.py:11-47 `integrate_tensor_2d`, the step right before the DLT in AlgebraicTriangulationNet
// (mvn/models/triangulation.py:164): softmax over each (h*w) map (or ReLU), marginal masses times arange(w) / arange(h)
// -> (x, y) in heat-map pixels, and the normalised maps as second output. In ReLU mode the coordinates ARE divided by
// the total mass (op.py:39-41) while the returned maps are not normalised (op.py:26).
// One CTA per map: a 96x96 map is 36 KB, so the second and third sweep are served by L1/L2; HBM sees 1 read + 1 write.
// `multiplier` folds the `heatmaps * heatmap_multiplier` of triangulation.py:164 into the exponent.
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kS2Threads = 256;

__device__ __forceinline__ float block_max(float v, float* s_red) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = s_red[0];
#pragma unroll
  for (int w = 1; w < kS2Threads / 32; ++w) r = fmaxf(r, s_red[w]);
  __syncthreads();
  return r;
}
// fixed-order block sum (deterministic)
__device__ __forceinline__ float block_sum(float v, float* s_red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = s_red[0];
#pragma unroll
  for (int w = 1; w < kS2Threads / 32; ++w) r += s_red[w];
  __syncthreads();
  return r;
}

template <bool SOFTMAX>
__global__ void __launch_bounds__(kS2Threads) k_sam2d_fwd(const float* __restrict__ hm, float* __restrict__ coords,
                                                          float* __restrict__ prob, int H, int W, float mult) {
  __shared__ float s_red[kS2Threads / 32];
  const int n = H * W;
  const float* row = hm + (size_t)blockIdx.x * n;
  float* orow = prob + (size_t)blockIdx.x * n;
  float m = 0.f;
  if (SOFTMAX) {
    float lm = -INFINITY;
    for (int i = threadIdx.x; i < n; i += kS2Threads) lm = fmaxf(lm, __ldg(row + i) * mult);
    m = block_max(lm, s_red);
  }
  const float L = mult * kLog2eF, shift = -m * kLog2eF;
  float s = 0.f, sx = 0.f, sy = 0.f;
  for (int i = threadIdx.x; i < n; i += kS2Threads) {
    const float x = __ldg(row + i);
    const float e = SOFTMAX ? fast_ex2(fmaf(x, L, shift)) : fmaxf(x * mult, 0.f);
    const int py = i / W, px = i - py * W;
    s += e;
    sx = fmaf(e, (float)px, sx);
    sy = fmaf(e, (float)py, sy);
    orow[i] = e;
  }
  s = block_sum(s, s_red);
  sx = block_sum(sx, s_red);
  sy = block_sum(sy, s_red);
  if (threadIdx.x == 0) {
    coords[(size_t)blockIdx.x * 2 + 0] = __fdiv_rn(sx, s);  // softmax: sum p*x ; relu: op.py:39-41 (0/0 = nan like torch)
    coords[(size_t)blockIdx.x * 2 + 1] = __fdiv_rn(sy, s);
  }
  if (SOFTMAX) {
    const float inv = __fdiv_rn(1.f, s);
    for (int i = threadIdx.x; i < n; i += kS2Threads) orow[i] *= inv;  // each thread re-reads its own writes
  }
}

//   softmax: g_i = mult * p_i * (G_i - dot),  G_i = gp_i + gx*x_i + gy*y_i,  dot = sum_k p_k G_k
//   relu   : g_i = mult * [p_i > 0] * (gp_i + (gx*(x_i - X) + gy*(y_i - Y)) / S),  S = sum_k p_k
template <bool SOFTMAX, bool HAS_GP>
__global__ void __launch_bounds__(kS2Threads) k_sam2d_bwd(const float* __restrict__ prob, const float* __restrict__ coords,
                                                          const float* __restrict__ gcoords,
                                                          const float* __restrict__ gprob, float* __restrict__ ghm,
                                                          int H, int W, float mult) {
  const int n = H * W;
  const float* prow = prob + (size_t)blockIdx.x * n;
  const float* grow = HAS_GP ? gprob + (size_t)blockIdx.x * n : nullptr;
  float* orow = ghm + (size_t)blockIdx.x * n;
  const float gx = __ldg(gcoords + (size_t)blockIdx.x * 2), gy = __ldg(gcoords + (size_t)blockIdx.x * 2 + 1);
  const float X = __ldg(coords + (size_t)blockIdx.x * 2), Y = __ldg(coords + (size_t)blockIdx.x * 2 + 1);
  // softmax: dot = sum_k p_k G_k, accumulated in fp64 over the saved probabilities. With heatmap_multiplier = 100 the
  // maps are nearly one-hot and G_i - dot cancels to ~1e-6 of its operands at the peak: forming it from the fp32
  // coordinates (gx*X + gy*Y) loses the gradient of every other pixel in that rounding error.
  // relu: S = sum_k p_k.
  // The saved probabilities sum to 1 only up to fp32 rounding; dividing by their fp64 sum makes `dot` the weighted mean
  // of exactly these p_k, so the peak's G - dot is sum_{k != peak} p_k (G_peak - G_k) with RELATIVE error only.
  __shared__ double s_red64[2][kS2Threads / 32];
  double red = 0.0, mass = 0.0;
  for (int i = threadIdx.x; i < n; i += kS2Threads) {
    const float p = __ldg(prow + i);
    mass += (double)p;
    if (SOFTMAX) {
      const int py = i / W, px = i - py * W;
      const double G = (double)(HAS_GP ? __ldg(grow + i) : 0.f) + (double)gx * px + (double)gy * py;
      red = fma((double)p, G, red);
    }
  }
  red = warp_sum(red);
  mass = warp_sum(mass);
  if ((threadIdx.x & 31) == 0) { s_red64[0][threadIdx.x >> 5] = red; s_red64[1][threadIdx.x >> 5] = mass; }
  __syncthreads();
  red = s_red64[0][0]; mass = s_red64[1][0];
#pragma unroll
  for (int w = 1; w < kS2Threads / 32; ++w) { red += s_red64[0][w]; mass += s_red64[1][w]; }
  const double dot = SOFTMAX ? red / mass : 0.0;
  const float invS = SOFTMAX ? 0.f : (float)(1.0 / mass);
  for (int i = threadIdx.x; i < n; i += kS2Threads) {
    const float p = __ldg(prow + i);
    const float gp = HAS_GP ? __ldg(grow + i) : 0.f;
    const int py = i / W, px = i - py * W;
    float g;
    if (SOFTMAX) g = mult * (p * (float)(((double)gp + (double)gx * px + (double)gy * py) - dot));
    else g = (p > 0.f) ? mult * (gp + (gx * ((float)px - X) + gy * ((float)py - Y)) * invS) : 0.f;
    orow[i] = g;
  }
}

cudaError_t launch_softargmax2d_fwd(const float* hm, float* coords, float* prob, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st) {
  if (BJ == 0) return cudaSuccess;
  if (softmax) k_sam2d_fwd<true><<<BJ, kS2Threads, 0, st>>>(hm, coords, prob, H, W, mult);
  else k_sam2d_fwd<false><<<BJ, kS2Threads, 0, st>>>(hm, coords, prob, H, W, mult);
  return cudaGetLastError();
}

cudaError_t launch_softargmax2d_bwd(const float* prob, const float* coords, const float* grad_coords,
                                    const float* grad_prob, float* grad_hm, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st) {
  if (BJ == 0 || H * W == 0) return cudaSuccess;
  const bool gp = grad_prob != nullptr;
#define LT_GO(S_, G_) k_sam2d_bwd<S_, G_><<<BJ, kS2Threads, 0, st>>>(prob, coords, grad_coords, grad_prob, grad_hm, H, W, mult)
  if (softmax) { if (gp) LT_GO(true, true); else LT_GO(true, false); }
  else         { if (gp) LT_GO(false, true); else LT_GO(false, false); }
#undef LT_GO
  return cudaGetLastError();
}

}  // namespace lt
