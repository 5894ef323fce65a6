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

// ------------------------------------------------------------------------------------------------------------------
// Fast path (W % 4 == 0, 16-byte aligned rows, map fits in shared memory): the map is read from HBM ONCE as float4
// into shared memory, every later sweep runs on that copy, and the result leaves as float4. A thread always re-reads
// the float4 slots it wrote itself, so the only barriers are the block reductions. (py, px) advance incrementally:
// no integer division in the loops (the scalar kernels above spend most of their issue slots on i / W).
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T block_sum_n(T v, T* s_red) {  // fixed order, any blockDim.x that is a multiple of 32
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
  __syncthreads();
  T r = s_red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) r += s_red[w];
  __syncthreads();
  return r;
}

struct RowWalk {  // position of float4 slot i4 inside an H x W map with W4 = W/4 slots per row
  int py, px4, dpy, dpx4, W4;
  __device__ __forceinline__ RowWalk(int i4, int stride4, int W4_) : W4(W4_) {
    py = i4 / W4; px4 = i4 - py * W4;
    dpy = stride4 / W4; dpx4 = stride4 - dpy * W4;
  }
  __device__ __forceinline__ void next() {
    py += dpy; px4 += dpx4;
    if (px4 >= W4) { px4 -= W4; ++py; }
  }
};

template <bool SOFTMAX>
__global__ void __launch_bounds__(1024) k_sam2d_fwd_v4(const float* __restrict__ hm, float* __restrict__ coords,
                                                       float* __restrict__ prob, int H, int W, float mult) {
  extern __shared__ __align__(16) float4 s_map[];
  __shared__ float s_red[32];
  const int n4 = (H * W) >> 2, W4 = W >> 2, nthr = blockDim.x;
  const float4* row = reinterpret_cast<const float4*>(hm + (size_t)blockIdx.x * H * W);
  float4* orow = reinterpret_cast<float4*>(prob + (size_t)blockIdx.x * H * W);
  float lm = -INFINITY;
  for (int i = threadIdx.x; i < n4; i += nthr) {
    const float4 v = ldg128_stream(reinterpret_cast<const float*>(row + i));
    s_map[i] = v;
    lm = fmaxf(fmaxf(lm, fmaxf(v.x * mult, v.y * mult)), fmaxf(v.z * mult, v.w * mult));
  }
  float m = 0.f;
  if (SOFTMAX) {
    lm = warp_max(lm);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = lm;
    __syncthreads();
    m = s_red[0];
    for (int w = 1; w < (nthr >> 5); ++w) m = fmaxf(m, s_red[w]);
    __syncthreads();
  }
  const float L = mult * kLog2eF, shift = -m * kLog2eF;  // same arithmetic as the scalar kernel
  float s = 0.f, sx = 0.f, sy = 0.f;
  RowWalk rw(threadIdx.x, nthr, W4);
  for (int i = threadIdx.x; i < n4; i += nthr, rw.next()) {
    float4 v = s_map[i];
    if (SOFTMAX) {
      v.x = fast_ex2(fmaf(v.x, L, shift)); v.y = fast_ex2(fmaf(v.y, L, shift));
      v.z = fast_ex2(fmaf(v.z, L, shift)); v.w = fast_ex2(fmaf(v.w, L, shift));
    } else {
      v.x = fmaxf(v.x * mult, 0.f); v.y = fmaxf(v.y * mult, 0.f);
      v.z = fmaxf(v.z * mult, 0.f); v.w = fmaxf(v.w * mult, 0.f);
    }
    s_map[i] = v;
    const float e = (v.x + v.y) + (v.z + v.w);
    const float q = fmaf(3.f, v.w, fmaf(2.f, v.z, v.y));  // sum_k k * e_k
    s += e;
    sx += fmaf((float)(4 * rw.px4), e, q);
    sy = fmaf((float)rw.py, e, sy);
  }
  s = block_sum_n(s, s_red);
  sx = block_sum_n(sx, s_red);
  sy = block_sum_n(sy, s_red);
  if (threadIdx.x == 0) {
    coords[(size_t)blockIdx.x * 2 + 0] = __fdiv_rn(sx, s);
    coords[(size_t)blockIdx.x * 2 + 1] = __fdiv_rn(sy, s);
  }
  const float inv = SOFTMAX ? __fdiv_rn(1.f, s) : 1.f;
  for (int i = threadIdx.x; i < n4; i += nthr) {
    float4 v = s_map[i];
    v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
    orow[i] = v;
  }
}

// softmax backward written around the fp64 centroid of the SAVED probabilities:
//   G_i - dot = (gp_i - GPbar) + gx*(x_i - Xbar) + gy*(y_i - Ybar),   Xbar = sum p x / sum p (fp64), ...
// x_i - Xbar is formed as (x_i - rint(Xbar)) - frac: an exact integer minus a small fp32, so the near-one-hot maps of
// heatmap_multiplier = 100 keep RELATIVE accuracy at the peak (see the scalar kernel for why that matters).
template <bool SOFTMAX, bool HAS_GP>
__global__ void __launch_bounds__(1024) k_sam2d_bwd_v4(const float* __restrict__ prob, const float* __restrict__ coords,
                                                       const float* __restrict__ gcoords,
                                                       const float* __restrict__ gprob, float* __restrict__ ghm,
                                                       int H, int W, float mult) {
  extern __shared__ __align__(16) float4 s_map[];
  __shared__ double s_red[32];
  const int n4 = (H * W) >> 2, W4 = W >> 2, nthr = blockDim.x;
  const float4* prow = reinterpret_cast<const float4*>(prob + (size_t)blockIdx.x * H * W);
  const float4* grow = HAS_GP ? reinterpret_cast<const float4*>(gprob + (size_t)blockIdx.x * H * W) : nullptr;
  float4* orow = reinterpret_cast<float4*>(ghm + (size_t)blockIdx.x * H * W);
  const float gx = __ldg(gcoords + (size_t)blockIdx.x * 2), gy = __ldg(gcoords + (size_t)blockIdx.x * 2 + 1);

  double mass = 0.0, SX = 0.0, SY = 0.0, SG = 0.0;
  {
    RowWalk rw(threadIdx.x, nthr, W4);
    for (int i = threadIdx.x; i < n4; i += nthr, rw.next()) {
      const float4 p = ldg128_stream(reinterpret_cast<const float*>(prow + i));
      s_map[i] = p;
      const double p0 = p.x, p1 = p.y, p2 = p.z, p3 = p.w;
      const double e = (p0 + p1) + (p2 + p3);
      mass += e;
      if (SOFTMAX) {
        SX += fma((double)(4 * rw.px4), e, fma(3.0, p3, fma(2.0, p2, p1)));
        SY = fma((double)rw.py, e, SY);
        if (HAS_GP) {
          const float4 g = __ldg(grow + i);
          SG = fma(p0, (double)g.x, fma(p1, (double)g.y, fma(p2, (double)g.z, fma(p3, (double)g.w, SG))));
        }
      }
    }
  }
  mass = block_sum_n(mass, s_red);
  float Xf = 0.f, Yf = 0.f, invS = 0.f;
  int Xi = 0, Yi = 0;
  double GPbar = 0.0;
  if (SOFTMAX) {
    SX = block_sum_n(SX, s_red);
    SY = block_sum_n(SY, s_red);
    if (HAS_GP) GPbar = block_sum_n(SG, s_red) / mass;
    const double Xbar = SX / mass, Ybar = SY / mass;
    const double xr = rint(Xbar), yr = rint(Ybar);
    Xi = (int)xr; Yi = (int)yr;
    Xf = (float)(Xbar - xr); Yf = (float)(Ybar - yr);
  } else {
    invS = (float)(1.0 / mass);
    Xf = __ldg(coords + (size_t)blockIdx.x * 2); Yf = __ldg(coords + (size_t)blockIdx.x * 2 + 1);  // X, Y of op.py:39-41
  }
  RowWalk rw(threadIdx.x, nthr, W4);
  for (int i = threadIdx.x; i < n4; i += nthr, rw.next()) {
    const float4 p = s_map[i];
    float4 gp = make_float4(0.f, 0.f, 0.f, 0.f);
    if (HAS_GP) gp = __ldg(grow + i);
    float4 g;
    if (SOFTMAX) {
      // every x offset is formed from the exact integer first: deriving t1..t3 from t0 (+gx, +2gx, ..) leaves the
      // rounding residue of gx*dx at the peak, where the true value is ~0
      const float cy = gy * ((float)(rw.py - Yi) - Yf);
      const int ix = 4 * rw.px4 - Xi;
      float t0 = fmaf(gx, (float)ix - Xf, cy), t1 = fmaf(gx, (float)(ix + 1) - Xf, cy);
      float t2 = fmaf(gx, (float)(ix + 2) - Xf, cy), t3 = fmaf(gx, (float)(ix + 3) - Xf, cy);
      if (HAS_GP) {
        t0 += (float)((double)gp.x - GPbar); t1 += (float)((double)gp.y - GPbar);
        t2 += (float)((double)gp.z - GPbar); t3 += (float)((double)gp.w - GPbar);
      }
      g.x = mult * (p.x * t0); g.y = mult * (p.y * t1); g.z = mult * (p.z * t2); g.w = mult * (p.w * t3);
    } else {
      const float cy = gy * ((float)rw.py - Yf);
      const int ix = 4 * rw.px4;
      g.x = (p.x > 0.f) ? mult * (gp.x + fmaf(gx, (float)ix - Xf, cy) * invS) : 0.f;
      g.y = (p.y > 0.f) ? mult * (gp.y + fmaf(gx, (float)(ix + 1) - Xf, cy) * invS) : 0.f;
      g.z = (p.z > 0.f) ? mult * (gp.z + fmaf(gx, (float)(ix + 2) - Xf, cy) * invS) : 0.f;
      g.w = (p.w > 0.f) ? mult * (gp.w + fmaf(gx, (float)(ix + 3) - Xf, cy) * invS) : 0.f;
    }
    orow[i] = g;
  }
}

constexpr size_t kS2MaxSmem = 200 * 1024;
static bool sam2d_fast_ok(const void* a, const void* b, const void* c, int H, int W) {
  const size_t bytes = (size_t)H * W * 4;
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  return (W % 4 == 0) && bytes <= kS2MaxSmem && al(a) && al(b) && al(c);
}
static int sam2d_threads(int n4) { return n4 <= 4096 ? 256 : 1024; }
template <typename K>
static cudaError_t sam2d_allow_smem(K kernel) {
  return smem_opt_in(reinterpret_cast<const void*>(kernel), kS2MaxSmem);
}

cudaError_t launch_softargmax2d_fwd(const float* hm, float* coords, float* prob, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st) {
  if (BJ == 0) return cudaSuccess;
  if (H * W > 0 && sam2d_fast_ok(hm, prob, nullptr, H, W)) {
    const size_t smem = (size_t)H * W * 4;
    const int nthr = sam2d_threads(H * W / 4);
    cudaError_t e = softmax ? sam2d_allow_smem(k_sam2d_fwd_v4<true>) : sam2d_allow_smem(k_sam2d_fwd_v4<false>);
    if (e != cudaSuccess) return e;
    if (softmax) k_sam2d_fwd_v4<true><<<BJ, nthr, smem, st>>>(hm, coords, prob, H, W, mult);
    else k_sam2d_fwd_v4<false><<<BJ, nthr, smem, st>>>(hm, coords, prob, H, W, mult);
    return cudaGetLastError();
  }
  if (softmax) k_sam2d_fwd<true><<<BJ, kS2Threads, 0, st>>>(hm, coords, prob, H, W, mult);
  else k_sam2d_fwd<false><<<BJ, kS2Threads, 0, st>>>(hm, coords, prob, H, W, mult);
  return cudaGetLastError();
}

cudaError_t launch_softargmax2d_bwd(const float* prob, const float* coords, const float* grad_coords,
                                    const float* grad_prob, float* grad_hm, int BJ, int H, int W, int softmax,
                                    float mult, cudaStream_t st) {
  if (BJ == 0 || H * W == 0) return cudaSuccess;
  const bool gp = grad_prob != nullptr;
  if (sam2d_fast_ok(prob, grad_hm, grad_prob, H, W)) {
    const size_t smem = (size_t)H * W * 4;
    const int nthr = sam2d_threads(H * W / 4);
#define LT_GO4(S_, G_)                                                                                              \
  do {                                                                                                              \
    cudaError_t e = sam2d_allow_smem(k_sam2d_bwd_v4<S_, G_>);                                                       \
    if (e != cudaSuccess) return e;                                                                                 \
    k_sam2d_bwd_v4<S_, G_><<<BJ, nthr, smem, st>>>(prob, coords, grad_coords, grad_prob, grad_hm, H, W, mult);      \
  } while (0)
    if (softmax) { if (gp) LT_GO4(true, true); else LT_GO4(true, false); }
    else         { if (gp) LT_GO4(false, true); else LT_GO4(false, false); }
#undef LT_GO4
    return cudaGetLastError();
  }
#define LT_GO(S_, G_) k_sam2d_bwd<S_, G_><<<BJ, kS2Threads, 0, st>>>(prob, coords, grad_coords, grad_prob, grad_hm, H, W, mult)
  if (softmax) { if (gp) LT_GO(true, true); else LT_GO(true, false); }
  else         { if (gp) LT_GO(false, true); else LT_GO(false, false); }
#undef LT_GO
  return cudaGetLastError();
}

}  // namespace lt
