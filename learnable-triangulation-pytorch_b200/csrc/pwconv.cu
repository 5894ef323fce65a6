// pwconv.cu — K7: the 1x1 convolution of `process_features` (mvn/models/triangulation.py:238-240, applied at :344-346
// right before op.unproject_heatmaps), producer-side layout fusion: it reads the backbone's channel-planar features
// (B·V, Cin=256, H, W) and writes K1's gather layout "hwcp" (layout.cu) directly, zero border and zero padded channels
// included, so K0 disappears from the volumetric step; its two backward kernels read the hwcp gradient map that K1's
// backward scatters into, so K0⁻¹ disappears too.
//
//   forward          out[n, y+1, x+1, c] = bias[c] + Σ_k W[c,k] · X[n,k,y,x]
//   backward (data)  dX[n,k,y,x]         = Σ_c W[c,k] · G[n, y+1, x+1, c]
//   backward (weight) dW[c,k] = Σ_{n,y,x} G[n,y+1,x+1,c] · X[n,k,y,x],   db[c] = Σ G[n,y+1,x+1,c]
//
// The work is GEMM shaped (M = pixels, N = 32, K = 256). At the reference's shape all three kernels run on the tcgen05
// tensor cores with a three-product TF32 split that keeps fp32 accuracy (namespace tc below; the parity bar is the fp32
// oracle and single-pass TF32 is 1e-3 away from it); shapes they do not cover run fp32 FMA chains in a fixed order as
// packed FFMA2 (k_pwconv_fwd / _bwd_data / _bwd_weight). Everything is deterministic (the weight gradient's per-CTA partial
// sums are merged in CTA order) and within 2e-6 of ATen's fp32 convolution. Bounds: each kernel reads or writes the
// (B·V,256,H,W) tensor once (HBM floor) and issues 2·Cin·Cout flops per pixel.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace lt {
namespace {

constexpr int kPx = 256;     // pixels per CTA tile; a thread owns two quads, 128 apart
constexpr int kCo = 32;      // output channels per CTA = one 128-byte texel of the gather layout
constexpr int kKc = 16;      // input channels per pipeline stage (forward)
constexpr int kStages = 4;
constexpr int kKn = 256;     // input channels resident per CTA (backward kernels)

__device__ __forceinline__ void cp_async16(float* smem, const float* g, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 16 : 0;  // src-size 0: nothing is read, the 16 bytes are zero filled
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(g), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async4(float* smem, const float* g, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int n = valid ? 4 : 0;
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(s), "l"(g), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// One reduction index r of out[px][n] += A[r][px] · Bm[r][n] for this thread's 8 pixels × 4 columns.
// acc[j][c]: register pair = pixels (2j, 2j+1) of column c -> 16 FFMA2 for 3 LDS.128 and 4 register copies.
__device__ __forceinline__ void mk_step(f2 (&acc)[4][4], const float* a_row, const float* b_row) {
  const float4 xa = *reinterpret_cast<const float4*>(a_row);
  const float4 xb = *reinterpret_cast<const float4*>(a_row + 128);
  const float4 w = *reinterpret_cast<const float4*>(b_row);
  const f2 x[4] = {pack2(xa.x, xa.y), pack2(xa.z, xa.w), pack2(xb.x, xb.y), pack2(xb.z, xb.w)};
  const f2 ww[4] = {pack2(w.x, w.x), pack2(w.y, w.y), pack2(w.z, w.z), pack2(w.w, w.w)};
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[j][c] = fma2(x[j], ww[c], acc[j][c]);
}

// Wt (Cin, Cp) = Wᵀ with zero-filled columns c >= Cout: 16-byte rows of 4 output channels for the forward's B operand.
__global__ void k_pw_transpose_weight(const float* __restrict__ Wm, float* __restrict__ Wt, int Cin, int Cout, int Cp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Cin * Cp) return;
  const int k = i / Cp, c = i - k * Cp;
  Wt[i] = c < Cout ? __ldg(Wm + (size_t)c * Cin + k) : 0.f;
}

// zero border of the gather layout: rows 0 and H+1, columns 0 and W+1 of every map (interior is the forward's job)
__global__ void __launch_bounds__(256) k_pw_zero_border(float* __restrict__ out, int Cp, int H, int W) {
  const int Hp = H + 2, Wp = W + 2, q = Cp / 4;
  float4* img = reinterpret_cast<float4*>(out + (size_t)blockIdx.x * Hp * Wp * Cp);
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const int nrow = Wp * q;  // float4 per padded row
  for (int i = threadIdx.x; i < 2 * nrow; i += blockDim.x)
    img[(size_t)(i < nrow ? 0 : Hp - 1) * nrow + (i < nrow ? i : i - nrow)] = z;
  for (int i = threadIdx.x; i < 2 * H * q; i += blockDim.x) {
    const int y = i / (2 * q), r = i - y * 2 * q;
    img[(size_t)(y + 1) * nrow + (r < q ? r : (Wp - 1) * q + (r - q))] = z;
  }
}

// ---- forward: grid (pixel tile, bv, 32-channel output block) ------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(256, 3)
    k_pwconv_fwd(const float* __restrict__ X, const float* __restrict__ Wt, const float* __restrict__ bias,
                 float* __restrict__ out, int Cin, int Cout, int Cp, int H, int W) {
  extern __shared__ __align__(16) float smem[];
  float* Xs = smem;                           // [kStages][kKc][kPx]
  float* Ws = smem + kStages * kKc * kPx;     // [kStages][kKc][kCo]
  const int HW = H * W;
  const int p0 = blockIdx.x * kPx, bv = blockIdx.y, co0 = blockIdx.z * kCo;
  const float* Xb = X + (size_t)bv * Cin * HW;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cq = lane & 7, pq = warp * 4 + (lane >> 3);
  const int nkt = (Cin + kKc - 1) / kKc;

  auto load_stage = [&](int kt, int slot) {
    const int k0 = kt * kKc;
    float* xs = Xs + slot * kKc * kPx;
    if (VEC) {
#pragma unroll
      for (int i = 0; i < kKc * kPx / 4 / 256; ++i) {
        const int idx = tid + i * 256, row = idx >> 6, c4 = (idx & 63) * 4;
        const bool ok = (k0 + row < Cin) && (p0 + c4 < HW);
        cp_async16(xs + row * kPx + c4, ok ? Xb + (size_t)(k0 + row) * HW + p0 + c4 : Xb, ok);
      }
    } else {
#pragma unroll
      for (int i = 0; i < kKc * kPx / 256; ++i) {
        const int idx = tid + i * 256, row = idx >> 8, c = idx & 255;
        const bool ok = (k0 + row < Cin) && (p0 + c < HW);
        cp_async4(xs + row * kPx + c, ok ? Xb + (size_t)(k0 + row) * HW + p0 + c : Xb, ok);
      }
    }
    if (tid < kKc * kCo / 4) {
      const int row = tid >> 3, c4 = (tid & 7) * 4;
      const bool ok = k0 + row < Cin;
      cp_async16(Ws + slot * kKc * kCo + row * kCo + c4, ok ? Wt + (size_t)(k0 + row) * Cp + co0 + c4 : Wt, ok);
    }
  };

  f2 acc[4][4];
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[j][c] = pack2(0.f, 0.f);

#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) {
    if (s < nkt) load_stage(s, s);
    cp_async_commit();
  }
  for (int kt = 0; kt < nkt; ++kt) {
    cp_async_wait<kStages - 2>();
    __syncthreads();
    {
      const int nx = kt + kStages - 1;
      if (nx < nkt) load_stage(nx, nx % kStages);
      cp_async_commit();
    }
    const int slot = kt % kStages;
    const float* xs = Xs + slot * kKc * kPx + pq * 4;
    const float* ws = Ws + slot * kKc * kCo + cq * 4;
#pragma unroll
    for (int k = 0; k < kKc; ++k) mk_step(acc, xs + k * kPx, ws + k * kCo);
  }
  cp_async_wait<0>();

  // epilogue: + bias, one 16-byte store per pixel (8 lanes = the 128-byte texel of the 32-channel block)
  float b[4] = {0.f, 0.f, 0.f, 0.f};
  if (bias != nullptr) {
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (co0 + cq * 4 + c < Cout) b[c] = __ldg(bias + co0 + cq * 4 + c);
  }
  const int Hp = H + 2, Wp = W + 2;
  float* ob = out + (size_t)bv * Hp * Wp * Cp + co0 + cq * 4;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float lo[4], hi[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) unpack2(acc[j][c], lo[c], hi[c]);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int p = p0 + (j >> 1) * 128 + pq * 4 + (j & 1) * 2 + h;
      if (p < HW) {
        const int y = p / W, x = p - y * W;
        const float* v = h ? hi : lo;
        *reinterpret_cast<float4*>(ob + ((size_t)(y + 1) * Wp + (x + 1)) * Cp) =
            make_float4(v[0] + b[0], v[1] + b[1], v[2] + b[2], v[3] + b[3]);
      }
    }
  }
}

// ---- backward w.r.t. the input: grid (pixel tile, bv, 256-channel input block); one 32-channel block of G per launch --
template <bool VEC>
__global__ void __launch_bounds__(256, 3)
    k_pwconv_bwd_data(const float* __restrict__ G, const float* __restrict__ Wm, float* __restrict__ dX, int Cin,
                      int Cout, int Cp, int H, int W, int co0, int accumulate) {
  extern __shared__ __align__(16) float smem[];
  float* Gs = smem;               // [kCo][kPx]  gradient tile, transposed to channel rows
  float* Wsm = smem + kCo * kPx;  // [kCo][kKn]  weight rows (input channel contiguous, as stored)
  const int HW = H * W, Hp = H + 2, Wp = W + 2;
  const int p0 = blockIdx.x * kPx, bv = blockIdx.y, kc0 = blockIdx.z * kKn;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int kq = lane & 7, pq = warp * 4 + (lane >> 3);
  {
    const int p = p0 + tid;
    const bool in = p < HW;
    const int y = in ? p / W : 0, x = in ? p - y * W : 0;
    const float* src = G + (((size_t)bv * Hp + y + 1) * Wp + x + 1) * Cp + co0;
#pragma unroll
    for (int c4 = 0; c4 < kCo / 4; ++c4) {
      const float4 g = in ? ldg128_stream(src + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
      Gs[(c4 * 4 + 0) * kPx + tid] = g.x;
      Gs[(c4 * 4 + 1) * kPx + tid] = g.y;
      Gs[(c4 * 4 + 2) * kPx + tid] = g.z;
      Gs[(c4 * 4 + 3) * kPx + tid] = g.w;
    }
    for (int idx = tid; idx < kCo * kKn; idx += 256) {
      const int c = idx >> 8, k = idx & (kKn - 1);
      Wsm[idx] = (co0 + c < Cout && kc0 + k < Cin) ? __ldg(Wm + (size_t)(co0 + c) * Cin + kc0 + k) : 0.f;
    }
  }
  __syncthreads();
  const int cmax = min(kCo, Cout - co0);
  const int nkb = min(kKn / 32, (Cin - kc0 + 31) / 32);
  for (int kb = 0; kb < nkb; ++kb) {
    f2 acc[4][4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[j][c] = pack2(0.f, 0.f);
    const float* gs = Gs + pq * 4;
    const float* ws = Wsm + kb * 32 + kq * 4;
#pragma unroll 8
    for (int c = 0; c < cmax; ++c) mk_step(acc, gs + c * kPx, ws + c * kKn);
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int k = kc0 + kb * 32 + kq * 4 + c;
      if (k >= Cin) continue;
      float* row = dX + ((size_t)bv * Cin + k) * HW;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int p = p0 + h * 128 + pq * 4;
        float v[4];
        unpack2(acc[2 * h][c], v[0], v[1]);
        unpack2(acc[2 * h + 1][c], v[2], v[3]);
        if (VEC && p + 3 < HW) {
          float4 o = make_float4(v[0], v[1], v[2], v[3]);
          if (accumulate) {
            const float4 old = *reinterpret_cast<const float4*>(row + p);
            o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
          }
          stg128_stream(row + p, o);
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i)
            if (p + i < HW) row[p + i] = accumulate ? row[p + i] + v[i] : v[i];
        }
      }
    }
  }
}

// ---- backward w.r.t. weight and bias: persistent workers over 16-pixel chunks, partial sums per worker -------------
constexpr int kWgPx = 16, kWgRow = kWgPx + 4, kWgStages = 4;
constexpr int kWgStageFloats = kKn * kWgRow + kWgPx * kCo;

template <bool VEC>
__global__ void __launch_bounds__(256, 2)
    k_pwconv_bwd_weight(const float* __restrict__ X, const float* __restrict__ G, float* __restrict__ part_w,
                        float* __restrict__ part_b, int Cin, int Cout, int Cp, int H, int W, int nchunk, int total,
                        int CinPad) {
  extern __shared__ __align__(16) float smem[];
  const int HW = H * W, Hp = H + 2, Wp = W + 2;
  const int kc0 = blockIdx.y * kKn, co0 = blockIdx.z * kCo;
  const int tid = threadIdx.x, kg = tid >> 3, cq = tid & 7;
  const int nw = gridDim.x, w = blockIdx.x;

  auto load_stage = [&](int t, int slot) {
    const int bv = t / nchunk, p0 = (t - bv * nchunk) * kWgPx;
    float* xs = smem + slot * kWgStageFloats;
    float* gs = xs + kKn * kWgRow;
    const float* Xb = X + ((size_t)bv * Cin + kc0) * HW + p0;
    if (VEC) {
#pragma unroll
      for (int i = 0; i < kKn * kWgPx / 4 / 256; ++i) {
        const int idx = tid + i * 256, row = idx >> 2, q = (idx & 3) * 4;
        const bool ok = (kc0 + row < Cin) && (p0 + q < HW);
        cp_async16(xs + row * kWgRow + q, ok ? Xb + (size_t)row * HW + q : X, ok);
      }
    } else {
#pragma unroll
      for (int i = 0; i < kKn * kWgPx / 256; ++i) {
        const int idx = tid + i * 256, row = idx >> 4, q = idx & 15;
        const bool ok = (kc0 + row < Cin) && (p0 + q < HW);
        cp_async4(xs + row * kWgRow + q, ok ? Xb + (size_t)row * HW + q : X, ok);
      }
    }
    if (tid < kWgPx * kCo / 4) {
      const int px = tid >> 3, c4 = (tid & 7) * 4;
      const int p = p0 + px;
      const bool ok = p < HW;
      const int y = ok ? p / W : 0, x = ok ? p - y * W : 0;
      cp_async16(gs + px * kCo + c4, G + (((size_t)bv * Hp + y + 1) * Wp + x + 1) * Cp + co0 + c4, ok);
    }
  };

  f2 acc[8][2];
#pragma unroll
  for (int r = 0; r < 8; ++r) acc[r][0] = acc[r][1] = pack2(0.f, 0.f);
  float bacc[4] = {0.f, 0.f, 0.f, 0.f};

  const int nt = w < total ? (total - w + nw - 1) / nw : 0;  // chunks of this worker: t = w + i*nw
#pragma unroll
  for (int s = 0; s < kWgStages - 1; ++s) {
    if (s < nt) load_stage(w + s * nw, s);
    cp_async_commit();
  }
  for (int it = 0; it < nt; ++it) {
    cp_async_wait<kWgStages - 2>();
    __syncthreads();
    {
      const int nx = it + kWgStages - 1;
      if (nx < nt) load_stage(w + nx * nw, nx % kWgStages);
      cp_async_commit();
    }
    const float* xs = smem + (it % kWgStages) * kWgStageFloats + kg * kWgRow;
    const float* gs = smem + (it % kWgStages) * kWgStageFloats + kKn * kWgRow + cq * 4;
#pragma unroll
    for (int q = 0; q < kWgPx / 4; ++q) {
      f2 g01[4], g23[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 g = *reinterpret_cast<const float4*>(gs + (q * 4 + i) * kCo);
        g01[i] = pack2(g.x, g.y);
        g23[i] = pack2(g.z, g.w);
        if (kg == 0) { bacc[0] += g.x; bacc[1] += g.y; bacc[2] += g.z; bacc[3] += g.w; }
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 x = *reinterpret_cast<const float4*>(xs + (32 * r) * kWgRow + q * 4);
        const float xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const f2 xx = pack2(xv[i], xv[i]);
          acc[r][0] = fma2(xx, g01[i], acc[r][0]);
          acc[r][1] = fma2(xx, g23[i], acc[r][1]);
        }
      }
    }
  }
  cp_async_wait<0>();
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    float4 o;
    unpack2(acc[r][0], o.x, o.y);
    unpack2(acc[r][1], o.z, o.w);
    *reinterpret_cast<float4*>(part_w + ((size_t)w * CinPad + kc0 + kg + 32 * r) * Cp + co0 + cq * 4) = o;
  }
  if (kg == 0 && blockIdx.y == 0)
    *reinterpret_cast<float4*>(part_b + (size_t)w * Cp + co0 + cq * 4) = make_float4(bacc[0], bacc[1], bacc[2], bacc[3]);
}

// dW[c][k] = Σ_w part_w[w][k][c] in worker order (8 interleaved slices, then a fixed tree): deterministic.
// grid = Cin + 1 rows: row k < Cin reduces weight row k, row Cin the bias. block = (32 channels, 8 slices).
__global__ void __launch_bounds__(256) k_pw_reduce_partials(const float* __restrict__ part_w, const float* __restrict__ part_b,
                                                            float* __restrict__ dW, float* __restrict__ db, int Cin, int Cout,
                                                            int Cp, int CinPad, int nw) {
  __shared__ float red[8][33];
  const int k = blockIdx.x, c0 = blockIdx.y * 32;
  const int c = threadIdx.x & 31, s = threadIdx.x >> 5;
  const bool is_bias = k == Cin;
  if (is_bias && db == nullptr) return;
  const float* src = is_bias ? part_b + c0 + c : part_w + (size_t)k * Cp + c0 + c;
  const size_t stride = is_bias ? (size_t)Cp : (size_t)CinPad * Cp;
  float a = 0.f;
  for (int i = s; i < nw; i += 8) a += __ldg(src + (size_t)i * stride);
  red[s][c] = a;
  __syncthreads();
  if (s == 0 && c0 + c < Cout) {
    const float t = ((red[0][c] + red[1][c]) + (red[2][c] + red[3][c])) + ((red[4][c] + red[5][c]) + (red[6][c] + red[7][c]));
    if (is_bias) db[c0 + c] = t;
    else dW[(size_t)(c0 + c) * Cin + k] = t;
  }
}


// ---- forward on the 5th-generation tensor cores: tcgen05.mma kind::tf32 with a three-product split --------------------
// TF32 keeps 10 mantissa bits, the parity bar is fp32. Every operand is split x = hi + lo (hi = the top 19 bits of x,
// lo = x - hi, exact in fp32) and D += hi·hi + hi·lo + lo·hi: the dropped lo·lo term is 2^-22 relative, the accumulators
// live in TMEM as fp32. Per CTA (one per SM, persistent over 128-pixel tiles):
//   warps 0-11  three producer groups of 128 threads, group g takes every third 32-channel stage: coalesced LDG of X
//               (a thread = one pixel), split, and tcgen05.st of the hi and lo rows STRAIGHT INTO TENSOR MEMORY — the A
//               operand never touches shared memory (an earlier version staged it there as UMMA canonical tiles: the STS
//               of hi+lo plus the tensor core's operand fetches kept the L1/shared data pipe 72 % busy and capped the
//               kernel at 170 us; measured, profiles/). The loads of a thread's next stage stay in flight in registers.
//   warp  12    one elected lane issues, per 8 input channels, one N = 64 tcgen05.mma (A_hi from TMEM x [W_hi | W_lo] from
//               shared memory: hi·hi and hi·lo in one instruction) and one N = 32 (A_lo x W_hi), then commits the stage's
//               TMEM columns back to the producers through an mbarrier
//   warps 13-16 epilogue: tcgen05.ld of the three 128 x 32 partial accumulators (a thread = one pixel), sum, + bias, four
//               32-byte stores = the 128-byte texel of the gather layout; accumulators are double buffered so tile t
//               drains under tile t+1.
// W (32 x Cin) is split once per CTA into two resident K-major tiles (UMMA canonical no-swizzle layout, core matrix =
// 8 channels x 16 bytes), W_lo right behind W_hi so that both form one 64-row B operand.
// TMEM map (512 columns x 128 lanes, lane = pixel of the tile): [0,192) two accumulator buffers of 96 columns
// (hi·hi | hi·lo | lo·hi); [192,512) ring of five A stages of 64 columns (32 channels hi | 32 channels lo).
namespace tc {
constexpr int kM = 128, kN = 32, kKs = 32, kStagesTc = 5;
constexpr int kGroups = 3, kProducerWarps = 4 * kGroups, kMmaWarp = kProducerWarps, kWarps = kProducerWarps + 5;
constexpr int kRawDepth = 3;   // raw stages in flight per producer group (16 KB each)
constexpr uint32_t kLboB = 128;                          // W tile: [channel group of 8][k4][8 channels][16 B]
constexpr uint32_t idesc_tf32(int n) { return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(kM >> 4) << 24); }
constexpr uint32_t kSpinLimit = 1u << 16;          // fast polls before backing off
constexpr uint64_t kWaitLimitNs = 20ull * 1000000000ull;
constexpr uint32_t kAccCols = 3 * kN, kRingCol0 = 2 * kAccCols, kStageCols = 2 * kKs, kTmemCols = 512;
static_assert(kRingCol0 + kStagesTc * kStageCols <= kTmemCols, "TMEM map");

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// bounded wait: a protocol error traps (the launch fails loudly) instead of hanging the GPU. The bound is TIME, not a
// poll count: after kSpinLimit fast polls the waiter backs off with nanosleep and only traps once the barrier has been
// stuck for kWaitLimitNs of wall clock (20 s), so a legitimately long stall (debugger, sanitizer, preemption) is not
// turned into a sticky context error.
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  const uint32_t a = smem_u32(b);
  uint64_t t0 = 0;
  for (uint32_t i = 0;; ++i) {
    uint32_t done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(a), "r"(parity) : "memory");
    if (done) return;
    if (i >= kSpinLimit) {
      if (t0 == 0) t0 = globaltimer_ns();
      __nanosleep(256);
      if ((i & 0x3ffu) == 0 && globaltimer_ns() - t0 > kWaitLimitNs) __trap();
    }
  }
}
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3ffffu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// D[tmem] (+)= A[tmem] · B[smem]ᵀ: A is 128 lanes x 8 columns of tf32, B a K-major descriptor
__device__ __forceinline__ void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* b) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void proxy_fence() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                 "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kWarps * 32, 1)
    k_pwconv_fwd_tc(const float* __restrict__ X, const float* __restrict__ Wm, const float* __restrict__ bias,
                    float* __restrict__ out, int Cin, int Cout, int Cp, int H, int W, int tiles_per_img, int ntiles) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* sW = sm;
  const uint32_t wbytes = (uint32_t)kN * Cin * 4;                 // one of the two W tiles
  uint64_t* bars = reinterpret_cast<uint64_t*>(sW + 2 * wbytes);
  uint64_t *full = bars, *empty = bars + kStagesTc, *tfull = bars + 2 * kStagesTc, *tempty = bars + 2 * kStagesTc + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStagesTc + 4);
  float* s_bias = reinterpret_cast<float*>(tmem_slot + 1);        // 32 floats, zero padded
  uint64_t* raw_bar = bars + 32;                                  // [kGroups][kRawDepth]
  unsigned char* sRaw = sW + 2 * wbytes + 512;                    // [kGroups][kRawDepth][32 rows][128 px] fp32
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int HW = H * W, nks = Cin / kKs;
  const uint32_t sboB = (uint32_t)(Cin / 4) * 128;

  if (tid == 0) {
    for (int s = 0; s < kStagesTc; ++s) { mbar_init(full + s, 4); mbar_init(empty + s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull + b, 1); mbar_init(tempty + b, 4); }
    for (int i = 0; i < kGroups * kRawDepth; ++i) mbar_init(raw_bar + i, 128);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < kN) s_bias[tid] = (bias != nullptr && tid < Cout) ? __ldg(bias + tid) : 0.f;
  for (int idx = tid; idx < kN * Cin; idx += kWarps * 32) {
    const int n = idx / Cin, k = idx - n * Cin;
    const float w = n < Cout ? __ldg(Wm + (size_t)n * Cin + k) : 0.f;
    const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
    const uint32_t off = (uint32_t)(n >> 3) * sboB + (uint32_t)(k >> 2) * kLboB + (n & 7) * 16 + (k & 3) * 4;
    *reinterpret_cast<float*>(sW + off) = hi;
    *reinterpret_cast<float*>(sW + wbytes + off) = w - hi;
  }
  proxy_fence();   // generic-proxy writes of W -> visible to the tensor core's async-proxy reads
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp < kProducerWarps) {
    // thread = pixel px = TMEM lane (warp & 3 is the lane quarter a warp may touch, and groups start at multiples of 4)
    const int grp = warp >> 2, px = tid & (kM - 1);
    const int total = my_tiles * nks;
    const uint32_t lane_base = tmem_base + kRingCol0 + ((uint32_t)((warp & 3) * 32) << 16);
    // Raw ring, private to the group: kRawDepth of ITS stages in flight as 16-byte cp.async (lanes along the 512-byte pixel
    // run of a channel row), each thread posting a no-increment arrive that fires when its copies have landed. Registers
    // hold one stage only; the depth of the prefetch no longer depends on them (the register-prefetch version was bound by
    // the latency of its loads at 148 us).
    float* const raw_g = reinterpret_cast<float*>(sRaw) + grp * kRawDepth * (kKs * kM);
    uint64_t* const raw_full = raw_bar + grp * kRawDepth;
    const int t128 = tid & 127;
    auto issue = [&](int n) {                       // n-th own stage of this group = global stage grp + n * kGroups
      const int j = grp + n * kGroups;
      if (j >= total) return;
      const int t = j / nks, ks = j - t * nks;
      const int tile = blockIdx.x + t * gridDim.x;
      const int bv = tile / tiles_per_img, p0 = (tile - bv * tiles_per_img) * kM;
      float* dst = raw_g + (n % kRawDepth) * (kKs * kM);
      const float* src = X + ((size_t)bv * Cin + ks * kKs) * HW + p0;
#pragma unroll
      for (int i = 0; i < kKs * kM / 4 / 128; ++i) {
        const int idx = t128 + 128 * i, row = idx >> 5, c4 = (idx & 31) * 4;
        const bool ok = p0 + c4 < HW;               // H*W % 4 == 0 (usable): a 16-byte piece is all in or all out
        cp_async16(dst + row * kM + c4, ok ? src + (size_t)row * HW + c4 : X, ok);
      }
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(raw_full + n % kRawDepth)) : "memory");
    };
#pragma unroll
    for (int n = 0; n < kRawDepth; ++n) issue(n);
    int n = 0;
    for (int j = grp; j < total; j += kGroups, ++n) {
      const int rs = n % kRawDepth;
      mbar_wait(raw_full + rs, (n / kRawDepth) & 1);
      float v[kKs];
      const float* rp = raw_g + rs * (kKs * kM) + px;
#pragma unroll
      for (int i = 0; i < kKs; ++i) v[i] = rp[i * kM];          // lanes = consecutive pixels: conflict free
      asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");   // the group has read the slot: refill it
      issue(n + kRawDepth);
      const int slot = j % kStagesTc;
      mbar_wait(empty + slot, ((j / kStagesTc) & 1) ^ 1);   // the MMAs that read this TMEM slot have completed
      tc_fence_after();
      const uint32_t t0 = lane_base + slot * kStageCols;
#pragma unroll
      for (int h = 0; h < kKs / 8; ++h) {
        uint32_t r[8];   // one 8-register staging set, reused for the lo half (hi is recomputed: one LOP3)
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = __float_as_uint(v[h * 8 + i]) & 0xffffe000u;
        tmem_st8(t0 + h * 8, r);
#pragma unroll
        for (int i = 0; i < 8; ++i)
          r[i] = __float_as_uint(v[h * 8 + i] - __uint_as_float(__float_as_uint(v[h * 8 + i]) & 0xffffe000u));
        tmem_st8(t0 + kKs + h * 8, r);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(full + slot);
    }
  } else if (warp == kMmaWarp) {
    // The whole warp walks the loop with warp-uniform values (TMEM base re-broadcast with a shuffle so that the compiler
    // can keep descriptors in uniform registers) and one elected lane issues. With the loop under `if (lane == 0)` every
    // UTCHMMA was wrapped in an ELECT / 4 x R2UR / branch sequence: ~300 cycles per 8-channel step, which alone bounded
    // the kernel at 165 us (measured).
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t leader;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(leader));
    int j = 0;
    const uint32_t w_addr = smem_u32(sW);
    for (int t = 0; t < my_tiles; ++t) {
      const int buf = t & 1;
      mbar_wait(tempty + buf, ((t >> 1) & 1) ^ 1);   // the epilogue has drained this accumulator buffer
      tc_fence_after();
      const uint32_t d = tb + buf * kAccCols;
      for (int ks = 0; ks < nks; ++ks, ++j) {
        const int slot = j % kStagesTc;
        mbar_wait(full + slot, (j / kStagesTc) & 1);
        tc_fence_after();
        const uint32_t a_hi = tb + kRingCol0 + slot * kStageCols, a_lo = a_hi + kKs;
        if (leader) {
#pragma unroll
          for (int k8 = 0; k8 < kKs / 8; ++k8) {
            const uint64_t b = umma_desc(w_addr + (uint32_t)((ks * kKs + k8 * 8) >> 2) * kLboB, kLboB, sboB);
            const uint32_t acc = (ks | k8) != 0 ? 1u : 0u;
            umma_tf32_ts(d, a_hi + k8 * 8, b, idesc_tf32(2 * kN), acc);        // columns 0-31 hi·hi, 32-63 hi·lo
            umma_tf32_ts(d + 2 * kN, a_lo + k8 * 8, b, idesc_tf32(kN), acc);   // columns 64-95 lo·hi
          }
          umma_commit(empty + slot);     // slot reusable once these MMAs have read it
        }
        __syncwarp();
      }
      if (leader) umma_commit(tfull + buf);   // accumulators complete
      __syncwarp();
    }
  } else {
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    const int Hp = H + 2, Wp = W + 2;
    for (int t = 0; t < my_tiles; ++t) {
      const int buf = t & 1;
      const int tile = blockIdx.x + t * gridDim.x;
      mbar_wait(tfull + buf, (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * kAccCols + ((uint32_t)(q * 32) << 16);
      float acc[kN];
#pragma unroll
      for (int h = 0; h < 2; ++h) {          // 16 columns at a time; the small products first, hi·hi last
        uint32_t r[16];
        tmem_ld16(taddr + kN + h * 16, r);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[h * 16 + i] = __uint_as_float(r[i]);
        tmem_ld16(taddr + 2 * kN + h * 16, r);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[h * 16 + i] += __uint_as_float(r[i]);
        tmem_ld16(taddr + h * 16, r);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[h * 16 + i] += __uint_as_float(r[i]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty + buf);
      const int bv = tile / tiles_per_img, p = (tile - bv * tiles_per_img) * kM + q * 32 + lane;
      if (p < HW) {
        const int y = p / W, x = p - y * W;
        float* dst = out + (((size_t)bv * Hp + y + 1) * Wp + x + 1) * Cp;
#pragma unroll
        for (int c8 = 0; c8 < kN / 8; ++c8) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = acc[c8 * 8 + i] + s_bias[c8 * 8 + i];
          stg256(dst + c8 * 8, v);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}

// ---- backward w.r.t. the input on the tensor cores -------------------------------------------------------------------
// dX[px][k] = Σ_c G[px][c] · W[c][k]: M = 128 pixels, N = Cin (<= 256), K = 32 output channels (zero padded). The same
// 3xTF32 split; here the accumulator is the big thing (128 x 256 fp32 = 256 TMEM columns, double buffered = all 512), so
// the small A operand (the pixel tile of the hwcp gradient map, already K-major: 128 contiguous bytes per pixel) goes
// through shared memory as UMMA canonical tiles and W^T (hi and lo, 64 KB) stays resident.
//   warps 0-3   producers (thread = pixel): 8 x LDG.128 of the pixel's texel one tile ahead, split, STS
//   warp  4     12 tcgen05.mma (N = Cin, K = 8) per tile into one accumulator: lo·hi, hi·lo, hi·hi
//   warps 5-20  two epilogue sets alternating tiles: a thread owns a pixel and half of the input channels; tcgen05.ld, then
//               one coalesced 128-byte store per channel and warp into the channel-planar gradient.
constexpr int kBdStages = 3, kBdProducerWarps = 4, kBdMmaWarp = 4, kBdEpiSets = 2, kBdWarps = 5 + 8 * kBdEpiSets;
constexpr uint32_t kBdLboA = (kM / 8) * 128, kBdSboA = 128, kBdTile = kM * kN * 4;   // A: [c4][pixel group][8 px][16 B]
constexpr uint32_t kBdLboB = 128, kBdSboB = (kN / 4) * 128;                          // W^T: [k group of 8][c4][8 k][16 B]

__device__ __forceinline__ void umma_tf32_ss(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
      "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kBdWarps * 32, 1)
    k_pwconv_bwd_data_tc(const float* __restrict__ G, const float* __restrict__ Wm, float* __restrict__ dX, int Cin,
                         int Cout, int Cp, int H, int W, int tiles_per_img, int ntiles) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* sA = sm;                                        // [kBdStages][hi 16 KB | lo 16 KB]
  unsigned char* sW = sm + kBdStages * 2 * kBdTile;              // W^T hi (Cin x 32) then lo
  const uint32_t wbytes = (uint32_t)Cin * kN * 4;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sW + 2 * wbytes);
  uint64_t *full = bars, *empty = bars + kBdStages, *tfull = bars + 2 * kBdStages, *tempty = bars + 2 * kBdStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kBdStages + 4);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int HW = H * W, Hp = H + 2, Wp = W + 2;

  if (tid == 0) {
    for (int s = 0; s < kBdStages; ++s) { mbar_init(full + s, 4); mbar_init(empty + s, 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(tfull + b, 1); mbar_init(tempty + b, 8); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kBdMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int idx = tid; idx < Cin * kN; idx += kBdWarps * 32) {   // idx = c * Cin + k: coalesced reads of W (Cout, Cin)
    const int c = idx / Cin, k = idx - c * Cin;
    const float w = c < Cout ? __ldg(Wm + (size_t)c * Cin + k) : 0.f;
    const float hi = __uint_as_float(__float_as_uint(w) & 0xffffe000u);
    const uint32_t off = (uint32_t)(k >> 3) * kBdSboB + (uint32_t)(c >> 2) * kBdLboB + (k & 7) * 16 + (c & 3) * 4;
    *reinterpret_cast<float*>(sW + off) = hi;
    *reinterpret_cast<float*>(sW + wbytes + off) = w - hi;
  }
  proxy_fence();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_tiles = (int)blockIdx.x < ntiles ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp < kBdProducerWarps) {
    const int px = tid & (kM - 1);
    unsigned char* const base0 = sA + (px >> 3) * kBdSboA + (px & 7) * 16;
    float4 v[kN / 4];
    auto load = [&](int t) {
      if (t >= my_tiles) return;
      const int tile = blockIdx.x + t * gridDim.x;
      const int bv = tile / tiles_per_img, p = (tile - bv * tiles_per_img) * kM + px;
      if (p < HW) {
        const int y = p / W, x = p - y * W;
        const float* src = G + (((size_t)bv * Hp + y + 1) * Wp + x + 1) * Cp;
#pragma unroll
        for (int q = 0; q < kN / 4; ++q) v[q] = ldg128_stream(src + q * 4);
      } else {
#pragma unroll
        for (int q = 0; q < kN / 4; ++q) v[q] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    };
    load(0);
    for (int t = 0; t < my_tiles; ++t) {
      const int slot = t % kBdStages;
      mbar_wait(empty + slot, ((t / kBdStages) & 1) ^ 1);
      unsigned char* const base = base0 + slot * 2 * kBdTile;
#pragma unroll
      for (int q = 0; q < kN / 4; ++q) {
        float4 hi, lo;
        hi.x = __uint_as_float(__float_as_uint(v[q].x) & 0xffffe000u); lo.x = v[q].x - hi.x;
        hi.y = __uint_as_float(__float_as_uint(v[q].y) & 0xffffe000u); lo.y = v[q].y - hi.y;
        hi.z = __uint_as_float(__float_as_uint(v[q].z) & 0xffffe000u); lo.z = v[q].z - hi.z;
        hi.w = __uint_as_float(__float_as_uint(v[q].w) & 0xffffe000u); lo.w = v[q].w - hi.w;
        *reinterpret_cast<float4*>(base + q * kBdLboA) = hi;
        *reinterpret_cast<float4*>(base + kBdTile + q * kBdLboA) = lo;
      }
      proxy_fence();
      __syncwarp();
      if (lane == 0) mbar_arrive(full + slot);
      load(t + 1);
    }
  } else if (warp == kBdMmaWarp) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t leader;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(leader));
    const uint32_t w_addr = smem_u32(sW), a_addr = smem_u32(sA);
    const uint32_t idesc = idesc_tf32(Cin);
    for (int t = 0; t < my_tiles; ++t) {
      const int buf = t & 1, slot = t % kBdStages;
      mbar_wait(tempty + buf, ((t >> 1) & 1) ^ 1);
      mbar_wait(full + slot, (t / kBdStages) & 1);
      tc_fence_after();
      if (leader) {
        const uint32_t d = tb + buf * 256;
#pragma unroll
        for (int k8 = 0; k8 < kN / 8; ++k8) {
          const uint64_t ah = umma_desc(a_addr + slot * 2 * kBdTile + k8 * 2 * kBdLboA, kBdLboA, kBdSboA);
          const uint64_t al = umma_desc(a_addr + slot * 2 * kBdTile + kBdTile + k8 * 2 * kBdLboA, kBdLboA, kBdSboA);
          const uint64_t bh = umma_desc(w_addr + k8 * 2 * kBdLboB, kBdLboB, kBdSboB);
          const uint64_t bl = umma_desc(w_addr + wbytes + k8 * 2 * kBdLboB, kBdLboB, kBdSboB);
          umma_tf32_ss(d, al, bh, idesc, k8 != 0 ? 1u : 0u);
          umma_tf32_ss(d, ah, bl, idesc, 1u);
          umma_tf32_ss(d, ah, bh, idesc, 1u);
        }
        umma_commit(empty + slot);
        umma_commit(tfull + buf);
      }
      __syncwarp();
    }
  } else {
    // Two epilogue sets of eight warps alternate tiles (set s drains accumulator buffer s): the tcgen05.ld phase of one
    // tile runs under the stores of the other. With one set the two phases added up (4.3 us per tile: 1.5 us without the
    // stores, 2.8 us of stores at HBM speed; timing experiments in DESIGN.md) and the kernel sat at 135 us.
    const int e = warp - (kBdMmaWarp + 1);            // 0..15
    const int set = e >> 3;
    const int q = warp & 3, half = (e >> 2) & 1;      // TMEM lane quarter, and which half of the input channels
    const int kper = (Cin + 1) / 2;                   // channels [half*kper, min(Cin, (half+1)*kper))
    for (int t = set; t < my_tiles; t += kBdEpiSets) {
      const int buf = t & 1;
      const int tile = blockIdx.x + t * gridDim.x;
      const int bv = tile / tiles_per_img, p = (tile - bv * tiles_per_img) * kM + q * 32 + lane;
      mbar_wait(tfull + buf, (t >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 256 + ((uint32_t)(q * 32) << 16);
      // Cin % 64 == 0 (bd_usable): each half is a whole number of 32-column loads, no per-channel bound check; the
      // store address advances by one plane per channel (a bound check + 64-bit multiply per store made this epilogue
      // 2600 instructions per warp and tile, and the kernel instruction bound at 235 us)
      float* q = dX + ((size_t)bv * Cin + half * kper) * HW + p;
      const size_t plane = (size_t)HW;
      for (int k0 = 0; k0 < kper; k0 += 32) {
        uint32_t r[32];
        tmem_ld32(taddr + half * kper + k0, r);
        if (p < HW) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            *q = __uint_as_float(r[i]);   // write-back: L2 merges a plane's neighbouring tiles
            q += plane;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty + buf);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kBdMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}
inline size_t bd_smem_bytes(int Cin) { return (size_t)kBdStages * 2 * kBdTile + 2 * (size_t)Cin * kN * 4 + 512; }
inline bool bd_usable(int Cin, int Cout) { return Cout <= kN && Cin % 64 == 0 && Cin >= 64 && Cin <= 256; }

// ---- backward w.r.t. the weight on the tensor cores ------------------------------------------------------------------
// dW[c][k] = Σ_p G[p][c] · X[k][p], computed as D[k][c]: M = 128 input channels per block (Cin / 128 blocks), N = 32 output
// channels, K = pixels. A = X rows (pixel-contiguous = K-major as stored), B = the gradient tile transposed to channel rows.
// Persistent CTAs walk 32-pixel chunks and keep their partial dW in TMEM until the end; the per-CTA partials go to the same
// workspace as the FFMA kernel's and are merged in CTA order by k_pw_reduce_partials (deterministic).
//   warps 0-15  two sets of 256 threads alternating chunks; a set fills its two raw stages with 16-byte cp.async (completion
//               counted by an mbarrier through cp.async.mbarrier.arrive.noinc), then converts, thread = input channel row = TMEM lane:
//               8 x LDS.128 of the row (144-byte pitch: conflict free),
//               split, tcgen05.st of hi and lo into the A ring in tensor memory; the same threads transpose + split the
//               gradient texels into the UMMA canonical B tile ([G_hi | G_lo] as 64 rows) and keep the bias-gradient sums
//   warp  16    per 8 pixels and channel block: one N = 64 MMA (X_hi x [G_hi | G_lo]) and one N = 32 (X_lo x G_hi)
// TMEM map: [0,192) accumulators of the two channel blocks (hi·hi | hi·lo | lo·hi), [192,448) two A slots of 2 x 64 columns.
constexpr int kBwPx = 32, kBwRawStages = 4, kBwSlots = 2, kBwSetWarps = 8, kBwSets = 2, kBwConvWarps = kBwSets * kBwSetWarps,
              kBwMmaWarp = kBwConvWarps, kBwWarps = kBwConvWarps + 1;
constexpr uint32_t kBwRowPitch = 144;                                   // raw X row: 128 B + 16 B pad
constexpr uint32_t kBwRawX = 256 * kBwRowPitch, kBwRawBytes = kBwRawX + kBwPx * 128;   // 36,864 + 4,096
// B tile [channel group of 8][p4][8 channels][16 B]: 64 rows x 32 pixels; 16 bytes of padding per channel group make the
// transposing 4-byte stores of a warp (8 channel quads x 4 pixels) hit 32 different banks (unpadded: 4-way conflicts)
constexpr uint32_t kBwLboB = 128, kBwSboB = (kBwPx / 4) * 128 + 16;
constexpr uint32_t kBwBTile = (2 * kN / 8) * kBwSboB;
constexpr uint32_t kBwAccCols = 3 * kN, kBwRing0 = 2 * kBwAccCols, kBwSlotCols = 2 * 2 * kBwPx;

__global__ void __launch_bounds__(kBwWarps * 32, 1)
    k_pwconv_bwd_weight_tc(const float* __restrict__ X, const float* __restrict__ G, float* __restrict__ part_w,
                           float* __restrict__ part_b, int Cin, int Cp, int H, int W, int chunks_per_img, int nchunks,
                           int CinPad) {
  extern __shared__ __align__(128) unsigned char sm[];
  unsigned char* sRaw = sm;                                              // [kBwRawStages][X rows | G texels]
  unsigned char* sB = sm + kBwRawStages * kBwRawBytes;                   // [kBwSlots][8 KB]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + kBwSlots * kBwBTile);
  uint64_t *full_raw = bars, *full_mma = bars + kBwRawStages, *empty_mma = full_mma + kBwSlots, *done = empty_mma + kBwSlots;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  float* s_db = reinterpret_cast<float*>(tmem_slot + 1);                 // 32 floats
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int HW = H * W, Hp = H + 2, Wp = W + 2;
  const int nblk = Cin / kM;                                            // 1 or 2 channel blocks

  if (tid == 0) {
    for (int s = 0; s < kBwRawStages; ++s) mbar_init(full_raw + s, kBwSetWarps * 32);
    for (int s = 0; s < kBwSlots; ++s) { mbar_init(full_mma + s, kBwSetWarps); mbar_init(empty_mma + s, 1); }
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kBwMmaWarp) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < kN) s_db[tid] = 0.f;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int my_chunks = (int)blockIdx.x < nchunks ? (nchunks - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == kBwMmaWarp) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    uint32_t leader;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(leader));
    const uint32_t b_addr = smem_u32(sB);
    for (int i = 0; i < my_chunks; ++i) {
      const int slot = i % kBwSlots;
      mbar_wait(full_mma + slot, (i / kBwSlots) & 1);
      tc_fence_after();
      if (leader) {
#pragma unroll
        for (int k8 = 0; k8 < kBwPx / 8; ++k8) {
          const uint64_t b = umma_desc(b_addr + slot * kBwBTile + k8 * 2 * kBwLboB, kBwLboB, kBwSboB);
          const uint32_t acc = (i | k8) != 0 ? 1u : 0u;
          for (int mb = 0; mb < nblk; ++mb) {
            const uint32_t d = tb + mb * kBwAccCols;
            const uint32_t a_hi = tb + kBwRing0 + slot * kBwSlotCols + mb * 2 * kBwPx + k8 * 8, a_lo = a_hi + kBwPx;
            umma_tf32_ts(d, a_hi, b, idesc_tf32(2 * kN), acc);        // columns 0-31 hi·hi, 32-63 hi·lo
            umma_tf32_ts(d + 2 * kN, a_lo, b, idesc_tf32(kN), acc);   // columns 64-95 lo·hi
          }
        }
        umma_commit(empty_mma + slot);
        if (i == my_chunks - 1) umma_commit(done);
      }
      __syncwarp();
    }
  } else {
    // Two converter sets of 256 threads alternate chunks (set s: chunks i = s, s+2, ...; raw stages s and s+2, A slot and B
    // tile s): the wait -> LDS -> split -> tcgen05.st -> fence -> arrive chain of one chunk takes about one chunk time at
    // HBM speed, so a single set bounded the kernel at 139 us (measured).
    // thread = input channel row `ct` (< Cin) = TMEM lane ct & 127 of channel block ct >> 7
    const int set = warp / kBwSetWarps, ct = tid & (kBwSetWarps * 32 - 1);
    const int mb = ct >> 7;
    const bool row_ok = ct < Cin;
    const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + kBwRing0 + mb * 2 * kBwPx;
    const int gp = ct >> 3, gc4 = ct & 7;           // gradient texel duty: pixel gp, channels 4*gc4..+3
    float bacc[4] = {0.f, 0.f, 0.f, 0.f};
    // raw ring fill: 16-byte cp.async, lanes along the pixels of a row (a warp instruction = four whole 128-byte lines);
    // each thread posts a no-increment arrive on the stage's mbarrier that fires when ITS copies have landed. (One warp
    // issuing 288 cp.async.bulk of 128 B per chunk instead: 8.8 us per chunk, measured.)
    auto issue = [&](int i) {
      if (i >= my_chunks) return;
      const int s = i % kBwRawStages;
      const int chunk = blockIdx.x + i * gridDim.x;
      const int bv = chunk / chunks_per_img, p0 = (chunk - bv * chunks_per_img) * kBwPx;
      float* raw = reinterpret_cast<float*>(sRaw + s * kBwRawBytes);
      const float* xs = X + ((size_t)bv * Cin + (ct >> 3)) * HW + p0 + (ct & 7) * 4;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int row = (ct >> 3) + 32 * j;
        if (row < Cin) cp_async16(raw + row * (kBwRowPitch / 4) + (ct & 7) * 4, xs + (size_t)(32 * j) * HW, true);
      }
      const int p = p0 + gp, y = p / W, x = p - y * W;   // HW % 32 == 0: every pixel of the chunk exists
      cp_async16(raw + kBwRawX / 4 + gp * 32 + gc4 * 4, G + (((size_t)bv * Hp + y + 1) * Wp + x + 1) * Cp + gc4 * 4, true);
      asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(full_raw + s)) : "memory");
    };
    issue(set); issue(set + kBwSets);
    for (int i = set; i < my_chunks; i += kBwSets) {
      const int s = i % kBwRawStages, slot = i % kBwSlots;   // slot == set
      mbar_wait(full_raw + s, (i / kBwRawStages) & 1);
      const unsigned char* raw = sRaw + s * kBwRawBytes;
      float4 xr[kBwPx / 4];
#pragma unroll
      for (int q = 0; q < kBwPx / 4; ++q)
        xr[q] = row_ok ? *reinterpret_cast<const float4*>(raw + ct * kBwRowPitch + q * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 g = *reinterpret_cast<const float4*>(raw + kBwRawX + gp * 128 + gc4 * 16);
      asm volatile("bar.sync %0, 256;" ::"r"(1 + set) : "memory");   // the set has the raw stage in registers: refill it
      issue(i + kBwRawStages);
      mbar_wait(empty_mma + slot, ((i / kBwSlots) & 1) ^ 1);
      tc_fence_after();
      const uint32_t t0 = lane_base + slot * kBwSlotCols;
#pragma unroll
      for (int q = 0; q < kBwPx / 8; ++q) {
        const float v[8] = {xr[2 * q].x, xr[2 * q].y, xr[2 * q].z, xr[2 * q].w,
                            xr[2 * q + 1].x, xr[2 * q + 1].y, xr[2 * q + 1].z, xr[2 * q + 1].w};
        uint32_t r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = __float_as_uint(v[j]) & 0xffffe000u;
        tmem_st8(t0 + q * 8, r);
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = __float_as_uint(v[j] - __uint_as_float(__float_as_uint(v[j]) & 0xffffe000u));
        tmem_st8(t0 + kBwPx + q * 8, r);
      }
      {
        // B tile: element (row n = channel (+32 for lo), pixel p) at (n>>3)*SBO + (p>>2)*LBO + (n&7)*16 + (p&3)*4
        unsigned char* bt = sB + slot * kBwBTile + (uint32_t)(gp >> 2) * kBwLboB + (gp & 3) * 4;
        const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int n = gc4 * 4 + j;
          const float hi = __uint_as_float(__float_as_uint(gv[j]) & 0xffffe000u);
          *reinterpret_cast<float*>(bt + (uint32_t)(n >> 3) * kBwSboB + (n & 7) * 16) = hi;
          *reinterpret_cast<float*>(bt + (uint32_t)((n + kN) >> 3) * kBwSboB + (n & 7) * 16) = gv[j] - hi;
          bacc[j] += gv[j];
        }
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      proxy_fence();
      __syncwarp();
      if (lane == 0) mbar_arrive(full_mma + slot);
    }
    // epilogue: this thread's dW row (input channel tid, all 32 output channels) from the three partial accumulators
    if (my_chunks > 0 && set == 0) {
      mbar_wait(done, 0);
      tc_fence_after();
    }
    if (row_ok && set == 0) {
      const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + mb * kBwAccCols;
      float* dst = part_w + ((size_t)blockIdx.x * CinPad + ct) * Cp;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        float acc[16];
        uint32_t r[16];
        if (my_chunks > 0) {
          tmem_ld16(taddr + kN + h * 16, r);
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = __uint_as_float(r[j]);
          tmem_ld16(taddr + 2 * kN + h * 16, r);
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] += __uint_as_float(r[j]);
          tmem_ld16(taddr + h * 16, r);
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] += __uint_as_float(r[j]);
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) acc[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 16; j += 4)
          *reinterpret_cast<float4*>(dst + h * 16 + j) = make_float4(acc[j], acc[j + 1], acc[j + 2], acc[j + 3]);
      }
    }
    // bias gradient: 32 threads share a channel quad; fixed-order tree through shared memory would need 1 KB, the sums
    // are added with shared atomics instead and the order across the CTA's 32 pixel lanes is fixed by the barrier below
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = bacc[j];                              // lanes with equal (lane & 7): xor-reduce over lane bits 3,4
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      bacc[j] = v;
    }
    asm volatile("bar.sync 3, 512;" ::: "memory");
    for (int w = 0; w < kBwConvWarps; ++w) {          // warps add in warp order: deterministic
      if (warp == w && lane < 8) {
#pragma unroll
        for (int j = 0; j < 4; ++j) s_db[lane * 4 + j] += bacc[j];
      }
      asm volatile("bar.sync 3, 512;" ::: "memory");
    }
    if (tid < kN) part_b[(size_t)blockIdx.x * Cp + tid] = s_db[tid];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kBwMmaWarp) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
  }
}
inline size_t bw_smem_bytes() { return (size_t)kBwRawStages * kBwRawBytes + kBwSlots * kBwBTile + 512; }
inline bool bw_usable(int Cin, int Cout, int HW) { return Cout <= kN && (Cin == 128 || Cin == 256) && HW % kBwPx == 0; }

inline size_t smem_bytes(int Cin) { return 2 * (size_t)kN * Cin * 4 + 512 + (size_t)kGroups * kRawDepth * kKs * kM * 4; }
inline bool usable(int Cin, int Cout, int HW) { return Cout <= kN && Cin % kKs == 0 && Cin >= kKs && Cin <= 256 && HW % 4 == 0; }
}  // namespace tc

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
inline bool aligned32(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 31) == 0; }
constexpr int kFwdSmem = kStages * (kKc * kPx + kKc * kCo) * 4;      // 73,728 B -> 3 CTAs per SM
constexpr int kBwdDataSmem = (kCo * kPx + kCo * kKn) * 4;            // 65,536 B -> 3 CTAs per SM
constexpr int kBwdWeightSmem = kWgStages * kWgStageFloats * 4;       // 90,112 B -> 2 CTAs per SM
// developer knob for A/B measurements (tools/bench_process_features.py): LT_B200_PWCONV=ffma keeps the forward on the
// FMA pipe. Read once; both paths compute the same fp32 result to 1e-6.
bool force_ffma() {
  static const bool v = [] { const char* e = getenv("LT_B200_PWCONV"); return e != nullptr && e[0] == 'f'; }();
  return v;
}

int weight_workers(int BV, int H, int W) {   // an upper bound for both weight-gradient kernels (the tcgen05 one uses <= 148)
  const long long total = (long long)BV * ((H * W + kWgPx - 1) / kWgPx);
  const long long workers = 2LL * sm_count();  // persistent: 2 CTAs on every SM
  return (int)(total < workers ? total : workers);
}

}  // namespace

size_t pwconv_ws_bytes(int BV, int Cin, int Cout, int H, int W) {
  const size_t Cp = (size_t)padded_channels(Cout);
  const size_t CinPad = (size_t)((Cin + kKn - 1) / kKn) * kKn;
  const size_t fwd = (size_t)Cin * Cp * 4;                                            // Wᵀ
  const size_t wgt = (size_t)weight_workers(BV, H, W) * (CinPad * Cp + Cp) * 4;       // partial sums
  return fwd > wgt ? fwd : wgt;
}

cudaError_t launch_pwconv_fwd(const float* x, const float* weight, const float* bias, float* out_hwcp, void* ws,
                              int BV, int Cin, int Cout, int H, int W, cudaStream_t st) {
  const int Cp = padded_channels(Cout), HW = H * W;
  if (tc::usable(Cin, Cout, HW) && aligned16(x) && aligned32(out_hwcp) && !force_ffma()) {   // epilogue: 32-byte stores
    k_pw_zero_border<<<BV, 256, 0, st>>>(out_hwcp, Cp, H, W);
    const int tiles_per_img = (HW + tc::kM - 1) / tc::kM, ntiles = BV * tiles_per_img;
    const int smem = (int)tc::smem_bytes(Cin);
    cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(&tc::k_pwconv_fwd_tc), (size_t)smem);
    if (e != cudaSuccess) return e;
    tc::k_pwconv_fwd_tc<<<ntiles < sm_count() ? ntiles : sm_count(), tc::kWarps * 32, smem, st>>>(x, weight, bias, out_hwcp, Cin, Cout,
                                                                                  Cp, H, W, tiles_per_img, ntiles);
    return cudaGetLastError();
  }
  float* Wt = static_cast<float*>(ws);
  k_pw_transpose_weight<<<(Cin * Cp + 255) / 256, 256, 0, st>>>(weight, Wt, Cin, Cout, Cp);
  k_pw_zero_border<<<BV, 256, 0, st>>>(out_hwcp, Cp, H, W);
  dim3 grid((unsigned)((HW + kPx - 1) / kPx), (unsigned)BV, (unsigned)(Cp / kCo));
  const bool vec = (HW % 4 == 0) && aligned16(x);
  auto kern = vec ? k_pwconv_fwd<true> : k_pwconv_fwd<false>;
  cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(kern), (size_t)kFwdSmem);
  if (e != cudaSuccess) return e;
  kern<<<grid, 256, kFwdSmem, st>>>(x, Wt, bias, out_hwcp, Cin, Cout, Cp, H, W);
  return cudaGetLastError();
}

cudaError_t launch_pwconv_bwd_data(const float* grad_hwcp, const float* weight, float* grad_x, int BV, int Cin,
                                   int Cout, int H, int W, cudaStream_t st) {
  const int Cp = padded_channels(Cout), HW = H * W;
  if (tc::bd_usable(Cin, Cout) && !force_ffma()) {
    const int tiles_per_img = (HW + tc::kM - 1) / tc::kM, ntiles = BV * tiles_per_img;
    const int smem = (int)tc::bd_smem_bytes(Cin);
    cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(&tc::k_pwconv_bwd_data_tc), (size_t)smem);
    if (e != cudaSuccess) return e;
    tc::k_pwconv_bwd_data_tc<<<ntiles < sm_count() ? ntiles : sm_count(), tc::kBdWarps * 32, smem, st>>>(grad_hwcp, weight, grad_x, Cin, Cout,
                                                                                         Cp, H, W, tiles_per_img, ntiles);
    return cudaGetLastError();
  }
  dim3 grid((unsigned)((HW + kPx - 1) / kPx), (unsigned)BV, (unsigned)((Cin + kKn - 1) / kKn));
  const bool vec = (HW % 4 == 0) && aligned16(grad_x);
  auto kern = vec ? k_pwconv_bwd_data<true> : k_pwconv_bwd_data<false>;
  cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(kern), (size_t)kBwdDataSmem);
  if (e != cudaSuccess) return e;
  // more than 32 output channels: one pass per 32-channel block of G, later passes add into grad_x
  for (int co0 = 0; co0 < Cout; co0 += kCo)
    kern<<<grid, 256, kBwdDataSmem, st>>>(grad_hwcp, weight, grad_x, Cin, Cout, Cp, H, W, co0, co0 > 0 ? 1 : 0);
  return cudaGetLastError();
}

cudaError_t launch_pwconv_bwd_weight(const float* x, const float* grad_hwcp, float* grad_weight, float* grad_bias,
                                     void* ws, int BV, int Cin, int Cout, int H, int W, cudaStream_t st) {
  const int Cp = padded_channels(Cout), HW = H * W;
  const int CinPad = ((Cin + kKn - 1) / kKn) * kKn;
  const int nchunk = (HW + kWgPx - 1) / kWgPx, total = BV * nchunk;
  const int nw = weight_workers(BV, H, W);
  if (tc::bw_usable(Cin, Cout, HW) && aligned16(x) && !force_ffma()) {
    const int chunks_per_img = HW / tc::kBwPx, nchunks = BV * chunks_per_img;
    const int nctas = nchunks < sm_count() ? nchunks : sm_count();
    float* pw = static_cast<float*>(ws);
    float* pb = pw + (size_t)nctas * CinPad * Cp;
    const int smem = (int)tc::bw_smem_bytes();
    cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(&tc::k_pwconv_bwd_weight_tc), (size_t)smem);
    if (e != cudaSuccess) return e;
    tc::k_pwconv_bwd_weight_tc<<<nctas, tc::kBwWarps * 32, smem, st>>>(x, grad_hwcp, pw, pb, Cin, Cp, H, W, chunks_per_img,
                                                                      nchunks, CinPad);
    dim3 rg((unsigned)(Cin + 1), (unsigned)(Cp / 32));
    k_pw_reduce_partials<<<rg, 256, 0, st>>>(pw, pb, grad_weight, grad_bias, Cin, Cout, Cp, CinPad, nctas);
    return cudaGetLastError();
  }
  float* part_w = static_cast<float*>(ws);
  float* part_b = part_w + (size_t)nw * CinPad * Cp;
  dim3 grid((unsigned)nw, (unsigned)(CinPad / kKn), (unsigned)(Cp / kCo));
  const bool vec = (HW % 4 == 0) && aligned16(x);
  auto kern = vec ? k_pwconv_bwd_weight<true> : k_pwconv_bwd_weight<false>;
  cudaError_t e = smem_opt_in(reinterpret_cast<const void*>(kern), (size_t)kBwdWeightSmem);
  if (e != cudaSuccess) return e;
  kern<<<grid, 256, kBwdWeightSmem, st>>>(x, grad_hwcp, part_w, part_b, Cin, Cout, Cp, H, W, nchunk, total, CinPad);
  dim3 rgrid((unsigned)(Cin + 1), (unsigned)(Cp / 32));
  k_pw_reduce_partials<<<rgrid, 256, 0, st>>>(part_w, part_b, grad_weight, grad_bias, Cin, Cout, Cp, CinPad, nw);
  return cudaGetLastError();
}

}  // namespace lt
