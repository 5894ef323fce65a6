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
// All three are fp32 FMA chains in a fixed order (k ascending; partial sums of the weight gradient merged in worker
// order): deterministic, and within 1e-6 of ATen's fp32 convolution. The work is GEMM shaped (M = pixels, N = 32,
// K = 256) but stays on the FMA pipe as packed FFMA2: the parity bar is the fp32 oracle, and single-pass TF32 tensor-core
// products are 1e-3 away from it (a 3xTF32 tcgen05 version is the listed follow-up). Bounds: forward and
// backward-data read or write the (B·V,256,H,W) tensor once (HBM floor) and issue 2·Cin·Cout flops per pixel (FFMA2 floor).
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

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
constexpr int kFwdSmem = kStages * (kKc * kPx + kKc * kCo) * 4;      // 73,728 B -> 3 CTAs per SM
constexpr int kBwdDataSmem = (kCo * kPx + kCo * kKn) * 4;            // 65,536 B -> 3 CTAs per SM
constexpr int kBwdWeightSmem = kWgStages * kWgStageFloats * 4;       // 90,112 B -> 2 CTAs per SM
constexpr int kWeightWorkers = 2 * 148;                              // persistent: 2 CTAs on each of the 148 SMs

int weight_workers(int BV, int H, int W) {
  const long long total = (long long)BV * ((H * W + kWgPx - 1) / kWgPx);
  return (int)(total < kWeightWorkers ? total : kWeightWorkers);
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
  float* Wt = static_cast<float*>(ws);
  k_pw_transpose_weight<<<(Cin * Cp + 255) / 256, 256, 0, st>>>(weight, Wt, Cin, Cout, Cp);
  k_pw_zero_border<<<BV, 256, 0, st>>>(out_hwcp, Cp, H, W);
  dim3 grid((unsigned)((HW + kPx - 1) / kPx), (unsigned)BV, (unsigned)(Cp / kCo));
  const bool vec = (HW % 4 == 0) && aligned16(x);
  auto kern = vec ? k_pwconv_fwd<true> : k_pwconv_fwd<false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem);
  if (e != cudaSuccess) return e;
  kern<<<grid, 256, kFwdSmem, st>>>(x, Wt, bias, out_hwcp, Cin, Cout, Cp, H, W);
  return cudaGetLastError();
}

cudaError_t launch_pwconv_bwd_data(const float* grad_hwcp, const float* weight, float* grad_x, int BV, int Cin,
                                   int Cout, int H, int W, cudaStream_t st) {
  const int Cp = padded_channels(Cout), HW = H * W;
  dim3 grid((unsigned)((HW + kPx - 1) / kPx), (unsigned)BV, (unsigned)((Cin + kKn - 1) / kKn));
  const bool vec = (HW % 4 == 0) && aligned16(grad_x);
  auto kern = vec ? k_pwconv_bwd_data<true> : k_pwconv_bwd_data<false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdDataSmem);
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
  float* part_w = static_cast<float*>(ws);
  float* part_b = part_w + (size_t)nw * CinPad * Cp;
  dim3 grid((unsigned)nw, (unsigned)(CinPad / kKn), (unsigned)(Cp / kCo));
  const bool vec = (HW % 4 == 0) && aligned16(x);
  auto kern = vec ? k_pwconv_bwd_weight<true> : k_pwconv_bwd_weight<false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdWeightSmem);
  if (e != cudaSuccess) return e;
  kern<<<grid, 256, kBwdWeightSmem, st>>>(x, grad_hwcp, part_w, part_b, Cin, Cout, Cp, H, W, nchunk, total, CinPad);
  dim3 rgrid((unsigned)(Cin + 1), (unsigned)(Cp / 32));
  k_pw_reduce_partials<<<rgrid, 256, 0, st>>>(part_w, part_b, grad_weight, grad_bias, Cin, Cout, Cp, CinPad, nw);
  return cudaGetLastError();
}

}  // namespace lt
