// layout.cu — K0: (B·V,C,H,W) channel-planar  <->  the gather layout "hwcp": (B·V, H+2, W+2, Cp) channels-last.
//
// The reference samples channel-planar maps with F.grid_sample (mvn/utils/op.py:114-134); every tap then touches
// C different cache lines. In hwcp
//   * one tap = one contiguous Cp*4-byte run (128 bytes = one line for the 32-channel features of
//     triangulation.py:239): one LDG.128 per lane in the gather, one 16-byte vector reduction per lane in the scatter;
//   * every map carries a one-texel zero border, so the four bilinear taps of any voxel that has at least one tap
//     inside the image are always addressable: grid_sample's zeros padding becomes data instead of per-tap bounds
//     logic in the hot loops (the gradient map uses the same shape; its border is scratch);
//   * channels >= C are zero filled (Cp = lt_padded_channels(C)).
// (A 16-channel-group variant of this layout, which makes every image row of a group one bulk-copy run for a
//  shared-memory-staged K1, was measured and set aside: see DESIGN.md "K1 tile experiment".)
// Pure HBM streaming: algorithmic bytes = BV*H*W*C*4 read + BV*(H+2)*(W+2)*Cp*4 written.
#include <mutex>

#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kPixTile = 64;   // consecutive padded pixels per CTA
constexpr int kChanTile = 32;  // channels per CTA (Cp is a multiple of 32)

// one CTA = (64 CONSECUTIVE padded pixels of one map in row-major order, bv, 32-channel block): a tile may straddle a row
// boundary, so every tile but the last of a map is full (per-row chunks left the second chunk of a 98-pixel row 47 %
// empty). smem tile [32][65]: the planar side moves runs of consecutive pixels per channel, the channels-last side 32
// consecutive floats per pixel (128 bytes).
template <bool TO_HWCP>
__global__ void __launch_bounds__(256) k_feat_layout(const float* __restrict__ src, float* __restrict__ dst, int C,
                                                     int Cp, int H, int W) {
  __shared__ float tile[kChanTile * (kPixTile + 1)];
  const int Hp = H + 2, Wp = W + 2;
  const int bv = blockIdx.y, c0 = blockIdx.z * kChanTile;
  const int P0 = blockIdx.x * kPixTile;
  const int npix = min(kPixTile, Hp * Wp - P0);
  const size_t hwcp_base = ((size_t)bv * Hp * Wp + P0) * Cp + c0;

  // planar side: thread -> pixel t%64, channels t/64 + 4k (no index arithmetic in the loop)
  // hwcp side  : thread -> channel quad t%8, pixels t/8 + 32k (one 16-byte vector per thread, 128 bytes per pixel)
  // tile[c*65 + p]: both access patterns are bank-conflict free (65 = 1 mod 32).
  const int pp = threadIdx.x & (kPixTile - 1), pc = threadIdx.x >> 6;
  const int hq = threadIdx.x & 7, hp = threadIdx.x >> 3;
  const int P = P0 + pp;
  const int yp = P / Wp, xp = P - yp * Wp;
  const int y = yp - 1, x = xp - 1;
  const bool px_in = pp < npix && y >= 0 && y < H && x >= 0 && x < W;
  const size_t planar = ((size_t)bv * C + c0 + pc) * H * W + (size_t)(px_in ? y : 0) * W + (px_in ? x : 0);
  // the channel-planar pointer walks 4 planes per step: kept as an opaque 64-bit pair and advanced by a constant (the
  // compiler otherwise re-derives (bv*C + c)*H*W + y*W + x with 64-bit multiplies for every one of the 8 accesses:
  // MEASURED with ncu's source view, 304 instructions per thread for 8 floats in and 8 out)
  const size_t plane4 = (size_t)4 * H * W;
  // the channels-last pointer of this thread's channel quad at pixel 0 of the tile (one of the two is used)
  float* hw_dst = dst + hwcp_base + 4 * hq;
  const float* hw_src = src + hwcp_base + 4 * hq;
  if (TO_HWCP) {
    asm volatile("" : "+l"(hw_dst));
    const float* sp = src + planar;
    asm volatile("" : "+l"(sp));
#pragma unroll
    for (int k = 0; k < kChanTile / 4; ++k) {
      const int c = pc + 4 * k;
      float v = 0.f;
      if (px_in && c0 + c < C) v = __ldg(sp);
      sp += plane4;
      tile[c * (kPixTile + 1) + pp] = v;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int p = hp + 32 * k;
      if (p < npix) {
        const float* t = tile + (4 * hq) * (kPixTile + 1) + p;
        const float4 v = make_float4(t[0], t[kPixTile + 1], t[2 * (kPixTile + 1)], t[3 * (kPixTile + 1)]);
        *reinterpret_cast<float4*>(hw_dst + (size_t)p * Cp) = v;
      }
    }
  } else {
    asm volatile("" : "+l"(hw_src));
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int p = hp + 32 * k;
      if (p < npix) {
        const float4 v = ldg128_stream(hw_src + (size_t)p * Cp);
        float* t = tile + (4 * hq) * (kPixTile + 1) + p;
        t[0] = v.x; t[kPixTile + 1] = v.y; t[2 * (kPixTile + 1)] = v.z; t[3 * (kPixTile + 1)] = v.w;
      }
    }
    __syncthreads();
    float* dp = dst + planar;
    asm volatile("" : "+l"(dp));
#pragma unroll
    for (int k = 0; k < kChanTile / 4; ++k) {
      const int c = pc + 4 * k;
      if (px_in && c0 + c < C) *dp = tile[c * (kPixTile + 1) + pp];
      dp += plane4;
    }
  }
}

int padded_channels(int C) {
  // 4 channels per lane and 8/16/32 lanes per voxel (unproject.cuh); above 128, blocks of 128 channels
  if (C <= 32) return 32;
  if (C <= 64) return 64;
  return ((C + 127) / 128) * 128;
}

size_t feat_layout_elems(int BV, int C, int H, int W) {
  if (BV <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  return (size_t)BV * padded_channels(C) * (size_t)(H + 2) * (size_t)(W + 2);
}

cudaError_t launch_feat_layout(const float* src, float* dst, int BV, int C, int H, int W, bool to_hwcp,
                               cudaStream_t st) {
  if (BV == 0 || C == 0 || H * W == 0) return cudaSuccess;
  const int Cp = padded_channels(C);
  const int ntile = ((H + 2) * (W + 2) + kPixTile - 1) / kPixTile;
  dim3 grid((unsigned)ntile, (unsigned)BV, (unsigned)(Cp / kChanTile));
  if (to_hwcp)
    k_feat_layout<true><<<grid, 256, 0, st>>>(src, dst, C, Cp, H, W);
  else
    k_feat_layout<false><<<grid, 256, 0, st>>>(src, dst, C, Cp, H, W);
  return cudaGetLastError();
}

// SM count of the current device, queried once per device (148 on B200): persistent kernels size their grids with it
int sm_count() {
  static int cached[64] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    cached[dev] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
  }
  return cached[dev];
}

// ---- one-time opt-in for more than 48 KB of dynamic shared memory (see common.cuh) ----
cudaError_t smem_opt_in(const void* func, size_t bytes) {
  struct Entry { const void* func; int dev; size_t bytes; };
  static Entry table[256];
  static int count = 0;
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  Entry* hit = nullptr;
  for (int i = 0; i < count; ++i)
    if (table[i].func == func && table[i].dev == dev) { hit = &table[i]; break; }
  if (hit != nullptr && hit->bytes >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  if (hit != nullptr) hit->bytes = bytes;
  else if (count < 256) table[count++] = Entry{func, dev, bytes};
  return cudaSuccess;
}

}  // namespace lt
