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
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kPixTile = 64;   // pixels (of one padded row) per CTA
constexpr int kChanTile = 32;  // channels per CTA (Cp is a multiple of 32)

// one CTA = (64 padded pixels of padded row yp, bv, 32-channel block). smem tile [32][65]: the planar side moves 64
// consecutive pixels per channel (256-byte runs), the channels-last side 32 consecutive floats per pixel (128 bytes).
template <bool TO_HWCP>
__global__ void __launch_bounds__(256) k_feat_layout(const float* __restrict__ src, float* __restrict__ dst, int C,
                                                     int Cp, int H, int W, int nchunk) {
  __shared__ float tile[kChanTile * (kPixTile + 1)];
  const int Hp = H + 2, Wp = W + 2;
  const int chunk = blockIdx.x % nchunk, yp = blockIdx.x / nchunk;
  const int bv = blockIdx.y, c0 = blockIdx.z * kChanTile;
  const int xp0 = chunk * kPixTile;
  const int npix = min(kPixTile, Wp - xp0);
  const int y = yp - 1;
  const bool row_in = (y >= 0 && y < H);
  const size_t planar_base = ((size_t)bv * C + c0) * H * W + (size_t)(row_in ? y : 0) * W;
  const size_t hwcp_base = (((size_t)bv * Hp + yp) * Wp + xp0) * Cp + c0;

  if (TO_HWCP) {
    for (int i = threadIdx.x; i < kChanTile * kPixTile; i += blockDim.x) {
      const int c = i / kPixTile, p = i % kPixTile;
      const int x = xp0 + p - 1;
      float v = 0.f;
      if (row_in && p < npix && x >= 0 && x < W && c0 + c < C) v = __ldg(src + planar_base + (size_t)c * H * W + x);
      tile[c * (kPixTile + 1) + p] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npix * kChanTile; i += blockDim.x) {
      const int p = i / kChanTile, c = i % kChanTile;
      dst[hwcp_base + (size_t)p * Cp + c] = tile[c * (kPixTile + 1) + p];
    }
  } else {
    if (!row_in) return;
    for (int i = threadIdx.x; i < npix * kChanTile; i += blockDim.x) {
      const int p = i / kChanTile, c = i % kChanTile;
      tile[c * (kPixTile + 1) + p] = __ldg(src + hwcp_base + (size_t)p * Cp + c);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kChanTile * kPixTile; i += blockDim.x) {
      const int c = i / kPixTile, p = i % kPixTile;
      const int x = xp0 + p - 1;
      if (p < npix && x >= 0 && x < W && c0 + c < C) dst[planar_base + (size_t)c * H * W + x] = tile[c * (kPixTile + 1) + p];
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
  const int nchunk = (W + 2 + kPixTile - 1) / kPixTile;
  dim3 grid((unsigned)(nchunk * (H + 2)), (unsigned)BV, (unsigned)(Cp / kChanTile));
  if (to_hwcp)
    k_feat_layout<true><<<grid, 256, 0, st>>>(src, dst, C, Cp, H, W, nchunk);
  else
    k_feat_layout<false><<<grid, 256, 0, st>>>(src, dst, C, Cp, H, W, nchunk);
  return cudaGetLastError();
}

}  // namespace lt
