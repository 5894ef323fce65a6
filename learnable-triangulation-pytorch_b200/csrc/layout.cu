// layout.cu — K0: (B·V,C,H,W) channel-planar  <->  (B·V,H,W,Cp) channels-last, zero padded.
//
// The reference samples channel-planar maps with F.grid_sample (mvn/utils/op.py:114-134); every tap
// then touches C different cache lines. Channels-last makes one tap one contiguous Cp*4-byte run, which
// is what the gather (K1 fwd) and the vector-reduction scatter (K1 bwd) are built around.
// Pure HBM streaming: algorithmic bytes = BV*H*W*(C+Cp)*4.
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kPixTile = 32;   // pixels per CTA tile
constexpr int kChanTile = 64;  // channels per CTA tile (Cp is a multiple of 8)

// smem tile [kChanTile][33]: the planar side moves 32 consecutive pixels per channel (128 B lines), the
// channels-last side up to 64 consecutive floats per pixel. +1 padding keeps both sides conflict free.
template <bool TO_HWC>
__global__ void __launch_bounds__(256) k_feat_layout(const float* __restrict__ src, float* __restrict__ dst,
                                                     int C, int Cp, int HW) {
  __shared__ float tile[kChanTile * 33];
  const int bv = blockIdx.y;
  const int p0 = blockIdx.x * kPixTile;
  const int c0 = blockIdx.z * kChanTile;
  const int npix = min(kPixTile, HW - p0);
  const int nch = min(kChanTile, Cp - c0);  // channels of this tile on the padded side
  const float* planar_src = src + ((size_t)bv * C + c0) * HW + p0;
  float* planar_dst = dst + ((size_t)bv * C + c0) * HW + p0;
  const float* hwc_src = src + ((size_t)bv * HW + p0) * Cp + c0;
  float* hwc_dst = dst + ((size_t)bv * HW + p0) * Cp + c0;

  if (TO_HWC) {
    for (int i = threadIdx.x; i < nch * kPixTile; i += blockDim.x) {
      int c = i / kPixTile, p = i % kPixTile;
      float v = 0.f;
      if (c0 + c < C && p < npix) v = __ldg(planar_src + (size_t)c * HW + p);
      tile[c * 33 + p] = v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npix * nch; i += blockDim.x) {
      int p = i / nch, c = i % nch;
      hwc_dst[(size_t)p * Cp + c] = tile[c * 33 + p];
    }
  } else {
    for (int i = threadIdx.x; i < npix * nch; i += blockDim.x) {
      int p = i / nch, c = i % nch;
      tile[c * 33 + p] = __ldg(hwc_src + (size_t)p * Cp + c);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < nch * kPixTile; i += blockDim.x) {
      int c = i / kPixTile, p = i % kPixTile;
      if (c0 + c < C && p < npix) planar_dst[(size_t)c * HW + p] = tile[c * 33 + p];
    }
  }
}

int padded_channels(int C) {
  // 4 channels per lane and 8/16/32 lanes per voxel (unproject.cuh); above 128, blocks of 128 channels
  if (C <= 32) return 32;
  if (C <= 64) return 64;
  return ((C + 127) / 128) * 128;
}

cudaError_t launch_feat_layout(const float* src, float* dst, int BV, int C, int H, int W, bool to_hwc,
                               cudaStream_t st) {
  if (BV == 0 || C == 0 || H * W == 0) return cudaSuccess;
  const int Cp = padded_channels(C);
  const int HW = H * W;
  dim3 grid((HW + kPixTile - 1) / kPixTile, BV, (Cp + kChanTile - 1) / kChanTile);
  if (to_hwc)
    k_feat_layout<true><<<grid, 256, 0, st>>>(src, dst, C, Cp, HW);
  else
    k_feat_layout<false><<<grid, 256, 0, st>>>(src, dst, C, Cp, HW);
  return cudaGetLastError();
}

}  // namespace lt
