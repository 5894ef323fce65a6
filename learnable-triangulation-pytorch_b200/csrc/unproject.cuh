// unproject.cuh — geometry shared by the K1 forward and backward kernels.
//
// Work decomposition (both directions)
//   * a lane owns kCPL = 8 consecutive channels; LPV = Cp/8 lanes (4 for the 32-channel features of
//     triangulation.py:239) cover one texel, so one tap is ONE contiguous Cp*4-byte read (LDG.E.256 per
//     lane, 1 L1 wavefront per 128 B) and a warp works on G = 32/LPV voxels at a time;
//   * a lane group walks a run of kRun = 8 consecutive voxels along the fastest volume axis, so that the
//     8 results of one channel leave as one full 32-byte sector (STG.E.256);
//   * a CTA (8 warps) covers a compact (4*G0)x(2*G1)x8 block of the volume (8x8x8 for C=32): the block's
//     image footprint (~17x17 texels per view at Human3.6M geometry) is what has to stay in L1;
//   * projection (mvn/utils/multiview.py:80-101 + op.py:121-130 + the index arithmetic of
//     F.grid_sample) is done ONCE per (voxel, view) by one lane ("phase A") and handed to the gather
//     lanes through a per-warp shared-memory descriptor table (offset + 4 tap weights).
#pragma once
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kRun = 8;           // voxels per lane-group run
constexpr int kCPL = 8;           // channels per lane
constexpr int kWarpsPerCta = 8;   // arranged 4 (axis 0) x 2 (axis 1)
constexpr int kViewChunk = 4;     // views resolved per phase-A round
constexpr int kMaxViewsSmem = 64; // projection matrices staged per CTA

constexpr int kOffMask = 0x0fffffff;
constexpr int kOffDx = 1 << 28;   // second column tap is a distinct texel
constexpr int kOffDy = 1 << 29;   // second row tap is a distinct texel

template <int LPV>
struct TileGeom {
  static constexpr int G = 32 / LPV;          // voxels (runs) in flight per warp
  static constexpr int G1 = G < 4 ? G : 4;    // runs along axis 1
  static constexpr int G0 = G / G1;           // runs along axis 0
  static constexpr int T0 = 4 * G0;           // CTA tile extent, axis 0
  static constexpr int T1 = 2 * G1;           // CTA tile extent, axis 1
  static constexpr int kSlots = G * kRun;     // descriptor slots per warp and view
};

struct UnprojArgs {
  const float* feat;   // (B,V,H,W,Cp)
  const float* proj;   // (B,V,3,4)
  const float* coord;  // (B,nvox,3)
  const float* conf;   // (B,V,C) or null
  float* out;          // fwd: (B,C,nvox)
  const float* gout;   // bwd: (B,C,nvox)
  float* gfeat;        // bwd: (B,V,H,W,Cp), pre-zeroed
  float* gconf_part;   // bwd: (B,tiles,V,C) per-CTA partial sums of grad_conf, or null
  int B, V, C, Cp, H, W, N0, N1, N2;
  int nt0, nt1, nt2, ncb;  // tile grid and channel blocks
  int align_corners;
  long long nvox;
};

// One (voxel, view) sampling descriptor. Mirrors, in order:
//   multiview.py:96   [x y z 1] @ P.t()   (k-sequential fp32 accumulation like an sgemm inner loop)
//   op.py:121-124     invalid = z<=0 ; u=x/z ; v=y/z
//   op.py:128-129     gx=2(u/H-.5) ; gy=2(v/W-.5)      (tensor/python-int = multiply by fp32 reciprocal on CUDA)
//   ATen grid_sampler_unnormalize + bilinear corner weights, zero padding.
// Returns off=-1 when the view contributes exactly 0 (behind camera, or all four taps outside).
__device__ __forceinline__ void make_desc(float x, float y, float z, const float4& r0, const float4& r1,
                                          const float4& r2, int H, int W, int Cp, float invH, float invW,
                                          int align_corners, int& off, float4& wts) {
  float px = x * r0.x; px = fmaf(y, r0.y, px); px = fmaf(z, r0.z, px); px += r0.w;
  float py = x * r1.x; py = fmaf(y, r1.y, py); py = fmaf(z, r1.z, py); py += r1.w;
  float pz = x * r2.x; pz = fmaf(y, r2.y, pz); pz = fmaf(z, r2.z, pz); pz += r2.w;
  off = -1;
  wts = make_float4(0.f, 0.f, 0.f, 0.f);
  if (pz <= 0.f) return;
  float u = __fdiv_rn(px, pz), v = __fdiv_rn(py, pz);
  float gx = 2.f * (u * invH - 0.5f);
  float gy = 2.f * (v * invW - 0.5f);
  float ix, iy;
  if (align_corners) {
    ix = ((gx + 1.f) * 0.5f) * (float)(W - 1);
    iy = ((gy + 1.f) * 0.5f) * (float)(H - 1);
  } else {
    ix = fmaf(gx + 1.f, (float)W, -1.f) * 0.5f;
    iy = fmaf(gy + 1.f, (float)H, -1.f) * 0.5f;
  }
  // anything outside (-1, W) x (-1, H) has no in-bounds tap (also catches inf/nan)
  if (!(ix > -1.f && ix < (float)W && iy > -1.f && iy < (float)H)) return;
  float fx0 = floorf(ix), fy0 = floorf(iy);
  float ax1 = (fx0 + 1.f) - ix, ax0 = ix - fx0;  // weights of column x0 and x0+1
  float ay1 = (fy0 + 1.f) - iy, ay0 = iy - fy0;
  int x0 = (int)fx0, y0 = (int)fy0;
  bool okx0 = x0 >= 0, okx1 = x0 + 1 < W, oky0 = y0 >= 0, oky1 = y0 + 1 < H;
  wts.x = (okx0 && oky0) ? ax1 * ay1 : 0.f;  // nw
  wts.y = (okx1 && oky0) ? ax0 * ay1 : 0.f;  // ne
  wts.z = (okx0 && oky1) ? ax1 * ay0 : 0.f;  // sw
  wts.w = (okx1 && oky1) ? ax0 * ay0 : 0.f;  // se
  int xb = okx0 ? x0 : 0, yb = oky0 ? y0 : 0;
  off = (yb * W + xb) * Cp;
  if (okx0 && okx1) off |= kOffDx;
  if (oky0 && oky1) off |= kOffDy;
}

void fill_unproj_tiling(UnprojArgs& a, int lpv);
int unproj_lpv(int Cp);

}  // namespace lt
