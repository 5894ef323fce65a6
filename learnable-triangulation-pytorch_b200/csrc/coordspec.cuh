// coordspec.cuh — coordinate volumes computed IN the consuming kernel.
//
// The reference builds `coord_volumes` (B,N0,N1,N2,3) per item in python (mvn/models/triangulation.py:306-341 with
// mvn/utils/volumetric.py:87-114) and every op then reads 12 bytes per voxel. K6 (coordvol.cu) builds the same tensor
// in one launch from an 18-float spec per item; the kernels that only CONSUME coordinates (K1 forward/backward, K2
// statistics/backward) can evaluate the same arithmetic per voxel instead of reading the tensor:
//   spec per item: pos[3] step[3] centre[3] rot[9] (row-major) = 18 floats, rounded to fp32 on the host exactly where
//   the reference's torch ops round their python scalars (volumetric.coord_volume_specs).
// `coord_of` is THE definition (K6 calls it too): un-contracted where ATen runs separate kernels, FMA chain for the
// K = 3 matmul, so a kernel that computes its coordinates gets bit-identical values to one that reads K6's tensor.
#pragma once
#include <cuda_runtime.h>

namespace lt {

constexpr int kCoordSpec = 18;

__device__ __forceinline__ void coord_of(const float* __restrict__ s, int i0, int i1, int i2, float (&o)[3]) {
  const float g0 = __fadd_rn(s[0], __fmul_rn(s[3], (float)i0));
  const float g1 = __fadd_rn(s[1], __fmul_rn(s[4], (float)i1));
  const float g2 = __fadd_rn(s[2], __fmul_rn(s[5], (float)i2));
  const float d0 = __fsub_rn(g0, s[6]), d1 = __fsub_rn(g1, s[7]), d2 = __fsub_rn(g2, s[8]);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    float acc = __fmul_rn(s[9 + 3 * k], d0);
    acc = __fmaf_rn(s[10 + 3 * k], d1, acc);
    acc = __fmaf_rn(s[11 + 3 * k], d2, acc);
    o[k] = __fadd_rn(acc, s[6 + k]);
  }
}

// where a kernel gets its voxel centres from: the tensor (coord != null) or the per-item spec (spec != null).
// transfer: the CMU -> Human3.6M axis swap of triangulation.py:336-339, out[i0,i1,i2] = cv[i0, i2, N1-1-i1].
struct CoordSrc {
  const float* coord;  // (B, nvox, 3) or null
  const float* spec;   // (B, 18) or null
  int N1, N2;          // volume extents of the two inner axes (spec mode: voxel index -> (i0, i1, i2))
  int transfer;
};

__device__ __forceinline__ void coord_of_voxel(const float* __restrict__ s, int transfer, int N1, int i0, int i1, int i2,
                                               float (&o)[3]) {
  if (transfer) coord_of(s, i0, i2, N1 - 1 - i1, o);
  else coord_of(s, i0, i1, i2, o);
}

}  // namespace lt
