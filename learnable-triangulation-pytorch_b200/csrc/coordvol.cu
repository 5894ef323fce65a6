// coordvol.cu — K6: coordinate volumes.
//
// Replaces the per-item python block of mvn/models/triangulation.py:306-341 (meshgrid, three strided assignments,
// subtract centre, 3x3 matmul with the rotation of mvn/utils/volumetric.py:87-114, add centre, optional CMU->Human3.6M
// axis swap: ~20 launches and a host round trip per batch item) with one launch for the whole batch.
// The arithmetic is the reference's, operation by operation, in fp32 and un-contracted where ATen runs separate kernels:
//   g_c   = pos_c + step_c * idx_c                 (triangulation.py:311-313: one multiply kernel, one add kernel)
//   d     = g - centre                             (:331)
//   r_k   = rot[k,0]*d_0 + rot[k,1]*d_1 + rot[k,2]*d_2   (:332 -> volumetric.py:108 `rot.mm(coord.t())`, K=3 FMA chain)
//   out_k = r_k + centre_k                         (:333)
// transfer (triangulation.py:336-339): out[i0,i1,i2] = cv[i0, i2, N1-1-i1] (permute(0,2,1,3), then reverse axis 1).
// Pure streaming write: 12 bytes per voxel.
#include "common.cuh"
#include "kernels.h"

namespace lt {

// spec per item: pos[3] step[3] centre[3] rot[9] (row-major) = 18 floats
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

// one thread = 4 consecutive voxels along the last axis = 12 consecutive floats = three 16-byte stores
__global__ void __launch_bounds__(256) k_coord_volume(const float* __restrict__ spec, float* __restrict__ out, int N0,
                                                      int N1, int N2, int transfer) {
  __shared__ float s[kCoordSpec];
  const int b = blockIdx.y;
  if (threadIdx.x < kCoordSpec) s[threadIdx.x] = __ldg(spec + (size_t)b * kCoordSpec + threadIdx.x);
  __syncthreads();
  const long long nvox = (long long)N0 * N1 * N2;
  const long long v0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
  if (v0 >= nvox) return;
  float* ob = out + ((size_t)b * nvox + v0) * 3;
  float f[12];
  int n = 0;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const long long v = v0 + e;
    if (v < nvox) {
      const int i2 = (int)(v % N2), i1 = (int)((v / N2) % N1), i0 = (int)(v / ((long long)N1 * N2));
      float o[3];
      if (transfer) coord_of(s, i0, i2, N1 - 1 - i1, o);
      else coord_of(s, i0, i1, i2, o);
      f[3 * e] = o[0]; f[3 * e + 1] = o[1]; f[3 * e + 2] = o[2];
      n = e + 1;
    }
  }
  if (n == 4 && (nvox & 3) == 0) {
#pragma unroll
    for (int q = 0; q < 3; ++q)
      stg128_stream(ob + 4 * q, make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]));
  } else {
    for (int k = 0; k < 3 * n; ++k) ob[k] = f[k];
  }
}

cudaError_t launch_coord_volumes(const float* spec, float* out, int B, int N0, int N1, int N2, int transfer,
                                 cudaStream_t st) {
  const long long nvox = (long long)N0 * N1 * N2;
  if (B == 0 || nvox == 0) return cudaSuccess;
  dim3 grid((unsigned)((nvox + 1023) / 1024), (unsigned)B);
  k_coord_volume<<<grid, 256, 0, st>>>(spec, out, N0, N1, N2, transfer);
  return cudaGetLastError();
}

}  // namespace lt
