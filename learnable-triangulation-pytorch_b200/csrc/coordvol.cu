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
#include "coordspec.cuh"
#include "kernels.h"

namespace lt {

// one thread = 4 consecutive voxels along the last axis = 12 consecutive floats = three 16-byte stores
__global__ void __launch_bounds__(256) k_coord_volume(const float* __restrict__ spec, float* __restrict__ out, int N0,
                                                      int N1, int N2, int transfer) {
  __shared__ float s[kCoordSpec];
  const int b = blockIdx.y;
  if (threadIdx.x < kCoordSpec) s[threadIdx.x] = __ldg(spec + (size_t)b * kCoordSpec + threadIdx.x);
  __syncthreads();
  const long long nvox = (long long)N0 * N1 * N2;
  const long long v0 = ((long long)blockIdx.x * 256 + threadIdx.x) * 4;
  float* ob = out + ((size_t)b * nvox + v0) * 3;
  // (i0, i1, i2) of the first voxel by 32-bit division (nvox < 2^31 is checked by the ABI), then carried increments:
  // 64-bit divisions per voxel made this kernel compute bound (28.7 us for 50 MB)
  unsigned t = (unsigned)v0;
  int i2 = (int)(t % (unsigned)N2);
  t /= (unsigned)N2;
  int i1 = (int)(t % (unsigned)N1), i0 = (int)(t / (unsigned)N1);
  // vector path: the CTA's 256 x 48 bytes are contiguous; stage them in shared memory so that every store instruction
  // writes 512 contiguous bytes (per-thread float4 stores at a 48-byte stride fill only a third of each line they touch)
  const bool vec = (nvox & 3) == 0 && ((long long)blockIdx.x * 1024 + 1024 <= nvox);
  __shared__ __align__(16) float s_out[256 * 12];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    if (v0 + e < nvox) {
      float o[3];
      if (transfer) coord_of(s, i0, i2, N1 - 1 - i1, o);
      else coord_of(s, i0, i1, i2, o);
      float* dst = vec ? s_out + threadIdx.x * 12 + 3 * e : ob + 3 * e;
      dst[0] = o[0]; dst[1] = o[1]; dst[2] = o[2];
      if (++i2 == N2) { i2 = 0; if (++i1 == N1) { i1 = 0; ++i0; } }
    }
  }
  if (vec) {   // block-uniform
    __syncthreads();
    float* cta_out = out + ((size_t)b * nvox + (size_t)blockIdx.x * 1024) * 3;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const int i4 = threadIdx.x + 256 * q;
      stg128_stream(cta_out + 4 * i4, *reinterpret_cast<const float4*>(s_out + 4 * i4));
    }
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
