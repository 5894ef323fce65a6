// red_paths.cu — which path retires a bilinear scatter (4 taps x 128 B per voxel-view) fastest on B200?
//   A: red.global.add.v4.f32 per lane (8 lanes = one 128-byte texel), what K1 backward does today
//   B: stage the 2x2 patch in shared memory, one lane issues two 256-byte cp.reduce.async.bulk (TMA reduce) per voxel
// Address pattern mimics K1: 64 maps of 98x98x32 floats, voxel (i0,i1,z) -> texel (1.45*i1+2, 1.45*z+2): 64 voxels along
// i0 collapse on one pixel like the depth direction does.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o red_paths red_paths.cu && ./red_paths
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int Wp = 98, Hp = 98, Cp = 32, MAPS = 64;
constexpr int NV = 64 * 64 * 64;  // voxels per map

__device__ __forceinline__ void tap_of(long long i, int& map, int& off, float4& w) {
  map = (int)(i / NV);
  const int v = (int)(i % NV);
  const int z = v & 63, i1 = (v >> 6) & 63, i0 = v >> 12;
  const float fx = 1.45f * z + 2.f + 0.01f * i0, fy = 1.45f * i1 + 2.f + 0.013f * i0;
  const int x0 = (int)fx, y0 = (int)fy;
  const float ax = fx - x0, ay = fy - y0;
  w = make_float4((1 - ax) * (1 - ay), ax * (1 - ay), (1 - ax) * ay, ax * ay);
  off = (y0 * Wp + x0) * Cp;
}

__global__ void __launch_bounds__(256) k_red_v4(float* g, long long n) {
  const int lane = threadIdx.x & 31, grp = lane >> 3, q = lane & 7;
  const long long warp = (blockIdx.x * 256LL + threadIdx.x) >> 5, nwarps = gridDim.x * 8LL;
  for (long long i = warp * 4 + grp; i < n; i += nwarps * 4) {
    int map, off; float4 w;
    tap_of(i, map, off, w);
    float* p = g + (size_t)map * Hp * Wp * Cp + off + q * 4;
    const float wt[4] = {w.x, w.y, w.z, w.w};
    const int d[4] = {0, Cp, Wp * Cp, Wp * Cp + Cp};
#pragma unroll
    for (int t = 0; t < 4; ++t)
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p + d[t]), "f"(wt[t]), "f"(wt[t] * 2), "f"(wt[t] * 3), "f"(wt[t] * 4) : "memory");
  }
}

// A': the same scatter as plain (non-atomic, racy) 16-byte stores: what the L2 write path takes without the reduction,
//     and A'': as four scalar red.global.add.f32 per lane and tap instead of one v4.
template <int MODE>
__global__ void __launch_bounds__(256) k_scatter_other(float* g, long long n) {
  const int lane = threadIdx.x & 31, grp = lane >> 3, q = lane & 7;
  const long long warp = (blockIdx.x * 256LL + threadIdx.x) >> 5, nwarps = gridDim.x * 8LL;
  for (long long i = warp * 4 + grp; i < n; i += nwarps * 4) {
    int map, off; float4 w;
    tap_of(i, map, off, w);
    float* p = g + (size_t)map * Hp * Wp * Cp + off + q * 4;
    const float wt[4] = {w.x, w.y, w.z, w.w};
    const int d[4] = {0, Cp, Wp * Cp, Wp * Cp + Cp};
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      if (MODE == 0) {
        asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p + d[t]), "f"(wt[t]), "f"(wt[t] * 2), "f"(wt[t] * 3), "f"(wt[t] * 4) : "memory");
      } else {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + d[t] + c), "f"(wt[t] * (c + 1)) : "memory");
      }
    }
  }
}

__global__ void __launch_bounds__(256) k_red_bulk(float* g, long long n) {
  // per warp: 2 buffers x 4 voxels x 2 rows x 64 floats
  __shared__ __align__(128) float s_buf[8][2][4][2][64];
  const int lane = threadIdx.x & 31, grp = lane >> 3, q = lane & 7, wib = threadIdx.x >> 5;
  const long long warp = (blockIdx.x * 256LL + threadIdx.x) >> 5, nwarps = gridDim.x * 8LL;
  int buf = 0;
  for (long long i0 = warp * 4; i0 < n; i0 += nwarps * 4, buf ^= 1) {
    const long long i = i0 + grp;
    int map = 0, off = 0; float4 w = make_float4(0, 0, 0, 0);
    if (i < n) tap_of(i, map, off, w);
    // the bulk reduce issued two iterations ago from this buffer must have finished READING it
    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
    __syncwarp();
    float* s = &s_buf[wib][buf][grp][0][0];
    *reinterpret_cast<float4*>(s + q * 4) = make_float4(w.x, w.x * 2, w.x * 3, w.x * 4);
    *reinterpret_cast<float4*>(s + 32 + q * 4) = make_float4(w.y, w.y * 2, w.y * 3, w.y * 4);
    *reinterpret_cast<float4*>(s + 64 + q * 4) = make_float4(w.z, w.z * 2, w.z * 3, w.z * 4);
    *reinterpret_cast<float4*>(s + 96 + q * 4) = make_float4(w.w, w.w * 2, w.w * 3, w.w * 4);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncwarp();
    if (q == 0 && i < n) {
      float* p = g + (size_t)map * Hp * Wp * Cp + off;
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(s);
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 256;" ::"l"(p), "r"(sa) : "memory");
      asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], 256;" ::"l"(p + Wp * Cp), "r"(sa + 256) : "memory");
    }
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
  }
  asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// C: the forward direction: 4 x LDG.128 per lane (one 128-byte texel per 8 lanes and tap), blended, one 16-byte store
//    per lane and voxel-view into a channels-last output (so the store side is cheap and coalesced).
__global__ void __launch_bounds__(256) k_gather_v4(const float* __restrict__ g, float* __restrict__ out, long long n) {
  const int lane = threadIdx.x & 31, grp = lane >> 3, q = lane & 7;
  const long long warp = (blockIdx.x * 256LL + threadIdx.x) >> 5, nwarps = gridDim.x * 8LL;
  for (long long i = warp * 4 + grp; i < n; i += nwarps * 4) {
    int map, off; float4 w;
    tap_of(i, map, off, w);
    const float* p = g + (size_t)map * Hp * Wp * Cp + off + q * 4;
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p + Cp));
    const float4 c = __ldg(reinterpret_cast<const float4*>(p + Wp * Cp)), d = __ldg(reinterpret_cast<const float4*>(p + Wp * Cp + Cp));
    float4 r;
    r.x = a.x * w.x + b.x * w.y + c.x * w.z + d.x * w.w; r.y = a.y * w.x + b.y * w.y + c.y * w.z + d.y * w.w;
    r.z = a.z * w.x + b.z * w.y + c.z * w.z + d.z * w.w; r.w = a.w * w.x + b.w * w.y + c.w * w.z + d.w * w.w;
    *reinterpret_cast<float4*>(out + (i % (4LL * NV)) * Cp + q * 4) = r;
  }
}

int main() {
  const size_t elems = (size_t)MAPS * Hp * Wp * Cp;
  float* g;
  cudaMalloc(&g, elems * 4);
  const long long n = (long long)MAPS * NV * 63 / 100;  // ~ the valid voxel-views of one cfg3 step
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int which = 0; which < 2; ++which) {
    for (int grid : {148 * 4, 148 * 8, 148 * 16}) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaMemset(g, 0, elems * 4);
        cudaEventRecord(a);
        if (which == 0) k_red_v4<<<grid, 256>>>(g, n); else k_red_bulk<<<grid, 256>>>(g, n);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      // checksum: total mass added = sum of weights * (1+2+3+4) * 8 lanes' worth
      float h[8];
      cudaMemcpy(h, g + (size_t)(5 * Wp + 7) * Cp, sizeof(h), cudaMemcpyDeviceToHost);
      printf("%s grid %5d: %.3f ms  %.1f GB/s payload  (err %s) probe %.4f %.4f %.4f %.4f\n", which ? "bulk-reduce" : "red.v4     ",
             grid, best, n * 512.0 / best * 1e-6, cudaGetErrorString(cudaGetLastError()), h[0], h[1], h[4], h[5]);
    }
  }
  for (int mode = 0; mode < 2; ++mode) {
    for (int grid : {148 * 8, 148 * 16}) {
      float best = 1e9f;
      for (int rep = 0; rep < 5; ++rep) {
        cudaMemset(g, 0, elems * 4);
        cudaEventRecord(a);
        if (mode == 0) k_scatter_other<0><<<grid, 256>>>(g, n); else k_scatter_other<1><<<grid, 256>>>(g, n);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (ms < best) best = ms;
      }
      printf("%s grid %5d: %.3f ms  %.1f GB/s payload  (err %s)\n", mode ? "red.f32 x4 " : "st.v4 (racy)", grid, best,
             n * 512.0 / best * 1e-6, cudaGetErrorString(cudaGetLastError()));
    }
  }
  float* out;
  cudaMalloc(&out, 4ull * NV * Cp * 4);
  for (int grid : {148 * 4, 148 * 8, 148 * 16, 148 * 32}) {
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
      cudaEventRecord(a);
      k_gather_v4<<<grid, 256>>>(g, out, n);
      cudaEventRecord(b);
      cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      if (ms < best) best = ms;
    }
    printf("gather.v4   grid %5d: %.3f ms  %.1f GB/s gathered  (err %s)\n", grid, best, n * 512.0 / best * 1e-6,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
