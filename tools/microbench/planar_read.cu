// planar_read.cu — how fast can B200 stream a channel-planar (BV, 256, 96, 96) fp32 tensor in the order a pixel-tiled
// 1x1 convolution (K7) needs it? A CTA owns a run of RUN pixels of one map and walks all 256 channel planes (stride
// 36,864 B) before moving to its next run; 148*k persistent CTAs. RUN = 128 px (512 B per plane visit) is K7's tcgen05
// tile, 256 px the FFMA tile. Compared with a linear read of the same bytes. Loads only (checksum), LDG.128 per lane.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o planar_read planar_read.cu && ./planar_read
#include <cstdio>
#include <cuda_runtime.h>

constexpr int BV = 64, C = 256, HW = 96 * 96;

// order 0: for each plane k: the CTA's RUN pixels (what K7 does). order 1: linear over the whole tensor.
template <int RUN, int DEPTH>
__global__ void __launch_bounds__(256) k_tiled(const float4* __restrict__ x, float* out, int ntiles) {
  constexpr int V4 = RUN / 4;                 // float4 per plane visit
  constexpr int KPI = 256 / V4;               // planes covered by one CTA-wide load instruction (>= 1)
  static_assert(V4 <= 256, "");
  const int t = threadIdx.x, col = t % V4, krow = t / V4;
  float s = 0.f;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int tpi = HW / RUN, bv = tile / tpi, p4 = (tile % tpi) * V4 + col;
    const float4* src = x + ((size_t)bv * C + krow) * (HW / 4) + p4;
    for (int k = 0; k < C; k += KPI * DEPTH) {
      float4 v[DEPTH];
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) v[d] = __ldg(src + (size_t)(k + d * KPI) * (HW / 4));
#pragma unroll
      for (int d = 0; d < DEPTH; ++d) s += v[d].x + v[d].y + v[d].z + v[d].w;
    }
  }
  if (s == 1.2345f) out[0] = s;
}

template <int DEPTH>
__global__ void __launch_bounds__(256) k_linear(const float4* __restrict__ x, float* out, size_t n4) {
  float s = 0.f;
  for (size_t i = (size_t)blockIdx.x * 256 * DEPTH + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256 * DEPTH) {
    float4 v[DEPTH];
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) v[d] = (i + d * 256 < n4) ? __ldg(x + i + d * 256) : make_float4(0, 0, 0, 0);
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) s += v[d].x + v[d].y + v[d].z + v[d].w;
  }
  if (s == 1.2345f) out[0] = s;
}

template <typename F>
float timeit(F f, int n = 20) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 3; ++i) f();
  cudaEventRecord(a);
  for (int i = 0; i < n; ++i) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / n * 1e3f;
}

int main() {
  const size_t n = (size_t)BV * C * HW;
  float *x, *out;
  cudaMalloc(&x, n * 4); cudaMalloc(&out, 4);
  cudaMemset(x, 0, n * 4);
  const double gb = n * 4 / 1e9;
  const float4* x4 = reinterpret_cast<const float4*>(x);
#define RUNCASE(RUN, DEPTH, CTAS)                                                                           \
  {                                                                                                         \
    const int ntiles = BV * (HW / RUN);                                                                     \
    float us = timeit([&] { k_tiled<RUN, DEPTH><<<148 * CTAS, 256>>>(x4, out, ntiles); });                  \
    printf("tiled  run=%4d px (%5d B per plane visit) depth=%d ctas/SM=%d : %7.1f us  %6.0f GB/s\n", RUN, RUN * 4, \
           DEPTH, CTAS, us, gb / us * 1e6);                                                                 \
  }
  RUNCASE(128, 8, 2) RUNCASE(128, 8, 4) RUNCASE(128, 16, 4)
  RUNCASE(256, 8, 2) RUNCASE(256, 8, 4) RUNCASE(256, 16, 4)
  RUNCASE(512, 8, 4) RUNCASE(1024, 8, 4) RUNCASE(1024, 16, 4)
  {
    float us = timeit([&] { k_linear<8><<<148 * 8, 256>>>(x4, out, n / 4); });
    printf("linear depth=8 : %7.1f us  %6.0f GB/s\n", us, gb / us * 1e6);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
