// common.cuh — shared device helpers for the lt_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "lt_b200 kernels are written for sm_100a (B200) only"
#endif

// The opt-in above 48 KB of dynamic shared memory is a per-(kernel, device) attribute: smem_opt_in() (layout.cu) issues
// the driver call the first time a launcher needs `bytes` for `func` on the current device (or needs more than it set
// before) instead of on every launch, and hands back its error code.
namespace lt {
cudaError_t smem_opt_in(const void* func, size_t bytes);
}
#define LT_SMEM_OPT_IN_ONCE(KERNEL, BYTES)                                                   \
  do {                                                                                       \
    cudaError_t lt_e_ = ::lt::smem_opt_in(reinterpret_cast<const void*>(&KERNEL), (BYTES));  \
    if (lt_e_ != cudaSuccess) return lt_e_;                                                  \
  } while (0)

namespace lt {

constexpr int kWarp = 32;

// ---- 256-bit global accesses (LDG.E.256 / STG.E.256, new on sm_100) and vector reductions --------
struct alignas(32) f8 { float v[8]; };

__device__ __forceinline__ f8 ldg256_nc(const float* p) {
  f8 r;
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]),
                 "=f"(r.v[6]), "=f"(r.v[7])
               : "l"(p));
  return r;
}
// streaming variant: data read exactly once (upstream gradients), keep it out of L1
__device__ __forceinline__ f8 ldg256_stream(const float* p) {
  f8 r;
  asm volatile("ld.global.nc.L1::no_allocate.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]),
                 "=f"(r.v[6]), "=f"(r.v[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg256(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]),
               "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
               : "memory");
}
__device__ __forceinline__ float4 ldg128_nc(const float* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}
__device__ __forceinline__ float4 ldg128_stream(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg128_stream(float* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
// REDG.E.ADD.F32x4: one 16-byte fire-and-forget reduction at L2 (sm_90+)
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// ---- packed fp32 (FFMA2 / FMUL2 / FADD2, new on sm_100): two IEEE fp32 results per issued instruction. Each half is
// the same correctly-rounded operation as the scalar form, so results are bit-identical; only issue slots are saved.
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 pack2(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f2 a, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}

// ---- fast transcendental helpers (MUFU): 1-2 ulp, used where the result feeds a normalised average ---------
__device__ __forceinline__ float fast_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr float kLog2eF = 1.4426950408889634f;

// ---- warp / block reductions ----------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// warp-wide maximum in ONE instruction: floats are mapped to integers with the same order (flip the magnitude bits of
// negatives) and reduced with REDUX.MAX.S32. -inf stays the smallest value; NaNs are not expected on these paths.
__device__ __forceinline__ float warp_max_redux(float v) {
  int i = __float_as_int(v);
  i ^= (i >> 31) & 0x7fffffff;
  i = __reduce_max_sync(0xffffffffu, i);
  i ^= (i >> 31) & 0x7fffffff;
  return __int_as_float(i);
}
// Four sums across the warp with 6 shuffles instead of 20: the first two butterfly steps also halve the number of values
// a lane carries (lanes with bit 4 clear keep a,b / set keep c,d; then bit 3 picks one of the pair). Every lane of the
// lane octet k = lane >> 3 returns the warp-wide total of value k (a, b, c, d for k = 0..3). Fixed order: deterministic.
__device__ __forceinline__ float warp_sum4(float a, float b, float c, float d) {
  const int lane = threadIdx.x & 31;
  const bool hi16 = lane & 16, hi8 = lane & 8;
  float k0 = hi16 ? c : a, k1 = hi16 ? d : b;
  const float s0 = hi16 ? a : c, s1 = hi16 ? b : d;
  k0 += __shfl_xor_sync(0xffffffffu, s0, 16);
  k1 += __shfl_xor_sync(0xffffffffu, s1, 16);
  float k = hi8 ? k1 : k0;
  const float s = hi8 ? k0 : k1;
  k += __shfl_xor_sync(0xffffffffu, s, 8);
  k += __shfl_xor_sync(0xffffffffu, k, 4);
  k += __shfl_xor_sync(0xffffffffu, k, 2);
  k += __shfl_xor_sync(0xffffffffu, k, 1);
  return k;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// online-softmax tuple (m, s, wx, wy, wz): s = sum e^(x-m), w = sum e^(x-m) * xyz
struct Lse5 {
  float m, s, x, y, z;
};
__device__ __forceinline__ Lse5 lse_identity() { return {-INFINITY, 0.f, 0.f, 0.f, 0.f}; }
__device__ __forceinline__ Lse5 lse_merge(const Lse5& a, const Lse5& b) {
  float m = fmaxf(a.m, b.m);
  if (m == -INFINITY) return lse_identity();
  float fa = expf(a.m - m), fb = expf(b.m - m);
  return {m, a.s * fa + b.s * fb, a.x * fa + b.x * fb, a.y * fa + b.y * fb, a.z * fa + b.z * fb};
}
__device__ __forceinline__ Lse5 lse_shfl_xor(const Lse5& a, int o) {
  return {__shfl_xor_sync(0xffffffffu, a.m, o), __shfl_xor_sync(0xffffffffu, a.s, o),
          __shfl_xor_sync(0xffffffffu, a.x, o), __shfl_xor_sync(0xffffffffu, a.y, o),
          __shfl_xor_sync(0xffffffffu, a.z, o)};
}
__device__ __forceinline__ Lse5 lse_warp_reduce(Lse5 a) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a = lse_merge(a, lse_shfl_xor(a, o));
  return a;
}

}  // namespace lt
