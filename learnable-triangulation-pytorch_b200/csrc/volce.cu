// volce.cu — K5: VolumetricCELoss, forward and backward.
//
// Replaces mvn/models/loss.py:52-80: per batch item, the distance of every voxel centre to every ground-truth joint
// (a (J, N^3) temporary), argmin, a device->host copy of the indices (loss.py:71 synchronises the stream), then a python
// loop of B*J scalar gathers:   loss = sum_{b,j} validity[b,j] * -log(p[b,j,argmin_i |c_i - k_bj|] + 1e-6) / (B*J).
// Here: one sweep over the coordinates (each voxel's xyz is fetched once and kept in registers for all J joints),
// no temporary, no synchronisation.
//   k_ce_argmin : grid (chunk, b). Per (joint, warp) one (distance, voxel) pair, smallest distance first, smallest
//                 index among equal distances (torch.argmin returns the first minimum on CPU).
//   k_ce_pick   : one warp per (b, j): merges the chunks, gathers p at the voxel, writes the term and the index.
//   k_ce_sum    : one CTA: adds the B*J terms in the reference's order (b outer, j inner, fp32) and divides by B*J.
// Distances use the reference's arithmetic, un-contracted: ((dx*dx + dy*dy) + dz*dz), then sqrt (two distinct sums can
// round to the same square root, and the FIRST of them must win).
// Backward: the gradient w.r.t. the volumes has B*J non-zeros: -validity / (p + 1e-6) / (B*J) * grad_loss.
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kCeThreads = 256;
constexpr int kCeEpt = 8;
constexpr int kCeChunk = kCeThreads * kCeEpt;  // 2048 voxels per CTA

struct DistIdx {
  float d;
  int i;
};
__device__ __forceinline__ DistIdx di_min(DistIdx a, DistIdx b) {
  // NaN distances never win (torch would propagate them; coordinates are finite on this path)
  return (b.d < a.d || (b.d == a.d && b.i < a.i)) ? b : a;
}
__device__ __forceinline__ DistIdx di_warp_min(DistIdx v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    DistIdx t;
    t.d = __shfl_xor_sync(0xffffffffu, v.d, o);
    t.i = __shfl_xor_sync(0xffffffffu, v.i, o);
    v = di_min(v, t);
  }
  return v;
}

static inline int ce_chunks(long long nvox) { return (int)((nvox + kCeChunk - 1) / kCeChunk); }

// ws: [(b*J + j)][chunk * kCeWarps + warp] (dist, idx)
// The square root of loss.py:66 is monotone, so the winner is found on the squared distances and only the voxels whose
// square could round to the same root as the warp's smallest one (within 5e-7 relative: two fp32 numbers with the same
// correctly rounded root differ by less than 2.4e-7) get an IEEE square root; among equal roots the lowest index wins,
// exactly what argmin over the reference's sqrt'ed distances returns. Reductions are single REDUX instructions on the
// bit patterns (non-negative floats order like unsigned integers); no shared memory, no barrier.
constexpr int kCeWarps = kCeThreads / 32;
__global__ void __launch_bounds__(kCeThreads) k_ce_argmin(const float* __restrict__ coord, const float* __restrict__ kp,
                                                          DistIdx* __restrict__ ws, int J, long long nvox) {
  const int b = blockIdx.y, chunk = blockIdx.x, nchunk = gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long base = (long long)chunk * kCeChunk;
  const float* cb = coord + (size_t)b * nvox * 3;
  // thread t owns voxels base + t*8 .. +7 (24 consecutive floats)
  float cx[kCeEpt], cy[kCeEpt], cz[kCeEpt];
  const long long v0 = base + (long long)threadIdx.x * kCeEpt;
  if ((nvox & 3) == 0 && (reinterpret_cast<uintptr_t>(coord) & 15u) == 0 && v0 + kCeEpt <= nvox) {  // 96 consecutive, 16-byte aligned bytes
    float f[kCeEpt * 3];
    const float4* p4 = reinterpret_cast<const float4*>(cb + v0 * 3);
#pragma unroll
    for (int q = 0; q < kCeEpt * 3 / 4; ++q) {
      const float4 t = __ldg(p4 + q);
      f[4 * q] = t.x; f[4 * q + 1] = t.y; f[4 * q + 2] = t.z; f[4 * q + 3] = t.w;
    }
#pragma unroll
    for (int e = 0; e < kCeEpt; ++e) { cx[e] = f[3 * e]; cy[e] = f[3 * e + 1]; cz[e] = f[3 * e + 2]; }
  } else {
#pragma unroll
    for (int e = 0; e < kCeEpt; ++e) {
      const bool in = v0 + e < nvox;
      cx[e] = in ? __ldg(cb + (v0 + e) * 3) : 0.f;
      cy[e] = in ? __ldg(cb + (v0 + e) * 3 + 1) : 0.f;
      cz[e] = in ? __ldg(cb + (v0 + e) * 3 + 2) : 0.f;
    }
  }
  const int nin = (int)max(0LL, min((long long)kCeEpt, nvox - v0));  // voxels of this thread inside the volume
  for (int j = 0; j < J; ++j) {
    const float* k = kp + ((size_t)b * J + j) * 3;
    const float kx = __ldg(k), ky = __ldg(k + 1), kz = __ldg(k + 2);
    float d2[kCeEpt];
    float m2 = INFINITY;
#pragma unroll
    for (int e = 0; e < kCeEpt; ++e) {
      const float dx = __fsub_rn(cx[e], kx), dy = __fsub_rn(cy[e], ky), dz = __fsub_rn(cz[e], kz);
      d2[e] = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
      if (e < nin) m2 = fminf(m2, d2[e]);       // NaN distances never win (coordinates are finite on this path)
    }
    const float wm2 = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(m2)));
    const float thr = wm2 * 1.0000005f;
    float bd = INFINITY;
    unsigned bi = 0xffffffffu;
#pragma unroll
    for (int e = 0; e < kCeEpt; ++e) {
      if (e < nin && d2[e] <= thr) {
        const float d = __fsqrt_rn(d2[e]);
        if (d < bd) { bd = d; bi = (unsigned)(v0 + e); }   // e ascending: strict '<' keeps the first
      }
    }
    const float wd = __uint_as_float(__reduce_min_sync(0xffffffffu, __float_as_uint(bd)));
    const unsigned wi = __reduce_min_sync(0xffffffffu, bd == wd ? bi : 0xffffffffu);
    if (lane == 0) {
      DistIdx r;
      r.d = wd;
      r.i = wi == 0xffffffffu ? 0x7fffffff : (int)wi;
      ws[((size_t)b * J + j) * nchunk * kCeWarps + (size_t)chunk * kCeWarps + warp] = r;
    }
  }
}

__global__ void __launch_bounds__(32) k_ce_pick(const DistIdx* __restrict__ ws, const float* __restrict__ prob,
                                                const float* __restrict__ validity, float* __restrict__ terms,
                                                int* __restrict__ idx_out, int nchunk, long long nvox) {
  const int bj = blockIdx.x, lane = threadIdx.x;
  DistIdx best = {INFINITY, 0x7fffffff};
  for (int c = lane; c < nchunk; c += 32) best = di_min(best, ws[(size_t)bj * nchunk + c]);
  best = di_warp_min(best);
  if (lane == 0) {
    const int i = best.i == 0x7fffffff ? 0 : best.i;
    const float p = __ldg(prob + (size_t)bj * nvox + i);
    terms[bj] = __ldg(validity + bj) * -logf(p + 1e-6f);  // loss.py:76
    idx_out[bj] = i;
  }
}

constexpr int kCeSumSmem = 8192;  // terms staged in shared memory (beyond that the tail is read from global, slowly)
__global__ void __launch_bounds__(256) k_ce_sum(const float* __restrict__ terms, float* __restrict__ loss, int BJ) {
  __shared__ float s_terms[kCeSumSmem];
  const int ns = min(BJ, kCeSumSmem);
  for (int i = threadIdx.x; i < ns; i += blockDim.x) s_terms[i] = terms[i];
  __syncthreads();
  if (threadIdx.x == 0) {
    float acc = 0.f;
    for (int i = 0; i < ns; ++i) acc += s_terms[i];  // python's `loss += ...` order (loss.py:74-77)
    for (int i = ns; i < BJ; ++i) acc += terms[i];
    loss[0] = __fdiv_rn(acc, (float)BJ);             // loss.py:80
  }
}

__global__ void k_ce_bwd(const float* __restrict__ prob, const int* __restrict__ idx, const float* __restrict__ validity,
                         const float* __restrict__ gloss, float* __restrict__ gvol, int BJ, long long nvox) {
  const int bj = blockIdx.x * blockDim.x + threadIdx.x;
  if (bj >= BJ) return;
  const size_t at = (size_t)bj * nvox + idx[bj];
  const float p = __ldg(prob + at);
  // d/dp [ v * -log(p + eps) / n ] = -v / (p + eps) / n, written the way autograd chains it: (g/n) * v * (-1/(p+eps))
  gvol[at] = __fdiv_rn(__ldg(gloss), (float)BJ) * __ldg(validity + bj) * -__fdiv_rn(1.f, p + 1e-6f);
}

// The same gradient kept SPARSE for K2's backward (no (B,J,nvox) tensor): vals = the B*J non-zeros, dots = p*vals =
// their contribution to the softmax-backward dot product; after lt_softargmax_bwd ran with these dots,
// k_ce_add_sparse adds the G_i term of the B*J voxels themselves.
__global__ void k_ce_bwd_sparse(const float* __restrict__ prob, const int* __restrict__ idx,
                                const float* __restrict__ validity, const float* __restrict__ gloss,
                                float* __restrict__ vals, float* __restrict__ dots, int BJ, long long nvox) {
  const int bj = blockIdx.x * blockDim.x + threadIdx.x;
  if (bj >= BJ) return;
  const float p = __ldg(prob + (size_t)bj * nvox + idx[bj]);
  const float v = __fdiv_rn(__ldg(gloss), (float)BJ) * __ldg(validity + bj) * -__fdiv_rn(1.f, p + 1e-6f);
  vals[bj] = v;
  dots[bj] = p * v;
}
__global__ void k_ce_add_sparse(const float* __restrict__ prob, const int* __restrict__ idx,
                                const float* __restrict__ vals, float* __restrict__ gvol, int BJ, long long nvox,
                                int softmax, float mult) {
  const int bj = blockIdx.x * blockDim.x + threadIdx.x;
  if (bj >= BJ) return;
  const size_t at = (size_t)bj * nvox + idx[bj];
  const float p = __ldg(prob + at);
  gvol[at] += softmax ? mult * (p * vals[bj]) : (p > 0.f ? mult * vals[bj] : 0.f);
}

cudaError_t launch_volce_bwd_sparse(const float* prob, const int* idx, const float* validity, const float* gloss,
                                    float* vals, float* dots, int B, int J, long long nvox, cudaStream_t st) {
  const int BJ = B * J;
  k_ce_bwd_sparse<<<(BJ + 127) / 128, 128, 0, st>>>(prob, idx, validity, gloss, vals, dots, BJ, nvox);
  return cudaGetLastError();
}
cudaError_t launch_softargmax_bwd_add_sparse(const float* prob, const int* idx, const float* vals, float* gvol, int B,
                                             int J, long long nvox, int softmax, float mult, cudaStream_t st) {
  const int BJ = B * J;
  k_ce_add_sparse<<<(BJ + 127) / 128, 128, 0, st>>>(prob, idx, vals, gvol, BJ, nvox, softmax, mult);
  return cudaGetLastError();
}

size_t volce_ws_bytes(int B, int J, long long nvox) {
  if (B <= 0 || J <= 0 || nvox <= 0) return 0;
  return (size_t)B * J * ce_chunks(nvox) * kCeWarps * sizeof(DistIdx) + (size_t)B * J * sizeof(float);
}

cudaError_t launch_volce_fwd(const float* coord, const float* prob, const float* kp, const float* validity, float* loss,
                             int* idx_out, void* ws, int B, int J, long long nvox, cudaStream_t st) {
  const int nchunk = ce_chunks(nvox), BJ = B * J;
  DistIdx* parts = reinterpret_cast<DistIdx*>(ws);
  float* terms = reinterpret_cast<float*>(parts + (size_t)BJ * nchunk * kCeWarps);
  k_ce_argmin<<<dim3((unsigned)nchunk, (unsigned)B), kCeThreads, 0, st>>>(coord, kp, parts, J, nvox);
  k_ce_pick<<<BJ, 32, 0, st>>>(parts, prob, validity, terms, idx_out, nchunk * kCeWarps, nvox);
  k_ce_sum<<<1, 256, 0, st>>>(terms, loss, BJ);
  return cudaGetLastError();
}

cudaError_t launch_volce_bwd(const float* prob, const int* idx, const float* validity, const float* gloss, float* gvol,
                             int B, int J, long long nvox, cudaStream_t st) {
  const int BJ = B * J;
  cudaError_t e = cudaMemsetAsync(gvol, 0, (size_t)BJ * nvox * sizeof(float), st);
  if (e != cudaSuccess) return e;
  k_ce_bwd<<<(BJ + 127) / 128, 128, 0, st>>>(prob, idx, validity, gloss, gvol, BJ, nvox);
  return cudaGetLastError();
}

}  // namespace lt
