// softargmax.cu — K2: volumetric soft-argmax in world coordinates, forward and backward.
//
// Replaces mvn/utils/op.py:84-96: softmax over each joint's flattened volume (or relu, un-normalised),
// then einsum("bnxyz,bxyzc->bnc") with the coordinate volume. Both outputs are produced: coordinates
// (B,J,3) and the normalised volume (B,J,nvox) that VolumetricCELoss indexes (mvn/models/loss.py:76).
//
// Structure (also the multi-GPU z-slab structure): the op is split where the one exchange
// of the sharded variant sits.
//   k_sam_stats   : grid (chunk, b). A CTA owns a chunk of voxels, keeps their xyz in registers and walks
//                   all J joints over them (coordinates are fetched once per voxel, not once per joint).
//                   Per (joint, warp) it emits one online-softmax tuple (m, S, Sx, Sy, Sz).
//   k_sam_combine : one warp per (b, joint) merges the tuples (fixed order => deterministic).
//   k_sam_final   : prob = e^(x-m)/S, streaming; coords = (Sx,Sy,Sz)/S.
// exp is evaluated as ex2(fma(x, mult*log2e, -m*log2e)): one FFMA + one MUFU per element; the shift is
// common to numerator and denominator so only the rounding of log2e (3e-8 * |x-m|) enters the result.
#include <stdlib.h>

#include "common.cuh"
#include "coordspec.cuh"
#include "kernels.h"

namespace lt {

constexpr int kSamThreads = 256;
constexpr int kSamEpt = 8;                           // voxels per thread: 2 groups of 4 (float4)
constexpr int kSamChunk = kSamThreads * kSamEpt;     // 2048 voxels per CTA
constexpr int kSamWarps = kSamThreads / 32;
constexpr float kLog2e = 1.4426950408889634f;

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

static inline int sam_chunks(long long nvox) { return (int)((nvox + kSamChunk - 1) / kSamChunk); }

// xyz of this thread's 16 voxels: group i covers voxels base + (i*256 + tid)*4 .. +3. Either read from the coordinate
// volume (128-bit loads when VEC) or evaluated from the item's 18-float spec (coordspec.cuh): same values bit for bit.
template <bool VEC, bool SPEC>
__device__ __forceinline__ void load_xyz16(const CoordSrc& cs, const float* s_spec, int b, long long base, long long nvox,
                                           float (&cx)[kSamEpt], float (&cy)[kSamEpt], float (&cz)[kSamEpt]) {
  if (SPEC) {
#pragma unroll
    for (int i = 0; i < kSamEpt / 4; ++i) {
      const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
      unsigned t = (unsigned)v0;               // nvox < 2^31 in spec mode (checked by the ABI)
      int i2 = (int)(t % (unsigned)cs.N2);
      t /= (unsigned)cs.N2;
      int i1 = (int)(t % (unsigned)cs.N1), i0 = (int)(t / (unsigned)cs.N1);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float o[3] = {0.f, 0.f, 0.f};
        if (v0 + e < nvox) coord_of_voxel(s_spec, cs.transfer, cs.N1, i0, i1, i2, o);
        cx[i * 4 + e] = o[0]; cy[i * 4 + e] = o[1]; cz[i * 4 + e] = o[2];
        if (++i2 == cs.N2) { i2 = 0; if (++i1 == cs.N1) { i1 = 0; ++i0; } }
      }
    }
    return;
  }
  const float* cb = cs.coord + (size_t)b * nvox * 3;
#pragma unroll
  for (int i = 0; i < kSamEpt / 4; ++i) {
    const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
    if (VEC && v0 < nvox) {  // rows are only 16-byte aligned when nvox % 4 == 0
      const float4* p = reinterpret_cast<const float4*>(cb + v0 * 3);
      float4 a = __ldg(p), bb = __ldg(p + 1), c = __ldg(p + 2);
      cx[i * 4 + 0] = a.x;  cy[i * 4 + 0] = a.y;  cz[i * 4 + 0] = a.z;
      cx[i * 4 + 1] = a.w;  cy[i * 4 + 1] = bb.x; cz[i * 4 + 1] = bb.y;
      cx[i * 4 + 2] = bb.z; cy[i * 4 + 2] = bb.w; cz[i * 4 + 2] = c.x;
      cx[i * 4 + 3] = c.y;  cy[i * 4 + 3] = c.z;  cz[i * 4 + 3] = c.w;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const long long v = v0 + e;
        const bool in = v < nvox;
        cx[i * 4 + e] = in ? __ldg(cb + v * 3) : 0.f;
        cy[i * 4 + e] = in ? __ldg(cb + v * 3 + 1) : 0.f;
        cz[i * 4 + e] = in ? __ldg(cb + v * 3 + 2) : 0.f;
      }
    }
  }
}
// the item's spec -> shared memory (no-op in tensor mode); every thread of the CTA must call it
template <bool SPEC>
__device__ __forceinline__ void stage_spec(const CoordSrc& cs, int b, float* s_spec) {
  if (SPEC) {
    if (threadIdx.x < kCoordSpec) s_spec[threadIdx.x] = __ldg(cs.spec + (size_t)b * kCoordSpec + threadIdx.x);
    __syncthreads();
  }
}


// ---------------------------------------------------------------------------------------------------
// pass 1: partial tuples. ws layout: [(b*J + j)][chunk*kSamWarps + warp][5]
// VEC: nvox % 4 == 0 (rows 16-byte aligned) -> 128-bit loads.
template <bool VEC, bool SOFTMAX, bool SPEC>
__device__ __forceinline__ void sam_stats_body(const float* __restrict__ vol, const CoordSrc& cs, float* s_spec,
                                               float* __restrict__ ws, int J, long long nvox, float mult, int b,
                                               int chunk, int nchunk) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long base = (long long)chunk * kSamChunk;
  stage_spec<SPEC>(cs, b, s_spec);

  float cx[kSamEpt], cy[kSamEpt], cz[kSamEpt];
  bool ok[kSamEpt / 4];
#pragma unroll
  for (int i = 0; i < kSamEpt / 4; ++i)
    ok[i] = VEC && (base + ((long long)i * kSamThreads + threadIdx.x) * 4 < nvox);
  load_xyz16<VEC, SPEC>(cs, s_spec, b, base, nvox, cx, cy, cz);

  const float L = mult * kLog2e;
  // rows are double-buffered in registers: the loads of joint j+1 are in flight while joint j is reduced
  auto load_row = [&](int j, float (&x)[kSamEpt]) {
    const float* row = vol + ((size_t)b * J + j) * nvox;
#pragma unroll
    for (int i = 0; i < kSamEpt / 4; ++i) {
      const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
      if (VEC && ok[i]) {
        float4 t = ldg128_stream(row + v0);
        x[i * 4] = t.x; x[i * 4 + 1] = t.y; x[i * 4 + 2] = t.z; x[i * 4 + 3] = t.w;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) x[i * 4 + e] = (v0 + e < nvox) ? __ldg(row + v0 + e) : -INFINITY;
      }
    }
  };
  auto reduce_row = [&](int j, const float (&x)[kSamEpt]) {
    float tot, m_out;
    if (SOFTMAX) {
      // shift by the warp-wide max of the scaled logits (mult > 0 is guaranteed by the caller); padding
      // voxels carry -inf and contribute e = 0.
      float ms = -INFINITY;
#pragma unroll
      for (int i = 0; i < kSamEpt; ++i) ms = fmaxf(ms, x[i] * mult);
      const float wm = warp_max_redux(ms);  // warp-wide shift: one rescale-free accumulation per warp
      float s = 0.f, sx = 0.f, sy = 0.f, sz = 0.f;
      if (wm > -INFINITY) {
        const float shift = -wm * kLog2e;
#pragma unroll
        for (int i = 0; i < kSamEpt; ++i) {
          const float e = (x[i] == -INFINITY) ? 0.f : ex2_approx(fmaf(x[i], L, shift));
          s += e;
          sx = fmaf(e, cx[i], sx);
          sy = fmaf(e, cy[i], sy);
          sz = fmaf(e, cz[i], sz);
        }
      }
      tot = warp_sum4(s, sx, sy, sz);   // lane octet k holds the total of (S, Sx, Sy, Sz)[k]
      m_out = wm;
    } else {
      float sx = 0.f, sy = 0.f, sz = 0.f;
#pragma unroll
      for (int i = 0; i < kSamEpt; ++i) {
        const float e = (x[i] == -INFINITY) ? 0.f : fmaxf(x[i] * mult, 0.f);
        sx = fmaf(e, cx[i], sx);
        sy = fmaf(e, cy[i], sy);
        sz = fmaf(e, cz[i], sz);
      }
      tot = warp_sum4(1.f / 32.f, sx, sy, sz);   // slot 1 of the tuple is the constant 1 (32 lanes x 1/32, exact)
      m_out = 0.f;
    }
    float* o = ws + (((size_t)b * J + j) * nchunk * kSamWarps + (size_t)chunk * kSamWarps + warp) * 5;
    if ((lane & 7) == 0) o[1 + (lane >> 3)] = tot;
    if (lane == 0) o[0] = m_out;
  };
  float xa[kSamEpt], xb[kSamEpt];
  if (J > 0) load_row(0, xa);
  for (int j = 0; j < J; j += 2) {
    if (j + 1 < J) load_row(j + 1, xb);
    reduce_row(j, xa);
    if (j + 1 < J) {
      if (j + 2 < J) load_row(j + 2, xa);
      reduce_row(j + 1, xb);
    }
  }
}

template <bool VEC, bool SOFTMAX, bool SPEC>
__global__ void __launch_bounds__(kSamThreads, 4) k_sam_stats(const float* __restrict__ vol, const CoordSrc cs,
                                                           float* __restrict__ ws, int J, long long nvox,
                                                           float mult) {
  __shared__ float s_spec[SPEC ? kCoordSpec : 1];
  sam_stats_body<VEC, SOFTMAX, SPEC>(vol, cs, s_spec, ws, J, nvox, mult, blockIdx.y, blockIdx.x, gridDim.x);
}

// merge P tuples per (b,j): one CTA each (a 256^3 volume has 65536 partial tuples per joint), threads stride over the
// tuples, then a fixed warp butterfly and a fixed-order merge of the warp results: deterministic.
constexpr int kCombThreads = 256;
template <bool SOFTMAX>
__global__ void __launch_bounds__(kCombThreads) k_sam_combine(const float* __restrict__ parts, float* __restrict__ stats,
                                                              int P, int BJ, size_t part_stride /* floats between
                                                              consecutive (b,j) */, size_t tuple_stride /* floats
                                                              between consecutive tuples of one (b,j) */) {
  __shared__ float s_part[kCombThreads / 32][5];
  const int bj = blockIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (bj >= BJ) return;
  const float* p = parts + (size_t)bj * part_stride;
  Lse5 acc = SOFTMAX ? lse_identity() : Lse5{0.f, 1.f, 0.f, 0.f, 0.f};
  if (SOFTMAX) {
    for (int i = threadIdx.x; i < P; i += kCombThreads) {
      const float* t = p + (size_t)i * tuple_stride;
      acc = lse_merge(acc, Lse5{t[0], t[1], t[2], t[3], t[4]});
    }
    acc = lse_warp_reduce(acc);
  } else {
    float x = 0.f, y = 0.f, z = 0.f;
    for (int i = threadIdx.x; i < P; i += kCombThreads) {
      const float* t = p + (size_t)i * tuple_stride;
      x += t[2]; y += t[3]; z += t[4];
    }
    acc = {0.f, 1.f, warp_sum(x), warp_sum(y), warp_sum(z)};
  }
  if (lane == 0) {
    s_part[warp][0] = acc.m; s_part[warp][1] = acc.s; s_part[warp][2] = acc.x; s_part[warp][3] = acc.y; s_part[warp][4] = acc.z;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    Lse5 r = {s_part[0][0], s_part[0][1], s_part[0][2], s_part[0][3], s_part[0][4]};
    for (int w = 1; w < kCombThreads / 32; ++w) {
      const Lse5 o = {s_part[w][0], s_part[w][1], s_part[w][2], s_part[w][3], s_part[w][4]};
      if (SOFTMAX) r = lse_merge(r, o);
      else { r.x += o.x; r.y += o.y; r.z += o.z; }
    }
    float* o = stats + (size_t)bj * 5;
    o[0] = r.m; o[1] = r.s; o[2] = r.x; o[3] = r.y; o[4] = r.z;
  }
}

// pass 2: prob = e^(mult*x - m)/S (softmax) or relu(mult*x); coords = (Sx,Sy,Sz)/S.
// (stats are read with ld.global.cg: in the fused kernel below they were written by another CTA of the same launch)
template <bool VEC, bool SOFTMAX, bool COHERENT>
__device__ __forceinline__ void sam_final_body(const float* __restrict__ vol, const float* stats,
                                               float* __restrict__ coords, float* __restrict__ prob, long long nvox,
                                               float mult, int bj, int chunk) {
  const float* st = stats + (size_t)bj * 5;
  auto ld = [](const float* p) { return COHERENT ? __ldcg(p) : __ldg(p); };  // __ldg: L1-resident broadcast
  const float m = ld(st), S = ld(st + 1);
  const float invS = SOFTMAX ? __fdiv_rn(1.f, S) : 1.f;
  if (chunk == 0 && threadIdx.x < 3 && coords != nullptr)
    coords[(size_t)bj * 3 + threadIdx.x] = SOFTMAX ? __fdiv_rn(ld(st + 2 + threadIdx.x), S) : ld(st + 2 + threadIdx.x);
  if (prob == nullptr) return;
  const float L = mult * kLog2e, shift = -m * kLog2e;
  const float* row = vol + (size_t)bj * nvox;
  float* orow = prob + (size_t)bj * nvox;
  const long long base = (long long)chunk * kSamChunk;
#pragma unroll
  for (int i = 0; i < kSamEpt / 4; ++i) {
    const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
    if (VEC) {
      if (v0 < nvox) {
        float4 t = ldg128_stream(row + v0), o;
        if (SOFTMAX) {
          o.x = ex2_approx(fmaf(t.x, L, shift)) * invS; o.y = ex2_approx(fmaf(t.y, L, shift)) * invS;
          o.z = ex2_approx(fmaf(t.z, L, shift)) * invS; o.w = ex2_approx(fmaf(t.w, L, shift)) * invS;
        } else {
          o.x = fmaxf(t.x * mult, 0.f); o.y = fmaxf(t.y * mult, 0.f);
          o.z = fmaxf(t.z * mult, 0.f); o.w = fmaxf(t.w * mult, 0.f);
        }
        stg128_stream(orow + v0, o);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (v0 + e < nvox) {
          const float t = __ldg(row + v0 + e);
          orow[v0 + e] = SOFTMAX ? ex2_approx(fmaf(t, L, shift)) * invS : fmaxf(t * mult, 0.f);
        }
    }
  }
}

template <bool VEC, bool SOFTMAX>
__global__ void __launch_bounds__(kSamThreads) k_sam_final(const float* __restrict__ vol,
                                                           const float* __restrict__ stats,
                                                           float* __restrict__ coords, float* __restrict__ prob,
                                                           long long nvox, float mult) {
  // reverse traversal: the rows the statistics pass read LAST are the ones still in L2, so they are re-read FIRST
  sam_final_body<VEC, SOFTMAX, false>(vol, stats, coords, prob, nvox, mult, gridDim.y - 1 - blockIdx.y,
                               gridDim.x - 1 - blockIdx.x);
}

// merge of one item's P tuples per joint by ONE CTA: warp <-> joint, lanes stride the tuples, fixed butterfly.
// Kept out of line: it runs once per item and must not cost the streaming loops of the fused kernel registers.
template <bool SOFTMAX>
__device__ __noinline__ void sam_merge_item(const float* ws, float* stats, int b, int J, int P) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = warp; j < J; j += kSamWarps) {   // warp <-> joint; lanes stride the tuples; fixed butterfly
    const float* p = ws + ((size_t)b * J + j) * P * 5;
    Lse5 acc = SOFTMAX ? lse_identity() : Lse5{0.f, 1.f, 0.f, 0.f, 0.f};
    for (int i = lane; i < P; i += 32) {
      const float* t = p + (size_t)i * 5;
      const Lse5 o = {__ldcg(t), __ldcg(t + 1), __ldcg(t + 2), __ldcg(t + 3), __ldcg(t + 4)};
      if (SOFTMAX) acc = lse_merge(acc, o);
      else { acc.x += o.x; acc.y += o.y; acc.z += o.z; }
    }
    if (SOFTMAX) acc = lse_warp_reduce(acc);
    else acc = {0.f, 1.f, warp_sum(acc.x), warp_sum(acc.y), warp_sum(acc.z)};
    if (lane == 0) {
      float* o = stats + ((size_t)b * J + j) * 5;
      o[0] = acc.m; o[1] = acc.s; o[2] = acc.x; o[3] = acc.y; o[4] = acc.z;
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// forward on one device as ONE persistent launch: the statistics of item b+1 run between the statistics and the
// normalisation pass of item b, so the second read of an item's volumes (J*nvox*4 bytes, 17.8 MB at 64^3) is served by
// L2 instead of HBM (the three-launch version re-reads 285 MB at cfg3: 1.39x the algorithmic bytes).
// Work units of 2048 voxels x all joints are handed out IN ORDER by a ticket counter:
//   S(0), S(1), F(0), S(2), F(1), ..., S(B-1), F(B-2), F(B-1)      (lag 1; LT_B200_K2LAG=n runs n items of S ahead)
// S = statistics (k_sam_stats' body); the CTA that finishes an item's last S unit merges the item's tuples in a fixed
// order (deterministic, same tree for every run) and raises the item's flag; F = normalisation (k_sam_final's body for
// every joint of the chunk) after the flag is up. A unit only ever waits for units with smaller tickets, which running
// CTAs already hold: no deadlock however few CTAs are resident.
// ctrl: [0] ticket, [1, 1+B) finished S units per item, [1+B, 1+2B) flags; zeroed by the launcher.
template <bool VEC, bool SOFTMAX, bool SPEC>
__global__ void __launch_bounds__(kSamThreads, 4) k_sam_fwd_fused(const float* __restrict__ vol, const CoordSrc cs,
                                                                  float* ws, float* stats, float* __restrict__ coords,
                                                                  float* __restrict__ prob, int* ctrl, int B, int J,
                                                                  long long nvox, float mult, int nchunk, int lag) {
  __shared__ float s_spec[SPEC ? kCoordSpec : 1];
  __shared__ int s_unit, s_last;
  const int total = 2 * B * nchunk, P = nchunk * kSamWarps;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* ws2 = ws + (size_t)B * J * P * 5;  // one tuple per (b, joint, chunk)
  for (;;) {
    __syncthreads();  // s_unit / s_spec of the previous unit are dead
    if (threadIdx.x == 0) s_unit = atomicAdd(&ctrl[0], 1);
    __syncthreads();
    const int unit = s_unit;
    if (unit >= total) return;
    const int g = unit / nchunk, chunk = unit % nchunk;
    // S(0..L), then F(k), S(L+1+k) alternating, then the last L+1 F groups (L = lag, clamped to B-1)
    bool is_stats;
    int b;
    if (g <= lag) { is_stats = true; b = g; }
    else {
      const int h = g - (lag + 1), mid = 2 * (B - lag - 1);
      if (h < mid) { is_stats = (h & 1) != 0; b = is_stats ? lag + 1 + h / 2 : h / 2; }
      else { is_stats = false; b = (B - lag - 1) + (h - mid); }
    }
    if (is_stats) {
      sam_stats_body<VEC, SOFTMAX, SPEC>(vol, cs, s_spec, ws, J, nvox, mult, b, chunk, nchunk);
      __syncthreads();
      // first level of the merge, by this CTA: its 8 warp tuples per joint -> one tuple per (joint, chunk) in ws2
      // (fixed butterfly over 8 lanes), so that the item's last CTA only merges nchunk tuples per joint
      for (int j = warp; j < J; j += kSamWarps) {
        const float* t = ws + (((size_t)b * J + j) * P + (size_t)chunk * kSamWarps + (lane & 7)) * 5;
        Lse5 acc = {__ldcg(t), __ldcg(t + 1), __ldcg(t + 2), __ldcg(t + 3), __ldcg(t + 4)};
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {
          const Lse5 other = lse_shfl_xor(acc, o);
          if (SOFTMAX) acc = lse_merge(acc, other);
          else { acc.x += other.x; acc.y += other.y; acc.z += other.z; }
        }
        if (lane == 0) {
          float* o2 = ws2 + (((size_t)b * J + j) * nchunk + chunk) * 5;
          o2[0] = acc.m; o2[1] = acc.s; o2[2] = acc.x; o2[3] = acc.y; o2[4] = acc.z;
        }
      }
      __threadfence();
      __syncthreads();
      if (threadIdx.x == 0) s_last = (atomicAdd(&ctrl[1 + b], 1) == nchunk - 1);
      __syncthreads();
      if (s_last) {
        __threadfence();
        sam_merge_item<SOFTMAX>(ws2, stats, b, J, nchunk);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) atomicExch(&ctrl[1 + B + b], 1);
      }
    } else {
      if (threadIdx.x == 0) {
        while (atomicAdd(&ctrl[1 + B + b], 0) == 0) __nanosleep(200);
        __threadfence();
      }
      __syncthreads();
      if (VEC) {
        // the chunk's rows of kFuseJ joints at a time: all their loads are in flight before the first exponential
        // (one joint per round trip left this unit latency-bound: 17 dependent L2 round trips per chunk)
        constexpr int kFuseJ = 4;
        const long long base = (long long)chunk * kSamChunk;
        const float L = mult * kLog2e;
        for (int j0 = 0; j0 < J; j0 += kFuseJ) {
          float4 t[kFuseJ][kSamEpt / 4];
#pragma unroll
          for (int jj = 0; jj < kFuseJ; ++jj) {
            const float* row = vol + ((size_t)b * J + min(j0 + jj, J - 1)) * nvox;
#pragma unroll
            for (int i = 0; i < kSamEpt / 4; ++i) {
              const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
              t[jj][i] = v0 < nvox ? ldg128_stream(row + v0) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
#pragma unroll
          for (int jj = 0; jj < kFuseJ; ++jj) {
            if (j0 + jj < J) {
              const int bj = b * J + j0 + jj;
              const float* st = stats + (size_t)bj * 5;
              const float m = __ldcg(st), S = __ldcg(st + 1);
              const float invS = SOFTMAX ? __fdiv_rn(1.f, S) : 1.f, shift = -m * kLog2e;
              if (chunk == 0 && threadIdx.x < 3)
                coords[(size_t)bj * 3 + threadIdx.x] =
                    SOFTMAX ? __fdiv_rn(__ldcg(st + 2 + threadIdx.x), S) : __ldcg(st + 2 + threadIdx.x);
              float* orow = prob + (size_t)bj * nvox;
#pragma unroll
              for (int i = 0; i < kSamEpt / 4; ++i) {
                const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
                if (v0 < nvox) {
                  const float4 x = t[jj][i];
                  float4 o;
                  if (SOFTMAX) {
                    o.x = ex2_approx(fmaf(x.x, L, shift)) * invS; o.y = ex2_approx(fmaf(x.y, L, shift)) * invS;
                    o.z = ex2_approx(fmaf(x.z, L, shift)) * invS; o.w = ex2_approx(fmaf(x.w, L, shift)) * invS;
                  } else {
                    o.x = fmaxf(x.x * mult, 0.f); o.y = fmaxf(x.y * mult, 0.f);
                    o.z = fmaxf(x.z * mult, 0.f); o.w = fmaxf(x.w * mult, 0.f);
                  }
                  stg128_stream(orow + v0, o);
                }
              }
            }
          }
        }
      } else {
        for (int j = 0; j < J; ++j)
          sam_final_body<VEC, SOFTMAX, true>(vol, stats, coords, prob, nvox, mult, b * J + j, chunk);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------
// backward
//   softmax: g_i = mult * p_i * (gp_i + gc.xyz_i - dot),  dot = sum_k p_k gp_k + gc.coords
//   relu   : g_i = mult * [p_i > 0] * (gp_i + gc.xyz_i)
// grid (chunk, b): coordinates in registers, loop over joints (same reuse as k_sam_stats).
template <bool VEC, bool SOFTMAX, bool HAS_GP, bool SPEC>
__global__ void __launch_bounds__(kSamThreads) k_sam_bwd(const float* __restrict__ prob,
                                                         const CoordSrc cs,
                                                         const float* __restrict__ coords_out,
                                                         const float* __restrict__ gcoords,
                                                         const float* __restrict__ gprob,
                                                         const float* __restrict__ dots, float* __restrict__ gvol,
                                                         int J, long long nvox, float mult) {
  __shared__ float s_spec[SPEC ? kCoordSpec : 1];
  const int b = blockIdx.y, chunk = blockIdx.x;
  const long long base = (long long)chunk * kSamChunk;
  stage_spec<SPEC>(cs, b, s_spec);
  float cx[kSamEpt], cy[kSamEpt], cz[kSamEpt];
  load_xyz16<VEC, SPEC>(cs, s_spec, b, base, nvox, cx, cy, cz);
  // rows (saved probabilities and, when present, the dense upstream gradient) are double-buffered in registers:
  // the loads of joint j+1 are in flight while joint j is computed and stored
  auto load_rows = [&](int j, float (&p)[kSamEpt], float (&gp)[kSamEpt]) {
    const size_t bj = (size_t)b * J + j;
    const float* prow = prob + bj * nvox;
    const float* grow = HAS_GP ? gprob + bj * nvox : nullptr;
#pragma unroll
    for (int i = 0; i < kSamEpt / 4; ++i) {
      const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
#pragma unroll
      for (int e = 0; e < 4; ++e) { p[i * 4 + e] = 0.f; gp[i * 4 + e] = 0.f; }
      if (VEC) {
        if (v0 < nvox) {
          float4 t = ldg128_stream(prow + v0);
          p[i * 4] = t.x; p[i * 4 + 1] = t.y; p[i * 4 + 2] = t.z; p[i * 4 + 3] = t.w;
          if (HAS_GP) {
            float4 u = ldg128_stream(grow + v0);
            gp[i * 4] = u.x; gp[i * 4 + 1] = u.y; gp[i * 4 + 2] = u.z; gp[i * 4 + 3] = u.w;
          }
        }
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (v0 + e < nvox) {
            p[i * 4 + e] = __ldg(prow + v0 + e);
            if (HAS_GP) gp[i * 4 + e] = __ldg(grow + v0 + e);
          }
      }
    }
  };
  auto emit = [&](int j, const float (&p)[kSamEpt], const float (&gp)[kSamEpt]) {
    const size_t bj = (size_t)b * J + j;
    const float gx = __ldg(gcoords + bj * 3), gy = __ldg(gcoords + bj * 3 + 1), gz = __ldg(gcoords + bj * 3 + 2);
    float dot = 0.f;
    if (SOFTMAX) {
      dot = gx * __ldg(coords_out + bj * 3) + gy * __ldg(coords_out + bj * 3 + 1) + gz * __ldg(coords_out + bj * 3 + 2);
      if (dots != nullptr) dot += __ldg(dots + bj);  // dense gp (k_sam_dot) or a sparse one (lt_volce_bwd_sparse)
    }
    float* orow = gvol + bj * nvox;
#pragma unroll
    for (int i = 0; i < kSamEpt / 4; ++i) {
      const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
      float o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float G = gp[i * 4 + e] + (gx * cx[i * 4 + e] + gy * cy[i * 4 + e] + gz * cz[i * 4 + e]);
        if (SOFTMAX) o[e] = mult * (p[i * 4 + e] * (G - dot));
        else o[e] = (p[i * 4 + e] > 0.f) ? mult * G : 0.f;
      }
      if (VEC) {
        if (v0 < nvox) stg128_stream(orow + v0, make_float4(o[0], o[1], o[2], o[3]));
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (v0 + e < nvox) orow[v0 + e] = o[e];
      }
    }
  };
  float pa[kSamEpt], pb[kSamEpt], ga[kSamEpt], gb[kSamEpt];
  if (J > 0) load_rows(0, pa, ga);
  for (int j = 0; j < J; j += 2) {
    if (j + 1 < J) load_rows(j + 1, pb, gb);
    emit(j, pa, ga);
    if (j + 1 < J) {
      if (j + 2 < J) load_rows(j + 2, pa, ga);
      emit(j + 1, pb, gb);
    }
  }
}

// partial sums of p*gp: ws[(bj)][chunk*kSamWarps + warp]
template <bool VEC>
__global__ void __launch_bounds__(kSamThreads) k_sam_dot(const float* __restrict__ prob,
                                                         const float* __restrict__ gprob, float* __restrict__ ws,
                                                         long long nvox) {
  const int bj = blockIdx.y, chunk = blockIdx.x, nchunk = gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float* prow = prob + (size_t)bj * nvox;
  const float* grow = gprob + (size_t)bj * nvox;
  const long long base = (long long)chunk * kSamChunk;
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kSamEpt / 4; ++i) {
    const long long v0 = base + ((long long)i * kSamThreads + threadIdx.x) * 4;
    if (VEC) {
      if (v0 < nvox) {
        float4 a = ldg128_nc(prow + v0), g = ldg128_nc(grow + v0);
        s = fmaf(a.x, g.x, s); s = fmaf(a.y, g.y, s); s = fmaf(a.z, g.z, s); s = fmaf(a.w, g.w, s);
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (v0 + e < nvox) s = fmaf(__ldg(prow + v0 + e), __ldg(grow + v0 + e), s);
    }
  }
  s = warp_sum(s);
  if (lane == 0) ws[(size_t)bj * nchunk * kSamWarps + (size_t)chunk * kSamWarps + warp] = s;
}

__global__ void k_sam_dot_combine(const float* __restrict__ ws, float* __restrict__ dots, int P, int BJ) {
  const int bj = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (bj >= BJ) return;
  float s = 0.f;
  for (int i = lane; i < P; i += 32) s += ws[(size_t)bj * P + i];
  s = warp_sum(s);
  if (lane == 0) dots[bj] = s;
}

// ---------------------------------------------------------------------------------------------------
static inline size_t sam_tuple_bytes(int B, int J, long long nvox) {
  // per-warp tuples + (fused forward) one pre-merged tuple per chunk
  return (((size_t)B * J * sam_chunks(nvox) * (kSamWarps + 1) * 5 * sizeof(float)) + 255) / 256 * 256;
}
size_t softargmax_ws_bytes(int B, int J, long long nvox) {
  return sam_tuple_bytes(B, J, nvox) + (size_t)(1 + 2 * B) * sizeof(int);  // + the fused forward's control words
}

// the 128-bit paths need 16-byte aligned bases on top of nvox % 4 == 0 (include/lt_b200.h "alignment")
static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

cudaError_t launch_softargmax_partial(const float* vol, const float* coord, float* stats, void* ws, int B,
                                      int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                      const CoordSpecArg* spec) {
  const CoordSrc cs{coord, spec ? spec->spec : nullptr, spec ? spec->N1 : 1, spec ? spec->N2 : 1,
                    spec ? spec->transfer : 0};
  if (B * J == 0) return cudaSuccess;
  if (nvox == 0) {  // empty volume: identity tuples
    return launch_softargmax_merge(nullptr, stats, 0, B * J, softmax, st);
  }
  const int nchunk = sam_chunks(nvox);
  const bool vec = (nvox % 4) == 0 && al16(vol) && (cs.spec != nullptr || al16(coord));
  dim3 grid(nchunk, B);
  float* w = static_cast<float*>(ws);
#define LT_GO(V_, S_)                                                                                         \
  do {                                                                                                        \
    if (cs.spec != nullptr) k_sam_stats<V_, S_, true><<<grid, kSamThreads, 0, st>>>(vol, cs, w, J, nvox, mult); \
    else k_sam_stats<V_, S_, false><<<grid, kSamThreads, 0, st>>>(vol, cs, w, J, nvox, mult);                 \
  } while (0)
  if (vec) { if (softmax) LT_GO(true, true); else LT_GO(true, false); }
  else     { if (softmax) LT_GO(false, true); else LT_GO(false, false); }
#undef LT_GO
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int P = nchunk * kSamWarps, BJ = B * J;
  if (softmax) k_sam_combine<true><<<BJ, kCombThreads, 0, st>>>(w, stats, P, BJ, (size_t)P * 5, 5);
  else k_sam_combine<false><<<BJ, kCombThreads, 0, st>>>(w, stats, P, BJ, (size_t)P * 5, 5);
  return cudaGetLastError();
}

// forward on one device: three launches, or (LT_B200_K2FWD=fused) the persistent fused kernel above
cudaError_t launch_softargmax_fwd(const float* vol, const float* coord, float* coords, float* prob, float* stats,
                                  void* ws, int B, int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                  const CoordSpecArg* spec) {
  // MEASURED at cfg3 (gpurun_out r2e): the fused launch runs 0.289 ms against 0.165 ms for the three launches (the merge
  // of an item's tuples by its last CTA sits on the critical path of a two-item-deep pipeline), so it is opt-in:
  // LT_B200_K2FWD=fused.
  // LT_B200_K2FWD = cluster | fused | (unset: the three launches); read per call so that tests can switch
  const char* mode = getenv("LT_B200_K2FWD");
  const bool split = !(mode != nullptr && mode[0] == 'f');
  if (mode != nullptr && mode[0] == 'c' && B * J > 0 && prob != nullptr && coords != nullptr && stats != nullptr &&
      softargmax_cluster_size(nvox) > 0 && al16(vol) && al16(prob) && (spec != nullptr || al16(coord)))
    return launch_softargmax_fwd_cluster(vol, coord, coords, prob, stats, B, J, nvox, softmax, mult, st, spec);
  if (split || B * J == 0 || nvox == 0 || prob == nullptr || coords == nullptr) {
    cudaError_t e = launch_softargmax_partial(vol, coord, stats, ws, B, J, nvox, softmax, mult, st, spec);
    if (e != cudaSuccess) return e;
    return launch_softargmax_finalize(vol, stats, coords, prob, B, J, nvox, softmax, mult, st);
  }
  const CoordSrc cs{coord, spec ? spec->spec : nullptr, spec ? spec->N1 : 1, spec ? spec->N2 : 1,
                    spec ? spec->transfer : 0};
  const int nchunk = sam_chunks(nvox);
  int* ctrl = reinterpret_cast<int*>(static_cast<char*>(ws) + sam_tuple_bytes(B, J, nvox));
  cudaError_t e = cudaMemsetAsync(ctrl, 0, (size_t)(1 + 2 * B) * sizeof(int), st);
  if (e != cudaSuccess) return e;
  const bool vec = (nvox % 4) == 0 && al16(vol) && al16(prob) && (cs.spec != nullptr || al16(coord));
  static const int lag_env = [] { const char* e = getenv("LT_B200_K2LAG"); return e ? atoi(e) : 1; }();
  const int lag = lag_env < 0 ? 0 : (lag_env > B - 1 ? B - 1 : lag_env);  // items of statistics ahead of the normalisation
  const long long total = 2LL * B * nchunk;
  const long long resident = 4LL * sm_count();
  const int grid = (int)(total < resident ? total : resident);
  float* w = static_cast<float*>(ws);
#define LT_GO(V_, S_)                                                                                                \
  do {                                                                                                               \
    if (cs.spec != nullptr)                                                                                          \
      k_sam_fwd_fused<V_, S_, true><<<grid, kSamThreads, 0, st>>>(vol, cs, w, stats, coords, prob, ctrl, B, J, nvox, \
                                                                  mult, nchunk, lag);                                \
    else                                                                                                             \
      k_sam_fwd_fused<V_, S_, false><<<grid, kSamThreads, 0, st>>>(vol, cs, w, stats, coords, prob, ctrl, B, J, nvox, \
                                                                   mult, nchunk, lag);                               \
  } while (0)
  if (vec) { if (softmax) LT_GO(true, true); else LT_GO(true, false); }
  else     { if (softmax) LT_GO(false, true); else LT_GO(false, false); }
#undef LT_GO
  return cudaGetLastError();
}

cudaError_t launch_softargmax_merge(const float* parts, float* merged, int R, int BJ, int softmax,
                                    cudaStream_t st) {
  if (BJ == 0) return cudaSuccess;
  // parts (R, BJ, 5): tuples of one (b,j) are BJ*5 floats apart
  if (softmax) k_sam_combine<true><<<BJ, kCombThreads, 0, st>>>(parts, merged, R, BJ, 5, (size_t)BJ * 5);
  else k_sam_combine<false><<<BJ, kCombThreads, 0, st>>>(parts, merged, R, BJ, 5, (size_t)BJ * 5);
  return cudaGetLastError();
}

cudaError_t launch_softargmax_finalize(const float* vol, const float* stats, float* coords, float* prob,
                                       int B, int J, long long nvox, int softmax, float mult,
                                       cudaStream_t st) {
  if (B * J == 0) return cudaSuccess;
  const int nchunk = nvox == 0 ? 1 : sam_chunks(nvox);
  const bool vec = (nvox % 4) == 0 && al16(vol) && al16(prob);
  dim3 grid(nchunk, B * J);
#define LT_GO(V_, S_) k_sam_final<V_, S_><<<grid, kSamThreads, 0, st>>>(vol, stats, coords, prob, nvox, mult)
  if (vec) { if (softmax) LT_GO(true, true); else LT_GO(true, false); }
  else     { if (softmax) LT_GO(false, true); else LT_GO(false, false); }
#undef LT_GO
  return cudaGetLastError();
}

cudaError_t launch_softargmax_bwd_dot(const float* prob, const float* grad_prob, float* dots, void* ws,
                                      int B, int J, long long nvox, cudaStream_t st) {
  const int BJ = B * J;
  if (BJ == 0) return cudaSuccess;
  if (nvox == 0) return cudaMemsetAsync(dots, 0, (size_t)BJ * sizeof(float), st);
  const int nchunk = sam_chunks(nvox);
  dim3 grid(nchunk, BJ);
  float* w = static_cast<float*>(ws);
  if (nvox % 4 == 0 && al16(prob) && al16(grad_prob)) k_sam_dot<true><<<grid, kSamThreads, 0, st>>>(prob, grad_prob, w, nvox);
  else k_sam_dot<false><<<grid, kSamThreads, 0, st>>>(prob, grad_prob, w, nvox);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  const int wpb = 4;
  k_sam_dot_combine<<<(BJ + wpb - 1) / wpb, wpb * 32, 0, st>>>(w, dots, nchunk * kSamWarps, BJ);
  return cudaGetLastError();
}

cudaError_t launch_softargmax_bwd(const float* prob, const float* coord, const float* coords,
                                  const float* grad_coords, const float* grad_prob, const float* dots,
                                  float* grad_vol, int B, int J, long long nvox, int softmax, float mult,
                                  cudaStream_t st, const CoordSpecArg* spec) {
  if (B * J == 0 || nvox == 0) return cudaSuccess;
  const CoordSrc cs{coord, spec ? spec->spec : nullptr, spec ? spec->N1 : 1, spec ? spec->N2 : 1,
                    spec ? spec->transfer : 0};
  dim3 grid(sam_chunks(nvox), B);
  const bool gp = grad_prob != nullptr;
  const bool vec = (nvox % 4) == 0 && al16(prob) && (cs.spec != nullptr || al16(coord)) && al16(grad_vol) &&
                   (!gp || al16(grad_prob));
#define LT_GO(V_, S_, G_)                                                                                             \
  do {                                                                                                                \
    if (cs.spec != nullptr)                                                                                           \
      k_sam_bwd<V_, S_, G_, true><<<grid, kSamThreads, 0, st>>>(prob, cs, coords, grad_coords, grad_prob, dots, grad_vol, \
                                                                J, nvox, mult);                                       \
    else                                                                                                              \
      k_sam_bwd<V_, S_, G_, false><<<grid, kSamThreads, 0, st>>>(prob, cs, coords, grad_coords, grad_prob, dots,      \
                                                                 grad_vol, J, nvox, mult);                            \
  } while (0)
  if (vec) {
    if (softmax) { if (gp) LT_GO(true, true, true); else LT_GO(true, true, false); }
    else         { if (gp) LT_GO(true, false, true); else LT_GO(true, false, false); }
  } else {
    if (softmax) { if (gp) LT_GO(false, true, true); else LT_GO(false, true, false); }
    else         { if (gp) LT_GO(false, false, true); else LT_GO(false, false, false); }
  }
#undef LT_GO
  return cudaGetLastError();
}

}  // namespace lt
