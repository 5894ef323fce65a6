// softargmax_cluster.cu — K2 forward with ONE read of the joint volumes (mvn/utils/op.py:84-96).
//
// The three-launch forward (softargmax.cu) reads every (b, joint) volume twice: once for the statistics, once to
// normalise (3S + K bytes of traffic for 2S + K algorithmic bytes; the second read cannot come from L2: 285 MB at
// BASELINE's cfg3 against 126 MB). Here a thread-block CLUSTER owns one (b, joint) volume and keeps it in the shared
// memory of its CTAs between the two passes:
//   * cluster of CS CTAs (CS = 1..16, the smallest power of two that leaves <= 64 KB per CTA, so that three CTAs share
//     an SM and one CTA's load overlaps another's arithmetic and stores); CTA r loads slice r of the volume with
//     cp.async.bulk onto an mbarrier (TMA engine: no register staging);
//   * pass 1 from shared memory: slice maximum, then (S, Sx, Sy, Sz) = sum e^(x-m) * (1, xyz); voxel centres are
//     evaluated from the item's 18-float spec (coordspec.cuh) or read from the coordinate tensor;
//   * ONE exchange: every CTA publishes its online-softmax tuple in its own shared memory, cluster barrier, every CTA
//     reads all CS tuples through distributed shared memory and merges them in rank order (deterministic; the same
//     lse_merge as the z-slab exchange of slab.py);
//   * pass 2 from shared memory: prob = e^(x-m)/S streamed out; rank 0 writes coordinates and the merged tuple.
// Traffic: S read + S written (+ K in tensor mode, through L2 once per joint). Applies when nvox % (4 CS) == 0 and the
// volume fits 16 x 64 KB (nvox <= 262144 = 64^3); larger volumes keep the two-pass kernels.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "common.cuh"
#include "coordspec.cuh"
#include "kernels.h"

namespace cg = cooperative_groups;

namespace lt {
namespace {

constexpr int kClThreads = 256;
constexpr int kClWarps = kClThreads / 32;
constexpr int kClMaxSlice = 16384;   // floats per CTA (64 KB)
constexpr float kLog2eC = 1.4426950408889634f;

__device__ __forceinline__ unsigned smem_u32c(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float ex2c(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void mbar_wait_c(unsigned mbar, unsigned parity) {
  for (;;) {
    unsigned done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
    if (done) return;
  }
}

struct ClSmem {
  unsigned long long full;
  float tup[5];                 // this CTA's tuple, read by the whole cluster
  float all[16][5];             // every rank's tuple, local copy
  float part[kClWarps][5];
  float spec[kCoordSpec];
  float bcast[2];
};

template <bool SOFTMAX, bool SPEC>
__global__ void __launch_bounds__(kClThreads, 3) k_sam_fwd_cluster(const float* __restrict__ vol, const CoordSrc cs,
                                                                 float* __restrict__ stats, float* __restrict__ coords,
                                                                 float* __restrict__ prob, int J, int nvox, int slice,
                                                                 float mult) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  ClSmem& S = *reinterpret_cast<ClSmem*>(smem_raw);
  float* sx = reinterpret_cast<float*>(smem_raw + ((sizeof(ClSmem) + 127) / 128) * 128);
  cg::cluster_group cluster = cg::this_cluster();
  const int CS = (int)cluster.num_blocks(), r = (int)cluster.block_rank();
  const int bj = blockIdx.y, b = bj / J;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int v_base = r * slice;
  const float* row = vol + (size_t)bj * nvox + v_base;

  // ---- load the slice: one elected thread arms the barrier and issues the bulk copies (16 KB each) ----
  const unsigned full = smem_u32c(&S.full);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(full) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(full), "r"((unsigned)slice * 4u) : "memory");
    for (int o = 0; o < slice; o += 4096) {
      const unsigned n = (unsigned)min(4096, slice - o) * 4u;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       smem_u32c(sx + o)), "l"(row + o), "r"(n), "r"(full) : "memory");
    }
  }
  if (SPEC && threadIdx.x < kCoordSpec) S.spec[threadIdx.x] = __ldg(cs.spec + (size_t)b * kCoordSpec + threadIdx.x);
  __syncthreads();   // the barrier is initialised (and the spec staged) before anybody waits
  mbar_wait_c(full, 0);

  const float L = mult * kLog2eC;
  const int ngroups = slice / 4;
  // ---- pass 1a: maximum of the scaled logits of the slice ----
  float m_loc = 0.f;
  if (SOFTMAX) {
    float ms = -INFINITY;
    for (int gi = threadIdx.x; gi < ngroups; gi += kClThreads) {
      const float4 t = *reinterpret_cast<const float4*>(sx + gi * 4);
      ms = fmaxf(fmaxf(ms, t.x * mult), fmaxf(fmaxf(t.y * mult, t.z * mult), t.w * mult));
    }
    ms = warp_max_redux(ms);
    if (lane == 0) S.part[warp][0] = ms;
    __syncthreads();
    if (threadIdx.x == 0) {
      float m = S.part[0][0];
      for (int w = 1; w < kClWarps; ++w) m = fmaxf(m, S.part[w][0]);
      S.bcast[0] = m;
    }
    __syncthreads();
    m_loc = S.bcast[0];
  }
  // ---- pass 1b: (S, Sx, Sy, Sz) of the slice ----
  {
    float s = 0.f, ax = 0.f, ay = 0.f, az = 0.f;
    const bool live = !SOFTMAX || m_loc > -INFINITY;      // a slice of -inf only contributes nothing
    const float shift = live ? -m_loc * kLog2eC : 0.f;
    const float* cb = SPEC ? nullptr : cs.coord + ((size_t)b * nvox + v_base) * 3;
    for (int gi = threadIdx.x; gi < ngroups; gi += kClThreads) {
      const float4 t = *reinterpret_cast<const float4*>(sx + gi * 4);
      const float xs[4] = {t.x, t.y, t.z, t.w};
      float cx[4], cy[4], cz[4];
#ifdef LT_K2C_NOCOORD   // timing experiment only: what the structure costs without any coordinate work
      if (true) {
#pragma unroll
        for (int e = 0; e < 4; ++e) { cx[e] = 1.f; cy[e] = 2.f; cz[e] = 3.f; }
      } else
#endif
      if (SPEC) {
        unsigned v = (unsigned)(v_base + gi * 4);
        int i2 = (int)(v % (unsigned)cs.N2);
        v /= (unsigned)cs.N2;
        int i1 = (int)(v % (unsigned)cs.N1), i0 = (int)(v / (unsigned)cs.N1);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float o[3];
          coord_of_voxel(S.spec, cs.transfer, cs.N1, i0, i1, i2, o);
          cx[e] = o[0]; cy[e] = o[1]; cz[e] = o[2];
          if (++i2 == cs.N2) { i2 = 0; if (++i1 == cs.N1) { i1 = 0; ++i0; } }
        }
      } else {
        const float4* p = reinterpret_cast<const float4*>(cb + (size_t)gi * 12);
        const float4 a = __ldg(p), bb = __ldg(p + 1), c = __ldg(p + 2);
        cx[0] = a.x;  cy[0] = a.y;  cz[0] = a.z;
        cx[1] = a.w;  cy[1] = bb.x; cz[1] = bb.y;
        cx[2] = bb.z; cy[2] = bb.w; cz[2] = c.x;
        cx[3] = c.y;  cy[3] = c.z;  cz[3] = c.w;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float w = SOFTMAX ? ((live && xs[e] > -INFINITY) ? ex2c(fmaf(xs[e], L, shift)) : 0.f) : fmaxf(xs[e] * mult, 0.f);
        s += w;
        ax = fmaf(w, cx[e], ax);
        ay = fmaf(w, cy[e], ay);
        az = fmaf(w, cz[e], az);
      }
    }
    const float tot = warp_sum4(s, ax, ay, az);   // lane octet k holds the warp total of value k
    if ((lane & 7) == 0) S.part[warp][1 + (lane >> 3)] = tot;
    __syncthreads();
    if (threadIdx.x == 0) {
      float q[4] = {0.f, 0.f, 0.f, 0.f};
      for (int w = 0; w < kClWarps; ++w)
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] += S.part[w][1 + k];
      S.tup[0] = SOFTMAX ? m_loc : 0.f;
      S.tup[1] = SOFTMAX ? q[0] : 1.f;
      S.tup[2] = q[1]; S.tup[3] = q[2]; S.tup[4] = q[3];
    }
  }
  // ---- the one exchange ----
  cluster.sync();
  if (threadIdx.x < CS * 5) {
    const int rr = threadIdx.x / 5, k = threadIdx.x % 5;
    const float* remote = cluster.map_shared_rank(S.tup, rr);
    S.all[rr][k] = remote[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    Lse5 acc = {S.all[0][0], S.all[0][1], S.all[0][2], S.all[0][3], S.all[0][4]};
    for (int rr = 1; rr < CS; ++rr) {
      const Lse5 o = {S.all[rr][0], S.all[rr][1], S.all[rr][2], S.all[rr][3], S.all[rr][4]};
      if (SOFTMAX) acc = lse_merge(acc, o);
      else { acc.x += o.x; acc.y += o.y; acc.z += o.z; }
    }
    S.bcast[0] = acc.m; S.bcast[1] = acc.s;
    if (r == 0) {
      float* o = stats + (size_t)bj * 5;
      o[0] = acc.m; o[1] = acc.s; o[2] = acc.x; o[3] = acc.y; o[4] = acc.z;
      if (coords != nullptr) {
        float* c = coords + (size_t)bj * 3;
        c[0] = SOFTMAX ? __fdiv_rn(acc.x, acc.s) : acc.x;
        c[1] = SOFTMAX ? __fdiv_rn(acc.y, acc.s) : acc.y;
        c[2] = SOFTMAX ? __fdiv_rn(acc.z, acc.s) : acc.z;
      }
    }
  }
  __syncthreads();
  // ---- pass 2: normalise the slice out of shared memory ----
  if (prob != nullptr) {
    const float m = S.bcast[0], Ssum = S.bcast[1];
    const float invS = SOFTMAX ? __fdiv_rn(1.f, Ssum) : 1.f;
    const float shift = -m * kLog2eC;
    float* orow = prob + (size_t)bj * nvox + v_base;
    for (int gi = threadIdx.x; gi < ngroups; gi += kClThreads) {
      const float4 t = *reinterpret_cast<const float4*>(sx + gi * 4);
      float4 o;
      if (SOFTMAX) {
        o.x = ex2c(fmaf(t.x, L, shift)) * invS; o.y = ex2c(fmaf(t.y, L, shift)) * invS;
        o.z = ex2c(fmaf(t.z, L, shift)) * invS; o.w = ex2c(fmaf(t.w, L, shift)) * invS;
      } else {
        o.x = fmaxf(t.x * mult, 0.f); o.y = fmaxf(t.y * mult, 0.f);
        o.z = fmaxf(t.z * mult, 0.f); o.w = fmaxf(t.w * mult, 0.f);
      }
      stg128_stream(orow + gi * 4, o);
    }
  }
  cluster.sync();   // nobody leaves while a neighbour may still be reading its tuple
}

}  // namespace

// cluster size for this volume, or 0 when the kernel does not apply
int softargmax_cluster_size(long long nvox) {
  if (nvox <= 0 || nvox > 16LL * kClMaxSlice) return 0;
  int csz = 1;
  while ((nvox + csz - 1) / csz > kClMaxSlice) csz *= 2;
  if (csz > 16 || nvox % (4LL * csz) != 0) return 0;
  return csz;
}

cudaError_t launch_softargmax_fwd_cluster(const float* vol, const float* coord, float* coords, float* prob, float* stats,
                                          int B, int J, long long nvox, int softmax, float mult, cudaStream_t st,
                                          const CoordSpecArg* spec) {
  const int csz = softargmax_cluster_size(nvox);
  if (csz == 0) return cudaErrorNotSupported;
  const CoordSrc cs{coord, spec ? spec->spec : nullptr, spec ? spec->N1 : 1, spec ? spec->N2 : 1, spec ? spec->transfer : 0};
  const int slice = (int)(nvox / csz);
  const size_t smem = ((sizeof(ClSmem) + 127) / 128) * 128 + (size_t)slice * 4;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)csz, (unsigned)(B * J), 1);
  cfg.blockDim = dim3(kClThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)csz; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
#define LT_GO(S_, P_)                                                                                                  \
  do {                                                                                                                 \
    LT_SMEM_OPT_IN_ONCE((k_sam_fwd_cluster<S_, P_>), smem);                                                            \
    if (csz > 8) { /* clusters of 16 are outside the portable range: opt in once per (function, device) */            \
      static bool np_done[64] = {};                                                                                    \
      int dev = 0;                                                                                                     \
      cudaError_t e = cudaGetDevice(&dev);                                                                             \
      if (e != cudaSuccess) return e;                                                                                  \
      if (dev < 0 || dev >= 64 || !np_done[dev]) {                                                                     \
        e = cudaFuncSetAttribute(k_sam_fwd_cluster<S_, P_>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);        \
        if (e != cudaSuccess) return e;                                                                                \
        if (dev >= 0 && dev < 64) np_done[dev] = true;                                                                 \
      }                                                                                                                \
    }                                                                                                                  \
    return cudaLaunchKernelEx(&cfg, k_sam_fwd_cluster<S_, P_>, vol, cs, stats, coords, prob, J, (int)nvox, slice, mult); \
  } while (0)
  if (cs.spec != nullptr) { if (softmax) LT_GO(true, true); else LT_GO(false, true); }
  else { if (softmax) LT_GO(true, false); else LT_GO(false, false); }
#undef LT_GO
}

}  // namespace lt
