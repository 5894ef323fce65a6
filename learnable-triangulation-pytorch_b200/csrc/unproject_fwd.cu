// unproject_fwd.cu — K1 forward: fused unproject + view aggregation.
//
// Replaces the B·V python iterations of mvn/utils/op.py:107-161 (project, mask, normalise, F.grid_sample,
// zero invalid, stack (V,C,N^3), aggregate) with one launch that never materialises per-view volumes.
// Algorithmic HBM bytes per batch item: F + K + O = V*H*W*Cp*4 + nvox*12 + C*nvox*4.
#include <stdlib.h>

#include "unproject.cuh"

namespace lt {

// AGG: lt_agg.  ALIGNED: N2 % 8 == 0 so every run is a whole, 32-byte aligned sector.
// kAggSoftmax is only launched here for V <= kViewChunk (all samples of a voxel in registers); more views go to
// k_unproject_fwd_softmax_manyviews below.
#ifndef LT_K1F_OPAQUE
#define LT_K1F_OPAQUE(AGG_) ((AGG_) == kAggSoftmax || (AGG_) == kAggConf)
#endif
#ifndef LT_K1F_CTAS
#define LT_K1F_CTAS 4
#endif
template <int LPV, int AGG, bool ALIGNED, bool SPEC>
__global__ void __launch_bounds__(kWarpsPerCta * 32, LT_K1F_CTAS) k_unproject_fwd(const KArgs<SPEC> a) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  constexpr int SLOTS = TG::kSlots;
  constexpr int CPT = LPV < kMaxLanesPerVoxel ? LPV * kCPL : 0;   // texel pitch: Cp == LPV * 4 unless channel-blocked

  // dynamic smem (> 48 KB): projection rows | tap weights | tap offsets | per-lane result slots (4 ch x 8 z steps)
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* s_proj = reinterpret_cast<float4*>(smem_raw);
  auto s_w = reinterpret_cast<float4(*)[kViewChunk][SLOTS]>(s_proj + kMaxViewsSmem * 3);
  auto s_acc = reinterpret_cast<float4(*)[kRun][32]>(&s_w[kWarpsPerCta][0][0]);
  auto s_off = reinterpret_cast<int(*)[kViewChunk][SLOTS]>(&s_acc[kWarpsPerCta][0][0]);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;
  const TileId T = decode_tile(a);

  const float* s_spec = SpecSmem<SPEC>::stage(a, T.b);
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)T.b * a.V * 3 + i);
  __syncthreads();

  const int i0_warp = T.t0 * kTile0 + (warp >> 1) * kWarpRuns0;
  const int i1_warp = T.t1 * kTile1 + (warp & 1) * kWarpRuns1;
  const int z0 = T.t2 * kRun;
  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = (a.W + 2) * a.Cp;
  const size_t viewStride = (size_t)(a.H + 2) * (a.W + 2) * a.Cp;
  const int c_lane = T.cb * (LPV * kCPL) + q * kCPL;  // first channel of this lane
  // the item's maps at this lane's channels, and the row below. Opaque (kept as 64-bit register pairs) where MEASURED
  // faster: softmax 0.497 -> 0.483 ms, conf 0.495 -> 0.481; sum and max lose 2.5 % with it (0.461 -> 0.473)
  const float* fb = a.feat + (size_t)T.b * a.V * viewStride + c_lane;
  const float* fb1 = fb + rowStride;
  if (LT_K1F_OPAQUE(AGG)) { fb = opaque_ptr(fb); fb1 = opaque_ptr(fb1); }

#pragma unroll 1
  for (int r = 0; r < TG::R; ++r) {
    const int i0 = i0_warp + TG::run0(r, g), i1 = i1_warp + TG::run1(r, g);
    const bool run_ok = (i0 < a.N0) && (i1 < a.N1);

#pragma unroll 1
    for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
      const int nv = min(kViewChunk, a.V - v0);
      __syncwarp();
      build_descs<LPV, SPEC>(a, T.b, i0_warp, i1_warp, z0, r, s_proj, s_w[warp], s_off[warp], v0, nv, -1, lane, invH, invW, s_spec);
      __syncwarp();

      float cw[kViewChunk][kCPL];
      if (AGG == kAggConf) {
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            const int c = c_lane + k;
            cw[vv][k] = (vv < nv && c < a.C) ? __ldg(a.conf + ((size_t)T.b * a.V + v0 + vv) * a.C + c) : 0.f;
          }
      }

      // z loop kept ROLLED (x2): with the 8 steps unrolled ptxas hoists every descriptor and tap load of the run
      // (~280 live registers, 600 B of spills at the 128 cap) and the SASS grows past the instruction cache.
      // The per-step results wait in this lane's private smem slots instead of 32 accumulator registers.
#pragma unroll 2
      for (int j = 0; j < kRun; ++j) {
        const int slot = j * G + g;
        float val[kCPL];
        if (AGG == kAggSoftmax) {
          float s[kViewChunk][kCPL];
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            const float fill = (vv < nv) ? 0.f : kNoView;  // zeroed views take part with 0, absent ones not at all
#pragma unroll
            for (int k = 0; k < kCPL; ++k) s[vv][k] = fill;
            if (vv < nv) {
              const int off = s_off[warp][vv][slot];
              if (off >= 0)
                gather4_rows<CPT>(fb, fb1, off, s_w[warp][vv][slot], a.Cp, s[vv]);
            }
          }
#pragma unroll
          for (int k = 0; k < kCPL; k += 2) {
            const float ca[4] = {s[0][k], s[1][k], s[2][k], s[3][k]};
            const float cb[4] = {s[0][k + 1], s[1][k + 1], s[2][k + 1], s[3][k + 1]};
            softmax_mix4_x2(ca, cb, val[k], val[k + 1]);
          }
        } else {
          if (v0 == 0) {
#pragma unroll
            for (int k = 0; k < kCPL; ++k) val[k] = (AGG == kAggMax) ? -INFINITY : 0.f;
          } else {
            const float4 p = s_acc[warp][j][lane];
            val[0] = p.x; val[1] = p.y; val[2] = p.z; val[3] = p.w;
          }
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            if (vv < nv) {
              float s[kCPL];
#pragma unroll
              for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
              const int off = s_off[warp][vv][slot];
              if (off >= 0) gather4_rows<CPT>(fb, fb1, off, s_w[warp][vv][slot], a.Cp, s);
#pragma unroll
              for (int k = 0; k < kCPL; ++k) {
                if (AGG == kAggSum) val[k] += s[k];
                else if (AGG == kAggConf) val[k] = fmaf(cw[vv][k], s[k], val[k]);
                else val[k] = fmaxf(val[k], s[k]);  // MAX: zeroed views take part with 0
              }
            }
          }
        }
        s_acc[warp][j][lane] = make_float4(val[0], val[1], val[2], val[3]);
      }
    }

    // ---------------- store: one 32-byte sector per (channel, run) ----------------
    if (run_ok) {
      float acc[kCPL][kRun];
#pragma unroll
      for (int j = 0; j < kRun; ++j) {
        const float4 p = s_acc[warp][j][lane];
        acc[0][j] = p.x; acc[1][j] = p.y; acc[2][j] = p.z; acc[3][j] = p.w;
      }
      float* ob = a.out + (size_t)T.b * a.C * a.nvox + ((long long)i0 * a.N1 + i1) * a.N2 + z0;
#pragma unroll
      for (int k = 0; k < kCPL; ++k) {
        const int c = c_lane + k;
        if (c < a.C) {
          float* p = ob + (size_t)c * a.nvox;
          if (ALIGNED) {
            stg256(p, acc[k]);
          } else {
#pragma unroll
            for (int j = 0; j < kRun; ++j)
              if (z0 + j < a.N2) p[j] = acc[k][j];
          }
        }
      }
    }
  }
}

// softmax over MORE than kViewChunk views: running (m, sum e, sum e*s) per channel; z step outermost so the state
// stays at 3 x 4 registers and phase A covers G voxels x nv views per round. Same tiling, same descriptors.
template <int LPV, bool ALIGNED, bool SPEC>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 2) k_unproject_fwd_softmax_manyviews(const KArgs<SPEC> a) {
  using TG = TileGeom<LPV>;
  constexpr int SLOTS = TG::kSlots;
  constexpr int CPT = LPV < kMaxLanesPerVoxel ? LPV * kCPL : 0;
  __shared__ float4 s_proj[kMaxViewsSmem * 3];
  __shared__ float4 s_w[kWarpsPerCta][kViewChunk][SLOTS];
  __shared__ int s_off[kWarpsPerCta][kViewChunk][SLOTS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;
  const TileId T = decode_tile(a);
  const float* s_spec = SpecSmem<SPEC>::stage(a, T.b);
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)T.b * a.V * 3 + i);
  __syncthreads();

  const int i0_warp = T.t0 * kTile0 + (warp >> 1) * kWarpRuns0;
  const int i1_warp = T.t1 * kTile1 + (warp & 1) * kWarpRuns1;
  const int z0 = T.t2 * kRun;
  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = (a.W + 2) * a.Cp;
  const size_t viewStride = (size_t)(a.H + 2) * (a.W + 2) * a.Cp;
  const int c_lane = T.cb * (LPV * kCPL) + q * kCPL;
  const float* fb = opaque_ptr(a.feat + (size_t)T.b * a.V * viewStride + c_lane);   // the item's maps, this lane's channels
  const float* fb1 = opaque_ptr(fb + rowStride);                                       // the row below

#pragma unroll 1
  for (int r = 0; r < TG::R; ++r) {
    const int i0 = i0_warp + TG::run0(r, g), i1 = i1_warp + TG::run1(r, g);
    const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
    float res[kCPL][kRun];
#pragma unroll
    for (int j = 0; j < kRun; ++j) {
      float m[kCPL], se[kCPL], sw[kCPL];
#pragma unroll
      for (int k = 0; k < kCPL; ++k) { m[k] = -INFINITY; se[k] = 0.f; sw[k] = 0.f; }
#pragma unroll 1
      for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
        const int nv = min(kViewChunk, a.V - v0);
        __syncwarp();
        build_descs<LPV, SPEC>(a, T.b, i0_warp, i1_warp, z0, r, s_proj, s_w[warp], s_off[warp], v0, nv, j, lane, invH, invW, s_spec);
        __syncwarp();
        for (int vv = 0; vv < nv; ++vv) {
          float s[kCPL];
#pragma unroll
          for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
          const int off = s_off[warp][vv][g];
          if (off >= 0) gather4_rows<CPT>(fb, fb1, off, s_w[warp][vv][g], a.Cp, s);
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            const float mn = fmaxf(m[k], s[k]);
            const float rs = fast_ex2((m[k] - mn) * kLog2eF), e = fast_ex2((s[k] - mn) * kLog2eF);  // ex2(-inf) = 0
            se[k] = fmaf(se[k], rs, e);
            sw[k] = fmaf(sw[k], rs, e * s[k]);
            m[k] = mn;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kCPL; ++k) res[k][j] = sw[k] * fast_rcp(se[k]);
    }
    if (run_ok) {
      float* ob = a.out + (size_t)T.b * a.C * a.nvox + ((long long)i0 * a.N1 + i1) * a.N2 + z0;
#pragma unroll
      for (int k = 0; k < kCPL; ++k) {
        const int c = c_lane + k;
        if (c < a.C) {
          float* p = ob + (size_t)c * a.nvox;
          if (ALIGNED) {
            stg256(p, res[k]);
          } else {
#pragma unroll
            for (int j = 0; j < kRun; ++j)
              if (z0 + j < a.N2) p[j] = res[k][j];
          }
        }
      }
    }
  }
}

template <int LPV, bool SPEC>
static cudaError_t launch_fwd_lpv(const UnprojArgs& ua, const CoordSpecArg* spec, int agg, dim3 grid, cudaStream_t st) {
  const KArgs<SPEC> a = make_kargs<SPEC>(ua, spec);
  const bool aligned = (a.N2 % kRun) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 31u) == 0;
  const dim3 block(kWarpsPerCta * 32);
  constexpr int SL = TileGeom<LPV>::kSlots;
  constexpr size_t smem = sizeof(float4) * kMaxViewsSmem * 3 + kWarpsPerCta * kViewChunk * SL * (sizeof(float4) + sizeof(int)) +
                          sizeof(float4) * kWarpsPerCta * kRun * 32;
#define LT_LAUNCH(AGG_)                                                           \
  do {                                                                            \
    if (aligned) {                                                                \
      LT_SMEM_OPT_IN_ONCE((k_unproject_fwd<LPV, AGG_, true, SPEC>), smem);        \
      k_unproject_fwd<LPV, AGG_, true, SPEC><<<grid, block, smem, st>>>(a);       \
    } else {                                                                      \
      LT_SMEM_OPT_IN_ONCE((k_unproject_fwd<LPV, AGG_, false, SPEC>), smem);       \
      k_unproject_fwd<LPV, AGG_, false, SPEC><<<grid, block, smem, st>>>(a);      \
    }                                                                             \
  } while (0)
  switch (agg) {
    case kAggSum: LT_LAUNCH(kAggSum); break;
    case kAggConf: LT_LAUNCH(kAggConf); break;
    case kAggMax: LT_LAUNCH(kAggMax); break;
    case kAggSoftmax:
      if (a.V <= kViewChunk) LT_LAUNCH(kAggSoftmax);
      else if (aligned) k_unproject_fwd_softmax_manyviews<LPV, true, SPEC><<<grid, block, 0, st>>>(a);
      else k_unproject_fwd_softmax_manyviews<LPV, false, SPEC><<<grid, block, 0, st>>>(a);
      break;
    default: return cudaErrorInvalidValue;
  }
#undef LT_LAUNCH
  return cudaGetLastError();
}

void fill_unproj_tiling(UnprojArgs& a, int lpv) {
  a.nt0 = (a.N0 + kTile0 - 1) / kTile0;
  a.nt1 = (a.N1 + kTile1 - 1) / kTile1;
  a.nt2 = (a.N2 + kRun - 1) / kRun;
  a.ncb = a.Cp / (lpv * kCPL);
}

int unproj_lpv(int Cp) { return Cp >= kMaxLanesPerVoxel * kCPL ? kMaxLanesPerVoxel : Cp / kCPL; }

cudaError_t launch_unproject_fwd(const float* feat_hwc, const float* proj, const float* coord,
                                 const float* conf, float* out, const UnprojShape& s, cudaStream_t st,
                                 const CoordSpecArg* spec) {
  UnprojArgs a{};
  a.feat = feat_hwc; a.proj = proj; a.coord = coord; a.conf = conf; a.out = out;
  a.B = s.B; a.V = s.V; a.C = s.C; a.Cp = padded_channels(s.C); a.H = s.H; a.W = s.W;
  a.N0 = s.N0; a.N1 = s.N1; a.N2 = s.N2; a.align_corners = s.align_corners;
  a.nvox = (long long)s.N0 * s.N1 * s.N2;
  if (s.B == 0 || s.C == 0 || a.nvox == 0) return cudaSuccess;
  const int lpv = unproj_lpv(a.Cp);
  fill_unproj_tiling(a, lpv);
  // LT_B200_K1FWD=tma: the persistent shared-memory-staged kernel (unproject_fwd_tma.cu) where it applies
  // (MEASURED at cfg3: 0.76 ms against 0.50 ms for the kernels below, so it is opt-in; read per call so that tests can switch)
  const char* k1fwd = getenv("LT_B200_K1FWD");
  if (k1fwd != nullptr && k1fwd[0] == 't' && unproject_fwd_tma_eligible(a)) return launch_unproject_fwd_tma(a, spec, s.agg, st);
  const long long nblk = (long long)a.B * a.ncb * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  dim3 grid((unsigned)nblk);
  // coordinates from the spec: instantiated for the reference's 32 channels; wider maps take the tensor (K6) instead
  if (spec != nullptr) return lpv == 8 ? launch_fwd_lpv<8, true>(a, spec, s.agg, grid, st) : cudaErrorNotSupported;
  switch (lpv) {
    case 8: return launch_fwd_lpv<8, false>(a, nullptr, s.agg, grid, st);
    case 16: return launch_fwd_lpv<16, false>(a, nullptr, s.agg, grid, st);
    case 32: return launch_fwd_lpv<32, false>(a, nullptr, s.agg, grid, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace lt
