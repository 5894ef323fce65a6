// unproject_bwd.cu — K1 backward: d(volume)/d(heatmaps) and d(volume)/d(vol_confidences).
//
// The reference gets this from autograd through B·V grid_sampler_2d_backward launches, CopySlices over the
// (V,C,N^3) temporaries and the softmax/mul/sum graph of mvn/utils/op.py:147-159. Here one launch
// re-derives the per-view samples (never stored by the forward), forms d out / d sample_v analytically
// and scatter-adds weight*grad into the channels-last gradient map with 16-byte vector reductions
// (REDG.E.ADD.F32x4): the 8 lanes of a voxel cover one 32-channel texel = four complete 32-byte sectors per
// instruction, which is what the L2 reduction unit (1 sector / cycle / slice) wants.
//   sum     : d s_v = g
//   conf    : d s_v = conf_v * g ;  d conf_v = sum_voxels g * s_v
//   softmax : out = sum_v s_v p_v, p = softmax_v(s)  =>  d s_v = g * p_v * (1 + s_v - out)
//   max     : d s_v = g for the first view attaining the maximum (zeroed views take part with value 0)
// Voxels behind the camera / outside the image get no gradient (op.py:138).
// Algorithmic HBM bytes per batch item: O (grad_out) + K + F (features, not for `sum`) + F (grad map).
#include "unproject.cuh"

namespace lt {


// scatter-add w_tap * d[0..3] into the four taps; taps with zero weight (outside the map) are skipped.
template <int CPT = 0>
__device__ __forceinline__ void scatter4_rows(float* __restrict__ row0, float* __restrict__ row1, int off, const float4& w,
                                              int Cp, const float (&d)[kCPL]) {
  float* p = row0 + off;
  float* p2 = row1 + off;
  const int dx = CPT > 0 ? CPT : Cp;
  const f2 dlo = pack2(d[0], d[1]), dhi = pack2(d[2], d[3]);
  float a0, a1, a2, a3;
  if (w.x != 0.f) {
    const f2 ws = pack2(w.x, w.x);
    unpack2(mul2(dlo, ws), a0, a1); unpack2(mul2(dhi, ws), a2, a3);
    red_add_v4(p, a0, a1, a2, a3);
  }
  if (w.y != 0.f) {
    const f2 ws = pack2(w.y, w.y);
    unpack2(mul2(dlo, ws), a0, a1); unpack2(mul2(dhi, ws), a2, a3);
    red_add_v4(p + dx, a0, a1, a2, a3);
  }
  if (w.z != 0.f) {
    const f2 ws = pack2(w.z, w.z);
    unpack2(mul2(dlo, ws), a0, a1); unpack2(mul2(dhi, ws), a2, a3);
    red_add_v4(p2, a0, a1, a2, a3);
  }
  if (w.w != 0.f) {
    const f2 ws = pack2(w.w, w.w);
    unpack2(mul2(dlo, ws), a0, a1); unpack2(mul2(dhi, ws), a2, a3);
    red_add_v4(p2 + dx, a0, a1, a2, a3);
  }
}

// common prologue of the backward kernels
struct BwdCtx {
  int warp, lane, g, q, i0_warp, i1_warp, z0, c_lane, rowStride;
  size_t viewStride;
  float invH, invW;
  const float* fb;    // the item's feature maps at this lane's channels; fb1 = one row further down
  const float* fb1;
  float* gb;          // the same positions in the gradient map
  float* gb1;
};

template <int LPV, bool SPEC, bool OPAQUE = true>
__device__ __forceinline__ BwdCtx bwd_prologue(const KArgs<SPEC>& a, const TileId& T, float4* s_proj) {
  BwdCtx c;
  c.warp = threadIdx.x >> 5; c.lane = threadIdx.x & 31;
  c.g = c.lane / LPV; c.q = c.lane % LPV;
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)T.b * a.V * 3 + i);
  __syncthreads();
  c.i0_warp = T.t0 * kTile0 + (c.warp >> 1) * kWarpRuns0;
  c.i1_warp = T.t1 * kTile1 + (c.warp & 1) * kWarpRuns1;
  c.z0 = T.t2 * kRun;
  c.invH = 1.f / (float)a.H; c.invW = 1.f / (float)a.W;
  c.rowStride = (a.W + 2) * a.Cp;
  c.viewStride = (size_t)(a.H + 2) * (a.W + 2) * a.Cp;
  c.c_lane = T.cb * (LPV * kCPL) + c.q * kCPL;
  c.fb = a.feat + (size_t)T.b * a.V * c.viewStride + c.c_lane;
  c.gb = a.gfeat + (size_t)T.b * a.V * c.viewStride + c.c_lane;
  if (OPAQUE) { c.fb = opaque_ptr(c.fb); c.gb = opaque_ptr(c.gb); }
  c.fb1 = c.fb + c.rowStride;
  c.gb1 = c.gb + c.rowStride;
  if (OPAQUE) { c.fb1 = opaque_ptr(c.fb1); c.gb1 = opaque_ptr(c.gb1); }
  return c;
}

// upstream gradient of this lane's 4 channels x 8 voxels: read once with 256-bit streaming loads, parked in the
// lane's private smem slots so that the z loop can stay rolled (see unproject_fwd.cu).
template <bool ALIGNED>
__device__ __forceinline__ void stage_grad_out(const UnprojArgs& a, int b, int c_lane, int i0, int i1, int z0,
                                               bool run_ok, float4 (*s_go)[32], int lane) {
  float go[kCPL][kRun];
  const float* gp = a.gout + (size_t)b * a.C * a.nvox + ((long long)i0 * a.N1 + i1) * a.N2 + z0;
#pragma unroll
  for (int k = 0; k < kCPL; ++k) {
    const int c = c_lane + k;
    if (run_ok && c < a.C) {
      if (ALIGNED) {
        const f8 t = ldg256_stream(gp + (size_t)c * a.nvox);
#pragma unroll
        for (int j = 0; j < kRun; ++j) go[k][j] = t.v[j];
      } else {
#pragma unroll
        for (int j = 0; j < kRun; ++j) go[k][j] = (z0 + j < a.N2) ? __ldg(gp + (size_t)c * a.nvox + j) : 0.f;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kRun; ++j) go[k][j] = 0.f;
    }
  }
#pragma unroll
  for (int j = 0; j < kRun; ++j) s_go[j][lane] = make_float4(go[0][j], go[1][j], go[2][j], go[3][j]);
}

// sum / conf for any V; softmax / max for V <= kViewChunk (all samples of a voxel live in registers).
#ifndef LT_K1B_BATCH
#define LT_K1B_BATCH 0   // softmax / max: views whose tap loads are in flight together (0 = one branch per view)
#endif
#ifndef LT_K1B_OPAQUE
#define LT_K1B_OPAQUE(AGG_) ((AGG_) == kAggSoftmax || (AGG_) == kAggMax)   // MEASURED: conf 1.051 vs 1.122 ms, sum 0.909 vs 0.926 without
#endif
#ifndef LT_K1B_CTAS
#define LT_K1B_CTAS 3
#endif
template <int LPV, int AGG, bool ALIGNED, bool SPEC>
__global__ void __launch_bounds__(kWarpsPerCta * 32, LT_K1B_CTAS) k_unproject_bwd(const KArgs<SPEC> a) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  constexpr int SLOTS = TG::kSlots;
  constexpr int CPT = LPV < kMaxLanesPerVoxel ? LPV * kCPL : 0;   // texel pitch: Cp == LPV * 4 unless channel-blocked
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* s_proj = reinterpret_cast<float4*>(smem_raw);
  auto s_w = reinterpret_cast<float4(*)[kViewChunk][SLOTS]>(s_proj + kMaxViewsSmem * 3);
  auto s_go = reinterpret_cast<float4(*)[kRun][32]>(&s_w[kWarpsPerCta][0][0]);
  auto s_off = reinterpret_cast<int(*)[kViewChunk][SLOTS]>(&s_go[kWarpsPerCta][0][0]);
  float* s_gc = reinterpret_cast<float*>(&s_off[kWarpsPerCta][0][0]);  // conf only: [warp][view][channel of the block]

  const TileId T = decode_tile(a);
  const float* s_spec = SpecSmem<SPEC>::stage(a, T.b);  // published by the barrier inside the prologue
  const BwdCtx c = bwd_prologue<LPV, SPEC, LT_K1B_OPAQUE(AGG)>(a, T, s_proj);
  const int warp = c.warp, lane = c.lane, g = c.g;
  const size_t ntile_rows = (size_t)a.nt0 * a.nt1 * a.nt2 * TG::R;

#pragma unroll 1
  for (int r = 0; r < TG::R; ++r) {
    const int i0 = c.i0_warp + TG::run0(r, g), i1 = c.i1_warp + TG::run1(r, g);
    const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
    stage_grad_out<ALIGNED>(a, T.b, c.c_lane, i0, i1, c.z0, run_ok, s_go[warp], lane);

#pragma unroll 1
    for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
      const int nv = min(kViewChunk, a.V - v0);
      __syncwarp();
      build_descs<LPV, SPEC>(a, T.b, c.i0_warp, c.i1_warp, c.z0, r, s_proj, s_w[warp], s_off[warp], v0, nv, -1, lane,
                       c.invH, c.invW, s_spec);
      __syncwarp();

      float cw[kViewChunk][kCPL];   // conf of this chunk            (conf only)
      float gcw[kViewChunk][kCPL];  // running d conf of this chunk  (conf only)
      if (AGG == kAggConf) {
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            const int ch = c.c_lane + k;
            cw[vv][k] = (vv < nv && ch < a.C) ? __ldg(a.conf + ((size_t)T.b * a.V + v0 + vv) * a.C + ch) : 0.f;
            gcw[vv][k] = 0.f;
          }
      }

#pragma unroll 2
      for (int j = 0; j < kRun; ++j) {
        const int slot = j * G + g;
        const float4 gj = s_go[warp][j][lane];
        const float go[kCPL] = {gj.x, gj.y, gj.z, gj.w};
        if (AGG == kAggSum || AGG == kAggConf) {
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            if (vv < nv) {
              const int off = s_off[warp][vv][slot];
              if (off >= 0) {
                const float4 w = s_w[warp][vv][slot];
                float d[kCPL];
                if (AGG == kAggConf) {
                  float s[kCPL];
                  gather4_rows<CPT>(c.fb, c.fb1, off, w, a.Cp, s);
#pragma unroll
                  for (int k = 0; k < kCPL; ++k) {
                    gcw[vv][k] = fmaf(go[k], s[k], gcw[vv][k]);
                    d[k] = cw[vv][k] * go[k];
                  }
                } else {
#pragma unroll
                  for (int k = 0; k < kCPL; ++k) d[k] = go[k];
                }
                scatter4_rows<CPT>(c.gb, c.gb1, off, w, a.Cp, d);
              }
            }
          }
        } else {
          // softmax / max over <= kViewChunk views, all samples in registers
          float s[kViewChunk][kCPL];
          int offs[kViewChunk];
#if LT_K1B_BATCH
          // The tap loads of LT_K1B_BATCH views are issued before the first blend. ncu's source view says that with a
          // branch per view 48 % of the stall samples are long-scoreboard waits at the first blend of each view, but
          // MEASURED at cfg3: 2 views / 2 CTAs per SM 1.220 ms, 4 views / 2 CTAs 1.164 ms, 2 views / 3 CTAs (100 bytes of
          // spills) 1.123 ms, against 1.036 ms for one branch per view at 3 CTAs: the registers the loads in flight need
          // cost more occupancy (or spills) than the overlap returns. Off by default.
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) offs[vv] = (vv < nv) ? s_off[warp][vv][slot] : -1;
#pragma unroll
          for (int vg = 0; vg < kViewChunk; vg += LT_K1B_BATCH) {
            float4 t[LT_K1B_BATCH][4];
#pragma unroll
            for (int u = 0; u < LT_K1B_BATCH; ++u)
              if (vg + u < nv) issue_taps<CPT>(c.fb, c.fb1, offs[vg + u], a.Cp, t[u]);
#pragma unroll
            for (int u = 0; u < LT_K1B_BATCH; ++u) {
              const int vv = vg + u;
              if (vv < nv) {
                blend_taps(t[u], s_w[warp][vv][slot], offs[vv], s[vv]);
              } else {
#pragma unroll
                for (int k = 0; k < kCPL; ++k) s[vv][k] = kNoView;
              }
            }
          }
#else
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            const float fill = (vv < nv) ? 0.f : kNoView;
#pragma unroll
            for (int k = 0; k < kCPL; ++k) s[vv][k] = fill;
            offs[vv] = -1;
            if (vv < nv) {
              offs[vv] = s_off[warp][vv][slot];
              if (offs[vv] >= 0)
                gather4_rows<CPT>(c.fb, c.fb1, offs[vv], s_w[warp][vv][slot], a.Cp, s[vv]);
            }
          }
#endif
          float d[kViewChunk][kCPL];
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            float m = s[0][k];
            int am = 0;
#pragma unroll
            for (int vv = 1; vv < kViewChunk; ++vv)
              if (s[vv][k] > m) { m = s[vv][k]; am = vv; }
            if (AGG == kAggSoftmax) {
              float e[kViewChunk], den = 0.f, num = 0.f;
#pragma unroll
              for (int vv = 0; vv < kViewChunk; ++vv) {
                e[vv] = fast_ex2((s[vv][k] - m) * kLog2eF);
                den += e[vv];
                num = fmaf(s[vv][k], e[vv], num);
              }
              const float inv = fast_rcp(den);
              const float o1 = 1.f - num * inv;  // 1 - out
              const float gi = go[k] * inv;
#pragma unroll
              for (int vv = 0; vv < kViewChunk; ++vv) d[vv][k] = gi * e[vv] * (o1 + s[vv][k]);
            } else {
#pragma unroll
              for (int vv = 0; vv < kViewChunk; ++vv) d[vv][k] = (vv == am) ? go[k] : 0.f;
            }
          }
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv)
            if (vv < nv && offs[vv] >= 0)
              scatter4_rows<CPT>(c.gb, c.gb1, offs[vv], s_w[warp][vv][slot], a.Cp, d[vv]);
        }
      }

      if (AGG == kAggConf && a.gconf_part != nullptr) {
        // d conf of this (round, view chunk): lane groups of a warp by shuffles, the warps of the CTA through shared
        // memory in warp order, then ONE plain store per (view, channel) into the row of this (tile, round): the rows are
        // summed in a fixed order by k_conf_reduce_tiles -> grad_conf is deterministic (no atomics anywhere).
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            float v = gcw[vv][k];
#pragma unroll
            for (int o = LPV; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (g == 0) s_gc[(warp * kViewChunk + vv) * (LPV * kCPL) + c.q * kCPL + k] = v;
          }
        __syncthreads();
        for (int i = threadIdx.x; i < nv * LPV * kCPL; i += blockDim.x) {
          const int vv = i / (LPV * kCPL), cc = i % (LPV * kCPL);
          float s = 0.f;
#pragma unroll
          for (int wi = 0; wi < kWarpsPerCta; ++wi) s += s_gc[(wi * kViewChunk + vv) * (LPV * kCPL) + cc];
          const int ch = T.cb * (LPV * kCPL) + cc;
          const size_t row = ((size_t)T.b * ntile_rows + (size_t)T.lin * TG::R + r);
          if (ch < a.C) a.gconf_part[(row * a.V + v0 + vv) * a.C + ch] = s;
        }
        __syncthreads();
      }
    }
    __syncwarp();  // s_go of this round is dead before the next round overwrites it (lane-private, but keep order)
  }
}

// softmax / max over MORE than kViewChunk views: per z step, pass 1 = running (max, argmax, sum e, sum e*s)
// over all views, pass 2 = recompute each view's sample and scatter its share.
template <int LPV, int AGG, bool ALIGNED, bool SPEC>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 2) k_unproject_bwd_manyviews(const KArgs<SPEC> a) {
  using TG = TileGeom<LPV>;
  constexpr int SLOTS = TG::kSlots;
  constexpr int CPT = LPV < kMaxLanesPerVoxel ? LPV * kCPL : 0;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float4* s_proj = reinterpret_cast<float4*>(smem_raw);
  auto s_w = reinterpret_cast<float4(*)[kViewChunk][SLOTS]>(s_proj + kMaxViewsSmem * 3);
  auto s_go = reinterpret_cast<float4(*)[kRun][32]>(&s_w[kWarpsPerCta][0][0]);
  auto s_off = reinterpret_cast<int(*)[kViewChunk][SLOTS]>(&s_go[kWarpsPerCta][0][0]);

  const TileId T = decode_tile(a);
  const float* s_spec = SpecSmem<SPEC>::stage(a, T.b);  // published by the barrier inside the prologue
  const BwdCtx c = bwd_prologue<LPV, SPEC>(a, T, s_proj);
  const int warp = c.warp, lane = c.lane, g = c.g;

#pragma unroll 1
  for (int r = 0; r < TG::R; ++r) {
    const int i0 = c.i0_warp + TG::run0(r, g), i1 = c.i1_warp + TG::run1(r, g);
    const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
    stage_grad_out<ALIGNED>(a, T.b, c.c_lane, i0, i1, c.z0, run_ok, s_go[warp], lane);
#pragma unroll 1
    for (int j = 0; j < kRun; ++j) {
      const float4 gj = s_go[warp][j][lane];
      const float go[kCPL] = {gj.x, gj.y, gj.z, gj.w};
      float m[kCPL], se[kCPL], sw[kCPL];
      int am[kCPL];
#pragma unroll
      for (int k = 0; k < kCPL; ++k) { m[k] = -INFINITY; se[k] = 0.f; sw[k] = 0.f; am[k] = 0; }
      // pass 1
      for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
        const int nv = min(kViewChunk, a.V - v0);
        __syncwarp();
        build_descs<LPV, SPEC>(a, T.b, c.i0_warp, c.i1_warp, c.z0, r, s_proj, s_w[warp], s_off[warp], v0, nv, j, lane,
                         c.invH, c.invW, s_spec);
        __syncwarp();
        for (int vv = 0; vv < nv; ++vv) {
          float s[kCPL];
#pragma unroll
          for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
          const int off = s_off[warp][vv][g];
          if (off >= 0) gather4_rows<CPT>(c.fb, c.fb1, off, s_w[warp][vv][g], a.Cp, s);
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            if (s[k] > m[k]) am[k] = v0 + vv;
            const float mn = fmaxf(m[k], s[k]);
            if (AGG == kAggSoftmax) {
              const float rs = fast_ex2((m[k] - mn) * kLog2eF), e = fast_ex2((s[k] - mn) * kLog2eF);
              se[k] = fmaf(se[k], rs, e);
              sw[k] = fmaf(sw[k], rs, e * s[k]);
            }
            m[k] = mn;
          }
        }
      }
      float o1[kCPL], gi[kCPL];
#pragma unroll
      for (int k = 0; k < kCPL; ++k) {
        const float inv = (AGG == kAggSoftmax) ? fast_rcp(se[k]) : 0.f;
        o1[k] = 1.f - sw[k] * inv;
        gi[k] = go[k] * inv;
      }
      // pass 2
      for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
        const int nv = min(kViewChunk, a.V - v0);
        __syncwarp();
        build_descs<LPV, SPEC>(a, T.b, c.i0_warp, c.i1_warp, c.z0, r, s_proj, s_w[warp], s_off[warp], v0, nv, j, lane,
                         c.invH, c.invW, s_spec);
        __syncwarp();
        for (int vv = 0; vv < nv; ++vv) {
          const int off = s_off[warp][vv][g];
          if (off < 0) continue;
          const float4 w = s_w[warp][vv][g];
          float d[kCPL];
          if (AGG == kAggSoftmax) {
            float s[kCPL];
            gather4_rows<CPT>(c.fb, c.fb1, off, w, a.Cp, s);
#pragma unroll
            for (int k = 0; k < kCPL; ++k) d[k] = gi[k] * fast_ex2((s[k] - m[k]) * kLog2eF) * (o1[k] + s[k]);
          } else {
            bool any = false;
#pragma unroll
            for (int k = 0; k < kCPL; ++k) {
              d[k] = (am[k] == v0 + vv) ? go[k] : 0.f;
              any |= (am[k] == v0 + vv);
            }
            if (!any) continue;
          }
          scatter4_rows<CPT>(c.gb, c.gb1, off, w, a.Cp, d);
        }
      }
    }
    __syncwarp();
  }
}

template <int LPV, bool SPEC>
static cudaError_t launch_bwd_lpv(const UnprojArgs& ua, const CoordSpecArg* spec, int agg, dim3 grid, cudaStream_t st) {
  const KArgs<SPEC> a = make_kargs<SPEC>(ua, spec);
  const bool aligned = (a.N2 % kRun) == 0 && (reinterpret_cast<uintptr_t>(a.gout) & 31u) == 0;
  const dim3 block(kWarpsPerCta * 32);
  constexpr int SL = TileGeom<LPV>::kSlots;
  constexpr size_t smem_base = sizeof(float4) * kMaxViewsSmem * 3 +
                               kWarpsPerCta * kViewChunk * SL * (sizeof(float4) + sizeof(int)) +
                               sizeof(float4) * kWarpsPerCta * kRun * 32;
  constexpr size_t smem_conf = sizeof(float) * kWarpsPerCta * kViewChunk * LPV * kCPL;  // grad_conf staging (conf only)
#define LT_LAUNCH(K, AGG_)                                                                    \
  do {                                                                                        \
    const size_t smem = smem_base + (AGG_ == kAggConf ? smem_conf : 0);                       \
    if (aligned) {                                                                            \
      LT_SMEM_OPT_IN_ONCE((K<LPV, AGG_, true, SPEC>), smem);                                  \
      K<LPV, AGG_, true, SPEC><<<grid, block, smem, st>>>(a);                                 \
    } else {                                                                                  \
      LT_SMEM_OPT_IN_ONCE((K<LPV, AGG_, false, SPEC>), smem);                                 \
      K<LPV, AGG_, false, SPEC><<<grid, block, smem, st>>>(a);                                \
    }                                                                                         \
  } while (0)
  switch (agg) {
    case kAggSum: LT_LAUNCH(k_unproject_bwd, kAggSum); break;
    case kAggConf: LT_LAUNCH(k_unproject_bwd, kAggConf); break;
    case kAggMax:
      if (a.V <= kViewChunk) LT_LAUNCH(k_unproject_bwd, kAggMax);
      else LT_LAUNCH(k_unproject_bwd_manyviews, kAggMax);
      break;
    case kAggSoftmax:
      if (a.V <= kViewChunk) LT_LAUNCH(k_unproject_bwd, kAggSoftmax);
      else LT_LAUNCH(k_unproject_bwd_manyviews, kAggSoftmax);
      break;
    default: return cudaErrorInvalidValue;
  }
#undef LT_LAUNCH
  return cudaGetLastError();
}

#define TILE_ROUNDS(lpv) ((kWarpRuns0 * kWarpRuns1) / (32 / (lpv)))
size_t unproject_bwd_ws_bytes(int B, int V, int C, int N0, int N1, int N2) {
  // one row of V*C partial sums per (item, 8x8x8 tile, round of the tile's warps): rounds = lanes per voxel / 4
  const int lpv_ = unproj_lpv(padded_channels(C));
  const size_t rows = (size_t)((N0 + kTile0 - 1) / kTile0) * ((N1 + kTile1 - 1) / kTile1) * ((N2 + kRun - 1) / kRun) *
                      TILE_ROUNDS(lpv_);
  const size_t scatter = (size_t)B * rows * V * C * sizeof(float);
  const UnprojShape s{B, V, C, 0, 0, N0, N1, N2, kAggConf, 0};
  const size_t tile = unproject_bwd_tile_ws_bytes(s);  // whichever kernel the launch picks, the buffer is big enough
  return scatter > tile ? scatter : tile;
}

cudaError_t launch_unproject_bwd(const float* grad_out, const float* feat_hwc, const float* proj,
                                 const float* coord, const float* conf, float* grad_feat_hwc,
                                 float* grad_conf, void* conf_ws, const UnprojShape& s, cudaStream_t st,
                                 const CoordSpecArg* spec) {
  UnprojArgs a{};
  a.feat = feat_hwc; a.proj = proj; a.coord = coord; a.conf = conf;
  a.gout = grad_out; a.gfeat = grad_feat_hwc;
  a.B = s.B; a.V = s.V; a.C = s.C; a.Cp = padded_channels(s.C); a.H = s.H; a.W = s.W;
  a.N0 = s.N0; a.N1 = s.N1; a.N2 = s.N2; a.align_corners = s.align_corners;
  a.nvox = (long long)s.N0 * s.N1 * s.N2;
  const bool want_gconf = (grad_conf != nullptr);
  const size_t nconf = (size_t)s.B * s.V * s.C;
  if (want_gconf && (s.agg != kAggConf || a.nvox == 0)) {
    cudaError_t e = cudaMemsetAsync(grad_conf, 0, nconf * sizeof(float), st);
    if (e != cudaSuccess) return e;
  }
  if (s.B == 0 || s.C == 0 || a.nvox == 0 || s.V == 0) return cudaSuccess;
  a.gconf_part = nullptr;
  if (unproject_bwd_tile_eligible(s, grad_out)) {
    // on-chip merged formulation (unproject_bwd_tile.cu); its grad_conf partials need no zero-fill
    if (want_gconf && s.agg == kAggConf) a.gconf_part = static_cast<float*>(conf_ws);
    return launch_unproject_bwd_tile(a, s, grad_conf, st, spec);
  }
  if (want_gconf && s.agg == kAggConf) a.gconf_part = static_cast<float*>(conf_ws);  // every row is fully written
  const int lpv = unproj_lpv(a.Cp);
  fill_unproj_tiling(a, lpv);
  const long long nblk = (long long)a.B * a.ncb * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  dim3 grid((unsigned)nblk);
  cudaError_t e = cudaErrorInvalidValue;
  if (spec != nullptr) {
    e = lpv == 8 ? launch_bwd_lpv<8, true>(a, spec, s.agg, grid, st) : cudaErrorNotSupported;
  } else {
    switch (lpv) {
      case 8: e = launch_bwd_lpv<8, false>(a, nullptr, s.agg, grid, st); break;
      case 16: e = launch_bwd_lpv<16, false>(a, nullptr, s.agg, grid, st); break;
      case 32: e = launch_bwd_lpv<32, false>(a, nullptr, s.agg, grid, st); break;
    }
  }
  if (e != cudaSuccess) return e;
  if (a.gconf_part != nullptr)
    e = launch_conf_reduce_rows(a.gconf_part, grad_conf, s.B, s.V * s.C, a.nt0 * a.nt1 * a.nt2 * TILE_ROUNDS(lpv), st);
  return e;
}

}  // namespace lt
