// unproject_bwd.cu — K1 backward: d(volume)/d(heatmaps) and d(volume)/d(vol_confidences).
//
// The reference gets this from autograd through B·V grid_sampler_2d_backward launches, CopySlices over the
// (V,C,N^3) temporaries and the softmax/mul/sum graph of mvn/utils/op.py:147-159. Here one launch
// re-derives the per-view samples (never stored by the forward), forms d out / d sample_v analytically
// and scatter-adds weight*grad into the channels-last gradient map with 16-byte vector reductions
// (REDG.E.ADD.F32x4): 4 lanes x 2 reductions cover one 32-channel texel.
//   sum     : d s_v = g
//   conf    : d s_v = conf_v * g ;  d conf_v = sum_voxels g * s_v
//   softmax : out = sum_v s_v p_v, p = softmax_v(s)  =>  d s_v = g * p_v * (1 + s_v - out)
//   max     : d s_v = g for the first view attaining the maximum (zeroed views take part with value 0)
// Voxels behind the camera / outside the image get no gradient (op.py:138).
// Algorithmic HBM bytes per batch item: O (grad_out) + K + F (features, not for `sum`) + F (grad map).
#include "unproject.cuh"

namespace lt {

constexpr int kConfSlots = 64;  // grad_conf partial-sum slots per batch item

__device__ __forceinline__ void gather8(const float* __restrict__ fv, int off, const float4& w, int Cp,
                                        int rowStride, float (&s)[kCPL]) {
  const float* p = fv + (off & kOffMask);
  const int dx = (off & kOffDx) ? Cp : 0;
  const int dy = (off & kOffDy) ? rowStride : 0;
  f8 t00 = ldg256_nc(p);
  f8 t01 = ldg256_nc(p + dx);
  f8 t10 = ldg256_nc(p + dy);
  f8 t11 = ldg256_nc(p + dy + dx);
#pragma unroll
  for (int k = 0; k < kCPL; ++k) {
    float a = t00.v[k] * w.x;
    a = fmaf(t01.v[k], w.y, a);
    a = fmaf(t10.v[k], w.z, a);
    a = fmaf(t11.v[k], w.w, a);
    s[k] = a;
  }
}

// scatter-add w_tap * d[0..7] into the four taps; taps with zero weight (outside the map) are skipped.
__device__ __forceinline__ void scatter8(float* __restrict__ gv, int off, const float4& w, int Cp,
                                         int rowStride, const float (&d)[kCPL]) {
  float* p = gv + (off & kOffMask);
  const int dx = (off & kOffDx) ? Cp : 0;
  const int dy = (off & kOffDy) ? rowStride : 0;
  if (w.x != 0.f) {
    red_add_v4(p, d[0] * w.x, d[1] * w.x, d[2] * w.x, d[3] * w.x);
    red_add_v4(p + 4, d[4] * w.x, d[5] * w.x, d[6] * w.x, d[7] * w.x);
  }
  if (w.y != 0.f) {
    red_add_v4(p + dx, d[0] * w.y, d[1] * w.y, d[2] * w.y, d[3] * w.y);
    red_add_v4(p + dx + 4, d[4] * w.y, d[5] * w.y, d[6] * w.y, d[7] * w.y);
  }
  if (w.z != 0.f) {
    red_add_v4(p + dy, d[0] * w.z, d[1] * w.z, d[2] * w.z, d[3] * w.z);
    red_add_v4(p + dy + 4, d[4] * w.z, d[5] * w.z, d[6] * w.z, d[7] * w.z);
  }
  if (w.w != 0.f) {
    red_add_v4(p + dy + dx, d[0] * w.w, d[1] * w.w, d[2] * w.w, d[3] * w.w);
    red_add_v4(p + dy + dx + 4, d[4] * w.w, d[5] * w.w, d[6] * w.w, d[7] * w.w);
  }
}

struct BwdTile {
  int b, cb, i0_warp, i1_warp, z0, i0, i1, tile_lin;
  bool run_ok;
  long long run_base;
};

template <int LPV>
__device__ __forceinline__ BwdTile decode_tile(const UnprojArgs& a, int warp, int g) {
  using TG = TileGeom<LPV>;
  BwdTile r;
  int t = blockIdx.x;
  const int t2 = t % a.nt2; t /= a.nt2;
  const int t1 = t % a.nt1; t /= a.nt1;
  const int t0 = t % a.nt0; t /= a.nt0;
  r.cb = t % a.ncb;
  r.b = t / a.ncb;
  r.tile_lin = (t0 * a.nt1 + t1) * a.nt2 + t2;
  r.i0_warp = t0 * TG::T0 + (warp >> 1) * TG::G0;
  r.i1_warp = t1 * TG::T1 + (warp & 1) * TG::G1;
  r.z0 = t2 * kRun;
  r.i0 = r.i0_warp + g / TG::G1;
  r.i1 = r.i1_warp + g % TG::G1;
  r.run_ok = (r.i0 < a.N0) && (r.i1 < a.N1);
  r.run_base = ((long long)r.i0 * a.N1 + r.i1) * a.N2 + r.z0;
  return r;
}

// phase A (same descriptors as the forward) for `nslots` voxel slots x nv views of this warp.
// zsel < 0: all 8 z steps (slot = j*G + g); zsel >= 0: only z step zsel (slot = g).
template <int LPV>
__device__ __forceinline__ void build_descs(const UnprojArgs& a, const BwdTile& T, const float4* s_proj,
                                            float4 (*s_w)[TileGeom<LPV>::kSlots],
                                            int (*s_off)[TileGeom<LPV>::kSlots], int v0, int nv, int zsel,
                                            int lane, float invH, float invW) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  const float* coord_b = a.coord + (size_t)T.b * a.nvox * 3;
  const int nslots = zsel < 0 ? TG::kSlots : G;
  for (int slot = lane; slot < nslots; slot += 32) {
    const int j = zsel < 0 ? slot / G : zsel, gg = slot % G;
    const int vi0 = T.i0_warp + gg / TG::G1, vi1 = T.i1_warp + gg % TG::G1, vz = T.z0 + j;
    const bool ok = (vi0 < a.N0) && (vi1 < a.N1) && (vz < a.N2);
    float x = 0.f, y = 0.f, z = 0.f;
    if (ok) {
      const float* cp = coord_b + (((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
      x = __ldg(cp); y = __ldg(cp + 1); z = __ldg(cp + 2);
    }
    for (int vv = 0; vv < nv; ++vv) {
      int off = -1;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok) {
        const int v = v0 + vv;
        float4 r0, r1, r2;
        if (v < kMaxViewsSmem) {
          r0 = s_proj[v * 3]; r1 = s_proj[v * 3 + 1]; r2 = s_proj[v * 3 + 2];
        } else {
          const float4* pp = reinterpret_cast<const float4*>(a.proj) + ((size_t)T.b * a.V + v) * 3;
          r0 = __ldg(pp); r1 = __ldg(pp + 1); r2 = __ldg(pp + 2);
        }
        make_desc(x, y, z, r0, r1, r2, a.H, a.W, a.Cp, invH, invW, a.align_corners, off, w);
      }
      s_w[vv][slot] = w;
      s_off[vv][slot] = off;
    }
  }
}

// sum / conf for any V; softmax / max for V <= kViewChunk (all samples of a voxel live in registers).
template <int LPV, int AGG, bool ALIGNED>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 1) k_unproject_bwd(const UnprojArgs a) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  constexpr int SLOTS = TG::kSlots;
  __shared__ float4 s_proj[kMaxViewsSmem * 3];
  __shared__ float4 s_w[kWarpsPerCta][kViewChunk][SLOTS];
  __shared__ int s_off[kWarpsPerCta][kViewChunk][SLOTS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;
  const BwdTile T = decode_tile<LPV>(a, warp, g);

  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)T.b * a.V * 3 + i);
  __syncthreads();

  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = a.W * a.Cp;
  const size_t viewStride = (size_t)a.H * a.W * a.Cp;
  const size_t lane_choff = (size_t)T.cb * (LPV * kCPL) + q * kCPL;
  const float* fb = a.feat + (size_t)T.b * a.V * viewStride + lane_choff;
  float* gb = a.gfeat + (size_t)T.b * a.V * viewStride + lane_choff;
  const int c_lane = (int)lane_choff;

  // upstream gradient of this lane's 8 channels x 8 voxels (read once: streaming loads)
  float go[kCPL][kRun];
  {
    const float* gp = a.gout + (size_t)T.b * a.C * a.nvox + T.run_base;
#pragma unroll
    for (int k = 0; k < kCPL; ++k) {
      const int c = c_lane + k;
      if (T.run_ok && c < a.C) {
        if (ALIGNED) {
          f8 t = ldg256_stream(gp + (size_t)c * a.nvox);
#pragma unroll
          for (int j = 0; j < kRun; ++j) go[k][j] = t.v[j];
        } else {
#pragma unroll
          for (int j = 0; j < kRun; ++j) go[k][j] = (T.z0 + j < a.N2) ? __ldg(gp + (size_t)c * a.nvox + j) : 0.f;
        }
      } else {
#pragma unroll
        for (int j = 0; j < kRun; ++j) go[k][j] = 0.f;
      }
    }
  }

  for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
    const int nv = min(kViewChunk, a.V - v0);
    __syncwarp();
    build_descs<LPV>(a, T, s_proj, s_w[warp], s_off[warp], v0, nv, -1, lane, invH, invW);
    __syncwarp();

    float cw[kViewChunk][kCPL];   // conf of this chunk            (conf only)
    float gcw[kViewChunk][kCPL];  // running d conf of this chunk  (conf only)
    if (AGG == kAggConf) {
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          const int c = c_lane + k;
          cw[vv][k] = (vv < nv && c < a.C) ? __ldg(a.conf + ((size_t)T.b * a.V + v0 + vv) * a.C + c) : 0.f;
          gcw[vv][k] = 0.f;
        }
    }

#pragma unroll
    for (int j = 0; j < kRun; ++j) {
      const int slot = j * G + g;
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
                gather8(fb + (size_t)(v0 + vv) * viewStride, off, w, a.Cp, rowStride, s);
#pragma unroll
                for (int k = 0; k < kCPL; ++k) {
                  gcw[vv][k] = fmaf(go[k][j], s[k], gcw[vv][k]);
                  d[k] = cw[vv][k] * go[k][j];
                }
              } else {
#pragma unroll
                for (int k = 0; k < kCPL; ++k) d[k] = go[k][j];
              }
              scatter8(gb + (size_t)(v0 + vv) * viewStride, off, w, a.Cp, rowStride, d);
            }
          }
        }
      } else {
        // softmax / max over <= kViewChunk views, all samples in registers
        float s[kViewChunk][kCPL];
        int offs[kViewChunk];
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv) {
#pragma unroll
          for (int k = 0; k < kCPL; ++k) s[vv][k] = 0.f;
          offs[vv] = -1;
          if (vv < nv) {
            offs[vv] = s_off[warp][vv][slot];
            if (offs[vv] >= 0)
              gather8(fb + (size_t)(v0 + vv) * viewStride, offs[vv], s_w[warp][vv][slot], a.Cp, rowStride, s[vv]);
          }
        }
        float d[kViewChunk][kCPL];
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          float m = s[0][k];
          int am = 0;
#pragma unroll
          for (int vv = 1; vv < kViewChunk; ++vv)
            if (vv < nv && s[vv][k] > m) { m = s[vv][k]; am = vv; }
          if (AGG == kAggSoftmax) {
            float e[kViewChunk], den = 0.f;
#pragma unroll
            for (int vv = 0; vv < kViewChunk; ++vv) {
              e[vv] = (vv < nv) ? expf(s[vv][k] - m) : 0.f;
              den += e[vv];
            }
            const float inv = __fdiv_rn(1.f, den);
            float o = 0.f;
#pragma unroll
            for (int vv = 0; vv < kViewChunk; ++vv) o = fmaf(s[vv][k], e[vv] * inv, o);
#pragma unroll
            for (int vv = 0; vv < kViewChunk; ++vv) d[vv][k] = go[k][j] * (e[vv] * inv) * (1.f + s[vv][k] - o);
          } else {
#pragma unroll
            for (int vv = 0; vv < kViewChunk; ++vv) d[vv][k] = (vv == am) ? go[k][j] : 0.f;
          }
        }
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
          if (vv < nv && offs[vv] >= 0)
            scatter8(gb + (size_t)(v0 + vv) * viewStride, offs[vv], s_w[warp][vv][slot], a.Cp, rowStride, d[vv]);
      }
    }

    if (AGG == kAggConf && a.gconf_part != nullptr) {
      // reduce d conf over the G lane groups of the warp, then one reduction per (view, channel) per warp
      // into one of kConfSlots partial buffers (summed by k_conf_reduce).
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          float v = gcw[vv][k];
#pragma unroll
          for (int o = LPV; o < 32; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          gcw[vv][k] = v;
        }
      if (g == 0) {
        const int slot = (T.tile_lin * kWarpsPerCta + warp) % kConfSlots;
        float* dst = a.gconf_part + (((size_t)T.b * kConfSlots + slot) * a.V + v0) * a.C;
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
          if (vv < nv)
#pragma unroll
            for (int k = 0; k < kCPL; ++k)
              if (c_lane + k < a.C) atomicAdd(dst + (size_t)vv * a.C + c_lane + k, gcw[vv][k]);
      }
    }
  }
}

// softmax / max over MORE than kViewChunk views: per z step, pass 1 = running (max, argmax, sum e, sum e*s)
// over all views, pass 2 = recompute each view's sample and scatter its share.
template <int LPV, int AGG, bool ALIGNED>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 1) k_unproject_bwd_manyviews(const UnprojArgs a) {
  using TG = TileGeom<LPV>;
  constexpr int SLOTS = TG::kSlots;
  __shared__ float4 s_proj[kMaxViewsSmem * 3];
  __shared__ float4 s_w[kWarpsPerCta][kViewChunk][SLOTS];
  __shared__ int s_off[kWarpsPerCta][kViewChunk][SLOTS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;
  const BwdTile T = decode_tile<LPV>(a, warp, g);
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)T.b * a.V * 3 + i);
  __syncthreads();

  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = a.W * a.Cp;
  const size_t viewStride = (size_t)a.H * a.W * a.Cp;
  const size_t lane_choff = (size_t)T.cb * (LPV * kCPL) + q * kCPL;
  const float* fb = a.feat + (size_t)T.b * a.V * viewStride + lane_choff;
  float* gb = a.gfeat + (size_t)T.b * a.V * viewStride + lane_choff;
  const int c_lane = (int)lane_choff;
  const float* gp = a.gout + (size_t)T.b * a.C * a.nvox + T.run_base;

  for (int j = 0; j < kRun; ++j) {
    float go[kCPL];
#pragma unroll
    for (int k = 0; k < kCPL; ++k) {
      const int c = c_lane + k;
      go[k] = (T.run_ok && c < a.C && T.z0 + j < a.N2) ? __ldg(gp + (size_t)c * a.nvox + j) : 0.f;
    }
    float m[kCPL], se[kCPL], sw[kCPL];
    int am[kCPL];
#pragma unroll
    for (int k = 0; k < kCPL; ++k) { m[k] = -INFINITY; se[k] = 0.f; sw[k] = 0.f; am[k] = 0; }
    // pass 1
    for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
      const int nv = min(kViewChunk, a.V - v0);
      __syncwarp();
      build_descs<LPV>(a, T, s_proj, s_w[warp], s_off[warp], v0, nv, j, lane, invH, invW);
      __syncwarp();
      for (int vv = 0; vv < nv; ++vv) {
        float s[kCPL];
#pragma unroll
        for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
        const int off = s_off[warp][vv][g];
        if (off >= 0) gather8(fb + (size_t)(v0 + vv) * viewStride, off, s_w[warp][vv][g], a.Cp, rowStride, s);
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          if (s[k] > m[k]) am[k] = v0 + vv;
          const float mn = fmaxf(m[k], s[k]);
          if (AGG == kAggSoftmax) {
            const float r = expf(m[k] - mn), e = expf(s[k] - mn);
            se[k] = fmaf(se[k], r, e);
            sw[k] = fmaf(sw[k], r, e * s[k]);
          }
          m[k] = mn;
        }
      }
    }
    float outv[kCPL], inv[kCPL];
#pragma unroll
    for (int k = 0; k < kCPL; ++k) {
      inv[k] = (AGG == kAggSoftmax) ? __fdiv_rn(1.f, se[k]) : 0.f;
      outv[k] = sw[k] * inv[k];
    }
    // pass 2
    for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
      const int nv = min(kViewChunk, a.V - v0);
      __syncwarp();
      build_descs<LPV>(a, T, s_proj, s_w[warp], s_off[warp], v0, nv, j, lane, invH, invW);
      __syncwarp();
      for (int vv = 0; vv < nv; ++vv) {
        const int off = s_off[warp][vv][g];
        if (off < 0) continue;
        const float4 w = s_w[warp][vv][g];
        float d[kCPL];
        if (AGG == kAggSoftmax) {
          float s[kCPL];
          gather8(fb + (size_t)(v0 + vv) * viewStride, off, w, a.Cp, rowStride, s);
#pragma unroll
          for (int k = 0; k < kCPL; ++k) d[k] = go[k] * (expf(s[k] - m[k]) * inv[k]) * (1.f + s[k] - outv[k]);
        } else {
          bool any = false;
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            d[k] = (am[k] == v0 + vv) ? go[k] : 0.f;
            any |= (am[k] == v0 + vv);
          }
          if (!any) continue;
        }
        scatter8(gb + (size_t)(v0 + vv) * viewStride, off, w, a.Cp, rowStride, d);
      }
    }
  }
}

// grad_conf[b,v,c] = sum over the kConfSlots partial buffers
__global__ void k_conf_reduce(const float* __restrict__ part, float* __restrict__ gconf, int VC, int total) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int b = i / VC, r = i % VC;
  float s = 0.f;
  for (int k = 0; k < kConfSlots; ++k) s += part[((size_t)b * kConfSlots + k) * VC + r];
  gconf[i] = s;
}

template <int LPV>
static cudaError_t launch_bwd_lpv(const UnprojArgs& a, int agg, dim3 grid, cudaStream_t st) {
  const bool aligned = (a.N2 % kRun) == 0;
  const dim3 block(kWarpsPerCta * 32);
#define LT_LAUNCH(K, AGG_)                                      \
  do {                                                          \
    if (aligned) K<LPV, AGG_, true><<<grid, block, 0, st>>>(a); \
    else K<LPV, AGG_, false><<<grid, block, 0, st>>>(a);        \
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

size_t unproject_bwd_ws_bytes(int B, int V, int C, int, int, int) {
  return (size_t)B * kConfSlots * V * C * sizeof(float);
}

cudaError_t launch_unproject_bwd(const float* grad_out, const float* feat_hwc, const float* proj,
                                 const float* coord, const float* conf, float* grad_feat_hwc,
                                 float* grad_conf, void* conf_ws, const UnprojShape& s, cudaStream_t st) {
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
  if (want_gconf && s.agg == kAggConf) {
    a.gconf_part = static_cast<float*>(conf_ws);
    cudaError_t e = cudaMemsetAsync(conf_ws, 0, unproject_bwd_ws_bytes(s.B, s.V, s.C, 0, 0, 0), st);
    if (e != cudaSuccess) return e;
  }
  const int lpv = unproj_lpv(a.Cp);
  fill_unproj_tiling(a, lpv);
  const long long nblk = (long long)a.B * a.ncb * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  dim3 grid((unsigned)nblk);
  cudaError_t e = cudaErrorInvalidValue;
  switch (lpv) {
    case 4: e = launch_bwd_lpv<4>(a, s.agg, grid, st); break;
    case 8: e = launch_bwd_lpv<8>(a, s.agg, grid, st); break;
    case 16: e = launch_bwd_lpv<16>(a, s.agg, grid, st); break;
    case 32: e = launch_bwd_lpv<32>(a, s.agg, grid, st); break;
  }
  if (e != cudaSuccess) return e;
  if (a.gconf_part != nullptr) {
    const int total = (int)nconf;
    k_conf_reduce<<<(total + 255) / 256, 256, 0, st>>>(a.gconf_part, grad_conf, s.V * s.C, total);
    e = cudaGetLastError();
  }
  return e;
}

}  // namespace lt
