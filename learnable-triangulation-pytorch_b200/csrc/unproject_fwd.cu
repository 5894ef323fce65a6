// unproject_fwd.cu — K1 forward: fused unproject + view aggregation.
//
// Replaces the B·V python iterations of mvn/utils/op.py:107-161 (project, mask, normalise, F.grid_sample,
// zero invalid, stack (V,C,N^3), aggregate) with one launch that never materialises per-view volumes.
// Algorithmic HBM bytes per batch item: F + K + O = V*H*W*Cp*4 + nvox*12 + C*nvox*4.
#include "unproject.cuh"

namespace lt {

// per-lane gather of one tap set: 4 x LDG.E.256 (8 channels each) and the bilinear blend, accumulated
// in the nw, ne, sw, se order of ATen's grid_sampler_2d.
__device__ __forceinline__ void sample8(const float* __restrict__ fv, int off, const float4& w, int Cp,
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

// AGG: lt_agg.  ALIGNED: N2 % 8 == 0 so every run is a whole, 32-byte aligned sector.
// kAggSoftmax is only launched here for V <= kViewChunk (exact two-pass softmax in registers); more
// views go to k_unproject_fwd_softmax_manyviews below.
template <int LPV, int AGG, bool ALIGNED>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 2) k_unproject_fwd(const UnprojArgs a) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  constexpr int SLOTS = TG::kSlots;

  __shared__ float4 s_proj[kMaxViewsSmem * 3];
  __shared__ float4 s_w[kWarpsPerCta][kViewChunk][SLOTS];
  __shared__ int s_off[kWarpsPerCta][kViewChunk][SLOTS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;

  // tile decode: blockIdx.x = ((b*ncb + cb)*nt0 + t0)*nt1*nt2 + t1*nt2 + t2
  int t = blockIdx.x;
  const int t2 = t % a.nt2; t /= a.nt2;
  const int t1 = t % a.nt1; t /= a.nt1;
  const int t0 = t % a.nt0; t /= a.nt0;
  const int cb = t % a.ncb;
  const int b = t / a.ncb;

  // projection matrices of this batch item -> smem (rows as float4)
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)b * a.V * 3 + i);
  __syncthreads();

  const int w0 = warp >> 1, w1 = warp & 1;
  const int i0_warp = t0 * TG::T0 + w0 * TG::G0;
  const int i1_warp = t1 * TG::T1 + w1 * TG::G1;
  const int z0 = t2 * kRun;
  // this lane group's run
  const int i0 = i0_warp + g / TG::G1, i1 = i1_warp + g % TG::G1;
  const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
  const long long run_base = ((long long)i0 * a.N1 + i1) * a.N2 + z0;

  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = a.W * a.Cp;
  const size_t viewStride = (size_t)a.H * a.W * a.Cp;
  const float* fb = a.feat + (size_t)b * a.V * viewStride + (size_t)cb * (LPV * kCPL) + q * kCPL;
  const float* coord_b = a.coord + (size_t)b * a.nvox * 3;
  const int c_lane = cb * (LPV * kCPL) + q * kCPL;  // first channel of this lane

  float acc[kCPL][kRun];
#pragma unroll
  for (int k = 0; k < kCPL; ++k)
#pragma unroll
    for (int j = 0; j < kRun; ++j) acc[k][j] = (AGG == kAggMax) ? -INFINITY : 0.f;

  for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
    const int nv = min(kViewChunk, a.V - v0);
    // ---------------- phase A: descriptors for SLOTS voxels x nv views ----------------
    __syncwarp();
    for (int slot = lane; slot < SLOTS; slot += 32) {
      const int j = slot / G, gg = slot % G;
      const int vi0 = i0_warp + gg / TG::G1, vi1 = i1_warp + gg % TG::G1, vz = z0 + j;
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
            const float4* pp = reinterpret_cast<const float4*>(a.proj) + ((size_t)b * a.V + v) * 3;
            r0 = __ldg(pp); r1 = __ldg(pp + 1); r2 = __ldg(pp + 2);
          }
          make_desc(x, y, z, r0, r1, r2, a.H, a.W, a.Cp, invH, invW, a.align_corners, off, w);
        }
        s_w[warp][vv][slot] = w;
        s_off[warp][vv][slot] = off;
      }
    }
    __syncwarp();

    // ---------------- phase B: gather + aggregate ----------------
    float cw[kViewChunk][kCPL];
    if (AGG == kAggConf) {
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          const int c = c_lane + k;
          cw[vv][k] = (vv < nv && c < a.C) ? __ldg(a.conf + ((size_t)b * a.V + v0 + vv) * a.C + c) : 0.f;
        }
    }

#pragma unroll
    for (int j = 0; j < kRun; ++j) {
      const int slot = j * G + g;
      if (AGG == kAggSoftmax) {
        // exact two-pass softmax over the (<= kViewChunk) views of this chunk
        float s[kViewChunk][kCPL];
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv) {
#pragma unroll
          for (int k = 0; k < kCPL; ++k) s[vv][k] = 0.f;
          if (vv < nv) {
            const int off = s_off[warp][vv][slot];
            if (off >= 0) sample8(fb + (size_t)(v0 + vv) * viewStride, off, s_w[warp][vv][slot], a.Cp, rowStride, s[vv]);
          }
        }
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          float m = s[0][k];
#pragma unroll
          for (int vv = 1; vv < kViewChunk; ++vv)
            if (vv < nv) m = fmaxf(m, s[vv][k]);
          float e[kViewChunk], den = 0.f;
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            e[vv] = (vv < nv) ? expf(s[vv][k] - m) : 0.f;
            den += e[vv];
          }
          float o = 0.f;
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv)
            if (vv < nv) o = fmaf(s[vv][k], __fdiv_rn(e[vv], den), o);
          acc[k][j] = o;
        }
      } else {
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv) {
          if (vv < nv) {
            float s[kCPL];
#pragma unroll
            for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
            const int off = s_off[warp][vv][slot];
            if (off >= 0) sample8(fb + (size_t)(v0 + vv) * viewStride, off, s_w[warp][vv][slot], a.Cp, rowStride, s);
#pragma unroll
            for (int k = 0; k < kCPL; ++k) {
              if (AGG == kAggSum) acc[k][j] += s[k];
              else if (AGG == kAggConf) acc[k][j] = fmaf(cw[vv][k], s[k], acc[k][j]);
              else acc[k][j] = fmaxf(acc[k][j], s[k]);  // MAX: zeroed views take part with 0
            }
          }
        }
      }
    }
  }

  // ---------------- store: one 32-byte sector per (channel, run) ----------------
  if (!run_ok) return;
  float* ob = a.out + (size_t)b * a.C * a.nvox + run_base;
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

// softmax over MORE than kViewChunk views: running (m, sum e, sum e*s) per channel; one z step at a time
// so that the state stays at 3*8 registers. Same tiling, same descriptors.
template <int LPV, bool ALIGNED>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 2) k_unproject_fwd_softmax_manyviews(const UnprojArgs a) {
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  constexpr int SLOTS = TG::kSlots;
  __shared__ float4 s_w[kWarpsPerCta][kViewChunk][SLOTS];
  __shared__ int s_off[kWarpsPerCta][kViewChunk][SLOTS];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane / LPV, q = lane % LPV;
  int t = blockIdx.x;
  const int t2 = t % a.nt2; t /= a.nt2;
  const int t1 = t % a.nt1; t /= a.nt1;
  const int t0 = t % a.nt0; t /= a.nt0;
  const int cb = t % a.ncb;
  const int b = t / a.ncb;
  const int w0 = warp >> 1, w1 = warp & 1;
  const int i0_warp = t0 * TG::T0 + w0 * TG::G0;
  const int i1_warp = t1 * TG::T1 + w1 * TG::G1;
  const int z0 = t2 * kRun;
  const int i0 = i0_warp + g / TG::G1, i1 = i1_warp + g % TG::G1;
  const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
  const long long run_base = ((long long)i0 * a.N1 + i1) * a.N2 + z0;
  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = a.W * a.Cp;
  const size_t viewStride = (size_t)a.H * a.W * a.Cp;
  const float* fb = a.feat + (size_t)b * a.V * viewStride + (size_t)cb * (LPV * kCPL) + q * kCPL;
  const float* coord_b = a.coord + (size_t)b * a.nvox * 3;
  const int c_lane = cb * (LPV * kCPL) + q * kCPL;

  // z step outermost: the running state is 3 x 8 registers and phase A covers G voxels x nv views.
  float res[kCPL][kRun];
#pragma unroll
  for (int j = 0; j < kRun; ++j) {
    float m[kCPL], se[kCPL], sw[kCPL];
#pragma unroll
    for (int k = 0; k < kCPL; ++k) { m[k] = -INFINITY; se[k] = 0.f; sw[k] = 0.f; }
    for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
      const int nv = min(kViewChunk, a.V - v0);
      __syncwarp();
      // phase A restricted to z step j: G voxels x nv views, one descriptor per lane at most
      for (int d = lane; d < G * nv; d += 32) {
        const int gg = d % G, vv = d / G;
        const int vi0 = i0_warp + gg / TG::G1, vi1 = i1_warp + gg % TG::G1, vz = z0 + j;
        int off = -1;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if ((vi0 < a.N0) && (vi1 < a.N1) && (vz < a.N2)) {
          const float* cp = coord_b + (((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
          const float4* pp = reinterpret_cast<const float4*>(a.proj) + ((size_t)b * a.V + v0 + vv) * 3;
          make_desc(__ldg(cp), __ldg(cp + 1), __ldg(cp + 2), __ldg(pp), __ldg(pp + 1), __ldg(pp + 2), a.H, a.W,
                    a.Cp, invH, invW, a.align_corners, off, w);
        }
        s_w[warp][vv][gg] = w;
        s_off[warp][vv][gg] = off;
      }
      __syncwarp();
      for (int vv = 0; vv < nv; ++vv) {
        float s[kCPL];
#pragma unroll
        for (int k = 0; k < kCPL; ++k) s[k] = 0.f;
        const int off = s_off[warp][vv][g];
        if (off >= 0) sample8(fb + (size_t)(v0 + vv) * viewStride, off, s_w[warp][vv][g], a.Cp, rowStride, s);
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          const float mn = fmaxf(m[k], s[k]);
          const float r = expf(m[k] - mn), e = expf(s[k] - mn);  // expf(-inf)=0 on the first view
          se[k] = fmaf(se[k], r, e);
          sw[k] = fmaf(sw[k], r, e * s[k]);
          m[k] = mn;
        }
      }
    }
#pragma unroll
    for (int k = 0; k < kCPL; ++k) res[k][j] = __fdiv_rn(sw[k], se[k]);
  }

  if (!run_ok) return;
  float* ob = a.out + (size_t)b * a.C * a.nvox + run_base;
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

template <int LPV>
static cudaError_t launch_fwd_lpv(const UnprojArgs& a, int agg, dim3 grid, cudaStream_t st) {
  const bool aligned = (a.N2 % kRun) == 0;
  const dim3 block(kWarpsPerCta * 32);
#define LT_LAUNCH(AGG_)                                                      \
  do {                                                                       \
    if (aligned) k_unproject_fwd<LPV, AGG_, true><<<grid, block, 0, st>>>(a); \
    else k_unproject_fwd<LPV, AGG_, false><<<grid, block, 0, st>>>(a);        \
  } while (0)
  switch (agg) {
    case kAggSum: LT_LAUNCH(kAggSum); break;
    case kAggConf: LT_LAUNCH(kAggConf); break;
    case kAggMax: LT_LAUNCH(kAggMax); break;
    case kAggSoftmax:
      if (a.V <= kViewChunk) LT_LAUNCH(kAggSoftmax);
      else if (aligned) k_unproject_fwd_softmax_manyviews<LPV, true><<<grid, block, 0, st>>>(a);
      else k_unproject_fwd_softmax_manyviews<LPV, false><<<grid, block, 0, st>>>(a);
      break;
    default: return cudaErrorInvalidValue;
  }
#undef LT_LAUNCH
  return cudaGetLastError();
}

void fill_unproj_tiling(UnprojArgs& a, int lpv) {
  const int G = 32 / lpv;
  const int G1 = G < 4 ? G : 4, G0 = G / G1;
  const int T0 = 4 * G0, T1 = 2 * G1;
  a.nt0 = (a.N0 + T0 - 1) / T0;
  a.nt1 = (a.N1 + T1 - 1) / T1;
  a.nt2 = (a.N2 + kRun - 1) / kRun;
  a.ncb = a.Cp / (lpv * kCPL);
}

int unproj_lpv(int Cp) { return Cp >= 256 ? 32 : Cp / kCPL; }

cudaError_t launch_unproject_fwd(const float* feat_hwc, const float* proj, const float* coord,
                                 const float* conf, float* out, const UnprojShape& s, cudaStream_t st) {
  UnprojArgs a{};
  a.feat = feat_hwc; a.proj = proj; a.coord = coord; a.conf = conf; a.out = out;
  a.B = s.B; a.V = s.V; a.C = s.C; a.Cp = padded_channels(s.C); a.H = s.H; a.W = s.W;
  a.N0 = s.N0; a.N1 = s.N1; a.N2 = s.N2; a.align_corners = s.align_corners;
  a.nvox = (long long)s.N0 * s.N1 * s.N2;
  if (s.B == 0 || s.C == 0 || a.nvox == 0) return cudaSuccess;
  const int lpv = unproj_lpv(a.Cp);
  fill_unproj_tiling(a, lpv);
  const long long nblk = (long long)a.B * a.ncb * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  dim3 grid((unsigned)nblk);
  switch (lpv) {
    case 4: return launch_fwd_lpv<4>(a, s.agg, grid, st);
    case 8: return launch_fwd_lpv<8>(a, s.agg, grid, st);
    case 16: return launch_fwd_lpv<16>(a, s.agg, grid, st);
    case 32: return launch_fwd_lpv<32>(a, s.agg, grid, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace lt
