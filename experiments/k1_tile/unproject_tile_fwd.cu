// unproject_tile_fwd.cu — K1 forward, shared-memory-staged variant (see unproject_tile.cuh for the machinery).
//
// Replaces the B·V python iterations of mvn/utils/op.py:107-161 for the common shapes: any C, any V for sum/conf,
// V <= 4 for softmax/max (more views keep a running state: unproject_fwd.cu). N2 % ZT == 0.
// Algorithmic HBM bytes per batch item: F + K + O = V*H*W*Cp*4 + nvox*12 + C*nvox*4 (boxes are re-read from L2).
#include <stdlib.h>

#include "unproject_tile.cuh"

namespace lt {

// softmax over <= 4 views of one channel of one voxel: out = sum_v s_v e^{s_v} / sum_v e^{s_v} (op.py:154-158).
// The maximum is moved to the front with a 3-comparator network and its exponential (exactly 1) is never evaluated:
// 3 MUFU.EX2 + 1 MUFU.RCP per channel. Absent views carry kNoView (e -> 0).
__device__ __forceinline__ float tile_softmax_mix4(float s0, float s1, float s2, float s3) {
  const float hi1 = fmaxf(s0, s1), lo1 = fminf(s0, s1), hi2 = fmaxf(s2, s3), lo2 = fminf(s2, s3);
  const float top = fmaxf(hi1, hi2), mid = fminf(hi1, hi2);
  const float nt = -top * kLog2eF;
  const float e1 = fast_ex2(fmaf(lo1, kLog2eF, nt));
  const float e2 = fast_ex2(fmaf(lo2, kLog2eF, nt));
  const float e3 = fast_ex2(fmaf(mid, kLog2eF, nt));
  const float den = ((1.f + e1) + e2) + e3;
  float num = fmaf(lo1, e1, top);
  num = fmaf(lo2, e2, num);
  num = fmaf(mid, e3, num);
  return num * fast_rcp(den);
}

// bilinear blend in ATen's nw, ne, sw, se order (weights already permuted with the taps by phase A); two channels per
// FFMA2: each half is an IEEE fma, so the result is bit-identical to the scalar chain.
__device__ __forceinline__ void blend4(const float4& t00, const float4& t01, const float4& t10, const float4& t11,
                                       const float4& w, float (&s)[4]) {
  const f2 wx = pack2(w.x, w.x), wy = pack2(w.y, w.y), wz = pack2(w.z, w.z), ww = pack2(w.w, w.w);
  f2 lo = mul2(pack2(t00.x, t00.y), wx), hi = mul2(pack2(t00.z, t00.w), wx);
  lo = fma2(pack2(t01.x, t01.y), wy, lo); hi = fma2(pack2(t01.z, t01.w), wy, hi);
  lo = fma2(pack2(t10.x, t10.y), wz, lo); hi = fma2(pack2(t10.z, t10.w), wz, hi);
  lo = fma2(pack2(t11.x, t11.y), ww, lo); hi = fma2(pack2(t11.z, t11.w), ww, hi);
  unpack2(lo, s[0], s[1]);
  unpack2(hi, s[2], s[3]);
}

// aggregate the samples of one view chunk for 4 channels: sum / conf -> partial sum, max -> max (zeroed views take part
// with 0), softmax -> the final mix (V <= 4). Views >= nv: softmax gets kNoView, the others are simply left out.
template <int AGG>
__device__ __forceinline__ float4 agg_chunk(const float (&s)[kViewChunk][4], int nv, const float (&cw)[kViewChunk][4]) {
  float r[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (AGG == kAggSoftmax) {
      r[k] = tile_softmax_mix4(s[0][k], nv > 1 ? s[1][k] : kNoView, nv > 2 ? s[2][k] : kNoView, nv > 3 ? s[3][k] : kNoView);
    } else if (AGG == kAggMax) {
      r[k] = s[0][k];
#pragma unroll
      for (int vv = 1; vv < kViewChunk; ++vv) r[k] = (vv < nv) ? fmaxf(r[k], s[vv][k]) : r[k];
    } else {
      r[k] = 0.f;
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv) {  // views >= nv sampled nothing: s == 0
        if (AGG == kAggConf) r[k] = fmaf(cw[vv][k], s[vv][k], r[k]);
        else r[k] += s[vv][k];
      }
    }
  }
  return make_float4(r[0], r[1], r[2], r[3]);
}

// this lane's voxel (column col, layer z), 4 channels, all views of the chunk, boxes in shared memory. Branch-free per
// lane: invalid voxels read an aliased address with zero weights; a view without any valid voxel is skipped CTA-wide.
template <int ZT, int AGG>
__device__ __forceinline__ float4 layer_smem(const TileCtx<ZT>& c, int z, int col, unsigned lane_pool, int nv,
                                             const int (&pitch)[kViewChunk], const float (&cw)[kViewChunk][4]) {
  using SM = TileSmem<ZT>;
  const unsigned didx = (unsigned)(z * kTileRuns + col);
  float s[kViewChunk][4];
#pragma unroll
  for (int vv = 0; vv < kViewChunk; ++vv) {
#pragma unroll
    for (int k = 0; k < 4; ++k) s[vv][k] = 0.f;
    if (pitch[vv] > 0) {
      const uint2 o = lds64u(c.doff + (vv * SM::kVox + didx) * 8u);
      const float4 w = lds128(c.dwts + (vv * SM::kVox + didx) * 16u);
      const unsigned tA = lane_pool + o.x, tB = lane_pool + o.y;
      const float4 t00 = lds128(tA), t01 = lds128(tB);
      const float4 t10 = lds128(tA + pitch[vv]), t11 = lds128(tB + pitch[vv]);
      blend4(t00, t01, t10, t11, w, s[vv]);
    }
  }
  return agg_chunk<AGG>(s, nv, cw);
}

// same, taps read from the g16p planes in global memory (slab whose boxes did not fit the pool even as one layer).
// Kept out of line: it is the slow path and must not cost the fast path registers or instruction cache.
template <int ZT, int AGG>
__device__ __noinline__ float4 layer_global(unsigned doff, unsigned dwts, int z, int col, const float* lane_plane0,
                                            size_t view_stride, int row_bytes, int nv, const float* conf_lane,
                                            int conf_stride, int nch) {
  using SM = TileSmem<ZT>;
  const unsigned didx = (unsigned)(z * kTileRuns + col);
  float s[kViewChunk][4], cw[kViewChunk][4];
#pragma unroll
  for (int vv = 0; vv < kViewChunk; ++vv) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      s[vv][k] = 0.f;
      cw[vv][k] = (AGG == kAggConf && vv < nv && k < nch) ? __ldg(conf_lane + (size_t)vv * conf_stride + k) : 0.f;
    }
    if (vv < nv) {
      const uint2 o = lds64u(doff + (vv * SM::kVox + didx) * 8u);
      const float4 w = lds128(dwts + (vv * SM::kVox + didx) * 16u);
      const char* p = reinterpret_cast<const char*>(lane_plane0 + vv * view_stride);
      const float4 t00 = ldg128_nc(reinterpret_cast<const float*>(p + o.x));
      const float4 t01 = ldg128_nc(reinterpret_cast<const float*>(p + o.y));
      const float4 t10 = ldg128_nc(reinterpret_cast<const float*>(p + o.x + row_bytes));
      const float4 t11 = ldg128_nc(reinterpret_cast<const float*>(p + o.y + row_bytes));
      blend4(t00, t01, t10, t11, w, s[vv]);
    }
  }
  return agg_chunk<AGG>(s, nv, cw);
}

// MULTI = false: V <= 4, one view chunk; the CTA walks all its channel groups with ONE phase A and stores each slab
//                 as soon as it is aggregated.
// MULTI = true : V > 4 (sum / conf): one channel group per CTA (gpc == 1), partial sums stay in registers across chunks.
template <int ZT, int AGG, bool MULTI>
__global__ void __launch_bounds__(kTileThreads, 2) k_unproject_fwd_tile(const TileArgs a) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TileCtx<ZT> c = tile_prologue<ZT>(a, smem_raw);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, q = lane & 3;
  const int col = warp * 8 + g;
  const int i0 = c.i0_base + warp, i1 = c.i1_base + g;  // column col = (col>>3, col&7)
  const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
  const unsigned lane_pool = c.pool + q * 16;
  const int Wp = a.W + 2;
  const size_t plane = (size_t)(a.H + 2) * Wp * kGroupCh;
  const size_t view_stride = plane * a.G;
  float* const out_col = a.out + (size_t)c.T.b * a.C * a.nvox + ((long long)i0 * a.N1 + i1) * a.N2 + c.z_base;

  float val[ZT][4];
#pragma unroll
  for (int j = 0; j < ZT; ++j)
#pragma unroll
    for (int k = 0; k < 4; ++k) val[j][k] = 0.f;

  int zext = ZT;
#pragma unroll 1
  for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
    const int nv = min(kViewChunk, a.V - v0);
#pragma unroll 1
    for (int zs = 0; zs < ZT;) {
      const int fits = tile_stage<ZT, true>(a, c, v0, nv, zs, zext);
      if (!fits && zext > 1) { zext >>= 1; continue; }
      int pitch[kViewChunk];
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv) pitch[vv] = (c.ctl->box[vv][0] >= 0) ? c.ctl->box[vv][1] * kTexelBytes : 0;
      const int nbuf = c.ctl->nbuf, fill_bytes = c.ctl->total_bytes;
#pragma unroll 1
      for (int gi = 0; gi < a.gpc; ++gi) {
        const int cg = c.T.cg + gi;
        const int c_lane = cg * kGroupCh + q * 4;
        const int buf = gi % nbuf;
        if (fits) tile_wait<ZT>(c, buf);
        const unsigned lane_pool_g = lane_pool + (unsigned)(buf * fill_bytes);
        float cw[kViewChunk][4];
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            cw[vv][k] = (AGG == kAggConf && vv < nv && c_lane + k < a.C)
                            ? __ldg(a.conf + ((size_t)c.T.b * a.V + v0 + vv) * a.C + c_lane + k) : 0.f;
        if (fits) {
#pragma unroll
          for (int j = 0; j < ZT; ++j) {
            if (j >= zs && j < zs + zext) {
              const float4 r = layer_smem<ZT, AGG>(c, j, col, lane_pool_g, nv, pitch, cw);
              if (MULTI) { val[j][0] += r.x; val[j][1] += r.y; val[j][2] += r.z; val[j][3] += r.w; }
              else { val[j][0] = r.x; val[j][1] = r.y; val[j][2] = r.z; val[j][3] = r.w; }
            }
          }
        } else {
          const float* lane_plane0 = a.feat + ((size_t)c.T.b * a.V + v0) * view_stride + (size_t)cg * plane + q * 4;
          const float* conf_lane = (AGG == kAggConf) ? a.conf + ((size_t)c.T.b * a.V + v0) * a.C + c_lane : nullptr;
#pragma unroll
          for (int j = 0; j < ZT; ++j) {
            if (j >= zs && j < zs + zext) {
              const float4 r = layer_global<ZT, AGG>(c.doff, c.dwts, j, col, lane_plane0, view_stride, Wp * kTexelBytes,
                                                     nv, conf_lane, a.C, a.C - c_lane);
              if (MULTI) { val[j][0] += r.x; val[j][1] += r.y; val[j][2] += r.z; val[j][3] += r.w; }
              else { val[j][0] = r.x; val[j][1] = r.y; val[j][2] = r.z; val[j][3] = r.w; }
            }
          }
        }
        if (fits && gi + nbuf < a.gpc) {
          __syncthreads();  // every warp is done with this buffer
          tile_fill<ZT>(a, c, v0, cg + nbuf, buf);
        }
        if (!MULTI && run_ok) {
          // store this slab of this channel group: full vectors when the slab is the whole block
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (c_lane + k < a.C) {
              float* p = out_col + (size_t)(c_lane + k) * a.nvox;
              if (zext == ZT) {
                float v[ZT];
#pragma unroll
                for (int j = 0; j < ZT; ++j) v[j] = val[j][k];
                if (ZT == 8) stg256(p, v);
                else stg128_stream(p, make_float4(v[0], v[1], v[2], v[3]));
              } else {
#pragma unroll
                for (int j = 0; j < ZT; ++j)
                  if (j >= zs && j < zs + zext) p[j] = val[j][k];
              }
            }
          }
        }
      }
      zs += zext;
    }
  }

  if (MULTI && run_ok) {
    const int c_lane = c.T.cg * kGroupCh + q * 4;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (c_lane + k < a.C) {
        float* p = out_col + (size_t)(c_lane + k) * a.nvox;
        float v[ZT];
#pragma unroll
        for (int j = 0; j < ZT; ++j) v[j] = val[j][k];
        if (ZT == 8) stg256(p, v);
        else stg128_stream(p, make_float4(v[0], v[1], v[2], v[3]));
      }
    }
  }
}

// ---- host side ------------------------------------------------------------------------------------------------------
int tile_env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && *e) ? atoi(e) : dflt;
}

bool tile_path_supported(int agg, int V, int N2, int zt) {
  if (tile_env_int("LT_K1_TILE", 0) == 0) return false;
  if (N2 % zt != 0 || V > kMaxViewsSmem) return false;
  if ((agg == kAggSoftmax || agg == kAggMax) && V > kViewChunk) return false;
  return true;
}

template <int ZT>
size_t tile_smem_bytes(int occ, int* pool_texels) {
  // dynamic shared memory per CTA for `occ` CTAs per SM: 228 KB per SM, 1 KB reserved per CTA
  const size_t per_cta = (size_t)(228 * 1024) / occ - 1024;
  const size_t cap = per_cta < (size_t)(227 * 1024) ? per_cta : (size_t)(227 * 1024);
  const int pool = (int)((cap - TileSmem<ZT>::off_pool) / kTexelBytes);
  *pool_texels = pool;
  return TileSmem<ZT>::bytes(pool);
}

template <int ZT, int AGG, bool MULTI>
static cudaError_t launch_fwd_tile_k(TileArgs& a, cudaStream_t st) {
  const int occ = tile_env_int("LT_K1_TILE_OCC", 2);
  const size_t smem = tile_smem_bytes<ZT>(occ, &a.pool_texels);
  a.gpc = MULTI ? 1 : a.G;
  a.nt0 = (a.N0 + 7) / 8; a.nt1 = (a.N1 + 7) / 8; a.nt2 = a.N2 / ZT;
  const long long nblk = (long long)a.B * (a.G / a.gpc) * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaFuncSetAttribute(k_unproject_fwd_tile<ZT, AGG, MULTI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)smem);
  if (e != cudaSuccess) return e;
  k_unproject_fwd_tile<ZT, AGG, MULTI><<<(unsigned)nblk, kTileThreads, smem, st>>>(a);
  return cudaGetLastError();
}

template <int ZT>
static cudaError_t launch_fwd_tile_zt(TileArgs& a, int agg, cudaStream_t st) {
  const bool multi = a.V > kViewChunk;
  switch (agg) {
    case kAggSum: return multi ? launch_fwd_tile_k<ZT, kAggSum, true>(a, st) : launch_fwd_tile_k<ZT, kAggSum, false>(a, st);
    case kAggConf: return multi ? launch_fwd_tile_k<ZT, kAggConf, true>(a, st) : launch_fwd_tile_k<ZT, kAggConf, false>(a, st);
    case kAggMax: return launch_fwd_tile_k<ZT, kAggMax, false>(a, st);
    case kAggSoftmax: return launch_fwd_tile_k<ZT, kAggSoftmax, false>(a, st);
  }
  return cudaErrorInvalidValue;
}

int tile_pick_zt(int N2) {
  const int want = tile_env_int("LT_K1_TILE_ZT", 4);
  if (want == 4 && N2 % 4 == 0) return 4;
  if (N2 % 8 == 0) return 8;
  if (N2 % 4 == 0) return 4;
  return 0;
}

cudaError_t launch_unproject_fwd_tile(const float* feat, const float* proj, const float* coord, const float* conf,
                                      float* out, const UnprojShape& s, cudaStream_t st) {
  TileArgs a{};
  a.feat = feat; a.proj = proj; a.coord = coord; a.conf = conf; a.out = out;
  a.B = s.B; a.V = s.V; a.C = s.C; a.G = padded_channels(s.C) / kGroupCh; a.H = s.H; a.W = s.W;
  a.N0 = s.N0; a.N1 = s.N1; a.N2 = s.N2; a.align_corners = s.align_corners;
  a.nvox = (long long)s.N0 * s.N1 * s.N2;
  const int zt = tile_pick_zt(s.N2);
  if (zt == 8) return launch_fwd_tile_zt<8>(a, s.agg, st);
  if (zt == 4) return launch_fwd_tile_zt<4>(a, s.agg, st);
  return cudaErrorInvalidValue;
}

}  // namespace lt
