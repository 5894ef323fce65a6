// unproject_bwd_tile.cu — K1 backward, "tile" formulation: the scatter of unproject_bwd.cu with the reductions of a
// voxel tile MERGED ON CHIP before they leave the SM.
//
// Why: unproject_bwd.cu sends every bilinear tap of every (voxel, view) to L2 as its own 128-byte reduction: 5.4 GB of
// reduction payload per launch at cfg3 for a 79 MB gradient map (72x), and the L2 retires ~6.3 TB/s of that
// (profiles/r1c_microbench_red_paths.txt) -> 0.85 ms no matter how the SM issues them. Neighbouring voxels of a tile hit
// the same texels (depth collapses onto a pixel), so the taps are merged here first.
//
// How (one CTA = one T0 x T1 x RUN voxel tile of one batch item, views in chunks of <= 4):
//   A  per (voxel, view): projection + grid_sample index arithmetic exactly as the scatter kernel (make_taps), the
//      pixel CELL (x0, y0) the voxel falls into becomes a sort key relative to the tile's projected bounding box (from
//      its 8 corners), counted into a per-view shared-memory histogram (native integer ATOMS) which also hands out
//      the rank among equal keys;
//   B  per voxel, as in the scatter kernel: re-gather every view's sample (4 x LDG.128 per lane, 8 lanes = one
//      32-channel texel), form d out / d sample_v (mvn/utils/op.py:147-159 differentiated by hand, see
//      unproject_bwd.cu), and park d_v (128 bytes per voxel and view) in shared memory instead of scattering it;
//   S  exclusive scan of the histograms -> position of every (voxel, view) in CELL-SORTED order (counting sort; the
//      order only decides how much merges, never the result);
//   C  "row walk": every lane group takes a contiguous chunk of the sorted list and keeps the FOUR corner texels of
//      the current cell in registers (16 floats per lane). Next record in the same cell: accumulate. In the cell to the
//      right: flush the left column (2 reductions), slide the right column over. Anything else: flush all four.
//      Flushes are red.global.add.v4.f32 into the hwcp gradient map, exactly like the scatter kernel's taps, so races
//      between lane groups / CTAs stay harmless and no footprint bounding box has to fit anywhere.
// On the synthetic Human3.6M geometry an 8x8x4 tile needs 1 reduction per 2.6 taps (tools/sim in DESIGN.md §3),
// i.e. ~2 GB instead of 5.4 GB of L2 reduction payload per launch.
//
// Taken for the reference's channel count (C <= 32 -> Cp == 32), N2 % 8 == 0, 32-byte aligned grad_out, and sum/conf
// with any number of views or max/softmax with <= 4 views; everything else runs unproject_bwd.cu. LT_B200_K1BWD=scatter
// forces the old kernel, LT_B200_K1BWD_TILE=<n> picks a tile configuration (A/B runs).
#include <stdlib.h>

#include "unproject.cuh"

namespace lt {

namespace {

constexpr int kLPV = 8;        // lanes per voxel (Cp == 32)
constexpr int kG = 32 / kLPV;  // voxels per warp instruction
constexpr int kCp = 32;

template <int T0_, int T1_, int RUN_, int NW_>
struct BtCfg {
  static constexpr int T0 = T0_, T1 = T1_, RUN = RUN_, NW = NW_;
  static constexpr int TV = T0 * T1 * RUN;  // voxels per tile
  static constexpr int NRUNS = T0 * T1;
  static constexpr int NG = NW * kG;        // lane groups per CTA
  static constexpr int ROUNDS = NRUNS / NG;
  static_assert(NRUNS % NG == 0 && ROUNDS >= 1, "runs must tile the lane groups");
  static_assert(RUN == 4 || RUN == 8, "run = one 128- or 256-bit load per channel");
  static constexpr int NB = TV >= 256 ? 1024 : 512;  // histogram bins per view (keys beyond share the last bin)
  static constexpr int NT = NW * 32;
  static constexpr int CTAS = TV >= 256 ? 1 : 2;
  // shared-memory map (bytes)
  static constexpr size_t oStage = 0;                                               // float4 [4][TV][8]
  static constexpr size_t oW = oStage + (size_t)kViewChunk * TV * kCp * 4;          // float4 [4][TV]
  static constexpr size_t oOff = oW + (size_t)kViewChunk * TV * 16;                 // int    [4][TV]
  static constexpr size_t oKr = oOff + (size_t)kViewChunk * TV * 4;                 // int    [4][TV]  key | rank << 16
  static constexpr size_t oHist = oKr + (size_t)kViewChunk * TV * 4;                // int    [4][NB]
  static constexpr size_t oIdx = oHist + (size_t)kViewChunk * NB * 4;               // u16    [4][TV]
  static constexpr size_t oProj = oIdx + (size_t)kViewChunk * TV * 2;               // float4 [4*3]
  static constexpr size_t oBox = oProj + (size_t)kViewChunk * 3 * 16;               // int4   [4]
  static constexpr size_t oN = oBox + (size_t)kViewChunk * 16;                      // int    [4]
  static constexpr size_t kBytes = oN + 32;
};

// approximate heat-map position of a point (bounding box of the sort keys only; the taps come from make_taps)
__device__ __forceinline__ bool approx_ixy(float x, float y, float z, const float4& r0, const float4& r1,
                                           const float4& r2, int H, int W, int align_corners, float& ix, float& iy) {
  const float px = fmaf(x, r0.x, fmaf(y, r0.y, fmaf(z, r0.z, r0.w)));
  const float py = fmaf(x, r1.x, fmaf(y, r1.y, fmaf(z, r1.z, r1.w)));
  const float pz = fmaf(x, r2.x, fmaf(y, r2.y, fmaf(z, r2.z, r2.w)));
  const float u = px / pz, v = py / pz;
  const float gx = 2.f * (u / (float)H - 0.5f), gy = 2.f * (v / (float)W - 0.5f);
  if (align_corners) {
    ix = (gx + 1.f) * 0.5f * (float)(W - 1);
    iy = (gy + 1.f) * 0.5f * (float)(H - 1);
  } else {
    ix = ((gx + 1.f) * (float)W - 1.f) * 0.5f;
    iy = ((gy + 1.f) * (float)H - 1.f) * 0.5f;
  }
  return pz > 0.f && fabsf(ix) < 1.0e6f && fabsf(iy) < 1.0e6f;  // also false for nan
}

__device__ __forceinline__ bool nz4(const float4& v) { return v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f; }
__device__ __forceinline__ void red4(float* p, const float4& v) {
  if (nz4(v)) red_add_v4(p, v.x, v.y, v.z, v.w);
}
__device__ __forceinline__ void fma4(float4& acc, float w, const float4& d) {
  acc.x = fmaf(w, d.x, acc.x); acc.y = fmaf(w, d.y, acc.y); acc.z = fmaf(w, d.z, acc.z); acc.w = fmaf(w, d.w, acc.w);
}

template <class CF, int AGG>
__global__ void __launch_bounds__(CF::NT, CF::CTAS) k_unproject_bwd_tile(const UnprojArgs a) {
  constexpr int TV = CF::TV, NB = CF::NB, NT = CF::NT, RUN = CF::RUN, NG = CF::NG;
  extern __shared__ __align__(128) unsigned char smem[];
  float4* s_stage = reinterpret_cast<float4*>(smem + CF::oStage);
  float4* s_w = reinterpret_cast<float4*>(smem + CF::oW);
  int* s_off = reinterpret_cast<int*>(smem + CF::oOff);
  int* s_kr = reinterpret_cast<int*>(smem + CF::oKr);
  int* s_hist = reinterpret_cast<int*>(smem + CF::oHist);
  unsigned short* s_idx = reinterpret_cast<unsigned short*>(smem + CF::oIdx);
  float4* s_proj = reinterpret_cast<float4*>(smem + CF::oProj);
  int4* s_box = reinterpret_cast<int4*>(smem + CF::oBox);
  int* s_n = reinterpret_cast<int*>(smem + CF::oN);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane / kLPV, q = lane % kLPV;
  int t = blockIdx.x;
  const int t2 = t % a.nt2; t /= a.nt2;
  const int t1 = t % a.nt1; t /= a.nt1;
  const int t0 = t % a.nt0;
  const int b = t / a.nt0;
  const int tile_lin = (t0 * a.nt1 + t1) * a.nt2 + t2;
  const int ntiles = a.nt0 * a.nt1 * a.nt2;

  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = (a.W + 2) * kCp;
  const size_t viewStride = (size_t)(a.H + 2) * (a.W + 2) * kCp;
  const int c_lane = q * kCPL;
  const float* fb = a.feat + (size_t)b * a.V * viewStride + c_lane;
  float* gb = a.gfeat + (size_t)b * a.V * viewStride + c_lane;
  const float* coord_b = a.coord + (size_t)b * a.nvox * 3;
  const int z0 = t2 * RUN;

#pragma unroll 1
  for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
    const int nv = min(kViewChunk, a.V - v0);
    // ---------------- P0: histograms, projection rows, key bounding boxes ----------------
    for (int i = tid; i < kViewChunk * NB / 4; i += NT) reinterpret_cast<int4*>(s_hist)[i] = make_int4(0, 0, 0, 0);
    if (tid < nv * 3) s_proj[tid] = __ldg(reinterpret_cast<const float4*>(a.proj) + ((size_t)b * a.V + v0) * 3 + tid);
    __syncthreads();
    if (warp == 0) {
      const int vv = lane >> 3, c = lane & 7;
      const int i0 = min(t0 * CF::T0 + ((c & 1) ? CF::T0 - 1 : 0), a.N0 - 1);
      const int i1 = min(t1 * CF::T1 + ((c & 2) ? CF::T1 - 1 : 0), a.N1 - 1);
      const int iz = min(z0 + ((c & 4) ? RUN - 1 : 0), a.N2 - 1);
      const float* cp = coord_b + (((long long)i0 * a.N1 + i1) * a.N2 + iz) * 3;
      float ix = 0.f, iy = 0.f;
      bool ok = false;
      if (vv < nv) ok = approx_ixy(__ldg(cp), __ldg(cp + 1), __ldg(cp + 2), s_proj[vv * 3], s_proj[vv * 3 + 1],
                                   s_proj[vv * 3 + 2], a.H, a.W, a.align_corners, ix, iy);
      float xmin = ix, xmax = ix, ymin = iy;
      int all_ok = ok ? 1 : 0;
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) {
        xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
        xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
        ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
        all_ok &= __shfl_xor_sync(0xffffffffu, all_ok, o);
      }
      if (c == 0) {
        int bx = -1, by = -1, bw = a.W + 1;
        if (all_ok) {
          bx = max(-1, min(a.W - 1, (int)floorf(xmin) - 1));
          by = max(-1, min(a.H - 1, (int)floorf(ymin) - 1));
          const int ex = max(bx, min(a.W - 1, (int)floorf(xmax) + 1));
          bw = ex - bx + 1;
        }
        s_box[vv] = make_int4(bx, by, bw, 0);
      }
    }
    __syncthreads();

    float cw[kViewChunk][kCPL];   // conf of this chunk            (conf only)
    float gcw[kViewChunk][kCPL];  // running d conf of this chunk  (conf only)
    if (AGG == kAggConf) {
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          const int ch = c_lane + k;
          cw[vv][k] = (vv < nv && ch < a.C) ? __ldg(a.conf + ((size_t)b * a.V + v0 + vv) * a.C + ch) : 0.f;
          gcw[vv][k] = 0.f;
        }
    }

#pragma unroll 1
    for (int r = 0; r < CF::ROUNDS; ++r) {
      const int run_w = r * NG + warp * kG;  // first of this warp's kG runs
      // ---------------- A: descriptors + sort keys of this warp's kG*RUN voxels x nv views ----------------
      for (int it = lane; it < kG * RUN * nv; it += 32) {
        const int vl = it % (kG * RUN), vv = it / (kG * RUN);
        const int run = run_w + vl / RUN, j = vl % RUN;
        const int vi0 = t0 * CF::T0 + run / CF::T1, vi1 = t1 * CF::T1 + run % CF::T1, vz = z0 + j;
        const int vox = run * RUN + j;
        int off = -1, kr = 0;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (vi0 < a.N0 && vi1 < a.N1 && vz < a.N2) {
          const float* cp = coord_b + (((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
          const float x = __ldg(cp), y = __ldg(cp + 1), z = __ldg(cp + 2);
          int x0, y0;
          if (make_taps(x, y, z, s_proj[vv * 3], s_proj[vv * 3 + 1], s_proj[vv * 3 + 2], a.H, a.W, invH, invW,
                        a.align_corners, x0, y0, w)) {
            off = ((y0 + 1) * (a.W + 2) + (x0 + 1)) * kCp;
            const int4 bx = s_box[vv];
            const int kx = min(max(x0 - bx.x, 0), bx.z - 1), ky = max(y0 - bx.y, 0);
            const int key = min(ky * bx.z + kx, NB - 1);
            const int rank = atomicAdd(&s_hist[vv * NB + key], 1);
            kr = key | (rank << 16);
          }
        }
        s_w[vv * TV + vox] = w;
        s_off[vv * TV + vox] = off;
        s_kr[vv * TV + vox] = kr;
      }
      __syncwarp();

      // ---------------- B: samples of every view, d out / d sample_v, parked in shared memory ----------------
      const int run = run_w + g;
      const int i0 = t0 * CF::T0 + run / CF::T1, i1 = t1 * CF::T1 + run % CF::T1;
      const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
      {
        // upstream gradient of this lane's 4 channels x RUN voxels, parked in the slots view 0's d will overwrite
        float go[kCPL][RUN];
        const float* gp = a.gout + (size_t)b * a.C * a.nvox + ((long long)i0 * a.N1 + i1) * a.N2 + z0;
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          const int c = c_lane + k;
          if (run_ok && c < a.C) {
            if (RUN == 8) {
              const f8 tt = ldg256_stream(gp + (size_t)c * a.nvox);
#pragma unroll
              for (int j = 0; j < RUN; ++j) go[k][j] = tt.v[j];
            } else {
              const float4 tt = ldg128_stream(gp + (size_t)c * a.nvox);
              go[k][0] = tt.x; go[k][1] = tt.y; go[k][2] = tt.z; go[k][3] = tt.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < RUN; ++j) go[k][j] = 0.f;
          }
        }
#pragma unroll
        for (int j = 0; j < RUN; ++j)
          s_stage[(run * RUN + j) * kLPV + q] = make_float4(go[0][j], go[1][j], go[2][j], go[3][j]);
      }

#pragma unroll 2
      for (int j = 0; j < RUN; ++j) {
        const int vox = run * RUN + j;
        const float4 gj = s_stage[vox * kLPV + q];
        const float go[kCPL] = {gj.x, gj.y, gj.z, gj.w};
        if (AGG == kAggSum || AGG == kAggConf) {
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            if (vv < nv) {
              const int off = s_off[vv * TV + vox];
              if (off >= 0) {
                float d[kCPL];
                if (AGG == kAggConf) {
                  float s[kCPL];
                  gather4(fb + (size_t)(v0 + vv) * viewStride, off, s_w[vv * TV + vox], kCp, rowStride, s);
#pragma unroll
                  for (int k = 0; k < kCPL; ++k) {
                    gcw[vv][k] = fmaf(go[k], s[k], gcw[vv][k]);
                    d[k] = cw[vv][k] * go[k];
                  }
                } else {
#pragma unroll
                  for (int k = 0; k < kCPL; ++k) d[k] = go[k];
                }
                s_stage[(vv * TV + vox) * kLPV + q] = make_float4(d[0], d[1], d[2], d[3]);
              }
            }
          }
        } else {
          float s[kViewChunk][kCPL];
          int offs[kViewChunk];
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            const float fill = (vv < nv) ? 0.f : kNoView;
#pragma unroll
            for (int k = 0; k < kCPL; ++k) s[vv][k] = fill;
            offs[vv] = -1;
            if (vv < nv) {
              offs[vv] = s_off[vv * TV + vox];
              if (offs[vv] >= 0)
                gather4(fb + (size_t)(v0 + vv) * viewStride, offs[vv], s_w[vv * TV + vox], kCp, rowStride, s[vv]);
            }
          }
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
              s_stage[(vv * TV + vox) * kLPV + q] = make_float4(d[vv][0], d[vv][1], d[vv][2], d[vv][3]);
        }
      }
    }
    __syncthreads();

    // ---------------- S: exclusive scan of each view's histogram (warp vv), then the sorted index lists ----------------
    if (warp < nv) {
      int* h = s_hist + warp * NB;
      int carry = 0;
#pragma unroll 4
      for (int i = 0; i < NB / 32; ++i) {
        const int v = h[i * 32 + lane];
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int n = __shfl_up_sync(0xffffffffu, incl, o);
          if (lane >= o) incl += n;
        }
        h[i * 32 + lane] = carry + incl - v;
        carry += __shfl_sync(0xffffffffu, incl, 31);
      }
      if (lane == 0) s_n[warp] = carry;
    }
    __syncthreads();
    for (int it = tid; it < nv * TV; it += NT) {
      const int off = s_off[it];
      if (off >= 0) {
        const int vv = it / TV, vox = it % TV, kr = s_kr[it];
        const int pos = s_hist[vv * NB + (kr & 0xffff)] + (kr >> 16);
        s_idx[vv * TV + pos] = (unsigned short)vox;
      }
    }
    // record r of the concatenated sorted lists belongs to view #(i in 1..3 : r >= b_i)
    const int b1 = s_n[0], b2 = b1 + (nv > 1 ? s_n[1] : 0), b3 = b2 + (nv > 2 ? s_n[2] : 0);
    const int n_total = b3 + (nv > 3 ? s_n[3] : 0);
    __syncthreads();

    // ---------------- C: row walk over the cell-sorted records ----------------
    {
      const int chunk = (n_total + NG - 1) / NG;
      const int gid = warp * kG + g;
      int rec = gid * chunk;
      const int rec_end = min(n_total, rec + chunk);
      float4 a00 = make_float4(0.f, 0.f, 0.f, 0.f), a01 = a00, a10 = a00, a11 = a00;
      int cur = -1, cur_v = 0;
      float* gv = gb;
      for (; rec < rec_end; ++rec) {
        const int vv = (rec >= b1) + (rec >= b2) + (rec >= b3);
        const int vox = s_idx[vv * TV + rec - (vv == 0 ? 0 : (vv == 1 ? b1 : (vv == 2 ? b2 : b3)))];
        const int off = s_off[vv * TV + vox];
        const float4 w = s_w[vv * TV + vox];
        const float4 d = s_stage[(vv * TV + vox) * kLPV + q];
        if (off != cur || vv != cur_v) {
          const bool shift = (vv == cur_v) && (off == cur + kCp);
          if (cur >= 0) {
            red4(gv + cur, a00);
            red4(gv + cur + rowStride, a10);
            if (!shift) {
              red4(gv + cur + kCp, a01);
              red4(gv + cur + rowStride + kCp, a11);
            }
          }
          if (shift && cur >= 0) {
            a00 = a01; a10 = a11;
          } else {
            a00 = make_float4(0.f, 0.f, 0.f, 0.f); a10 = a00;
          }
          a01 = make_float4(0.f, 0.f, 0.f, 0.f); a11 = a01;
          cur = off; cur_v = vv;
          gv = gb + (size_t)(v0 + vv) * viewStride;
        }
        fma4(a00, w.x, d); fma4(a01, w.y, d); fma4(a10, w.z, d); fma4(a11, w.w, d);
      }
      if (cur >= 0) {
        red4(gv + cur, a00);
        red4(gv + cur + kCp, a01);
        red4(gv + cur + rowStride, a10);
        red4(gv + cur + rowStride + kCp, a11);
      }
    }

    // ---------------- grad_conf: fixed-order reduction of the CTA, one partial row per (item, tile) ----------------
    if (AGG == kAggConf && a.gconf_part != nullptr) {
      __syncthreads();  // stage is dead: reuse it as [warp][vv][32 channels]
      float* s_gc = reinterpret_cast<float*>(s_stage);
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) {
          float v = gcw[vv][k];
          v += __shfl_xor_sync(0xffffffffu, v, 8);
          v += __shfl_xor_sync(0xffffffffu, v, 16);
          if (g == 0) s_gc[(warp * kViewChunk + vv) * kCp + c_lane + k] = v;
        }
      __syncthreads();
      if (tid < nv * kCp) {
        const int vv = tid / kCp, c = tid % kCp;
        float s = 0.f;
        for (int wi = 0; wi < CF::NW; ++wi) s += s_gc[(wi * kViewChunk + vv) * kCp + c];
        if (c < a.C) a.gconf_part[(((size_t)b * ntiles + tile_lin) * a.V + v0 + vv) * a.C + c] = s;
      }
    }
    __syncthreads();  // next view chunk rewrites every table
  }
}

// grad_conf[b, v, c] = sum over tiles of the per-CTA partials, in tile order (deterministic).
// grid (ceil(VC/32), B), block (32, 8): lane = (v, c) column, the 8 rows split the tiles in a fixed interleave.
__global__ void __launch_bounds__(256) k_conf_reduce_tiles(const float* __restrict__ part, float* __restrict__ gconf,
                                                           int VC, int ntiles) {
  __shared__ float s[8][33];
  const int col = blockIdx.x * 32 + threadIdx.x, b = blockIdx.y;
  float acc = 0.f;
  if (col < VC)
    for (int t = threadIdx.y; t < ntiles; t += 8) acc += part[((size_t)b * ntiles + t) * VC + col];
  s[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && col < VC) {
    float r = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) r += s[k][threadIdx.x];
    gconf[(size_t)b * VC + col] = r;
  }
}

using CfgA = BtCfg<8, 8, 4, 16>;  // 256 voxels, 512 threads, 1 CTA/SM
using CfgB = BtCfg<8, 4, 4, 8>;   // 128 voxels, 256 threads, 2 CTAs/SM
using CfgC = BtCfg<8, 4, 8, 8>;   // 256 voxels as 8-voxel runs (256-bit grad_out loads), 256 threads, 1 CTA/SM
using CfgD = BtCfg<4, 4, 8, 4>;   // 128 voxels as 8-voxel runs, 128 threads, 2 CTAs/SM

int tile_cfg_id() {
  const char* e = getenv("LT_B200_K1BWD_TILE");
  const int id = e ? atoi(e) : 0;
  return (id >= 0 && id <= 3) ? id : 0;
}

template <class CF>
void tile_dims(const UnprojShape& s, int& nt0, int& nt1, int& nt2) {
  nt0 = (s.N0 + CF::T0 - 1) / CF::T0;
  nt1 = (s.N1 + CF::T1 - 1) / CF::T1;
  nt2 = (s.N2 + CF::RUN - 1) / CF::RUN;
}

long long tile_count(const UnprojShape& s) {
  int n0, n1, n2;
  switch (tile_cfg_id()) {
    case 1: tile_dims<CfgB>(s, n0, n1, n2); break;
    case 2: tile_dims<CfgC>(s, n0, n1, n2); break;
    case 3: tile_dims<CfgD>(s, n0, n1, n2); break;
    default: tile_dims<CfgA>(s, n0, n1, n2); break;
  }
  return (long long)n0 * n1 * n2;
}

template <class CF, int AGG>
cudaError_t launch_one(UnprojArgs& a, const UnprojShape& s, cudaStream_t st) {
  tile_dims<CF>(s, a.nt0, a.nt1, a.nt2);
  a.ncb = 1;
  const long long nblk = (long long)a.B * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  LT_SMEM_OPT_IN_ONCE((k_unproject_bwd_tile<CF, AGG>), CF::kBytes);
  k_unproject_bwd_tile<CF, AGG><<<(unsigned)nblk, CF::NT, CF::kBytes, st>>>(a);
  return cudaGetLastError();
}

template <class CF>
cudaError_t launch_agg(UnprojArgs& a, const UnprojShape& s, cudaStream_t st) {
  switch (s.agg) {
    case kAggSum: return launch_one<CF, kAggSum>(a, s, st);
    case kAggConf: return launch_one<CF, kAggConf>(a, s, st);
    case kAggMax: return launch_one<CF, kAggMax>(a, s, st);
    case kAggSoftmax: return launch_one<CF, kAggSoftmax>(a, s, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

bool unproject_bwd_tile_eligible(const UnprojShape& s, const float* grad_out) {
  const char* e = getenv("LT_B200_K1BWD");
  if (e && e[0] == 's') return false;  // "scatter"
  if (padded_channels(s.C) != kCp) return false;
  if (s.N2 % 8 != 0 || (reinterpret_cast<uintptr_t>(grad_out) & 31u) != 0) return false;
  if ((s.agg == kAggMax || s.agg == kAggSoftmax) && s.V > kViewChunk) return false;
  if (s.H + 2 > 2047 || s.W + 2 > 2047) return false;  // texel offsets stay far inside int32
  return true;
}

size_t unproject_bwd_tile_ws_bytes(const UnprojShape& s) {
  return (size_t)s.B * (size_t)tile_count(s) * s.V * s.C * sizeof(float);
}

cudaError_t launch_unproject_bwd_tile(UnprojArgs& a, const UnprojShape& s, float* grad_conf, cudaStream_t st) {
  cudaError_t e;
  switch (tile_cfg_id()) {
    case 1: e = launch_agg<CfgB>(a, s, st); break;
    case 2: e = launch_agg<CfgC>(a, s, st); break;
    case 3: e = launch_agg<CfgD>(a, s, st); break;
    default: e = launch_agg<CfgA>(a, s, st); break;
  }
  if (e != cudaSuccess) return e;
  if (a.gconf_part != nullptr) {
    const int VC = s.V * s.C;
    k_conf_reduce_tiles<<<dim3((VC + 31) / 32, s.B), dim3(32, 8), 0, st>>>(a.gconf_part, grad_conf, VC,
                                                                          a.nt0 * a.nt1 * a.nt2);
    e = cudaGetLastError();
  }
  return e;
}

}  // namespace lt
