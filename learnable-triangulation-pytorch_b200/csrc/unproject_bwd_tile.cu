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
  static constexpr int NB = 512;             // histogram bins per view (keys beyond share the last bin)
  static constexpr int PER = NB / 32;        // bins per lane in the scan
  static constexpr int HS = 32 * (PER + 1);  // padded pitch of one view's histogram
  __host__ __device__ static constexpr int hslot(int key) { return (key / PER) * (PER + 1) + key % PER; }
  static constexpr int NT = NW * 32;
  static constexpr int CTAS = TV >= 256 ? 1 : 2;
  // shared-memory map (bytes). "rec" arrays hold the (voxel, view) records of all views in CELL-SORTED order, view after
  // view; the voxel-order weights and key|rank words are only needed until the records are written and live in the
  // stage (which is not touched before that).
  static constexpr size_t oStage = 0;                                               // float4 [4][TV][8]  s, then d
  static constexpr size_t oRecW = oStage + (size_t)kViewChunk * TV * kCp * 4;       // float4 [4*TV]  tap weights
  static constexpr size_t oRecM = oRecW + (size_t)kViewChunk * TV * 16;             // int2   [4*TV]  (off | view, slot)
  static constexpr size_t oOff = oRecM + (size_t)kViewChunk * TV * 8;               // int    [4][TV] voxel order
  static constexpr size_t oHist = oOff + (size_t)kViewChunk * TV * 4;               // int    [4][HS]
  static constexpr size_t oProj = oHist + (size_t)kViewChunk * HS * 4;              // float4 [4*3]
  static constexpr size_t oBox = oProj + (size_t)kViewChunk * 3 * 16;               // int4   [4]
  static constexpr size_t oN = oBox + (size_t)kViewChunk * 16;                      // int    [4]
  static constexpr size_t kBytes = oN + 32;
  static constexpr size_t oW = oStage;                                              // float4 [4][TV] voxel order (temp)
  static constexpr size_t oKr = oStage + (size_t)kViewChunk * TV * 16;              // int    [4][TV] key | rank << 16
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

__device__ __forceinline__ void red4(float* p, const float4& v) { red_add_v4(p, v.x, v.y, v.z, v.w); }

// softmax aggregation over <= 4 views, derivative w.r.t. every view's sample, two channels per call:
//   out = sum_v s_v p_v, p = softmax_v(s)  =>  d s_v = g * p_v * (1 + s_v - out)        (mvn/utils/op.py:154-159)
// absent views carry kNoView (e -> 0 -> d = 0). Same operation order as the scalar form in unproject_bwd.cu.
__device__ __forceinline__ void softmax_bwd4_x2(f2 s0, f2 s1, f2 s2, f2 s3, f2 g, f2 (&d)[4]) {
  float a0, b0, a1, b1, a2, b2, a3, b3;
  unpack2(s0, a0, b0); unpack2(s1, a1, b1); unpack2(s2, a2, b2); unpack2(s3, a3, b3);
  const float ma = fmaxf(fmaxf(a0, a1), fmaxf(a2, a3)), mb = fmaxf(fmaxf(b0, b1), fmaxf(b2, b3));
  const f2 L2 = pack2(kLog2eF, kLog2eF);
  const f2 nm = mul2(pack2(-ma, -mb), L2);
  const f2 sv[4] = {s0, s1, s2, s3};
  f2 e[4];
  f2 den = pack2(0.f, 0.f), num = pack2(0.f, 0.f);
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    float ta, tb;
    unpack2(fma2(sv[v], L2, nm), ta, tb);
    e[v] = pack2(fast_ex2(ta), fast_ex2(tb));
    den = add2(den, e[v]);
    num = fma2(sv[v], e[v], num);
  }
  float da, db;
  unpack2(den, da, db);
  const f2 inv = pack2(fast_rcp(da), fast_rcp(db));
  const f2 one = pack2(1.f, 1.f), neg = pack2(-1.f, -1.f);
  const f2 o1 = fma2(mul2(num, inv), neg, one);  // 1 - out
  const f2 gi = mul2(g, inv);
#pragma unroll
  for (int v = 0; v < 4; ++v) d[v] = mul2(mul2(gi, e[v]), add2(o1, sv[v]));
}
__device__ __forceinline__ void fma4(float4& acc, float w, const float4& d) {
  acc.x = fmaf(w, d.x, acc.x); acc.y = fmaf(w, d.y, acc.y); acc.z = fmaf(w, d.z, acc.z); acc.w = fmaf(w, d.w, acc.w);
}

template <bool SPEC>
__device__ __forceinline__ void stage_spec_if(const KArgs<SPEC>&, int, float*) {}
template <>
__device__ __forceinline__ void stage_spec_if<true>(const KArgs<true>& a, int b, float* s) { stage_unproj_spec(a, b, s); }

template <class CF, int AGG, bool SPEC>
__global__ void __launch_bounds__(CF::NT, CF::CTAS) k_unproject_bwd_tile(const KArgs<SPEC> a) {
  constexpr int TV = CF::TV, NB = CF::NB, NT = CF::NT, RUN = CF::RUN, NG = CF::NG;
  extern __shared__ __align__(128) unsigned char smem[];
  float4* s_stage = reinterpret_cast<float4*>(smem + CF::oStage);
  float4* s_w = reinterpret_cast<float4*>(smem + CF::oW);
  int* s_off = reinterpret_cast<int*>(smem + CF::oOff);
  int* s_kr = reinterpret_cast<int*>(smem + CF::oKr);
  int* s_hist = reinterpret_cast<int*>(smem + CF::oHist);
  float4* s_rec_w = reinterpret_cast<float4*>(smem + CF::oRecW);
  int2* s_rec_m = reinterpret_cast<int2*>(smem + CF::oRecM);
  float4* s_proj = reinterpret_cast<float4*>(smem + CF::oProj);
  int4* s_box = reinterpret_cast<int4*>(smem + CF::oBox);
  int* s_n = reinterpret_cast<int*>(smem + CF::oN);

  __shared__ float s_spec_mem[SPEC ? kCoordSpec : 1];
  const float* s_spec = s_spec_mem;
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
  const int viewStrideI = (a.H + 2) * (a.W + 2) * kCp;
  const int c_lane = q * kCPL;
  const float* fb = a.feat + (size_t)b * a.V * viewStride + c_lane;
  float* gb = a.gfeat + (size_t)b * a.V * viewStride + c_lane;
  const float* coord_b = a.coord + (size_t)b * a.nvox * 3;
  const int z0 = t2 * RUN;

#pragma unroll 1
  for (int v0 = 0; v0 < a.V; v0 += kViewChunk) {
    const int nv = min(kViewChunk, a.V - v0);
    // ---------------- P0: histograms, projection rows, key bounding boxes ----------------
    for (int i = tid; i < kViewChunk * CF::HS / 4; i += NT) reinterpret_cast<int4*>(s_hist)[i] = make_int4(0, 0, 0, 0);
    if (tid < nv * 3) s_proj[tid] = __ldg(reinterpret_cast<const float4*>(a.proj) + ((size_t)b * a.V + v0) * 3 + tid);
    stage_spec_if<SPEC>(a, b, s_spec_mem);
    __syncthreads();
    if (warp == 0) {
      const int vv = lane >> 3, c = lane & 7;
      const int i0 = min(t0 * CF::T0 + ((c & 1) ? CF::T0 - 1 : 0), a.N0 - 1);
      const int i1 = min(t1 * CF::T1 + ((c & 2) ? CF::T1 - 1 : 0), a.N1 - 1);
      const int iz = min(z0 + ((c & 4) ? RUN - 1 : 0), a.N2 - 1);
      float cx, cy, cz;
      voxel_xyz<SPEC>(a, coord_b, s_spec, i0, i1, iz, cx, cy, cz);
      float ix = 0.f, iy = 0.f;
      bool ok = false;
      if (vv < nv) ok = approx_ixy(cx, cy, cz, s_proj[vv * 3], s_proj[vv * 3 + 1], s_proj[vv * 3 + 2], a.H, a.W,
                                   a.align_corners, ix, iy);
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

    // ---------------- A: descriptors + sort keys of every (voxel, view) of the tile ----------------
    // thread <-> voxel in memory order (coalesced coordinate loads); the rank among equal keys comes from the atomic
    for (int it = tid; it < nv * TV; it += NT) {
      const int vox = it % TV, vv = it / TV;
      const int run = vox / RUN, j = vox % RUN;
      const int vi0 = t0 * CF::T0 + run / CF::T1, vi1 = t1 * CF::T1 + run % CF::T1, vz = z0 + j;
      int off = -1, kr = 0;
      float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
      if (vi0 < a.N0 && vi1 < a.N1 && vz < a.N2) {
        float x, y, z;
        voxel_xyz<SPEC>(a, coord_b, s_spec, vi0, vi1, vz, x, y, z);
        int x0, y0;
        if (make_taps(x, y, z, s_proj[vv * 3], s_proj[vv * 3 + 1], s_proj[vv * 3 + 2], a.H, a.W, invH, invW,
                      a.align_corners, x0, y0, w)) {
          off = ((y0 + 1) * (a.W + 2) + (x0 + 1)) * kCp;
          const int4 bx = s_box[vv];
          const int kx = min(max(x0 - bx.x, 0), bx.z - 1), ky = max(y0 - bx.y, 0);
          const int key = min(ky * bx.z + kx, NB - 1);
          const int rank = atomicAdd(&s_hist[vv * CF::HS + CF::hslot(key)], 1);
          kr = key | (rank << 16);
        }
      }
      s_w[it] = w;
      s_off[it] = off;
      s_kr[it] = kr;
    }
    __syncthreads();

    // ---------------- S: exclusive scan of each view's histogram (warp vv; a lane owns PER consecutive bins, the
    // padded pitch keeps both this walk and the random atomics above free of bank conflicts) ----------------
    if (warp < nv) {
      int* h = s_hist + warp * CF::HS + lane * (CF::PER + 1);
      int vals[CF::PER];
      int sum = 0;
#pragma unroll
      for (int i = 0; i < CF::PER; ++i) { vals[i] = h[i]; }
#pragma unroll
      for (int i = 0; i < CF::PER; ++i) { const int t_ = vals[i]; vals[i] = sum; sum += t_; }
      int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += n;
      }
      const int excl = incl - sum;
#pragma unroll
      for (int i = 0; i < CF::PER; ++i) h[i] = vals[i] + excl;
      if (lane == 31) s_n[warp] = incl;
    }
    __syncthreads();
    // records in cell-sorted order, view after view: (texel offset | view, stage slot) + the four tap weights
    const int b1 = s_n[0], b2 = b1 + (nv > 1 ? s_n[1] : 0), b3 = b2 + (nv > 2 ? s_n[2] : 0);
    const int n_total = b3 + (nv > 3 ? s_n[3] : 0);
    for (int it = tid; it < nv * TV; it += NT) {
      const int off = s_off[it];
      if (off >= 0) {
        const int vv = it / TV, kr = s_kr[it];
        const int pos = (vv == 0 ? 0 : (vv == 1 ? b1 : (vv == 2 ? b2 : b3))) +
                        s_hist[vv * CF::HS + CF::hslot(kr & 0xffff)] + (kr >> 16);
        // offset relative to the chunk's first map (view stride folded in: one 32-bit index per tap, see build_descs);
        // a multiple of Cp = 32, so the low bits are free for the view
        s_rec_m[pos] = make_int2((off + vv * viewStrideI) | vv, it);
        s_rec_w[pos] = s_w[it];
      }
    }
    __syncthreads();
    const int chunk = (n_total + NG - 1) / NG;
    const int gid = warp * kG + g;
    const int rec_begin = min(n_total, gid * chunk), rec_end = min(n_total, rec_begin + chunk);
    const float* fbv = opaque_ptr(fb + (size_t)v0 * viewStride);   // opaque 64-bit pairs: see unproject.cuh
    const float* fbv1 = opaque_ptr(fbv + rowStride);
    float* gbv = opaque_ptr(gb + (size_t)v0 * viewStride);
    float* gbv1 = opaque_ptr(gbv + rowStride);

    // ---------------- G: samples, gathered in CELL-SORTED order: consecutive records read the same or neighbouring
    // texels, so all but the first touch of a texel hit L1 ----------------
    if (AGG != kAggSum) {
      int rec = rec_begin;
#pragma unroll 1
      for (; rec + 1 < rec_end; rec += 2) {
        const int2 ma = s_rec_m[rec], mb = s_rec_m[rec + 1];
        float sa[kCPL], sb[kCPL];
        gather4_rows<kCp>(fbv, fbv1, ma.x & ~3, s_rec_w[rec], kCp, sa);
        gather4_rows<kCp>(fbv, fbv1, mb.x & ~3, s_rec_w[rec + 1], kCp, sb);
        s_stage[ma.y * kLPV + q] = make_float4(sa[0], sa[1], sa[2], sa[3]);
        s_stage[mb.y * kLPV + q] = make_float4(sb[0], sb[1], sb[2], sb[3]);
      }
      if (rec < rec_end) {
        const int2 ma = s_rec_m[rec];
        float sa[kCPL];
        gather4_rows<kCp>(fbv, fbv1, ma.x & ~3, s_rec_w[rec], kCp, sa);
        s_stage[ma.y * kLPV + q] = make_float4(sa[0], sa[1], sa[2], sa[3]);
      }
      __syncthreads();
    }

    // ---------------- D: d out / d sample_v per voxel (mvn/utils/op.py:147-159 differentiated by hand), in place ----
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
      const int run = r * NG + gid;
      const int i0 = t0 * CF::T0 + run / CF::T1, i1 = t1 * CF::T1 + run % CF::T1;
      const bool run_ok = (i0 < a.N0) && (i1 < a.N1);
      float go[kCPL][RUN];  // upstream gradient of this lane's 4 channels x RUN voxels
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
      for (int j = 0; j < RUN; ++j) {
        const int vox = run * RUN + j;
        float4 s4[kViewChunk];
        bool ok[kViewChunk];
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv) {
          const float fill = (vv < nv) ? 0.f : kNoView;  // zeroed views take part with 0, absent ones not at all
          s4[vv] = make_float4(fill, fill, fill, fill);
          ok[vv] = (vv < nv) && (s_off[vv * TV + vox] >= 0);
          if (AGG != kAggSum && ok[vv]) s4[vv] = s_stage[(vv * TV + vox) * kLPV + q];
        }
        float4 d4[kViewChunk];
        if (AGG == kAggSum) {
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) d4[vv] = make_float4(go[0][j], go[1][j], go[2][j], go[3][j]);
        } else if (AGG == kAggConf) {
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            const float sv[kCPL] = {s4[vv].x, s4[vv].y, s4[vv].z, s4[vv].w};
#pragma unroll
            for (int k = 0; k < kCPL; ++k) gcw[vv][k] = fmaf(go[k][j], sv[k], gcw[vv][k]);  // s == 0 where unseen
            d4[vv] = make_float4(cw[vv][0] * go[0][j], cw[vv][1] * go[1][j], cw[vv][2] * go[2][j], cw[vv][3] * go[3][j]);
          }
        } else if (AGG == kAggSoftmax) {
          // two channels per packed instruction (FFMA2 / FMUL2 / FADD2: each half the same IEEE operation as the
          // scalar form): d_v = g * e_v / den * (1 - out + s_v), e_v = exp(s_v - max)
          f2 dl[kViewChunk], dh[kViewChunk];
          softmax_bwd4_x2(pack2(s4[0].x, s4[0].y), pack2(s4[1].x, s4[1].y), pack2(s4[2].x, s4[2].y),
                          pack2(s4[3].x, s4[3].y), pack2(go[0][j], go[1][j]), dl);
          softmax_bwd4_x2(pack2(s4[0].z, s4[0].w), pack2(s4[1].z, s4[1].w), pack2(s4[2].z, s4[2].w),
                          pack2(s4[3].z, s4[3].w), pack2(go[2][j], go[3][j]), dh);
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) {
            unpack2(dl[vv], d4[vv].x, d4[vv].y);
            unpack2(dh[vv], d4[vv].z, d4[vv].w);
          }
        } else {
          float dd[kViewChunk][kCPL];
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            const float sk[kViewChunk] = {(&s4[0].x)[k], (&s4[1].x)[k], (&s4[2].x)[k], (&s4[3].x)[k]};
            float m = sk[0];
            int am = 0;
#pragma unroll
            for (int vv = 1; vv < kViewChunk; ++vv)
              if (sk[vv] > m) { m = sk[vv]; am = vv; }
#pragma unroll
            for (int vv = 0; vv < kViewChunk; ++vv) dd[vv][k] = (vv == am) ? go[k][j] : 0.f;
          }
#pragma unroll
          for (int vv = 0; vv < kViewChunk; ++vv) d4[vv] = make_float4(dd[vv][0], dd[vv][1], dd[vv][2], dd[vv][3]);
        }
#pragma unroll
        for (int vv = 0; vv < kViewChunk; ++vv)
          if (ok[vv]) s_stage[(vv * TV + vox) * kLPV + q] = d4[vv];
      }
    }
    __syncthreads();

    // ---------------- C: row walk over the cell-sorted records ----------------
    {
      float4 a00 = make_float4(0.f, 0.f, 0.f, 0.f), a01 = a00, a10 = a00, a11 = a00;
      int cur = -64;      // (offset | view) of the cell in the registers; -64 = none (never one cell left of anything)
      float* pc = gbv;    // its north-west texel in the gradient map, and the texel below
      float* pc1 = gbv1;
#pragma unroll 1
      for (int rec = rec_begin; rec < rec_end; ++rec) {
        const int2 m = s_rec_m[rec];
        const float4 w = s_rec_w[rec];
        const float4 d = s_stage[m.y * kLPV + q];
        if (m.x != cur) {
          const bool shift = (m.x == cur + kCp);  // same view (low bits), next cell to the right in the same row
          if (cur >= 0) {
            red4(pc, a00);
            red4(pc1, a10);
            if (!shift) {
              red4(pc + kCp, a01);
              red4(pc1 + kCp, a11);
            }
          }
          if (shift) {
            a00 = a01; a10 = a11;
          } else {
            a00 = make_float4(0.f, 0.f, 0.f, 0.f); a10 = a00;
          }
          a01 = make_float4(0.f, 0.f, 0.f, 0.f); a11 = a01;
          cur = m.x;
          pc = gbv + (m.x & ~3);
          pc1 = gbv1 + (m.x & ~3);
        }
        fma4(a00, w.x, d); fma4(a01, w.y, d); fma4(a10, w.z, d); fma4(a11, w.w, d);
      }
      if (cur >= 0) {
        red4(pc, a00);
        red4(pc + kCp, a01);
        red4(pc1, a10);
        red4(pc1 + kCp, a11);
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
cudaError_t launch_one(UnprojArgs& a, const UnprojShape& s, cudaStream_t st, const CoordSpecArg* spec) {
  tile_dims<CF>(s, a.nt0, a.nt1, a.nt2);
  a.ncb = 1;
  const long long nblk = (long long)a.B * a.nt0 * a.nt1 * a.nt2;
  if (nblk > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  if (spec != nullptr) {
    LT_SMEM_OPT_IN_ONCE((k_unproject_bwd_tile<CF, AGG, true>), CF::kBytes);
    k_unproject_bwd_tile<CF, AGG, true><<<(unsigned)nblk, CF::NT, CF::kBytes, st>>>(make_kargs<true>(a, spec));
  } else {
    LT_SMEM_OPT_IN_ONCE((k_unproject_bwd_tile<CF, AGG, false>), CF::kBytes);
    k_unproject_bwd_tile<CF, AGG, false><<<(unsigned)nblk, CF::NT, CF::kBytes, st>>>(make_kargs<false>(a, nullptr));
  }
  return cudaGetLastError();
}

template <class CF>
cudaError_t launch_agg(UnprojArgs& a, const UnprojShape& s, cudaStream_t st, const CoordSpecArg* spec) {
  switch (s.agg) {
    case kAggSum: return launch_one<CF, kAggSum>(a, s, st, spec);
    case kAggConf: return launch_one<CF, kAggConf>(a, s, st, spec);
    case kAggMax: return launch_one<CF, kAggMax>(a, s, st, spec);
    case kAggSoftmax: return launch_one<CF, kAggSoftmax>(a, s, st, spec);
  }
  return cudaErrorInvalidValue;
}

}  // namespace

bool unproject_bwd_tile_eligible(const UnprojShape& s, const float* grad_out) {
  // MEASURED (profiles/r2a_k1bwd_tile_ncu.txt): at cfg3 this formulation cuts the L2 reduction sectors 2.5x but runs
  // 1.37 ms against 1.05 ms for the scatter kernel (issue-bound: 2.6 k instructions per warp at 16 warps per SM), so it
  // is opt-in: LT_B200_K1BWD=tile.
  const char* e = getenv("LT_B200_K1BWD");
  if (!(e && e[0] == 't')) return false;
  if (padded_channels(s.C) != kCp) return false;
  if (s.N2 % 8 != 0 || (reinterpret_cast<uintptr_t>(grad_out) & 31u) != 0) return false;
  if ((s.agg == kAggMax || s.agg == kAggSoftmax) && s.V > kViewChunk) return false;
  if (s.H + 2 > 2047 || s.W + 2 > 2047) return false;  // texel offsets stay far inside int32
  return true;
}

size_t unproject_bwd_tile_ws_bytes(const UnprojShape& s) {
  return (size_t)s.B * (size_t)tile_count(s) * s.V * s.C * sizeof(float);
}

cudaError_t launch_unproject_bwd_tile(UnprojArgs& a, const UnprojShape& s, float* grad_conf, cudaStream_t st,
                                      const CoordSpecArg* spec) {
  cudaError_t e;
  switch (tile_cfg_id()) {
    case 1: e = launch_agg<CfgB>(a, s, st, spec); break;
    case 2: e = launch_agg<CfgC>(a, s, st, spec); break;
    case 3: e = launch_agg<CfgD>(a, s, st, spec); break;
    default: e = launch_agg<CfgA>(a, s, st, spec); break;
  }
  if (e != cudaSuccess) return e;
  if (a.gconf_part != nullptr) e = launch_conf_reduce_rows(a.gconf_part, grad_conf, s.B, s.V * s.C, a.nt0 * a.nt1 * a.nt2, st);
  return e;
}

// grad_conf (B, V*C) = sum of the `nrows` partial rows per item, in row order (shared with the scatter kernel)
cudaError_t launch_conf_reduce_rows(const float* part, float* grad_conf, int B, int VC, int nrows, cudaStream_t st) {
  k_conf_reduce_tiles<<<dim3((VC + 31) / 32, B), dim3(32, 8), 0, st>>>(part, grad_conf, VC, nrows);
  return cudaGetLastError();
}

}  // namespace lt
