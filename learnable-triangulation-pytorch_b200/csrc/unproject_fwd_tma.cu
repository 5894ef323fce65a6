// unproject_fwd_tma.cu — K1 forward as a persistent, warp-specialised kernel whose gather reads shared memory.
//
// Same function as k_unproject_fwd (unproject_fwd.cu; mvn/utils/op.py:107-161) for the reference's own shape class
// (32-channel features, <= 4 views, N2 % 8 == 0); everything else keeps the tag-cached kernel.
//
//   * one CTA per SM, resident for the whole launch; it walks the tiles  blockIdx.x, + gridDim.x, ...  so that the SMs
//     work on neighbouring tiles of one batch item at any time (its 4.7 MB of feature maps stay in L2);
//   * a tile is 4 x 4 x 8 voxels. Its footprint in the feature map of view v is the bounding box of the projections of
//     its 8 corner voxels (a projective map sends the cuboid's convex hull to the hull of the corner images; the box is
//     widened by 1/64 pixel against fp32 rounding). Mean 143 texels x 128 B per view at Human3.6M geometry;
//   * warp 9 ("box warp") computes the four boxes of the tiles ahead and queues them; warp 8 ("issue warp") reserves
//     space for them in a ring of shared memory (variable-size regions: boxes range from 0 to ~500 texels) and moves
//     every box row with ONE cp.async.bulk (TMA engine, UBLKCP) that completes on the tile's `full` mbarrier: no LSU
//     instruction, no L1 tag, no miss. Up to kSlots tiles are in flight; the ring is released tile by tile through
//     `empty` mbarriers. (MEASURED: with both jobs in one warp that warp's ~700 dependent instructions per tile, at one
//     issue per ~8 cycles, bounded the kernel at 0.95 ms whether or not anything was copied or gathered.)
//   * warps 0..7 ("consumers") own 4 runs x 4 z steps = 16 voxels of every tile. Phase A: one lane per (voxel, view
//     pair) computes the taps (make_taps: the SAME arithmetic as every other K1 kernel; the voxel's coordinates were
//     requested one tile earlier with cp.async) and, once `full` has completed, translates them into shared-memory
//     addresses inside the boxes. Gather: 8 lanes x LDS.128 = the 128-byte texel of one voxel = one conflict-free
//     wavefront, 4 taps per (voxel, view), predicated (an absent view costs no wavefront), all 16 loads of a z step in
//     flight together, FFMA2 blend, aggregation over the views in registers, one STG.128 per channel (4 voxels);
//   * a tap that is not inside its box (a corner voxel behind the camera: no box; a box that does not fit the ring;
//     rounding beyond the guard band) is fetched from global memory instead (same descriptor, negative encoding): the
//     result never depends on the box being right, only the speed does.
#include <stdio.h>
#include <stdlib.h>

#include "unproject.cuh"

namespace lt {
namespace {

constexpr int kTmaConsumers = 8;                       // gather warps
constexpr int kTmaThreads = (kTmaConsumers + 2) * 32;  // + the issue warp and the box warp
constexpr int kBoxQ = 4;
#ifndef LT_TMA_VGROUP
#define LT_TMA_VGROUP 2    // views whose 4 tap loads are in flight together in the gather
#endif
#ifndef LT_TMA_CPASYNC
#define LT_TMA_CPASYNC 0   // 1: voxel centres of the next tile through cp.async + shared memory instead of registers
#endif                               // boxes queued between the two
constexpr int kTmaT0 = 4, kTmaT1 = 4, kTmaT2 = 8;      // tile extents (voxels)
constexpr int kTmaSlots = 4;                           // tiles in flight (descriptor slots / mbarrier pairs)
constexpr int kTexelBytes = 128;                       // Cp == 32 floats
constexpr int kVoxPerWarp = 16;                        // 4 runs x 4 z steps
constexpr unsigned kRingBytes = 209u * 1024u;          // staging ring (1672 texels)
constexpr float kBoxGuard = 1.f / 64.f;                // pixels

struct TileDesc {          // producer -> consumers, one per slot
  int b, i0, i1, z0;       // item and voxel origin of the tile
  int bx0[4], by0[4];      // padded texel coordinates of the staged box origin
  int bx1[4], by1[4];      // inclusive upper corner; bx1 < bx0 = nothing staged for this view
  unsigned base[4];        // shared-memory byte address of the box origin
  unsigned pitch[4];       // bytes per staged row
};

struct BoxEntry {          // box warp -> issue warp
  int b, i0, i1, z0;
  int4 box[4];             // padded texel coordinates (x0, y0, x1, y1), inclusive; x1 < x0 = nothing to stage
};

struct TmaSmem {
  unsigned long long full[kTmaSlots];
  unsigned long long empty[kTmaSlots];
  unsigned long long bfull[kBoxQ];
  unsigned long long bempty[kBoxQ];
  BoxEntry boxq[kBoxQ];
  unsigned vstart[kTmaSlots];   // producer bookkeeping: virtual ring address of the slot's first region
  TileDesc desc[kTmaSlots];
  float4 proj[kTmaConsumers][kViewChunk * 3];         // each warp's copy of the current item's projection rows
  float4 xyz[kTmaConsumers][2][kVoxPerWarp];          // voxel centres of the next tile, filled by cp.async (tensor path)
  float4 w[kTmaConsumers][kViewChunk][kVoxPerWarp];   // per-warp tap table, voxel index = z step * 4 + run
  int off[kTmaConsumers][kViewChunk][kVoxPerWarp];
};

template <bool SPEC>
struct TmaArgs {
  KArgs<SPEC> k;
  int mt0, mt1, mt2;   // tiles per axis
  int ntiles;          // B * mt0 * mt1 * mt2
  int dbg;             // LT_B200_TMA_DBG (timing experiments, results wrong): 1 = no copies issued, 2 = no gather
#ifdef LT_TMA_PROF
  unsigned long long* prof;   // [8] cycle counters summed over CTAs (see PROF_ADD)
#endif
};

#ifdef LT_TMA_PROF
#define PROF_T0() const long long prof_t0_ = clock64()
#define PROF_ADD(acc) acc += clock64() - prof_t0_
#else
#define PROF_T0() do {} while (0)
#define PROF_ADD(acc) do {} while (0)
#endif

// ---- PTX wrappers ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned mbar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned mbar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// bounded wait: a protocol error becomes a trap after 20 s (time-based, as in pwconv.cu) instead of a hung GPU
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity) {
  unsigned long long t0 = 0;
  for (unsigned i = 0;; ++i) {
    unsigned done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(mbar), "r"(parity) : "memory");
    if (done) return;
    if (i >= (1u << 14)) {
      if (t0 == 0) t0 = globaltimer_ns();
      __nanosleep(128);
      if ((i & 0x3ffu) == 0 && globaltimer_ns() - t0 > 20000000000ull) __trap();
    }
  }
}
// one contiguous global run -> shared memory through the TMA engine; bytes % 16 == 0, both sides 16-byte aligned
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar) : "memory");
}
__device__ __forceinline__ float4 lds128(unsigned addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
// predicated on off >= 0: no data-pipe wavefront for a voxel whose view is absent; the result is then UNDEFINED (the
// caller selects 0 after the blend: pre-zeroing 64 registers per z step cost 30 % of the gather's instructions)
__device__ __forceinline__ float4 lds128_if(unsigned addr, int off) {
  float4 r;
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %5, 0;\n\t@p ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr), "r"(off));
  return r;
}
__device__ __forceinline__ void cp_async4(unsigned dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ bool elect_one() {
  unsigned leader;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(leader));
  return leader != 0;
}
__device__ __forceinline__ void stg128_cs(float* p, float a, float b, float c, float d) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// heat-map position of a point, for the bounding boxes only (the taps come from make_taps)
__device__ __forceinline__ bool box_ixy(float x, float y, float z, const float4& r0, const float4& r1, const float4& r2,
                                        int H, int W, int align_corners, float& ix, float& iy) {
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

template <bool SPEC>
__device__ __forceinline__ void tile_xyz(const KArgs<SPEC>& a, int b, int vi0, int vi1, int vz, float& x, float& y, float& z);
template <>
__device__ __forceinline__ void tile_xyz<false>(const KArgs<false>& a, int b, int vi0, int vi1, int vz, float& x, float& y,
                                                float& z) {
  const float* cp = a.coord + ((size_t)b * a.nvox + ((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
  x = __ldg(cp); y = __ldg(cp + 1); z = __ldg(cp + 2);
}
template <>
__device__ __forceinline__ void tile_xyz<true>(const KArgs<true>& a, int b, int vi0, int vi1, int vz, float& x, float& y,
                                               float& z) {
  float o[3];
  coord_of_voxel(a.spec + (size_t)b * kCoordSpec, a.transfer, a.N1, vi0, vi1, vz, o);   // 18 floats, L1-resident
  x = o[0]; y = o[1]; z = o[2];
}

// tile index -> (b, t0, t1, t2), advanced by the grid stride without divisions (digits with carries)
struct TileWalk {
  int b, t0, t1, t2;
  int sb, s0, s1, s2;
  template <bool SPEC>
  __device__ __forceinline__ void init(const TmaArgs<SPEC>& A, int first, int stride) {
    int t = first;
    t2 = t % A.mt2; t /= A.mt2; t1 = t % A.mt1; t /= A.mt1; t0 = t % A.mt0; b = t / A.mt0;
    t = stride;
    s2 = t % A.mt2; t /= A.mt2; s1 = t % A.mt1; t /= A.mt1; s0 = t % A.mt0; sb = t / A.mt0;
  }
  template <bool SPEC>
  __device__ __forceinline__ void next(const TmaArgs<SPEC>& A) {
    t2 += s2; int c = t2 >= A.mt2 ? 1 : 0; t2 -= c * A.mt2;
    t1 += s1 + c; c = t1 >= A.mt1 ? 1 : 0; t1 -= c * A.mt1;
    t0 += s0 + c; c = t0 >= A.mt0 ? 1 : 0; t0 -= c * A.mt0;
    b += sb + c;
  }
};

// ---------------------------------------------------------------------------------------------------------------------
// box warp: bounding boxes of the tiles ahead -> queue. Lane = (view, corner voxel).
// ---------------------------------------------------------------------------------------------------------------------
template <bool SPEC>
__device__ __forceinline__ void corner_xyz(const TmaArgs<SPEC>& A, const TileWalk& T, int pc, float& x, float& y, float& z) {
  const KArgs<SPEC>& a = A.k;
  const int ci0 = min(T.t0 * kTmaT0 + ((pc & 4) ? kTmaT0 - 1 : 0), a.N0 - 1);
  const int ci1 = min(T.t1 * kTmaT1 + ((pc & 2) ? kTmaT1 - 1 : 0), a.N1 - 1);
  const int cz = min(T.t2 * kTmaT2 + ((pc & 1) ? kTmaT2 - 1 : 0), a.N2 - 1);
  tile_xyz<SPEC>(a, T.b, ci0, ci1, cz, x, y, z);
}

template <bool SPEC>
__device__ __forceinline__ void box_warp(const TmaArgs<SPEC>& A, TmaSmem& S, int lane) {
  const KArgs<SPEC>& a = A.k;
  const int pv = lane >> 3, pc = lane & 7;
  TileWalk T;
  T.init(A, blockIdx.x, gridDim.x);
  float cx = 0.f, cy = 0.f, cz = 0.f;
  if (blockIdx.x < A.ntiles) corner_xyz(A, T, pc, cx, cy, cz);
  float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0;
  int pb = -1;
  int seq = 0;
  long long pw0 = 0;
  const long long pstart = clock64();
  for (int tile = blockIdx.x; tile < A.ntiles; tile += gridDim.x, ++seq) {
    const int b = T.b, i0 = T.t0 * kTmaT0, i1 = T.t1 * kTmaT1, z0 = T.t2 * kTmaT2;
    const float x = cx, y = cy, z = cz;
    // the next tile's corner coordinates are requested now and used one iteration later
    T.next(A);
    if (tile + (int)gridDim.x < A.ntiles) corner_xyz(A, T, pc, cx, cy, cz);
    if (b != pb && pv < a.V) {
      const float4* pp = reinterpret_cast<const float4*>(a.proj) + ((size_t)b * a.V + pv) * 3;
      r0 = __ldg(pp); r1 = __ldg(pp + 1); r2 = __ldg(pp + 2);
    }
    pb = b;
    float ix = 0.f, iy = 0.f;
    int all_ok = 0;
    if (pv < a.V) all_ok = box_ixy(x, y, z, r0, r1, r2, a.H, a.W, a.align_corners, ix, iy) ? 1 : 0;
    float xmin = ix, xmax = ix, ymin = iy, ymax = iy;
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      xmin = fminf(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
      xmax = fmaxf(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
      ymin = fminf(ymin, __shfl_xor_sync(0xffffffffu, ymin, o));
      ymax = fmaxf(ymax, __shfl_xor_sync(0xffffffffu, ymax, o));
      all_ok &= __shfl_xor_sync(0xffffffffu, all_ok, o);
    }
    // padded texel coordinates: tap (x0, y0) reads padded columns x0+1, x0+2 and rows y0+1, y0+2; x0 in [-1, W-1]
    int4 bx = make_int4(0, 0, -1, -1);
    if (all_ok) {
      const float lo_x = fminf(fmaxf(xmin - kBoxGuard, -2.f), (float)a.W + 1.f), hi_x = fminf(fmaxf(xmax + kBoxGuard, -2.f), (float)a.W + 1.f);
      const float lo_y = fminf(fmaxf(ymin - kBoxGuard, -2.f), (float)a.H + 1.f), hi_y = fminf(fmaxf(ymax + kBoxGuard, -2.f), (float)a.H + 1.f);
      bx.x = max((int)floorf(lo_x), -1) + 1;
      bx.z = min((int)floorf(hi_x), a.W - 1) + 2;
      bx.y = max((int)floorf(lo_y), -1) + 1;
      bx.w = min((int)floorf(hi_y), a.H - 1) + 2;
      if (bx.z < bx.x + 1 || bx.w < bx.y + 1) { bx.z = bx.x - 1; bx.w = bx.y - 1; }  // entirely outside the image
    }
    const int q = seq % kBoxQ;
    { PROF_T0(); mbar_wait(smem_u32(&S.bempty[q]), ((unsigned)(seq / kBoxQ) & 1u) ^ 1u); PROF_ADD(pw0); }
    BoxEntry& E = S.boxq[q];
    if (pc == 0) E.box[pv] = bx;
    if (lane == 0) { E.b = b; E.i0 = i0; E.i1 = i1; E.z0 = z0; }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&S.bfull[q]));
  }
#ifdef LT_TMA_PROF
  if (lane == 0 && A.prof) { atomicAdd(A.prof + 0, (unsigned long long)pw0); atomicAdd(A.prof + 1, (unsigned long long)(clock64() - pstart)); }
#endif
}

// ---------------------------------------------------------------------------------------------------------------------
// issue warp: ring space -> descriptor -> bulk copies. All 32 lanes run the same control flow with the same values.
// ---------------------------------------------------------------------------------------------------------------------
template <bool SPEC>
__device__ __forceinline__ void issue_warp(const TmaArgs<SPEC>& A, TmaSmem& S, unsigned ring_addr, int lane) {
  const KArgs<SPEC>& a = A.k;
  const size_t viewStride = (size_t)(a.H + 2) * (a.W + 2) * a.Cp;
  unsigned vhead = 0, phead = 0;            // virtual (wrapping) and physical ring head
  int oldest = 0;                           // oldest tile still owning ring space
  int seq = 0;
  long long pw0 = 0, pw1 = 0;
  const long long pstart = clock64();
  for (int tile = blockIdx.x; tile < A.ntiles; tile += gridDim.x, ++seq) {
    const int q = seq % kBoxQ;
    { PROF_T0(); mbar_wait(smem_u32(&S.bfull[q]), (unsigned)(seq / kBoxQ) & 1u); PROF_ADD(pw0); }
    const BoxEntry& E = S.boxq[q];
    const int b = E.b, i0 = E.i0, i1 = E.i1, z0 = E.z0;
    int X0[4], Y0[4], X1[4], Y1[4];
    unsigned bytes[4], total = 0;
#pragma unroll
    for (int v = 0; v < 4; ++v) {
      const int4 bx = E.box[v];
      X0[v] = bx.x; Y0[v] = bx.y; X1[v] = bx.z; Y1[v] = bx.w;
      bytes[v] = (X1[v] >= X0[v]) ? (unsigned)((X1[v] - X0[v] + 1) * (Y1[v] - Y0[v] + 1)) * kTexelBytes : 0u;
      total += bytes[v];
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&S.bempty[q]));
    // a tile that cannot fit the ring at all: its largest boxes are not staged (those views gather from global memory)
    while (total > kRingBytes) {
      int big = 0;
#pragma unroll
      for (int v = 1; v < 4; ++v)
        if (bytes[v] > bytes[big]) big = v;
#pragma unroll
      for (int v = 0; v < 4; ++v)
        if (v == big) { total -= bytes[v]; bytes[v] = 0; X1[v] = X0[v] - 1; Y1[v] = Y0[v] - 1; }
    }
    // ---- ring space: regions are contiguous per view; wait for the oldest tiles until the new span fits ----
    unsigned base[4], vstart = 0, vh = 0, ph = 0;
    for (;;) {
      vh = vhead; ph = phead;
      bool first = true;
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        base[v] = 0;
        if (bytes[v] != 0) {
          if (ph + bytes[v] > kRingBytes) { vh += kRingBytes - ph; ph = 0; }
          if (first) { vstart = vh; first = false; }
          base[v] = ph;
          ph += bytes[v]; vh += bytes[v];
        }
      }
      if (first) vstart = vh;
      while (oldest < seq && (seq - oldest >= kTmaSlots || vh - S.vstart[oldest % kTmaSlots] > kRingBytes)) {
        { PROF_T0(); mbar_wait(smem_u32(&S.empty[oldest % kTmaSlots]), (unsigned)(oldest / kTmaSlots) & 1u); PROF_ADD(pw1); }
        ++oldest;
      }
      if (oldest == seq && vh - vstart > kRingBytes) {  // alone, but the skips at the ring's end make it overlap itself
        vhead += kRingBytes - phead; phead = 0;
        continue;
      }
      break;
    }
    vhead = vh; phead = ph;
    const int slot = seq % kTmaSlots;
    const unsigned full = smem_u32(&S.full[slot]);
    // ---- publish the descriptor, arm the barrier, issue the copies ----
    TileDesc& D = S.desc[slot];
    if (lane == 0) { D.b = b; D.i0 = i0; D.i1 = i1; D.z0 = z0; S.vstart[slot] = vstart; }
    if (lane < 4) {
#pragma unroll
      for (int v = 0; v < 4; ++v)
        if (lane == v) {
          D.bx0[v] = X0[v]; D.by0[v] = Y0[v]; D.bx1[v] = X1[v]; D.by1[v] = Y1[v];
          D.base[v] = ring_addr + base[v];
          D.pitch[v] = (unsigned)(X1[v] - X0[v] + 1) * kTexelBytes;
        }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive_expect_tx(full, A.dbg == 1 ? 0u : total);
    __syncwarp();
    // one bulk copy per box row, issued from warp-uniform code by one elected lane. (MEASURED: with one row per lane the
    // compiler serialises the lanes of the uniform-datapath instruction UBLKCP in an ELECT / R2UR loop, ~70 cycles per
    // copy = 3.8 k cycles per tile; the values are re-broadcast with shuffles so that it can prove them uniform.)
    const bool leader = elect_one();
    if (A.dbg != 1) {
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        const int hv = __shfl_sync(0xffffffffu, bytes[v] ? (Y1[v] - Y0[v] + 1) : 0, 0);
        const unsigned rowb = __shfl_sync(0xffffffffu, (unsigned)(X1[v] - X0[v] + 1) * kTexelBytes, 0);
        const unsigned dst0 = __shfl_sync(0xffffffffu, ring_addr + base[v], 0);
        const unsigned long long src0 = (unsigned long long)(a.feat + ((size_t)b * a.V + v) * viewStride +
                                                             ((size_t)Y0[v] * (a.W + 2) + X0[v]) * a.Cp);
        const unsigned lo = __shfl_sync(0xffffffffu, (unsigned)src0, 0), hi = __shfl_sync(0xffffffffu, (unsigned)(src0 >> 32), 0);
        const char* src = reinterpret_cast<const char*>(((unsigned long long)hi << 32) | lo);
        const unsigned gpitch = (unsigned)(a.W + 2) * a.Cp * 4u;
        for (int r = 0; r < hv; ++r)
          if (leader) bulk_g2s(dst0 + (unsigned)r * rowb, src + (size_t)r * gpitch, rowb, full);
      }
    }
  }
#ifdef LT_TMA_PROF
  if (lane == 0 && A.prof) { atomicAdd(A.prof + 2, (unsigned long long)pw0); atomicAdd(A.prof + 3, (unsigned long long)pw1); atomicAdd(A.prof + 4, (unsigned long long)(clock64() - pstart)); }
#endif
}

// ---------------------------------------------------------------------------------------------------------------------
// consumer warps
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void blend4(const float4& t00, const float4& t01, const float4& t10, const float4& t11,
                                       const float4& w, float (&s)[kCPL]) {
  const f2 wx = pack2(w.x, w.x), wy = pack2(w.y, w.y), wz = pack2(w.z, w.z), ww = pack2(w.w, w.w);
  f2 lo = mul2(pack2(t00.x, t00.y), wx), hi = mul2(pack2(t00.z, t00.w), wx);
  lo = fma2(pack2(t01.x, t01.y), wy, lo); hi = fma2(pack2(t01.z, t01.w), wy, hi);
  lo = fma2(pack2(t10.x, t10.y), wz, lo); hi = fma2(pack2(t10.z, t10.w), wz, hi);
  lo = fma2(pack2(t11.x, t11.y), ww, lo); hi = fma2(pack2(t11.z, t11.w), ww, hi);
  unpack2(lo, s[0], s[1]);
  unpack2(hi, s[2], s[3]);
}

template <int AGG>
__device__ __forceinline__ void aggregate(const float (&s)[kViewChunk][kCPL], const float (&cw)[kViewChunk][kCPL], int nv,
                                          float (&val)[kCPL]) {
  if (AGG == kAggSoftmax) {
#pragma unroll
    for (int k = 0; k < kCPL; k += 2) {
      const float ca[4] = {s[0][k], s[1][k], s[2][k], s[3][k]};
      const float cb[4] = {s[0][k + 1], s[1][k + 1], s[2][k + 1], s[3][k + 1]};
      softmax_mix4_x2(ca, cb, val[k], val[k + 1]);
    }
  } else {
#pragma unroll
    for (int k = 0; k < kCPL; ++k) {
      float r = (AGG == kAggMax) ? -INFINITY : 0.f;
#pragma unroll
      for (int v = 0; v < kViewChunk; ++v)
        if (v < nv) {
          if (AGG == kAggSum) r += s[v][k];
          else if (AGG == kAggConf) r = fmaf(cw[v][k], s[v][k], r);
          else r = fmaxf(r, s[v][k]);
        }
      val[k] = r;
    }
  }
}

template <int AGG, bool SPEC>
__device__ __forceinline__ void consumer(const TmaArgs<SPEC>& A, TmaSmem& S, int warp, int lane) {
  const KArgs<SPEC>& a = A.k;
  // gather role: lane = (run g of the warp's round, channel quad q)
  // A warp owns 2 of the tile's 16 runs; lane groups 0/1 take their first 4 z steps, groups 2/3 the last 4, so that the
  // two halves of a 32-byte output sector leave in the SAME store instruction (MEASURED: with the z halves in different
  // warps every STG.128 was 32 half-sector requests and the four stores of a tile took ~1000 cycles).
  const int g = lane >> 3, q = lane & 7;
  const int gR = 2 * warp + (g & 1), gd0 = gR >> 2, gd1 = gR & 3, gzh = g >> 1;
  // phase-A role: lane = (voxel, view pair), voxel = z step * 4 + lane group
  const int voxA = lane & 15, vpA = lane >> 4;
  const int gA = voxA & 3, jA = voxA >> 2;
  const int aR = 2 * warp + (gA & 1), ad0 = aR >> 2, ad1 = aR & 3, azh = gA >> 1;
  const float invH = 1.f / (float)a.H, invW = 1.f / (float)a.W;
  const int rowStride = (a.W + 2) * a.Cp;
  const size_t viewStride = (size_t)(a.H + 2) * (a.W + 2) * a.Cp;
  const int nv = a.V;
  float4 (*s_w)[kVoxPerWarp] = S.w[warp];
  int (*s_off)[kVoxPerWarp] = S.off[warp];
  const unsigned qb = (unsigned)q * 16u;

  float cw[kViewChunk][kCPL];
#pragma unroll
  for (int v = 0; v < kViewChunk; ++v)
#pragma unroll
    for (int k = 0; k < kCPL; ++k) cw[v][k] = 0.f;
  // projection rows of the current item: in shared memory, not in 24 registers (a spill in this loop is a local-memory
  // access with ~24 KB of L1 left beside the ring: MEASURED long-scoreboard stalls of 3.7 cycles per issue)
  float4* P = S.proj[warp];
  int pb = -1;

  // Voxel centres (tensor path): the first view-pair lane of each voxel requests them one tile ahead with cp.async into
  // the warp's staging rows: no register holds them across the gather. (MEASURED: as prefetched registers they were
  // spilled, and the spill store waits for the load: the DRAM latency sat on every tile.)
  TileWalk T;
  T.init(A, blockIdx.x, gridDim.x);
  float4 (*s_xyz)[kVoxPerWarp] = S.xyz[warp];
#if LT_TMA_CPASYNC
  if (!SPEC && vpA == 0) {
    const int vi0 = T.t0 * kTmaT0 + ad0, vi1 = T.t1 * kTmaT1 + ad1, vz = T.t2 * kTmaT2 + azh * 4 + jA;
    if (blockIdx.x < A.ntiles && vi0 < a.N0 && vi1 < a.N1) {
      const float* cp = a.coord + ((size_t)T.b * a.nvox + ((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
      const unsigned dst = smem_u32(&s_xyz[0][voxA]);
      cp_async4(dst, cp); cp_async4(dst + 4, cp + 1); cp_async4(dst + 8, cp + 2);
    }
  }
#else
  float nx = 0.f, ny = 0.f, nz = 0.f;   // next tile's voxel centre, in registers
  if (!SPEC) {
    const int vi0 = T.t0 * kTmaT0 + ad0, vi1 = T.t1 * kTmaT1 + ad1, vz = T.t2 * kTmaT2 + azh * 4 + jA;
    if (blockIdx.x < A.ntiles && vi0 < a.N0 && vi1 < a.N1) tile_xyz<SPEC>(a, T.b, vi0, vi1, vz, nx, ny, nz);
  }
#endif
  int seq = 0;
  long long pw0 = 0, pA = 0, pG = 0, pS = 0;
  const long long pstart = clock64();
  for (int tile = blockIdx.x; tile < A.ntiles; tile += gridDim.x, ++seq) {
    const int slot = seq % kTmaSlots;
    const int b = T.b, i0 = T.t0 * kTmaT0 + gd0, vi1 = T.t1 * kTmaT1 + gd1, zb = T.t2 * kTmaT2 + gzh * 4;
    const int a0 = T.t0 * kTmaT0 + ad0, a1 = T.t1 * kTmaT1 + ad1, az = T.t2 * kTmaT2 + azh * 4 + jA;   // phase-A voxel
#ifdef LT_TMA_PROF
    const long long ps0 = clock64();
#endif
    if (b != pb) {
      pb = b;
      __syncwarp();
      if (lane < nv * 3) P[lane] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)b * a.V * 3 + lane);
      __syncwarp();
      if (AGG == kAggConf) {
#pragma unroll
        for (int v = 0; v < kViewChunk; ++v)
#pragma unroll
          for (int k = 0; k < kCPL; ++k) {
            const int c = q * kCPL + k;
            cw[v][k] = (v < nv && c < a.C) ? __ldg(a.conf + ((size_t)b * a.V + v) * a.C + c) : 0.f;
          }
      }
    }
    // ---- phase A, arithmetic part: the taps of (voxel voxA, views 2*vpA, 2*vpA+1); needs no staged data ----
    int tx[2], ty[2];
    float4 tw[2];
    bool tv[2];
    {
      const bool okA = a0 < a.N0 && a1 < a.N1;
      float x = 0.f, y = 0.f, z = 0.f;
      if (SPEC) {
        if (okA) tile_xyz<SPEC>(a, b, a0, a1, az, x, y, z);
      } else {
#if LT_TMA_CPASYNC
        cp_async_wait_all();
        __syncwarp();
        const float4 c = s_xyz[seq & 1][voxA];
        x = c.x; y = c.y; z = c.z;
#else
        x = nx; y = ny; z = nz;
#endif
      }
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        tw[e] = make_float4(0.f, 0.f, 0.f, 0.f);
        tx[e] = 0; ty[e] = 0;
        tv[e] = okA && (vpA * 2 + e < nv) &&
                make_taps(x, y, z, P[(vpA * 2 + e) * 3], P[(vpA * 2 + e) * 3 + 1], P[(vpA * 2 + e) * 3 + 2], a.H, a.W, invH, invW,
                          a.align_corners, tx[e], ty[e], tw[e]);
      }
    }
    // next tile's voxels: requested now, read after this tile's gather
    T.next(A);
#if LT_TMA_CPASYNC
    if (!SPEC && vpA == 0) {
      const int vi0 = T.t0 * kTmaT0 + ad0, vi1 = T.t1 * kTmaT1 + ad1, vz = T.t2 * kTmaT2 + azh * 4 + jA;
      if (tile + (int)gridDim.x < A.ntiles && vi0 < a.N0 && vi1 < a.N1) {
        const float* cp = a.coord + ((size_t)T.b * a.nvox + ((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
        const unsigned dst = smem_u32(&s_xyz[(seq + 1) & 1][voxA]);
        cp_async4(dst, cp); cp_async4(dst + 4, cp + 1); cp_async4(dst + 8, cp + 2);
      }
    }
#else
    if (!SPEC) {
      const int vi0 = T.t0 * kTmaT0 + ad0, vi1 = T.t1 * kTmaT1 + ad1, vz = T.t2 * kTmaT2 + azh * 4 + jA;
      nx = 0.f; ny = 0.f; nz = 0.f;
      if (tile + (int)gridDim.x < A.ntiles && vi0 < a.N0 && vi1 < a.N1) tile_xyz<SPEC>(a, T.b, vi0, vi1, vz, nx, ny, nz);
    }
#endif

#ifdef LT_TMA_PROF
    const long long ps1 = clock64();
    pA += ps1 - ps0;
#endif
    { PROF_T0(); mbar_wait(smem_u32(&S.full[slot]), (unsigned)(seq / kTmaSlots) & 1u); PROF_ADD(pw0); }
#ifdef LT_TMA_PROF
    const long long ps2 = clock64();
#endif
    const TileDesc& D = S.desc[slot];
    // ---- phase A, address part: inside the staged box -> shared-memory address, else global offset ----
    bool direct = false;
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int v = vpA * 2 + e;
      int off = -1;
      if (tv[e]) {
        const int px = tx[e] + 1, py = ty[e] + 1;
        if (px >= D.bx0[v] && px + 1 <= D.bx1[v] && py >= D.by0[v] && py + 1 <= D.by1[v]) {
          off = (int)(D.base[v] + (unsigned)(py - D.by0[v]) * D.pitch[v] + (unsigned)(px - D.bx0[v]) * kTexelBytes);
        } else {
          off = -2 - (py * (a.W + 2) + px) * a.Cp;   // not staged: float offset inside the view's map
          direct = true;
        }
      }
      s_w[v][voxA] = tw[e];
      s_off[v][voxA] = off;
    }
    unsigned pitch[kViewChunk];
#pragma unroll
    for (int v = 0; v < kViewChunk; ++v) pitch[v] = D.pitch[v];
    const bool slow = __any_sync(0xffffffffu, direct);   // also orders the table writes before the reads below

    // ---- gather + aggregate: 4 z steps of run g ----
    float val[4][kCPL];
    if (A.dbg == 2) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int k = 0; k < kCPL; ++k) val[j][k] = s_w[0][j * 4 + g].x;
    } else if (!slow) {
      // every tap of the warp is staged: branch-free, all loads of a z step in flight together
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int sl = j * 4 + g;
        float s[kViewChunk][kCPL];
        int off[kViewChunk];
        float4 w[kViewChunk], t[kViewChunk][4];
#pragma unroll
        for (int v = 0; v < kViewChunk; ++v) {
          off[v] = (v < nv) ? s_off[v][sl] : -1;
          w[v] = (v < nv) ? s_w[v][sl] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int vp = 0; vp < kViewChunk; vp += LT_TMA_VGROUP) {   // loads of LT_TMA_VGROUP views in flight together
#pragma unroll
          for (int v = vp; v < vp + LT_TMA_VGROUP; ++v) {
            if (v < nv) {
              const unsigned p = (unsigned)off[v] + qb;
              t[v][0] = lds128_if(p, off[v]);
              t[v][1] = lds128_if(p + kTexelBytes, off[v]);
              t[v][2] = lds128_if(p + pitch[v], off[v]);
              t[v][3] = lds128_if(p + pitch[v] + kTexelBytes, off[v]);
            }
          }
#pragma unroll
          for (int v = vp; v < vp + LT_TMA_VGROUP; ++v) {
            if (v < nv) {
              blend4(t[v][0], t[v][1], t[v][2], t[v][3], w[v], s[v]);
#pragma unroll
              for (int k = 0; k < kCPL; ++k) s[v][k] = off[v] >= 0 ? s[v][k] : 0.f;   // zeroed views take part with 0
            } else {
#pragma unroll
              for (int k = 0; k < kCPL; ++k) s[v][k] = (AGG == kAggSoftmax) ? kNoView : 0.f;
            }
          }
        }
        aggregate<AGG>(s, cw, nv, val[j]);
      }
    } else {
      // some tap of the warp is not staged (no box, box too large, rounding beyond the guard band): per-tap choice
#pragma unroll 1
      for (int j = 0; j < 4; ++j) {
        const int sl = j * 4 + g;
        float s[kViewChunk][kCPL];
#pragma unroll
        for (int v = 0; v < kViewChunk; ++v) {
          const float fill = (AGG == kAggSoftmax && v >= nv) ? kNoView : 0.f;
#pragma unroll
          for (int k = 0; k < kCPL; ++k) s[v][k] = fill;
          if (v < nv) {
            const int off = s_off[v][sl];
            if (off >= 0) {
              const unsigned p = (unsigned)off + qb;
              const float4 t00 = lds128(p), t01 = lds128(p + kTexelBytes);
              const float4 t10 = lds128(p + pitch[v]), t11 = lds128(p + pitch[v] + kTexelBytes);
              blend4(t00, t01, t10, t11, s_w[v][sl], s[v]);
            } else if (off < -1) {
              gather4(a.feat + ((size_t)b * a.V + v) * viewStride + q * kCPL, -2 - off, s_w[v][sl], a.Cp, rowStride, s[v]);
            }
          }
        }
        float r[kCPL];
        aggregate<AGG>(s, cw, nv, r);
#pragma unroll
        for (int jj = 0; jj < 4; ++jj)
          if (jj == j) {
#pragma unroll
            for (int k = 0; k < kCPL; ++k) val[jj][k] = r[k];
          }
      }
    }
    // the staged boxes and this warp's tap table are consumed: hand the slot back before the stores
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&S.empty[slot]));
#ifdef LT_TMA_PROF
    const long long ps3 = clock64();
    pG += ps3 - ps2;
#endif

    if (i0 < a.N0 && vi1 < a.N1) {
      float* ob = a.out + (size_t)b * a.C * a.nvox + ((long long)i0 * a.N1 + vi1) * a.N2 + zb;
#pragma unroll
      for (int k = 0; k < kCPL; ++k) {
        const int c = q * kCPL + k;
        if (c < a.C) stg128_cs(ob + (size_t)c * a.nvox, val[0][k], val[1][k], val[2][k], val[3][k]);
      }
    }
#ifdef LT_TMA_PROF
    pS += clock64() - ps3;
#endif
  }
#ifdef LT_TMA_PROF
  if (lane == 0 && A.prof) { atomicAdd(A.prof + 7, (unsigned long long)pA); atomicAdd(A.prof + 8, (unsigned long long)pG); atomicAdd(A.prof + 9, (unsigned long long)pS); }
  if (lane == 0 && A.prof) { atomicAdd(A.prof + 5, (unsigned long long)pw0); atomicAdd(A.prof + 6, (unsigned long long)(clock64() - pstart)); }
#endif
}

template <int AGG, bool SPEC>
__global__ void __launch_bounds__(kTmaThreads, 1) k_unproject_fwd_tma(const TmaArgs<SPEC> A) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  TmaSmem& S = *reinterpret_cast<TmaSmem*>(smem_raw);
  const unsigned ring_addr = (smem_u32(smem_raw) + (unsigned)sizeof(TmaSmem) + 127u) & ~127u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kTmaSlots; ++s) {
      mbar_init(smem_u32(&S.full[s]), 1);
      mbar_init(smem_u32(&S.empty[s]), kTmaConsumers);
    }
#pragma unroll
    for (int s = 0; s < kBoxQ; ++s) {
      mbar_init(smem_u32(&S.bfull[s]), 1);
      mbar_init(smem_u32(&S.bempty[s]), 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (warp < kTmaConsumers) consumer<AGG, SPEC>(A, S, warp, lane);
  else if (warp == kTmaConsumers) issue_warp<SPEC>(A, S, ring_addr, lane);
  else box_warp<SPEC>(A, S, lane);
}

constexpr size_t kTmaSmemBytes = sizeof(TmaSmem) + 128 + kRingBytes;
static_assert(kTmaSmemBytes <= 227 * 1024, "shared-memory budget of one CTA on sm_100");

template <bool SPEC>
cudaError_t launch_tma(const UnprojArgs& ua, const CoordSpecArg* spec, int agg, cudaStream_t st) {
  TmaArgs<SPEC> A;
  A.k = make_kargs<SPEC>(ua, spec);
  A.mt0 = (ua.N0 + kTmaT0 - 1) / kTmaT0;
  A.mt1 = (ua.N1 + kTmaT1 - 1) / kTmaT1;
  A.mt2 = ua.N2 / kTmaT2;
  const long long nt = (long long)ua.B * A.mt0 * A.mt1 * A.mt2;
  if (nt > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
  A.ntiles = (int)nt;
  const char* dbg = getenv("LT_B200_TMA_DBG");
  A.dbg = dbg ? atoi(dbg) : 0;
#ifdef LT_TMA_PROF
  static unsigned long long* prof = nullptr;
  static int calls = 0;
  if (!prof) { cudaMalloc(&prof, 128); }
  cudaMemsetAsync(prof, 0, 128, st);
  A.prof = prof;
#endif
  const int grid = (int)(nt < sm_count() ? nt : sm_count());
#define LT_GO(AGG_)                                                                          \
  do {                                                                                       \
    LT_SMEM_OPT_IN_ONCE((k_unproject_fwd_tma<AGG_, SPEC>), kTmaSmemBytes);                   \
    k_unproject_fwd_tma<AGG_, SPEC><<<grid, kTmaThreads, kTmaSmemBytes, st>>>(A);            \
  } while (0)
  switch (agg) {
    case kAggSum: LT_GO(kAggSum); break;
    case kAggConf: LT_GO(kAggConf); break;
    case kAggMax: LT_GO(kAggMax); break;
    case kAggSoftmax: LT_GO(kAggSoftmax); break;
    default: return cudaErrorInvalidValue;
  }
#undef LT_GO
#ifdef LT_TMA_PROF
  if (++calls == 20) {
    unsigned long long h[16];
    cudaStreamSynchronize(st);
    cudaMemcpy(h, prof, 128, cudaMemcpyDeviceToHost);
    const double tiles = (double)nt, ctas = grid;
    fprintf(stderr, "[tma prof] per tile, cycles: box wait_bempty %.0f total %.0f | issue wait_bfull %.0f wait_space %.0f total %.0f | consumer(avg of %d warps) wait_full %.0f total %.0f = phaseA %.0f + wait + addr/gather %.0f + stores %.0f\n",
            h[0] / tiles, h[1] / tiles, h[2] / tiles, h[3] / tiles, h[4] / tiles, kTmaConsumers, h[5] / tiles / kTmaConsumers, h[6] / tiles / kTmaConsumers, h[7] / tiles / kTmaConsumers, h[8] / tiles / kTmaConsumers, h[9] / tiles / kTmaConsumers);
    (void)ctas;
  }
#endif
  return cudaGetLastError();
}

}  // namespace

bool unproject_fwd_tma_eligible(const UnprojArgs& a) {
  return a.Cp == 32 && a.V >= 1 && a.V <= kViewChunk && a.N2 % kTmaT2 == 0 && a.N2 > 0 &&
         (reinterpret_cast<uintptr_t>(a.out) & 15u) == 0 && (reinterpret_cast<uintptr_t>(a.feat) & 15u) == 0 &&
         (long long)(a.H + 2) * (a.W + 2) * a.Cp < 0x3fffffffLL;
}

cudaError_t launch_unproject_fwd_tma(const UnprojArgs& a, const CoordSpecArg* spec, int agg, cudaStream_t st) {
  return spec != nullptr ? launch_tma<true>(a, spec, agg, st) : launch_tma<false>(a, nullptr, agg, st);
}

}  // namespace lt
