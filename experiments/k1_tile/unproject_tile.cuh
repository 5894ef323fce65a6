// unproject_tile.cuh — the shared-memory-staged ("tile") variant of K1: common machinery of forward and backward.
//
// One CTA = one 8 x 8 x ZT block of voxels x ONE 16-channel group of the g16p feature layout (layout.cu).
//   phase A  every thread projects its voxels through the (<= 4) views of the current view chunk once
//            (make_taps: projection + grid_sample index arithmetic), the per-view bounding boxes of the taps are
//            reduced with redux.sync + shared atomics, and the boxes are packed into a shared-memory pool;
//   load     each image row of each box is ONE contiguous run in g16p, moved by one cp.async.bulk (TMA engine,
//            completion on an mbarrier) issued by one thread per row: no LSU traffic, no L1 tags, no misses;
//   gather   a warp owns 8 voxel columns and walks z; 4 lanes x float4 cover the 64-byte texel of one voxel.
//            The two voxels that share a quarter-warp read opposite halves of the 128-byte bank window:
//            phase A swaps the x-order of the odd voxel's taps (and weights) whenever the parities collide, so
//            every LDS.128 wavefront is conflict free;
//   when the boxes of a block do not fit the pool the block is processed in thinner z slabs (ZT -> ZT/2 -> .. 1);
//   if even one layer does not fit, that slab samples global memory directly (slow path, same descriptors).
#pragma once
#include "unproject.cuh"

namespace lt {

constexpr int kTileThreads = 256;
constexpr int kTileRuns = 64;       // 8 x 8 voxel columns per CTA
constexpr int kTexelBytes = kGroupCh * 4;  // 64
constexpr unsigned kDescInvalid = 0xffffffffu;

struct TileArgs {
  const float* feat;   // g16p (B*V, G, H+2, W+2, 16)
  const float* proj;   // (B,V,3,4)
  const float* coord;  // (B,nvox,3)
  const float* conf;   // (B,V,C) or null
  float* out;          // fwd: (B,C,nvox)
  const float* gout;   // bwd: (B,C,nvox)
  float* gfeat;        // bwd: g16p, pre-zeroed
  float* gconf_part;   // bwd: partial sums of grad_conf, or null
  int B, V, C, G, H, W, N0, N1, N2;
  int nt0, nt1, nt2;
  int align_corners;
  int pool_texels;     // capacity of the box pool in 64-byte texels
  int gpc;             // channel groups handled by one CTA (G: phase A shared by all groups; 1: one group per CTA)
  long long nvox;
};

// ---- PTX wrappers -------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
}
// every thread of the CTA arrives once per fill; threads that issue copies add their bytes first (expect_tx without
// arrive), so the phase cannot complete before all copies have been issued and have landed: no CTA barrier needed
__device__ __forceinline__ void mbar_expect_tx(unsigned mbar, unsigned bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned mbar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(mbar), "r"(parity)
      : "memory");
}
// one contiguous global run -> shared memory through the TMA engine (UBLKCP), bytes % 16 == 0, both sides 16-B aligned
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned mbar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(mbar)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ float4 lds128(unsigned addr) {
  float4 r;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
  return r;
}
__device__ __forceinline__ uint2 lds64u(unsigned addr) {
  uint2 r;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(addr));
  return r;
}

// ---- packed fp32 (FFMA2 / FMUL2 / FADD2 on sm_100): two IEEE fp32 lanes per instruction ----------------------------
struct f2 { unsigned long long v; };
__device__ __forceinline__ f2 pack2(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(f2 a, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
  return r;
}
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}

// ---- shared-memory carve-up -----------------------------------------------------------------------------------------
// [ proj 64*3 float4 | coord 64*(ZT*3+1) float | ctl ints | mbar | desc offs uint2[4][64*ZT] | desc wts float4[4][64*ZT]
//   | pool ]
struct TileCtl {
  int bb[kViewChunk][4];    // xmin, ymin, xmax, ymax of the north-west taps (valid voxels only)
  int box[kViewChunk][4];   // byte offset inside the pool (or -1: no valid voxel), width in texels, xlo, ylo
  int rows_before[kViewChunk + 1];  // prefix sum of box heights (row -> copy assignment)
  int fits;                 // 1: boxes fit the pool, 0: they do not
  int total_bytes;          // bytes of one fill (all boxes of one channel group)
  int nbuf;                 // 2: the pool holds two fills (double-buffered channel groups), else 1
  int pad[1];
};

template <int ZT>
struct TileSmem {
  static constexpr int kVox = kTileRuns * ZT;
  static constexpr int kCoordPitch = ZT * 3 + 1;  // odd: conflict-free column reads
  static constexpr size_t off_proj = 0;
  static constexpr size_t off_coord = off_proj + sizeof(float4) * kMaxViewsSmem * 3;
  static constexpr size_t off_ctl = off_coord + ((sizeof(float) * kTileRuns * kCoordPitch + 15) / 16) * 16;
  static constexpr size_t off_mbar = off_ctl + ((sizeof(TileCtl) + 15) / 16) * 16;
  static constexpr size_t off_doff = off_mbar + 16;  // two mbarriers
  static constexpr size_t off_dwts = off_doff + sizeof(uint2) * kViewChunk * kVox;
  static constexpr size_t off_pool = ((off_dwts + sizeof(float4) * kViewChunk * kVox + 127) / 128) * 128;
  static size_t bytes(int pool_texels) { return off_pool + (size_t)pool_texels * kTexelBytes; }
};

struct TileId3 {
  int b, cg, t0, t1, t2;
};
// blockIdx.x = (((b*(G/gpc) + cg/gpc)*nt0 + t0)*nt1 + t1)*nt2 + t2 ; cg = first group of this CTA
__device__ __forceinline__ TileId3 decode_tile3(const TileArgs& a) {
  TileId3 r;
  int t = blockIdx.x;
  r.t2 = t % a.nt2; t /= a.nt2;
  r.t1 = t % a.nt1; t /= a.nt1;
  r.t0 = t % a.nt0; t /= a.nt0;
  const int ngrp = a.G / a.gpc;
  r.cg = (t % ngrp) * a.gpc;
  r.b = t / ngrp;
  return r;
}


// ---- per-CTA context ------------------------------------------------------------------------------------------------
template <int ZT>
struct TileCtx {
  using SM = TileSmem<ZT>;
  unsigned char* smem;
  float4* s_proj;
  float* s_coord;
  TileCtl* ctl;
  unsigned mbar;       // shared-space address of mbarrier 0 (mbarrier 1 at +8)
  unsigned doff, dwts; // shared-space addresses of the descriptor arrays
  unsigned pool;       // shared-space address of the box pool
  TileId3 T;
  int i0_base, i1_base, z_base;
  float invH, invW;
  unsigned phase[2];   // mbarrier phase parities
};

template <int ZT>
__device__ __forceinline__ TileCtx<ZT> tile_prologue(const TileArgs& a, unsigned char* smem) {
  using SM = TileSmem<ZT>;
  TileCtx<ZT> c;
  c.smem = smem;
  c.s_proj = reinterpret_cast<float4*>(smem + SM::off_proj);
  c.s_coord = reinterpret_cast<float*>(smem + SM::off_coord);
  c.ctl = reinterpret_cast<TileCtl*>(smem + SM::off_ctl);
  c.mbar = smem_u32(smem + SM::off_mbar);
  c.doff = smem_u32(smem + SM::off_doff);
  c.dwts = smem_u32(smem + SM::off_dwts);
  c.pool = smem_u32(smem + SM::off_pool);
  c.T = decode_tile3(a);
  c.i0_base = c.T.t0 * 8; c.i1_base = c.T.t1 * 8; c.z_base = c.T.t2 * ZT;
  c.invH = 1.f / (float)a.H; c.invW = 1.f / (float)a.W;
  c.phase[0] = 0; c.phase[1] = 0;
  // projection rows of this batch item
  for (int i = threadIdx.x; i < a.V * 3 && i < kMaxViewsSmem * 3; i += blockDim.x)
    c.s_proj[i] = __ldg(reinterpret_cast<const float4*>(a.proj) + (size_t)c.T.b * a.V * 3 + i);
  // voxel coordinates of the block: 64 columns x ZT*3 contiguous floats each (columns outside the volume -> 0)
  const float* cb = a.coord + (size_t)c.T.b * a.nvox * 3;
  for (int i = threadIdx.x; i < kTileRuns * ZT * 3; i += blockDim.x) {
    const int run = i / (ZT * 3), r = i % (ZT * 3);
    const int i0 = c.i0_base + (run >> 3), i1 = c.i1_base + (run & 7);
    float v = 0.f;
    if (i0 < a.N0 && i1 < a.N1) v = __ldg(cb + (((long long)i0 * a.N1 + i1) * a.N2 + c.z_base) * 3 + r);
    c.s_coord[run * SM::kCoordPitch + r] = v;
  }
  if (threadIdx.x == 0) {
    mbar_init(c.mbar, kTileThreads);
    mbar_init(c.mbar + 8, kTileThreads);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  return c;
}

// Fill pool buffer `buf` with the packed boxes of channel group cg: one contiguous image row of one box per issuing
// thread (cp.async.bulk, completion on mbarrier `buf`). Every thread of the CTA calls it and arrives once. The caller
// guarantees that no thread still reads that buffer (CTA barrier) before the call.
template <int ZT>
__device__ __forceinline__ void tile_fill(const TileArgs& a, const TileCtx<ZT>& c, int v0, int cg, int buf) {
  const int Hp = a.H + 2, Wp = a.W + 2;
  const int total_rows = c.ctl->rows_before[kViewChunk];
  const unsigned mbar = c.mbar + 8u * buf;
  const unsigned dst0 = c.pool + (unsigned)(buf * c.ctl->total_bytes);
  if (threadIdx.x < total_rows) fence_proxy_async();
  for (int r = threadIdx.x; r < total_rows; r += kTileThreads) {
    int vv = 0;
#pragma unroll
    for (int t = 1; t < kViewChunk; ++t) vv += (r >= c.ctl->rows_before[t]) ? 1 : 0;
    const int row = r - c.ctl->rows_before[vv];
    const int w = c.ctl->box[vv][1], xlo = c.ctl->box[vv][2], ylo = c.ctl->box[vv][3];
    const float* src = a.feat + ((((size_t)c.T.b * a.V + v0 + vv) * a.G + cg) * Hp + (ylo + row)) * Wp * kGroupCh +
                       (size_t)xlo * kGroupCh;
    mbar_expect_tx(mbar, (unsigned)(w * kTexelBytes));
    bulk_g2s(dst0 + c.ctl->box[vv][0] + row * w * kTexelBytes, src, (unsigned)(w * kTexelBytes), mbar);
  }
  mbar_arrive(mbar);
}

template <int ZT>
__device__ __forceinline__ void tile_wait(TileCtx<ZT>& c, int buf) {
  mbar_wait(c.mbar + 8u * buf, c.phase[buf]);
  c.phase[buf] ^= 1;
}

// Phase A + box packing + bulk loads + descriptor table for the z layers [zs, zs+zext) of the block and the views
// [v0, v0+nv). Returns 1 when the boxes are in shared memory, 0 when they did not fit (descriptors then address the
// g16p group plane in global memory: byte offsets relative to the plane of the view). Every thread of the CTA calls it.
//   descriptor of (view vv, layer z, column run):  index = z*64 + run
//     offs.x / offs.y : byte offset of the FIRST / SECOND x-tap of the upper row (second row = + pitch of the box)
//     wts             : weights in the order (first, second) x (upper, lower); all zero for an invalid voxel, whose
//                       offsets alias its quarter-warp partner (or the box origin / the zero border of the plane):
//                       always readable, never a bank conflict, so the gather loops carry no validity branches
template <int ZT, bool TRY_SMEM>
__device__ __forceinline__ int tile_stage(const TileArgs& a, TileCtx<ZT>& c, int v0, int nv, int zs, int zext) {
  using SM = TileSmem<ZT>;
  constexpr int kMaxPer = (kTileRuns * ZT) / kTileThreads;  // voxels per thread when zext == ZT
  const int tid = threadIdx.x, lane = tid & 31;
  const int nvox_slab = kTileRuns * zext;

  // ---- projection of this thread's voxels (registers) ----
  int xy[kMaxPer][kViewChunk];      // x0+1 | (y0+1) << 16, or -1
  float4 wt[kMaxPer][kViewChunk];
  if (tid < kViewChunk * 4) (&c.ctl->bb[0][0])[tid] = (tid & 2) ? -0x7fffffff : 0x7fffffff;  // min slots / max slots
  __syncthreads();  // also: every reader of the previous slab's pool / descriptors is done
#pragma unroll
  for (int k = 0; k < kMaxPer; ++k) {
    const int vid = tid + k * kTileThreads;
    const bool in_slab = vid < nvox_slab;
    const int zl = zs + vid / kTileRuns, run = vid % kTileRuns;
    const int i0 = c.i0_base + (run >> 3), i1 = c.i1_base + (run & 7);
    const bool ok = in_slab && i0 < a.N0 && i1 < a.N1 && (c.z_base + zl) < a.N2;
    float x = 0.f, y = 0.f, z = 0.f;
    if (ok) {
      const float* cp = c.s_coord + run * SM::kCoordPitch + zl * 3;
      x = cp[0]; y = cp[1]; z = cp[2];
    }
#pragma unroll
    for (int vv = 0; vv < kViewChunk; ++vv) {
      xy[k][vv] = -1;
      wt[k][vv] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (vv < nv && ok) {
        const int v = v0 + vv;
        const float4 r0 = c.s_proj[v * 3], r1 = c.s_proj[v * 3 + 1], r2 = c.s_proj[v * 3 + 2];  // V <= kMaxViewsSmem
        int x0, y0;
        if (make_taps<false>(x, y, z, r0, r1, r2, a.H, a.W, c.invH, c.invW, a.align_corners, x0, y0, wt[k][vv]))
          xy[k][vv] = (x0 + 1) | ((y0 + 1) << 16);
      }
    }
  }
  // ---- bounding boxes: thread -> warp (redux) -> CTA (shared atomics) ----
#pragma unroll
  for (int vv = 0; vv < kViewChunk; ++vv) {
    if (vv < nv) {
      int xmin = 0x7fffffff, ymin = 0x7fffffff, xmax = -0x7fffffff, ymax = -0x7fffffff;
#pragma unroll
      for (int k = 0; k < kMaxPer; ++k)
        if (xy[k][vv] >= 0) {
          const int px = xy[k][vv] & 0xffff, py = xy[k][vv] >> 16;
          xmin = min(xmin, px); xmax = max(xmax, px); ymin = min(ymin, py); ymax = max(ymax, py);
        }
      xmin = __reduce_min_sync(0xffffffffu, xmin); ymin = __reduce_min_sync(0xffffffffu, ymin);
      xmax = __reduce_max_sync(0xffffffffu, xmax); ymax = __reduce_max_sync(0xffffffffu, ymax);
      if (lane == 0 && xmax >= 0) {
        atomicMin(&c.ctl->bb[vv][0], xmin); atomicMin(&c.ctl->bb[vv][1], ymin);
        atomicMax(&c.ctl->bb[vv][2], xmax); atomicMax(&c.ctl->bb[vv][3], ymax);
      }
    }
  }
  __syncthreads();
  // ---- pack the boxes (one thread; <= 4 views) ----
  if (tid == 0) {
    int used = 0, rows = 0;
    c.ctl->rows_before[0] = 0;
    for (int vv = 0; vv < kViewChunk; ++vv) {
      int base = -1, w = 0, h = 0, xlo = 0, ylo = 0;
      if (vv < nv && c.ctl->bb[vv][2] >= 0) {
        xlo = c.ctl->bb[vv][0]; ylo = c.ctl->bb[vv][1];
        w = c.ctl->bb[vv][2] - xlo + 2;   // taps x0 and x0+1 (padded coordinates: always inside the plane)
        h = c.ctl->bb[vv][3] - ylo + 2;
        base = used * kTexelBytes;
        used += w * h;
      }
      c.ctl->box[vv][0] = base; c.ctl->box[vv][1] = w; c.ctl->box[vv][2] = xlo; c.ctl->box[vv][3] = ylo;
      rows += h;
      c.ctl->rows_before[vv + 1] = rows;
    }
    const int fits = TRY_SMEM && used <= a.pool_texels;
    c.ctl->fits = fits;
    c.ctl->total_bytes = used * kTexelBytes;
    c.ctl->nbuf = (fits && a.gpc > 1 && 2 * used <= a.pool_texels) ? 2 : 1;
  }
  __syncthreads();
  const int fits = c.ctl->fits;
  const int Wp = a.W + 2;
  if (fits) {  // first fill(s): channel group 0 (and 1 when the pool is double-buffered); the caller waits
    tile_fill<ZT>(a, c, v0, c.T.cg, 0);
    if (c.ctl->nbuf == 2) tile_fill<ZT>(a, c, v0, c.T.cg + 1, 1);
  }
  // ---- descriptor table ----
#pragma unroll
  for (int k = 0; k < kMaxPer; ++k) {
    const int vid = tid + k * kTileThreads;
    if (vid < nvox_slab) {
      const int didx = (zs * kTileRuns) + vid;  // layer-major: z*64 + run
#pragma unroll
      for (int vv = 0; vv < kViewChunk; ++vv) {
        if (vv < nv) {
          const int base = c.ctl->box[vv][0], w = c.ctl->box[vv][1], xlo = c.ctl->box[vv][2], ylo = c.ctl->box[vv][3];
          unsigned offA, offB;
          float4 wq = wt[k][vv];
          const bool valid = xy[k][vv] >= 0;
          if (fits) {
            offA = valid ? (unsigned)(base + (((xy[k][vv] >> 16) - ylo) * w + ((xy[k][vv] & 0xffff) - xlo)) * kTexelBytes)
                         : (unsigned)max(base, 0);
            // quarter-warp partner = the voxel of column run^1 in the same layer = lane^4 ... (run = vid%64, so
            // partner thread = tid^1): exchange the address of the first tap
            const unsigned pA = __shfl_xor_sync(0xffffffffu, offA, 1);
            const bool pvalid = __shfl_xor_sync(0xffffffffu, valid ? 1 : 0, 1) != 0;
            if (!valid) {
              offA = pvalid ? pA : offA;  // alias the partner: same address = broadcast
              offB = offA;
            } else {
              offB = offA + kTexelBytes;
              // odd column of the pair swaps its x order when both first taps sit in the same 64-byte bank half
              if ((tid & 1) && pvalid && (((pA ^ offA) & kTexelBytes) == 0) && pA != offA) {
                const unsigned t = offA; offA = offB; offB = t;
                wq = make_float4(wq.y, wq.x, wq.w, wq.z);
              }
            }
          } else {
            // global mode: byte offsets inside the g16p plane of (view, group)
            offA = valid ? (unsigned)(((xy[k][vv] >> 16) * Wp + (xy[k][vv] & 0xffff)) * kTexelBytes) : 0u;
            offB = valid ? offA + kTexelBytes : 0u;
          }
          uint2* dO = reinterpret_cast<uint2*>(c.smem + SM::off_doff) + vv * SM::kVox + didx;
          float4* dW = reinterpret_cast<float4*>(c.smem + SM::off_dwts) + vv * SM::kVox + didx;
          *dO = make_uint2(offA, offB);
          *dW = wq;
        }
      }
    }
  }
  __syncthreads();  // descriptors visible
  return fits;
}

}  // namespace lt
