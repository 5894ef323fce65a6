// unproject.cuh — geometry shared by the K1 forward and backward kernels.
//
// Work decomposition (both directions)
//   * a lane owns kCPL = 4 consecutive channels; LPV = Cp/4 lanes (8 for the 32-channel features of
//     triangulation.py:239) cover one texel, so one tap is ONE contiguous Cp*4-byte access per lane group:
//     LDG.E.128 per lane = one full 128-byte line per voxel in the forward gather, and one
//     REDG.E.ADD.F32x4 per lane = four complete 32-byte sectors per voxel in the backward scatter (the L2
//     reduction unit retires one sector per cycle per slice, so half-filled sectors would halve its rate);
//   * a lane group walks a run of kRun = 8 consecutive voxels along the fastest volume axis, so that the
//     8 results of one channel leave as one full 32-byte sector (STG.E.256);
//   * a warp owns a 2x4 block of runs and works on G = 32/LPV of them per round (R = 8/G rounds);
//   * a CTA (8 warps, 4x2) therefore always covers a compact 8x8x8 block of the volume: its image footprint
//     (~17x17 texels per view at Human3.6M geometry) is what has to stay in L1;
//   * projection (mvn/utils/multiview.py:80-101 + op.py:121-130 + the index arithmetic of
//     F.grid_sample) is done ONCE per (voxel, view) by one lane ("phase A") and handed to the gather
//     lanes through a per-warp shared-memory descriptor table (offset + 4 tap weights).
#pragma once
#include "common.cuh"
#include "coordspec.cuh"
#include "kernels.h"

namespace lt {

constexpr int kRun = 8;           // voxels per lane-group run
constexpr int kCPL = 4;           // channels per lane
constexpr int kWarpsPerCta = 8;   // arranged 4 (axis 0) x 2 (axis 1)
constexpr int kWarpRuns0 = 2, kWarpRuns1 = 4;  // runs owned by a warp: 2 (axis 0) x 4 (axis 1)
constexpr int kTile0 = 4 * kWarpRuns0, kTile1 = 2 * kWarpRuns1;  // CTA tile 8 x 8 (x 8 along the run axis)
constexpr int kViewChunk = 4;     // views resolved per phase-A round
constexpr int kMaxViewsSmem = 32; // projection matrices staged per CTA
constexpr int kMaxLanesPerVoxel = 32;  // => channel blocks of 128


template <int LPV>
struct TileGeom {
  static constexpr int G = 32 / LPV;                              // runs in flight per warp
  static constexpr int R = (kWarpRuns0 * kWarpRuns1) / G;         // rounds per warp
  static constexpr int kSlots = G * kRun;                         // descriptor slots per warp, view and round
  // run index (0..7) of lane group g in round r -> offsets inside the warp's 2x4 block
  __device__ static __forceinline__ int run0(int r, int g) { return (r * G + g) / kWarpRuns1; }
  __device__ static __forceinline__ int run1(int r, int g) { return (r * G + g) % kWarpRuns1; }
};

struct UnprojArgs {
  const float* feat;   // (B,V,H,W,Cp)
  const float* proj;   // (B,V,3,4)
  const float* coord;  // (B,nvox,3)
  const float* conf;   // (B,V,C) or null
  float* out;          // fwd: (B,C,nvox)
  const float* gout;   // bwd: (B,C,nvox)
  float* gfeat;        // bwd: (B,V,H,W,Cp), pre-zeroed
  float* gconf_part;   // bwd: (B,tiles,V,C) per-CTA partial sums of grad_conf, or null
  int B, V, C, Cp, H, W, N0, N1, N2;
  int nt0, nt1, nt2, ncb;  // tile grid and channel blocks
  int align_corners;
  long long nvox;
};

// Kernel argument block. The SPEC instantiations (voxel centres computed from the per-item spec, coordspec.cuh) carry
// two more fields; the tensor-path kernels keep exactly UnprojArgs (MEASURED: the two extra fields in a shared struct
// were enough to make ptxas spill in the 64-register forward kernel: 0.497 -> 0.525 ms at cfg3).
template <bool SPEC>
struct KArgs : UnprojArgs {};
template <>
struct KArgs<true> : UnprojArgs {
  const float* spec;   // (B,18): `coord` is not read
  int transfer;        // CMU -> Human3.6M axis swap
};
template <bool SPEC>
inline KArgs<SPEC> make_kargs(const UnprojArgs& a, const CoordSpecArg* spec);
template <>
inline KArgs<false> make_kargs<false>(const UnprojArgs& a, const CoordSpecArg*) {
  KArgs<false> k;
  static_cast<UnprojArgs&>(k) = a;
  return k;
}
template <>
inline KArgs<true> make_kargs<true>(const UnprojArgs& a, const CoordSpecArg* spec) {
  KArgs<true> k;
  static_cast<UnprojArgs&>(k) = a;
  k.spec = spec->spec;
  k.transfer = spec->transfer;
  return k;
}
__device__ __forceinline__ int kargs_transfer(const KArgs<false>&) { return 0; }
__device__ __forceinline__ int kargs_transfer(const KArgs<true>& a) { return a.transfer; }

// voxel centre (vi0, vi1, vz) of item b: read from the coordinate volume, or evaluated from the item's spec staged in
// shared memory (bit-identical to K6's tensor)
// SPEC is a compile-time switch: the tensor path must not carry the spec path's registers (measured: a run-time branch
// cost the 64-register forward kernel spills and 8 % of its time).
template <bool SPEC>
__device__ __forceinline__ void voxel_xyz(const KArgs<SPEC>& a, const float* coord_b, const float* s_spec, int vi0, int vi1,
                                          int vz, float& x, float& y, float& z) {
  if (SPEC) {
    float o[3];
    coord_of_voxel(s_spec, kargs_transfer(a), a.N1, vi0, vi1, vz, o);
    x = o[0]; y = o[1]; z = o[2];
  } else {
    const float* cp = coord_b + (((long long)vi0 * a.N1 + vi1) * a.N2 + vz) * 3;
    x = __ldg(cp); y = __ldg(cp + 1); z = __ldg(cp + 2);
  }
}
// the item's spec -> shared memory; call before the CTA-wide barrier that publishes s_proj
__device__ __forceinline__ void stage_unproj_spec(const KArgs<true>& a, int b, float* s_spec) {
  if (threadIdx.x < kCoordSpec) s_spec[threadIdx.x] = __ldg(a.spec + (size_t)b * kCoordSpec + threadIdx.x);
}

// One (voxel, view) tap set. Mirrors, in order:
//   multiview.py:96   [x y z 1] @ P.t()   (k-sequential fp32 accumulation like an sgemm inner loop)
//   op.py:121-124     invalid = z<=0 ; u=x/z ; v=y/z
//   op.py:128-129     gx=2(u/H-.5) ; gy=2(v/W-.5)      (tensor/python-int = multiply by fp32 reciprocal on CUDA)
//   ATen grid_sampler_unnormalize + bilinear corner weights, zero padding.
// Returns false when the view contributes exactly 0 (behind camera, or all four taps outside); otherwise
// x0 in [-1, W-1], y0 in [-1, H-1] is the north-west tap and wts = (nw, ne, sw, se) with out-of-image taps zeroed.
// FAST_DIV: u = px * (1/pz) with a Newton-refined MUFU.RCP instead of two IEEE divisions. MEASURED on B200: the 64^3
// softmax volume then differs from the reference by 2.6e-5 relative (bar: 1e-5), so nothing uses it; kept as a record.
template <bool FAST_DIV = false>
__device__ __forceinline__ bool make_taps(float x, float y, float z, const float4& r0, const float4& r1,
                                          const float4& r2, int H, int W, float invH, float invW, int align_corners,
                                          int& x0, int& y0, float4& wts) {
  float px = x * r0.x; px = fmaf(y, r0.y, px); px = fmaf(z, r0.z, px); px += r0.w;
  float py = x * r1.x; py = fmaf(y, r1.y, py); py = fmaf(z, r1.z, py); py += r1.w;
  float pz = x * r2.x; pz = fmaf(y, r2.y, pz); pz = fmaf(z, r2.z, pz); pz += r2.w;
  x0 = 0; y0 = 0;
  wts = make_float4(0.f, 0.f, 0.f, 0.f);
  if (pz <= 0.f) return false;
  float u, v;
  if (FAST_DIV) {
    float r = fast_rcp(pz);
    r = fmaf(fmaf(-pz, r, 1.f), r, r);
    u = px * r; v = py * r;
  } else {
    u = __fdiv_rn(px, pz); v = __fdiv_rn(py, pz);
  }
  float gx = 2.f * (u * invH - 0.5f);
  float gy = 2.f * (v * invW - 0.5f);
  float ix, iy;
  if (align_corners) {
    ix = ((gx + 1.f) * 0.5f) * (float)(W - 1);
    iy = ((gy + 1.f) * 0.5f) * (float)(H - 1);
  } else {
    ix = fmaf(gx + 1.f, (float)W, -1.f) * 0.5f;
    iy = fmaf(gy + 1.f, (float)H, -1.f) * 0.5f;
  }
  // anything outside (-1, W) x (-1, H) has no in-bounds tap (also catches inf/nan)
  if (!(ix > -1.f && ix < (float)W && iy > -1.f && iy < (float)H)) return false;
  float fx0 = floorf(ix), fy0 = floorf(iy);
  float ax1 = (fx0 + 1.f) - ix, ax0 = ix - fx0;  // weights of column x0 and x0+1
  float ay1 = (fy0 + 1.f) - iy, ay0 = iy - fy0;
  x0 = (int)fx0; y0 = (int)fy0;
  bool okx0 = x0 >= 0, okx1 = x0 + 1 < W, oky0 = y0 >= 0, oky1 = y0 + 1 < H;
  wts.x = (okx0 && oky0) ? ax1 * ay1 : 0.f;  // nw
  wts.y = (okx1 && oky0) ? ax0 * ay1 : 0.f;  // ne
  wts.z = (okx0 && oky1) ? ax1 * ay0 : 0.f;  // sw
  wts.w = (okx1 && oky1) ? ax0 * ay0 : 0.f;  // se
  return true;
}

// descriptor: float offset of the north-west tap inside one hwcp map (padded texel (y0+1, x0+1); all four taps exist
// thanks to the zero border), or -1.
__device__ __forceinline__ void make_desc(float x, float y, float z, const float4& r0, const float4& r1,
                                          const float4& r2, int H, int W, int Cp, float invH, float invW,
                                          int align_corners, int& off, float4& wts) {
  int x0, y0;
  const bool ok = make_taps(x, y, z, r0, r1, r2, H, W, invH, invW, align_corners, x0, y0, wts);
  off = ok ? ((y0 + 1) * (W + 2) + (x0 + 1)) * Cp : -1;
}

constexpr float kNoView = -1.0e30f;  // sample value of a view that does not exist (softmax weight underflows to 0)

struct TileId {
  int b, cb, t0, t1, t2, lin;
};
// blockIdx.x = ((b*ncb + cb)*nt0 + t0)*nt1*nt2 + t1*nt2 + t2
__device__ __forceinline__ TileId decode_tile(const UnprojArgs& a) {
  TileId r;
  int t = blockIdx.x;
  r.t2 = t % a.nt2; t /= a.nt2;
  r.t1 = t % a.nt1; t /= a.nt1;
  r.t0 = t % a.nt0; t /= a.nt0;
  r.cb = t % a.ncb;
  r.b = t / a.ncb;
  r.lin = (r.t0 * a.nt1 + r.t1) * a.nt2 + r.t2;
  return r;
}

// phase A: descriptors of this warp's round-r voxels x nv views -> per-warp smem table.
// zsel < 0: all 8 z steps, slot = j*G + g (G*8 slots, one per lane for C=32); zsel >= 0: only that z step, slot = g.
template <int LPV, bool SPEC = false>
__device__ __forceinline__ void build_descs(const KArgs<SPEC>& a, int b, int i0_warp, int i1_warp, int z0, int r,
                                            const float4* s_proj, float4 (*s_w)[TileGeom<LPV>::kSlots],
                                            int (*s_off)[TileGeom<LPV>::kSlots], int v0, int nv, int zsel, int lane,
                                            float invH, float invW, const float* s_spec) {
  // The descriptor offset is relative to the ITEM's first map: view v's maps start v * viewStride floats further on.
  // (MEASURED with ncu's source view: as `base + (size_t)v * viewStride + off` per tap, 64-bit address arithmetic was
  // 17-35 instructions per (voxel, view) in the gather loops, half of their instruction count.)
  const int viewStrideI = (a.H + 2) * (a.W + 2) * a.Cp;
  using TG = TileGeom<LPV>;
  constexpr int G = TG::G;
  const float* coord_b = a.coord + (size_t)b * a.nvox * 3;
  const int nslots = zsel < 0 ? TG::kSlots : G;
  for (int slot = lane; slot < nslots; slot += 32) {
    const int j = zsel < 0 ? slot / G : zsel, gg = slot % G;
    const int vi0 = i0_warp + TG::run0(r, gg), vi1 = i1_warp + TG::run1(r, gg), vz = z0 + j;
    const bool ok = (vi0 < a.N0) && (vi1 < a.N1) && (vz < a.N2);
    float x = 0.f, y = 0.f, z = 0.f;
    if (ok) voxel_xyz<SPEC>(a, coord_b, s_spec, vi0, vi1, vz, x, y, z);
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
        if (off >= 0) off += v * viewStrideI;
      }
      s_w[vv][slot] = w;
      s_off[vv][slot] = off;
    }
  }
}

// an opaque copy of a pointer: the compiler keeps it as a 64-bit register pair instead of re-deriving it from the kernel
// argument plus a 64-bit index at every use (two extra instructions per address in the gather loops)
template <typename T>
__device__ __forceinline__ T* opaque_ptr(T* p) {
  asm volatile("" : "+l"(p));
  return p;
}

// the same with both row pointers prepared by the caller (once per CTA): address arithmetic = 2 x (LEA, LEA.HI.X)
template <int CPT = 0>
__device__ __forceinline__ void gather4_rows(const float* __restrict__ row0, const float* __restrict__ row1, int off,
                                             const float4& w, int Cp, float (&s)[kCPL]) {
  const int cp = CPT > 0 ? CPT : Cp;
  const float* p = row0 + off;
  const float* p2 = row1 + off;
  const float4 t00 = ldg128_nc(p);
  const float4 t01 = ldg128_nc(p + cp);
  const float4 t10 = ldg128_nc(p2);
  const float4 t11 = ldg128_nc(p2 + cp);
  // two channels per FFMA2 (the weight operand is broadcast by the instruction): same nw, ne, sw, se chain per channel
  const f2 wx = pack2(w.x, w.x), wy = pack2(w.y, w.y), wz = pack2(w.z, w.z), ww = pack2(w.w, w.w);
  f2 lo = mul2(pack2(t00.x, t00.y), wx), hi = mul2(pack2(t00.z, t00.w), wx);
  lo = fma2(pack2(t01.x, t01.y), wy, lo); hi = fma2(pack2(t01.z, t01.w), wy, hi);
  lo = fma2(pack2(t10.x, t10.y), wz, lo); hi = fma2(pack2(t10.z, t10.w), wz, hi);
  lo = fma2(pack2(t11.x, t11.y), ww, lo); hi = fma2(pack2(t11.z, t11.w), ww, hi);
  unpack2(lo, s[0], s[1]);
  unpack2(hi, s[2], s[3]);
}

// The two halves of gather4_rows for callers that want the loads of SEVERAL views in flight before the first blend
// (a branch per view serialises them: one L2 round trip per view instead of one per voxel). The loads are predicated on
// off >= 0 and leave t undefined otherwise; blend_taps then returns exactly 0 for that view.
template <int CPT = 0>
__device__ __forceinline__ void issue_taps(const float* __restrict__ row0, const float* __restrict__ row1, int off, int Cp,
                                           float4 (&t)[4]) {
  const int cp = CPT > 0 ? CPT : Cp;
  const float* p = row0 + off;
  const float* p2 = row1 + off;
#define LT_LDG_IF(dst, ptr)                                                                                         \
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ge.s32 p, %5, 0;\n\t@p ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];\n\t}"    \
               : "=f"(dst.x), "=f"(dst.y), "=f"(dst.z), "=f"(dst.w) : "l"(ptr), "r"(off))
  LT_LDG_IF(t[0], p);
  LT_LDG_IF(t[1], p + cp);
  LT_LDG_IF(t[2], p2);
  LT_LDG_IF(t[3], p2 + cp);
#undef LT_LDG_IF
}
__device__ __forceinline__ void blend_taps(const float4 (&t)[4], const float4& w, int off, float (&s)[kCPL]) {
  const f2 wx = pack2(w.x, w.x), wy = pack2(w.y, w.y), wz = pack2(w.z, w.z), ww = pack2(w.w, w.w);
  f2 lo = mul2(pack2(t[0].x, t[0].y), wx), hi = mul2(pack2(t[0].z, t[0].w), wx);
  lo = fma2(pack2(t[1].x, t[1].y), wy, lo); hi = fma2(pack2(t[1].z, t[1].w), wy, hi);
  lo = fma2(pack2(t[2].x, t[2].y), wz, lo); hi = fma2(pack2(t[2].z, t[2].w), wz, hi);
  lo = fma2(pack2(t[3].x, t[3].y), ww, lo); hi = fma2(pack2(t[3].z, t[3].w), ww, hi);
  unpack2(lo, s[0], s[1]);
  unpack2(hi, s[2], s[3]);
#pragma unroll
  for (int k = 0; k < kCPL; ++k) s[k] = off >= 0 ? s[k] : 0.f;
}

// per-lane gather of one tap set: 4 x LDG.E.128 (4 channels each; the LPV lanes of a voxel read one full texel)
// and the bilinear blend, accumulated in the nw, ne, sw, se order of ATen's grid_sampler_2d.
// CPT > 0: the texel pitch is a compile-time constant (an immediate offset of the load instead of a 64-bit add).
template <int CPT = 0>
__device__ __forceinline__ void gather4(const float* __restrict__ fv, int off, const float4& w, int Cp,
                                        int rowStride, float (&s)[kCPL]) {
  gather4_rows<CPT>(fv, fv + rowStride, off, w, Cp, s);
}

// shared-memory home of the item's spec: exists only in the SPEC instantiations (a dummy array in the tensor-path
// kernels was enough to tip the 64-register forward kernel into spilling)
template <bool SPEC>
struct SpecSmem {
  __device__ static __forceinline__ const float* stage(const KArgs<false>&, int) { return nullptr; }
};
template <>
struct SpecSmem<true> {
  __device__ static __forceinline__ const float* stage(const KArgs<true>& a, int b) {
    __shared__ float s_spec[kCoordSpec];
    stage_unproj_spec(a, b, s_spec);
    return s_spec;
  }
};

// softmax over <= 4 views of a channel of one voxel: out = sum_v s_v e^{s_v} / sum_v e^{s_v}, two channels per call.
// Absent views carry kNoView (e -> 0). ex2.approx / rcp.approx are good to ~2 ulp, far inside the 1e-5 bar (round 1
// started with expf + IEEE divisions at 97 % SFU load, then a 3-comparator sorting network that skipped the
// exponential of the maximum; this form trades one more MUFU for half the issue slots).
// Packed forms (FFMA2 / FMUL2 / FADD2) wherever the ISA has one: shift by the maximum of the four views (3 FMNMX per
// channel), all four exponentials evaluated (the SFU has room: 30 % busy), numerator and denominator accumulated as
// pairs. ~15 issue slots per channel instead of ~25.
__device__ __forceinline__ void softmax_mix4_x2(const float (&a)[4], const float (&b)[4], float& out_a, float& out_b) {
  const float ma = fmaxf(fmaxf(a[0], a[1]), fmaxf(a[2], a[3]));
  const float mb = fmaxf(fmaxf(b[0], b[1]), fmaxf(b[2], b[3]));
  const f2 L2 = pack2(kLog2eF, kLog2eF);
  const f2 nm = mul2(pack2(-ma, -mb), L2);            // -m*log2e per channel
  f2 num = pack2(0.f, 0.f), den = pack2(0.f, 0.f);
#pragma unroll
  for (int v = 0; v < 4; ++v) {
    const f2 sv = pack2(a[v], b[v]);
    float ta, tb;
    unpack2(fma2(sv, L2, nm), ta, tb);               // (s - m)*log2e, <= 0
    const f2 e = pack2(fast_ex2(ta), fast_ex2(tb));  // absent views (kNoView) underflow to exactly 0
    den = add2(den, e);
    num = fma2(sv, e, num);
  }
  float na, nb, da, db;
  unpack2(num, na, nb);
  unpack2(den, da, db);
  out_a = na * fast_rcp(da);
  out_b = nb * fast_rcp(db);
}

void fill_unproj_tiling(UnprojArgs& a, int lpv);
int unproj_lpv(int Cp);

}  // namespace lt
