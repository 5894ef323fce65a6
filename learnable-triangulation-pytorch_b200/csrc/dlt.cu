// dlt.cu — K3: batched confidence-weighted DLT triangulation, forward and backward.
//
// Replaces the B·J python iterations of mvn/utils/multiview.py:162-174, each of which builds the 2V x 4
// system A (multiview.py:150-152) and calls torch.svd (multiview.py:154), by one launch with one warp per
// (batch item, joint). Lanes own views (v = lane, lane+32).
//
// Numerics: the columns of A differ by ~1e3 in scale (column 3 carries K·t), so forming AᵀA in fp32 is
// useless and even the reference's own fp32 svd is only within ~0.3 mm of its fp64 run.
// Here the rows are formed in fp64 from the fp32 inputs, A is reduced to its 4x4 triangular factor R by a
// modified Gram-Schmidt QR whose inner products are warp all-reductions (RᵀR = AᵀA without squaring the
// condition number), and R's right singular vectors come from one-sided Jacobi sweeps in fp64 registers.
// X = h[:3]/h[3] for the singular vector h of the smallest singular value (sign cancels, multiview.py:156).
//
// Backward (implicit differentiation of the smallest eigenpair of M = AᵀA, equal to SvdBackward restricted
// to the last right singular vector):  w = sum_{k!=3} v_k (v_k.g_h)/(s3^2 - s_k^2),
// grad_A = (A h) wᵀ + (A w) hᵀ, then through the row construction to points and confidences. Each lane
// finishes its own views: no cross-lane traffic in the backward.
#include "common.cuh"
#include "kernels.h"

namespace lt {

constexpr int kDltWarps = 4;       // problems per CTA
constexpr int kDltViewsPerLane = 2;  // V <= 64
constexpr int kDltSaved = 20;      // 16 (V, row-major, column 3 = h) + 4 singular values

struct DltRows {
  double a[2 * kDltViewsPerLane][4];
};

// rows of this lane's views, fp64:  conf * (pt * P[2,:] - P[r,:])
__device__ __forceinline__ void dlt_build_rows(const float* __restrict__ proj_b, const float* __restrict__ pts_b,
                                               const float* __restrict__ conf_b, int V, int J, int j, int lane,
                                               DltRows& R, double (&cf)[kDltViewsPerLane],
                                               double (&px)[kDltViewsPerLane], double (&py)[kDltViewsPerLane]) {
#pragma unroll
  for (int kv = 0; kv < kDltViewsPerLane; ++kv) {
    const int v = lane + 32 * kv;
    cf[kv] = 0.0; px[kv] = 0.0; py[kv] = 0.0;
#pragma unroll
    for (int c = 0; c < 4; ++c) { R.a[2 * kv][c] = 0.0; R.a[2 * kv + 1][c] = 0.0; }
    if (v < V) {
      const float* P = proj_b + (size_t)v * 12;
      const double x = (double)__ldg(pts_b + ((size_t)v * J + j) * 2);
      const double y = (double)__ldg(pts_b + ((size_t)v * J + j) * 2 + 1);
      const double cw = conf_b ? (double)__ldg(conf_b + (size_t)v * J + j) : 1.0;
      cf[kv] = cw; px[kv] = x; py[kv] = y;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const double p0 = (double)__ldg(P + c), p1 = (double)__ldg(P + 4 + c), p2 = (double)__ldg(P + 8 + c);
        R.a[2 * kv][c] = cw * (x * p2 - p0);
        R.a[2 * kv + 1][c] = cw * (y * p2 - p1);
      }
    }
  }
}

// One-sided Jacobi SVD of a 4x4 matrix G (columns rotated in place), accumulating V. On return the
// column of V belonging to the smallest singular value has been swapped into column 3.
__device__ __forceinline__ void jacobi_svd4(double (&G)[4][4], double (&Vm)[4][4], double (&sig)[4]) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) Vm[i][c] = (i == c) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    double off = 0.0;
#pragma unroll
    for (int p = 0; p < 3; ++p) {
#pragma unroll
      for (int q = p + 1; q < 4; ++q) {
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          al = fma(G[i][p], G[i][p], al);
          be = fma(G[i][q], G[i][q], be);
          ga = fma(G[i][p], G[i][q], ga);
        }
        const double lim = 1e-16 * sqrt(al * be);
        if (fabs(ga) > lim && ga != 0.0) {
          off = fmax(off, fabs(ga) / fmax(sqrt(al * be), 1e-300));
          const double zeta = (be - al) / (2.0 * ga);
          const double t = copysign(1.0, zeta) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          const double c = 1.0 / sqrt(1.0 + t * t), s = c * t;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const double gp = G[i][p], gq = G[i][q];
            G[i][p] = c * gp - s * gq;
            G[i][q] = s * gp + c * gq;
            const double vp = Vm[i][p], vq = Vm[i][q];
            Vm[i][p] = c * vp - s * vq;
            Vm[i][q] = s * vp + c * vq;
          }
        }
      }
    }
    if (off < 1e-15) break;
  }
  int imin = 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    double n = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) n = fma(G[i][c], G[i][c], n);
    sig[c] = sqrt(n);
  }
#pragma unroll
  for (int c = 1; c < 4; ++c)
    if (sig[c] < sig[imin]) imin = c;
  // swap column imin <-> 3 (static indexing only)
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    if (c == imin) {
#pragma unroll
      for (int i = 0; i < 4; ++i) { const double t = Vm[i][c]; Vm[i][c] = Vm[i][3]; Vm[i][3] = t; }
      const double t = sig[c]; sig[c] = sig[3]; sig[3] = t;
    }
  }
}

__global__ void __launch_bounds__(kDltWarps * 32) k_dlt_fwd(const float* __restrict__ proj,
                                                            const float* __restrict__ pts,
                                                            const float* __restrict__ conf, float* __restrict__ out,
                                                            double* __restrict__ saved, int BJ, int V, int J) {
  const int prob = blockIdx.x * kDltWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (prob >= BJ) return;
  const int b = prob / J, j = prob % J;
  DltRows R;
  double cf[kDltViewsPerLane], px[kDltViewsPerLane], py[kDltViewsPerLane];
  dlt_build_rows(proj + (size_t)b * V * 12, pts + (size_t)b * V * J * 2, conf ? conf + (size_t)b * V * J : nullptr,
                 V, J, j, lane, R, cf, px, py);

  // modified Gram-Schmidt: R factor (upper triangular) with warp all-reduced inner products
  double T[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) T[i][c] = 0.0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    double d[4];
#pragma unroll
    for (int c = k; c < 4; ++c) {
      double s = 0.0;
#pragma unroll
      for (int r = 0; r < 2 * kDltViewsPerLane; ++r) s = fma(R.a[r][k], R.a[r][c], s);
      d[c] = warp_sum(s);
    }
    const double rkk = sqrt(fmax(d[k], 0.0));
    T[k][k] = rkk;
    if (rkk > 0.0) {
#pragma unroll
      for (int c = k + 1; c < 4; ++c) {
        T[k][c] = d[c] / rkk;
        const double f = d[c] / d[k];
#pragma unroll
        for (int r = 0; r < 2 * kDltViewsPerLane; ++r) R.a[r][c] = fma(-f, R.a[r][k], R.a[r][c]);
      }
    }
  }

  double Vm[4][4], sig[4];
  jacobi_svd4(T, Vm, sig);

  if (lane == 0) {
    const double h3 = Vm[3][3];
    out[(size_t)prob * 3 + 0] = (float)(Vm[0][3] / h3);
    out[(size_t)prob * 3 + 1] = (float)(Vm[1][3] / h3);
    out[(size_t)prob * 3 + 2] = (float)(Vm[2][3] / h3);
  }
  if (saved != nullptr && lane < kDltSaved) {
    // lanes 0..15 write V row-major, lanes 16..19 the singular values (static register indexing)
    double val = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (lane == i * 4 + c) val = Vm[i][c];
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (lane == 16 + c) val = sig[c];
    saved[(size_t)prob * kDltSaved + lane] = val;
  }
}

__global__ void __launch_bounds__(kDltWarps * 32) k_dlt_bwd(const float* __restrict__ gout,
                                                            const float* __restrict__ proj,
                                                            const float* __restrict__ pts,
                                                            const float* __restrict__ conf,
                                                            const double* __restrict__ saved,
                                                            float* __restrict__ gpts, float* __restrict__ gconf,
                                                            int BJ, int V, int J) {
  const int prob = blockIdx.x * kDltWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (prob >= BJ) return;
  const int b = prob / J, j = prob % J;
  const float* proj_b = proj + (size_t)b * V * 12;
  DltRows R;
  double cf[kDltViewsPerLane], px[kDltViewsPerLane], py[kDltViewsPerLane];
  dlt_build_rows(proj_b, pts + (size_t)b * V * J * 2, conf ? conf + (size_t)b * V * J : nullptr, V, J, j, lane, R,
                 cf, px, py);

  const double* sv = saved + (size_t)prob * kDltSaved;
  double Vm[4][4], lam[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) Vm[i][c] = sv[i * 4 + c];
#pragma unroll
  for (int c = 0; c < 4; ++c) lam[c] = sv[16 + c] * sv[16 + c];

  const double h[4] = {Vm[0][3], Vm[1][3], Vm[2][3], Vm[3][3]};
  const double g0 = (double)__ldg(gout + (size_t)prob * 3), g1 = (double)__ldg(gout + (size_t)prob * 3 + 1),
               g2 = (double)__ldg(gout + (size_t)prob * 3 + 2);
  // X = h[:3]/h3
  double gh[4];
  gh[0] = g0 / h[3]; gh[1] = g1 / h[3]; gh[2] = g2 / h[3];
  gh[3] = -(g0 * h[0] + g1 * h[1] + g2 * h[2]) / (h[3] * h[3]);
  double w[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const double den = lam[3] - lam[k];
    double dp = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) dp = fma(Vm[i][k], gh[i], dp);
    const double f = (den != 0.0) ? dp / den : 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = fma(Vm[i][k], f, w[i]);
  }

#pragma unroll
  for (int kv = 0; kv < kDltViewsPerLane; ++kv) {
    const int v = lane + 32 * kv;
    if (v >= V) continue;
    const float* P = proj_b + (size_t)v * 12;
    double gx = 0.0, gy = 0.0, gc = 0.0;
    double Ah[2] = {0.0, 0.0}, Aw[2] = {0.0, 0.0};
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        Ah[r] = fma(R.a[2 * kv + r][c], h[c], Ah[r]);
        Aw[r] = fma(R.a[2 * kv + r][c], w[c], Aw[r]);
      }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const double p0 = (double)__ldg(P + c), p1 = (double)__ldg(P + 4 + c), p2 = (double)__ldg(P + 8 + c);
      const double gA0 = Ah[0] * w[c] + Aw[0] * h[c];
      const double gA1 = Ah[1] * w[c] + Aw[1] * h[c];
      gx = fma(gA0, p2, gx);
      gy = fma(gA1, p2, gy);
      gc = fma(gA0, px[kv] * p2 - p0, gc);
      gc = fma(gA1, py[kv] * p2 - p1, gc);
    }
    gpts[(((size_t)b * V + v) * J + j) * 2] = (float)(cf[kv] * gx);
    gpts[(((size_t)b * V + v) * J + j) * 2 + 1] = (float)(cf[kv] * gy);
    if (gconf != nullptr) gconf[((size_t)b * V + v) * J + j] = (float)gc;
  }
}

cudaError_t launch_dlt_fwd(const float* proj, const float* pts, const float* conf, float* out, double* saved,
                           int B, int V, int J, cudaStream_t st) {
  const int BJ = B * J;
  if (BJ == 0) return cudaSuccess;
  k_dlt_fwd<<<(BJ + kDltWarps - 1) / kDltWarps, kDltWarps * 32, 0, st>>>(proj, pts, conf, out, saved, BJ, V, J);
  return cudaGetLastError();
}

cudaError_t launch_dlt_bwd(const float* grad_out, const float* proj, const float* pts, const float* conf,
                           const double* saved, float* grad_pts, float* grad_conf, int B, int V, int J,
                           cudaStream_t st) {
  const int BJ = B * J;
  if (BJ == 0 || V == 0) return cudaSuccess;
  k_dlt_bwd<<<(BJ + kDltWarps - 1) / kDltWarps, kDltWarps * 32, 0, st>>>(grad_out, proj, pts, conf, saved, grad_pts,
                                                                         grad_conf, BJ, V, J);
  return cudaGetLastError();
}

}  // namespace lt
