/*
 * lt_b200.h — C ABI of the B200 (sm_100a) geometric hot path of learnable-triangulation.
 *
 * The reference (blufzzz/learnable-triangulation-pytorch) has no FFI layer: the hot path is three
 * module-level Python functions. This header is the boundary a binding would bind instead of them:
 *
 *   lt_unproject_*   replaces  mvn/utils/op.py:99-163     unproject_heatmaps (+ multiview.py:80-101 projection)
 *   lt_softargmax_*  replaces  mvn/utils/op.py:84-96      integrate_tensor_3d_with_coordinates
 *   lt_softargmax2d_* replaces mvn/utils/op.py:11-47      integrate_tensor_2d
 *   lt_volce_*        replaces mvn/models/loss.py:52-80   VolumetricCELoss.forward
 *   lt_coord_volumes  replaces mvn/models/triangulation.py:306-341  (coordinate-volume construction)
 *   lt_pwconv_*       replaces mvn/models/triangulation.py:238-240,344-346  (process_features 1x1 convolution)
 *   lt_dlt_*         replaces  mvn/utils/multiview.py:132-174  triangulate_batch_of_points /
 *                                                              triangulate_point_from_multiple_views_linear_torch
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous fp32 (fp64 where the type says so) on the current
 *     CUDA device; the library never allocates, never synchronises and never touches host memory;
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *   - inputs are borrowed and never written; outputs are fully overwritten unless stated otherwise;
 *   - return value: 0 on success, a cudaError_t (>0) if a launch failed, LT_ERR_INVALID (-1) for
 *     bad arguments. lt_last_error() returns a thread-local message for the last non-zero return.
 *     Nothing throws across this boundary.
 *   - there is no CPU implementation behind any entry point.
 *   - alignment: every pointer must be aligned to its element type (4 bytes for float/int, 8 for double). Pointers
 *     from cudaMalloc / torch's allocator are 256-byte aligned and take the vector paths; a pointer that is only
 *     element-aligned (a view with an odd storage offset) is legal and takes scalar paths: the launchers test the
 *     16/32-byte alignment of every tensor they access with vector instructions. Exceptions, which are REQUIRED to
 *     be 16-byte aligned and rejected with LT_ERR_INVALID otherwise: proj (rows are read as float4), the hwcp
 *     buffers of K0/K1/K7 (texels are read and reduced as float4) and every workspace.
 */
#ifndef LT_B200_H
#define LT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LT_ABI_VERSION 2   /* 2: + the *_spec entry points (coordinates computed in the consuming kernel) */
#define LT_ERR_INVALID (-1)

/* view aggregation, op.py:147-161 (`conf` covers every method name starting with "conf") */
enum lt_agg { LT_AGG_SUM = 0, LT_AGG_CONF = 1, LT_AGG_MAX = 2, LT_AGG_SOFTMAX = 3 };

int lt_abi_version(void);
const char *lt_last_error(void);

/* ---- K0: feature layout ------------------------------------------------------------------------
 * The reference hands (B,V,C,H,W) channel-planar maps to F.grid_sample (op.py:114-134). The gather
 * kernels want one tap = one contiguous run of channels, so K0 re-tiles to the library's gather
 * layout "hwcp": (B·V, H+2, W+2, Cp) channels-last, Cp = lt_padded_channels(C) >= C (zero filled),
 * with a one-texel ZERO border around every map (grid_sample's zeros padding as data instead of
 * per-tap bounds logic). Callers treat it as an opaque buffer of lt_feat_layout_elems() floats.
 * lt_feat_to_chw is the inverse, used for grad_heatmaps; it drops border and padding. */
int lt_padded_channels(int C);
size_t lt_feat_layout_elems(int BV, int C, int H, int W);
int lt_feat_to_hwc(const float *src_chw, float *dst_hwcp, int BV, int C, int H, int W, void *stream);
int lt_feat_to_chw(const float *src_hwcp, float *dst_chw, int BV, int C, int H, int W, void *stream);

/* ---- K1: fused unproject + view aggregation  (op.py:99-163) ------------------------------------
 * feat_hwc: (B*V) maps in the hwcp layout of K0   proj (B,V,3,4)   coord (B,N0,N1,N2,3)   conf (B,V,C) or NULL
 * out (B,C,N0,N1,N2).  align_corners: 0 = F.grid_sample default of torch>=1.3 (what the reference
 * runs as today), 1 = torch 1.0.1 behaviour (requirements.txt:20).
 * Semantics: p=[x y z 1]·Pᵀ; voxels with p.z<=0 contribute 0 for that view (op.py:121,138);
 * gx=2(u/H-.5), gy=2(v/W-.5) (op.py:128-129, H for x and W for y as written); bilinear, zero padding;
 * aggregation per enum lt_agg, zeroed views take part in max/softmax with value 0. */
int lt_unproject_fwd(const float *feat_hwc, const float *proj, const float *coord, const float *conf,
                     float *out, int B, int V, int C, int H, int W, int N0, int N1, int N2,
                     int agg, int align_corners, void *stream);

/* Backward of the above w.r.t. heatmaps and vol_confidences (proj/coord never need grad:
 * triangulation.py:277,283). grad_feat_hwc (hwcp layout, lt_feat_layout_elems floats) MUST be zero-filled by the caller: the
 * kernel scatter-adds into it with vector reductions (non-deterministic summation order).
 * grad_conf (B,V,C) may be NULL (required non-NULL only for LT_AGG_CONF if wanted); it is
 * fully overwritten. conf_ws: scratch of lt_unproject_bwd_ws_bytes() bytes (may be NULL when
 * grad_conf is NULL). */
size_t lt_unproject_bwd_ws_bytes(int B, int V, int C, int N0, int N1, int N2);
int lt_unproject_bwd(const float *grad_out, const float *feat_hwc, const float *proj,
                     const float *coord, const float *conf, float *grad_feat_hwc, float *grad_conf,
                     void *conf_ws, int B, int V, int C, int H, int W, int N0, int N1, int N2,
                     int agg, int align_corners, void *stream);

/* ---- K2: volumetric soft-argmax  (op.py:84-96) -------------------------------------------------
 * vol (B,J,nvox) scaled by `multiplier` on the fly (triangulation.py:353 does volumes*multiplier
 * before the call; pass 1.0 for the verbatim signature).  coord (B,nvox,3).
 * softmax=1: prob=softmax over nvox; softmax=0: prob=relu(vol) un-normalised (op.py:91).
 * coords (B,J,3) = sum_i prob_i * coord_i (op.py:94).  prob (B,J,nvox).
 *
 * The op is split where the z-slab sharded variant needs its one exchange:
 *   partial : per (b,j) tuple stats[5] = {m, S=sum e^(x-m), sum e^(x-m)*xyz} over THIS buffer
 *             (softmax=0: {0, 1, sum relu(x)*xyz}).  ws: lt_softargmax_ws_bytes() scratch.
 *   (multi-GPU only: merge the tuples of all slabs — python does it with one all_gather)
 *   finalize: prob_i = e^(x_i-m)/S and coords = stats[2:5]/S from (possibly merged) stats.
 * lt_softargmax_fwd = partial + finalize on one device; stats (B,J,5) is kept for backward/debug. */
size_t lt_softargmax_ws_bytes(int B, int J, long long nvox);
int lt_softargmax_partial(const float *vol, const float *coord, float *stats, void *ws, int B, int J,
                          long long nvox, int softmax, float multiplier, void *stream);
int lt_softargmax_finalize(const float *vol, const float *stats, float *coords, float *prob, int B,
                           int J, long long nvox, int softmax, float multiplier, void *stream);
int lt_softargmax_fwd(const float *vol, const float *coord, float *coords, float *prob, float *stats,
                      void *ws, int B, int J, long long nvox, int softmax, float multiplier,
                      void *stream);
/* log-sum-exp merge of R slabs' tuples: parts (R,B*J,5) -> merged (B*J,5); deterministic order. */
int lt_softargmax_merge(const float *parts, float *merged, int R, int BJ, int softmax, void *stream);

/* lt_softargmax_fwd without caller-owned scratch, for callers that cannot allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernels and bit-identical results as lt_softargmax_fwd; `stats` and `ws` live in one
 * static device array of LT_NOALLOC_SCRATCH_BYTES inside the library (every GPU that loads the library holds it until
 * exit).  Shapes whose scratch (256-rounded B*J*5*4 + lt_softargmax_ws_bytes(B, J, nvox)) exceeds that maximum are
 * rejected with LT_ERR_INVALID and a message; 8 MiB covers BASELINE config 3 (16 x 17 joints at 64^3: 6.3 MB)
 * and one item x 17 joints at 128^3 (3.1 MB), not 256^3 (25 MB for one item).  Calls no cudaMalloc / cudaFree / allocating routine.  Every call shares the array
 * and finds the last call's contents in it (nothing in it needs clearing between calls): calls that overlap on the
 * GPU - unordered streams, or two host threads on one device - race; calls ordered on one stream are safe. */
#define LT_NOALLOC_SCRATCH_BYTES (8u << 20)
int lt_softargmax_fwd_noalloc(const float *vol, const float *coord, float *coords, float *prob, int B, int J,
                              long long nvox, int softmax, float multiplier, void *stream);

/* Backward.  grad_coords (B,J,3); grad_prob (B,J,nvox) or NULL (the dense gradient that arrives
 * through the second return value, loss.py:76); prob = saved second output; coords = saved first
 * output (global coordinates, also in the slab-sharded case); grad_vol (B,J,nvox) includes the
 * `multiplier` factor.  softmax=1: g_i = p_i*(G_i - sum_k p_k G_k), G_i = gp_i + gc·xyz_i.
 * dots (B,J) or NULL = sum_k p_k*gp_k: lt_softargmax_bwd_dot fills it from a dense grad_prob over THIS buffer
 * (slab-sharded callers all-reduce it between the two calls); with grad_prob == NULL it carries a sparse
 * upstream gradient (see lt_volce_bwd_sparse). Required when softmax && grad_prob. */
int lt_softargmax_bwd_dot(const float *prob, const float *grad_prob, float *dots, void *ws, int B,
                          int J, long long nvox, void *stream);
int lt_softargmax_bwd(const float *prob, const float *coord, const float *coords,
                      const float *grad_coords, const float *grad_prob, const float *dots,
                      float *grad_vol, int B, int J, long long nvox, int softmax, float multiplier,
                      void *stream);

/* ---- K4: 2-D soft-argmax of joint heat maps  (op.py:11-47 integrate_tensor_2d) -------------------
 * hm (BJ, H, W) scaled by `multiplier` on the fly (triangulation.py:164 does heatmaps*heatmap_multiplier
 * before the call).  softmax=1: prob = softmax over H*W, coords (BJ,2) = (sum p*x, sum p*y) in heat-map
 * pixels.  softmax=0: prob = relu(hm) un-normalised (op.py:26) while coords = (sum p*x, sum p*y)/sum p
 * (op.py:39-41; 0/0 = nan as in the reference).  The step that feeds lt_dlt_fwd in the algebraic net. */
int lt_softargmax2d_fwd(const float *hm, float *coords, float *prob, int BJ, int H, int W, int softmax,
                        float multiplier, void *stream);
/* grad_coords (BJ,2); grad_prob (BJ,H,W) or NULL; prob/coords = saved outputs; grad_hm includes `multiplier`. */
int lt_softargmax2d_bwd(const float *prob, const float *coords, const float *grad_coords,
                        const float *grad_prob, float *grad_hm, int BJ, int H, int W, int softmax,
                        float multiplier, void *stream);

/* ---- K6: coordinate volumes  (mvn/models/triangulation.py:306-341, mvn/utils/volumetric.py:87-114) ----
 * spec (B, LT_COORD_SPEC) fp32 per item: pos[3] (cuboid corner), step[3] (side/(N-1)), centre[3] (base point),
 * rot[9] (row-major rotation about the base point).  out (B,N0,N1,N2,3):
 *   out[b,i0,i1,i2,:] = rot * ((pos + step*(i0,i1,i2)) - centre) + centre   in the reference's fp32 operation order.
 * transfer=1 applies the CMU->Human3.6M swap of triangulation.py:336-339: out[i0,i1,i2] = cv[i0,i2,N1-1-i1]. */
#define LT_COORD_SPEC 18
int lt_coord_volumes(const float *spec, float *out, int B, int N0, int N1, int N2, int transfer,
                     void *stream);

/* ---- K1 / K2 with the coordinates computed IN the kernel   ------------------------------------
 * The reference builds coord_volumes in python (triangulation.py:280-341) and every op reads 12 bytes per voxel of it.
 * These variants take the K6 spec (B, LT_COORD_SPEC) + `transfer` INSTEAD of the coordinate tensor and evaluate the same
 * fp32 arithmetic per voxel (one definition, shared with lt_coord_volumes): outputs are bit-identical to running the
 * tensor variants on lt_coord_volumes' output.  The volume extents are explicit (K2's tensor variants only need nvox):
 * nvox = N0*N1*N2 <= 2^31-1, transfer needs N1 == N2.  Everything else as documented for the tensor variants. */
int lt_unproject_fwd_spec(const float *feat_hwc, const float *proj, const float *coord_spec, int transfer,
                          const float *conf, float *out, int B, int V, int C, int H, int W, int N0, int N1, int N2,
                          int agg, int align_corners, void *stream);
int lt_unproject_bwd_spec(const float *grad_out, const float *feat_hwc, const float *proj, const float *coord_spec,
                          int transfer, const float *conf, float *grad_feat_hwc, float *grad_conf, void *conf_ws,
                          int B, int V, int C, int H, int W, int N0, int N1, int N2, int agg, int align_corners,
                          void *stream);
int lt_softargmax_partial_spec(const float *vol, const float *coord_spec, int transfer, float *stats, void *ws, int B,
                               int J, int N0, int N1, int N2, int softmax, float multiplier, void *stream);
int lt_softargmax_fwd_spec(const float *vol, const float *coord_spec, int transfer, float *coords, float *prob,
                           float *stats, void *ws, int B, int J, int N0, int N1, int N2, int softmax,
                           float multiplier, void *stream);
int lt_softargmax_bwd_spec(const float *prob, const float *coord_spec, int transfer, const float *coords,
                           const float *grad_coords, const float *grad_prob, const float *dots, float *grad_vol,
                           int B, int J, int N0, int N1, int N2, int softmax, float multiplier, void *stream);

/* ---- K5: volumetric cross-entropy loss  (mvn/models/loss.py:52-80 VolumetricCELoss) --------------
 * coord (B,nvox,3)  prob (B,J,nvox) = second output of lt_softargmax_fwd  kp_gt (B,J,3)  validity (B,J)
 * loss (1) = sum_{b,j} validity * -log(prob[b,j,i*] + 1e-6) / (B*J),  i* = first argmin_i |coord_i - kp_gt_bj|
 * (fp32 distances in the reference's operation order, loss.py:66-70).  idx (B,J) int32 receives i* for the
 * backward.  ws: lt_volce_ws_bytes bytes of scratch.  No host synchronisation (loss.py:71 copies the indices to
 * the host).  B*J > 0 is required: the reference divides by n_losses = B*J (loss.py:80). */
size_t lt_volce_ws_bytes(int B, int J, long long nvox);
int lt_volce_fwd(const float *coord, const float *prob, const float *kp_gt, const float *validity,
                 float *loss, int *idx, void *ws, int B, int J, long long nvox, void *stream);
/* grad_loss (1) on the device; grad_prob (B,J,nvox) is fully overwritten: zeros except the B*J voxels idx. */
int lt_volce_bwd(const float *prob, const int *idx, const float *validity, const float *grad_loss,
                 float *grad_prob, int B, int J, long long nvox, void *stream);

/* The same gradient kept sparse, for callers that own both ops (no (B,J,nvox) gradient tensor between K5 and K2):
 *   1. lt_volce_bwd_sparse        -> vals (B,J) = the non-zeros, dots (B,J) = prob[idx]*vals
 *   2. lt_softargmax_bwd(..., grad_prob = NULL, dots, ...)   (dots without grad_prob = sparse upstream gradient)
 *   3. lt_softargmax_bwd_add_sparse adds the B*J voxels' own G_i term into grad_vol. */
int lt_volce_bwd_sparse(const float *prob, const int *idx, const float *validity,
                        const float *grad_loss, float *vals, float *dots, int B, int J, long long nvox,
                        void *stream);
int lt_softargmax_bwd_add_sparse(const float *prob, const int *idx, const float *vals, float *grad_vol,
                                 int B, int J, long long nvox, int softmax, float multiplier,
                                 void *stream);

/* ---- K7: process_features, the 1x1 convolution in front of K1  (mvn/models/triangulation.py:238-240, 344-346) ----
 * The reference applies nn.Conv2d(256, 32, 1) to the backbone's features (B*V, Cin, H, W) and hands the result
 * to unproject_heatmaps. lt_pwconv_fwd computes out[n,c,y,x] = bias[c] + sum_k weight[c,k] * x[n,k,y,x] and writes it
 * DIRECTLY in the gather layout of K0 (out_hwcp: lt_feat_layout_elems(BV, Cout, H, W) floats, zero border and zero
 * padded channels included), so lt_feat_to_hwc is not needed between the two.  weight (Cout, Cin) row-major = the
 * (Cout, Cin, 1, 1) tensor of nn.Conv2d; bias (Cout) or NULL.  fp32 FMA chain over k ascending (deterministic).
 * Backward: grad_hwcp is the gradient map lt_unproject_bwd scatters into (same layout; border and padded channels are
 * ignored).  grad_x (BV, Cin, H, W) fully overwritten; grad_weight (Cout, Cin) and grad_bias (Cout, may be NULL) fully
 * overwritten, summed in a fixed order (deterministic).  ws: lt_pwconv_ws_bytes() bytes of scratch, 16-byte aligned
 * (forward and backward-weight; may be shared, contents are not kept between calls).
 * At the reference's shape (Cout <= 32, Cin = 128 or 256, H*W a multiple of 32, x 16-byte and out_hwcp 32-byte aligned)
 * all three run on the tcgen05 tensor cores with a three-product TF32 split: operands are truncated to hi (19 bits) +
 * lo and the lo*lo product is dropped, measured 1-2e-6 relative (max-norm) against an fp64 convolution, i.e. the level
 * of an fp32 FMA chain but NOT bit-identical to one; other shapes and alignments run fp32 FMA kernels (same results to
 * 2e-6). */
size_t lt_pwconv_ws_bytes(int BV, int Cin, int Cout, int H, int W);
int lt_pwconv_fwd(const float *x, const float *weight, const float *bias, float *out_hwcp, void *ws,
                  int BV, int Cin, int Cout, int H, int W, void *stream);
int lt_pwconv_bwd_data(const float *grad_hwcp, const float *weight, float *grad_x, int BV, int Cin,
                       int Cout, int H, int W, void *stream);
int lt_pwconv_bwd_weight(const float *x, const float *grad_hwcp, float *grad_weight, float *grad_bias,
                         void *ws, int BV, int Cin, int Cout, int H, int W, void *stream);

/* ---- K3: batched confidence-weighted DLT  (multiview.py:132-174) -------------------------------
 * proj (B,V,3,4)  pts (B,V,J,2)  conf (B,V,J) or NULL (=> ones, multiview.py:147-148)
 * out (B,J,3) fp32.  saved (B,J,LT_DLT_SAVED) fp64: right singular vectors + singular values of
 * A (2V x 4), kept for backward.  Rows A[v,r,:] = conf_v*(pt[v,r]*P[v,2,:] - P[v,r,:]); the solve is
 * a QR of A followed by a one-sided Jacobi SVD of the 4x4 R in fp64; X = h[:3]/h[3] with h the right
 * singular vector of the smallest singular value. V <= LT_DLT_MAX_VIEWS. */
#define LT_DLT_SAVED 20
#define LT_DLT_MAX_VIEWS 64
int lt_dlt_fwd(const float *proj, const float *pts, const float *conf, float *out, double *saved,
               int B, int V, int J, void *stream);
/* grad_pts (B,V,J,2), grad_conf (B,V,J) or NULL; fully overwritten. */
int lt_dlt_bwd(const float *grad_out, const float *proj, const float *pts, const float *conf,
               const double *saved, float *grad_pts, float *grad_conf, int B, int V, int J,
               void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LT_B200_H */
