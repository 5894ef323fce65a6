"""Drop-in replacements for the batched linear (DLT) triangulation of ``mvn.utils.multiview`` (reference file
``mvn/utils/multiview.py``): same names, arguments, return values and exceptions; one kernel launch instead of
B*J python iterations with one ``torch.svd`` each.

    triangulate_batch_of_points(proj_matricies_batch, points_batch, confidences_batch=None)
        reference: mvn/utils/multiview.py:162-174
    triangulate_point_from_multiple_views_linear_torch(proj_matricies, points, confidences=None)
        reference: mvn/utils/multiview.py:132-159
"""
import torch

from . import _lib
from ._nvtx import annotate
from .op import _require_cuda, _f32c, _stream

MAX_VIEWS = 64  # LT_DLT_MAX_VIEWS
_SAVED = 20     # LT_DLT_SAVED


class _DLT(torch.autograd.Function):
    @staticmethod
    def forward(ctx, proj, pts, conf):
        L = _lib.lib()
        B, V, J = pts.shape[:3]
        with torch.cuda.device(pts.device):
            st = _stream()
            proj_c, pts_c = _f32c(proj), _f32c(pts)
            conf_c = _f32c(conf) if conf is not None else None
            out = torch.zeros((B, J, 3), device=pts.device, dtype=torch.float32)  # multiview.py:164: always fp32
            saved = torch.empty((B, J, _SAVED), device=pts.device, dtype=torch.float64)
            _lib.check(L.lt_dlt_fwd(_lib.ptr(proj_c), _lib.ptr(pts_c), _lib.ptr(conf_c), _lib.ptr(out),
                                    _lib.ptr(saved), B, V, J, st))
        ctx.save_for_backward(proj_c, pts_c, conf_c, saved)
        ctx.meta = (B, V, J, pts.dtype, conf.dtype if conf is not None else None)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_out):
        L = _lib.lib()
        proj_c, pts_c, conf_c, saved = ctx.saved_tensors
        B, V, J, pts_dtype, conf_dtype = ctx.meta
        need_pts = ctx.needs_input_grad[1]
        need_conf = ctx.needs_input_grad[2] and conf_c is not None
        if not (need_pts or need_conf):
            return None, None, None
        with torch.cuda.device(pts_c.device):
            st = _stream()
            go = _f32c(g_out)
            gpts = torch.zeros((B, V, J, 2), device=pts_c.device, dtype=torch.float32)
            gconf = torch.zeros((B, V, J), device=pts_c.device, dtype=torch.float32) if need_conf else None
            _lib.check(L.lt_dlt_bwd(_lib.ptr(go), _lib.ptr(proj_c), _lib.ptr(pts_c), _lib.ptr(conf_c),
                                    _lib.ptr(saved), _lib.ptr(gpts), _lib.ptr(gconf), B, V, J, st))
        return None, (gpts.to(pts_dtype) if need_pts else None), (gconf.to(conf_dtype) if need_conf else None)


@annotate("lt_b200.triangulate_batch_of_points")
def triangulate_batch_of_points(proj_matricies_batch, points_batch, confidences_batch=None):
    """Confidence-weighted DLT of every (batch item, joint) (reference: multiview.py:162-174).

    proj_matricies_batch (B,V,3,4), points_batch (B,V,J,2), confidences_batch (B,V,J) or None (= ones).
    Returns (B,J,3) float32. Differentiable w.r.t. points and confidences (the Algebraic net trains through it).
    """
    _require_cuda("proj_matricies_batch", proj_matricies_batch)
    _require_cuda("points_batch", points_batch)
    if confidences_batch is not None:
        _require_cuda("confidences_batch", confidences_batch)
    if points_batch.dim() != 4 or points_batch.shape[-1] != 2:
        raise RuntimeError("points_batch must be (B,V,J,2), got {}".format(tuple(points_batch.shape)))
    B, V, J = points_batch.shape[:3]
    if B > 0 and J > 0:
        # multiview.py:143 asserts per item that the two view counts agree
        assert proj_matricies_batch.shape[1] == V
    if tuple(proj_matricies_batch.shape[-2:]) != (3, 4) or proj_matricies_batch.shape[0] < B:
        raise RuntimeError("proj_matricies_batch must be (B,V,3,4), got {}".format(tuple(proj_matricies_batch.shape)))
    if V < 1 or V > MAX_VIEWS:
        raise RuntimeError("lt_b200 DLT supports 1..{} views, got {}".format(MAX_VIEWS, V))
    conf = None
    if confidences_batch is not None:
        conf = confidences_batch[:B].reshape(B, V, J)
    from . import torch_ops
    if torch_ops.available():
        return torch.ops.lt_b200.triangulate_batch_of_points(proj_matricies_batch[:B], points_batch, conf)
    return _DLT.apply(proj_matricies_batch[:B], points_batch, conf)


def triangulate_point_from_multiple_views_linear_torch(proj_matricies, points, confidences=None):
    """One DLT solve (reference: multiview.py:132-159): proj_matricies (V,3,4), points (V,2), confidences (V,) or None.
    Returns (3,) float32."""
    assert len(proj_matricies) == len(points)
    conf = confidences.reshape(1, -1, 1) if confidences is not None else None
    out = triangulate_batch_of_points(proj_matricies.unsqueeze(0), points.reshape(1, -1, 1, 2), conf)
    return out[0, 0]
