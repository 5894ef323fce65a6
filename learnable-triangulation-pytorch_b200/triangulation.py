"""Producer-side layout fusion for the volumetric net (reference file ``mvn/models/triangulation.py``).

``VolumetricTriangulationNet`` passes the backbone's 256-channel features through
``self.process_features = nn.Sequential(nn.Conv2d(256, 32, 1))`` (triangulation.py:238-240, applied at :344-346) and
hands the result to ``op.unproject_heatmaps`` (:349). Here that 1x1 convolution is kernel K7 (``csrc/pwconv.cu``): it
writes K1's gather layout directly, so the layout pass K0 and its inverse drop out of the training step, and its
backward reads the gradient map K1's backward scatters into.

    process_features(features, weight, bias)      functional form
    ProcessFeatures(in_channels=256, out_channels=32)
        drop-in for the reference's ``nn.Sequential(nn.Conv2d(256, 32, 1))``: same parameters, same ``state_dict`` keys
        (``0.weight``, ``0.bias``), so checkpoints of the reference load unchanged
    use_fused_process_features(model)
        switches ``model.process_features`` of an existing (reference) net over in place

The value returned is an ordinary float32 tensor of shape (B*V, Cout, H, W) whose strides walk the interior of the
gather-layout buffer; ``.view(batch_size, n_views, ...)`` (triangulation.py:346) keeps that, and ``op.unproject_heatmaps``
recognises it. Any other consumer sees a normal (non-contiguous) tensor. CUDA only, no CPU path.
"""
import torch
from torch import nn

from . import _lib
from ._nvtx import annotate
from . import op as _op


class _ProcessFeatures(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        L = _lib.lib()
        BV, Cin, H, W = x.shape
        Cout = weight.shape[0]
        with torch.cuda.device(x.device):
            st = _op._stream()
            xc = _op._f32c(x)
            wc = _op._f32c(weight).reshape(Cout, Cin)
            bc = _op._f32c(bias) if bias is not None else None
            feat = torch.empty((L.lt_feat_layout_elems(BV, Cout, H, W),), device=x.device, dtype=torch.float32)
            if BV > 0:
                ws = torch.empty((max(L.lt_pwconv_ws_bytes(BV, Cin, Cout, H, W), 4) // 4,), device=x.device,
                                 dtype=torch.float32)
                _lib.check(L.lt_pwconv_fwd(_lib.ptr(xc), _lib.ptr(wc), _lib.ptr(bc), _lib.ptr(feat), _lib.ptr(ws),
                                           BV, Cin, Cout, H, W, st))
            _op._hwcp_register(feat, 'feat')
        ctx.save_for_backward(xc, wc)
        ctx.meta = (BV, Cin, Cout, H, W, x.dtype, weight.dtype, tuple(weight.shape),
                    bias.dtype if bias is not None else None)
        out = _op._hwcp_view(feat, (BV,), Cout, H, W)
        return out if x.dtype == torch.float32 else out.to(x.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_out):
        L = _lib.lib()
        xc, wc = ctx.saved_tensors
        BV, Cin, Cout, H, W, x_dtype, w_dtype, w_shape, b_dtype = ctx.meta
        need_x, need_w, need_b = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2] and b_dtype
        gx = gw = gb = None
        with torch.cuda.device(xc.device):
            st = _op._stream()
            g = _op._hwcp_base(grad_out, 'grad')  # K1's backward left its gradient map in the gather layout
            if g is None and BV > 0:
                go = _op._f32c(grad_out)
                g = torch.empty((L.lt_feat_layout_elems(BV, Cout, H, W),), device=xc.device, dtype=torch.float32)
                _lib.check(L.lt_feat_to_hwc(_lib.ptr(go), _lib.ptr(g), BV, Cout, H, W, st))
            if need_x:
                gx = torch.empty((BV, Cin, H, W), device=xc.device, dtype=torch.float32)
                _lib.check(L.lt_pwconv_bwd_data(_lib.ptr(g), _lib.ptr(wc), _lib.ptr(gx), BV, Cin, Cout, H, W, st))
                gx = gx.to(x_dtype)
            if need_w or need_b:
                gw = torch.empty((Cout, Cin), device=xc.device, dtype=torch.float32)
                gb = torch.empty((Cout,), device=xc.device, dtype=torch.float32) if need_b else None
                ws = torch.empty((max(L.lt_pwconv_ws_bytes(BV, Cin, Cout, H, W), 4) // 4,), device=xc.device,
                                 dtype=torch.float32)
                _lib.check(L.lt_pwconv_bwd_weight(_lib.ptr(xc), _lib.ptr(g), _lib.ptr(gw), _lib.ptr(gb), _lib.ptr(ws),
                                                  BV, Cin, Cout, H, W, st))
                gw = gw.reshape(w_shape).to(w_dtype) if need_w else None
                gb = gb.to(b_dtype) if need_b else None
        return gx, gw, gb


@annotate("lt_b200.process_features")
def process_features(features, weight, bias=None):
    """``F.conv2d(features, weight, bias)`` for a 1x1 kernel (the ``process_features`` step, triangulation.py:238-240,
    344-346), written straight into the layout ``op.unproject_heatmaps`` gathers from.

    features (B*V, Cin, H, W); weight (Cout, Cin, 1, 1) or (Cout, Cin); bias (Cout) or None. Returns (B*V, Cout, H, W)
    in the dtype of ``features``. Differentiable w.r.t. all three."""
    _op._require_cuda("features", features)
    _op._require_cuda("weight", weight)
    if features.dim() != 4:
        raise RuntimeError("Expected 4D (batched) input to conv2d, but got input of size: {}".format(
            list(features.shape)))
    if weight.dim() == 4 and tuple(weight.shape[2:]) != (1, 1):
        raise RuntimeError("process_features is a 1x1 convolution; got a kernel of size {}".format(
            tuple(weight.shape[2:])))
    if weight.dim() not in (2, 4) or weight.shape[1] != features.shape[1]:
        raise RuntimeError("weight {} does not match {} input channels".format(tuple(weight.shape), features.shape[1]))
    if bias is not None:
        _op._require_cuda("bias", bias)
        if tuple(bias.shape) != (weight.shape[0],):
            raise RuntimeError("bias {} does not match {} output channels".format(tuple(bias.shape), weight.shape[0]))
    return _ProcessFeatures.apply(features, weight, bias)


class ProcessFeatures(nn.Sequential):
    """Drop-in for ``nn.Sequential(nn.Conv2d(256, 32, 1))`` (triangulation.py:238-240) running kernel K7."""

    def __init__(self, in_channels=256, out_channels=32):
        super().__init__(nn.Conv2d(in_channels, out_channels, 1))

    def forward(self, x):
        conv = self[0]
        return process_features(x, conv.weight, conv.bias)


def use_fused_process_features(model):
    """Make ``model.process_features`` (an ``nn.Sequential`` holding one 1x1 ``nn.Conv2d``, as the reference builds it)
    run K7, keeping its parameters, hooks and ``state_dict`` keys. Returns the model."""
    seq = model.process_features
    if not (isinstance(seq, nn.Sequential) and len(seq) == 1 and isinstance(seq[0], nn.Conv2d)
            and seq[0].kernel_size == (1, 1) and seq[0].stride == (1, 1) and seq[0].padding == (0, 0)
            and seq[0].groups == 1 and seq[0].dilation == (1, 1)):
        raise TypeError("process_features is not nn.Sequential(nn.Conv2d(cin, cout, 1))")
    seq.__class__ = ProcessFeatures
    return model
