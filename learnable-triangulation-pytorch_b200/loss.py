"""Drop-in replacement for ``mvn.models.loss.VolumetricCELoss`` (reference file ``mvn/models/loss.py:52-80``): same
class name, same ``forward(coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity)``, same
scalar result. The reference builds a (J, N^3) distance tensor per batch item, copies the argmin indices to the host
(``loss.py:71``: a stream synchronisation per item) and gathers B*J scalars in a python loop; here the whole loss is
three small launches behind ``lt_volce_fwd`` with no synchronisation and no temporary. CUDA only, no CPU path.
"""
import torch
from torch import nn

from . import _lib
from ._nvtx import annotate
from .op import _f32c, _require_cuda, _stream


class _VolumetricCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, coord_volumes, volumes_pred, keypoints_gt, validity):
        L = _lib.lib()
        B, J = volumes_pred.shape[:2]
        nvox = volumes_pred[0, 0].numel()
        with torch.cuda.device(volumes_pred.device):
            dev = volumes_pred.device
            prob = _f32c(volumes_pred)
            coord, kp, val = _f32c(coord_volumes), _f32c(keypoints_gt), _f32c(validity)
            loss = torch.empty((1,), device=dev, dtype=torch.float32)
            idx = torch.empty((B, J), device=dev, dtype=torch.int32)
            ws = torch.empty((L.lt_volce_ws_bytes(B, J, nvox),), device=dev, dtype=torch.uint8)
            _lib.check(L.lt_volce_fwd(_lib.ptr(coord), _lib.ptr(prob), _lib.ptr(kp), _lib.ptr(val), _lib.ptr(loss),
                                      _lib.ptr(idx), _lib.ptr(ws), B, J, nvox, _stream()))
        ctx.save_for_backward(prob, idx, val)
        ctx.meta = (tuple(volumes_pred.shape), volumes_pred.dtype)
        return loss[0].to(volumes_pred.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_loss):
        if not ctx.needs_input_grad[1]:
            return None, None, None, None
        L = _lib.lib()
        prob, idx, val = ctx.saved_tensors
        shape, dtype = ctx.meta
        B, J = shape[:2]
        with torch.cuda.device(prob.device):
            g = _f32c(g_loss).reshape(1)
            gvol = torch.empty(shape, device=prob.device, dtype=torch.float32)
            _lib.check(L.lt_volce_bwd(_lib.ptr(prob), _lib.ptr(idx), _lib.ptr(val), _lib.ptr(g), _lib.ptr(gvol), B, J,
                                      prob[0, 0].numel(), _stream()))
        return None, gvol.to(dtype), None, None


@annotate("lt_b200.volumetric_ce_loss")
def volumetric_ce_loss(coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity):
    """``-log(p + 1e-6)`` of the predicted volume at the voxel nearest to each ground-truth joint, weighted by the
    joint's validity and averaged over ALL B*J joints (reference: loss.py:56-80).

    coord_volumes_batch (B,N0,N1,N2,3); volumes_batch_pred (B,J,N0,N1,N2) (the second output of
    ``integrate_tensor_3d_with_coordinates``); keypoints_gt (B,J,3); keypoints_binary_validity (B,J,1).
    Returns a 0-dim tensor; gradient flows into ``volumes_batch_pred`` only (as in the reference: the argmin is detached
    at loss.py:71 and the other inputs are ground truth).
    """
    for name, t in (("coord_volumes_batch", coord_volumes_batch), ("volumes_batch_pred", volumes_batch_pred),
                    ("keypoints_gt", keypoints_gt), ("keypoints_binary_validity", keypoints_binary_validity)):
        _require_cuda(name, t)
    B, J = volumes_batch_pred.shape[:2]
    if B * J == 0:
        # `loss / n_losses` with loss = 0.0 and n_losses = 0 (loss.py:57-58,80)
        raise ZeroDivisionError("float division by zero")
    if tuple(coord_volumes_batch.shape) != (B,) + tuple(volumes_batch_pred.shape[2:]) + (3,):
        raise RuntimeError("coord_volumes_batch {} does not match volumes_batch_pred {}".format(
            tuple(coord_volumes_batch.shape), tuple(volumes_batch_pred.shape)))
    if tuple(keypoints_gt.shape) != (B, J, 3):
        raise RuntimeError("keypoints_gt must be (B, J, 3), got {}".format(tuple(keypoints_gt.shape)))
    validity = keypoints_binary_validity.reshape(B, J, -1)[:, :, 0]   # `validity[0]` of loss.py:76
    return _VolumetricCE.apply(coord_volumes_batch, volumes_batch_pred, keypoints_gt, validity)


class VolumetricCELoss(nn.Module):
    """Same constructor and forward signature as the reference class (loss.py:52-55)."""

    def __init__(self):
        super().__init__()

    def forward(self, coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity):
        return volumetric_ce_loss(coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity)


class _SoftArgmaxWithCE(torch.autograd.Function):
    """K2 forward + K5 forward; one backward in which the CE gradient stays sparse (B*J values) instead of becoming a
    (B,J,N^3) tensor that K2's backward would have to read."""

    @staticmethod
    def forward(ctx, volumes, coord, keypoints_gt, validity, softmax, multiplier):
        L = _lib.lib()
        B, J = volumes.shape[:2]
        nvox = volumes[0, 0].numel()
        with torch.cuda.device(volumes.device):
            st, dev = _stream(), volumes.device
            vol, coord_c = _f32c(volumes), _f32c(coord)
            kp, val = _f32c(keypoints_gt), _f32c(validity)
            coords = torch.empty((B, J, 3), device=dev, dtype=torch.float32)
            prob = torch.empty(vol.shape, device=dev, dtype=torch.float32)
            stats = torch.empty((B, J, 5), device=dev, dtype=torch.float32)
            ws = torch.empty((max(L.lt_softargmax_ws_bytes(B, J, nvox), L.lt_volce_ws_bytes(B, J, nvox)),), device=dev,
                             dtype=torch.uint8)
            _lib.check(L.lt_softargmax_fwd(_lib.ptr(vol), _lib.ptr(coord_c), _lib.ptr(coords), _lib.ptr(prob),
                                           _lib.ptr(stats), _lib.ptr(ws), B, J, nvox, softmax, multiplier, st))
            loss = torch.empty((1,), device=dev, dtype=torch.float32)
            idx = torch.empty((B, J), device=dev, dtype=torch.int32)
            _lib.check(L.lt_volce_fwd(_lib.ptr(coord_c), _lib.ptr(prob), _lib.ptr(kp), _lib.ptr(val), _lib.ptr(loss),
                                      _lib.ptr(idx), _lib.ptr(ws), B, J, nvox, st))
        ctx.save_for_backward(prob, coord_c, coords, idx, val)
        ctx.meta = (B, J, nvox, softmax, multiplier, volumes.dtype)
        ctx.set_materialize_grads(False)
        return coords.to(volumes.dtype), prob.to(volumes.dtype), loss[0].to(volumes.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_coords, g_prob, g_loss):
        if not ctx.needs_input_grad[0]:
            return (None,) * 6
        L = _lib.lib()
        prob, coord_c, coords, idx, val = ctx.saved_tensors
        B, J, nvox, softmax, multiplier, dtype = ctx.meta
        with torch.cuda.device(prob.device):
            st, dev = _stream(), prob.device
            gvol = torch.empty_like(prob)
            gc = _f32c(g_coords) if g_coords is not None else torch.zeros((B, J, 3), device=dev, dtype=torch.float32)
            gp = _f32c(g_prob) if g_prob is not None else None
            vals = dots = None
            if g_loss is not None:
                gl = _f32c(g_loss).reshape(1)
                if gp is None:
                    vals = torch.empty((B, J), device=dev, dtype=torch.float32)
                    dots = torch.empty((B, J), device=dev, dtype=torch.float32)
                    _lib.check(L.lt_volce_bwd_sparse(_lib.ptr(prob), _lib.ptr(idx), _lib.ptr(val), _lib.ptr(gl),
                                                     _lib.ptr(vals), _lib.ptr(dots), B, J, nvox, st))
                else:  # a dense gradient arrives through the volumes as well: fold the B*J values into it
                    dense = torch.empty_like(prob)
                    _lib.check(L.lt_volce_bwd(_lib.ptr(prob), _lib.ptr(idx), _lib.ptr(val), _lib.ptr(gl),
                                              _lib.ptr(dense), B, J, nvox, st))
                    gp = gp + dense
            if gp is not None and softmax:
                dots = torch.empty((B, J), device=dev, dtype=torch.float32)
                ws = torch.empty((L.lt_softargmax_ws_bytes(B, J, nvox),), device=dev, dtype=torch.uint8)
                _lib.check(L.lt_softargmax_bwd_dot(_lib.ptr(prob), _lib.ptr(gp), _lib.ptr(dots), _lib.ptr(ws), B, J,
                                                   nvox, st))
            _lib.check(L.lt_softargmax_bwd(_lib.ptr(prob), _lib.ptr(coord_c), _lib.ptr(coords), _lib.ptr(gc),
                                           _lib.ptr(gp), _lib.ptr(dots) if softmax else None, _lib.ptr(gvol), B, J,
                                           nvox, softmax, multiplier, st))
            if vals is not None:
                _lib.check(L.lt_softargmax_bwd_add_sparse(_lib.ptr(prob), _lib.ptr(idx), _lib.ptr(vals),
                                                          _lib.ptr(gvol), B, J, nvox, softmax, multiplier, st))
        return gvol.to(dtype), None, None, None, None, None


@annotate("lt_b200.integrate_tensor_3d_with_coordinates_and_ce")
def integrate_tensor_3d_with_coordinates_and_ce(volumes, coord_volumes, keypoints_gt, keypoints_binary_validity,
                                                softmax=True, *, multiplier=1.0):
    """``integrate_tensor_3d_with_coordinates`` (op.py:84-96) and ``VolumetricCELoss`` (loss.py:52-80) of its second
    output in one autograd node. Returns ``(coordinates (B,J,3), volumes (B,J,N0,N1,N2), ce_loss ())``, each equal to
    what the two separate calls return; the difference is the backward, where the loss gradient reaches the soft-argmax
    as B*J numbers instead of a dense (B,J,N^3) tensor (train.py:218-231 is the call pattern this serves)."""
    for name, t in (("volumes", volumes), ("coord_volumes", coord_volumes), ("keypoints_gt", keypoints_gt),
                    ("keypoints_binary_validity", keypoints_binary_validity)):
        _require_cuda(name, t)
    if volumes.dim() != 5 or tuple(coord_volumes.shape) != (volumes.shape[0],) + tuple(volumes.shape[2:]) + (3,):
        raise RuntimeError("coord_volumes {} does not match volumes {}".format(
            tuple(coord_volumes.shape), tuple(volumes.shape)))
    B, J = volumes.shape[:2]
    if B * J == 0 or volumes[0, 0].numel() == 0:
        raise ZeroDivisionError("float division by zero")
    if tuple(keypoints_gt.shape) != (B, J, 3):
        raise RuntimeError("keypoints_gt must be (B, J, 3), got {}".format(tuple(keypoints_gt.shape)))
    multiplier = float(multiplier)
    if not multiplier > 0.0:
        volumes, multiplier = volumes * multiplier, 1.0
    validity = keypoints_binary_validity.reshape(B, J, -1)[:, :, 0]
    return _SoftArgmaxWithCE.apply(volumes, coord_volumes, keypoints_gt, validity, 1 if softmax else 0, multiplier)
