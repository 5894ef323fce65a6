"""Drop-in replacement for ``mvn.models.loss.VolumetricCELoss`` (reference file ``mvn/models/loss.py:52-80``): same
class name, same ``forward(coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity)``, same
scalar result. The reference builds a (J, N^3) distance tensor per batch item, copies the argmin indices to the host
(``loss.py:71``: a stream synchronisation per item) and gathers B*J scalars in a python loop; here the whole loss is
three small launches behind ``lt_volce_fwd`` with no synchronisation and no temporary. CUDA only, no CPU path.
"""
import torch
from torch import nn

from . import _lib
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
