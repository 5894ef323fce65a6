"""z-slab sharded volumetric soft-argmax (BASELINE.json config 5): the volume is split into contiguous
slabs along its first spatial axis, one slab per rank; the only exchange is the per-(batch item, joint) online-softmax
tuple ``(m, S, Sx, Sy, Sz)`` — 5 floats — gathered once with ``torch.distributed.all_gather_into_tensor`` (NCCL over
NVLink on the GPU box, gloo in the CPU tests) and merged in a fixed rank order (deterministic).

The reference has no counterpart (it holds the whole ``(V,C,N^3)`` stack on one device, mvn/utils/op.py:111); the maths
is the same softmax/einsum of op.py:84-96 evaluated blockwise.

Backward needs no collective unless a dense gradient arrives through the second output: then the scalar
``sum_k p_k gp_k`` per (b, joint) is all-reduced.
"""
import torch
import torch.distributed as dist

from . import _lib
from .op import _f32c, _require_cuda, _stream


class CudaSlabOps:
    """The three device steps of the sharded op, on this rank's slab, through the C ABI."""

    @staticmethod
    def partial(vol, coord, softmax, multiplier):
        L = _lib.lib()
        B, J = vol.shape[:2]
        nvox = vol.shape[2] * vol.shape[3] * vol.shape[4]
        with torch.cuda.device(vol.device):
            stats = torch.empty((B, J, 5), device=vol.device, dtype=torch.float32)
            nbytes = L.lt_softargmax_ws_bytes(B, J, nvox)
            ws = torch.empty((max(nbytes, 4) // 4,), device=vol.device, dtype=torch.float32)
            _lib.check(L.lt_softargmax_partial(_lib.ptr(vol), _lib.ptr(coord), _lib.ptr(stats), _lib.ptr(ws), B, J,
                                               nvox, softmax, multiplier, _stream()))
        return stats

    @staticmethod
    def merge(parts, softmax):
        L = _lib.lib()
        R, B, J = parts.shape[:3]
        with torch.cuda.device(parts.device):
            merged = torch.empty((B, J, 5), device=parts.device, dtype=torch.float32)
            _lib.check(L.lt_softargmax_merge(_lib.ptr(parts), _lib.ptr(merged), R, B * J, softmax, _stream()))
        return merged

    @staticmethod
    def finalize(vol, stats, softmax, multiplier):
        L = _lib.lib()
        B, J = vol.shape[:2]
        nvox = vol.shape[2] * vol.shape[3] * vol.shape[4]
        with torch.cuda.device(vol.device):
            coords = torch.zeros((B, J, 3), device=vol.device, dtype=torch.float32)
            prob = torch.empty_like(vol)
            _lib.check(L.lt_softargmax_finalize(_lib.ptr(vol), _lib.ptr(stats), _lib.ptr(coords), _lib.ptr(prob), B, J,
                                                nvox, softmax, multiplier, _stream()))
        return coords, prob

    @staticmethod
    def backward(prob, coord, coords, g_coords, g_prob, dots, softmax, multiplier):
        L = _lib.lib()
        B, J = prob.shape[:2]
        nvox = prob.shape[2] * prob.shape[3] * prob.shape[4]
        with torch.cuda.device(prob.device):
            gvol = torch.empty_like(prob)
            _lib.check(L.lt_softargmax_bwd(_lib.ptr(prob), _lib.ptr(coord), _lib.ptr(coords), _lib.ptr(g_coords),
                                           _lib.ptr(g_prob), _lib.ptr(dots), _lib.ptr(gvol), B, J, nvox, softmax,
                                           multiplier, _stream()))
        return gvol

    @staticmethod
    def dot(prob, g_prob):
        L = _lib.lib()
        B, J = prob.shape[:2]
        nvox = prob.shape[2] * prob.shape[3] * prob.shape[4]
        with torch.cuda.device(prob.device):
            dots = torch.empty((B, J), device=prob.device, dtype=torch.float32)
            nbytes = L.lt_softargmax_ws_bytes(B, J, nvox)
            ws = torch.empty((max(nbytes, 4) // 4,), device=prob.device, dtype=torch.float32)
            _lib.check(L.lt_softargmax_bwd_dot(_lib.ptr(prob), _lib.ptr(g_prob), _lib.ptr(dots), _lib.ptr(ws), B, J,
                                               nvox, _stream()))
        return dots


def slab_bounds(n0, world_size, rank):
    """[lo, hi) of rank's slab along the first spatial axis: contiguous, sizes differ by at most one."""
    base, rem = divmod(n0, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def exchange_stats(local_stats, group=None):
    """all_gather of the (B,J,5) tuples -> (R,B,J,5) in rank order. The one collective of the forward."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local_stats.unsqueeze(0)
    # 1-D in, 1-D out: the layout both NCCL and gloo accept for all_gather_into_tensor
    flat = local_stats.contiguous().view(-1)
    out = torch.empty((world * flat.numel(),), device=flat.device, dtype=flat.dtype)
    dist.all_gather_into_tensor(out, flat, group=group)
    gathered = out.view((world,) + tuple(local_stats.shape))
    return gathered


class _ShardedSoftArgmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, volumes, coord, softmax, multiplier, group, ops):
        vol = _f32c(volumes)
        coord_c = _f32c(coord)
        local = ops.partial(vol, coord_c, softmax, multiplier)
        merged = ops.merge(exchange_stats(local, group), softmax)
        coords, prob = ops.finalize(vol, merged, softmax, multiplier)
        ctx.save_for_backward(prob, coord_c, coords)
        ctx.meta = (softmax, multiplier, group, ops, volumes.dtype)
        ctx.set_materialize_grads(False)
        return coords.to(volumes.dtype), prob.to(volumes.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_coords, g_prob):
        prob, coord_c, coords = ctx.saved_tensors
        softmax, multiplier, group, ops, dtype = ctx.meta
        if not ctx.needs_input_grad[0]:
            return None, None, None, None, None, None
        gc = _f32c(g_coords) if g_coords is not None else torch.zeros_like(coords)
        gp = _f32c(g_prob) if g_prob is not None else None
        dots = None
        if gp is not None and softmax:
            dots = ops.dot(prob, gp)
            if dist.is_initialized() and dist.get_world_size(group) > 1:
                dist.all_reduce(dots, op=dist.ReduceOp.SUM, group=group)
        gvol = ops.backward(prob, coord_c, coords, gc, gp, dots, softmax, multiplier)
        return gvol.to(dtype), None, None, None, None, None


def integrate_tensor_3d_with_coordinates_sharded(volumes_slab, coord_volumes_slab, softmax=True, *, multiplier=1.0,
                                                 group=None, ops=None):
    """Soft-argmax over a volume whose first spatial axis is sharded across the ranks of ``group``.

    volumes_slab (B,J,n0_local,N1,N2) and coord_volumes_slab (B,n0_local,N1,N2,3) are THIS rank's slab. Returns
    ``(coordinates (B,J,3), prob_slab)``: the coordinates are global (identical on every rank), the normalised volume
    stays sharded. ``ops`` is the device backend (default: the CUDA kernels); the CPU tests inject the oracle's
    restatement here to exercise the exchange logic under gloo.
    """
    if ops is None:
        ops = CudaSlabOps
        _require_cuda("volumes_slab", volumes_slab)
        _require_cuda("coord_volumes_slab", coord_volumes_slab)
    if volumes_slab.dim() != 5 or coord_volumes_slab.dim() != 5 or \
            tuple(coord_volumes_slab.shape[1:4]) != tuple(volumes_slab.shape[2:]):
        raise RuntimeError("slab shapes do not match: {} vs {}".format(
            tuple(volumes_slab.shape), tuple(coord_volumes_slab.shape)))
    multiplier = float(multiplier)
    if not multiplier > 0.0:
        volumes_slab = volumes_slab * multiplier
        multiplier = 1.0
    return _ShardedSoftArgmax.apply(volumes_slab, coord_volumes_slab, 1 if softmax else 0, multiplier, group, ops)
