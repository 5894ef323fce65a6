"""Drop-in replacements for the two volumetric functions of ``mvn.utils.op`` (reference file
``mvn/utils/op.py``): same names, same positional/keyword arguments, same return values, same exceptions.
Everything runs in hand-written sm_100a kernels behind the C ABI of ``include/lt_b200.h``; there is no CPU path.

    unproject_heatmaps(heatmaps, proj_matricies, coord_volumes, volume_aggregation_method='sum', vol_confidences=None)
        reference: mvn/utils/op.py:99-163
    integrate_tensor_3d_with_coordinates(volumes, coord_volumes, softmax=True)
        reference: mvn/utils/op.py:84-96
    integrate_tensor_2d(heatmaps, softmax=True)
        reference: mvn/utils/op.py:11-47
    integrate_tensor_3d(volumes, softmax=True)
        reference: mvn/utils/op.py:50-81
"""
import torch

from . import _lib
from ._nvtx import annotate
from .coordspec import CoordSpec

_AGG = {"sum": 0, "max": 2, "softmax": 3}  # 'conf*' -> 1, see _agg_code


def _agg_code(method):
    # op.py:147-161: startswith('conf') first, then exact names, else ValueError with this exact text
    if method.startswith("conf"):
        return 1
    if method in _AGG:
        return _AGG[method]
    raise ValueError("Unknown volume_aggregation_method: {}".format(method))


def _require_cuda(name, t):
    if not torch.is_tensor(t):
        raise TypeError("{} must be a torch tensor".format(name))
    if not t.is_cuda:
        raise RuntimeError(
            "lt_b200: {} is on {}; these ops only exist as CUDA kernels (no CPU fallback)".format(name, t.device))


def _f32c(t):
    return t.detach().to(torch.float32).contiguous()


def _stream():
    return torch.cuda.current_stream().cuda_stream


# ---- tensors that already live in the gather layout -----------------------------------------------------------------
# triangulation.process_features writes K1's gather layout "hwcp" directly and returns a strided (B*V,C,H,W) view of its
# interior; K1's backward hands its gradient map back the same way. The fact "this storage holds a hwcp buffer" is an
# attribute ON THE STORAGE OBJECT (torch keeps one python object per live storage, shared by every view / detach() of
# it), so it disappears with the buffer: no registry, no addresses, nothing to prune, nothing shared between threads.
# A tensor is taken as "already hwcp" only if its storage carries the tag AND its strides/offset are exactly the interior
# view (also after the `.view(batch_size, n_views, ...)` of triangulation.py:346). Anything else takes K0.
_HWCP_ATTR = "_lt_b200_hwcp"  # -> (kind, number of floats); kind 'feat' has a zero border, 'grad' a scratch border


def _hwcp_register(base, kind):
    setattr(base.untyped_storage(), _HWCP_ATTR, (kind, base.numel()))


def _hwcp_view(base, lead, C, H, W):
    """Interior of the 1-D gather-layout buffer `base` as a strided (*lead, C, H, W) tensor (no copy)."""
    Cp = _lib.lib().lt_padded_channels(C)
    Wp, S = W + 2, (H + 2) * (W + 2) * Cp
    strides, acc = [], S
    for n in reversed(lead):
        strides.insert(0, acc)
        acc *= n
    return base.as_strided(tuple(lead) + (C, H, W), tuple(strides) + (1, Wp * Cp, Cp), (Wp + 1) * Cp)


def _hwcp_base(t, kind):
    """The 1-D gather-layout buffer behind `t` (*lead, C, H, W) if `t` is exactly its interior view, else None."""
    if t.dtype != torch.float32 or not t.is_cuda or t.dim() < 4 or t.numel() == 0:
        return None
    st = t.untyped_storage()
    entry = getattr(st, _HWCP_ATTR, None)
    if entry is None or entry[0] != kind:
        return None
    *lead, C, H, W = t.shape
    Cp = _lib.lib().lt_padded_channels(C)
    Wp, S = W + 2, (H + 2) * (W + 2) * Cp
    n = 1
    for d in lead:
        n *= d
    if entry[1] != n * S or st.nbytes() != n * S * 4 or t.storage_offset() != (Wp + 1) * Cp:
        return None
    want, acc = [], S
    for d in reversed(lead):
        want.insert(0, acc)
        acc *= d
    want += [1, Wp * Cp, Cp]
    if any(sz != 1 and a != b for sz, a, b in zip(t.shape, t.stride(), want)):
        return None
    return torch.empty(0, device=t.device, dtype=torch.float32).set_(st, 0, (n * S,), (1,))


class _Unproject(torch.autograd.Function):
    @staticmethod
    def forward(ctx, heatmaps, proj, coord, conf, agg, align_corners):
        L = _lib.lib()
        B, V, C, H, W = heatmaps.shape
        N0, N1, N2 = coord.shape[1:4]
        spec = coord if isinstance(coord, CoordSpec) else None   # coordinates computed in the kernels (coordspec.py)
        with torch.cuda.device(heatmaps.device):
            st = _stream()
            proj_c = _f32c(proj)
            coord_c = spec.specs if spec is not None else _f32c(coord)
            conf_c = _f32c(conf) if conf is not None else None
            feat = _hwcp_base(heatmaps, 'feat')  # written in the gather layout by K7 (process_features): no K0
            from_hwcp = feat is not None
            if not from_hwcp:
                hm = _f32c(heatmaps)
                # K0: gather layout (channels-last, padded channels, zero border) of the feature maps; opaque to this side
                feat = torch.empty((L.lt_feat_layout_elems(B * V, C, H, W),), device=hm.device, dtype=torch.float32)
                _lib.check(L.lt_feat_to_hwc(_lib.ptr(hm), _lib.ptr(feat), B * V, C, H, W, st))
            out = torch.empty((B, C, N0, N1, N2), device=feat.device, dtype=torch.float32)
            if spec is not None:
                _lib.check(L.lt_unproject_fwd_spec(_lib.ptr(feat), _lib.ptr(proj_c), _lib.ptr(coord_c),
                                                   1 if spec.transfer else 0, _lib.ptr(conf_c), _lib.ptr(out), B, V, C,
                                                   H, W, N0, N1, N2, agg, align_corners, st))
            else:
                _lib.check(L.lt_unproject_fwd(_lib.ptr(feat), _lib.ptr(proj_c), _lib.ptr(coord_c), _lib.ptr(conf_c),
                                              _lib.ptr(out), B, V, C, H, W, N0, N1, N2, agg, align_corners, st))
        # with features that already are the gather layout `feat` aliases the caller's tensor: save that tensor as well so
        # that autograd's version check catches an in-place edit between forward and backward, as it would in the reference
        ctx.save_for_backward(feat, proj_c, coord_c, conf_c, heatmaps if from_hwcp else None)
        ctx.meta = (B, V, C, H, W, N0, N1, N2, agg, align_corners, heatmaps.dtype,
                    conf.dtype if conf is not None else None, from_hwcp,
                    None if spec is None else (1 if spec.transfer else 0))
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_out):
        L = _lib.lib()
        feat, proj_c, coord_c, conf_c, _ = ctx.saved_tensors
        B, V, C, H, W, N0, N1, N2, agg, align_corners, hm_dtype, conf_dtype, from_hwcp, spec_transfer = ctx.meta
        need_hm, need_conf = ctx.needs_input_grad[0], ctx.needs_input_grad[3] and conf_c is not None
        g_hm = g_conf = None
        if not (need_hm or need_conf):
            return None, None, None, None, None, None
        with torch.cuda.device(feat.device):
            st = _stream()
            go = _f32c(grad_out)
            gfeat = torch.zeros_like(feat)  # the kernel scatter-adds into it
            gconf = ws = None
            if need_conf:
                gconf = torch.empty((B, V, C), device=feat.device, dtype=torch.float32)
                nbytes = L.lt_unproject_bwd_ws_bytes(B, V, C, N0, N1, N2)
                ws = torch.empty((max(nbytes, 4) // 4,), device=feat.device, dtype=torch.float32)
            if spec_transfer is not None:
                _lib.check(L.lt_unproject_bwd_spec(_lib.ptr(go), _lib.ptr(feat), _lib.ptr(proj_c), _lib.ptr(coord_c),
                                                   spec_transfer, _lib.ptr(conf_c), _lib.ptr(gfeat), _lib.ptr(gconf),
                                                   _lib.ptr(ws), B, V, C, H, W, N0, N1, N2, agg, align_corners, st))
            else:
                _lib.check(L.lt_unproject_bwd(_lib.ptr(go), _lib.ptr(feat), _lib.ptr(proj_c), _lib.ptr(coord_c),
                                              _lib.ptr(conf_c), _lib.ptr(gfeat), _lib.ptr(gconf), _lib.ptr(ws),
                                              B, V, C, H, W, N0, N1, N2, agg, align_corners, st))
            if need_hm and from_hwcp:
                # the producer (K7's backward) reads the gradient map in place: hand its interior back as a view
                _hwcp_register(gfeat, 'grad')
                g_hm = _hwcp_view(gfeat, (B, V), C, H, W)
            elif need_hm:
                g_hm = torch.empty((B, V, C, H, W), device=feat.device, dtype=torch.float32)
                _lib.check(L.lt_feat_to_chw(_lib.ptr(gfeat), _lib.ptr(g_hm), B * V, C, H, W, st))
                g_hm = g_hm.to(hm_dtype)
            if need_conf:
                g_conf = gconf.to(conf_dtype)
        return g_hm, None, None, g_conf, None, None


@annotate("lt_b200.unproject_heatmaps")
def unproject_heatmaps(heatmaps, proj_matricies, coord_volumes, volume_aggregation_method='sum',
                       vol_confidences=None, *, align_corners=None):
    """Lift V feature maps per batch item into one aggregated 3-D volume (reference: op.py:99-163).

    heatmaps (B,V,C,H,W), proj_matricies (B,V,3,4), coord_volumes (B,N0,N1,N2,3), vol_confidences (B,V,C) or None.
    Returns (B,C,N0,N1,N2) float32, always (the reference allocates its output with the default dtype, op.py:104).
    ``align_corners`` (keyword only): None/False = what ``F.grid_sample`` does by default on torch >= 1.3, i.e. what
    the reference computes today; True = the behaviour under its pinned torch 1.0.1.
    ``coord_volumes`` may also be a ``lt_b200.CoordSpec`` (18 floats per item): the voxel centres are then computed
    inside the kernels, bit-identical to passing ``spec.materialize()``, and the tensor never exists.
    Differentiable w.r.t. heatmaps and vol_confidences.
    """
    agg = _agg_code(volume_aggregation_method)
    if agg == 1 and vol_confidences is None:
        # the reference fails on `vol_confidences[batch_i]` (op.py:148)
        raise TypeError("'NoneType' object is not subscriptable")
    _require_cuda("heatmaps", heatmaps)
    _require_cuda("proj_matricies", proj_matricies)
    if not isinstance(coord_volumes, CoordSpec):
        _require_cuda("coord_volumes", coord_volumes)
    if heatmaps.dim() != 5:
        raise RuntimeError("heatmaps must be (B,V,C,H,W), got {}".format(tuple(heatmaps.shape)))
    B, V = heatmaps.shape[:2]
    if coord_volumes.dim() != 5 or coord_volumes.shape[-1] != 3 or coord_volumes.shape[0] < B:
        raise RuntimeError("coord_volumes must be (B,N0,N1,N2,3), got {}".format(tuple(coord_volumes.shape)))
    if tuple(proj_matricies.shape[-2:]) != (3, 4) or proj_matricies.shape[0] < B or proj_matricies.shape[1] < V:
        raise RuntimeError("proj_matricies must be (B,V,3,4), got {}".format(tuple(proj_matricies.shape)))
    if V == 0:
        raise RuntimeError("unproject_heatmaps needs at least one view")
    conf = None
    if agg == 1:
        _require_cuda("vol_confidences", vol_confidences)
        conf = vol_confidences[:B].reshape(B, V, heatmaps.shape[2])
    # the reference only ever indexes [batch_i, view_i] of these (op.py:108,118)
    proj = proj_matricies[:B, :V]
    coord = coord_volumes[:B]
    from . import torch_ops
    if torch_ops.available() and _hwcp_base(heatmaps, 'feat') is None:
        # the torch extension: one dispatcher call, C++ autograd node (torch.ops.lt_b200.unproject_heatmaps)
        if isinstance(coord, CoordSpec):
            return torch.ops.lt_b200.unproject_heatmaps(heatmaps, proj, coord.specs, agg, conf, bool(align_corners),
                                                        list(coord.size), coord.transfer)
        return torch.ops.lt_b200.unproject_heatmaps(heatmaps, proj, coord, agg, conf, bool(align_corners), [], False)
    # features already in the gather layout (written by K7, triangulation.process_features), or no extension built:
    # the ctypes binding of the same kernels
    return _Unproject.apply(heatmaps, proj, coord, conf, agg, 1 if align_corners else 0)


class _SoftArgmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, volumes, coord, softmax, multiplier):
        L = _lib.lib()
        B, J = volumes.shape[:2]
        N0, N1, N2 = (int(v) for v in volumes.shape[2:])
        nvox = N0 * N1 * N2
        spec = coord if isinstance(coord, CoordSpec) else None   # coordinates computed in the kernels (coordspec.py)
        with torch.cuda.device(volumes.device):
            st = _stream()
            vol = _f32c(volumes)
            coord_c = spec.specs if spec is not None else _f32c(coord)
            coords = torch.zeros((B, J, 3), device=vol.device, dtype=torch.float32)
            prob = torch.empty(vol.shape, device=vol.device, dtype=torch.float32)
            stats = torch.empty((B, J, 5), device=vol.device, dtype=torch.float32)
            if nvox > 0:
                nbytes = L.lt_softargmax_ws_bytes(B, J, nvox)
                ws = torch.empty((max(nbytes, 4) // 4,), device=vol.device, dtype=torch.float32)
                if spec is not None:
                    _lib.check(L.lt_softargmax_fwd_spec(_lib.ptr(vol), _lib.ptr(coord_c), 1 if spec.transfer else 0,
                                                        _lib.ptr(coords), _lib.ptr(prob), _lib.ptr(stats), _lib.ptr(ws),
                                                        B, J, N0, N1, N2, softmax, multiplier, st))
                else:
                    _lib.check(L.lt_softargmax_fwd(_lib.ptr(vol), _lib.ptr(coord_c), _lib.ptr(coords), _lib.ptr(prob),
                                                   _lib.ptr(stats), _lib.ptr(ws), B, J, nvox, softmax, multiplier, st))
        ctx.save_for_backward(prob, coord_c, coords)
        ctx.meta = (B, J, nvox, softmax, multiplier, volumes.dtype,
                    None if spec is None else (1 if spec.transfer else 0, N0, N1, N2))
        ctx.set_materialize_grads(False)
        return coords.to(volumes.dtype), prob.to(volumes.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_coords, g_prob):
        L = _lib.lib()
        prob, coord_c, coords = ctx.saved_tensors
        B, J, nvox, softmax, multiplier, dtype, spec_meta = ctx.meta
        if not ctx.needs_input_grad[0]:
            return None, None, None, None
        with torch.cuda.device(prob.device):
            st = _stream()
            gvol = torch.empty_like(prob)
            if nvox == 0 or B * J == 0:
                return gvol.to(dtype), None, None, None
            gc = _f32c(g_coords) if g_coords is not None else torch.zeros((B, J, 3), device=prob.device,
                                                                          dtype=torch.float32)
            gp = _f32c(g_prob) if g_prob is not None else None
            dots = None
            if gp is not None and softmax:
                dots = torch.empty((B, J), device=prob.device, dtype=torch.float32)
                nbytes = L.lt_softargmax_ws_bytes(B, J, nvox)
                ws = torch.empty((max(nbytes, 4) // 4,), device=prob.device, dtype=torch.float32)
                _lib.check(L.lt_softargmax_bwd_dot(_lib.ptr(prob), _lib.ptr(gp), _lib.ptr(dots), _lib.ptr(ws),
                                                   B, J, nvox, st))
            if spec_meta is not None:
                _lib.check(L.lt_softargmax_bwd_spec(_lib.ptr(prob), _lib.ptr(coord_c), spec_meta[0], _lib.ptr(coords),
                                                    _lib.ptr(gc), _lib.ptr(gp), _lib.ptr(dots), _lib.ptr(gvol), B, J,
                                                    spec_meta[1], spec_meta[2], spec_meta[3], softmax, multiplier, st))
            else:
                _lib.check(L.lt_softargmax_bwd(_lib.ptr(prob), _lib.ptr(coord_c), _lib.ptr(coords), _lib.ptr(gc),
                                               _lib.ptr(gp), _lib.ptr(dots), _lib.ptr(gvol), B, J, nvox, softmax,
                                               multiplier, st))
        return gvol.to(dtype), None, None, None


@annotate("lt_b200.integrate_tensor_3d_with_coordinates")
def integrate_tensor_3d_with_coordinates(volumes, coord_volumes, softmax=True, *, multiplier=1.0):
    """Soft-argmax of each joint volume in world coordinates (reference: op.py:84-96).

    volumes (B,J,N0,N1,N2), coord_volumes (B,N0,N1,N2,3). Returns ``(coordinates (B,J,3), volumes (B,J,N0,N1,N2))``
    where the second value is the softmax-normalised volume (softmax=True) or relu(volumes) un-normalised
    (softmax=False, op.py:91). ``multiplier`` (keyword only, default 1.0) fuses the ``volumes * volume_multiplier``
    that the caller does before the call (triangulation.py:353). Gradient flows from both outputs into ``volumes``.
    """
    _require_cuda("volumes", volumes)
    if not isinstance(coord_volumes, CoordSpec):   # a CoordSpec: the centres are computed inside the kernels
        _require_cuda("coord_volumes", coord_volumes)
    if volumes.dim() != 5:
        raise RuntimeError("volumes must be (B,J,N0,N1,N2), got {}".format(tuple(volumes.shape)))
    B = volumes.shape[0]
    if coord_volumes.dim() != 5 or coord_volumes.shape[-1] != 3 or coord_volumes.shape[0] != B or \
            tuple(coord_volumes.shape[1:4]) != tuple(volumes.shape[2:]):
        # torch.einsum("bnxyz, bxyzc -> bnc") would raise RuntimeError on mismatching operands (op.py:94)
        raise RuntimeError("coord_volumes {} does not match volumes {}".format(
            tuple(coord_volumes.shape), tuple(volumes.shape)))
    multiplier = float(multiplier)
    if not multiplier > 0.0:  # the kernels fold a positive scale into the exponent; anything else is pre-scaled
        volumes = volumes * multiplier
        multiplier = 1.0
    from . import torch_ops
    if torch_ops.available():
        if isinstance(coord_volumes, CoordSpec):
            return torch.ops.lt_b200.integrate_tensor_3d_with_coordinates(volumes, coord_volumes.specs, bool(softmax),
                                                                          multiplier, True, coord_volumes.transfer)
        return torch.ops.lt_b200.integrate_tensor_3d_with_coordinates(volumes, coord_volumes, bool(softmax), multiplier,
                                                                      False, False)
    return _SoftArgmax.apply(volumes, coord_volumes, 1 if softmax else 0, multiplier)


@annotate("lt_b200.integrate_tensor_3d")
def integrate_tensor_3d(volumes, softmax=True):
    """Soft-argmax of each joint volume in VOXEL-INDEX coordinates (reference: op.py:50-81; not used by the reference's
    nets, kept so that every ``integrate_*`` of ``mvn.utils.op`` has its counterpart here).

    volumes (B,J,N0,N1,N2). Returns ``(coordinates (B,J,3) as (x, y, z) indices, volumes)``: softmax=True gives the
    softmax-normalised volume and its expected index; softmax=False gives relu(volumes) un-normalised while the
    coordinates are divided by the total mass (op.py:73-76; 0/0 = nan as in the reference). Built from the same
    kernels: the index grid is a coordinate volume with unit step (K6), the integration is K2."""
    from . import volumetric
    _require_cuda("volumes", volumes)
    if volumes.dim() != 5:
        raise ValueError("not enough values to unpack (expected 5, got {})".format(volumes.dim())
                         if volumes.dim() < 5 else "too many values to unpack (expected 5)")
    B = volumes.shape[0]
    spec = torch.zeros((B, volumetric.COORD_SPEC), dtype=torch.float32)
    spec[:, 3:6] = 1.0                                   # step 1 along every axis from the origin
    spec[:, 9], spec[:, 13], spec[:, 17] = 1.0, 1.0, 1.0   # identity rotation: coordinates are the exact indices
    with torch.cuda.device(volumes.device):
        grid = volumetric.coord_volumes_from_specs(spec.to(volumes.device, non_blocking=True), tuple(volumes.shape[2:]))
    coords, vol = integrate_tensor_3d_with_coordinates(volumes, grid.to(volumes.dtype), softmax=softmax)
    if not softmax:
        coords = coords / vol.flatten(2).sum(dim=2, keepdim=True)
    return coords, vol


class _SoftArgmax2D(torch.autograd.Function):
    @staticmethod
    def forward(ctx, heatmaps, softmax, multiplier):
        L = _lib.lib()
        B, J, H, W = heatmaps.shape
        with torch.cuda.device(heatmaps.device):
            hm = _f32c(heatmaps)
            coords = torch.zeros((B, J, 2), device=hm.device, dtype=torch.float32)
            prob = torch.empty(hm.shape, device=hm.device, dtype=torch.float32)
            _lib.check(L.lt_softargmax2d_fwd(_lib.ptr(hm), _lib.ptr(coords), _lib.ptr(prob), B * J, H, W, softmax,
                                             multiplier, _stream()))
        ctx.save_for_backward(prob, coords)
        ctx.meta = (B, J, H, W, softmax, multiplier, heatmaps.dtype)
        ctx.set_materialize_grads(False)
        return coords.to(heatmaps.dtype), prob.to(heatmaps.dtype)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g_coords, g_prob):
        L = _lib.lib()
        prob, coords = ctx.saved_tensors
        B, J, H, W, softmax, multiplier, dtype = ctx.meta
        if not ctx.needs_input_grad[0]:
            return None, None, None
        with torch.cuda.device(prob.device):
            ghm = torch.empty_like(prob)
            if B * J * H * W > 0:
                gc = _f32c(g_coords) if g_coords is not None else torch.zeros((B, J, 2), device=prob.device,
                                                                              dtype=torch.float32)
                gp = _f32c(g_prob) if g_prob is not None else None
                _lib.check(L.lt_softargmax2d_bwd(_lib.ptr(prob), _lib.ptr(coords), _lib.ptr(gc), _lib.ptr(gp),
                                                 _lib.ptr(ghm), B * J, H, W, softmax, multiplier, _stream()))
        return ghm.to(dtype), None, None


@annotate("lt_b200.integrate_tensor_2d")
def integrate_tensor_2d(heatmaps, softmax=True, *, multiplier=1.0):
    """Soft-argmax of joint heat maps in heat-map pixels (reference: op.py:11-47).

    heatmaps (B,J,H,W). Returns ``(coordinates (B,J,2) as (x, y), heatmaps (B,J,H,W))`` where the second value is the
    softmax-normalised map (softmax=True) or relu(heatmaps) un-normalised (softmax=False; the coordinates are then
    divided by the total mass, op.py:39-41). ``multiplier`` (keyword only, default 1.0) fuses the
    ``heatmaps * heatmap_multiplier`` of the caller (triangulation.py:164). Gradient flows from both outputs.
    """
    _require_cuda("heatmaps", heatmaps)
    if heatmaps.dim() != 4:
        # `batch_size, n_heatmaps, h, w = heatmaps.shape` (op.py:21) raises ValueError on any other rank
        raise ValueError("not enough values to unpack (expected 4, got {})".format(heatmaps.dim())
                         if heatmaps.dim() < 4 else "too many values to unpack (expected 4)")
    multiplier = float(multiplier)
    if not multiplier > 0.0:
        heatmaps = heatmaps * multiplier
        multiplier = 1.0
    return _SoftArgmax2D.apply(heatmaps, 1 if softmax else 0, multiplier)
