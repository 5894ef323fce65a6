"""GPU: the on-chip merged K1 backward (csrc/unproject_bwd_tile.cu, opt-in with LT_B200_K1BWD=tile) in every tile
configuration against the oracle and against the scatter kernel (csrc/unproject_bwd.cu, the default), plus the BASELINE
configs that round 1 only benchmarked: cfg4's item (31 views, 64^3) and cfg5's 256-deep slabs.

Reference semantics: autograd through mvn/utils/op.py:107-161. Tolerances (BASELINE.json): 1e-5 relative on the volume,
1e-4 relative on gradients (the reductions are unordered); "relative" = conftest.rel_err (max-norm)."""
import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu
VOL_TOL, GRAD_TOL = 1e-5, 1e-4


def _grads(d, method, gout, conf):
    hm = d["heatmaps"].cuda().requires_grad_(True)
    vc = d["vol_confidences"].cuda().requires_grad_(True) if conf else None
    out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), method, vc)
    out.backward(gout)
    return out.detach(), hm.grad, (vc.grad if conf else None)


@pytest.mark.parametrize("tile", ["0", "1", "2", "3"])
@pytest.mark.parametrize("method", ["sum", "conf", "max", "softmax"])
@pytest.mark.parametrize("shape", [
    dict(B=2, V=4, C=32, H=24, W=24, N=16),              # several tiles per axis, every view chunk full
    dict(B=1, V=3, C=17, H=12, W=20, N=(9, 10, 16)),     # padded channels, ragged tiles along axes 0/1, H != W
    dict(B=1, V=1, C=32, H=16, W=16, N=(4, 4, 8)),       # one view, one partial tile
    dict(B=1, V=4, C=32, H=48, W=48, N=(16, 16, 32)),    # larger footprints per tile
])
def test_tile_backward_against_oracle_and_scatter(tile, method, shape, monkeypatch):
    conf = "conf" if method == "conf" else None
    d = synthetic.make_unproject_inputs(seed=11, train=True, conf=conf, inside_view=True, **shape)
    hm_ref = d["heatmaps"].clone().requires_grad_(True)
    vc_ref = d["vol_confidences"].clone().requires_grad_(True) if conf else None
    ref = ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"], d["coord_volumes"], method, vc_ref)
    gout = torch.randn(ref.shape, generator=torch.Generator().manual_seed(12))
    ref.backward(gout)

    monkeypatch.setenv("LT_B200_K1BWD", "tile")
    monkeypatch.setenv("LT_B200_K1BWD_TILE", tile)
    out, g_tile, gc_tile = _grads(d, method, gout.cuda(), conf)
    monkeypatch.setenv("LT_B200_K1BWD", "scatter")
    _, g_scat, gc_scat = _grads(d, method, gout.cuda(), conf)

    assert rel_err(out.cpu(), ref) < VOL_TOL
    assert rel_err(g_tile.cpu(), hm_ref.grad) < GRAD_TOL
    assert rel_err(g_tile, g_scat) < GRAD_TOL
    if conf:
        assert rel_err(gc_tile.cpu(), vc_ref.grad) < GRAD_TOL
        assert rel_err(gc_tile, gc_scat) < GRAD_TOL


@pytest.mark.parametrize("tile", ["0", "1"])
def test_tile_backward_many_views_in_chunks(tile, monkeypatch):
    """sum/conf walk the views in chunks of four through the same tile (tables rebuilt per chunk)."""
    monkeypatch.setenv("LT_B200_K1BWD", "tile")
    monkeypatch.setenv("LT_B200_K1BWD_TILE", tile)
    d = synthetic.make_unproject_inputs(B=1, V=9, C=32, H=16, W=16, N=(8, 8, 16), seed=3, train=True, conf="conf",
                                        dome=True, inside_view=True)
    for method in ("sum", "conf"):
        hm_ref = d["heatmaps"].clone().requires_grad_(True)
        vc_ref = d["vol_confidences"].clone().requires_grad_(True)
        ref = ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"], d["coord_volumes"], method, vc_ref)
        gout = torch.randn(ref.shape, generator=torch.Generator().manual_seed(4))
        ref.backward(gout)
        _, g, gc = _grads(d, method, gout.cuda(), True)
        assert rel_err(g.cpu(), hm_ref.grad) < GRAD_TOL
        if method == "conf":
            assert rel_err(gc.cpu(), vc_ref.grad) < GRAD_TOL


def test_tile_backward_grad_conf_is_deterministic(monkeypatch):
    """grad_conf of the tile kernel is a fixed-order reduction (per-CTA partials, summed in tile order): bit-equal runs."""
    monkeypatch.setenv("LT_B200_K1BWD", "tile")
    d = synthetic.make_unproject_inputs(B=2, V=4, C=32, H=32, W=32, N=32, seed=7, train=True, conf="conf")
    gout = torch.randn(2, 32, 32, 32, 32, generator=torch.Generator().manual_seed(8)).cuda()
    runs = [_grads(d, "conf", gout, True)[2] for _ in range(3)]
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])


@pytest.mark.parametrize("shape", [dict(C=32, V=4, N=32), dict(C=40, V=6, N=(8, 8, 16)), dict(C=130, V=2, N=(8, 8, 8))])
def test_scatter_backward_grad_conf_is_deterministic(shape, monkeypatch):
    """The default (scatter) kernel as well: warps of a CTA are summed through shared memory in warp order, CTAs and
    rounds leave one partial row each, rows are summed in row order; no atomics on the grad_conf path. Every
    lanes-per-voxel instantiation (32 / 64 / 256 padded channels), more than one view chunk."""
    monkeypatch.setenv("LT_B200_K1BWD", "scatter")
    d = synthetic.make_unproject_inputs(B=2, H=24, W=24, seed=17, train=True, conf="conf", inside_view=True, **shape)
    n = shape["N"] if isinstance(shape["N"], tuple) else (shape["N"],) * 3
    gout = torch.randn((2, shape["C"]) + n, generator=torch.Generator().manual_seed(8)).cuda()
    runs = [_grads(d, "conf", gout, True)[2] for _ in range(3)]
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2])
    hm_ref = d["heatmaps"].clone().requires_grad_(True)
    vc_ref = d["vol_confidences"].clone().requires_grad_(True)
    ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"], d["coord_volumes"], "conf", vc_ref).backward(gout.cpu())
    assert rel_err(runs[0].cpu(), vc_ref.grad) < GRAD_TOL


def test_degenerate_key_boxes_still_exact(monkeypatch):
    """A camera INSIDE the cuboid makes tile corners fall behind it: the sort keys fall back to the whole image and get
    clamped into the last histogram bin. The order only decides how much merges, never the result."""
    monkeypatch.setenv("LT_B200_K1BWD", "tile")
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=32, seed=21, train=True, inside_view=True,
                                        side=6000.0)
    gout = torch.randn(1, 32, 32, 32, 32, generator=torch.Generator().manual_seed(2)).cuda()
    hm_ref = d["heatmaps"].cuda().requires_grad_(True)
    ref = ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), "softmax")
    ref.backward(gout)
    out, g, _ = _grads(d, "softmax", gout, None)
    assert rel_err(out, ref) < VOL_TOL
    assert rel_err(g, hm_ref.grad) < GRAD_TOL


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE config 4: B=8, 31 views, 32 ch 96x96, 64^3 — one item at its own shape, every aggregation, both gradients
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("method", ["conf", "softmax", "max", "sum"])
def test_cfg4_item_31_views_full_size(method):
    """The many-view code paths (view chunks in k_unproject_fwd / the tile backward, the running-state kernels for
    max/softmax over more than four views) at cfg4's own shape against the reference algorithm run with CUDA tensors
    on the same device (oracle restatement, same ATen calls)."""
    d = synthetic.make_unproject_inputs(B=1, V=31, C=32, H=96, W=96, N=64, seed=1, train=True, conf="conf", dome=True)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    gout = torch.randn(1, 32, 64, 64, 64, generator=torch.Generator().manual_seed(5)).cuda()
    hm_ref = d["heatmaps"].cuda().requires_grad_(True)
    vc_ref = d["vol_confidences"].cuda().requires_grad_(True)
    ref = ref_ops.unproject_heatmaps(hm_ref, proj, coord, method, vc_ref)
    ref.backward(gout)
    hm = d["heatmaps"].cuda().requires_grad_(True)
    vc = d["vol_confidences"].cuda().requires_grad_(True)
    out = lt_b200.unproject_heatmaps(hm, proj, coord, method, vc)
    out.backward(gout)
    assert rel_err(out, ref) < VOL_TOL
    if method == "max":
        # 8.4 M (voxel, channel) arg-max decisions over 31 candidates each: a handful of them are ties to within the
        # fp32 rounding of the samples (1e-7 relative), where our sample and ATen's pick different views and the whole
        # upstream gradient of that voxel moves to another map. Isolated, inherent to `max`; everything else must agree.
        diff = (hm.grad - hm_ref.grad).abs()
        bad = int((diff > GRAD_TOL * hm_ref.grad.abs().max()).sum().item())
        assert bad <= 64, "%d gradient entries off (a flipped tie moves 8 entries)" % bad
    else:
        assert rel_err(hm.grad, hm_ref.grad) < GRAD_TOL
    if method == "conf":
        assert rel_err(vc.grad, vc_ref.grad) < GRAD_TOL


# ---------------------------------------------------------------------------------------------------------------------
# BASELINE config 5: 256^3 — K1 on a 256-deep slab, K2 on the full 256^3 volume (16.7 M voxels per joint)
# ---------------------------------------------------------------------------------------------------------------------
def test_cfg5_unproject_on_a_256_deep_slab():
    """K1 forward/backward on a (256, 64, 64) slab: 1 M voxels per item as cfg5's ranks see them at 4 GPUs (the long
    first axis, 128 tiles along it), against the reference algorithm on the same device."""
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=(256, 64, 64), seed=2, train=True)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    gout = torch.randn(1, 32, 256, 64, 64, generator=torch.Generator().manual_seed(6)).cuda()
    hm_ref = d["heatmaps"].cuda().requires_grad_(True)
    ref = ref_ops.unproject_heatmaps(hm_ref, proj, coord, "softmax")
    ref.backward(gout)
    hm = d["heatmaps"].cuda().requires_grad_(True)
    out = lt_b200.unproject_heatmaps(hm, proj, coord, "softmax")
    out.backward(gout)
    assert rel_err(out, ref) < VOL_TOL
    assert rel_err(hm.grad, hm_ref.grad) < GRAD_TOL


def test_cfg5_softargmax_full_256_cubed():
    """K2 forward/backward at B=1, J=17, 256^3 (nvox = 16.7 M: the 64-bit index paths, 8192 chunks per item) against
    the fp64 reference algorithm on the same device: coordinates to 1e-3 mm, probabilities 1e-5, gradient 1e-4."""
    s = synthetic.make_softargmax_inputs(B=1, J=17, N=256, seed=3, peaked=True, train=True)
    vol, coord = s["volumes"].cuda(), s["coord_volumes"].cuda()
    gc = torch.randn(1, 17, 3, generator=torch.Generator().manual_seed(7)).cuda()
    v = vol.clone().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(v, coord)
    (g,) = torch.autograd.grad(coords, v, gc)
    v64 = vol.double().requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(v64, coord.double())
    (g_ref,) = torch.autograd.grad(c_ref, v64, gc.double())
    assert (coords.double() - c_ref).abs().max().item() < 1e-3
    assert rel_err(prob, p_ref.detach()) < 1e-5
    assert rel_err(g, g_ref) < GRAD_TOL
