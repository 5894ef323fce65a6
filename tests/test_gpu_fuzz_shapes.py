"""GPU: seeded random shapes for every op against the oracle (the reference's own ATen calls on the CPU). The fixed cases
elsewhere pin the known corners; this sweep walks the kernel dispatch space — channel counts around the lane-width
boundaries (32 / 64 / 128), view counts around the view chunk (4), volume extents around the tile (8) and vector (4, 8)
widths, image extents that are not multiples of anything — so that a shape nobody thought of still has a witness.
Same tolerances as the fixed tests (BASELINE.json)."""
import numpy as np
import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu


def _draw_unproject(seed):
    r = np.random.RandomState(1000 + seed)
    C = int(r.choice([1, 3, 17, 31, 32, 33, 48, 64, 65, 100, 128, 129]))
    V = int(r.choice([1, 2, 3, 4, 5, 7, 8, 9]))
    big = C > 64 or V > 5
    return dict(B=int(r.randint(1, 3)), V=V, C=C, H=int(r.randint(5, 21)), W=int(r.randint(5, 21)),
                N=tuple(int(v) for v in r.randint(1, 8 if big else 12, size=3)),
                method=str(r.choice(["sum", "conf", "conf_norm", "max", "softmax"])),
                align_corners=bool(r.randint(0, 2)), dome=bool(r.randint(0, 2)), inside_view=bool(r.randint(0, 2)))


@pytest.mark.parametrize("seed", range(24))
def test_unproject_random_shape(seed):
    cfg = _draw_unproject(seed)
    method, ac = cfg.pop("method"), cfg.pop("align_corners")
    conf = method if method.startswith("conf") else None
    d = synthetic.make_unproject_inputs(seed=seed, train=True, conf=conf, **cfg)
    hm_ref = d["heatmaps"].clone().requires_grad_(True)
    vc_ref = d["vol_confidences"].clone().requires_grad_(True) if conf else None
    ref = ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"], d["coord_volumes"], method, vc_ref, align_corners=ac)
    gout = torch.randn(ref.shape, generator=torch.Generator().manual_seed(seed))
    ref.backward(gout)
    hm = d["heatmaps"].cuda().requires_grad_(True)
    vc = d["vol_confidences"].cuda().requires_grad_(True) if conf else None
    out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), method, vc,
                                     align_corners=ac)
    tag = "%s %s ac=%s" % (cfg, method, ac)
    assert out.shape == ref.shape, tag
    if ref.abs().max().item() > 0:
        assert rel_err(out.cpu(), ref) < 1e-5, tag
    else:                                           # every view behind the camera / outside the image
        assert out.abs().max().item() == 0.0, tag
    out.backward(gout.cuda())
    if hm_ref.grad.abs().max().item() > 0:
        assert rel_err(hm.grad.cpu(), hm_ref.grad) < 1e-4, tag
    else:
        assert hm.grad.abs().max().item() == 0.0, tag
    if conf and vc_ref.grad.abs().max().item() > 0:
        assert rel_err(vc.grad.cpu(), vc_ref.grad) < 1e-4, tag


@pytest.mark.parametrize("seed", range(12))
def test_soft_argmax_3d_random_shape(seed):
    r = np.random.RandomState(2000 + seed)
    B, J = int(r.randint(1, 4)), int(r.randint(1, 20))
    N = tuple(int(v) for v in r.randint(1, 14, size=3))
    softmax = bool(r.randint(0, 2))
    mult = float(r.choice([1.0, 0.5, 3.0]))
    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=seed, train=True)
    v64 = (d["volumes"].double() * mult).requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(v64, d["coord_volumes"].double(), softmax=softmax)
    gen = torch.Generator().manual_seed(seed)
    gc, gp = torch.randn(c_ref.shape, generator=gen), torch.randn(p_ref.shape, generator=gen)
    (g_ref,) = torch.autograd.grad([c_ref, p_ref], v64, [gc.double(), gp.double()])
    vol = d["volumes"].cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, d["coord_volumes"].cuda(), softmax=softmax,
                                                                multiplier=mult)
    (gv,) = torch.autograd.grad([coords, prob], vol, [gc.cuda(), gp.cuda()])
    tag = "B=%d J=%d N=%s softmax=%s mult=%s" % (B, J, N, softmax, mult)
    tol = max(1e-3, 4 * 1.2e-7 * c_ref.abs().max().item())
    assert (coords.double().cpu() - c_ref).abs().max().item() < tol, tag
    assert rel_err(prob.cpu(), p_ref) < 1e-5, tag
    assert rel_err(gv.cpu(), g_ref * mult) < 1e-4, tag


@pytest.mark.parametrize("seed", range(12))
def test_soft_argmax_2d_random_shape(seed):
    r = np.random.RandomState(3000 + seed)
    B, J, H, W = int(r.randint(1, 4)), int(r.randint(1, 20)), int(r.randint(1, 40)), int(r.randint(1, 40))
    softmax = bool(r.randint(0, 2))
    mult = float(r.choice([1.0, 10.0, 100.0])) if softmax else 1.0
    gen = torch.Generator().manual_seed(seed)
    hm0 = torch.randn(B, J, H, W, generator=gen)
    if not softmax:
        hm0 = hm0 + 0.5
        hm0[:, :, 0, 0] = 1.0                      # every map keeps some mass (0/0 is its own test)
    h64 = (hm0.double() * mult).requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_2d(h64, softmax)
    gc, gp = torch.randn(c_ref.shape, generator=gen), torch.randn(p_ref.shape, generator=gen)
    (g_ref,) = torch.autograd.grad([c_ref, p_ref], h64, [gc.double(), gp.double()])
    hm = hm0.cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_2d(hm, softmax, multiplier=mult)
    (gh,) = torch.autograd.grad([coords, prob], hm, [gc.cuda(), gp.cuda()])
    tag = "B=%d J=%d H=%d W=%d softmax=%s mult=%s" % (B, J, H, W, softmax, mult)
    assert (coords.double().cpu() - c_ref).abs().max().item() < 1e-4, tag
    assert rel_err(prob.cpu(), p_ref) < 1e-5, tag
    assert rel_err(gh.cpu(), g_ref * mult) < 1e-4, tag


@pytest.mark.parametrize("seed", range(8))
def test_dlt_random_shape(seed):
    r = np.random.RandomState(4000 + seed)
    B, V, J = int(r.randint(1, 6)), int(r.choice([2, 3, 4, 5, 8, 17, 33, 64])), int(r.randint(1, 20))
    with_conf = bool(r.randint(0, 2))
    t = synthetic.make_dlt_inputs(B=B, V=V, J=J, seed=seed, with_conf=with_conf, dome=V > 8)
    conf = t["confidences"]
    ref = ref_ops.triangulate_batch_of_points(t["proj_matricies"].double(), t["points"].double(),
                                              conf.double() if with_conf else None, out_dtype=None)
    out = lt_b200.triangulate_batch_of_points(t["proj_matricies"].cuda(), t["points"].cuda(),
                                              conf.cuda() if with_conf else None)
    assert (out.double().cpu() - ref).abs().max().item() < 1e-2, "B=%d V=%d J=%d conf=%s" % (B, V, J, with_conf)
