"""GPU: K3 (triangulate_batch_of_points) through the C ABI. Tolerance (BASELINE.json): 1e-2 mm against the float64
run of the reference (its own fp32 run is only within ~0.3 mm of that), 1e-4 relative on gradients."""
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden, rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", golden_names("dlt_"))
def test_golden_forward_backward(name):
    g = load_golden(name)
    conf = g.get("confidences")
    pts = g["points"].cuda().requires_grad_(True)
    cf = conf.cuda().requires_grad_(True) if conf is not None else None
    out = lt_b200.triangulate_batch_of_points(g["proj_matricies"].cuda(), pts, cf)
    assert out.shape == g["out64"].shape and out.dtype == torch.float32
    err64 = (out.cpu().double() - g["out64"]).abs().max().item()
    err32 = (out.cpu() - g["out32"]).abs().max().item()
    ref_gap = (g["out32"].double() - g["out64"]).abs().max().item()
    # fp32 output rounding of coordinates up to ~8000 mm is 5e-4 mm
    assert err64 < 1e-2, (err64, err32, ref_gap)
    assert err32 < ref_gap + 1e-2
    out.backward(g["grad_out"].cuda())
    assert rel_err(pts.grad.cpu(), g["grad_points64"]) < 1e-4
    if cf is not None:
        assert rel_err(cf.grad.cpu(), g["grad_confidences64"]) < 1e-4


@pytest.mark.parametrize("shape", [dict(B=64, V=4, J=17), dict(B=8, V=31, J=17, dome=True), dict(B=2, V=2, J=3),
                                   dict(B=1, V=40, J=2, dome=True), dict(B=5, V=3, J=1, with_conf=False)])
def test_against_float64_oracle(shape):
    d = synthetic.make_dlt_inputs(seed=4, **shape)
    conf = d["confidences"]
    p64 = d["points"].double().requires_grad_(True)
    c64 = conf.double().requires_grad_(True) if conf is not None else None
    ref = ref_ops.triangulate_batch_of_points(d["proj_matricies"].double(), p64, c64, out_dtype=None)
    gout = torch.randn(ref.shape, generator=torch.Generator().manual_seed(2))
    ref.backward(gout.double())
    pts = d["points"].cuda().requires_grad_(True)
    cf = conf.cuda().requires_grad_(True) if conf is not None else None
    out = lt_b200.triangulate_batch_of_points(d["proj_matricies"].cuda(), pts, cf)
    assert (out.cpu().double() - ref).abs().max().item() < 1e-2
    out.backward(gout.cuda())
    assert rel_err(pts.grad.cpu(), p64.grad) < 1e-4
    if cf is not None:
        assert rel_err(cf.grad.cpu(), c64.grad) < 1e-4


def test_noise_free_projections_recover_the_points():
    d = synthetic.make_dlt_inputs(B=16, V=4, J=17, seed=8, noise_px=0.0)
    out = lt_b200.triangulate_batch_of_points(d["proj_matricies"].cuda(), d["points"].cuda(), d["confidences"].cuda())
    assert (out.cpu() - d["points_3d"]).abs().max().item() < 0.05
    # confidences do not matter when all rays meet in one point
    out2 = lt_b200.triangulate_batch_of_points(d["proj_matricies"].cuda(), d["points"].cuda())
    assert (out2 - out).abs().max().item() < 0.05


def test_single_point_wrapper_and_errors():
    d = synthetic.make_dlt_inputs(B=1, V=4, J=2, seed=9)
    P, pts, conf = d["proj_matricies"].cuda(), d["points"].cuda(), d["confidences"].cuda()
    batch = lt_b200.triangulate_batch_of_points(P, pts, conf)
    one = lt_b200.triangulate_point_from_multiple_views_linear_torch(P[0], pts[0, :, 1], conf[0, :, 1])
    assert one.shape == (3,) and torch.equal(one, batch[0, 1])
    with pytest.raises(AssertionError):  # multiview.py:143
        lt_b200.triangulate_batch_of_points(P[:, :3], pts, conf)
    with pytest.raises(RuntimeError):
        big = synthetic.make_dlt_inputs(B=1, V=65, J=1, seed=1, dome=True)
        lt_b200.triangulate_batch_of_points(big["proj_matricies"].cuda(), big["points"].cuda())
    empty = lt_b200.triangulate_batch_of_points(P[:0], pts[:0], conf[:0])
    assert empty.shape == (0, 2, 3)
