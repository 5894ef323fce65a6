"""GPU: K0+K1 (unproject_heatmaps) through the C ABI against (a) the golden vectors of the unmodified reference,
(b) the oracle on seeded inputs, (c) analytic known answers, (d) stock torch on the same GPU at full size.

Tolerances (BASELINE.json): 1e-5 relative on the volume, 1e-4 relative on gradients (atomics reorder the sums);
"relative" = max|ours-ref| / max|ref| (conftest.rel_err)."""
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden, rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu
VOL_TOL, GRAD_TOL = 1e-5, 1e-4


def _cuda(t):
    return None if t is None else t.cuda()


@pytest.mark.parametrize("name", golden_names("unproject_"))
def test_golden_forward_backward(name):
    g = load_golden(name)
    hm = g["heatmaps"].cuda().requires_grad_(True)
    vc = g.get("vol_confidences")
    vc = vc.cuda().requires_grad_(True) if vc is not None else None
    out = lt_b200.unproject_heatmaps(hm, g["proj_matricies"].cuda(), g["coord_volumes"].cuda(), g["method"], vc,
                                     align_corners=bool(g["align_corners"]))
    assert out.shape == g["out"].shape and out.dtype == torch.float32 and out.is_cuda
    assert rel_err(out.cpu(), g["out"]) < VOL_TOL
    out.backward(g["grad_out"].cuda())
    assert rel_err(hm.grad.cpu(), g["grad_heatmaps"]) < GRAD_TOL
    if vc is not None:
        assert rel_err(vc.grad.cpu(), g["grad_vol_confidences"]) < GRAD_TOL


@pytest.mark.parametrize("method", ["sum", "conf", "conf_norm", "max", "softmax"])
@pytest.mark.parametrize("shape", [
    dict(B=2, V=4, C=32, H=24, W=24, N=16),                 # fast path, aligned runs, several tiles per axis
    dict(B=1, V=3, C=17, H=12, W=20, N=(9, 10, 11)),        # joint heat maps: padded channels, ragged tiles, H != W
    dict(B=1, V=2, C=40, H=16, W=16, N=(4, 6, 16)),         # 64-channel instantiation (8 lanes per voxel)
    dict(B=1, V=9, C=32, H=16, W=16, N=8, dome=True),       # more than one view chunk -> many-view kernels
    dict(B=1, V=2, C=130, H=8, W=8, N=(3, 4, 8)),           # 256-channel instantiation (32 lanes per voxel)
])
def test_against_oracle(method, shape):
    conf = method if method.startswith("conf") else None
    d = synthetic.make_unproject_inputs(seed=5, train=True, conf=conf, inside_view=True, **shape)
    hm_ref = d["heatmaps"].clone().requires_grad_(True)
    vc_ref = d["vol_confidences"].clone().requires_grad_(True) if conf else None
    ref = ref_ops.unproject_heatmaps(hm_ref, d["proj_matricies"], d["coord_volumes"], method, vc_ref)
    gout = torch.randn(ref.shape, generator=torch.Generator().manual_seed(9))
    ref.backward(gout)

    hm = d["heatmaps"].cuda().requires_grad_(True)
    vc = d["vol_confidences"].cuda().requires_grad_(True) if conf else None
    out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), method, vc)
    assert rel_err(out.cpu(), ref) < VOL_TOL
    out.backward(gout.cuda())
    assert rel_err(hm.grad.cpu(), hm_ref.grad) < GRAD_TOL
    if conf:
        assert rel_err(vc.grad.cpu(), vc_ref.grad) < GRAD_TOL


def test_constant_map_counts_valid_views():
    """sum of a constant map == constant * number of views that see the voxel (fully inside their image)."""
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=16, seed=2, side=600.0)
    hm = torch.full_like(d["heatmaps"], 0.5).cuda()
    out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), "sum")
    assert synthetic.inside_fraction(d["proj_matricies"], d["coord_volumes"], 96, 96) > 0.999
    assert torch.allclose(out, torch.full_like(out, 2.0), atol=1e-5)


def test_behind_camera_is_exactly_zero_and_gets_no_gradient():
    d = synthetic.make_unproject_inputs(B=1, V=1, C=32, H=16, W=16, N=8, seed=1, inside_view=True)
    pts = torch.cat([d["coord_volumes"][0].reshape(-1, 3), torch.ones(512, 1)], 1)
    behind = (pts @ d["proj_matricies"][0, 0, 2]) <= 0
    assert behind.any()
    hm = (d["heatmaps"].abs() + 1.0).cuda().requires_grad_(True)
    out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), "sum")
    flat = out[0].reshape(32, -1).cpu()
    assert (flat[:, behind] == 0).all()
    # gradient only flows from voxels in front of the camera
    gout = torch.zeros_like(out)
    gout[0].view(32, -1)[:, behind.cuda()] = 1.0
    out.backward(gout)
    assert hm.grad.abs().max().item() == 0.0


def test_empty_and_degenerate_inputs():
    d = synthetic.make_unproject_inputs(B=1, V=2, C=4, H=8, W=8, N=4)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    out = lt_b200.unproject_heatmaps(d["heatmaps"][:0].cuda(), proj[:0], coord[:0], "sum")
    assert out.shape == (0, 4, 4, 4, 4)
    out = lt_b200.unproject_heatmaps(d["heatmaps"].cuda(), proj, coord[:, :0], "softmax")
    assert out.shape == (1, 4, 0, 4, 4)
    # non-contiguous / double inputs are accepted like the reference accepts them; output stays fp32 (op.py:104)
    hm = d["heatmaps"].cuda().double().transpose(3, 4).contiguous().transpose(3, 4)
    out = lt_b200.unproject_heatmaps(hm, proj.double(), coord.double(), "sum")
    ref = ref_ops.unproject_heatmaps(d["heatmaps"], d["proj_matricies"], d["coord_volumes"], "sum")
    assert out.dtype == torch.float32 and rel_err(out.cpu(), ref) < VOL_TOL


def test_full_size_against_stock_torch_on_gpu():
    """BASELINE config 1/3 shape (64^3, 4 views, 32 ch, 96x96): the reference algorithm (oracle restatement, same
    ATen calls) run with CUDA tensors on the same device vs the fused kernels; all four aggregations + gradients."""
    d = synthetic.make_unproject_inputs(B=2, V=4, C=32, H=96, W=96, N=64, seed=0, train=True, conf="conf")
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    gout = torch.randn(2, 32, 64, 64, 64, generator=torch.Generator().manual_seed(3)).cuda()
    for method in ("softmax", "sum", "conf", "max"):
        hm_ref = d["heatmaps"].cuda().requires_grad_(True)
        vc_ref = d["vol_confidences"].cuda().requires_grad_(True)
        ref = ref_ops.unproject_heatmaps(hm_ref, proj, coord, method, vc_ref)
        ref.backward(gout)
        hm = d["heatmaps"].cuda().requires_grad_(True)
        vc = d["vol_confidences"].cuda().requires_grad_(True)
        out = lt_b200.unproject_heatmaps(hm, proj, coord, method, vc)
        out.backward(gout)
        assert rel_err(out, ref) < VOL_TOL, method
        assert rel_err(hm.grad, hm_ref.grad) < GRAD_TOL, method
        if method == "conf":
            assert rel_err(vc.grad, vc_ref.grad) < GRAD_TOL


def test_linearity_in_the_features():
    """size-independent property at full size: sum/conf aggregation is linear in the feature maps."""
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=64, seed=4, train=True, conf="conf")
    proj, coord, vc = d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), d["vol_confidences"].cuda()
    a = d["heatmaps"].cuda()
    b = torch.randn(a.shape, generator=torch.Generator().manual_seed(8)).cuda()
    for method in ("sum", "conf"):
        fa = lt_b200.unproject_heatmaps(a, proj, coord, method, vc)
        fb = lt_b200.unproject_heatmaps(b, proj, coord, method, vc)
        fab = lt_b200.unproject_heatmaps(a + 2.0 * b, proj, coord, method, vc)
        assert rel_err(fab, fa + 2.0 * fb) < 1e-5
