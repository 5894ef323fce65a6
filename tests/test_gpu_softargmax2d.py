"""GPU: K4 (integrate_tensor_2d, mvn/utils/op.py:11-47) through the C ABI against golden vectors generated from the
unmodified reference, the float64 oracle at the Algebraic net's shapes, analytic known answers and the DLT it feeds.
Tolerances: 1e-4 px on coordinates (fp32 ulp at 96 px is 7.6e-6), 1e-5 relative on the maps, 1e-4 relative on gradients."""
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden, rel_err
from oracle import ref_ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", golden_names("softargmax2d_"))
def test_golden_forward_backward(name):
    g = load_golden(name)
    sm = bool(g["softmax"])
    hm = g["heatmaps"].cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_2d(hm, sm)
    assert coords.shape == g["coords"].shape and prob.shape == g["prob"].shape
    assert (coords.cpu().double() - g["coords64"]).abs().max().item() < 1e-4
    assert rel_err(prob.cpu(), g["prob64"]) < 1e-5
    (gh,) = torch.autograd.grad(coords, hm, g["grad_coords"].cuda(), retain_graph=True)
    assert rel_err(gh.cpu(), g["grad_heatmaps_coords_only64"]) < 1e-4
    (gh,) = torch.autograd.grad([coords, prob], hm, [g["grad_coords"].cuda(), g["grad_prob"].cuda()])
    assert rel_err(gh.cpu(), g["grad_heatmaps_both64"]) < 1e-4


@pytest.mark.parametrize("softmax", [True, False])
# (.., 96, 96) / (.., 6, 8): float4 shared-memory kernels; (.., 224, 224): the same with 1024 threads; (.., 13, 31) and
# (1,1,1,1): scalar kernels (W % 4 != 0); (.., 256, 256): scalar kernels (map larger than the shared-memory budget)
@pytest.mark.parametrize("shape", [(64, 17, 96, 96), (3, 5, 13, 31), (1, 1, 1, 1), (2, 2, 6, 8), (2, 3, 224, 224),
                                   (1, 2, 256, 256)])
def test_against_float64_oracle(softmax, shape):
    gen = torch.Generator().manual_seed(7)
    hm0 = torch.randn(*shape, generator=gen)
    mult = 100.0 if softmax else 1.0                      # heatmap_multiplier of human36m_alg.yaml
    if not softmax and hm0.numel() == 1:
        hm0 = hm0.abs() + 0.1                             # an all-negative map is 0/0 in ReLU mode (checked below)
    h64 = (hm0.double() * mult).requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_2d(h64, softmax)
    gc = torch.randn(c_ref.shape, generator=gen)
    (g_ref,) = torch.autograd.grad(c_ref, h64, gc.double(), retain_graph=True)
    hm = hm0.cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_2d(hm, softmax, multiplier=mult)
    (gh,) = torch.autograd.grad(coords, hm, gc.cuda(), retain_graph=True)
    # 1e-4 px at the reference's 96x96 maps; fp32 sums over the 50-65 k pixels of the large maps get 2e-6 relative
    assert (coords.double().cpu() - c_ref).abs().max().item() < max(1e-4, 2e-6 * c_ref.abs().max().item())
    assert rel_err(prob.cpu(), p_ref) < 1e-5
    assert rel_err(gh.cpu(), g_ref * mult) < 1e-4     # d/d(unscaled input) = mult * d/d(scaled input)
    gp = torch.randn(p_ref.shape, generator=gen)        # gradient through both outputs
    (g_ref2,) = torch.autograd.grad([c_ref, p_ref], h64, [gc.double(), gp.double()])
    (gh2,) = torch.autograd.grad([coords, prob], hm, [gc.cuda(), gp.cuda()])
    assert rel_err(gh2.cpu(), g_ref2 * mult) < 1e-4
    # the keyword equals pre-scaling on the caller side (triangulation.py:164)
    # (rounding hm*100 to fp32 first moves each logit by up to ~2.4e-5, i.e. each probability by that relative amount:
    #  with two near-equal peaks ~90 px apart that is ~5e-4 px: these tolerances are the caller-side rounding, not the kernel)
    c2, p2 = lt_b200.integrate_tensor_2d(hm.detach() * mult, softmax)
    assert (c2 - coords).abs().max().item() < 2e-3 and rel_err(p2, prob.detach()) < 2e-4


def test_relu_mode_with_no_mass_is_nan_like_the_reference():
    # op.py:39-41 divides by the total mass without a guard: torch gives nan, so do we
    hm = -torch.rand(1, 2, 5, 7) - 0.1
    c_ref, _ = ref_ops.integrate_tensor_2d(hm.double(), False)
    coords, prob = lt_b200.integrate_tensor_2d(hm.cuda(), False)
    assert torch.isnan(c_ref).all() and torch.isnan(coords).all() and (prob == 0).all()


def test_known_answers_and_errors():
    # a sharp peak -> its pixel; a uniform map -> the centre of the pixel grid ((W-1)/2, (H-1)/2)
    hm = torch.zeros(1, 2, 9, 12)
    hm[0, 0, 6, 4] = 50.0
    coords, prob = lt_b200.integrate_tensor_2d(hm.cuda())
    assert torch.allclose(coords[0, 0].cpu(), torch.tensor([4.0, 6.0]), atol=1e-4)
    assert torch.allclose(coords[0, 1].cpu(), torch.tensor([5.5, 4.0]), atol=1e-4)
    assert abs(prob[0, 1].sum().item() - 1.0) < 1e-5
    empty, _ = lt_b200.integrate_tensor_2d(torch.zeros(0, 17, 8, 8).cuda())
    assert empty.shape == (0, 17, 2)
    with pytest.raises(ValueError):
        lt_b200.integrate_tensor_2d(torch.zeros(2, 8, 8).cuda())


def test_algebraic_tail_matches_the_oracle():
    """integrate_tensor_2d -> heat-map-to-image scaling -> DLT, as AlgebraicTriangulationNet.forward does
    (triangulation.py:164-188), against the oracle's two functions chained in float64."""
    from lt_b200 import synthetic
    t = synthetic.make_dlt_inputs(B=4, V=4, J=17, seed=3)
    B, V, J = 4, 4, 17
    gen = torch.Generator().manual_seed(11)
    hm0 = torch.randn(B * V, J, 96, 96, generator=gen) * 0.1
    pts = (t["points"].reshape(B * V, J, 2) / 4.0).round().long().clamp(2, 93)     # 384 px image -> 96 px heat map
    for i in range(B * V):
        for j in range(J):
            hm0[i, j, pts[i, j, 1], pts[i, j, 0]] += 1.0
    k_ref, _ = ref_ops.integrate_tensor_2d(hm0.double() * 100.0, True)
    x_ref = ref_ops.triangulate_batch_of_points(t["proj_matricies"].double(), k_ref.view(B, V, J, 2) * 4.0,
                                                t["confidences"].double(), out_dtype=None)
    k, _ = lt_b200.integrate_tensor_2d(hm0.cuda(), True, multiplier=100.0)
    x = lt_b200.triangulate_batch_of_points(t["proj_matricies"].cuda(), k.view(B, V, J, 2) * 4.0, t["confidences"].cuda())
    # 1e-4 px of key-point error moves the triangulated point by ~1e-3 mm at this geometry
    assert (x.double().cpu() - x_ref).abs().max().item() < 2e-2
