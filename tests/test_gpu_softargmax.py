"""GPU: K2 (integrate_tensor_3d_with_coordinates) through the C ABI against golden vectors, the float64 oracle,
analytic known answers and the slab-split entry points. Tolerances (BASELINE.json): 1e-3 mm on coordinates,
1e-5 relative on the normalised volume, 1e-4 relative on gradients."""
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden, rel_err
from lt_b200 import slab, synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu


def coord_tol(ref):
    # un-normalised relu sums reach 1e4..1e5 mm where one fp32 ulp already exceeds 1e-3
    return max(1e-3, 4 * 1.2e-7 * ref.abs().max().item())


@pytest.mark.parametrize("name", golden_names("softargmax_"))
def test_golden_forward_backward(name):
    g = load_golden(name)
    sm = bool(g["softmax"])
    vol = g["volumes"].cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, g["coord_volumes"].cuda(), softmax=sm)
    assert coords.shape == g["coords"].shape and prob.shape == g["prob"].shape
    assert (coords.cpu().double() - g["coords64"]).abs().max().item() < coord_tol(g["coords64"])
    assert rel_err(prob.cpu(), g["prob64"]) < 1e-5
    (gv,) = torch.autograd.grad(coords, vol, g["grad_coords"].cuda(), retain_graph=True)
    assert rel_err(gv.cpu(), g["grad_volumes_coords_only64"]) < 1e-4
    (gv,) = torch.autograd.grad([coords, prob], vol, [g["grad_coords"].cuda(), g["grad_prob"].cuda()])
    assert rel_err(gv.cpu(), g["grad_volumes_both64"]) < 1e-4


@pytest.mark.parametrize("softmax", [True, False])
@pytest.mark.parametrize("shape", [dict(B=2, J=17, N=32), dict(B=1, J=5, N=(7, 9, 11)), dict(B=3, J=1, N=(2, 2, 2)),
                                   dict(B=1, J=2, N=(33, 20, 20))])
@pytest.mark.parametrize("peaked", [False, True])
def test_against_float64_oracle(softmax, shape, peaked):
    d = synthetic.make_softargmax_inputs(seed=6, train=True, peaked=peaked, **shape)
    v64 = d["volumes"].double().requires_grad_(True)
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(v64, d["coord_volumes"].double(), softmax)
    gen = torch.Generator().manual_seed(1)
    gc, gp = torch.randn(c_ref.shape, generator=gen), torch.randn(p_ref.shape, generator=gen)
    (g_ref,) = torch.autograd.grad([c_ref, p_ref], v64, [gc.double(), gp.double()])
    vol = d["volumes"].cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, d["coord_volumes"].cuda(), softmax)
    assert (coords.cpu().double() - c_ref).abs().max().item() < coord_tol(c_ref)
    assert rel_err(prob.cpu(), p_ref) < 1e-5
    (gv,) = torch.autograd.grad([coords, prob], vol, [gc.cuda(), gp.cuda()])
    assert rel_err(gv.cpu(), g_ref) < 1e-4


def test_multiplier_keyword_equals_prescaled_input():
    d = synthetic.make_softargmax_inputs(B=1, J=3, N=16, seed=2)
    vol, coord = d["volumes"].cuda(), d["coord_volumes"].cuda()
    for mult in (0.37, 3.0, -1.5):
        c1, p1 = lt_b200.integrate_tensor_3d_with_coordinates(vol * mult, coord)
        v = vol.clone().requires_grad_(True)
        c2, p2 = lt_b200.integrate_tensor_3d_with_coordinates(v, coord, multiplier=mult)
        assert (c1 - c2).abs().max().item() < 1e-3 and rel_err(p2, p1) < 1e-5
        (g2,) = torch.autograd.grad(c2, v, torch.ones_like(c2))
        v1 = vol.clone().requires_grad_(True)
        c1, _ = lt_b200.integrate_tensor_3d_with_coordinates(v1 * mult, coord)
        (g1,) = torch.autograd.grad(c1, v1, torch.ones_like(c1))
        assert rel_err(g2, g1) < 1e-4


def test_known_answers():
    d = synthetic.make_softargmax_inputs(B=2, J=4, N=16, seed=3, train=True)
    coord = d["coord_volumes"].cuda()
    # uniform volume -> centroid of the coordinate volume; probabilities 1/nvox
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(torch.zeros(2, 4, 16, 16, 16, device="cuda"), coord)
    centroid = coord.double().reshape(2, -1, 3).mean(1).float()
    assert (coords - centroid.unsqueeze(1)).abs().max().item() < 1e-3
    assert torch.allclose(prob, torch.full_like(prob, 1.0 / 4096), rtol=1e-5)
    # one dominant voxel -> its coordinate; probabilities sum to one
    vol = torch.zeros(2, 4, 16, 16, 16, device="cuda")
    vol[:, :, 3, 7, 11] = 200.0
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord)
    assert (coords - coord[:, 3, 7, 11].unsqueeze(1)).abs().max().item() < 1e-3
    assert (prob.flatten(2).sum(2) - 1).abs().max().item() < 1e-5
    # relu mode: no normalisation (op.py:91), negative voxels contribute nothing
    vol = -torch.ones(2, 4, 16, 16, 16, device="cuda")
    vol[:, :, 1, 2, 3] = 2.0
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord, softmax=False)
    assert (coords - 2.0 * coord[:, 1, 2, 3].unsqueeze(1)).abs().max().item() < 1e-2
    assert prob.sum().item() == pytest.approx(2.0 * 8)


def test_empty_inputs():
    coord = torch.zeros(0, 4, 4, 4, 3, device="cuda")
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(torch.zeros(0, 3, 4, 4, 4, device="cuda"), coord)
    assert coords.shape == (0, 3, 3) and prob.shape == (0, 3, 4, 4, 4)
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(torch.zeros(2, 0, 4, 4, 4, device="cuda"),
                                                                torch.zeros(2, 4, 4, 4, 3, device="cuda"))
    assert coords.shape == (2, 0, 3)


@pytest.mark.parametrize("softmax", [True, False])
def test_slab_pieces_on_one_device(softmax):
    """partial / merge / finalize on 1..4 slabs of one GPU == unsharded result (the N-GPU data path, minus NCCL)."""
    d = synthetic.make_softargmax_inputs(B=2, J=17, N=(20, 16, 16), seed=7, train=True, peaked=True)
    vol, coord = d["volumes"].cuda(), d["coord_volumes"].cuda()
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(d["volumes"].double(), d["coord_volumes"].double(),
                                                                softmax)
    for R in (1, 2, 3, 4):
        spans = [slab.slab_bounds(20, R, r) for r in range(R)]
        parts = torch.stack([slab.CudaSlabOps.partial(vol[:, :, a:b].contiguous(), coord[:, a:b].contiguous(),
                                                      int(softmax), 1.0) for a, b in spans])
        merged = slab.CudaSlabOps.merge(parts, int(softmax))
        probs = []
        for a, b in spans:
            coords, p = slab.CudaSlabOps.finalize(vol[:, :, a:b].contiguous(), merged, int(softmax), 1.0)
            probs.append(p)
        assert (coords.cpu().double() - c_ref).abs().max().item() < coord_tol(c_ref)
        assert rel_err(torch.cat(probs, 2).cpu(), p_ref) < 1e-5


def test_full_size_properties():
    """BASELINE shape (B=4 of the 16, J=17, 64^3): sums to one, coordinates inside the cuboid, matches stock torch
    run on the same GPU, and the gradient through coordinates sums to zero per joint (softmax Jacobian)."""
    d = synthetic.make_softargmax_inputs(B=4, J=17, N=64, seed=0, train=True)
    vol, coord = d["volumes"].cuda().requires_grad_(True), d["coord_volumes"].cuda()
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord)
    assert (prob.flatten(2).sum(2) - 1).abs().max().item() < 1e-4
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(vol.detach().double(), coord.double())
    assert (coords.double() - c_ref).abs().max().item() < 1e-3
    assert rel_err(prob, p_ref) < 1e-5
    (gv,) = torch.autograd.grad(coords, vol, torch.randn_like(coords))
    assert gv.flatten(2).sum(2).abs().max().item() < 1e-3 * gv.abs().max().item() * 64


@pytest.mark.parametrize("name", golden_names("softargmax3dindex_"))
def test_index_coordinate_variant_golden(name):
    """integrate_tensor_3d (op.py:50-81): K6 unit grid + K2, against the unmodified reference in fp32 and fp64"""
    g = load_golden(name)
    sm = bool(g["softmax"])
    vol = g["volumes"].cuda().requires_grad_(True)
    coords, prob = lt_b200.integrate_tensor_3d(vol, sm)
    assert coords.shape == g["coords"].shape and prob.shape == g["prob"].shape
    assert (coords.cpu().double() - g["coords64"]).abs().max().item() < 1e-4       # voxel units
    assert rel_err(prob.cpu(), g["prob64"]) < 1e-5
    (gv,) = torch.autograd.grad(coords, vol, g["grad_coords"].cuda())
    assert rel_err(gv.cpu(), g["grad_volumes64"]) < 1e-4


def test_index_coordinate_variant_known_answers():
    vol = torch.full((1, 2, 4, 5, 6), -30.0)
    vol[0, 0, 3, 1, 4] = 30.0
    vol[0, 1] = 0.0
    coords, prob = lt_b200.integrate_tensor_3d(vol.cuda())
    assert torch.allclose(coords[0, 0].cpu(), torch.tensor([3.0, 1.0, 4.0]), atol=1e-4)
    assert torch.allclose(coords[0, 1].cpu(), torch.tensor([1.5, 2.0, 2.5]), atol=1e-4)   # uniform -> centre of the grid
    relu_coords, relu_vol = lt_b200.integrate_tensor_3d(vol.cuda(), softmax=False)
    assert torch.allclose(relu_coords[0, 0].cpu(), torch.tensor([3.0, 1.0, 4.0]), atol=1e-4)
    assert torch.isnan(relu_coords[0, 1]).all() and float(relu_vol[0, 1].abs().max()) == 0.0   # 0/0 like the reference
    with pytest.raises(ValueError):
        lt_b200.integrate_tensor_3d(vol[0].cuda())
