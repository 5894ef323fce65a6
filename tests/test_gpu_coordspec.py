"""GPU: coordinates computed INSIDE K1 / K2 from the 18-float spec (lt_b200.CoordSpec, the lt_*_spec entry points,
SURVEY.md §8f row 1) against the same ops fed with the coordinate tensor K6 builds from the same spec, and against the
reference's own construction (mvn/models/triangulation.py:306-341 restated in oracle.ref_ops.build_coord_volume).

Bar: bit-identical to the tensor path wherever the kernel is deterministic (K1 forward, all of K2: one definition of the
arithmetic, csrc/coordspec.cuh); K1's gradients to 1e-4 (unordered reductions); against the reference's coordinates
exactly for un-rotated items and to 1 ulp of the coordinates for rotated ones (the K6 tests)."""
import numpy as np
import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic, volumetric

pytestmark = pytest.mark.gpu


def _spec(B, size, seed, rotated=True, transfer=False):
    rng = np.random.RandomState(seed)
    base = rng.normal(0.0, 100.0, size=(B, 3))
    thetas = rng.uniform(0.0, 2.0 * np.pi, size=B) if rotated else np.zeros(B)
    specs = volumetric.coord_volume_specs(base, 2500.0, size, thetas).cuda()
    return lt_b200.CoordSpec(specs, size, transfer), base


@pytest.mark.parametrize("method", ["sum", "conf", "max", "softmax"])
@pytest.mark.parametrize("case", [
    dict(size=(16, 16, 16), V=4, rotated=True, transfer=False),
    dict(size=(9, 12, 12), V=3, rotated=True, transfer=True),      # ragged tiles + the CMU -> Human3.6M swap
    dict(size=(8, 8, 13), V=2, rotated=False, transfer=False),     # N2 % 8 != 0: scalar store path
    dict(size=(8, 8, 8), V=9, rotated=True, transfer=False),       # view chunks / running-state kernels
])
def test_unproject_with_spec_equals_tensor_path(method, case):
    B, C, H, W = 2, 32, 24, 24
    cs, base = _spec(B, case["size"], 3, case["rotated"], case["transfer"])
    coord = cs.materialize()
    proj = synthetic.make_cameras(B, case["V"], base, 7, (H, W), dome=case["V"] > 4, inside_view=True).cuda()
    g = torch.Generator().manual_seed(5)
    feats = torch.randn(B, case["V"], C, H, W, generator=g)
    conf = torch.rand(B, case["V"], C, generator=g).cuda() if method == "conf" else None
    gout = torch.randn((B, C) + case["size"], generator=g).cuda()
    outs = []
    for cv in (coord, cs):
        hm = feats.cuda().requires_grad_(True)
        vc = conf.clone().requires_grad_(True) if conf is not None else None
        out = lt_b200.unproject_heatmaps(hm, proj, cv, method, vc)
        out.backward(gout)
        outs.append((out.detach(), hm.grad, vc.grad if vc is not None else None))
    assert torch.equal(outs[0][0], outs[1][0])                      # same arithmetic, same bits
    assert rel_err(outs[1][1], outs[0][1]) < 1e-4
    if conf is not None:
        assert rel_err(outs[1][2], outs[0][2]) < 1e-4


@pytest.mark.parametrize("tile", ["0", "1", "2", "3"])
def test_tile_backward_with_spec(tile, monkeypatch):
    monkeypatch.setenv("LT_B200_K1BWD", "tile")
    monkeypatch.setenv("LT_B200_K1BWD_TILE", tile)
    B, V, C, H, W, size = 1, 4, 32, 32, 32, (16, 16, 16)
    cs, base = _spec(B, size, 9)
    proj = synthetic.make_cameras(B, V, base, 2, (H, W)).cuda()
    g = torch.Generator().manual_seed(6)
    feats, gout = torch.randn(B, V, C, H, W, generator=g), torch.randn((B, C) + size, generator=g).cuda()
    grads = []
    for cv in (cs.materialize(), cs):
        hm = feats.cuda().requires_grad_(True)
        lt_b200.unproject_heatmaps(hm, proj, cv, "softmax").backward(gout)
        grads.append(hm.grad)
    assert rel_err(grads[1], grads[0]) < 1e-4


@pytest.mark.parametrize("softmax", [True, False])
@pytest.mark.parametrize("case", [
    dict(size=(16, 16, 16), rotated=True, transfer=False),
    dict(size=(7, 10, 10), rotated=True, transfer=True),            # nvox % 4 != 0: scalar loads
    dict(size=(64, 64, 64), rotated=False, transfer=False),         # the reference's volume, evaluation (theta = 0)
])
def test_softargmax_with_spec_is_bit_identical(softmax, case):
    B, J = 2, 17
    cs, _ = _spec(B, case["size"], 4, case["rotated"], case["transfer"])
    coord = cs.materialize()
    g = torch.Generator().manual_seed(8)
    vol = torch.randn((B, J) + case["size"], generator=g).cuda()
    gc, gp = torch.randn(B, J, 3, generator=g).cuda(), (torch.randn((B, J) + case["size"], generator=g) * 1e-3).cuda()
    res = []
    for cv in (coord, cs):
        v = vol.clone().requires_grad_(True)
        coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(v, cv, softmax=softmax)
        (gv,) = torch.autograd.grad([coords, prob], v, [gc, gp])
        res.append((coords.detach(), prob.detach(), gv))
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a, b)


def test_spec_path_against_the_reference_construction():
    """End to end against the oracle: coordinates built like triangulation.py:306-333 on the CPU, ops of the oracle,
    vs the ops fed with nothing but the spec. Un-rotated (evaluation): the coordinates are exact, so the BASELINE
    tolerances apply unchanged."""
    from oracle import ref_ops
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=32, W=32, N=16, seed=0, train=False)
    ref = ref_ops.unproject_heatmaps(d["heatmaps"], d["proj_matricies"], d["coord_volumes"], "softmax")
    specs = volumetric.coord_volume_specs(d["base_points"], 2500.0, 16, d["thetas"]).cuda()
    cs = lt_b200.CoordSpec(specs, 16)
    assert torch.equal(cs.materialize().cpu(), d["coord_volumes"])
    out = lt_b200.unproject_heatmaps(d["heatmaps"].cuda(), d["proj_matricies"].cuda(), cs, "softmax")
    assert rel_err(out.cpu(), ref) < 1e-5
    s = synthetic.make_softargmax_inputs(B=1, J=17, N=16, seed=0)
    c_ref, p_ref = ref_ops.integrate_tensor_3d_with_coordinates(s["volumes"].double(), d["coord_volumes"].double())
    coords, prob = lt_b200.integrate_tensor_3d_with_coordinates(s["volumes"].cuda(), cs)
    assert (coords.double().cpu() - c_ref).abs().max().item() < 1e-3 and rel_err(prob.cpu(), p_ref) < 1e-5


def test_coordspec_argument_checks():
    cs, _ = _spec(2, (8, 8, 8), 1)
    assert tuple(cs.shape) == (2, 8, 8, 8, 3) and cs.dim() == 5 and tuple(cs[:1].shape) == (1, 8, 8, 8, 3)
    with pytest.raises(RuntimeError):
        lt_b200.CoordSpec(torch.zeros(2, 18), 8)                          # not on the device
    with pytest.raises(RuntimeError):
        lt_b200.CoordSpec(torch.zeros(2, 17).cuda(), 8)
    with pytest.raises(RuntimeError):
        lt_b200.CoordSpec(torch.zeros(2, 18).cuda(), (4, 6, 8), True)     # the axis swap needs N1 == N2
    with pytest.raises(RuntimeError):                                     # volume / spec shape mismatch
        lt_b200.integrate_tensor_3d_with_coordinates(torch.zeros(2, 3, 4, 8, 8).cuda(), cs)
