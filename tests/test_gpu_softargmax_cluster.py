"""GPU: K2 forward with one read per volume (csrc/softargmax_cluster.cu, opt-in with LT_B200_K2FWD=cluster): a
thread-block cluster keeps one (b, joint) volume in shared memory between the statistics and the normalisation pass.
Against the fp64 oracle (mvn/utils/op.py:84-96) and against the default three-launch path: coordinates 1e-3 mm,
probabilities 1e-5 relative (the bars of tests/test_gpu_softargmax.py). Cluster sizes 1..16 (the smallest power of two
that leaves <= 64 KB per CTA), both modes (softmax / ReLU), tensor and in-kernel coordinates; shapes it does not cover
(nvox % (4 x cluster size) != 0, nvox > 64^3) must fall through to the default path unchanged."""
import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic, volumetric
from oracle import ref_ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("softmax", [True, False])
@pytest.mark.parametrize("N", [(8, 8, 8), (16, 16, 32), (32, 32, 32), (40, 40, 40), (64, 64, 64), (64, 64, 32)])
def test_cluster_forward_against_oracle_and_default(N, softmax, monkeypatch):
    B, J = 2, 5
    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=31, peaked=True, train=True)
    vol, coord = d["volumes"].cuda(), d["coord_volumes"].cuda()
    c64, p64 = ref_ops.integrate_tensor_3d_with_coordinates(d["volumes"].double(), d["coord_volumes"].double(),
                                                            softmax=softmax)
    monkeypatch.setenv("LT_B200_K2FWD", "cluster")
    c1, p1 = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord, softmax=softmax)
    monkeypatch.delenv("LT_B200_K2FWD")
    c0, p0 = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord, softmax=softmax)
    tol_c = max(1e-3, 5e-7 * c64.abs().max().item())   # ReLU mode: un-normalised sums, relative fp32 accuracy
    assert (c1.double().cpu() - c64).abs().max().item() < tol_c
    assert rel_err(p1.cpu(), p64) < 1e-5
    assert (c1 - c0).abs().max().item() < tol_c and rel_err(p1, p0) < 1e-5


def test_cluster_forward_with_in_kernel_coordinates_and_gradients(monkeypatch):
    B, J, N = 2, 17, (32, 32, 32)
    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=32, peaked=True, train=True)
    rng_base = d["coord_volumes"].reshape(B, -1, 3).mean(1).numpy()        # cuboid centres
    specs = volumetric.coord_volume_specs(rng_base, 2500.0, N, None).cuda()
    spec = lt_b200.CoordSpec(specs, N)
    vol = d["volumes"].cuda().requires_grad_(True)
    gc = torch.randn(B, J, 3, generator=torch.Generator().manual_seed(1)).cuda()
    monkeypatch.setenv("LT_B200_K2FWD", "cluster")
    c1, p1 = lt_b200.integrate_tensor_3d_with_coordinates(vol, spec, softmax=True)
    (g1,) = torch.autograd.grad(c1, vol, gc)
    monkeypatch.delenv("LT_B200_K2FWD")
    c0, p0 = lt_b200.integrate_tensor_3d_with_coordinates(vol, spec, softmax=True)
    (g0,) = torch.autograd.grad(c0, vol, gc)
    assert (c1 - c0).abs().max().item() < 1e-3 and rel_err(p1, p0) < 1e-5 and rel_err(g1, g0) < 1e-4


def test_shapes_outside_its_domain_use_the_default_path(monkeypatch):
    for N in [(5, 6, 7), (9, 9, 9), (72, 64, 64)]:           # nvox % 4 != 0; odd; larger than 16 x 64 KB
        d = synthetic.make_softargmax_inputs(B=1, J=3, N=N, seed=33, peaked=True, train=True)
        vol, coord = d["volumes"].cuda(), d["coord_volumes"].cuda()
        monkeypatch.setenv("LT_B200_K2FWD", "cluster")
        c1, p1 = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord, softmax=True)
        monkeypatch.delenv("LT_B200_K2FWD")
        c0, p0 = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord, softmax=True)
        assert torch.equal(c1, c0) and torch.equal(p1, p0)
