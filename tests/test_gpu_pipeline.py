"""GPU: the geometric part of VolumetricTriangulationNet.forward (mvn/models/triangulation.py:280-355) and the training
loss of train.py:218-231 as one autograd graph — coordinate volumes (K6) -> unproject (K0+K1) -> a 1x1x1 conv standing
in for the V2V net (stock torch on both sides) -> soft-argmax (K2) -> L1 keypoint loss + volumetric CE (K5) -> backward
into the feature maps and the conv weights — against the same chain built from the oracle's restatements on the CPU."""
import numpy as np
import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("agg,fused_ce", [("softmax", True), ("softmax", False), ("conf", False), ("sum", True)])
def test_volumetric_forward_and_loss_chain(agg, fused_ce):
    B, V, C, J, N = 2, 4, 32, 17, 16
    d = synthetic.make_unproject_inputs(B=B, V=V, C=C, H=24, W=24, N=N, seed=3, train=True,
                                        conf="conf" if agg == "conf" else None)
    gen = torch.Generator().manual_seed(11)
    w0 = torch.randn(J, C, 1, 1, 1, generator=gen) * 0.3
    kp_gt = torch.from_numpy(d["base_points"]).float().view(B, 1, 3) + 150.0 * torch.randn(B, J, 3, generator=gen)
    val = (torch.rand(B, J, 1, generator=gen) > 0.2).float()
    mult, ce_w = 1.5, 0.01

    def chain(dev, ours):
        feat = d["heatmaps"].to(dev).requires_grad_(True)
        w = w0.to(dev).requires_grad_(True)
        proj = d["proj_matricies"].to(dev)
        vc = d["vol_confidences"].to(dev) if agg == "conf" else None
        if ours:
            coord = lt_b200.build_coord_volumes(d["base_points"], 2500.0, N, d["thetas"])
            vol = lt_b200.unproject_heatmaps(feat, proj, coord, agg, vc)
        else:
            coord = torch.stack([ref_ops.build_coord_volume(d["base_points"][b], 2500.0, N, d["thetas"][b], [0, 0, 1])
                                 for b in range(B)])
            vol = ref_ops.unproject_heatmaps(feat, proj, coord, agg, vc)
        x = torch.nn.functional.conv3d(vol, w)                                   # V2V stand-in, stock torch
        g, t = kp_gt.to(dev), val.to(dev)
        if ours and fused_ce:
            kp, prob, ce = lt_b200.integrate_tensor_3d_with_coordinates_and_ce(x, coord, g, t, multiplier=mult)
        elif ours:
            kp, prob = lt_b200.integrate_tensor_3d_with_coordinates(x, coord, multiplier=mult)
            ce = lt_b200.VolumetricCELoss()(coord, prob, g, t)
        else:
            kp, prob = ref_ops.integrate_tensor_3d_with_coordinates(x * mult, coord, softmax=True)
            ce = ref_ops.volumetric_ce_loss(coord, prob, g, t)
        l1 = torch.sum(torch.abs(g - kp) * t) / max(1, torch.sum(t).item())     # KeypointsMAELoss, loss.py:31-37
        total = l1 + ce_w * ce
        total.backward()
        return kp.detach().cpu(), total.item(), feat.grad.cpu(), w.grad.cpu(), vol.detach().cpu()

    saved = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False       # the stand-in (forward AND backward) must not add TF32 noise
    try:
        ours = chain("cuda", True)
    finally:
        torch.backends.cudnn.allow_tf32 = saved
    ref = chain("cpu", False)
    assert rel_err(ours[4], ref[4]) < 1e-5                                       # unprojected volume
    assert (ours[0] - ref[0]).abs().max().item() < 2e-2                          # keypoints, mm (fp32 chain on both sides)
    assert abs(ours[1] - ref[1]) <= 1e-4 * abs(ref[1])
    assert rel_err(ours[2], ref[2]) < 2e-4 and rel_err(ours[3], ref[3]) < 2e-4   # gradients through the whole chain
