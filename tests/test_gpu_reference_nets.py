"""GPU: the reference's own nets, UNMODIFIED, running on the B200 through the drop-ins.

`oracle/fetch_ref.py` places a verbatim copy of the reference's `mvn` package under `oracle/_ref/` (git-ignored test
infrastructure that travels with the gpurun snapshot; in the build container `/root/reference` itself is used). The
tests build `VolumetricTriangulationNet` (mvn/models/triangulation.py:203-355) and `AlgebraicTriangulationNet`
(:132-200) with a ResNet-18 backbone and random weights on CUDA, run forward + backward on one seeded synthetic batch
  (a) stock   — every op the reference's own torch code,
  (b) patched — `lt_b200.patch_reference()` (+ `use_fused_process_features` for the volumetric net),
and compare key points and parameter gradients. Nothing in the reference's files is touched: the nets resolve
`op.unproject_heatmaps` (:349), `op.integrate_tensor_3d_with_coordinates` (:353), `op.integrate_tensor_2d` (:164),
`multiview.triangulate_batch_of_points` (:188) and `self.process_features` (:345) at call time.

Tolerances: key points 2e-2 mm; parameter gradients 1e-4 relative, max-norm, over the whole gradient and per tensor
(see _grad_distance for how cancelling sums are treated). The algebraic net is judged against the stock net run in fp64: the reference's
fp32 `torch.svd` is itself 0.3-0.5 mm from its fp64 result, which is reported next to ours."""
import copy
import os
import sys
from types import SimpleNamespace as NS

import numpy as np
import pytest
import torch

import lt_b200
from conftest import ROOT

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.join(ROOT, "oracle"))
from fetch_ref import ref_path  # noqa: E402

REF = ref_path()
needs_ref = pytest.mark.skipif(REF is None, reason="no copy of the reference: run `python oracle/fetch_ref.py` where "
                                                   "/root/reference exists (build() does)")


def _cfg(volume_size=32, agg="softmax"):
    return NS(model=NS(backbone=NS(num_joints=17, num_layers=18, style="simple", init_weights=False, checkpoint=None,
                                   alg_confidences=False, vol_confidences=False),
                       volume_aggregation_method=agg, volume_softmax=True, volume_multiplier=1.0,
                       volume_size=volume_size, cuboid_side=2500.0, kind="mpii", use_gt_pelvis=False,
                       heatmap_softmax=True, heatmap_multiplier=100.0, use_confidences=True))


def _batch(B, V, img, seed):
    """Cameras on a ring looking at the pelvis (mvn.utils.multiview.Camera objects, as the data loader hands them
    over), random images, pelvis positions."""
    from mvn.utils.multiview import Camera
    rng = np.random.RandomState(seed)
    pelvis = rng.normal(0.0, 100.0, size=(B, 17, 3))
    cams = []
    for v in range(V):
        row = []
        for b in range(B):
            az = 2.0 * np.pi * v / V + 0.3 * b
            c = pelvis[b, 6] + np.array([4500.0 * np.cos(az), 4500.0 * np.sin(az), 1500.0])
            fwd = pelvis[b, 6] - c
            fwd /= np.linalg.norm(fwd)
            right = np.cross(fwd, [0.0, 0.0, 1.0])
            right /= np.linalg.norm(right)
            R = np.stack([right, np.cross(fwd, right), fwd])
            f = 2.5 * img  # the whole cuboid in view
            K = np.array([[f, 0.0, img / 2.0], [0.0, f, img / 2.0], [0.0, 0.0, 1.0]])
            row.append(Camera(R, (-R @ c)[:, None], K))
        cams.append(row)
    images = torch.randn(B, V, 3, img, img, generator=torch.Generator().manual_seed(seed))
    return images, {"cameras": cams, "pred_keypoints_3d": pelvis}


def _run(net, images, batch, target, proj=None):
    """forward + backward of a LINEAR loss sum(w * key points), w fixed: d loss / d key points is the same in every arm,
    so the parameter gradients compare the backward paths at (almost) the same linearisation point. (An L1 loss flips
    signs on the 1e-2 mm the arms differ by; an L2 loss scales its gradient by (kp - target), i.e. turns the allowed
    2e-2 mm key-point difference into a 2e-4 relative gradient difference that says nothing about the ops.)"""
    net.zero_grad(set_to_none=True)
    out = net(images, proj, batch)
    kp = out[0]
    w = torch.randn(kp.shape, generator=torch.Generator().manual_seed(99), dtype=torch.float64).to(kp)
    loss = (kp * w).sum() + 0.0 * target.sum()
    loss.backward()
    grads = {n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None}
    return kp.detach().clone(), grads


def _grad_distance(ga, gb, floor=1e-2):
    """Two distances between the parameter gradients of two runs of the same net:
      * global: max |ga - gb| over ALL parameters / the largest gradient entry of the net (BASELINE's "relative on
        gradients", max-norm, taken over the whole gradient);
      * worst tensor: max-norm relative per parameter tensor, with the denominator floored at `floor` x the largest
        entry of the net. Tensors far below that are sums that cancel (the output bias of V2V in front of a softmax
        has an analytically ZERO gradient; measured 2.5e-7 of the scale in both arms): their fp32 value is rounding
        noise of the activations x upstream products, in the stock net as much as in the patched one."""
    assert ga.keys() == gb.keys() and len(ga) > 10
    scale = max(v.abs().max().item() for v in gb.values())
    worst, glob = 0.0, 0.0
    for n in gb:
        diff = (ga[n].double() - gb[n].double()).abs().max().item()
        glob = max(glob, diff / scale)
        worst = max(worst, diff / max(gb[n].abs().max().item(), floor * scale))
    return glob, worst


def _compare_grads(ga, gb, tol, what="", noise=(0.0, 0.0)):
    """`noise` = the same two distances between the STOCK net and the STOCK net fed with images perturbed by one fp32
    ulp (_one_ulp): how far fp32 rounding alone moves this net's gradients. The patched net must stay within
    max(tol, 4 x that) of the stock one: a drop-in may differ from ATen by rounding, not by more."""
    glob, worst = _grad_distance(ga, gb)
    print("parameter gradients %s: global %.3g, worst tensor %.3g (relative, max-norm); one-ulp input noise of the "
          "stock net: %.3g, %.3g" % (what, glob, worst, noise[0], noise[1]))
    assert glob < max(tol, 4.0 * noise[0]) and worst < max(tol, 4.0 * noise[1]), \
        "parameter gradients differ: global %.3g, worst tensor %.3g (noise %.3g, %.3g)" % (glob, worst, noise[0], noise[1])
    return worst


def _one_ulp(images, seed=123):
    """every pixel moved by one unit in the last place, sign at random"""
    sign = torch.randint(0, 2, images.shape, generator=torch.Generator().manual_seed(seed)).to(images) * 2.0 - 1.0
    return images * (1.0 + sign * 2.0 ** -23)


@pytest.fixture
def reference_on_path():
    sys.path.insert(0, REF)
    was_tf32 = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False       # both arms in true fp32: differences come from the swapped ops only
    torch.backends.cuda.matmul.allow_tf32 = False
    det = torch.backends.cudnn.deterministic
    torch.backends.cudnn.deterministic = True
    try:
        yield
    finally:
        lt_b200.unpatch_reference()
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = was_tf32
        torch.backends.cudnn.deterministic = det
        sys.path.remove(REF)


@needs_ref
@pytest.mark.parametrize("agg", ["softmax", "conf_norm"])
def test_volumetric_net_unmodified_stock_vs_patched(reference_on_path, agg):
    from mvn.models.triangulation import VolumetricTriangulationNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = VolumetricTriangulationNet(_cfg(32, agg), device=dev).to(dev).eval()
    images, batch = _batch(B=2, V=4, img=128, seed=1)
    images = images.to(dev)
    target = torch.from_numpy(batch["pred_keypoints_3d"]).float().to(dev)

    kp_ref, g_ref = _run(net, images, batch, target)
    kp_ulp, g_ulp = _run(net, _one_ulp(images), batch, target)  # conditioning of the stock net itself
    noise = _grad_distance(g_ulp, g_ref)
    print("stock net, images moved by one ulp: key points %.3g mm" % (kp_ulp - kp_ref).abs().max().item())

    lt_b200.patch_reference()
    kp_ops, g_ops = _run(net, images, batch, target)          # the three ops swapped, nn.Conv2d process_features kept
    fused = copy.deepcopy(net)
    lt_b200.use_fused_process_features(fused)                   # K7 writes the gather layout, K0 / K0^-1 leave the step
    kp_fused, g_fused = _run(fused, images, batch, target)
    lt_b200.unpatch_reference()

    for kp in (kp_ops, kp_fused):
        assert kp.shape == (2, 17, 3)
        assert (kp - kp_ref).abs().max().item() < 2e-2         # mm
    _compare_grads(g_ops, g_ref, 1e-4, 'ops patched vs stock', noise)
    _compare_grads(g_fused, g_ref, 1e-4, 'ops + fused process_features vs stock', noise)


@needs_ref
def test_volumetric_net_training_mode_random_rotation(reference_on_path):
    """train(): the reference draws a rotation per item with numpy (triangulation.py:318) and batch norm uses batch
    statistics; same numpy seed in both arms."""
    from mvn.models.triangulation import VolumetricTriangulationNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(2)
    net = VolumetricTriangulationNet(_cfg(32, "softmax"), device=dev).to(dev).train()
    images, batch = _batch(B=2, V=3, img=128, seed=3)
    images = images.to(dev)
    target = torch.from_numpy(batch["pred_keypoints_3d"]).float().to(dev)
    state = copy.deepcopy(net.state_dict())
    np.random.seed(7)
    kp_ref, g_ref = _run(net, images, batch, target)
    net.load_state_dict(state)                                  # batch-norm running statistics back to the start
    np.random.seed(7)
    _, g_ulp = _run(net, _one_ulp(images), batch, target)
    noise = _grad_distance(g_ulp, g_ref)
    net.load_state_dict(state)
    lt_b200.patch_reference()
    np.random.seed(7)
    kp_ops, g_ops = _run(net, images, batch, target)
    lt_b200.unpatch_reference()
    assert (kp_ops - kp_ref).abs().max().item() < 2e-2
    _compare_grads(g_ops, g_ref, 1e-4, 'train(), ops patched vs stock', noise)


@needs_ref
def test_algebraic_net_unmodified_stock_vs_patched(reference_on_path):
    from mvn.models.triangulation import AlgebraicTriangulationNet
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    net = AlgebraicTriangulationNet(_cfg(), device=dev).to(dev).eval()
    images, batch = _batch(B=2, V=4, img=128, seed=5)
    images = images.to(dev)
    # image-space projection matrices, as train.py:166 builds them from the same cameras
    proj = torch.stack([torch.stack([torch.from_numpy(c.projection) for c in row], 0) for row in batch["cameras"]], 0)
    proj = proj.transpose(1, 0).float().to(dev)
    target = torch.from_numpy(batch["pred_keypoints_3d"]).float().to(dev)

    kp32, g32 = _run(net, images, batch, target, proj)
    net64 = copy.deepcopy(net).double()
    kp64, g64 = _run(net64, images.double(), batch, target.double(), proj.double())

    lt_b200.patch_reference()
    kp_ops, g_ops = _run(net, images, batch, target, proj)
    lt_b200.unpatch_reference()

    err_ours = (kp_ops.double() - kp64.double()).abs().max().item()
    err_ref32 = (kp32.double() - kp64.double()).abs().max().item()
    print("algebraic net key points vs the fp64 run of the reference: ours %.3g mm, reference fp32 %.3g mm"
          % (err_ours, err_ref32))
    assert kp_ops.shape == (2, 17, 3)
    assert err_ours < max(2e-2, err_ref32)
    # gradients against the fp64 run; the fp32 reference is only as close as its SVD lets it be
    (glob_ours, worst_ours), (glob_ref, worst_ref) = _grad_distance(g_ops, g64), _grad_distance(g32, g64)
    print("algebraic net parameter gradients vs fp64 (global / worst tensor): ours %.3g / %.3g, reference fp32 %.3g / %.3g"
          % (glob_ours, worst_ours, glob_ref, worst_ref))
    # the fp32 backbone in front of the ops is shared by both arms: ours must be as close to fp64 as the stock fp32 run
    assert glob_ours < max(1e-4, 2.0 * glob_ref) and worst_ours < max(1e-4, 2.0 * worst_ref)
