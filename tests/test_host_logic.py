"""CPU: host-side behaviour of the drop-in functions that does not need a device: signatures, error conventions
(SURVEY.md §8b), synthetic generators, slab bounds, reference patching."""
import inspect
import os
import sys

import pytest
import torch

import lt_b200
from lt_b200 import slab, synthetic


def test_signatures_match_reference():
    sig = inspect.signature(lt_b200.op.unproject_heatmaps)
    names = list(sig.parameters)
    assert names[:5] == ["heatmaps", "proj_matricies", "coord_volumes", "volume_aggregation_method", "vol_confidences"]
    assert sig.parameters["volume_aggregation_method"].default == "sum"
    assert sig.parameters["vol_confidences"].default is None
    assert all(sig.parameters[n].kind is inspect.Parameter.KEYWORD_ONLY for n in names[5:])
    sig = inspect.signature(lt_b200.op.integrate_tensor_3d_with_coordinates)
    assert list(sig.parameters)[:3] == ["volumes", "coord_volumes", "softmax"] and sig.parameters["softmax"].default is True
    sig = inspect.signature(lt_b200.op.integrate_tensor_2d)
    assert list(sig.parameters)[:2] == ["heatmaps", "softmax"] and sig.parameters["softmax"].default is True
    assert sig.parameters["multiplier"].kind is inspect.Parameter.KEYWORD_ONLY
    sig = inspect.signature(lt_b200.multiview.triangulate_batch_of_points)
    assert list(sig.parameters) == ["proj_matricies_batch", "points_batch", "confidences_batch"]
    sig = inspect.signature(lt_b200.loss.VolumetricCELoss.forward)
    assert list(sig.parameters) == ["self", "coord_volumes_batch", "volumes_batch_pred", "keypoints_gt",
                                    "keypoints_binary_validity"]
    sig = inspect.signature(lt_b200.multiview.triangulate_point_from_multiple_views_linear_torch)
    assert list(sig.parameters) == ["proj_matricies", "points", "confidences"]


def test_error_conventions():
    d = synthetic.make_unproject_inputs(B=1, V=2, C=4, H=8, W=8, N=4)
    with pytest.raises(ValueError, match="Unknown volume_aggregation_method: mean"):
        lt_b200.unproject_heatmaps(d["heatmaps"], d["proj_matricies"], d["coord_volumes"], "mean")
    with pytest.raises(TypeError):  # reference: vol_confidences[batch_i] on None (op.py:148)
        lt_b200.unproject_heatmaps(d["heatmaps"], d["proj_matricies"], d["coord_volumes"], "conf_norm")
    # no CPU fallback: CPU tensors are refused, loudly
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lt_b200.unproject_heatmaps(d["heatmaps"], d["proj_matricies"], d["coord_volumes"], "sum")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lt_b200.integrate_tensor_3d_with_coordinates(torch.zeros(1, 2, 4, 4, 4), d["coord_volumes"])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lt_b200.integrate_tensor_2d(torch.zeros(1, 2, 4, 4))
    dl = synthetic.make_dlt_inputs(B=1, V=3, J=2)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        lt_b200.triangulate_batch_of_points(dl["proj_matricies"], dl["points"], dl["confidences"])
    with pytest.raises(AssertionError):  # multiview.py:143
        lt_b200.triangulate_point_from_multiple_views_linear_torch(dl["proj_matricies"][0], dl["points"][0, :2, 0])


def test_synthetic_generators_are_seeded_and_shaped():
    a = synthetic.make_unproject_inputs(B=2, V=4, C=8, H=24, W=24, N=8, seed=3, train=True, conf="conf_norm")
    b = synthetic.make_unproject_inputs(B=2, V=4, C=8, H=24, W=24, N=8, seed=3, train=True, conf="conf_norm")
    for k in ("heatmaps", "proj_matricies", "coord_volumes", "vol_confidences"):
        assert torch.equal(a[k], b[k])
    assert a["coord_volumes"].shape == (2, 8, 8, 8, 3) and a["proj_matricies"].shape == (2, 4, 3, 4)
    assert torch.allclose(a["vol_confidences"].sum(1), torch.ones(2, 8), atol=1e-6)
    # the cuboid is 2500 mm wide and centred on the base point
    cv = a["coord_volumes"][0].reshape(-1, 3)
    assert abs((cv.max(0).values - cv.min(0).values)[2].item() - 2500.0) < 1e-2
    frac = synthetic.inside_fraction(a["proj_matricies"], a["coord_volumes"], 24, 24)
    assert 0.4 < frac < 0.8
    inside = synthetic.make_unproject_inputs(B=1, V=2, C=1, H=8, W=8, N=8, seed=1, inside_view=True)
    pts = torch.cat([inside["coord_volumes"][0].reshape(-1, 3), torch.ones(512, 1)], 1)
    depth = pts @ inside["proj_matricies"][0, 0, 2]
    assert (depth <= 0).any() and (depth > 0).any()  # exercises the z <= 0 branch (op.py:121)


def test_slab_bounds_partition():
    for n0, world in ((256, 8), (64, 4), (7, 3), (5, 8)):
        spans = [slab.slab_bounds(n0, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n0
        assert all(a[1] == b[0] for a, b in zip(spans[:-1], spans[1:]))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


@pytest.mark.skipif(not os.path.isdir("/root/reference/mvn"), reason="the reference tree only exists in the build container")
def test_patch_reference_swaps_the_three_call_sites():
    sys.path.insert(0, "/root/reference")
    try:
        import mvn.utils.multiview as ref_mv
        import mvn.utils.op as ref_op
        orig = ref_op.unproject_heatmaps
        done = lt_b200.patch_reference()
        assert "mvn.utils.op.unproject_heatmaps" in done
        assert ref_op.unproject_heatmaps is lt_b200.op.unproject_heatmaps
        assert ref_op.integrate_tensor_3d_with_coordinates is lt_b200.op.integrate_tensor_3d_with_coordinates
        assert ref_mv.triangulate_batch_of_points is lt_b200.multiview.triangulate_batch_of_points
        assert ref_op.integrate_tensor_2d is lt_b200.op.integrate_tensor_2d
        assert ref_op.integrate_tensor_3d is lt_b200.op.integrate_tensor_3d
        import mvn.models.loss as ref_loss
        ref_ce = inspect.signature(lt_b200.VolumetricCELoss.forward)
        assert ref_loss.VolumetricCELoss is lt_b200.loss.VolumetricCELoss
        lt_b200.unpatch_reference()
        assert ref_op.unproject_heatmaps is orig
        # the drop-in class keeps the reference's forward signature (loss.py:56)
        assert list(ref_ce.parameters) == list(inspect.signature(ref_loss.VolumetricCELoss.forward).parameters)
        assert ref_loss.VolumetricCELoss is not lt_b200.loss.VolumetricCELoss
    finally:
        sys.path.remove("/root/reference")


def test_coord_volume_specs_round_like_the_reference():
    """position/step/centre/rotation are rounded to fp32 where the reference's torch ops round their python scalars"""
    import numpy as np
    from lt_b200 import volumetric
    base = np.array([[10.123456789, -20.5, 1000.000123]])
    spec = volumetric.coord_volume_specs(base, 2500.0, 64, thetas=[0.3], axis=(0, 0, 1))
    assert spec.shape == (1, 18) and spec.dtype == torch.float32
    assert np.array_equal(spec[0, 0:3].numpy(), (base[0] - 1250.0).astype(np.float32))
    assert np.array_equal(spec[0, 3:6].numpy(), np.full(3, 2500.0 / 63).astype(np.float32))
    assert np.array_equal(spec[0, 6:9].numpy(), base[0].astype(np.float32))
    R = spec[0, 9:].view(3, 3).double().numpy()
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-6) and np.allclose(R[2], [0, 0, 1], atol=1e-7)
    assert np.isclose(R[0, 0], np.cos(0.3), atol=1e-7)
    with pytest.raises(RuntimeError):
        volumetric.build_coord_volumes(base, 2500.0, 8, device="cpu")


def test_use_fused_process_features_keeps_parameters_and_state_dict():
    """Host side of K7's drop-in (no GPU needed): the module of an existing net is switched in place, parameters, their
    identity and the state_dict keys of the reference (`process_features.0.weight/.bias`) are unchanged; anything that is
    not `nn.Sequential(nn.Conv2d(cin, cout, 1))` is refused; CPU tensors fail loudly (no fallback)."""
    import torch
    from torch import nn

    import lt_b200

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.process_features = nn.Sequential(nn.Conv2d(256, 32, 1))   # mvn/models/triangulation.py:238-240

    net = Net()
    before = {k: v.clone() for k, v in net.state_dict().items()}
    wid = id(net.process_features[0].weight)
    assert lt_b200.use_fused_process_features(net) is net
    assert isinstance(net.process_features, lt_b200.ProcessFeatures)
    assert list(net.state_dict()) == ["process_features.0.weight", "process_features.0.bias"]
    assert id(net.process_features[0].weight) == wid
    for k, v in net.state_dict().items():
        assert torch.equal(v, before[k])
    fresh = Net()
    fresh.load_state_dict(net.state_dict())                              # checkpoints travel both ways
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        net.process_features(torch.zeros(1, 256, 4, 4))

    class Other(nn.Module):
        def __init__(self):
            super().__init__()
            self.process_features = nn.Sequential(nn.Conv2d(256, 32, 3, padding=1))

    with pytest.raises(TypeError):
        lt_b200.use_fused_process_features(Other())
    with pytest.raises(RuntimeError):
        lt_b200.process_features(torch.zeros(1, 8, 4, 4), torch.zeros(4, 8, 3, 3))


@pytest.mark.skipif(not os.path.isdir("/root/reference/mvn"), reason="the reference tree only exists in the build container")
def test_reference_net_forward_reaches_the_drop_ins_at_its_own_call_sites():
    """Wiring evidence for SURVEY 8(f) row 4 with the UNMODIFIED reference `VolumetricTriangulationNet.forward`: built here on
    the CPU (ResNet-18 backbone, random weights, a two-view 64 x 64 batch with real `multiview.Camera` objects), patched,
    and run. There is no GPU in this container and no CPU fallback in the product, so the forward must die with the
    product's own error, raised from the reference's own call site: `self.process_features(features)` (triangulation.py:345)
    once `use_fused_process_features` has switched the module, `op.unproject_heatmaps(...)` (:349) with the ops patched only.
    (On the GPU box the reference does not exist; the numerics of the same chain are covered by tests/test_gpu_pipeline.py and
    tests/test_gpu_process_features.py::test_chain_into_unproject.)"""
    import traceback
    from types import SimpleNamespace as NS

    import numpy as np

    sys.path.insert(0, "/root/reference")
    try:
        from mvn.models.triangulation import VolumetricTriangulationNet
        from mvn.utils.multiview import Camera
        cfg = NS(model=NS(backbone=NS(num_joints=17, num_layers=18, style="simple", init_weights=False, checkpoint=None,
                                      alg_confidences=False, vol_confidences=False),
                          volume_aggregation_method="softmax", volume_softmax=True, volume_multiplier=1.0, volume_size=8,
                          cuboid_side=2500.0, kind="mpii", use_gt_pelvis=False, heatmap_softmax=True,
                          heatmap_multiplier=100.0))
        torch.manual_seed(0)
        net = VolumetricTriangulationNet(cfg, device="cpu").eval()
        K = np.array([[1145.0, 0, 32.0], [0, 1145.0, 32.0], [0, 0, 1.0]])
        cams = [[Camera(np.eye(3), np.array([[0.0], [0.0], [5000.0 + 100 * v]]), K)] for v in range(2)]
        batch = {"cameras": cams, "pred_keypoints_3d": np.zeros((1, 17, 3))}
        images = torch.randn(1, 2, 3, 64, 64)

        def failing_site():
            with pytest.raises(RuntimeError, match="no CPU fallback") as ei:
                with torch.no_grad():
                    net(images, None, batch)
            frames = traceback.extract_tb(ei.tb)
            ref = [f for f in frames if f.filename.endswith("mvn/models/triangulation.py")]
            ours = [f for f in frames if "learnable-triangulation-pytorch_b200" in f.filename]
            assert ref and ours
            return ref[-1].line, os.path.basename(ours[0].filename)

        lt_b200.patch_reference()
        try:
            line, mod = failing_site()
            assert "op.unproject_heatmaps(" in line and mod == "op.py"
            lt_b200.use_fused_process_features(net)
            line, mod = failing_site()
            assert "self.process_features(features)" in line and mod == "triangulation.py"
            # the algebraic net: first drop-in on its path is the 2-D soft-argmax (triangulation.py:164)
            from mvn.models.triangulation import AlgebraicTriangulationNet
            cfg.model.use_confidences = False   # the confidence head pools 64 x 64 inputs away
            net = AlgebraicTriangulationNet(cfg, device="cpu").eval()
            line, mod = failing_site()
            assert "op.integrate_tensor_2d(" in line and mod == "op.py"
        finally:
            lt_b200.unpatch_reference()
    finally:
        sys.path.remove("/root/reference")


def test_nvtx_ranges_are_opt_in(monkeypatch):
    """LT_B200_NVTX=1 wraps the public ops in `lt_b200.<op>` ranges; without it the decorator returns the function itself
    (no wrapper frame on the hot path)."""
    import importlib
    from lt_b200 import _nvtx
    calls = []

    def f(x):
        return x + 1
    monkeypatch.setattr(_nvtx, "ENABLED", False)
    assert _nvtx.annotate("lt_b200.f")(f) is f
    monkeypatch.setattr(_nvtx, "ENABLED", True)
    monkeypatch.setattr(torch.cuda.nvtx, "range_push", lambda name: calls.append(("push", name)))
    monkeypatch.setattr(torch.cuda.nvtx, "range_pop", lambda: calls.append(("pop",)))
    g = _nvtx.annotate("lt_b200.f")(f)
    assert g is not f and g(1) == 2 and g.__name__ == "f"
    assert calls == [("push", "lt_b200.f"), ("pop",)]
    with pytest.raises(ZeroDivisionError):            # the range is closed when the op raises
        _nvtx.annotate("lt_b200.h")(lambda: 1 / 0)()
    assert calls[-2:] == [("push", "lt_b200.h"), ("pop",)]
    # the public ops carry the decorator (identity by default)
    import lt_b200
    assert lt_b200.unproject_heatmaps.__name__ == "unproject_heatmaps"
    importlib.reload(_nvtx)
