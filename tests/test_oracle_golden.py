"""CPU: the oracle (oracle/ref_ops.py) against the golden vectors produced by the unmodified reference
(oracle/make_golden.py). This is what pins the oracle."""
import numpy as np
import pytest
import torch

from conftest import golden_names, load_golden, rel_err
from oracle import ref_ops


@pytest.mark.parametrize("name", golden_names("unproject_"))
def test_unproject_matches_reference(name):
    g = load_golden(name)
    hm = g["heatmaps"].clone().requires_grad_(True)
    vc = g.get("vol_confidences")
    if vc is not None:
        vc = vc.clone().requires_grad_(True)
    out = ref_ops.unproject_heatmaps(hm, g["proj_matricies"], g["coord_volumes"], g["method"], vc,
                                     align_corners=bool(g["align_corners"]))
    assert out.shape == g["out"].shape and out.dtype == torch.float32
    # same ATen calls in the same order: bit-exact on the same machine, 1e-6 leaves room for other CPUs
    assert rel_err(out, g["out"]) < 1e-6
    out.backward(g["grad_out"])
    assert rel_err(hm.grad, g["grad_heatmaps"]) < 1e-5
    if vc is not None:
        assert rel_err(vc.grad, g["grad_vol_confidences"]) < 1e-5


@pytest.mark.parametrize("name", golden_names("unproject_"))
def test_unproject_np64_referee(name):
    """The torch-free float64 restatement agrees with the reference to fp32 rounding: grid_sample's semantics
    (bilinear, zeros padding, align_corners switch, H/W swap of op.py:128-129) are understood, not just called."""
    g = load_golden(name)
    out = ref_ops.unproject_heatmaps_np64(g["heatmaps"].numpy(), g["proj_matricies"].numpy(),
                                          g["coord_volumes"].numpy(), g["method"],
                                          g["vol_confidences"].numpy() if "vol_confidences" in g else None,
                                          align_corners=bool(g["align_corners"]))
    assert rel_err(torch.from_numpy(out), g["out"]) < 2e-5


def coord_tol(ref):
    """1e-3 mm (BASELINE.json) for softmax coordinates; un-normalised relu sums reach 1e4..1e5 where one fp32 ulp is
    already > 1e-3, so the bar there is 4 ulp of the largest magnitude."""
    return max(1e-3, 4 * 1.2e-7 * ref.abs().max().item())


@pytest.mark.parametrize("name", golden_names("softargmax_"))
def test_softargmax_matches_reference(name):
    g = load_golden(name)
    vol = g["volumes"].clone().requires_grad_(True)
    coords, prob = ref_ops.integrate_tensor_3d_with_coordinates(vol, g["coord_volumes"], softmax=bool(g["softmax"]))
    assert (coords - g["coords"]).abs().max().item() < coord_tol(g["coords"])  # mm
    assert rel_err(prob, g["prob"]) < 1e-6
    (gv,) = torch.autograd.grad([coords, prob], vol, [g["grad_coords"], g["grad_prob"]])
    assert rel_err(gv, g["grad_volumes_both"]) < 1e-5


@pytest.mark.parametrize("name", golden_names("softargmax_"))
def test_softargmax_slab_pieces(name):
    """partial -> merge -> finalize over 1, 2 and 3 slabs reproduces the unsharded reference result."""
    g = load_golden(name)
    vol, coord, sm = g["volumes"], g["coord_volumes"], bool(g["softmax"])
    n0 = vol.shape[2]
    for R in (1, 2, 3):
        cuts = np.linspace(0, n0, R + 1).astype(int)
        parts = torch.stack([ref_ops.softargmax_partial(vol[:, :, a:b], coord[:, a:b], sm)
                             for a, b in zip(cuts[:-1], cuts[1:])])
        merged = ref_ops.softargmax_merge(parts, sm)
        probs = []
        for a, b in zip(cuts[:-1], cuts[1:]):
            coords, p = ref_ops.softargmax_finalize(vol[:, :, a:b], merged, sm)
            probs.append(p)
        assert (coords.double() - g["coords64"]).abs().max().item() < coord_tol(g["coords64"])
        assert rel_err(torch.cat(probs, dim=2), g["prob64"]) < 1e-5


@pytest.mark.parametrize("name", golden_names("softargmax2d_"))
def test_softargmax2d_matches_reference(name):
    g = load_golden(name)
    hm = g["heatmaps"].clone().requires_grad_(True)
    coords, prob = ref_ops.integrate_tensor_2d(hm, bool(g["softmax"]))
    assert torch.allclose(coords, g["coords"], rtol=1e-6, atol=1e-6) and torch.allclose(prob, g["prob"], rtol=1e-6, atol=1e-9)
    (gh,) = torch.autograd.grad([coords, prob], hm, [g["grad_coords"], g["grad_prob"]])
    assert rel_err(gh, g["grad_heatmaps_both"]) < 1e-6
    c64, p64 = ref_ops.integrate_tensor_2d(g["heatmaps"].double(), bool(g["softmax"]))
    assert torch.allclose(c64, g["coords64"], rtol=1e-12, atol=1e-12) and torch.allclose(p64, g["prob64"], rtol=1e-12, atol=1e-15)


@pytest.mark.parametrize("name", golden_names("softargmax3dindex_"))
def test_softargmax3d_index_matches_reference(name):
    g = load_golden(name)
    v = g["volumes"].clone().requires_grad_(True)
    coords, vol = ref_ops.integrate_tensor_3d(v, bool(g["softmax"]))
    assert torch.allclose(coords, g["coords"], rtol=1e-6, atol=1e-6) and torch.allclose(vol, g["prob"], rtol=1e-6, atol=1e-9)
    (gv,) = torch.autograd.grad(coords, v, g["grad_coords"])
    assert rel_err(gv, g["grad_volumes"]) < 1e-6


@pytest.mark.parametrize("name", golden_names("coordvol_"))
def test_coord_volume_matches_reference(name):
    g = load_golden(name)
    for b in range(g["base_points"].shape[0]):
        cv = ref_ops.build_coord_volume(g["base_points"][b].numpy(), g["side"], g["volume_size"], g["thetas"][b].item(),
                                        g["axis"].tolist(), bool(g["transfer"]))
        assert torch.equal(cv, g["coord_volumes"][b])          # same statements on the same BLAS: same bits
    # theta = 0 keeps the grid axis-aligned: item 0 spans exactly [base - side/2, base + side/2] per axis
    cv0 = g["coord_volumes"][0]
    lo, hi = cv0.flatten(0, 2).min(0).values.double(), cv0.flatten(0, 2).max(0).values.double()
    assert torch.allclose((lo + hi) / 2, g["base_points"][0], atol=1e-3) and torch.allclose(hi - lo, torch.full((3,), g["side"], dtype=torch.float64), atol=1e-3)


@pytest.mark.parametrize("name", golden_names("volce_"))
def test_volumetric_ce_matches_reference(name):
    g = load_golden(name)
    prob = g["prob"].clone().requires_grad_(True)
    loss = ref_ops.volumetric_ce_loss(g["coord_volumes"], prob, g["keypoints_gt"], g["validity"])
    assert loss.item() == g["loss"]                                   # same statements, same bits
    (gp,) = torch.autograd.grad(loss, prob)
    assert torch.equal(gp, g["grad_prob"])
    assert torch.equal(ref_ops.volumetric_ce_nearest_voxels(g["coord_volumes"], g["keypoints_gt"]), g["nearest"])
    assert int((g["grad_prob"] != 0).sum()) == int(g["validity"].sum())  # one non-zero per valid joint
    l64 = ref_ops.volumetric_ce_loss(g["coord_volumes"].double(), g["prob"].double(), g["keypoints_gt"].double(),
                                     g["validity"].double())
    # the golden fp64 loss was taken on fp64 softmax volumes; these are the fp32 ones widened (1e-7 relative apart)
    assert abs(l64.item() - g["loss64"]) < 1e-6


@pytest.mark.parametrize("name", golden_names("dlt_"))
def test_dlt_matches_reference(name):
    g = load_golden(name)
    conf = g.get("confidences")
    out32 = ref_ops.triangulate_batch_of_points(g["proj_matricies"], g["points"], conf)
    assert out32.dtype == torch.float32
    out64 = ref_ops.triangulate_batch_of_points(g["proj_matricies"].double(), g["points"].double(),
                                                conf.double() if conf is not None else None, out_dtype=None)
    # fp64 oracle vs fp64 reference: the 1e-2 mm bar of BASELINE.json (they agree to ~1e-7 mm)
    assert (out64 - g["out64"]).abs().max().item() < 1e-4
    # fp32 runs are only as good as fp32 svd of an ill-scaled system (~0.3 mm between precisions)
    assert (out32.double() - g["out64"]).abs().max().item() < 5.0
    assert (g["out32"].double() - g["out64"]).abs().max().item() < 5.0


def test_dlt_noise_free_recovers_points():
    from lt_b200 import synthetic
    d = synthetic.make_dlt_inputs(B=4, V=4, J=6, seed=5, noise_px=0.0)
    out = ref_ops.triangulate_batch_of_points(d["proj_matricies"].double(), d["points"].double(),
                                              d["confidences"].double(), out_dtype=None)
    assert (out - d["points_3d"].double()).abs().max().item() < 0.05  # fp32 rounding of the 2-D points only


@pytest.mark.parametrize("name", golden_names("pwconv_"))
def test_process_features_matches_reference(name):
    """oracle.process_features == the `process_features` module of a real reference VolumetricTriangulationNet
    (triangulation.py:238-240) applied as in forward() (:344-346), values and all three gradients."""
    g = load_golden(name)
    f5 = g["features"].clone().requires_grad_(True)
    w = g["weight"].clone().requires_grad_(True)
    b = g["bias"].clone().requires_grad_(True)
    y = ref_ops.process_features(f5.view(-1, *f5.shape[2:]), w, b)
    y5 = y.view(f5.shape[0], f5.shape[1], *y.shape[1:])
    assert y5.shape == g["out"].shape and rel_err(y5, g["out"]) < 1e-6
    y5.backward(g["grad_out"])
    assert rel_err(f5.grad, g["grad_features"]) < 1e-6
    assert rel_err(w.grad, g["grad_weight"]) < 1e-5
    assert rel_err(b.grad, g["grad_bias"]) < 1e-5
