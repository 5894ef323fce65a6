"""GPU: K6 (coordinate volumes, mvn/models/triangulation.py:306-341) through the C ABI against golden vectors made with
the reference's own statements, the oracle at the BASELINE size, and geometric properties. The reference rotates with
an sgemm whose K=3 accumulation order is the BLAS's business: one fp32 ulp at |x| < 4096 mm (4.9e-4) is the bar for
rotated items, bit equality for the un-rotated item of every batch (theta = 0: the evaluation path)."""
import numpy as np
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden
from oracle import ref_ops

pytestmark = pytest.mark.gpu
ULP = 4.9e-4


@pytest.mark.parametrize("name", golden_names("coordvol_"))
def test_golden(name):
    g = load_golden(name)
    out = lt_b200.build_coord_volumes(g["base_points"].numpy(), g["side"], g["volume_size"], g["thetas"].numpy(),
                                      g["axis"].tolist(), bool(g["transfer"]))
    assert out.shape == g["coord_volumes"].shape and out.dtype == torch.float32
    assert (out.cpu() - g["coord_volumes"]).abs().max().item() <= ULP
    assert torch.equal(out[0].cpu(), g["coord_volumes"][0])      # theta = 0


@pytest.mark.parametrize("n,axis,transfer", [(64, (0, 0, 1), False), (64, (0, 1, 0), True), (7, (0, 0, 1), False)])
def test_against_oracle_at_baseline_size(n, axis, transfer):
    rng = np.random.RandomState(3)
    base = rng.normal(0.0, 200.0, size=(4, 3))
    thetas = np.concatenate([[0.0], rng.uniform(0, 2 * np.pi, size=3)])
    out = lt_b200.build_coord_volumes(base, 2500.0, n, thetas, axis, transfer).cpu()
    for b in range(4):
        ref = ref_ops.build_coord_volume(base[b], 2500.0, n, thetas[b], axis, transfer)
        assert (out[b] - ref).abs().max().item() <= ULP
    # rotation about the base point: every voxel keeps its distance to the rotation axis through the base point
    c = torch.from_numpy(base).float().view(4, 1, 1, 1, 3)
    a = torch.tensor(axis, dtype=torch.float32)
    rel = out - c
    along = (rel * a).sum(-1)
    assert torch.allclose(along[0], along[1], atol=2e-3)      # the component along the axis does not depend on theta
    radial = (rel - along.unsqueeze(-1) * a).norm(dim=-1)
    assert torch.allclose(radial[0].flatten().sort().values, radial[2].flatten().sort().values, atol=5e-3)


def test_feeds_the_unprojection_and_errors():
    from lt_b200 import synthetic
    d = synthetic.make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=16, seed=0, train=False)
    # synthetic.make_unproject_inputs builds its coordinates on the host with the same formulas (theta = 0)
    base = d["coord_volumes"][0].flatten(0, 2).double().mean(0).numpy()
    cv = lt_b200.build_coord_volumes(base[None], 2500.0, 16)
    assert (cv.cpu() - d["coord_volumes"]).abs().max().item() < 2e-3
    vol = lt_b200.unproject_heatmaps(d["heatmaps"].cuda(), d["proj_matricies"].cuda(), cv, "softmax")
    assert vol.shape == (1, 32, 16, 16, 16) and torch.isfinite(vol).all()
    assert lt_b200.build_coord_volumes(np.zeros((0, 3)), 2500.0, 8).shape == (0, 8, 8, 8, 3)
    with pytest.raises(RuntimeError):
        lt_b200.build_coord_volumes(base[None], 2500.0, 8, device="cpu")
    with pytest.raises(RuntimeError):
        lt_b200.build_coord_volumes(base[None], 2500.0, (4, 5, 6), transfer_cmu_to_human36m=True)
