"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported in place from /root/reference) on small
seeded inputs. Run in the authoring container only (the reference does not travel to the GPU box):

    python oracle/make_golden.py

Every file stores the inputs next to the reference's outputs (and autograd gradients), so the fixtures do not depend
on the synthetic generators staying bit-stable. TEST INFRASTRUCTURE ONLY.
"""
import os
import sys
import warnings

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("LT_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

import lt_b200  # noqa: E402  (only for lt_b200.synthetic, host-side generators)
from mvn.utils import multiview as ref_mv  # noqa: E402
from mvn.utils import op as ref_op  # noqa: E402

syn = lt_b200.synthetic
OUT = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")


def npz(name, **arrays):
    arrays = {k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in arrays.items()
              if v is not None}
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **arrays)
    print("%-40s %8.1f KiB" % (name, os.path.getsize(path) / 1024))


def run_unproject(tag, method, align_corners, **kw):
    conf_kind = method if method.startswith("conf") else None
    d = syn.make_unproject_inputs(conf=conf_kind, **kw)
    hm = d["heatmaps"].clone().requires_grad_(True)
    vc = d["vol_confidences"].clone().requires_grad_(True) if d["vol_confidences"] is not None else None
    saved = ref_op.F.grid_sample
    if align_corners:  # what the reference computed under its pinned torch 1.0.1 (requirements.txt:20)
        ref_op.F.grid_sample = lambda inp, grid: saved(inp, grid, align_corners=True)
    try:
        out = ref_op.unproject_heatmaps(hm, d["proj_matricies"], d["coord_volumes"], method, vc)
        g = torch.Generator().manual_seed(kw.get("seed", 0) + 77)
        gout = torch.randn(out.shape, generator=g)
        out.backward(gout)
        # the same call with float64 inputs (per-view samples still rounded to the fp32 buffers of op.py:104,111)
        out64 = ref_op.unproject_heatmaps(d["heatmaps"].double(), d["proj_matricies"].double(),
                                          d["coord_volumes"].double(), method,
                                          d["vol_confidences"].double() if vc is not None else None)
    finally:
        ref_op.F.grid_sample = saved
    npz("unproject_%s" % tag, heatmaps=d["heatmaps"], proj_matricies=d["proj_matricies"],
        coord_volumes=d["coord_volumes"], vol_confidences=d["vol_confidences"], out=out, out64=out64, grad_out=gout,
        grad_heatmaps=hm.grad, grad_vol_confidences=(vc.grad if vc is not None else None),
        align_corners=np.array(int(align_corners)), method=np.array(method))


def run_softargmax(tag, softmax, **kw):
    d = syn.make_softargmax_inputs(**kw)
    for dt, suffix in ((torch.float32, ""), (torch.float64, "64")):
        vol = d["volumes"].to(dt).clone().requires_grad_(True)
        coords, prob = ref_op.integrate_tensor_3d_with_coordinates(vol, d["coord_volumes"].to(dt), softmax=softmax)
        g = torch.Generator().manual_seed(kw.get("seed", 0) + 78)
        gc = torch.randn(coords.shape, generator=g)
        gp = torch.randn(prob.shape, generator=g)
        (gv_c,) = torch.autograd.grad(coords, vol, gc.to(dt), retain_graph=True)          # through coordinates only
        (gv_b,) = torch.autograd.grad([coords, prob], vol, [gc.to(dt), gp.to(dt)])        # through both outputs
        if suffix == "":
            res = dict(volumes=d["volumes"], coord_volumes=d["coord_volumes"], coords=coords, prob=prob, grad_coords=gc,
                       grad_prob=gp, grad_volumes_coords_only=gv_c, grad_volumes_both=gv_b)
        else:
            res.update(coords64=coords, prob64=prob, grad_volumes_coords_only64=gv_c, grad_volumes_both64=gv_b)
    npz("softargmax_%s" % tag, softmax=np.array(int(softmax)), **res)


def run_softargmax3d_index(tag, softmax, B, J, N, seed):
    g = torch.Generator().manual_seed(seed)
    v0 = torch.randn(B, J, *N, generator=g)
    res = {}
    for dt, suffix in ((torch.float32, ""), (torch.float64, "64")):
        v = v0.to(dt).clone().requires_grad_(True)
        coords, vol = ref_op.integrate_tensor_3d(v, softmax)
        g2 = torch.Generator().manual_seed(seed + 82)
        gc = torch.randn(coords.shape, generator=g2)
        (gv,) = torch.autograd.grad(coords, v, gc.to(dt))
        if suffix == "":
            res.update(volumes=v0, coords=coords, prob=vol, grad_coords=gc, grad_volumes=gv)
        else:
            res.update(coords64=coords, prob64=vol, grad_volumes64=gv)
    npz("softargmax3dindex_%s" % tag, softmax=np.array(int(softmax)), **res)


def run_softargmax2d(tag, softmax, B, J, H, W, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    hm0 = torch.randn(B, J, H, W, generator=g) * scale
    res = {}
    for dt, suffix in ((torch.float32, ""), (torch.float64, "64")):
        hm = hm0.to(dt).clone().requires_grad_(True)
        coords, prob = ref_op.integrate_tensor_2d(hm, softmax)
        g2 = torch.Generator().manual_seed(seed + 80)
        gc = torch.randn(coords.shape, generator=g2)
        gp = torch.randn(prob.shape, generator=g2)
        (g_c,) = torch.autograd.grad(coords, hm, gc.to(dt), retain_graph=True)
        (g_b,) = torch.autograd.grad([coords, prob], hm, [gc.to(dt), gp.to(dt)])
        if suffix == "":
            res.update(heatmaps=hm0, coords=coords, prob=prob, grad_coords=gc, grad_prob=gp,
                       grad_heatmaps_coords_only=g_c, grad_heatmaps_both=g_b)
        else:
            res.update(coords64=coords, prob64=prob, grad_heatmaps_coords_only64=g_c, grad_heatmaps_both64=g_b)
    npz("softargmax2d_%s" % tag, softmax=np.array(int(softmax)), **res)


def run_coordvol(tag, n, seed, axis, transfer):
    """The reference's own pieces where they are callable (Cuboid3D, rotate_coord_volume from mvn/utils/volumetric.py)
    around the inline statements of triangulation.py:306-313,331-339, which only exist inside forward()."""
    from mvn.utils import volumetric as ref_vol
    rng = np.random.RandomState(seed)
    B = 3
    base = rng.normal(0.0, 300.0, size=(B, 3))
    thetas = np.concatenate([[0.0], rng.uniform(0.0, 2 * np.pi, size=B - 1)])
    side = 2500.0
    vols = []
    for b in range(B):
        sides = np.array([side, side, side])
        position = base[b] - sides / 2
        ref_vol.Cuboid3D(position, sides)
        xxx, yyy, zzz = torch.meshgrid(torch.arange(n), torch.arange(n), torch.arange(n))
        grid = torch.stack([xxx, yyy, zzz], dim=-1).type(torch.float).reshape((-1, 3))
        gc = torch.zeros_like(grid)
        gc[:, 0] = position[0] + (sides[0] / (n - 1)) * grid[:, 0]
        gc[:, 1] = position[1] + (sides[1] / (n - 1)) * grid[:, 1]
        gc[:, 2] = position[2] + (sides[2] / (n - 1)) * grid[:, 2]
        cv = gc.reshape(n, n, n, 3)
        center = torch.from_numpy(base[b]).type(torch.float)
        cv = cv - center
        cv = ref_vol.rotate_coord_volume(cv, thetas[b], axis)
        cv = cv + center
        if transfer:
            cv = cv.permute(0, 2, 1, 3)
            cv = cv.index_select(1, torch.arange(cv.shape[1] - 1, -1, -1).long())
        vols.append(cv)
    npz("coordvol_%s" % tag, base_points=base, thetas=thetas, axis=np.array(axis), side=np.array(side),
        volume_size=np.array(n), transfer=np.array(int(transfer)), coord_volumes=torch.stack(vols))


def run_volce(tag, B, J, N, seed, spread):
    """The unmodified VolumetricCELoss (mvn/models/loss.py:52-80) on the softmaxed volumes of the unmodified
    integrate_tensor_3d_with_coordinates; ground-truth joints = predicted joints + N(0, spread mm)."""
    from mvn.models.loss import VolumetricCELoss
    d = syn.make_softargmax_inputs(B=B, J=J, N=N, seed=seed, train=True, peaked=True)
    g = torch.Generator().manual_seed(seed + 81)
    res = {}
    for dt, suffix in ((torch.float32, ""), (torch.float64, "64")):
        vol = d["volumes"].to(dt).clone().requires_grad_(True)
        coords, prob = ref_op.integrate_tensor_3d_with_coordinates(vol, d["coord_volumes"].to(dt), softmax=True)
        if suffix == "":
            kp_gt = (coords.detach() + spread * torch.randn(coords.shape, generator=g)).float()
            validity = (torch.rand(B, J, 1, generator=g) > 0.25).float()
            validity[0, 0, 0] = 0.0
        prob_leaf = prob.detach().clone().requires_grad_(True)
        loss = VolumetricCELoss()(d["coord_volumes"].to(dt), prob_leaf, kp_gt.to(dt), validity.to(dt))
        (gp,) = torch.autograd.grad(loss, prob_leaf)
        if suffix == "":
            res.update(coord_volumes=d["coord_volumes"], prob=prob.detach(), keypoints_gt=kp_gt, validity=validity,
                       loss=loss.detach(), grad_prob=gp,
                       nearest=ref_nearest(d["coord_volumes"], kp_gt))
        else:
            res.update(loss64=loss.detach(), grad_prob64=gp)
    npz("volce_%s" % tag, **res)


def ref_nearest(coord_volumes, kp_gt):
    # the argmin statements of loss.py:63-71, kept as flat indices
    out = []
    for b in range(coord_volumes.shape[0]):
        dists = torch.sqrt(((coord_volumes[b].unsqueeze(0) - kp_gt[b].unsqueeze(1).unsqueeze(1).unsqueeze(1)) ** 2).sum(-1))
        out.append(torch.argmin(dists.view(dists.shape[0], -1), dim=-1))
    return torch.stack(out)


def run_dlt(tag, **kw):
    d = syn.make_dlt_inputs(**kw)
    P, pts, conf = d["proj_matricies"], d["points"], d["confidences"]
    pts32 = pts.clone().requires_grad_(True)
    conf32 = conf.clone().requires_grad_(True) if conf is not None else None
    out32 = ref_mv.triangulate_batch_of_points(P, pts32, conf32)
    g = torch.Generator().manual_seed(kw.get("seed", 0) + 79)
    gout = torch.randn(out32.shape, generator=g)
    out32.backward(gout)
    # float64 run of the reference's per-point solver (the batched wrapper rounds to fp32, multiview.py:164)
    B, V, J = pts.shape[:3]
    pts64 = pts.double().clone().requires_grad_(True)
    conf64 = conf.double().clone().requires_grad_(True) if conf is not None else None
    rows = []
    for b in range(B):
        for j in range(J):
            c = conf64[b, :, j] if conf64 is not None else torch.ones(V, dtype=torch.float64)
            rows.append(ref_mv.triangulate_point_from_multiple_views_linear_torch(P[b].double(), pts64[b, :, j], c))
    out64 = torch.stack(rows).view(B, J, 3)
    out64.backward(gout.double())
    npz("dlt_%s" % tag, proj_matricies=P, points=pts, confidences=conf, points_3d=d["points_3d"], out32=out32,
        out64=out64, grad_out=gout, grad_points32=pts32.grad, grad_points64=pts64.grad,
        grad_confidences32=(conf32.grad if conf32 is not None else None),
        grad_confidences64=(conf64.grad if conf64 is not None else None))


def run_process_features(tag, BV, H, W, seed, views=None):
    """`process_features` of an actual reference VolumetricTriangulationNet (triangulation.py:238-240), applied the way
    forward() applies it (:344-346), with the gradients autograd gives for a seeded upstream gradient."""
    from types import SimpleNamespace as NS
    from mvn.models.triangulation import VolumetricTriangulationNet
    cfg = NS(model=NS(backbone=NS(num_joints=17, num_layers=18, style="simple", init_weights=False, checkpoint=None,
                                  alg_confidences=False, vol_confidences=False),
                      volume_aggregation_method="softmax", volume_softmax=True, volume_multiplier=1.0, volume_size=64,
                      cuboid_side=2500.0, kind="mpii", use_gt_pelvis=False, heatmap_softmax=True,
                      heatmap_multiplier=100.0))
    torch.manual_seed(seed)
    net = VolumetricTriangulationNet(cfg, device="cpu")
    conv = net.process_features[0]
    gen = torch.Generator().manual_seed(seed + 1)
    views = views or 1
    feats = torch.randn(BV // views, views, conv.in_channels, H, W, generator=gen).requires_grad_(True)
    x = feats.view(-1, *feats.shape[2:])                               # triangulation.py:344
    y = net.process_features(x)                                        # :345
    y5 = y.view(BV // views, views, *y.shape[1:])                      # :346
    gy = torch.randn(y5.shape, generator=gen)
    y5.backward(gy)
    npz("pwconv_" + tag, features=feats.detach(), weight=conv.weight.detach(), bias=conv.bias.detach(), out=y5.detach(),
        grad_out=gy, grad_features=feats.grad, grad_weight=conv.weight.grad, grad_bias=conv.bias.grad)


def main():
    os.makedirs(OUT, exist_ok=True)
    if os.environ.get("LT_GOLDEN_ONLY") == "pwconv":
        run_process_features("small", BV=4, H=6, W=10, seed=81, views=2)
        run_process_features("ragged", BV=3, H=5, W=7, seed=82, views=3)      # H*W % 4 != 0: the 4-byte copy path
        return
    torch.set_num_threads(4)
    small = dict(B=2, V=3, C=6, H=10, W=12, N=(5, 6, 7), seed=11, train=True, inside_view=True, side=2500.0)
    for method in ("sum", "conf", "conf_norm", "max", "softmax"):
        run_unproject("small_%s" % method, method, False, **small)
    run_unproject("small_softmax_ac", "softmax", True, **small)
    run_unproject("small_sum_ac", "sum", True, **small)
    # 32-channel features on an 8^3 grid: the fast kernel instantiation (C=32 -> 4 lanes per voxel, aligned runs)
    c32 = dict(B=1, V=4, C=32, H=16, W=16, N=8, seed=12, train=True, side=2500.0)
    for method in ("sum", "conf", "max", "softmax"):
        run_unproject("c32_%s" % method, method, False, **c32)
    # more views than one chunk (kViewChunk = 4): exercises the many-view softmax/max paths
    many = dict(B=1, V=7, C=8, H=16, W=16, N=(4, 8, 8), seed=13, train=True, dome=True, side=2500.0)
    for method in ("sum", "conf", "max", "softmax"):
        run_unproject("many_%s" % method, method, False, **many)
    run_softargmax("small_softmax", True, B=2, J=3, N=(5, 6, 7), seed=21, train=True)
    run_softargmax("small_relu", False, B=2, J=3, N=(5, 6, 7), seed=22, train=True)
    run_softargmax("peaked_softmax", True, B=1, J=3, N=12, seed=23, peaked=True, train=True)
    run_softargmax3d_index("small_softmax", True, B=2, J=3, N=(5, 6, 8), seed=71)
    run_softargmax3d_index("small_relu", False, B=2, J=3, N=(4, 7, 5), seed=72)
    run_softargmax2d("small_softmax", True, B=2, J=3, H=7, W=9, seed=41)
    run_softargmax2d("small_relu", False, B=2, J=3, H=7, W=9, seed=42)
    run_softargmax2d("sharp_softmax", True, B=1, J=4, H=24, W=24, seed=43, scale=100.0)   # heatmap_multiplier = 100
    run_coordvol("mpii", 8, 61, [0, 0, 1], False)
    run_coordvol("coco", 6, 62, [0, 1, 0], False)
    run_coordvol("cmu_transfer", 8, 63, [0, 0, 1], True)
    run_volce("small", B=2, J=3, N=(5, 6, 7), seed=51, spread=150.0)
    run_volce("cube12", B=2, J=4, N=12, seed=52, spread=60.0)
    run_dlt("conf", B=3, V=4, J=5, seed=31)
    run_dlt("noconf", B=2, V=3, J=4, seed=32, with_conf=False)
    run_dlt("manyview", B=1, V=31, J=6, seed=33, dome=True)
    run_dlt("twoview", B=2, V=2, J=3, seed=34)
    run_process_features("small", BV=4, H=6, W=10, seed=81, views=2)
    run_process_features("ragged", BV=3, H=5, W=7, seed=82, views=3)      # H*W % 4 != 0: the 4-byte copy path


if __name__ == "__main__":
    main()
