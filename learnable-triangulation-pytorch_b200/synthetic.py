"""Seeded synthetic Human3.6M-shaped inputs for the hot path. Host-side only (numpy/torch CPU):
tests, bench.py and the golden-vector script all draw their inputs from here so that the CPU oracle and the GPU
kernels see identical bits.

Geometry follows what the reference feeds the ops:
  * cameras: pin-hole ``P = K [R|t]`` like ``Camera.projection`` (mvn/utils/multiview.py:37-43), K already rescaled to
    heat-map pixels (``update_after_resize``, multiview.py:24-35; called at triangulation.py:272-275);
  * coord_volumes: built exactly like mvn/models/triangulation.py:306-333 (cuboid of ``side`` mm around the base
    point, N points per axis, rotated by theta about the vertical axis through the base point).
"""
import math

import numpy as np
import torch


def _gen(seed):
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    return g


def look_at_camera(center, target, focal, cx, cy):
    """3x4 projection of a camera at ``center`` looking at ``target`` (world z is up, image y points down)."""
    center = np.asarray(center, dtype=np.float64)
    fwd = np.asarray(target, dtype=np.float64) - center
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd], axis=0)
    t = -R @ center
    K = np.array([[focal, 0.0, cx], [0.0, focal, cy], [0.0, 0.0, 1.0]])
    return K @ np.concatenate([R, t[:, None]], axis=1)


def make_cameras(B, V, base_points, seed, heatmap_hw=(96, 96), focal=240.0, dome=False, inside_view=False):
    """(B,V,3,4) float32 projection matrices in heat-map pixel units.

    Ring of V cameras, radius U(4000,5500) mm, 1500 mm above the base point's ground plane, looking at the base
    point; ``dome`` spreads them over three elevations (many-view, Panoptic-like). ``inside_view`` replaces view 0
    of every item by a camera INSIDE the cuboid so that part of the volume lies behind it (z <= 0 branch, op.py:121).
    """
    rng = np.random.RandomState(seed)
    H, W = heatmap_hw
    P = np.zeros((B, V, 3, 4), dtype=np.float64)
    for b in range(B):
        phase = rng.uniform(0.0, 2.0 * math.pi)
        for v in range(V):
            radius = rng.uniform(4000.0, 5500.0)
            az = phase + 2.0 * math.pi * v / V + rng.normal(0.0, 0.05)
            height = 1500.0 + rng.normal(0.0, 100.0)
            if dome:
                height = (500.0, 1500.0, 3000.0)[v % 3] + rng.normal(0.0, 100.0)
            c = base_points[b] + np.array([radius * math.cos(az), radius * math.sin(az), height])
            if inside_view and v == 0:
                c = base_points[b] + np.array([300.0, -200.0, 100.0])
            P[b, v] = look_at_camera(c, base_points[b], focal * W / 96.0, W / 2.0, H / 2.0)
    return torch.from_numpy(P).float()


def make_coord_volumes(B, N, base_points, thetas, side=2500.0):
    """(B,N0,N1,N2,3) float32 world coordinates, same arithmetic as triangulation.py:306-333 (kind='mpii': rotation
    about the z axis). N may be an int or a (N0,N1,N2) tuple; a tuple gives a cuboid with the same per-axis formula."""
    if isinstance(N, int):
        N = (N, N, N)
    out = torch.zeros(B, N[0], N[1], N[2], 3)
    for b in range(B):
        base = np.asarray(base_points[b], dtype=np.float64)
        sides = np.array([side, side, side])
        position = base - sides / 2
        xxx, yyy, zzz = torch.meshgrid(torch.arange(N[0]), torch.arange(N[1]), torch.arange(N[2]), indexing="ij")
        grid = torch.stack([xxx, yyy, zzz], dim=-1).type(torch.float).reshape((-1, 3))
        gc = torch.zeros_like(grid)
        for a in range(3):
            gc[:, a] = position[a] + (sides[a] / max(N[a] - 1, 1)) * grid[:, a]
        cv = gc.reshape(N[0], N[1], N[2], 3)
        center = torch.from_numpy(base).type(torch.float)
        th = float(thetas[b])
        # volumetric.get_rotation_matrix(axis=[0,0,1], theta) (volumetric.py:87-99), Euler-Rodrigues form
        a_ = math.cos(th / 2.0)
        d_ = -math.sin(th / 2.0)
        rot = np.array([[a_ * a_ - d_ * d_, 2 * a_ * d_, 0.0], [-2 * a_ * d_, a_ * a_ - d_ * d_, 0.0], [0.0, 0.0, a_ * a_ + d_ * d_]])
        rot = torch.from_numpy(rot).type(torch.float)
        cv = cv - center
        cv = rot.mm(cv.view(-1, 3).t()).t().view(N[0], N[1], N[2], 3)
        cv = cv + center
        out[b] = cv
    return out


def make_unproject_inputs(B=1, V=4, C=32, H=96, W=96, N=64, seed=0, train=False, conf=None, dome=False,
                          inside_view=False, side=2500.0):
    """Inputs of ``unproject_heatmaps``. conf: None | 'conf' | 'conf_norm'. train=True draws a rotation per item."""
    g = _gen(seed)
    rng = np.random.RandomState(seed + 1000)
    base = rng.normal(0.0, 100.0, size=(B, 3))
    thetas = rng.uniform(0.0, 2.0 * math.pi, size=B) if train else np.zeros(B)
    feats = torch.randn(B, V, C, H, W, generator=g)
    proj = make_cameras(B, V, base, seed + 2000, (H, W), dome=dome, inside_view=inside_view)
    coord = make_coord_volumes(B, N, base, thetas, side)
    vc = None
    if conf is not None:
        vc = torch.rand(B, V, C, generator=g)
        if conf == "conf_norm":
            vc = vc / vc.sum(dim=1, keepdim=True)
    return {"heatmaps": feats, "proj_matricies": proj, "coord_volumes": coord, "vol_confidences": vc,
            "base_points": base, "thetas": thetas}


def make_softargmax_inputs(B=1, J=17, N=64, seed=0, peaked=False, train=False, side=2500.0):
    """Inputs of ``integrate_tensor_3d_with_coordinates``: N(0,1) volumes, optionally +10 at one random voxel."""
    g = _gen(seed)
    rng = np.random.RandomState(seed + 1000)
    if isinstance(N, int):
        N = (N, N, N)
    base = rng.normal(0.0, 100.0, size=(B, 3))
    thetas = rng.uniform(0.0, 2.0 * math.pi, size=B) if train else np.zeros(B)
    vol = torch.randn(B, J, *N, generator=g)
    if peaked:
        flat = vol.view(B, J, -1)
        idx = torch.randint(0, flat.shape[-1], (B, J), generator=g)
        flat.scatter_add_(2, idx.unsqueeze(-1), torch.full((B, J, 1), 10.0))
    coord = make_coord_volumes(B, N, base, thetas, side)
    return {"volumes": vol, "coord_volumes": coord}


def make_dlt_inputs(B=64, V=4, J=17, seed=0, noise_px=2.0, with_conf=True, dome=False):
    """Inputs of ``triangulate_batch_of_points``: image-space cameras (384x384, f=960), joints ~ pelvis + N(0,400 mm),
    projected + N(0, noise_px) pixel noise; confidences U(0,1) normalised over views + 1e-5 (triangulation.py:173-174).
    Also returns the noise-free 3-D points."""
    g = _gen(seed)
    rng = np.random.RandomState(seed + 1000)
    base = rng.normal(0.0, 100.0, size=(B, 3))
    proj = make_cameras(B, V, base, seed + 2000, (384, 384), focal=240.0, dome=dome)  # focal scales with W/96 -> 960
    X = torch.from_numpy(base).float().unsqueeze(1) + 400.0 * torch.randn(B, J, 3, generator=g)
    Xh = torch.cat([X, torch.ones(B, J, 1)], dim=-1).double()
    p = torch.einsum("bvrc,bjc->bvjr", proj.double(), Xh)
    pts = (p[..., :2] / p[..., 2:3]).float() + noise_px * torch.randn(B, V, J, 2, generator=g)
    conf = None
    if with_conf:
        conf = torch.rand(B, V, J, generator=g)
        conf = conf / conf.sum(dim=1, keepdim=True) + 1e-5
    return {"proj_matricies": proj, "points": pts, "confidences": conf, "points_3d": X}


def inside_fraction(proj, coord, H, W):
    """Fraction of (voxel, view) pairs whose projection falls inside the heat map with positive depth (diagnostic)."""
    B, V = proj.shape[:2]
    pts = coord.reshape(B, -1, 3).double()
    ph = torch.cat([pts, torch.ones(B, pts.shape[1], 1, dtype=torch.float64)], dim=-1)
    p = torch.einsum("bvrc,bnc->bvnr", proj.double(), ph)
    z = p[..., 2]
    u, v = p[..., 0] / z, p[..., 1] / z
    ok = (z > 0) & (u >= 0) & (u < H) & (v >= 0) & (v < W)
    return ok.double().mean().item()
