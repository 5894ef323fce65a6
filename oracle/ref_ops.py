"""CPU oracle for the lt_b200 hot path.  TEST INFRASTRUCTURE ONLY.

This module restates, on the CPU with stock torch/numpy, the algorithm of the three reference functions that the CUDA
kernels replace. Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it; the product (``learnable-triangulation-pytorch_b200/``) never does.

Parity pinning: the reference (blufzzz/learnable-triangulation-pytorch) ships NO tests and NO golden vectors
(SURVEY.md §4), so the oracle is pinned against the reference ITSELF: ``oracle/make_golden.py`` imports the unmodified
reference from /root/reference, runs it on small seeded inputs and commits inputs+outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every function below against those vectors.

Third-party arithmetic the reference relies on (not under /root/reference): ``F.grid_sample``, ``F.softmax``,
``torch.einsum``, ``torch.svd`` from PyTorch (reference pin torch==1.0.1, requirements.txt:20; this image has 2.11).
``unproject_heatmaps`` below calls the same ``F.grid_sample`` as the reference does (op.py:134); ``bilinear_sample_np``
is an independent float64 restatement of its published algorithm (bilinear, zeros padding, both align_corners modes).
"""
import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------------------------------------
# unproject + aggregate  (reference: mvn/utils/op.py:99-163, mvn/utils/multiview.py:80-101)
# --------------------------------------------------------------------------------------------------------------------
def project_homogeneous(proj, pts):
    """[x y z 1] @ P^T for one view: proj (3,4), pts (n,3) -> (n,3).  multiview.py:46-60,96."""
    ones = torch.ones((pts.shape[0], 1), dtype=pts.dtype, device=pts.device)
    return torch.cat([pts, ones], dim=1) @ proj.t()


def sample_view(fmap, proj, pts, align_corners=False):
    """One view's per-voxel samples: fmap (C,H,W), proj (3,4), pts (n,3) -> (C,n).  op.py:114-138."""
    C, H, W = fmap.shape
    hom = project_homogeneous(proj, pts)
    behind = hom[:, 2] <= 0.0                                   # op.py:121
    depth = torch.where(hom[:, 2] == 0.0, torch.ones_like(hom[:, 2]), hom[:, 2])  # op.py:123
    uv = hom[:, :2] / depth.unsqueeze(1)                        # op.py:124 / multiview.py:75
    # op.py:128-129: x is scaled by heatmap_shape[0] (=H) and y by heatmap_shape[1] (=W), as written
    gx = 2 * (uv[:, 0] / H - 0.5)
    gy = 2 * (uv[:, 1] / W - 0.5)
    grid = torch.stack([gx, gy], dim=1).view(1, -1, 1, 2)       # op.py:133
    out = F.grid_sample(fmap.unsqueeze(0), grid, mode="bilinear", padding_mode="zeros",
                        align_corners=align_corners)            # op.py:134 (defaults; align_corners per torch version)
    out = out.view(C, -1).clone()
    out[:, behind] = 0.0                                        # op.py:138
    return out


def aggregate_views(stack, method, conf=None):
    """stack (V,C,n) -> (C,n).  op.py:147-161 (conf* / sum / max / softmax, else ValueError)."""
    if method.startswith("conf"):
        return (stack * conf.view(conf.shape[0], conf.shape[1], 1)).sum(0)
    if method == "sum":
        return stack.sum(0)
    if method == "max":
        return stack.max(0)[0]
    if method == "softmax":
        w = F.softmax(stack.reshape(stack.shape[0], -1), dim=0).view_as(stack)
        return (stack * w).sum(0)
    raise ValueError("Unknown volume_aggregation_method: {}".format(method))


def unproject_heatmaps(heatmaps, proj_matricies, coord_volumes, volume_aggregation_method="sum", vol_confidences=None,
                       align_corners=False, out_dtype=None):
    """heatmaps (B,V,C,H,W), proj (B,V,3,4), coord (B,N0,N1,N2,3) -> (B,C,N0,N1,N2).

    Same operation sequence as the reference (one grid_sample per (b, v), stack, aggregate). Runs in the dtype of the
    inputs; ``out_dtype=torch.float32`` reproduces the reference's fp32 result buffers (op.py:104,111) when fed fp64.
    """
    B, V, C = heatmaps.shape[:3]
    vshape = tuple(coord_volumes.shape[1:4])
    if not (volume_aggregation_method.startswith("conf") or volume_aggregation_method in ("sum", "max", "softmax")):
        raise ValueError("Unknown volume_aggregation_method: {}".format(volume_aggregation_method))
    outs = []
    for b in range(B):
        pts = coord_volumes[b].reshape(-1, 3)
        stack = torch.stack([sample_view(heatmaps[b, v], proj_matricies[b, v], pts, align_corners) for v in range(V)])
        if out_dtype is not None:
            stack = stack.to(out_dtype)
        conf = None
        if volume_aggregation_method.startswith("conf"):
            conf = vol_confidences[b].to(stack.dtype)
        outs.append(aggregate_views(stack, volume_aggregation_method, conf).view(C, *vshape))
    if not outs:
        return torch.zeros((0, C) + vshape, dtype=out_dtype or heatmaps.dtype)
    return torch.stack(outs)


def bilinear_sample_np(fmap, gx, gy, align_corners=False):
    """Independent float64 restatement of grid_sample(bilinear, zeros): fmap (C,H,W), gx/gy (n,) in [-1,1] -> (C,n)."""
    fmap = np.asarray(fmap, dtype=np.float64)
    C, H, W = fmap.shape
    gx, gy = np.asarray(gx, np.float64), np.asarray(gy, np.float64)
    if align_corners:
        ix, iy = (gx + 1) / 2 * (W - 1), (gy + 1) / 2 * (H - 1)
    else:
        ix, iy = ((gx + 1) * W - 1) / 2, ((gy + 1) * H - 1) / 2
    x0, y0 = np.floor(ix), np.floor(iy)
    out = np.zeros((C, gx.shape[0]))
    for dy in (0, 1):
        for dx in (0, 1):
            xx, yy = x0 + dx, y0 + dy
            w = (1 - np.abs(ix - xx)) * (1 - np.abs(iy - yy))
            ok = (xx >= 0) & (xx < W) & (yy >= 0) & (yy < H)
            xi, yi = np.clip(xx, 0, W - 1).astype(int), np.clip(yy, 0, H - 1).astype(int)
            out += np.where(ok, w, 0.0)[None, :] * fmap[:, yi, xi]
    return out


def unproject_heatmaps_np64(heatmaps, proj_matricies, coord_volumes, method="sum", vol_confidences=None,
                            align_corners=False):
    """All-float64 numpy evaluation of the same maths (no torch kernels involved): the high-precision referee."""
    hm = np.asarray(heatmaps, np.float64)
    P = np.asarray(proj_matricies, np.float64)
    cv = np.asarray(coord_volumes, np.float64)
    B, V, C, H, W = hm.shape
    vshape = cv.shape[1:4]
    out = np.zeros((B, C) + tuple(vshape))
    for b in range(B):
        pts = cv[b].reshape(-1, 3)
        hom_pts = np.concatenate([pts, np.ones((pts.shape[0], 1))], axis=1)
        stack = np.zeros((V, C, pts.shape[0]))
        for v in range(V):
            h = hom_pts @ P[b, v].T
            behind = h[:, 2] <= 0
            z = np.where(h[:, 2] == 0, 1.0, h[:, 2])
            gx, gy = 2 * (h[:, 0] / z / H - 0.5), 2 * (h[:, 1] / z / W - 0.5)
            s = bilinear_sample_np(hm[b, v], gx, gy, align_corners)
            s[:, behind] = 0.0
            stack[v] = s
        if method.startswith("conf"):
            agg = (stack * np.asarray(vol_confidences, np.float64)[b][:, :, None]).sum(0)
        elif method == "sum":
            agg = stack.sum(0)
        elif method == "max":
            agg = stack.max(0)
        elif method == "softmax":
            e = np.exp(stack - stack.max(0, keepdims=True))
            agg = (stack * e / e.sum(0, keepdims=True)).sum(0)
        else:
            raise ValueError("Unknown volume_aggregation_method: {}".format(method))
        out[b] = agg.reshape((C,) + tuple(vshape))
    return out


# --------------------------------------------------------------------------------------------------------------------
# volumetric soft-argmax  (reference: mvn/utils/op.py:84-96)
# --------------------------------------------------------------------------------------------------------------------
def integrate_tensor_3d_with_coordinates(volumes, coord_volumes, softmax=True):
    B, J = volumes.shape[:2]
    flat = volumes.reshape(B, J, -1)
    flat = F.softmax(flat, dim=2) if softmax else F.relu(flat)         # op.py:88-91 (relu: no normalisation)
    coords = torch.bmm(flat, coord_volumes.reshape(B, -1, 3).to(flat.dtype))   # == einsum("bnxyz,bxyzc->bnc"), op.py:94
    return coords, flat.view_as(volumes)


def softargmax_partial(volumes, coord_volumes, softmax=True, multiplier=1.0):
    """Slab-local tuple (m, S, Sx, Sy, Sz) per (b, joint) -> (B,J,5). CPU stand-in of lt_softargmax_partial."""
    B, J = volumes.shape[:2]
    x = volumes.reshape(B, J, -1).to(torch.float64) * multiplier
    xyz = coord_volumes.reshape(B, -1, 3).to(torch.float64)
    if softmax:
        m = x.max(dim=2).values if x.shape[2] > 0 else torch.full((B, J), -float("inf"), dtype=torch.float64)
        e = torch.exp(x - m.unsqueeze(-1)) if x.shape[2] > 0 else x
        S = e.sum(2)
    else:
        m = torch.zeros(B, J, dtype=torch.float64)
        e = torch.relu(x)
        S = torch.ones(B, J, dtype=torch.float64)
    W = torch.bmm(e, xyz)
    return torch.cat([m.unsqueeze(-1), S.unsqueeze(-1), W], dim=2).to(torch.float32)


def softargmax_merge(parts, softmax=True):
    """(R,B,J,5) -> (B,J,5): log-sum-exp merge in rank order. CPU stand-in of lt_softargmax_merge."""
    parts = parts.to(torch.float64)
    if not softmax:
        out = parts[0].clone()
        out[..., 2:] = parts[..., 2:].sum(0)
        return out.to(torch.float32)
    m = parts[..., 0].max(dim=0).values
    scale = torch.exp(parts[..., 0] - m.unsqueeze(0))
    scale = torch.where(torch.isnan(scale), torch.zeros_like(scale), scale)
    S = (parts[..., 1] * scale).sum(0)
    W = (parts[..., 2:] * scale.unsqueeze(-1)).sum(0)
    return torch.cat([m.unsqueeze(-1), S.unsqueeze(-1), W], dim=-1).to(torch.float32)


def softargmax_finalize(volumes, stats, softmax=True, multiplier=1.0):
    """prob slab and global coordinates from merged stats. CPU stand-in of lt_softargmax_finalize."""
    st = stats.to(torch.float64)
    x = volumes.to(torch.float64) * multiplier
    if softmax:
        prob = torch.exp(x - st[..., 0].view(*st.shape[:2], 1, 1, 1)) / st[..., 1].view(*st.shape[:2], 1, 1, 1)
        coords = st[..., 2:] / st[..., 1:2]
    else:
        prob = torch.relu(x)
        coords = st[..., 2:]
    return coords.to(torch.float32), prob.to(torch.float32)


class CpuSlabOps:
    """Oracle-backed device steps for lt_b200.slab (used by the gloo tests to drive the exchange logic on CPU)."""

    @staticmethod
    def partial(vol, coord, softmax, multiplier):
        return softargmax_partial(vol, coord, bool(softmax), multiplier)

    @staticmethod
    def merge(parts, softmax):
        return softargmax_merge(parts, bool(softmax))

    @staticmethod
    def finalize(vol, stats, softmax, multiplier):
        return softargmax_finalize(vol, stats, bool(softmax), multiplier)

    @staticmethod
    def dot(prob, g_prob):
        return (prob.double() * g_prob.double()).flatten(2).sum(2).float()

    @staticmethod
    def backward(prob, coord, coords, g_coords, g_prob, dots, softmax, multiplier):
        B, J = prob.shape[:2]
        G = torch.einsum("bjc,bxyzc->bjxyz", g_coords.double(), coord.double())
        if g_prob is not None:
            G = G + g_prob.double()
        if softmax:
            dot = (g_coords.double() * coords.double()).sum(-1)
            if dots is not None:
                dot = dot + dots.double()
            return (multiplier * prob.double() * (G - dot.view(B, J, 1, 1, 1))).float()
        return (multiplier * (prob > 0).double() * G).float()


def integrate_tensor_3d(volumes, softmax=True):
    """Voxel-index variant (reference: op.py:50-81): softmax or relu over the flattened volume, marginal masses per axis
    times arange, relu mode divides by the total mass; returns (coordinates (B,J,3), volumes)."""
    B, J, nx, ny, nz = volumes.shape
    flat = volumes.reshape(B, J, -1)
    flat = F.softmax(flat, dim=2) if softmax else F.relu(flat)
    vol = flat.reshape(B, J, nx, ny, nz)
    mass = (vol.sum(dim=3).sum(dim=3), vol.sum(dim=2).sum(dim=3), vol.sum(dim=2).sum(dim=2))
    out = []
    for m, n in zip(mass, (nx, ny, nz)):
        c = (m * torch.arange(n).type(torch.float).to(m.device).to(m.dtype)).sum(dim=2, keepdim=True)
        if not softmax:
            c = c / m.sum(dim=2, keepdim=True)
        out.append(c)
    return torch.cat(out, dim=2).reshape(B, J, 3), vol


# --------------------------------------------------------------------------------------------------------------------
# 2-D soft-argmax of joint heat maps  (reference: mvn/utils/op.py:11-47)
# --------------------------------------------------------------------------------------------------------------------
def integrate_tensor_2d(heatmaps, softmax=True):
    """heatmaps (B,J,H,W) -> ((B,J,2) as (x, y), (B,J,H,W)). Same operation sequence as the reference: softmax or relu
    over the flattened map (op.py:23-27), marginal masses (op.py:31-32) times arange (op.py:34-35); relu mode divides
    the coordinates by the total mass (op.py:39-41) and returns the un-normalised relu maps."""
    B, J, H, W = heatmaps.shape
    flat = heatmaps.reshape(B, J, -1)
    flat = F.softmax(flat, dim=2) if softmax else F.relu(flat)
    hm = flat.reshape(B, J, H, W)
    mass_x, mass_y = hm.sum(dim=2), hm.sum(dim=3)
    x = (mass_x * torch.arange(W, dtype=hm.dtype)).sum(dim=2, keepdim=True)
    y = (mass_y * torch.arange(H, dtype=hm.dtype)).sum(dim=2, keepdim=True)
    if not softmax:
        x = x / mass_x.sum(dim=2, keepdim=True)
        y = y / mass_y.sum(dim=2, keepdim=True)
    return torch.cat((x, y), dim=2), hm


# --------------------------------------------------------------------------------------------------------------------
# coordinate volumes  (reference: mvn/models/triangulation.py:306-341 + mvn/utils/volumetric.py:87-114)
# --------------------------------------------------------------------------------------------------------------------
def process_features(features, weight, bias=None):
    """The `process_features` step of VolumetricTriangulationNet: `nn.Sequential(nn.Conv2d(256, 32, 1))`
    (mvn/models/triangulation.py:238-240) applied to the (B*V, 256, H, W) backbone features (:344-346)."""
    if weight.dim() == 2:
        weight = weight[:, :, None, None]
    return F.conv2d(features, weight, bias)


def build_coord_volume(base_point, cuboid_side, volume_size, theta, axis, transfer_cmu_to_human36m=False,
                       device="cpu"):
    """One item, the same torch statements as the reference block: meshgrid -> position + step * index per axis
    (triangulation.py:306-313), minus centre, rot.mm(coord.t()).t(), plus centre (:331-333, volumetric.py:103-114),
    optional permute + reversed index_select (:336-339). base_point: float64 numpy (3,)."""
    n = volume_size
    sides = np.array([cuboid_side, cuboid_side, cuboid_side])
    position = base_point - sides / 2
    xxx, yyy, zzz = torch.meshgrid(torch.arange(n, device=device), torch.arange(n, device=device),
                                   torch.arange(n, device=device), indexing="ij")
    grid = torch.stack([xxx, yyy, zzz], dim=-1).type(torch.float).reshape((-1, 3))
    grid_coord = torch.zeros_like(grid)
    for c in range(3):
        grid_coord[:, c] = position[c] + (sides[c] / (n - 1)) * grid[:, c]
    cv = grid_coord.reshape(n, n, n, 3)
    center = torch.from_numpy(np.asarray(base_point)).type(torch.float).to(device)
    a = np.asarray(axis, dtype=np.float64)
    a = a / np.sqrt(np.dot(a, a))
    qa = np.cos(theta / 2.0)
    qb, qc, qd = -a * np.sin(theta / 2.0)
    rot = np.array([[qa * qa + qb * qb - qc * qc - qd * qd, 2 * (qb * qc + qa * qd), 2 * (qb * qd - qa * qc)],
                    [2 * (qb * qc - qa * qd), qa * qa + qc * qc - qb * qb - qd * qd, 2 * (qc * qd + qa * qb)],
                    [2 * (qb * qd + qa * qc), 2 * (qc * qd - qa * qb), qa * qa + qd * qd - qb * qb - qc * qc]])
    rot = torch.from_numpy(rot).type(torch.float).to(device)
    cv = cv - center
    cv = rot.mm(cv.view(-1, 3).t()).t().view(n, n, n, 3)
    cv = cv + center
    if transfer_cmu_to_human36m:
        cv = cv.permute(0, 2, 1, 3)
        cv = cv.index_select(1, torch.arange(cv.shape[1] - 1, -1, -1).long().to(device))
    return cv.contiguous()


# --------------------------------------------------------------------------------------------------------------------
# volumetric cross-entropy loss  (reference: mvn/models/loss.py:52-80)
# --------------------------------------------------------------------------------------------------------------------
def volumetric_ce_loss(coord_volumes_batch, volumes_batch_pred, keypoints_gt, keypoints_binary_validity):
    """Same statements as VolumetricCELoss.forward: per batch item the (J, N^3) distances of every voxel centre to
    every ground-truth joint (loss.py:63-69), argmin -> host indices (loss.py:71-72), then validity * -log(p + 1e-6)
    per joint in a python loop (loss.py:74-77), divided by the number of (b, j) pairs (loss.py:80)."""
    loss, n_losses = 0.0, 0
    for b in range(volumes_batch_pred.shape[0]):
        cv = coord_volumes_batch[b].unsqueeze(0)
        kp = keypoints_gt[b].unsqueeze(1).unsqueeze(1).unsqueeze(1)
        dists = torch.sqrt(((cv - kp) ** 2).sum(-1))
        dists = dists.view(dists.shape[0], -1)
        flat = torch.argmin(dists, dim=-1).detach().cpu().numpy()
        idx = np.stack(np.unravel_index(flat, volumes_batch_pred.shape[-3:]), axis=1)
        for j, ix in enumerate(idx):
            loss = loss + keypoints_binary_validity[b, j][0] * (-torch.log(volumes_batch_pred[b, j, ix[0], ix[1], ix[2]] + 1e-6))
            n_losses += 1
    return loss / n_losses


def volumetric_ce_nearest_voxels(coord_volumes_batch, keypoints_gt):
    """(B, J) flat index of the voxel nearest to each joint: the argmin of loss.py:66-71 on its own."""
    out = []
    for b in range(coord_volumes_batch.shape[0]):
        d = torch.sqrt(((coord_volumes_batch[b].unsqueeze(0) - keypoints_gt[b].view(-1, 1, 1, 1, 3)) ** 2).sum(-1))
        out.append(torch.argmin(d.view(d.shape[0], -1), dim=-1))
    return torch.stack(out)


# --------------------------------------------------------------------------------------------------------------------
# batched DLT  (reference: mvn/utils/multiview.py:132-174)
# --------------------------------------------------------------------------------------------------------------------
def dlt_system(proj, pts, conf=None):
    """A (B,J,2V,4) with rows conf_v*(pt[v,r]*P[v,2,:] - P[v,r,:]).  multiview.py:147-152."""
    B, V, J = pts.shape[:3]
    if conf is None:
        conf = torch.ones(B, V, J, dtype=pts.dtype, device=pts.device)
    P = proj.to(pts.dtype)
    A = P[:, :, None, 2:3, :] * pts.unsqueeze(-1)          # (B,V,J,2,4)
    A = A - P[:, :, None, :2, :]
    A = A * conf.view(B, V, J, 1, 1)
    return A.permute(0, 2, 1, 3, 4).reshape(B, J, 2 * V, 4)


def triangulate_batch_of_points(proj_matricies_batch, points_batch, confidences_batch=None, out_dtype=torch.float32):
    """(B,V,3,4),(B,V,J,2),(B,V,J)|None -> (B,J,3). SVD of A, last right singular vector, dehomogenise
    (multiview.py:154-157). Runs in the dtype of ``points_batch``; the reference returns fp32 (multiview.py:164)."""
    assert proj_matricies_batch.shape[1] == points_batch.shape[1]
    A = dlt_system(proj_matricies_batch, points_batch, confidences_batch)
    _, _, vh = torch.linalg.svd(A, full_matrices=False)
    h = -vh[..., 3, :]
    X = h[..., :3] / h[..., 3:4]
    return X.to(out_dtype) if out_dtype is not None else X
