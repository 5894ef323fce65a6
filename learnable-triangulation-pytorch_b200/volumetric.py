"""Coordinate volumes for the volumetric net: the block of ``VolumetricTriangulationNet.forward`` that builds
``coord_volumes`` (reference ``mvn/models/triangulation.py:280-341`` with ``mvn/utils/volumetric.py:87-114``) as ONE
launch for the whole batch instead of a python loop of ~20 small ops per item. CUDA only.

The reference does this inline, so there is no function to mirror; the arguments are the quantities that block reads:
the per-item base point (pelvis), ``cuboid_side``, ``volume_size``, the rotation angle/axis and the
``transfer_cmu_to_human36m`` flag.
"""
import numpy as np
import torch

from . import _lib
from ._nvtx import annotate
from .op import _stream

from .coordspec import COORD_SPEC, CoordSpec  # noqa: F401


def rotation_matrix(axis, theta):
    """Rotation by ``theta`` about ``axis`` as the reference defines it (volumetric.py:87-100, Euler-Rodrigues with
    the unit quaternion q = (cos(theta/2), -axis*sin(theta/2))), written in closed form:
    R = (w^2 - |v|^2) I + 2 v v^T - 2 w [v]_x   for q = (w, v). float64."""
    n = np.asarray(axis, dtype=np.float64)
    n = n / np.linalg.norm(n)
    w = np.cos(0.5 * theta)
    v = -np.sin(0.5 * theta) * n
    cross = np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])
    return (w * w - v.dot(v)) * np.eye(3) + 2.0 * np.outer(v, v) - 2.0 * w * cross


def coord_volume_specs(base_points, cuboid_side, volume_size, thetas=None, axis=(0, 0, 1)):
    """Host side: (B, 18) float32 = [position, step, centre, rotation] per item, each rounded to fp32 exactly where the
    reference's torch ops round them (python/numpy scalars become fp32 operands of fp32 tensor ops)."""
    base_points = np.asarray(base_points, dtype=np.float64).reshape(-1, 3)
    B = base_points.shape[0]
    sizes = np.broadcast_to(np.asarray(volume_size, dtype=np.int64), (3,))
    thetas = np.zeros(B) if thetas is None else np.broadcast_to(np.asarray(thetas, dtype=np.float64), (B,))
    sides = np.array([cuboid_side] * 3, dtype=np.float64)                # triangulation.py:298
    spec = np.zeros((B, COORD_SPEC), dtype=np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        step = sides / (sizes - 1)                                       # triangulation.py:311-313
    for b in range(B):
        spec[b, 0:3] = base_points[b] - sides / 2                        # triangulation.py:299
        spec[b, 3:6] = step
        spec[b, 6:9] = base_points[b]                                    # triangulation.py:328
        spec[b, 9:18] = rotation_matrix(axis, thetas[b]).reshape(-1)     # volumetric.py:106-107
    return torch.from_numpy(spec)


@annotate("lt_b200.coord_volumes_from_specs")
def coord_volumes_from_specs(specs, volume_size, transfer_cmu_to_human36m=False):
    """The launch on its own: ``specs`` (B,18) fp32 ALREADY ON THE DEVICE (``coord_volume_specs`` + a pinned,
    non-blocking copy in a data-loading pipeline) -> (B,N0,N1,N2,3). No host work, no synchronisation."""
    if not (torch.is_tensor(specs) and specs.is_cuda):
        raise RuntimeError("lt_b200: specs must be a CUDA tensor (no CPU fallback)")
    if specs.dim() != 2 or specs.shape[1] != COORD_SPEC:
        raise RuntimeError("specs must be (B, {}), got {}".format(COORD_SPEC, tuple(specs.shape)))
    L = _lib.lib()
    N0, N1, N2 = (int(v) for v in np.broadcast_to(np.asarray(volume_size, dtype=np.int64), (3,)))
    B = specs.shape[0]
    with torch.cuda.device(specs.device):
        spec_d = specs.detach().to(torch.float32).contiguous()
        out = torch.empty((B, N0, N1, N2, 3), device=specs.device, dtype=torch.float32)
        _lib.check(L.lt_coord_volumes(_lib.ptr(spec_d), _lib.ptr(out), B, N0, N1, N2,
                                      1 if transfer_cmu_to_human36m else 0, _stream()))
    return out


@annotate("lt_b200.build_coord_volumes")
def build_coord_volumes(base_points, cuboid_side, volume_size, thetas=None, axis=(0, 0, 1),
                        transfer_cmu_to_human36m=False, device="cuda"):
    """(B, N0, N1, N2, 3) fp32 world coordinates of the voxel centres.

    base_points (B,3) array-like (pelvis, mm); cuboid_side float (mm); volume_size int or (N0,N1,N2); thetas None
    (evaluation: 0.0, triangulation.py:320) or B angles (training: uniform in [0, 2pi), :318); axis (0,0,1) for
    ``kind == 'mpii'``, (0,1,0) for ``'coco'`` (:322-325); transfer_cmu_to_human36m as in :336-339."""
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("lt_b200: build_coord_volumes only exists as a CUDA kernel (no CPU fallback)")
    N0, N1, N2 = (int(v) for v in np.broadcast_to(np.asarray(volume_size, dtype=np.int64), (3,)))
    spec = coord_volume_specs(base_points, cuboid_side, (N0, N1, N2), thetas, axis)
    with torch.cuda.device(device):
        return coord_volumes_from_specs(spec.to(device, non_blocking=True), (N0, N1, N2), transfer_cmu_to_human36m)
