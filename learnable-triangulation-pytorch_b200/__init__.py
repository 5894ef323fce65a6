"""lt_b200 — B200 (sm_100a) kernels for the geometric hot path of learnable-triangulation-pytorch.

The directory is called ``learnable-triangulation-pytorch_b200``; import it as ``lt_b200`` (the ``lt_b200.py`` shim at
the repository root registers it under that name).

    lt_b200.op.unproject_heatmaps                      <- mvn/utils/op.py:99
    lt_b200.op.integrate_tensor_3d_with_coordinates    <- mvn/utils/op.py:84
    lt_b200.op.integrate_tensor_2d                     <- mvn/utils/op.py:11
    lt_b200.op.integrate_tensor_3d                     <- mvn/utils/op.py:50
    lt_b200.multiview.triangulate_batch_of_points      <- mvn/utils/multiview.py:162
    lt_b200.loss.VolumetricCELoss                      <- mvn/models/loss.py:52
    lt_b200.volumetric.build_coord_volumes             <- mvn/models/triangulation.py:280-341 (inline block)
    lt_b200.CoordSpec                                  the same block as 18 floats per item, evaluated inside K1/K2
    lt_b200.triangulation.ProcessFeatures              <- mvn/models/triangulation.py:238-240,344-346 (1x1 conv, K7)
    lt_b200.slab.integrate_tensor_3d_with_coordinates_sharded   (z-slab multi-GPU variant, one small exchange)
    lt_b200.patch_reference()                          monkey-patches the names above into ``mvn``
"""
from . import _lib, coordspec, loss, multiview, op, slab, synthetic, torch_ops, triangulation, volumetric  # noqa: F401
from .coordspec import CoordSpec  # noqa: F401
from .volumetric import build_coord_volumes  # noqa: F401
from .loss import VolumetricCELoss, integrate_tensor_3d_with_coordinates_and_ce, volumetric_ce_loss  # noqa: F401
from .multiview import triangulate_batch_of_points, triangulate_point_from_multiple_views_linear_torch  # noqa: F401
from .op import integrate_tensor_2d, integrate_tensor_3d, integrate_tensor_3d_with_coordinates, unproject_heatmaps  # noqa: F401
from .patch import patch_reference, unpatch_reference  # noqa: F401
from .triangulation import ProcessFeatures, process_features, use_fused_process_features  # noqa: F401

__all__ = [
    "op", "multiview", "slab", "synthetic", "unproject_heatmaps", "integrate_tensor_3d_with_coordinates",
    "integrate_tensor_2d", "integrate_tensor_3d", "loss", "VolumetricCELoss", "volumetric_ce_loss",
    "integrate_tensor_3d_with_coordinates_and_ce", "volumetric", "build_coord_volumes", "CoordSpec", "coordspec",
    "triangulate_batch_of_points", "triangulate_point_from_multiple_views_linear_torch", "patch_reference",
    "unpatch_reference", "triangulation", "ProcessFeatures", "process_features", "use_fused_process_features",
]
