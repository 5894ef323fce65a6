"""``CoordSpec``: a batch of coordinate volumes DESCRIBED (18 floats per item) instead of stored (12 bytes per voxel).

``VolumetricTriangulationNet.forward`` builds ``coord_volumes`` (B,N0,N1,N2,3) per item in python (reference
``mvn/models/triangulation.py:280-341`` with ``mvn/utils/volumetric.py:87-114``) and hands the tensor to
``op.unproject_heatmaps`` (:349) and ``op.integrate_tensor_3d_with_coordinates`` (:353). Every voxel centre is
``rot @ ((corner + step * index) - centre) + centre``: the drop-in ops accept a ``CoordSpec`` wherever they take
``coord_volumes`` and evaluate that expression inside K1 / K2 (``lt_*_spec`` entry points, csrc/coordspec.cuh) in the
reference's fp32 operation order, so the results are bit-identical to passing the tensor that ``materialize()`` (K6)
builds from the same spec, without the tensor ever existing."""
import numpy as np
import torch

COORD_SPEC = 18


class CoordSpec:
    """specs (B,18) fp32 CUDA tensor = [corner, step, centre, rotation row-major] per item (build it on the host with
    ``volumetric.coord_volume_specs``), volume_size int or (N0,N1,N2), transfer = the CMU->Human3.6M axis swap of
    triangulation.py:336-339. Quacks like the tensor where the ops look: ``shape``, ``device``, ``dim()``, ``[:B]``."""

    def __init__(self, specs, volume_size, transfer_cmu_to_human36m=False):
        if not (torch.is_tensor(specs) and specs.is_cuda):
            raise RuntimeError("lt_b200: specs must be a CUDA tensor (no CPU fallback)")
        if specs.dim() != 2 or specs.shape[1] != COORD_SPEC:
            raise RuntimeError("specs must be (B, {}), got {}".format(COORD_SPEC, tuple(specs.shape)))
        self.specs = specs.detach().to(torch.float32).contiguous()
        self.size = tuple(int(v) for v in np.broadcast_to(np.asarray(volume_size, dtype=np.int64), (3,)))
        self.transfer = bool(transfer_cmu_to_human36m)
        if self.transfer and self.size[1] != self.size[2]:
            raise RuntimeError("the CMU->Human3.6M axis swap needs N1 == N2 (triangulation.py:337)")
        if self.size[0] * self.size[1] * self.size[2] > 0x7fffffff:
            raise RuntimeError("more than 2^31 voxels per item")

    @property
    def shape(self):
        return torch.Size((self.specs.shape[0],) + self.size + (3,))

    @property
    def device(self):
        return self.specs.device

    @property
    def is_cuda(self):
        return True

    def dim(self):
        return 5

    def __getitem__(self, idx):
        if not isinstance(idx, slice):
            raise TypeError("CoordSpec only supports slicing along the batch axis")
        return CoordSpec(self.specs[idx], self.size, self.transfer)

    def slab(self, lo, hi):
        """The spec of the sub-volume [lo, hi) along the first volume axis (z-slab sharding): the corner moves by
        lo steps. NOT bit-identical to slicing the full tensor (corner + step*lo is rounded once more); use it where
        the slabs define the volume."""
        s = self.specs.clone()
        s[:, 0] = s[:, 0] + s[:, 3] * float(lo)
        return CoordSpec(s, (hi - lo,) + self.size[1:], self.transfer)

    def materialize(self):
        """The (B,N0,N1,N2,3) tensor itself (K6), for consumers that need one."""
        from . import volumetric
        return volumetric.coord_volumes_from_specs(self.specs, self.size, self.transfer)
