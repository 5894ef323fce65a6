"""``torch.ops.lt_b200.*`` — the thin torch extension over the C ABI (torch_ext/lt_b200_torch.cpp).

The reference's three hot-path functions as dispatcher ops with C++ autograd, the raw forward / backward launches they
consist of as ops of their own, and (here) the fake-tensor registrations that give torch.compile / export the output
shapes. The extension contains no kernel: it allocates with ATen, takes torch's current stream and calls the ``lt_*``
entry points of liblt_b200.so, which python loads first (RTLD_GLOBAL) so that the extension needs no link line.

    build()      compile torch_ext/lt_b200_torch.cpp in-tree with torch.utils.cpp_extension (g++, no nvcc needed)
    available()  load it if it has been built; False otherwise (the python ops then go through ctypes: same kernels)
"""
import os
import threading

import torch

from . import _lib

_HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(_HERE, "torch_ext", "lt_b200_torch.cpp")
BUILD_DIR = os.path.join(_HERE, "torch_ext", "build")
SO_PATH = os.path.join(BUILD_DIR, "lt_b200_torch.so")

_lock = threading.Lock()
_state = {"loaded": None}


def build(verbose=False):
    """g++ -> torch_ext/build/lt_b200_torch.so (in-tree: it travels with the gpurun snapshot). Returns the path."""
    from torch.utils import cpp_extension
    os.makedirs(BUILD_DIR, exist_ok=True)
    _lib.lib()  # the extension's lt_* symbols resolve against the already loaded library
    cpp_extension.load(name="lt_b200_torch", sources=[SRC], build_directory=BUILD_DIR, with_cuda=True,
                       extra_cflags=["-O2", "-std=c++17"], is_python_module=False, verbose=verbose)
    _state["loaded"] = None
    return SO_PATH


def available():
    """True once torch.ops.lt_b200 is usable. Never builds (the driver's build() does); disabled with
    LT_B200_TORCH_EXT=0 (A/B against the ctypes binding)."""
    if _state["loaded"] is not None:
        return _state["loaded"]
    with _lock:
        if _state["loaded"] is not None:
            return _state["loaded"]
        ok = False
        if os.environ.get("LT_B200_TORCH_EXT", "1") != "0" and os.path.exists(SO_PATH) and os.path.exists(_lib.LIB_PATH):
            try:
                _lib.lib()
                if not hasattr(torch.ops.lt_b200, "unproject_heatmaps"):
                    torch.ops.load_library(SO_PATH)
                _register_fakes()
                ok = True
            except Exception as e:  # noqa: BLE001 - stale build against another torch: say so, use the ctypes binding
                import warnings
                warnings.warn("lt_b200: torch extension %s not usable (%s); using the ctypes binding" % (SO_PATH, e))
        _state["loaded"] = ok
    return ok


_fakes_done = False


def _register_fakes():
    """Output shapes of the raw ops for fake tensors (torch.compile / export trace them as opaque calls)."""
    global _fakes_done
    if _fakes_done:
        return
    _fakes_done = True
    L = _lib.lib()
    reg = torch.library.register_fake

    @reg("lt_b200::feat_to_hwc")
    def _(heatmaps):
        BV, C, H, W = heatmaps.shape
        return heatmaps.new_empty((L.lt_feat_layout_elems(BV, C, H, W),))

    @reg("lt_b200::feat_to_chw")
    def _(hwcp, BV, C, H, W):
        return hwcp.new_empty((BV, C, H, W))

    @reg("lt_b200::unproject_fwd")
    def _(feat_hwc, proj, coord, conf, B, V, C, H, W, N0, N1, N2, agg, align_corners, coord_is_spec, transfer):
        return feat_hwc.new_empty((B, C, N0, N1, N2))

    @reg("lt_b200::unproject_bwd")
    def _(grad_out, feat_hwc, proj, coord, conf, B, V, C, H, W, N0, N1, N2, agg, align_corners, coord_is_spec, transfer,
          need_conf):
        return torch.empty_like(feat_hwc), feat_hwc.new_empty((B, V, C) if (need_conf and conf is not None) else (0,))

    @reg("lt_b200::softargmax3d_fwd")
    def _(volumes, coord, softmax, multiplier, coord_is_spec, transfer):
        return volumes.new_empty(volumes.shape[:2] + (3,)), torch.empty_like(volumes)

    @reg("lt_b200::softargmax3d_bwd")
    def _(prob, coord, coords, grad_coords, grad_prob, softmax, multiplier, coord_is_spec, transfer):
        return torch.empty_like(prob)

    @reg("lt_b200::dlt_fwd")
    def _(proj, points, conf):
        B, V, J = points.shape[:3]
        return points.new_empty((B, J, 3)), points.new_empty((B, J, 20), dtype=torch.float64)

    @reg("lt_b200::dlt_bwd")
    def _(grad_out, proj, points, conf, saved, need_conf):
        B, V, J = points.shape[:3]
        return points.new_empty((B, V, J, 2)), points.new_empty((B, V, J) if (need_conf and conf is not None) else (0,))
