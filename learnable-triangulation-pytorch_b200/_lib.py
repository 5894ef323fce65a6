"""ctypes binding of liblt_b200.so — the C ABI declared in include/lt_b200.h.

There is no CPU implementation behind this module: if the shared library is missing the import of any op
fails with a RuntimeError that says how to build it (``python -c "import __graft_entry__ as g; g.build()"``
or ``make -C learnable-triangulation-pytorch_b200/csrc``).
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# LT_B200_LIB: developer knob for A/B-ing differently built libraries on the GPU box (tools/gpu_variants.sh)
LIB_PATH = os.environ.get("LT_B200_LIB") or os.path.join(_HERE, "liblt_b200.so")
CSRC_DIR = os.path.join(_HERE, "csrc")

_c_float_p = ctypes.c_void_p  # device pointers travel as integers
_vp = ctypes.c_void_p
_i = ctypes.c_int
_ll = ctypes.c_longlong
_f = ctypes.c_float

# name -> (restype, argtypes); must list every symbol of include/lt_b200.h (tests/test_abi.py checks it)
SIGNATURES = {
    "lt_abi_version": (_i, []),
    "lt_last_error": (ctypes.c_char_p, []),
    "lt_padded_channels": (_i, [_i]),
    "lt_feat_layout_elems": (ctypes.c_size_t, [_i, _i, _i, _i]),
    "lt_feat_to_hwc": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "lt_feat_to_chw": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "lt_unproject_fwd": (_i, [_vp, _vp, _vp, _vp, _vp] + [_i] * 10 + [_vp]),
    "lt_unproject_bwd_ws_bytes": (ctypes.c_size_t, [_i] * 6),
    "lt_unproject_bwd": (_i, [_vp] * 8 + [_i] * 10 + [_vp]),
    "lt_softargmax_ws_bytes": (ctypes.c_size_t, [_i, _i, _ll]),
    "lt_softargmax_partial": (_i, [_vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f, _vp]),
    "lt_softargmax_finalize": (_i, [_vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f, _vp]),
    "lt_softargmax_fwd": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f, _vp]),
    "lt_softargmax_merge": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "lt_softargmax_fwd_noalloc": (_i, [_vp, _vp, _vp, _vp, _i, _i, _ll, _i, _f, _vp]),
    "lt_softargmax_bwd_dot": (_i, [_vp, _vp, _vp, _vp, _i, _i, _ll, _vp]),
    "lt_softargmax_bwd": (_i, [_vp] * 7 + [_i, _i, _ll, _i, _f, _vp]),
    "lt_unproject_fwd_spec": (_i, [_vp, _vp, _vp, _i, _vp, _vp] + [_i] * 10 + [_vp]),
    "lt_unproject_bwd_spec": (_i, [_vp] * 4 + [_i] + [_vp] * 4 + [_i] * 10 + [_vp]),
    "lt_softargmax_partial_spec": (_i, [_vp, _vp, _i, _vp, _vp] + [_i] * 6 + [_f, _vp]),
    "lt_softargmax_fwd_spec": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _vp] + [_i] * 6 + [_f, _vp]),
    "lt_softargmax_bwd_spec": (_i, [_vp, _vp, _i] + [_vp] * 5 + [_i] * 6 + [_f, _vp]),
    "lt_softargmax2d_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _f, _vp]),
    "lt_softargmax2d_bwd": (_i, [_vp] * 5 + [_i, _i, _i, _i, _f, _vp]),
    "lt_coord_volumes": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "lt_volce_ws_bytes": (ctypes.c_size_t, [_i, _i, _ll]),
    "lt_volce_fwd": (_i, [_vp] * 7 + [_i, _i, _ll, _vp]),
    "lt_volce_bwd": (_i, [_vp] * 5 + [_i, _i, _ll, _vp]),
    "lt_volce_bwd_sparse": (_i, [_vp] * 6 + [_i, _i, _ll, _vp]),
    "lt_softargmax_bwd_add_sparse": (_i, [_vp] * 4 + [_i, _i, _ll, _i, _f, _vp]),
    "lt_pwconv_ws_bytes": (ctypes.c_size_t, [_i] * 5),
    "lt_pwconv_fwd": (_i, [_vp] * 5 + [_i] * 5 + [_vp]),
    "lt_pwconv_bwd_data": (_i, [_vp] * 3 + [_i] * 5 + [_vp]),
    "lt_pwconv_bwd_weight": (_i, [_vp] * 5 + [_i] * 5 + [_vp]),
    "lt_dlt_fwd": (_i, [_vp] * 5 + [_i, _i, _i, _vp]),
    "lt_dlt_bwd": (_i, [_vp] * 7 + [_i, _i, _i, _vp]),
}

_lock = threading.Lock()
_lib = None


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into liblt_b200.so (nvcc cross-compiles without a GPU)."""
    jobs = str(min(8, os.cpu_count() or 1))
    proc = subprocess.run(["make", "-C", CSRC_DIR, "-j", jobs], capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout[-4000:])
        print(proc.stderr[-4000:])
    if proc.returncode != 0:
        raise RuntimeError("building liblt_b200.so failed (see output above)")
    return LIB_PATH


def lib():
    """The loaded library. Raises RuntimeError (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "lt_b200: CUDA library %s is missing and there is no CPU fallback. Build it with "
                "`make -C %s` (needs nvcc, sm_100a)." % (LIB_PATH, CSRC_DIR))
        # RTLD_GLOBAL: the torch extension (torch_ops.py) resolves its lt_* symbols against this library at load time
        handle = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        if handle.lt_abi_version() != 2:
            raise RuntimeError("lt_b200: ABI version mismatch")
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().lt_last_error().decode("utf-8", "replace")
        raise RuntimeError("lt_b200: %s (code %d)" % (msg, rc))


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
