"""CPU: the C-ABI library loads and exports exactly what include/lt_b200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

import lt_b200
from lt_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "lt_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lt_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_expected_entry_points():
    syms = header_symbols()
    for s in ("lt_unproject_fwd", "lt_unproject_bwd", "lt_softargmax_fwd", "lt_softargmax_bwd", "lt_softargmax_partial",
              "lt_softargmax_finalize", "lt_softargmax_merge", "lt_dlt_fwd", "lt_dlt_bwd", "lt_feat_to_hwc"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        _lib.build()
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for s in header_symbols():
        assert hasattr(handle, s), "missing export: " + s


def test_python_binding_covers_header():
    assert sorted(_lib.SIGNATURES) == header_symbols()


def test_host_side_entry_points_without_gpu():
    L = _lib.lib()
    assert L.lt_abi_version() == 2
    assert L.lt_padded_channels(32) == 32 and L.lt_padded_channels(17) == 32
    assert L.lt_padded_channels(33) == 64 and L.lt_padded_channels(256) == 256 and L.lt_padded_channels(300) == 384
    assert L.lt_softargmax_ws_bytes(16, 17, 64 ** 3) == 16 * 17 * 128 * 9 * 5 * 4 + (1 + 2 * 16) * 4   # 2048-voxel chunks, 8 warps, 5 floats (+ the fused forward's ticket / per-item counters and flags)
    # gather layout: (BV, H+2, W+2, Cp) floats
    assert L.lt_feat_layout_elems(8, 32, 96, 96) == 8 * 98 * 98 * 32 and L.lt_feat_layout_elems(2, 17, 5, 7) == 2 * 7 * 9 * 32
    # bad arguments come back as LT_ERR_INVALID with a message, nothing is launched
    rc = L.lt_dlt_fwd(None, None, None, None, None, 1, 0, 1, None)
    assert rc == -1 and b"LT_DLT_MAX_VIEWS" in L.lt_last_error()
    rc = L.lt_unproject_fwd(None, None, None, None, None, 1, 4, 32, 96, 96, 64, 64, 64, 7, 0, None)
    assert rc == -1 and b"aggregation" in L.lt_last_error()
    with pytest.raises(RuntimeError):
        _lib.check(rc)
    # tap descriptors are 32-bit offsets from an item's first map: V * (H+2) * (W+2) * Cp must stay below 2^31
    rc = L.lt_unproject_fwd(None, None, None, None, None, 1, 64, 32, 1022, 1022, 8, 8, 8, 0, 0, None)
    assert rc == -1 and b"bad shape" in L.lt_last_error()
    rc = L.lt_unproject_bwd(None, None, None, None, None, None, None, None, 1, 64, 32, 1022, 1022, 8, 8, 8, 0, 0, None)
    assert rc == -1 and b"bad shape" in L.lt_last_error()


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/liblt_b200.so")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.lib()
