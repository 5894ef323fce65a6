"""`lt_softargmax_fwd_noalloc` (include/lt_b200.h): K2 forward with its scratch in a static device array inside the
library, for callers that cannot allocate. It has no counterpart in the reference,
so it is checked against the existing entry point `lt_softargmax_fwd` — same kernels, caller-owned scratch
— through the C ABI, on consecutive calls that share the array (the second one finds the first one's contents in it),
and against the fp64 oracle. Tolerances as in tests/test_gpu_softargmax.py: coordinates 1e-3 mm, probabilities 1e-5
relative (max-norm).

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import pytest
import torch

from conftest import rel_err
from lt_b200 import _lib, synthetic

SCRATCH_BYTES = 8 << 20  # LT_NOALLOC_SCRATCH_BYTES


def _st():
    return torch.cuda.current_stream().cuda_stream


def _inputs(B, J, N, seed, peaked=True):
    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=seed, peaked=peaked, train=True)
    return d["volumes"].cuda().contiguous(), d["coord_volumes"].cuda().contiguous()


def _existing(L, vol, coord, softmax, mult):
    """the existing entry point, caller-owned scratch"""
    B, J = vol.shape[:2]
    nvox = vol[0, 0].numel()
    coords = torch.full((B, J, 3), float("nan"), device="cuda")
    prob = torch.full_like(vol, float("nan"))
    stats = torch.empty(B, J, 5, device="cuda")
    ws = torch.empty(max(L.lt_softargmax_ws_bytes(B, J, nvox), 16), device="cuda", dtype=torch.uint8)
    _lib.check(L.lt_softargmax_fwd(vol.data_ptr(), coord.data_ptr(), coords.data_ptr(), prob.data_ptr(),
                                   stats.data_ptr(), ws.data_ptr(), B, J, nvox, softmax, mult, _st()))
    return coords, prob


def _noalloc(L, vol, coord, softmax, mult, coords=None, prob=None):
    B, J = vol.shape[:2]
    nvox = vol[0, 0].numel()
    coords = torch.full((B, J, 3), float("nan"), device="cuda") if coords is None else coords
    prob = torch.full_like(vol, float("nan")) if prob is None else prob
    _lib.check(L.lt_softargmax_fwd_noalloc(vol.data_ptr(), coord.data_ptr(), coords.data_ptr(), prob.data_ptr(),
                                           B, J, nvox, softmax, mult, _st()))
    return coords, prob


def _same(got, want, what):
    (c, p), (c0, p0) = got, want
    torch.cuda.synchronize()
    assert not torch.isnan(c).any() and not torch.isnan(p).any(), what + ": output not fully written"
    assert (c - c0).abs().max().item() <= 1e-3, what + ": coordinates differ from lt_softargmax_fwd"
    assert rel_err(p, p0) <= 1e-5, what + ": probabilities differ from lt_softargmax_fwd"


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_existing_entry_point():
    L = _lib.lib()
    # call 1: 68 (item, joint) pairs at 32^3 on the vector paths
    vol1, coord1 = _inputs(4, 17, (32, 32, 32), seed=11)
    want1 = _existing(L, vol1, coord1, 1, 1.0)
    got1 = _noalloc(L, vol1, coord1, 1, 1.0)
    # call 2, straight after, no synchronisation in between: smaller, ragged (scalar paths), other mode and multiplier;
    # it finds call 1's tuples in the shared array and must not be affected by them
    vol2, coord2 = _inputs(2, 5, (5, 6, 7), seed=12, peaked=False)
    got2 = _noalloc(L, vol2, coord2, 0, 1.0)
    want2 = _existing(L, vol2, coord2, 0, 1.0)
    _same(got1, want1, "call 1")
    _same(got2, want2, "call 2")
    # call 3 repeats call 1 after the array held another shape's data: same answer again, bit for bit
    got3 = _noalloc(L, vol1, coord1, 1, 1.0)
    _same(got3, want1, "call 3")
    assert torch.equal(got3[0], got1[0]) and torch.equal(got3[1], got1[1])
    # and the fp64 oracle (the same bar as tests/test_gpu_softargmax.py)
    from oracle import ref_ops
    c64, p64 = ref_ops.integrate_tensor_3d_with_coordinates(vol1.double().cpu(), coord1.double().cpu(), softmax=True)
    assert (got3[0].double().cpu() - c64).abs().max().item() <= 1e-3
    assert rel_err(got3[1].cpu(), p64) <= 1e-5


@pytest.mark.gpu
def test_full_size_item_and_no_allocation_between_calls():
    """cfg3's item shape (17 joints at 64^3, 0.4 MB of the array) twice in a row; neither torch's allocator nor the
    driver hands out memory during the second call."""
    L = _lib.lib()
    vol, coord = _inputs(1, 17, (64, 64, 64), seed=13)
    want = _existing(L, vol, coord, 1, 2.0)
    coords = torch.full((1, 17, 3), float("nan"), device="cuda")
    prob = torch.full_like(vol, float("nan"))
    _noalloc(L, vol, coord, 1, 2.0, coords, prob)           # first call on this device: takes the symbol's address
    torch.cuda.synchronize()
    torch_before, (free_before, _) = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()
    coords.fill_(float("nan"))
    prob.fill_(float("nan"))
    _noalloc(L, vol, coord, 1, 2.0, coords, prob)
    torch.cuda.synchronize()
    torch_after, (free_after, _) = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()
    _same((coords, prob), want, "second call")
    assert torch_after == torch_before
    print("driver free memory around the second call: %d -> %d bytes" % (free_before, free_after))
    assert free_after >= free_before, "device memory was allocated during lt_softargmax_fwd_noalloc"


@pytest.mark.gpu
def test_captures_into_a_cuda_graph_and_replays():
    """what the entry point is for: under stream capture nothing may allocate; outputs and inputs are the caller's"""
    L = _lib.lib()
    vol, coord = _inputs(2, 17, (16, 16, 16), seed=14)
    coords = torch.zeros(2, 17, 3, device="cuda")
    prob = torch.zeros_like(vol)
    _noalloc(L, vol, coord, 1, 1.0, coords, prob)           # outside capture once (address look-up, module load)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _noalloc(L, vol, coord, 1, 1.0, coords, prob)
    for seed in (15, 16):                                    # new data in the same buffers, two replays in a row
        v, _ = _inputs(2, 17, (16, 16, 16), seed=seed)
        vol.copy_(v)
        coords.fill_(float("nan"))
        prob.fill_(float("nan"))
        g.replay()
        _same((coords, prob), _existing(L, vol, coord, 1, 1.0), "replay with seed %d" % seed)


def test_rejects_shapes_beyond_the_static_array():
    """argument checking only, nothing is launched (runs without a GPU): 17 joints at 256^3 need 25 MB of scratch"""
    L = _lib.lib()
    B, J, nvox = 1, 17, 256 ** 3
    assert 256 + L.lt_softargmax_ws_bytes(B, J, nvox) > SCRATCH_BYTES
    fake = 4096  # non-null, never dereferenced: the size check comes first
    rc = L.lt_softargmax_fwd_noalloc(fake, fake, fake, fake, B, J, nvox, 1, 1.0, None)
    msg = L.lt_last_error().decode()
    assert rc == -1 and "LT_NOALLOC_SCRATCH_BYTES" in msg and "lt_softargmax_fwd" in msg
    with pytest.raises(RuntimeError):
        _lib.check(rc)
    # the largest configuration of the bench that fits: cfg3, 16 x 17 joints at 64^3
    assert 16 * 17 * 5 * 4 + 256 + L.lt_softargmax_ws_bytes(16, 17, 64 ** 3) <= SCRATCH_BYTES
    # same argument errors as lt_softargmax_fwd
    assert L.lt_softargmax_fwd_noalloc(fake, fake, fake, fake, 1, 1, 8, 1, 0.0, None) == -1
    assert b"multiplier" in L.lt_last_error()
    assert L.lt_softargmax_fwd_noalloc(None, None, None, None, 1, 1, 8, 1, 1.0, None) == -1
    assert b"null" in L.lt_last_error()
    assert L.lt_softargmax_fwd_noalloc(None, None, None, None, 0, 17, 8, 1, 1.0, None) == 0   # empty batch: no-op
