"""GPU: K7, the `process_features` 1x1 convolution (mvn/models/triangulation.py:238-240, 344-346) writing K1's gather
layout directly, through the C ABI (lt_pwconv_*) and the drop-in module, against the golden vectors of the unmodified
reference module and the oracle restatement. Tolerances: output 1e-5 relative (the bar of the unprojected volume it
feeds), gradients 1e-5 (BASELINE.json allows 1e-4: these sums are fixed-order FMA chains, not atomics)."""
import pytest
import torch

import lt_b200
from conftest import golden_names, load_golden, rel_err
from lt_b200 import _lib, op as lt_op, synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu


def _ours(feats, w, b):
    x = feats.reshape(-1, *feats.shape[-3:])
    y = lt_b200.process_features(x, w, b)
    return y


@pytest.mark.parametrize("name", golden_names("pwconv_"))
def test_golden(name):
    g = load_golden(name)
    f5 = g["features"].cuda().requires_grad_(True)
    w = g["weight"].cuda().requires_grad_(True)
    b = g["bias"].cuda().requires_grad_(True)
    y = lt_b200.process_features(f5.view(-1, *f5.shape[2:]), w, b)
    y5 = y.view(f5.shape[0], f5.shape[1], *y.shape[1:])                # triangulation.py:346 on the strided result
    assert y5.shape == g["out"].shape and y5.dtype == torch.float32
    assert rel_err(y5.cpu(), g["out"]) < 1e-5
    y5.backward(g["grad_out"].cuda())
    assert rel_err(f5.grad.cpu(), g["grad_features"]) < 1e-5
    assert rel_err(w.grad.cpu(), g["grad_weight"]) < 1e-5
    assert rel_err(b.grad.cpu(), g["grad_bias"]) < 1e-5


SHAPES = [  # BV, Cin, Cout, H, W, bias
    (2, 256, 32, 24, 24, True),     # the reference's channel counts
    (3, 40, 17, 9, 7, True),        # ragged everything: H*W % 4 != 0, Cout < 32 (zero padded channels), Cin % 16 != 0
    (1, 300, 40, 10, 12, False),    # Cin > 256 (two resident blocks), Cout > 32 (two texel blocks, accumulate pass)
    (2, 16, 130, 6, 6, True),       # Cp = 256
    (1, 3, 1, 1, 1, True),          # a single pixel
    (5, 256, 32, 19, 23, True),     # pixel tile tail
    # dispatch corners of the tcgen05 kernels (forward: Cout <= 32, Cin % 32 == 0, Cin <= 256, H*W % 4 == 0; backward-data:
    # Cin % 64 == 0; backward-weight: Cin in {128, 256}, H*W % 32 == 0)
    (3, 128, 32, 16, 16, True),     # one 128-channel block in the weight gradient, N = 128 accumulator in backward-data
    (2, 64, 20, 8, 8, True),        # Cin = 64: tensor-core forward and backward-data, FFMA weight gradient; zero-padded Cout
    (1, 256, 7, 12, 12, False),     # H*W = 144: partial 128-pixel tile on the tensor cores, not a multiple of 32 -> FFMA dW
    (70, 256, 32, 8, 8, True),      # more maps than tiles per map: every CTA walks several maps, half-empty tiles
    (2, 256, 32, 40, 48, True),     # 15 tiles per map (H*W = 1920), more tiles than CTAs would get in one wave at B*V = 64
    (1, 32, 32, 4, 8, True),        # one 32-channel stage, a single 32-pixel chunk
]


@pytest.mark.parametrize("BV,Cin,Cout,H,W,with_bias", SHAPES)
def test_oracle_shapes(BV, Cin, Cout, H, W, with_bias):
    gen = torch.Generator().manual_seed(BV * 1000 + Cin + Cout)
    x0 = torch.randn(BV, Cin, H, W, generator=gen)
    w0 = torch.randn(Cout, Cin, 1, 1, generator=gen) / Cin ** 0.5
    b0 = torch.randn(Cout, generator=gen) if with_bias else None
    gy = torch.randn(BV, Cout, H, W, generator=gen)

    def run(fn, dev, dtype):
        x = x0.to(dev, dtype).requires_grad_(True)
        w = w0.to(dev, dtype).requires_grad_(True)
        b = b0.to(dev, dtype).requires_grad_(True) if with_bias else None
        y = fn(x, w, b)
        y.backward(gy.to(dev, dtype))
        return [t.detach().cpu().double() for t in (y, x.grad, w.grad) + ((b.grad,) if with_bias else ())]

    ref = run(ref_ops.process_features, "cpu", torch.float64)
    ours = run(lt_b200.process_features, "cuda", torch.float32)
    for name, a, r in zip(("out", "grad_x", "grad_w", "grad_b"), ours, ref):
        assert a.shape == r.shape, name
        assert rel_err(a, r) < 1e-5, name


def test_gather_layout_written_in_place():
    """The result IS the gather layout: zero border, zero padded channels, interior = conv; unproject_heatmaps takes it
    without K0 and returns its gradient map the same way (no K0 inverse)."""
    L = _lib.lib()
    BV, Cin, Cout, H, W = 4, 64, 17, 10, 12
    gen = torch.Generator().manual_seed(5)
    x = torch.randn(BV, Cin, H, W, generator=gen).cuda()
    w = torch.randn(Cout, Cin, 1, 1, generator=gen).cuda()
    b = torch.randn(Cout, generator=gen).cuda()
    y = lt_b200.process_features(x, w, b)
    base = lt_op._hwcp_base(y, 'feat')
    assert base is not None and base.numel() == L.lt_feat_layout_elems(BV, Cout, H, W)
    Cp = L.lt_padded_channels(Cout)
    full = base.view(BV, H + 2, W + 2, Cp)
    assert full[:, 0].abs().max() == 0 and full[:, -1].abs().max() == 0
    assert full[:, :, 0].abs().max() == 0 and full[:, :, -1].abs().max() == 0
    assert full[..., Cout:].abs().max() == 0
    # K0 of the dense result gives the same buffer
    dense = y.contiguous()
    k0 = torch.empty_like(base)
    _lib.check(L.lt_feat_to_hwc(dense.data_ptr(), k0.data_ptr(), BV, Cout, H, W, None))
    torch.cuda.synchronize()
    assert torch.equal(k0, base)
    # views keep the fast path, copies and slices lose it (and fall back to K0)
    assert lt_op._hwcp_base(y.view(2, 2, Cout, H, W), 'feat') is not None
    assert lt_op._hwcp_base(dense, 'feat') is None
    assert lt_op._hwcp_base(y[1:], 'feat') is None
    assert lt_op._hwcp_base(y, 'grad') is None


@pytest.mark.parametrize("agg", ["softmax", "conf", "sum"])
def test_chain_into_unproject(agg, monkeypatch):
    """process_features -> .view(B, V, ...) -> unproject_heatmaps as in forward() (:344-349): same volume and same
    gradients into the backbone features, the conv weight and bias as the oracle chain, with K0 never launched."""
    B, V, C, N, Cin = 2, 4, 32, 12, 48
    d = synthetic.make_unproject_inputs(B=B, V=V, C=C, H=20, W=24, N=N, seed=9, train=True,
                                        conf="conf" if agg == "conf" else None)
    gen = torch.Generator().manual_seed(17)
    x0 = torch.randn(B, V, Cin, 20, 24, generator=gen)
    w0 = torch.randn(C, Cin, 1, 1, generator=gen) / Cin ** 0.5
    b0 = torch.randn(C, generator=gen) * 0.1
    gout = torch.randn(B, C, N, N, N, generator=gen)

    def chain(dev, ours, dtype):
        x = x0.to(dev, dtype).requires_grad_(True)
        w = w0.to(dev, dtype).requires_grad_(True)
        b = b0.to(dev, dtype).requires_grad_(True)
        proj, coord = d["proj_matricies"].to(dev, dtype), d["coord_volumes"].to(dev, dtype)
        vc = d["vol_confidences"].to(dev, dtype) if agg == "conf" else None
        f = x.view(-1, *x.shape[2:])
        f = (lt_b200.process_features if ours else ref_ops.process_features)(f, w, b)
        f = f.view(B, V, *f.shape[1:])
        vol = (lt_b200.unproject_heatmaps if ours else ref_ops.unproject_heatmaps)(f, proj, coord, agg, vc)
        vol.backward(gout.to(dev, vol.dtype))
        return [t.detach().cpu() for t in (vol, x.grad, w.grad, b.grad)]

    calls = {"n": 0}
    L = _lib.lib()
    real_to_hwc, real_to_chw = L.lt_feat_to_hwc, L.lt_feat_to_chw

    class Counting:
        def __init__(self, fn):
            self.fn = fn

        def __call__(self, *a):
            calls["n"] += 1
            return self.fn(*a)

    monkeypatch.setattr(L, "lt_feat_to_hwc", Counting(real_to_hwc), raising=False)
    monkeypatch.setattr(L, "lt_feat_to_chw", Counting(real_to_chw), raising=False)
    ours = chain("cuda", True, torch.float32)
    assert calls["n"] == 0, "the layout passes K0 / K0 inverse ran although the features were written as hwcp"
    ref = chain("cpu", False, torch.float32)
    assert rel_err(ours[0], ref[0]) < 1e-5
    for name, a, r in zip(("grad_x", "grad_w", "grad_b"), ours[1:], ref[1:]):
        assert rel_err(a, r) < 1e-4, name


def test_full_size_against_fp32_torch_on_gpu():
    """cfg3's shape (B*V = 64 maps of 256 x 96 x 96 -> 32 channels) against the oracle run with CUDA tensors, TF32 off."""
    BV, Cin, Cout, H, W = 64, 256, 32, 96, 96
    gen = torch.Generator(device="cuda").manual_seed(3)
    x0 = torch.randn(BV, Cin, H, W, device="cuda", generator=gen)
    w0 = torch.randn(Cout, Cin, 1, 1, device="cuda", generator=gen) / 16
    b0 = torch.randn(Cout, device="cuda", generator=gen)
    gy = torch.randn(BV, Cout, H, W, device="cuda", generator=gen)
    saved = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
    try:
        res = []
        for fn in (lt_b200.process_features, ref_ops.process_features):
            x, w, b = (t.clone().requires_grad_(True) for t in (x0, w0, b0))
            y = fn(x, w, b)
            y.backward(gy)
            res.append((y.detach(), x.grad, w.grad, b.grad))
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = saved
    for name, a, r in zip(("out", "grad_x", "grad_w", "grad_b"), *res):
        assert rel_err(a.cpu(), r.cpu()) < 2e-5, name
    # fixed summation order: the weight gradient is bit-reproducible
    x, w, b = (t.clone().requires_grad_(True) for t in (x0, w0, b0))
    lt_b200.process_features(x, w, b).backward(gy)
    assert torch.equal(w.grad, res[0][2]) and torch.equal(b.grad, res[0][3]) and torch.equal(x.grad, res[0][1])


def test_guard_bands_and_empty_batch():
    """Every output of the three entry points carved out of a sentinel-filled buffer at a ragged shape."""
    L = _lib.lib()
    BV, Cin, Cout, H, W = 3, 37, 19, 7, 9
    gen = torch.Generator().manual_seed(23)
    x = torch.randn(BV, Cin, H, W, generator=gen).cuda()
    w = torch.randn(Cout, Cin, generator=gen).cuda()
    b = torch.randn(Cout, generator=gen).cuda()
    pad, S = 64, 7.25
    n_out = L.lt_feat_layout_elems(BV, Cout, H, W)
    ws = torch.empty(max(L.lt_pwconv_ws_bytes(BV, Cin, Cout, H, W), 4) // 4, device="cuda")

    def carve(n):
        buf = torch.full((n + 2 * pad,), S, device="cuda")
        return buf, buf[pad:pad + n]

    buf_o, out = carve(n_out)
    _lib.check(L.lt_pwconv_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(), ws.data_ptr(), BV, Cin, Cout, H, W, None))
    buf_x, gx = carve(x.numel())
    _lib.check(L.lt_pwconv_bwd_data(out.data_ptr(), w.data_ptr(), gx.data_ptr(), BV, Cin, Cout, H, W, None))
    buf_w, gw = carve(w.numel())
    buf_b, gb = carve(Cout)
    _lib.check(L.lt_pwconv_bwd_weight(x.data_ptr(), out.data_ptr(), gw.data_ptr(), gb.data_ptr(), ws.data_ptr(), BV, Cin, Cout, H, W, None))
    torch.cuda.synchronize()
    for buf, payload in ((buf_o, out), (buf_x, gx), (buf_w, gw), (buf_b, gb)):
        assert (buf[:pad] == S).all() and (buf[-pad:] == S).all()
        assert not (payload == S).any()
    # empty batch: nothing launched, zero weight gradient
    y = lt_b200.process_features(x[:0].clone().requires_grad_(True), w.clone().requires_grad_(True), b)
    assert y.shape == (0, Cout, H, W)
    assert L.lt_pwconv_fwd(None, None, None, None, None, 1, 0, 32, 8, 8, None) == -1


def test_forward_with_a_16_byte_aligned_output_buffer():
    """The tensor-core forward stores 32 bytes per instruction; an output buffer that is only 16-byte aligned (legal per
    include/lt_b200.h) must take the other kernel and give the same values."""
    L = _lib.lib()
    BV, Cin, Cout, H, W = 2, 256, 32, 16, 16
    gen = torch.Generator().manual_seed(31)
    x = torch.randn(BV, Cin, H, W, generator=gen).cuda()
    w = (torch.randn(Cout, Cin, generator=gen) / 16).cuda()
    b = torch.randn(Cout, generator=gen).cuda()
    n = L.lt_feat_layout_elems(BV, Cout, H, W)
    ws = torch.empty(max(L.lt_pwconv_ws_bytes(BV, Cin, Cout, H, W), 4) // 4, device="cuda")
    buf = torch.zeros(n + 16, device="cuda")
    outs = []
    for shift in (0, 4):                        # 0: 32-byte aligned (tcgen05 path), 4 floats: 16-byte aligned only
        out = buf[shift:shift + n]
        assert out.data_ptr() % 32 == (16 if shift else 0)
        _lib.check(L.lt_pwconv_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), out.data_ptr(), ws.data_ptr(), BV, Cin, Cout,
                                   H, W, None))
        torch.cuda.synchronize()
        outs.append(out.clone())
    assert rel_err(outs[1].cpu(), outs[0].cpu()) < 2e-6
    assert outs[0].abs().max() > 0


def test_drop_in_module_keeps_state_dict():
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.Conv2d(256, 32, 1)).cuda()          # triangulation.py:238-240
    ours = lt_b200.ProcessFeatures().cuda()
    ours.load_state_dict(ref.state_dict())
    assert list(ours.state_dict()) == list(ref.state_dict()) == ["0.weight", "0.bias"]
    x = torch.randn(2, 256, 12, 12, device="cuda")
    saved = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        assert rel_err(ours(x).cpu(), ref(x).cpu()) < 1e-5

        class Net(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.process_features = torch.nn.Sequential(torch.nn.Conv2d(256, 32, 1))

        net = Net().cuda()
        want = net.process_features(x)
        lt_b200.use_fused_process_features(net)
        assert isinstance(net.process_features, lt_b200.ProcessFeatures)
        assert rel_err(net.process_features(x).cpu(), want.cpu()) < 1e-5
    finally:
        torch.backends.cudnn.allow_tf32 = saved


def _fuzz_cases(n=14, seed=20261105):
    import random
    rnd = random.Random(seed)
    cases = []
    for _ in range(n):
        cin = rnd.choice([32, 64, 96, 128, 192, 256, 256, 256, 24, 300])
        cout = rnd.choice([32, 32, 17, 8, 33])
        h, w = rnd.randint(1, 20), rnd.choice([4, 8, 12, 16, 5, 9])
        cases.append((rnd.randint(1, 5), cin, cout, h, w, rnd.random() < 0.7))
    return cases


@pytest.mark.parametrize("BV,Cin,Cout,H,W,with_bias", _fuzz_cases())
def test_random_shapes_over_the_dispatch_space(BV, Cin, Cout, H, W, with_bias):
    """Seeded random shapes across the tensor-core / FFMA dispatch boundaries of all three kernels."""
    test_oracle_shapes(BV, Cin, Cout, H, W, with_bias)


def test_in_place_edit_of_the_features_is_caught_like_in_the_reference():
    """unproject_heatmaps reads the gather-layout buffer behind the features in place; editing them between forward and
    backward must raise autograd's version error (as it does for the reference's grid_sample), not change the gradient."""
    B, V, N = 1, 2, 8
    d = synthetic.make_unproject_inputs(B=B, V=V, C=32, H=16, W=16, N=N, seed=2, train=True)
    x = torch.randn(B * V, 64, 16, 16, device="cuda", requires_grad=True)
    w = torch.randn(32, 64, 1, 1, device="cuda", requires_grad=True)
    f = lt_b200.process_features(x, w, None)
    f5 = f.view(B, V, *f.shape[1:])
    vol = lt_b200.unproject_heatmaps(f5, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), "softmax")
    with pytest.raises(RuntimeError, match="(?i)inplace"):   # at the edit itself or at backward, depending on the torch version
        with torch.no_grad():
            f5.mul_(2.0)
        vol.sum().backward()
