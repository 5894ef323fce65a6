#!/usr/bin/env python
"""Measurement of K7 (process_features 1x1 convolution written in the gather layout) at BASELINE
config 3's shape: B*V = 64 maps, 256 -> 32 channels, 96 x 96. Each kernel through the C ABI with CUDA events on the
launching stream, next to the reference's statement (`nn.Conv2d(256, 32, 1)` = F.conv2d, stock cuDNN on the same GPU) with TF32 off (the fp32 result the parity bar is) and on (cuDNN's
default). One JSON line. Inputs (604 MB) exceed L2, so no flush between iterations."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lt_b200  # noqa: E402


def timed(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3  # us


def main():
    BV, Cin, Cout, H, W = 64, 256, 32, 96, 96
    peak = 6545.6
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    L, st = lt_b200._lib.lib(), torch.cuda.current_stream().cuda_stream
    gen = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(BV, Cin, H, W, device="cuda", generator=gen)
    w = torch.randn(Cout, Cin, device="cuda", generator=gen) / 16
    b = torch.randn(Cout, device="cuda", generator=gen)
    n_hwcp = L.lt_feat_layout_elems(BV, Cout, H, W)
    out = torch.empty(n_hwcp, device="cuda")
    g = torch.randn(n_hwcp, device="cuda", generator=gen)
    gx, gw, gb = torch.empty_like(x), torch.empty_like(w), torch.empty_like(b)
    ws = torch.empty(L.lt_pwconv_ws_bytes(BV, Cin, Cout, H, W) // 4, device="cuda")
    P = [t.data_ptr() for t in (x, w, b, out, g, gx, gw, gb, ws)]
    xp, wp, bp, op, gp, gxp, gwp, gbp, wsp = P
    fwd = timed(lambda: L.lt_pwconv_fwd(xp, wp, bp, op, wsp, BV, Cin, Cout, H, W, st))
    bwd_d = timed(lambda: L.lt_pwconv_bwd_data(gp, wp, gxp, BV, Cin, Cout, H, W, st))
    bwd_w = timed(lambda: L.lt_pwconv_bwd_weight(xp, gp, gwp, gbp, wsp, BV, Cin, Cout, H, W, st))
    k0 = timed(lambda: L.lt_feat_to_hwc(gxp, op, BV, Cout, H, W, st))  # the layout pass K7 makes unnecessary

    xb, ob = x.numel() * 4, n_hwcp * 4
    flops = 2.0 * BV * H * W * Cin * Cout
    w4 = w.view(Cout, Cin, 1, 1)
    res = {"workload": "cfg3 shape: B*V=64, 256->32 channels, 96x96", "hbm_peak_gbs": peak, "flops_per_call": flops}
    for name, us, nbytes in (("fwd", fwd, xb + ob), ("bwd_data", bwd_d, xb + ob), ("bwd_weight", bwd_w, xb + ob)):
        res["k7_%s_us" % name] = us
        res["k7_%s_algorithmic_gbs" % name] = nbytes / us * 1e-3
        res["k7_%s_frac_of_hbm_peak" % name] = nbytes / us * 1e-3 / peak
        res["k7_%s_tflops_fp32" % name] = flops / us * 1e-6
    res["k0_layout_pass_us"] = k0

    def stock(tf32):
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = tf32
        xr, wr, br = x.clone().requires_grad_(True), w4.clone().requires_grad_(True), b.clone().requires_grad_(True)
        gy = torch.randn(BV, Cout, H, W, device="cuda")
        f = timed(lambda: torch.nn.functional.conv2d(xr, wr, br), n=10, warm=3)

        def fb():
            xr.grad = wr.grad = br.grad = None
            torch.nn.functional.conv2d(xr, wr, br).backward(gy)
        return f, timed(fb, n=10, warm=3)
    f32 = stock(False)
    tf32 = stock(True)
    res.update({"stock_conv2d_fp32_fwd_us": f32[0], "stock_conv2d_fp32_fwd_bwd_us": f32[1],
                "stock_conv2d_tf32_fwd_us": tf32[0], "stock_conv2d_tf32_fwd_bwd_us": tf32[1],
                "k7_fwd_bwd_us": fwd + bwd_d + bwd_w})

    # through the public op, chained into unproject (no K0 either way) vs. dense conv + K0
    xr = x.clone().requires_grad_(True)
    wr, br = w4.clone().requires_grad_(True), b.clone().requires_grad_(True)

    def public():
        xr.grad = wr.grad = br.grad = None
        y = lt_b200.process_features(xr, wr, br)
        y.backward(y.detach())   # dense gradient: includes the K0 fallback of the backward
    res["k7_public_op_fwd_bwd_us_dense_upstream_gradient"] = timed(public, n=10, warm=3)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
