#!/usr/bin/env python
"""K4 + K3: the geometric tail of AlgebraicTriangulationNet (mvn/models/triangulation.py:164-188) at BASELINE config 2
(B=64, V=4, J=17, 96x96 heat maps): 2-D soft-argmax with heatmap_multiplier=100 -> image pixels -> weighted DLT.
Prints one JSON line: us per launch of K4 forward / backward (CUDA events on the launching stream; the 160 MB of maps
exceed the 126 MB L2), their HBM roofline fraction (algorithmic bytes: one read + one write of
the maps per direction), the whole tail forward+backward through the public ops, and the same tail in stock ATen ops
on the same GPU beside it."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lt_b200  # noqa: E402
from lt_b200 import synthetic  # noqa: E402


def hbm_peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6548.5, "fallback"


def timed(fn, n=100, warm=10):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3  # us


def stock_integrate_2d(hm, mult):
    # the ATen op sequence of mvn/utils/op.py:11-47 (softmax branch), written out with torch calls
    b, j, h, w = hm.shape
    p = torch.nn.functional.softmax((hm * mult).reshape(b, j, -1), dim=2).reshape(b, j, h, w)
    mx, my = p.sum(dim=2), p.sum(dim=3)
    x = (mx * torch.arange(w, device=hm.device, dtype=hm.dtype)).sum(dim=2, keepdim=True)
    y = (my * torch.arange(h, device=hm.device, dtype=hm.dtype)).sum(dim=2, keepdim=True)
    return torch.cat((x, y), dim=2), p


def main():
    dev = torch.device("cuda:0")
    B, V, J, H, W, mult = 64, 4, 17, 96, 96, 100.0
    t = synthetic.make_dlt_inputs(B=B, V=V, J=J, seed=0)
    P, conf = t["proj_matricies"].to(dev), t["confidences"].to(dev)
    gen = torch.Generator().manual_seed(0)
    hm = torch.randn(B * V, J, H, W, generator=gen).to(dev)
    lib = lt_b200._lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    coords = torch.empty(B * V, J, 2, device=dev)
    prob = torch.empty_like(hm)
    gh = torch.empty_like(hm)
    gc = torch.randn(B * V, J, 2, device=dev)
    BJ = B * V * J

    def k4_fwd():
        rc = lib.lt_softargmax2d_fwd(hm.data_ptr(), coords.data_ptr(), prob.data_ptr(), BJ, H, W, 1, mult, st)
        assert rc == 0

    def k4_bwd():
        rc = lib.lt_softargmax2d_bwd(prob.data_ptr(), coords.data_ptr(), gc.data_ptr(), None, gh.data_ptr(), BJ, H, W,
                                     1, mult, st)
        assert rc == 0

    fwd_us, bwd_us = timed(k4_fwd), timed(k4_bwd)
    nbytes = 2 * hm.numel() * 4
    peak, src = hbm_peak_gbs()

    scale = 4.0  # heat-map pixels -> image pixels (384 / 96, triangulation.py:181-184)

    def tail(int2d, dlt, hm_in):
        kp, _ = int2d(hm_in)
        kp = kp.view(B, V, J, 2) * scale
        return dlt(P, kp, conf)

    hm_g = hm.clone().requires_grad_(True)
    g3 = torch.randn(B, J, 3, device=dev)

    def ours():
        out = tail(lambda x: lt_b200.integrate_tensor_2d(x, True, multiplier=mult),
                   lt_b200.triangulate_batch_of_points, hm_g)
        torch.autograd.grad(out, hm_g, g3)

    def stock_front():  # stock ATen soft-argmax forward+backward only (the reference DLT is a python loop of 1088 svds)
        kp, _ = stock_integrate_2d(hm_g, mult)
        torch.autograd.grad(kp, hm_g, gc)

    def ours_front():
        kp, _ = lt_b200.integrate_tensor_2d(hm_g, True, multiplier=mult)
        torch.autograd.grad(kp, hm_g, gc)

    c_stock, _ = stock_integrate_2d(hm, mult)
    k4_fwd()
    err_px = (coords - c_stock).abs().max().item()
    print(json.dumps({
        "workload": "cfg2 algebraic tail: B=64 V=4 J=17 96x96 heat maps, heatmap_multiplier=100",
        "k4_fwd_us": fwd_us, "k4_bwd_us": bwd_us, "k4_algorithmic_bytes_per_direction": nbytes,
        "k4_fwd_gbs": nbytes / fwd_us * 1e-3, "k4_bwd_gbs": nbytes / bwd_us * 1e-3,
        "hbm_peak_gbs": peak, "hbm_peak_source": src,
        "k4_fwd_frac": nbytes / fwd_us * 1e-3 / peak, "k4_bwd_frac": nbytes / bwd_us * 1e-3 / peak,
        "tail_fwd_bwd_us_public_ops": timed(ours, n=50),
        "softargmax2d_fwd_bwd_us_public_op": timed(ours_front, n=50),
        "softargmax2d_fwd_bwd_us_stock_aten_same_gpu": timed(stock_front, n=50),
        "max_coord_diff_px_vs_stock_aten": err_px,
    }))


if __name__ == "__main__":
    main()
