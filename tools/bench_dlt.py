#!/usr/bin/env python
"""K3 (batched DLT, BASELINE config 2: B=64, V=4, J=17 -> 1088 solves): us per launch and joints/s for forward and
forward+backward through the public op, parity against the oracle in fp64, and the oracle's own CPU time beside it.
Latency-bound (77 KB of I/O): no roofline fraction is claimed for it."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lt_b200  # noqa: E402
from lt_b200 import synthetic  # noqa: E402
from oracle import ref_ops  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    t = synthetic.make_dlt_inputs(B=64, V=4, J=17, seed=0)
    P, pts, conf = (t[k].to(dev) for k in ("proj_matricies", "points", "confidences"))
    ref64 = ref_ops.triangulate_batch_of_points(t["proj_matricies"].double(), t["points"].double(),
                                                t["confidences"].double(), out_dtype=None)
    ref32 = ref_ops.triangulate_batch_of_points(t["proj_matricies"], t["points"], t["confidences"])
    out = lt_b200.triangulate_batch_of_points(P, pts, conf)
    err64 = (out.double().cpu() - ref64).abs().max().item()
    err32 = (ref32.double() - ref64).abs().max().item()

    def timed(fn, n=200):
        for _ in range(20):
            fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(n):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / n * 1e3  # us

    fwd_us = timed(lambda: lt_b200.triangulate_batch_of_points(P, pts, conf))
    pg, cg = pts.clone().requires_grad_(True), conf.clone().requires_grad_(True)
    g = torch.randn(64, 17, 3, device=dev)

    def fb():
        o = lt_b200.triangulate_batch_of_points(P, pg, cg)
        torch.autograd.grad(o, (pg, cg), g)

    fb_us = timed(fb)
    t0 = time.perf_counter()
    for _ in range(3):
        ref_ops.triangulate_batch_of_points(t["proj_matricies"], t["points"], t["confidences"])
    cpu_us = (time.perf_counter() - t0) / 3 * 1e6
    print(json.dumps({"workload": "cfg2: DLT B=64 V=4 J=17 (1088 solves)", "fwd_us_per_call": fwd_us,
                      "fwd_bwd_us_per_call": fb_us, "joints_per_s_fwd": 1088 / fwd_us * 1e6,
                      "max_err_mm_vs_fp64_reference": err64, "fp32_reference_err_mm_vs_fp64": err32,
                      "oracle_cpu_batched_svd_us": cpu_us, "cpu_threads": torch.get_num_threads(),
                      "note": "public python op incl. torch allocation + ctypes launch overhead; kernel itself is ~us"}))


if __name__ == "__main__":
    main()
