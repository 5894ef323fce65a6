#!/usr/bin/env python
"""Measurement for the SURVEY.md 8(f) rows built after the headline path, at BASELINE config 3's shape (B=16, J=17,
64^3): K6 (coordinate volumes), K5 (VolumetricCELoss) as a drop-in and fused behind the soft-argmax, each next to the
reference's own statements run in stock ATen ops on the same GPU (the oracle restatement with CUDA tensors: tools/ may
use it as the checker/baseline, the product never does). CUDA events on the launching stream; one JSON line."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lt_b200  # noqa: E402
from lt_b200 import synthetic  # noqa: E402
from oracle import ref_ops  # noqa: E402


def timed(fn, n=50, warm=5):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3  # us


def main():
    B, J, N = 16, 17, 64
    peak = 6545.6
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    rng = np.random.RandomState(0)
    base = rng.normal(0.0, 100.0, size=(B, 3))
    thetas = rng.uniform(0.0, 2 * np.pi, size=B)
    spec = lt_b200.volumetric.coord_volume_specs(base, 2500.0, N, thetas).cuda()
    L, st = lt_b200._lib.lib(), torch.cuda.current_stream().cuda_stream
    out6 = torch.empty(B, N, N, N, 3, device="cuda")
    k6_us = timed(lambda: L.lt_coord_volumes(spec.data_ptr(), out6.data_ptr(), B, N, N, N, 0, st), n=200)  # C ABI
    k6_public_us = timed(lambda: lt_b200.build_coord_volumes(base, 2500.0, N, thetas), n=20)

    def stock_coord():
        return torch.stack([ref_ops.build_coord_volume(base[b], 2500.0, N, thetas[b], [0, 0, 1], False, "cuda")
                            for b in range(B)])
    stock_k6_us = timed(stock_coord, n=5, warm=2)
    coord = lt_b200.volumetric.coord_volumes_from_specs(spec, N)
    k6_bytes = coord.numel() * 4

    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=1, train=True)
    vol = d["volumes"].cuda()
    kp0, prob = lt_b200.integrate_tensor_3d_with_coordinates(vol, coord)
    gen = torch.Generator().manual_seed(2)
    kp_gt = (kp0.cpu() + 60.0 * torch.randn(B, J, 3, generator=gen)).cuda()
    val = torch.ones(B, J, 1, device="cuda")

    k5_fwd_us = timed(lambda: lt_b200.volumetric_ce_loss(coord, prob, kp_gt, val))
    nvox = N ** 3
    loss1, idx = torch.empty(1, device="cuda"), torch.empty(B, J, dtype=torch.int32, device="cuda")
    ws = torch.empty(L.lt_volce_ws_bytes(B, J, nvox), dtype=torch.uint8, device="cuda")
    val2 = val.view(B, J).contiguous()
    k5_raw_us = timed(lambda: L.lt_volce_fwd(coord.data_ptr(), prob.data_ptr(), kp_gt.data_ptr(), val2.data_ptr(),
                                             loss1.data_ptr(), idx.data_ptr(), ws.data_ptr(), B, J, nvox, st), n=200)
    g1, gdense = torch.ones(1, device="cuda"), torch.empty(B, J, nvox, device="cuda")
    vals, dots = torch.empty(B, J, device="cuda"), torch.empty(B, J, device="cuda")
    k5_bwd_dense_us = timed(lambda: L.lt_volce_bwd(prob.data_ptr(), idx.data_ptr(), val2.data_ptr(), g1.data_ptr(),
                                                   gdense.data_ptr(), B, J, nvox, st), n=200)
    k5_bwd_sparse_us = timed(lambda: L.lt_volce_bwd_sparse(prob.data_ptr(), idx.data_ptr(), val2.data_ptr(),
                                                           g1.data_ptr(), vals.data_ptr(), dots.data_ptr(), B, J, nvox,
                                                           st), n=200)
    stock_k5_fwd_us = timed(lambda: ref_ops.volumetric_ce_loss(coord, prob, kp_gt, val), n=3, warm=1)
    ours = lt_b200.volumetric_ce_loss(coord, prob, kp_gt, val).item()
    ref = ref_ops.volumetric_ce_loss(coord, prob, kp_gt, val).item()

    gcoords = torch.randn(B, J, 3, device="cuda")
    gce = torch.tensor(0.01, device="cuda")   # made once: a per-step host scalar would add a pageable H2D copy

    def chained():   # drop-in: two autograd nodes, dense (B,J,N^3) gradient between them
        v = vol.detach().requires_grad_(True)
        kp, p = lt_b200.integrate_tensor_3d_with_coordinates(v, coord)
        ce = lt_b200.volumetric_ce_loss(coord, p, kp_gt, val)
        torch.autograd.backward([kp, ce], [gcoords, gce])

    def fused():     # one node, the loss gradient stays B*J numbers
        v = vol.detach().requires_grad_(True)
        kp, p, ce = lt_b200.integrate_tensor_3d_with_coordinates_and_ce(v, coord, kp_gt, val)
        torch.autograd.backward([kp, ce], [gcoords, gce])

    def no_ce():     # soft-argmax forward+backward alone, for scale
        v = vol.detach().requires_grad_(True)
        kp, p = lt_b200.integrate_tensor_3d_with_coordinates(v, coord)
        torch.autograd.backward([kp], [gcoords])

    def stock_chained():
        v = vol.detach().requires_grad_(True)
        kp, p = ref_ops.integrate_tensor_3d_with_coordinates(v, coord, softmax=True)
        ce = ref_ops.volumetric_ce_loss(coord, p, kp_gt, val)
        torch.autograd.backward([kp, ce], [gcoords, gce])

    print(json.dumps({
        "workload": "cfg3 shape: B=16, J=17, 64^3",
        "k6_coord_volumes_us_c_abi": k6_us, "k6_bytes": k6_bytes, "k6_gbs": k6_bytes / k6_us * 1e-3,
        "k6_frac_of_hbm_peak": k6_bytes / k6_us * 1e-3 / peak, "k6_public_call_us_incl_host_spec": k6_public_us,
        "coord_volumes_stock_aten_python_loop_us": stock_k6_us,
        "k5_ce_forward_us_c_abi": k5_raw_us, "k5_ce_forward_us_public_op": k5_fwd_us,
        "k5_ce_backward_dense_us_c_abi": k5_bwd_dense_us, "k5_ce_backward_sparse_us_c_abi": k5_bwd_sparse_us,
        "ce_forward_stock_aten_us": stock_k5_fwd_us,
        "ce_value_ours": ours, "ce_value_stock": ref,
        "softargmax_fwd_bwd_us": timed(no_ce, n=20),
        "softargmax_plus_ce_fwd_bwd_us_drop_in_dense_gradient": timed(chained, n=20),
        "softargmax_plus_ce_fwd_bwd_us_fused_sparse_gradient": timed(fused, n=20),
        "softargmax_plus_ce_fwd_bwd_us_stock_aten": timed(stock_chained, n=3, warm=1),
    }))


if __name__ == "__main__":
    main()
