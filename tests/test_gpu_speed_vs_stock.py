"""GPU: the fused path against the reference algorithm run in stock ATen ops ON THE SAME B200 ("also time
the reference on the B200 as the stock-torch-on-same-GPU comparison"). The reference tree does not travel to the GPU
box, so the stock arm is the oracle restatement (oracle/ref_ops.py: the same torch calls in the same order) with CUDA
tensors. One step = K1 forward + K2 forward + both backwards at the BASELINE config-3 item shape (4 views, 32 channels,
96x96 maps, 64^3 voxels, 17 joints). Asserts parity of the step's outputs and that the fused path is faster; the
measured ratio is written to gpurun_out/stock_torch_gpu.json when that directory exists."""
import json
import os

import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic
from oracle import ref_ops

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _timed(fn, n, warm=2):
    for _ in range(warm):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def test_step_is_faster_than_stock_torch_on_the_same_gpu():
    B = 4
    d = synthetic.make_unproject_inputs(B=B, V=4, C=32, H=96, W=96, N=64, seed=0, train=True)
    s = synthetic.make_softargmax_inputs(B=B, J=17, N=64, seed=1, train=True)
    feat = d["heatmaps"].cuda().requires_grad_(True)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    vol = s["volumes"].cuda().requires_grad_(True)
    gout = torch.randn(B, 32, 64, 64, 64, generator=torch.Generator().manual_seed(5)).cuda()
    gcoords = torch.randn(B, 17, 3, generator=torch.Generator().manual_seed(6)).cuda()

    def step(unproject, integrate):
        out = unproject(feat, proj, coord, "softmax")
        (gf,) = torch.autograd.grad(out, feat, gout)
        kp, _ = integrate(vol, coord, softmax=True)
        (gv,) = torch.autograd.grad(kp, vol, gcoords)
        return out, gf, kp, gv

    ours = step(lt_b200.unproject_heatmaps, lt_b200.integrate_tensor_3d_with_coordinates)
    stock = step(ref_ops.unproject_heatmaps, ref_ops.integrate_tensor_3d_with_coordinates)
    assert rel_err(ours[0], stock[0]) < 1e-5 and rel_err(ours[1], stock[1]) < 1e-4
    # stock fp32 (a bmm over 262144 terms) is itself ~3e-3 mm away from fp64; the 1e-3 mm bar is checked against the
    # fp64 oracle in test_gpu_softargmax.py, here the two fp32 results only have to agree to fp32-bmm accuracy
    assert (ours[2] - stock[2]).abs().max().item() < 2e-2 and rel_err(ours[3], stock[3]) < 1e-4
    del ours, stock

    ms_ours = _timed(lambda: step(lt_b200.unproject_heatmaps, lt_b200.integrate_tensor_3d_with_coordinates), 20)
    ms_stock = _timed(lambda: step(ref_ops.unproject_heatmaps, ref_ops.integrate_tensor_3d_with_coordinates), 5)
    rec = {"workload": "cfg3 item shape, B=4: K1 fwd+bwd (softmax aggregation) + K2 fwd+bwd through the public ops",
           "volumes_per_s_fused": B / ms_ours * 1e3, "volumes_per_s_stock_aten_same_gpu": B / ms_stock * 1e3,
           "ms_per_step_fused": ms_ours, "ms_per_step_stock_aten": ms_stock, "speedup": ms_stock / ms_ours}
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        with open(os.path.join(out_dir, "stock_torch_gpu.json"), "w") as f:
            json.dump(rec, f)
    print(json.dumps(rec))
    assert ms_ours < ms_stock, rec
