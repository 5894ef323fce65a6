"""GPU: the torch extension over the C ABI (torch.ops.lt_b200.*, torch_ext/lt_b200_torch.cpp): the three differentiable
ops against the ctypes binding of the same kernels (forward bit-identical, gradients to 1e-4 where reductions are
unordered, bit-identical otherwise), schema + fake-tensor registration of the raw ops (torch.library.opcheck), and the
host-side cost of a public call."""
import time

import pytest
import torch

import lt_b200
from conftest import rel_err
from lt_b200 import synthetic, torch_ops
from lt_b200.multiview import _DLT
from lt_b200.op import _SoftArgmax, _Unproject

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch_ops.available(), reason="torch extension not built (build() does it)")]


@pytest.mark.parametrize("method,agg", [("sum", 0), ("conf", 1), ("max", 2), ("softmax", 3)])
def test_unproject_op_equals_ctypes_binding(method, agg):
    d = synthetic.make_unproject_inputs(B=2, V=4, C=32, H=24, W=24, N=16, seed=3, train=True, conf="conf",
                                        inside_view=True)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    gout = torch.randn(2, 32, 16, 16, 16, generator=torch.Generator().manual_seed(4)).cuda()
    res = []
    for use_ext in (True, False):
        hm = d["heatmaps"].cuda().requires_grad_(True)
        vc = d["vol_confidences"].cuda().requires_grad_(True) if agg == 1 else None
        if use_ext:
            out = torch.ops.lt_b200.unproject_heatmaps(hm, proj, coord, agg, vc, False, [], False)
        else:
            out = _Unproject.apply(hm, proj, coord, vc, agg, 0)
        out.backward(gout)
        res.append((out.detach(), hm.grad, vc.grad if vc is not None else None))
    assert torch.equal(res[0][0], res[1][0])
    assert rel_err(res[0][1], res[1][1]) < 1e-4
    if agg == 1:
        assert rel_err(res[0][2], res[1][2]) < 1e-4
    # and it is what the public function dispatches to
    out = lt_b200.unproject_heatmaps(d["heatmaps"].cuda(), proj, coord, method, d["vol_confidences"].cuda())
    assert torch.equal(out, res[0][0])
    node = lt_b200.unproject_heatmaps(d["heatmaps"].cuda().requires_grad_(True), proj, coord, method,
                                      d["vol_confidences"].cuda()).grad_fn
    assert "UnprojectFn" in node.name()          # the C++ autograd node of the extension, not a python Function


@pytest.mark.parametrize("softmax", [True, False])
def test_softargmax_op_equals_ctypes_binding(softmax):
    s = synthetic.make_softargmax_inputs(B=2, J=17, N=16, seed=1, peaked=True, train=True)
    coord = s["coord_volumes"].cuda()
    g = torch.Generator().manual_seed(2)
    gc, gp = torch.randn(2, 17, 3, generator=g).cuda(), (torch.randn(2, 17, 16, 16, 16, generator=g) * 1e-3).cuda()
    res = []
    for use_ext in (True, False):
        v = s["volumes"].cuda().requires_grad_(True)
        if use_ext:
            coords, prob = torch.ops.lt_b200.integrate_tensor_3d_with_coordinates(v, coord, softmax, 1.0, False, False)
        else:
            coords, prob = _SoftArgmax.apply(v, coord, 1 if softmax else 0, 1.0)
        (gv,) = torch.autograd.grad([coords, prob], v, [gc, gp])
        (gv_c,) = torch.autograd.grad([_SoftArgmax.apply(v, coord, 1 if softmax else 0, 1.0)[0]], v, [gc]) \
            if not use_ext else torch.autograd.grad(
                [torch.ops.lt_b200.integrate_tensor_3d_with_coordinates(v, coord, softmax, 1.0, False, False)[0]], v, [gc])
        res.append((coords.detach(), prob.detach(), gv, gv_c))
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a, b)          # K2 is deterministic: the two bindings launch the same kernels


def test_dlt_op_equals_ctypes_binding_and_is_cheap_to_call():
    t = synthetic.make_dlt_inputs(B=64, V=4, J=17, seed=0)
    proj, conf = t["proj_matricies"].cuda(), t["confidences"].cuda()
    gx = torch.randn(64, 17, 3, generator=torch.Generator().manual_seed(5)).cuda()
    res = []
    for fn in (torch.ops.lt_b200.triangulate_batch_of_points, _DLT.apply):
        pts = t["points"].cuda().requires_grad_(True)
        cf = conf.clone().requires_grad_(True)
        x = fn(proj, pts, cf)
        x.backward(gx)
        res.append((x.detach(), pts.grad, cf.grad))
    for a, b in zip(res[0], res[1]):
        assert torch.equal(a, b)
    # host-side cost of the public forward call (python checks + dispatcher + allocation + launch), no sync inside
    pts = t["points"].cuda()
    for _ in range(20):
        lt_b200.triangulate_batch_of_points(proj, pts, conf)
    torch.cuda.synchronize()
    n = 200
    t0 = time.perf_counter()
    for _ in range(n):
        lt_b200.triangulate_batch_of_points(proj, pts, conf)
    torch.cuda.synchronize()
    us = (time.perf_counter() - t0) / n * 1e6
    print("triangulate_batch_of_points public call, cfg2 shape: %.1f us" % us)
    assert us < 60.0, us    # 70 us with the ctypes binding in round 1; bar of the round-2 review: <= 35 us


def test_raw_ops_schema_and_fake_tensors():
    d = synthetic.make_unproject_inputs(B=1, V=2, C=32, H=16, W=16, N=8, seed=0)
    hm, proj, coord = d["heatmaps"].cuda(), d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    feat = torch.ops.lt_b200.feat_to_hwc(hm.view(2, 32, 16, 16))
    tests = ("test_schema", "test_faketensor")
    torch.library.opcheck(torch.ops.lt_b200.feat_to_hwc.default, (hm.view(2, 32, 16, 16),), test_utils=tests)
    args = (feat, proj, coord, None, 1, 2, 32, 16, 16, 8, 8, 8, 3, False, False, False)
    torch.library.opcheck(torch.ops.lt_b200.unproject_fwd.default, args, test_utils=tests)
    vol = torch.ops.lt_b200.unproject_fwd(*args)
    torch.library.opcheck(torch.ops.lt_b200.unproject_bwd.default,
                          (torch.randn_like(vol), feat, proj, coord, None, 1, 2, 32, 16, 16, 8, 8, 8, 3, False, False,
                           False, False), test_utils=tests)
    x = torch.randn(1, 17, 8, 8, 8, device="cuda")
    torch.library.opcheck(torch.ops.lt_b200.softargmax3d_fwd.default, (x, coord, True, 1.0, False, False), test_utils=tests)
    t = synthetic.make_dlt_inputs(B=2, V=4, J=17, seed=0)
    torch.library.opcheck(torch.ops.lt_b200.dlt_fwd.default,
                          (t["proj_matricies"].cuda(), t["points"].cuda(), t["confidences"].cuda()), test_utils=tests)


def test_cpu_tensors_are_refused_with_the_librarys_message():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        torch.ops.lt_b200.triangulate_batch_of_points(torch.zeros(1, 2, 3, 4), torch.zeros(1, 2, 3, 2), None)
