"""GPU: the public ops can be recorded by torch.cuda.graph (forward AND backward: they enqueue on torch's current stream,
allocate through the caching allocator and never synchronise) and a replay on new contents of the static input buffers
reproduces the eager results. This is what bench.py's e2e arm does every step."""
import pytest
import torch

import lt_b200
from lt_b200 import synthetic

pytestmark = pytest.mark.gpu


def test_volumetric_step_is_capturable_and_replays_on_new_inputs():
    d = [synthetic.make_unproject_inputs(B=2, V=4, C=32, H=32, W=32, N=16, seed=s, train=True) for s in (0, 1)]
    static = {k: d[0][k].cuda().clone() for k in ("heatmaps", "proj_matricies")}
    spec = lt_b200.volumetric.coord_volume_specs(d[0]["base_points"], 2500.0, 16, d[0]["thetas"]).cuda()
    x = torch.randn(2, 17, 16, 16, 16, device="cuda")
    gvol = torch.randn(2, 32, 16, 16, 16, device="cuda")
    out = {}

    def step():
        feat = static["heatmaps"].detach().requires_grad_(True)
        xs = x.detach().requires_grad_(True)
        coord = lt_b200.volumetric.coord_volumes_from_specs(spec, 16)
        vol = lt_b200.unproject_heatmaps(feat, static["proj_matricies"], coord, "softmax")
        kp, prob, ce = lt_b200.integrate_tensor_3d_with_coordinates_and_ce(
            xs, coord, torch.zeros(2, 17, 3, device="cuda"), torch.ones(2, 17, 1, device="cuda"))
        loss = kp.abs().mean() + 0.1 * ce
        torch.autograd.backward([loss, vol], [None, gvol])
        out.update(vol=vol.detach(), kp=kp.detach(), loss=loss.detach(), gfeat=feat.grad, gx=xs.grad)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    captured = dict(out)                       # tensors living in the graph's pool
    for i in (1, 0):
        static["heatmaps"].copy_(d[i]["heatmaps"])
        static["proj_matricies"].copy_(d[i]["proj_matricies"])
        spec.copy_(lt_b200.volumetric.coord_volume_specs(d[i]["base_points"], 2500.0, 16, d[i]["thetas"]))
        graph.replay()
        torch.cuda.synchronize()
        got = {k: v.clone() for k, v in captured.items()}
        step()                                  # eager, same inputs
        torch.cuda.synchronize()
        for k in ("vol", "kp", "loss", "gx"):
            assert torch.equal(got[k], out[k]), k
        # the scatter-add of K1's backward orders its atomics differently from run to run
        assert (got["gfeat"] - out["gfeat"]).abs().max().item() <= 1e-4 * out["gfeat"].abs().max().item()


def test_algebraic_tail_is_capturable():
    t = synthetic.make_dlt_inputs(B=4, V=4, J=17, seed=0)
    P, conf = t["proj_matricies"].cuda(), t["confidences"].cuda()
    hm = torch.randn(16, 17, 24, 24, device="cuda")
    out = {}

    def step():
        h = hm.detach().requires_grad_(True)
        kp, _ = lt_b200.integrate_tensor_2d(h, True, multiplier=100.0)
        X = lt_b200.triangulate_batch_of_points(P, kp.view(4, 4, 17, 2) * 16.0, conf)
        X.square().mean().backward()
        out.update(X=X.detach(), g=h.grad)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    captured = dict(out)
    hm.normal_()
    graph.replay()
    torch.cuda.synchronize()
    got = {k: v.clone() for k, v in captured.items()}
    step()
    torch.cuda.synchronize()
    assert torch.equal(got["X"], out["X"]) and torch.equal(got["g"], out["g"])


def test_process_features_into_unproject_is_capturable():
    """K7 -> .view -> K1 (no layout pass in either direction) recorded once, replayed on new backbone features: the tcgen05
    kernels (256 -> 32 channels) launch on the capturing stream, and the gather-layout registry keys on storage addresses,
    which a graph's private pool keeps stable."""
    B, V, N = 2, 4, 12
    d = synthetic.make_unproject_inputs(B=B, V=V, C=32, H=32, W=32, N=N, seed=4, train=True)
    proj, coord = d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    gen = torch.Generator().manual_seed(9)
    xs = [torch.randn(B * V, 256, 32, 32, generator=gen).cuda() for _ in range(2)]
    static_x = xs[0].clone()
    net = lt_b200.ProcessFeatures().cuda()
    gvol = torch.randn(B, 32, N, N, N, device="cuda")
    out = {}

    def step():
        x = static_x.detach().requires_grad_(True)
        for p in net.parameters():
            p.grad = None
        f = net(x)
        f = f.view(B, V, *f.shape[1:])
        vol = lt_b200.unproject_heatmaps(f, proj, coord, "softmax")
        vol.backward(gvol)
        out.update(vol=vol.detach(), gx=x.grad, gw=net[0].weight.grad, gb=net[0].bias.grad)

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    captured = dict(out)
    for i in (1, 0):
        static_x.copy_(xs[i])
        graph.replay()
        torch.cuda.synchronize()
        got = {k: v.clone() for k, v in captured.items()}
        step()
        torch.cuda.synchronize()
        assert torch.equal(got["vol"], out["vol"])
        for k in ("gx", "gw", "gb"):      # downstream of K1's atomically ordered scatter
            assert (got[k] - out[k]).abs().max().item() <= 1e-4 * out[k].abs().max().item(), k
