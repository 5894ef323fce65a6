"""GPU: every output pointer handed to the C ABI is carved out of a larger buffer whose margins hold a sentinel; after
the launch the margins must be untouched (no out-of-bounds write at ragged sizes: nvox % 4 != 0, N2 % 8 != 0, W % 4 != 0,
C not a multiple of the lane width). compute-sanitizer is not available on this pool; this is the bounds check the
kernels get instead, on the shapes where vector paths switch to scalar tails."""
import pytest
import torch

import lt_b200
from lt_b200 import _lib, synthetic

pytestmark = pytest.mark.gpu

GUARD = 1024          # floats on each side
SENTINEL = -7.25e30


class Guarded:
    """`n` floats of output in the middle of a sentinel-filled buffer; 256-byte aligned like a torch allocation."""

    def __init__(self, n, dtype=torch.float32, fill=None):
        self.n = n
        self.buf = torch.full((n + 2 * GUARD,), SENTINEL, device="cuda", dtype=torch.float32)
        self.mid = self.buf[GUARD:GUARD + n]
        if fill is not None:
            self.mid.fill_(fill)
        self.dtype = dtype

    def ptr(self):
        return self.mid.data_ptr()

    def view(self, *shape):
        return (self.mid.view(torch.int32) if self.dtype == torch.int32 else self.mid).view(*shape)

    def check(self, name):
        torch.cuda.synchronize()
        lo, hi = self.buf[:GUARD], self.buf[GUARD + self.n:]
        assert bool((lo == SENTINEL).all()) and bool((hi == SENTINEL).all()), "out-of-bounds write around " + name
        if self.n:
            assert not bool((self.mid == SENTINEL).any()), name + " was not fully written"


def _st():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("C,H,W,N,V,agg", [(32, 96, 96, (8, 8, 8), 4, 3), (17, 13, 31, (5, 6, 7), 3, 1),
                                           (40, 10, 12, (3, 9, 10), 5, 3), (130, 8, 9, (4, 4, 6), 2, 2),
                                           (32, 16, 16, (8, 8, 12), 4, 0)])
def test_layout_and_unproject(C, H, W, N, V, agg):
    L = _lib.lib()
    B = 2
    d = synthetic.make_unproject_inputs(B=B, V=V, C=C, H=H, W=W, N=N, seed=2, train=True, conf="conf")
    hm, proj, coord = d["heatmaps"].cuda(), d["proj_matricies"].cuda(), d["coord_volumes"].cuda()
    conf = d["vol_confidences"].cuda()
    nvox = N[0] * N[1] * N[2]
    feat = Guarded(L.lt_feat_layout_elems(B * V, C, H, W))
    _lib.check(L.lt_feat_to_hwc(hm.data_ptr(), feat.ptr(), B * V, C, H, W, _st()))
    feat.check("hwcp features")
    out = Guarded(B * C * nvox)
    _lib.check(L.lt_unproject_fwd(feat.ptr(), proj.data_ptr(), coord.data_ptr(), conf.data_ptr(), out.ptr(), B, V, C, H,
                                  W, N[0], N[1], N[2], agg, 0, _st()))
    out.check("unprojected volume")
    ref = lt_b200.unproject_heatmaps(hm, proj, coord, ["sum", "conf", "max", "softmax"][agg], conf)
    assert torch.equal(out.view(B, C, *N), ref)          # the guarded call is the same computation as the public op
    gout = torch.randn(B, C, *N, device="cuda")
    gfeat = Guarded(feat.n, fill=0.0)
    gconf = Guarded(B * V * C)
    ws = torch.empty((max(L.lt_unproject_bwd_ws_bytes(B, V, C, *N), 4),), device="cuda", dtype=torch.uint8)
    _lib.check(L.lt_unproject_bwd(gout.data_ptr(), feat.ptr(), proj.data_ptr(), coord.data_ptr(), conf.data_ptr(),
                                  gfeat.ptr(), gconf.ptr(), ws.data_ptr(), B, V, C, H, W, N[0], N[1], N[2], agg, 0, _st()))
    gfeat.check("hwcp gradient")
    if agg == 1:
        gconf.check("grad_conf")
    ghm = Guarded(hm.numel())
    _lib.check(L.lt_feat_to_chw(gfeat.ptr(), ghm.ptr(), B * V, C, H, W, _st()))
    ghm.check("grad_heatmaps")


@pytest.mark.parametrize("N,softmax", [((8, 8, 8), 1), ((5, 6, 7), 1), ((3, 9, 11), 0), ((16, 16, 17), 1)])
def test_soft_argmax_3d_and_ce(N, softmax):
    L = _lib.lib()
    B, J = 2, 5
    d = synthetic.make_softargmax_inputs(B=B, J=J, N=N, seed=3, train=True)
    vol, coord = d["volumes"].cuda(), d["coord_volumes"].cuda()
    nvox = N[0] * N[1] * N[2]
    coords, prob, stats = Guarded(B * J * 3), Guarded(B * J * nvox), Guarded(B * J * 5)
    ws = torch.empty((max(L.lt_softargmax_ws_bytes(B, J, nvox), L.lt_volce_ws_bytes(B, J, nvox)),), device="cuda",
                     dtype=torch.uint8)
    _lib.check(L.lt_softargmax_fwd(vol.data_ptr(), coord.data_ptr(), coords.ptr(), prob.ptr(), stats.ptr(), ws.data_ptr(),
                                   B, J, nvox, softmax, 1.0, _st()))
    for g, name in ((coords, "coordinates"), (prob, "probabilities"), (stats, "statistics")):
        g.check(name)
    gc, gp = torch.randn(B, J, 3, device="cuda"), torch.randn(B, J, nvox, device="cuda")
    dots = Guarded(B * J)
    _lib.check(L.lt_softargmax_bwd_dot(prob.ptr(), gp.data_ptr(), dots.ptr(), ws.data_ptr(), B, J, nvox, _st()))
    dots.check("dots")
    gvol = Guarded(B * J * nvox)
    _lib.check(L.lt_softargmax_bwd(prob.ptr(), coord.data_ptr(), coords.ptr(), gc.data_ptr(), gp.data_ptr(), dots.ptr(),
                                   gvol.ptr(), B, J, nvox, softmax, 1.0, _st()))
    gvol.check("grad_volumes")
    # K5 on the same probabilities
    kp = coords.view(B, J, 3).clone()
    val = torch.ones(B, J, device="cuda")
    p_in = prob.view(B, J, nvox).abs() + 1e-3
    loss, idx = Guarded(1), Guarded(B * J, dtype=torch.int32)
    _lib.check(L.lt_volce_fwd(coord.data_ptr(), p_in.data_ptr(), kp.data_ptr(), val.data_ptr(), loss.ptr(), idx.ptr(),
                                ws.data_ptr(), B, J, nvox, _st()))
    loss.check("loss")
    idx.check("nearest voxels")
    assert int(idx.view(B, J).min()) >= 0 and int(idx.view(B, J).max()) < nvox
    gl = torch.ones(1, device="cuda")
    gprob, vals, sd = Guarded(B * J * nvox), Guarded(B * J), Guarded(B * J)
    _lib.check(L.lt_volce_bwd(p_in.data_ptr(), idx.ptr(), val.data_ptr(), gl.data_ptr(), gprob.ptr(), B, J, nvox, _st()))
    gprob.check("loss gradient")
    _lib.check(L.lt_volce_bwd_sparse(p_in.data_ptr(), idx.ptr(), val.data_ptr(), gl.data_ptr(), vals.ptr(), sd.ptr(), B,
                                     J, nvox, _st()))
    vals.check("sparse values")
    sd.check("sparse dots")
    _lib.check(L.lt_softargmax_bwd_add_sparse(p_in.data_ptr(), idx.ptr(), vals.ptr(), gvol.ptr(), B, J, nvox, softmax,
                                              1.0, _st()))
    gvol.check("grad_volumes after the sparse add")


@pytest.mark.parametrize("H,W,softmax", [(96, 96, 1), (13, 31, 1), (7, 12, 0), (1, 1, 1), (224, 224, 1), (5, 4, 0)])
def test_soft_argmax_2d(H, W, softmax):
    L = _lib.lib()
    BJ = 7
    hm = torch.randn(BJ, H, W, device="cuda")
    coords, prob = Guarded(BJ * 2), Guarded(BJ * H * W)
    _lib.check(L.lt_softargmax2d_fwd(hm.data_ptr(), coords.ptr(), prob.ptr(), BJ, H, W, softmax, 3.0, _st()))
    coords.check("2-D coordinates")
    prob.check("2-D maps")
    gc, gp = torch.randn(BJ, 2, device="cuda"), torch.randn(BJ, H, W, device="cuda")
    for gp_ptr in (None, gp.data_ptr()):
        ghm = Guarded(BJ * H * W)
        _lib.check(L.lt_softargmax2d_bwd(prob.ptr(), coords.ptr(), gc.data_ptr(), gp_ptr, ghm.ptr(), BJ, H, W, softmax,
                                         3.0, _st()))
        ghm.check("grad_heatmaps (2-D)")


@pytest.mark.parametrize("N,transfer", [((8, 8, 8), 0), ((5, 6, 7), 0), ((3, 7, 7), 1), ((64, 64, 64), 1)])
def test_coord_volumes(N, transfer):
    L = _lib.lib()
    spec = lt_b200.volumetric.coord_volume_specs([[1.0, 2.0, 3.0], [-40.0, 5.0, 900.0]], 2500.0, N, [0.0, 1.1]).cuda()
    out = Guarded(2 * N[0] * N[1] * N[2] * 3)
    _lib.check(L.lt_coord_volumes(spec.data_ptr(), out.ptr(), 2, N[0], N[1], N[2], transfer, _st()))
    out.check("coordinate volumes")


@pytest.mark.parametrize("V,J,with_conf", [(4, 17, True), (2, 3, False), (31, 5, True)])
def test_dlt(V, J, with_conf):
    L = _lib.lib()
    B = 3
    t = synthetic.make_dlt_inputs(B=B, V=V, J=J, seed=1, with_conf=with_conf, dome=V > 8)
    P, pts = t["proj_matricies"].cuda(), t["points"].cuda()
    conf = t["confidences"].cuda() if with_conf else None
    out = Guarded(B * J * 3)
    saved = torch.empty(B, J, 20, device="cuda", dtype=torch.float64)
    _lib.check(L.lt_dlt_fwd(P.data_ptr(), pts.data_ptr(), _lib.ptr(conf), out.ptr(), saved.data_ptr(), B, V, J, _st()))
    out.check("triangulated points")
    g = torch.randn(B, J, 3, device="cuda")
    gpts, gconf = Guarded(B * V * J * 2), Guarded(B * V * J)
    _lib.check(L.lt_dlt_bwd(g.data_ptr(), P.data_ptr(), pts.data_ptr(), _lib.ptr(conf), saved.data_ptr(), gpts.ptr(),
                            gconf.ptr() if with_conf else None, B, V, J, _st()))
    gpts.check("grad_points")
    if with_conf:
        gconf.check("grad_confidences")
