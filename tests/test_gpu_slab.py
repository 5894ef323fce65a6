"""GPU, >= 2 devices, NCCL: the z-slab sharded soft-argmax (lt_b200.slab, BASELINE.json config 5) with the CUDA kernels
on real ranks against the unsharded op on one device and the fp64 oracle.

One process per GPU (torch.multiprocessing spawn, rendezvous on 127.0.0.1). Every rank computes its slab; the only
forward collective is the all_gather of the (B,J,5) tuples, the only backward collective the all-reduce of B*J scalars
when a dense gradient arrives on the second output. Skipped on boxes with one GPU (run with `gpurun --gpus 2`).

Asserted: coordinates identical on every rank and within 1e-3 mm of the fp64 oracle; probability / gradient slabs
within 1e-5 / 1e-4 (max-norm relative) of the unsharded CUDA op and of the oracle. Maths: mvn/utils/op.py:84-96
evaluated blockwise."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _inputs(N, softmax):
    from lt_b200 import synthetic
    s = synthetic.make_softargmax_inputs(B=2, J=17, N=N, seed=4, peaked=True, train=True)
    g = torch.Generator().manual_seed(5)
    gc = torch.randn(2, 17, 3, generator=g)
    gp = torch.randn(s["volumes"].shape, generator=g) * 1e-3
    return s["volumes"], s["coord_volumes"], gc, gp


def _worker(rank, world, port, N, softmax, with_gp, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        import lt_b200  # noqa: F401
        from lt_b200 import slab
        vol_all, coord_all, gc, gp_all = _inputs(N, softmax)
        n0 = vol_all.shape[2]
        lo, hi = slab.slab_bounds(n0, world, rank)
        vol = vol_all[:, :, lo:hi].contiguous().to(dev).requires_grad_(True)
        coord = coord_all[:, lo:hi].contiguous().to(dev)
        coords, prob = slab.integrate_tensor_3d_with_coordinates_sharded(vol, coord, softmax=softmax)
        if with_gp:
            (gv,) = torch.autograd.grad([coords, prob], vol, [gc.to(dev), gp_all[:, :, lo:hi].contiguous().to(dev)])
        else:
            (gv,) = torch.autograd.grad([coords], vol, [gc.to(dev)])
        torch.cuda.synchronize()
        q.put((rank, lo, hi, coords.detach().cpu().numpy().copy(), prob.detach().cpu().numpy().copy(),
               gv.cpu().numpy().copy()))
    finally:
        dist.destroy_process_group()


def _cases():
    """(world, softmax, with_gp): the full cross on two ranks; three ranks (unequal slabs) and every GPU of the box with
    the softmax + dense-gradient case (each case spawns `world` processes: keep the 8-GPU bill small)."""
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    cases = [(2, sm, gp) for sm in (True, False) for gp in (False, True)]
    for w in sorted({3, n} - {2}):
        if 2 < w <= n:
            cases.append((w, True, True))
    return cases


@pytest.mark.parametrize("world,softmax,with_gp", _cases())
def test_sharded_softargmax_on_real_ranks(world, softmax, with_gp):
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d CUDA devices" % world)
    import lt_b200
    from oracle import ref_ops
    N = (37, 16, 24)  # slabs of unequal depth; 24 % 8 == 0 keeps the vector paths, 37 does not divide by any world
    vol_all, coord_all, gc, gp_all = _inputs(N, softmax)

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, N, softmax, with_gp, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=300) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0

    # unsharded op on one device + fp64 oracle
    dev = torch.device("cuda", 0)
    v1 = vol_all.to(dev).requires_grad_(True)
    c1, p1 = lt_b200.integrate_tensor_3d_with_coordinates(v1, coord_all.to(dev), softmax=softmax)
    outs, grads = ([c1, p1], [gc.to(dev), gp_all.to(dev)]) if with_gp else ([c1], [gc.to(dev)])
    (g1,) = torch.autograd.grad(outs, v1, grads)
    v64 = vol_all.double().requires_grad_(True)
    c64, p64 = ref_ops.integrate_tensor_3d_with_coordinates(v64, coord_all.double(), softmax=softmax)
    outs, grads = ([c64, p64], [gc.double(), gp_all.double()]) if with_gp else ([c64], [gc.double()])
    (g64,) = torch.autograd.grad(outs, v64, grads)

    tol_c = max(1e-3, 5e-7 * c64.abs().max().item())   # ReLU mode: un-normalised sums, relative fp32 accuracy
    first = torch.from_numpy(res[0][3])
    for rank, lo, hi, coords, prob, gv in res:
        coords, prob, gv = torch.from_numpy(coords), torch.from_numpy(prob), torch.from_numpy(gv)
        assert torch.equal(coords, first)                                        # global and identical on every rank
        assert (coords.double() - c64.detach()).abs().max().item() < tol_c
        assert (coords - c1.detach().cpu()).abs().max().item() < tol_c
        pmax, gmax = p64.abs().max().item(), g64.abs().max().item()
        assert (prob.double() - p64.detach()[:, :, lo:hi]).abs().max().item() <= 1e-5 * pmax
        assert (prob - p1.detach().cpu()[:, :, lo:hi]).abs().max().item() <= 1e-5 * pmax
        assert (gv.double() - g64[:, :, lo:hi]).abs().max().item() <= 1e-4 * gmax
        assert (gv - g1.cpu()[:, :, lo:hi]).abs().max().item() <= 1e-4 * gmax
