"""CPU, world_size 2 and 3 over gloo: the exchange logic of the z-slab sharded soft-argmax (lt_b200.slab) — slab
bounds, all_gather of the 5-float tuples, rank-ordered merge, global coordinates on every rank, backward with and
without the dense second-output gradient. The device steps are stood in for by the oracle's CPU restatement
(oracle.ref_ops.CpuSlabOps); on the GPU box the same driver runs with the CUDA kernels (tests/test_gpu_slab.py)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_golden


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, name, with_gp, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import lt_b200
        from lt_b200 import slab
        from oracle.ref_ops import CpuSlabOps
        g = load_golden(name)
        lo, hi = slab.slab_bounds(g["volumes"].shape[2], world, rank)
        vol = g["volumes"][:, :, lo:hi].clone().requires_grad_(True)
        coord = g["coord_volumes"][:, lo:hi]
        coords, prob = slab.integrate_tensor_3d_with_coordinates_sharded(
            vol, coord, softmax=bool(g["softmax"]), ops=CpuSlabOps)
        gp = g["grad_prob"][:, :, lo:hi] if with_gp else None
        outs, grads = ([coords, prob], [g["grad_coords"], gp]) if with_gp else ([coords], [g["grad_coords"]])
        (gv,) = torch.autograd.grad(outs, vol, grads)
        # plain numpy through the queue: torch tensors travel by fd-sharing, which races with worker exit
        q.put((rank, lo, hi, coords.detach().numpy().copy(), prob.detach().numpy().copy(), gv.numpy().copy()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("name", ["softargmax_small_softmax", "softargmax_small_relu"])
@pytest.mark.parametrize("with_gp", [False, True])
def test_sharded_softargmax_matches_unsharded_reference(world, name, with_gp):
    g = load_golden(name)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, name, with_gp, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    res = [(r, lo, hi, torch.from_numpy(c), torch.from_numpy(p_), torch.from_numpy(g_)) for r, lo, hi, c, p_, g_ in res]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tol = max(1e-3, 5e-7 * g["coords64"].abs().max().item())
    for rank, lo, hi, coords, prob, gv in res:
        assert (coords.double() - g["coords64"]).abs().max().item() < tol          # global, identical on every rank
        assert torch.equal(coords, res[0][3])
        ref_p = g["prob64"][:, :, lo:hi]
        assert (prob.double() - ref_p).abs().max().item() <= 1e-5 * g["prob64"].abs().max().item()
        ref_g = (g["grad_volumes_both64"] if with_gp else g["grad_volumes_coords_only64"])[:, :, lo:hi]
        assert (gv.double() - ref_g).abs().max().item() <= 1e-4 * ref_g.abs().max().item() + 1e-12
