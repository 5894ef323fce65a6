"""Host->device ceiling of the e2e arm, N ranks at once: every rank copies the SAME pinned buffer
bench.py's e2e arm copies per step (cfg3: 75.5 MB) to its GPU, nothing else runs. One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 \
        tools/microbench/h2d_ranks.py [--mb 75.5] [--iters 200]

Rank 0 prints one JSON line: per-rank and aggregate GB/s (CUDA events on the copy stream, barrier on both sides, max
over ranks), for one cudaMemcpyAsync per step and for the four separate copies round 1's arm issued."""
import argparse
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=75.5)
    ap.add_argument("--iters", type=int, default=200)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = int(args.mb * 1e6 / 4)
    host = torch.empty(n, dtype=torch.float32).pin_memory()
    host.normal_()
    dst = [torch.empty(n, device=dev) for _ in range(3)]
    stream = torch.cuda.Stream(device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(pieces):
        cuts = [n * i // pieces for i in range(pieces + 1)]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            for i in range(5):
                dst[i % 3].copy_(host, non_blocking=True)
        barrier()
        with torch.cuda.stream(stream):
            a.record(stream)
            for i in range(args.iters):
                d = dst[i % 3]
                for p in range(pieces):
                    d[cuts[p]:cuts[p + 1]].copy_(host[cuts[p]:cuts[p + 1]], non_blocking=True)
            b.record(stream)
        barrier()
        ms = torch.tensor([a.elapsed_time(b)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return n * 4 * args.iters / (ms.item() * 1e-3) / 1e9

    one, four = run(1), run(4)
    if rank == 0:
        print(json.dumps({"ranks": world, "bytes_per_copy": n * 4, "iters": args.iters,
                          "one_memcpy_per_step": {"gbs_per_rank": round(one, 2), "gbs_aggregate": round(one * world, 1)},
                          "four_memcpys_per_step": {"gbs_per_rank": round(four, 2), "gbs_aggregate": round(four * world, 1)},
                          "steps_per_s_ceiling_per_rank": round(one * 1e9 / (n * 4), 1),
                          "volumes_per_s_ceiling_all_ranks_cfg3": round(one * 1e9 / (n * 4) * 16 * world, 0)}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
