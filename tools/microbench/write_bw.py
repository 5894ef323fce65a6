"""How fast does this B200 absorb a write-only stream (what K7's backward-data kernel is: 604 MB written, 79 MB read)?
fill_ / zero_ of the (64, 256, 96, 96) fp32 tensor, a read-only reduction and a copy, CUDA events. One JSON line."""
import json
import torch


def timed(fn, n=20):
    for _ in range(3):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


x = torch.empty(64, 256, 96, 96, device="cuda")
y = torch.empty_like(x)
nb = x.numel() * 4
res = {"bytes": nb}
for name, fn, mult in (("fill", lambda: x.fill_(1.0), 1), ("zero", lambda: x.zero_(), 1), ("read_sum", lambda: x.sum(), 1),
                       ("copy", lambda: y.copy_(x), 2)):
    us = timed(fn)
    res[name + "_us"] = round(us, 1)
    res[name + "_gbs"] = round(mult * nb / us * 1e-3, 1)
print(json.dumps(res))
