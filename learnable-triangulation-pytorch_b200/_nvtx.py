"""NVTX ranges around the public ops (the reference has no tracing hooks at all).

Off by default and free when off: `annotate` returns the function unchanged unless LT_B200_NVTX=1 was set before the
package was imported. With it, every public op pushes a range named `lt_b200.<op>` on the calling thread, so that
Nsight Systems / `ncu --nvtx` group the K0..K7 launches (and the allocator / autograd work around them) per reference
call site. The dispatcher ops of the torch extension (`torch.ops.lt_b200.*`, forward AND the C++ autograd backward
nodes) are additionally labelled by `torch.autograd.profiler.emit_nvtx()` like any ATen op, with no help from here."""
import functools
import os

import torch

ENABLED = os.environ.get("LT_B200_NVTX", "0") not in ("", "0")


def annotate(name):
    def deco(fn):
        if not ENABLED:
            return fn

        @functools.wraps(fn)
        def wrapped(*args, **kwargs):
            torch.cuda.nvtx.range_push(name)
            try:
                return fn(*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapped
    return deco
