#!/usr/bin/env python
"""Soak run for K1 (GPU box): many seeded random shapes and camera placements through the default kernels, the
TMA-staged forward (LT_B200_K1FWD=tma) and the on-chip merged backward (LT_B200_K1BWD=tile), against the oracle run with
CUDA tensors on the same device (the reference's own ATen calls). The pytest suite pins the known corners; this walks the
space the opt-in kernels' fallbacks depend on (boxes that do not exist, boxes larger than the ring, cameras inside the
cuboid, domes, ragged tiles) with several hundred cases. Prints one line per failure and a summary; exit code = failures.

The bar on the volume is BASELINE's 1e-5 (max-norm relative) wherever fp32 can hold it. With strongly non-square maps
(H = 96, W = 20: the reference normalises x by H and y by W, op.py:128-129) the reference's OWN fp32 runs sit 2e-5..6e-5
from its fp64 run and its CPU and CUDA runs 1.2e-5 apart (tools/diag/k1_precision.py); there the bar is the distance of
the fp32 oracle from the fp64 oracle (`cond`): a kernel passes if it is within max(1e-5, 1.25 x cond) of the fp32 oracle
and within cond + 1e-5 of the fp64 one.

    python tools/soak_k1.py [cases=300] [seed0=0]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lt_b200  # noqa: E402
from lt_b200 import synthetic  # noqa: E402
from oracle import ref_ops  # noqa: E402


def rel(a, b):
    den = b.abs().max().item()
    return (a.double() - b.double()).abs().max().item() / (den if den > 0 else 1.0)


def draw(seed):
    r = np.random.RandomState(7000 + seed)
    return dict(B=int(r.randint(1, 3)), V=int(r.choice([1, 2, 3, 4])), C=int(r.choice([17, 32])),
                H=int(r.choice([12, 24, 48, 96])), W=int(r.choice([12, 20, 48, 96])),
                N=(int(r.randint(1, 25)), int(r.randint(1, 25)), int(r.choice([8, 16, 24, 32]))),
                dome=bool(r.randint(0, 2)), inside_view=bool(r.randint(0, 3) == 0)), \
        str(r.choice(["sum", "conf", "max", "softmax"])), float(r.choice([1500.0, 2500.0, 6000.0]))


def run(d, method, conf, env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        hm = d["heatmaps"].cuda().requires_grad_(True)
        vc = d["vol_confidences"].cuda().requires_grad_(True) if conf else None
        out = lt_b200.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), method, vc)
        out.backward(d["gout"])
        return out.detach(), hm.grad, (vc.grad if conf else None)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 300
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    fails, hard, worst = 0, 0, [0.0, 0.0, 0.0, 0.0]
    for seed in range(seed0, seed0 + cases):
        cfg, method, side = draw(seed)
        conf = "conf" if method == "conf" else None
        d = synthetic.make_unproject_inputs(seed=seed, train=True, conf=conf, side=side, **cfg)
        hm = d["heatmaps"].cuda().requires_grad_(True)
        vc = d["vol_confidences"].cuda().requires_grad_(True) if conf else None
        ref = ref_ops.unproject_heatmaps(hm, d["proj_matricies"].cuda(), d["coord_volumes"].cuda(), method, vc)
        t64 = ref_ops.unproject_heatmaps(hm.detach().double(), d["proj_matricies"].cuda().double(),
                                         d["coord_volumes"].cuda().double(), method, vc.detach().double() if conf else None)
        cond = rel(ref.detach(), t64)              # how far fp32 arithmetic itself is from the truth on this case
        bar = max(1e-5, 1.25 * cond)
        d["gout"] = torch.randn(ref.shape, generator=torch.Generator().manual_seed(seed)).cuda()
        ref.backward(d["gout"])
        base = run(d, method, conf, {})
        tma = run(d, method, conf, {"LT_B200_K1FWD": "tma"})
        tile = run(d, method, conf, {"LT_B200_K1BWD": "tile", "LT_B200_K1BWD_TILE": str(seed % 4)})
        e = [rel(base[0], ref.detach()), rel(base[1], hm.grad), rel(tma[0], ref.detach()), rel(tile[1], hm.grad)]
        if conf:
            e[1] = max(e[1], rel(base[2], vc.grad))
            e[3] = max(e[3], rel(tile[2], vc.grad))
        worst = [max(a, b) for a, b in zip(worst, e)]
        hard = hard + (1 if cond > 8e-6 else 0)
        bad = (e[0] > bar or e[1] > 1e-4 or e[2] > bar or e[3] > 1e-4 or rel(tma[0], base[0]) > 2e-6 or
               rel(base[0], t64) > cond + 1e-5 or rel(tma[0], t64) > cond + 1e-5)
        if bad:
            fails += 1
            print("FAIL seed %d %s %s side=%g: fwd %.2e bwd %.2e tma-fwd %.2e tile-bwd %.2e tma-vs-default %.2e"
                  % (seed, cfg, method, side, e[0], e[1], e[2], e[3], rel(tma[0], base[0])))
    print("soak: %d cases (%d of them ill-conditioned in fp32: the fp32 oracle itself > 8e-6 from the fp64 oracle), %d "
          "failures; worst relative errors vs the fp32 oracle: default fwd %.2e, default bwd %.2e (bar 1e-4), tma fwd %.2e, "
          "tile bwd %.2e (1e-4)" % (cases, hard, fails, *worst))
    return fails


if __name__ == "__main__":
    sys.exit(min(main(), 100))
