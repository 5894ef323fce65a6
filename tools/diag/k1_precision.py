#!/usr/bin/env python
"""Where does a > 1e-5 difference between K1 forward and the oracle-on-CUDA come from? For the soak cases that exceeded
the bar: the fp64 CPU run of the reference's arithmetic is the truth; printed are the distances to it of (a) the fp32 CPU
ATen run (what the golden vectors pin), (b) the fp32 CUDA ATen run (what the soak compared with), (c) our kernel, and the
pairwise distances. max-norm relative to max |truth|."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import lt_b200
from lt_b200 import synthetic
from oracle import ref_ops
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from soak_k1 import draw

def rel(a, b, den):
    return (a.double().cpu() - b.double().cpu()).abs().max().item() / den

for seed in [int(s) for s in sys.argv[1:]] or [7, 72, 103, 106, 185, 248]:
    cfg, method, side = draw(seed)
    conf = "conf" if method == "conf" else None
    d = synthetic.make_unproject_inputs(seed=seed, train=True, conf=conf, side=side, **cfg)
    hm, P, cv, vc = d["heatmaps"], d["proj_matricies"], d["coord_volumes"], d["vol_confidences"] if conf else None
    t64 = ref_ops.unproject_heatmaps(hm.double(), P.double(), cv.double(), method, vc.double() if conf else None)
    cpu = ref_ops.unproject_heatmaps(hm, P, cv, method, vc)
    gpu = ref_ops.unproject_heatmaps(hm.cuda(), P.cuda(), cv.cuda(), method, vc.cuda() if conf else None)
    with torch.no_grad():
        ours = lt_b200.unproject_heatmaps(hm.cuda(), P.cuda(), cv.cuda(), method, vc.cuda() if conf else None)
    den = t64.abs().max().item()
    print("seed %d %s H=%d W=%d: vs fp64: cpu-aten %.2e  cuda-aten %.2e  ours %.2e | ours-cpu %.2e  ours-cuda %.2e  cpu-cuda %.2e"
          % (seed, method, cfg["H"], cfg["W"], rel(cpu, t64, den), rel(gpu, t64, den), rel(ours, t64, den),
             rel(ours, cpu, den), rel(ours, gpu, den), rel(cpu, gpu, den)))
