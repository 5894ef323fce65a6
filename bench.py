#!/usr/bin/env python
"""bench.py — volumes/s, forward+backward, of the fused unproject+aggregate (K0+K1) and volumetric soft-argmax (K2)
path on synthetic Human3.6M-shaped data (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg1|cfg4|cfg5] [--impl ours|reference]

One "step" = one pass of the hot path over one batch: K0 (layout) + K1 fwd + K2 fwd + K2 bwd + K1 bwd + K0 inverse.
`value`  : device-resident arm, direct C-ABI launches on preallocated buffers, CUDA events, max over ranks.
`e2e`    : the public python API (lt_b200.unproject_heatmaps / integrate_tensor_3d_with_coordinates under autograd) fed
           from PINNED HOST tensors every step (H2D inside the timed region: feature maps, projection matrices and the
           cuboid specs; the voxel coordinates are evaluated from the specs inside the kernels, lt_b200.CoordSpec) with the loss +
           coordinates read back.
`roofline`: per-kernel CUDA-event durations taken inside the same timed loop against MEASURED_PEAKS.json.
`cpu_baseline`: the oracle port of the reference's torch CPU path, timed on this box's host cores (rank 0, N=1).
--impl reference prints the same line for the CPU reference arm (oracle port; the reference is pure python and its
tree does not travel to the GPU box).
"""
import argparse
import json
import os
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# ceilings of the two access patterns K1 is made of, measured in isolation on this pool's B200
# (tools/microbench/red_paths.cu, profiles/r1c_microbench_red_paths.txt): fp32 reduction payload the L2 retires per
# second (the same whether issued as red.global.add.v4.f32 or as TMA bulk reduce), and LDG.128 tap-gather bytes.
L2_RED_PEAK_GBS = 6352.0
L1_GATHER_PEAK_GBS = 18537.0

WORKLOADS = {
    # name: B per GPU (weak) or total (strong), V, C, H, W, N, J, aggregation, scaling
    "cfg3": dict(B=16, V=4, C=32, H=96, W=96, N=64, J=17, agg="softmax", scaling="weak",
                 desc="volumetric training step: B=16/GPU, 4 views, 32ch 96x96, 64^3, softmax aggregation, fwd+bwd"),
    "cfg1": dict(B=1, V=4, C=32, H=96, W=96, N=64, J=17, agg="softmax", scaling="weak",
                 desc="H36M eval shape: B=1, 4 views, 32ch 96x96, 64^3, softmax aggregation, fwd+bwd"),
    "cfg4": dict(B=8, V=31, C=32, H=96, W=96, N=64, J=17, agg="conf", scaling="strong",
                 desc="many-view stress: B=8 total (batch-sharded), 31 views, conf aggregation, fwd+bwd"),
    "cfg5": dict(B=1, V=4, C=32, H=96, W=96, N=256, J=17, agg="softmax", scaling="strong",
                 desc="high-resolution volume: B=1, 4 views, 256^3 split into z-slabs along the first volume axis; K1 fwd+bwd slab-local, K2 fwd with ONE NCCL all_gather of the (B,J,5) soft-argmax tuples at world > 1, K2 bwd slab-local"),
}


def algorithmic_bytes(w, B):
    """SURVEY.md §8d / BASELINE.md §3 per-kernel compulsory HBM bytes for B batch items."""
    F = w["V"] * w["C"] * w["H"] * w["W"] * 4
    n0 = w.get("N0", w["N"])
    nvox = n0 * w["N"] * w["N"]
    K, O, S = nvox * 12, w["C"] * nvox * 4, w["J"] * nvox * 4
    return {
        "k0_to_hwc": B * 2 * F,
        "k1_unproject_fwd": B * (F + K + O),
        "k2_softargmax_fwd": B * (S + K + S),
        "k2_softargmax_bwd": B * (S + K + S),
        "k1_unproject_bwd": B * (O + K + (0 if w["agg"] == "sum" else F) + F),
        "k0_to_chw": B * 2 * F,
    }


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons of this rank's GPU, sampled through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40,
                     "hw_power_brake_slowdown": 0x80, "sw_power_cap": 0x4}
            while True:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    bits = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.reasons.update(k for k, m in names.items() if bits & m)
                if self._stop_evt.wait(0.02):
                    break
        except Exception as e:  # NVML missing: report that instead of inventing clocks
            self.reasons.add("nvml_unavailable:%s" % type(e).__name__)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ---------------------------------------------------------------------------------------------------------------------
# device-resident arm: direct C-ABI launches on preallocated buffers
# ---------------------------------------------------------------------------------------------------------------------
STAGES = ["k0_to_hwc", "k1_unproject_fwd", "k2_softargmax_fwd", "k2_softargmax_bwd", "k1_unproject_bwd", "k0_to_chw"]


def launches_per_step(w, B, nvox):
    """Kernels of ours enqueued by one step (K2 forward = partial + combine + finalize)."""
    return {"k0_to_hwc": 1, "k1_unproject_fwd": 1, "k2_softargmax_fwd": 3, "k2_softargmax_bwd": 1,
            "k1_unproject_bwd": 1, "k0_to_chw": 1}


class DeviceStep:
    """Preallocated buffers + one method that enqueues the whole step (optionally with an event after each stage)."""

    def __init__(self, w, B, dev, seed, slab_world=1):
        import lt_b200
        self.slab_world = slab_world
        from lt_b200 import _lib, synthetic
        self.L, self.lib, self.w, self.B, self.dev = _lib.lib(), _lib, w, B, dev
        conf = "conf" if w["agg"] == "conf" else None
        N = w["N"] if "N0" not in w else (w["N0"], w["N"], w["N"])
        d = synthetic.make_unproject_inputs(B=B, V=w["V"], C=w["C"], H=w["H"], W=w["W"], N=N, seed=seed, train=True,
                                            conf=conf, dome=w["V"] > 8)
        g = torch.Generator().manual_seed(seed + 5)
        self.host = d
        self.feat = d["heatmaps"].to(dev)
        self.proj = d["proj_matricies"].to(dev)
        self.coord = d["coord_volumes"].to(dev)
        self.conf = d["vol_confidences"].to(dev) if conf else None
        self.N = tuple(self.coord.shape[1:4])
        nvox = self.N[0] * self.N[1] * self.N[2]
        self.nvox = nvox
        self.feat_hwc = torch.empty((self.L.lt_feat_layout_elems(B * w["V"], w["C"], w["H"], w["W"]),), device=dev)
        self.vol = torch.empty((B, w["C"]) + self.N, device=dev)
        gd = torch.Generator(device=dev).manual_seed(seed + 6)
        self.gvol = torch.randn((B, w["C"]) + self.N, generator=gd, device=dev)     # upstream grad of K1's output
        self.x = torch.randn((B, w["J"]) + self.N, generator=gd, device=dev)        # V2V output stand-in, N(0,1)
        self.coords = torch.empty((B, w["J"], 3), device=dev)
        self.prob = torch.empty_like(self.x)
        self.stats = torch.empty((B, w["J"], 5), device=dev)
        self.parts = torch.empty((slab_world, B, w["J"], 5), device=dev)
        self.gcoords = torch.randn((B, w["J"], 3), generator=g).to(dev)
        self.gx = torch.empty_like(self.x)
        self.gfeat_hwc = torch.empty_like(self.feat_hwc)
        self.gfeat = torch.empty_like(self.feat)
        self.gconf = torch.empty_like(self.conf) if conf else None
        self.Cp = self.L.lt_padded_channels(w["C"])
        ws_bytes = max(self.L.lt_softargmax_ws_bytes(B, w["J"], nvox),
                       self.L.lt_unproject_bwd_ws_bytes(B, w["V"], w["C"], *self.N), 4)
        self.ws = torch.empty((ws_bytes // 4,), device=dev)
        self.agg = {"sum": 0, "conf": 1, "max": 2, "softmax": 3}[w["agg"]]

    def valid_voxel_views(self):
        try:
            return _valid_voxel_views(self.coord, self.proj, self.w["H"], self.w["W"])
        except Exception:
            return None

    def run(self, events=None):
        L, p, chk, w, B = self.L, self.lib.ptr, self.lib.check, self.w, self.B
        st = torch.cuda.current_stream().cuda_stream
        N0, N1, N2 = self.N

        def mark(i):
            if events is not None:
                events[i].record()

        mark(0)
        chk(L.lt_feat_to_hwc(p(self.feat), p(self.feat_hwc), B * w["V"], w["C"], w["H"], w["W"], st))
        mark(1)
        chk(L.lt_unproject_fwd(p(self.feat_hwc), p(self.proj), p(self.coord), p(self.conf), p(self.vol), B, w["V"],
                               w["C"], w["H"], w["W"], N0, N1, N2, self.agg, 0, st))
        mark(2)
        if self.slab_world > 1:
            # z-slab sharded soft-argmax (cfg5): slab-local tuples, ONE small all_gather over NCCL, fixed-order merge
            import torch.distributed as dist
            chk(L.lt_softargmax_partial(p(self.x), p(self.coord), p(self.stats), p(self.ws), B, w["J"], self.nvox, 1,
                                        1.0, st))
            dist.all_gather_into_tensor(self.parts.view(-1), self.stats.view(-1))
            chk(L.lt_softargmax_merge(p(self.parts), p(self.stats), self.slab_world, B * w["J"], 1, st))
            chk(L.lt_softargmax_finalize(p(self.x), p(self.stats), p(self.coords), p(self.prob), B, w["J"], self.nvox,
                                         1, 1.0, st))
        else:
            chk(L.lt_softargmax_fwd(p(self.x), p(self.coord), p(self.coords), p(self.prob), p(self.stats), p(self.ws),
                                    B, w["J"], self.nvox, 1, 1.0, st))
        mark(3)
        chk(L.lt_softargmax_bwd(p(self.prob), p(self.coord), p(self.coords), p(self.gcoords), None, None, p(self.gx),
                                B, w["J"], self.nvox, 1, 1.0, st))
        mark(4)
        self.gfeat_hwc.zero_()
        chk(L.lt_unproject_bwd(p(self.gvol), p(self.feat_hwc), p(self.proj), p(self.coord), p(self.conf),
                               p(self.gfeat_hwc), p(self.gconf), p(self.ws), B, w["V"], w["C"], w["H"], w["W"], N0, N1,
                               N2, self.agg, 0, st))
        mark(5)
        chk(L.lt_feat_to_chw(p(self.gfeat_hwc), p(self.gfeat), B * w["V"], w["C"], w["H"], w["W"], st))
        mark(6)


def _valid_voxel_views(coord, proj, H, W):
    """Number of (voxel, view) pairs with at least one bilinear tap inside the image: op.py:117-130 + the unnormalise
    step of grid_sample (align_corners=False), evaluated with torch ops once at set-up (not timed)."""
    B = coord.shape[0]
    total = 0
    for b in range(B):
        pts = torch.cat([coord[b].reshape(-1, 3), torch.ones(coord[b].numel() // 3, 1, device=coord.device)], dim=1)
        for v in range(proj.shape[1]):
            p = pts @ proj[b, v].t()
            z = p[:, 2]
            ix = ((2.0 * (p[:, 0] / z / H - 0.5) + 1.0) * W - 1.0) * 0.5
            iy = ((2.0 * (p[:, 1] / z / W - 0.5) + 1.0) * H - 1.0) * 0.5
            total += int(((z > 0) & (ix > -1) & (ix < W) & (iy > -1) & (iy < H)).sum().item())
    return total


class E2EStep:
    """Public-API arm: pinned host inputs -> H2D -> unproject (+ its backward) and soft-argmax
    -> L1 loss (+ its backward) -> loss + coordinates back to the host. The same work as one step of the device arm.

    The host inputs are what the reference's forward receives from its data loader: feature maps, projection
    matrices and the cuboid of every item (base point, side, rotation: 18 floats). Like the reference
    (triangulation.py:280-341 builds `coord_volumes` on the device), nothing per voxel is shipped over PCIe: the ops get
    a lt_b200.CoordSpec and evaluate the voxel centres in-kernel (bit-identical to building the tensor with K6).

    Every step copies ITS OWN inputs from pinned host memory inside the timed region. Like a DataLoader with
    pin_memory + non_blocking prefetch, the copy of step i+1 is issued on a second stream while step i computes
    (two device input sets, events in both directions); the ops themselves run on torch's current stream through
    lt_b200.unproject_heatmaps / lt_b200.integrate_tensor_3d_with_coordinates under autograd."""

    DEPTH = 3   # device input sets: the one being computed on + two prefetched (prefetch depth 2)

    def __init__(self, dstep):
        h, w = dstep.host, dstep.w
        self.w, self.dev = w, dstep.dev
        self.x, self.gvol = dstep.x, dstep.gvol
        from lt_b200 import volumetric
        self.vol_size = tuple(h["coord_volumes"].shape[1:4])   # (N0/ranks, N, N) for the slab-sharded cfg5
        specs = volumetric.coord_volume_specs(h["base_points"], 2500.0, self.vol_size, h["thetas"])
        host = {"feat": h["heatmaps"], "proj": h["proj_matricies"], "spec": specs,
                "target": torch.zeros(dstep.B, w["J"], 3)}
        if h["vol_confidences"] is not None:
            host["conf"] = h["vol_confidences"]
        # ONE pinned staging buffer and ONE device buffer per input set: a step's inputs travel as a single
        # cudaMemcpyAsync (every tensor starts on a 256-byte boundary inside it); the named tensors are views.
        offs, total = {}, 0
        for k, v in host.items():
            offs[k] = total
            total += (v.numel() + 63) // 64 * 64
        self.h_flat = torch.empty(total, dtype=torch.float32).pin_memory()
        self.host = {}
        for k, v in host.items():
            self.host[k] = self.h_flat[offs[k]:offs[k] + v.numel()].view(v.shape)
            self.host[k].copy_(v)
        self.d_flat = [torch.empty(total, device=self.dev) for _ in range(self.DEPTH)]
        self.dbuf = [{k: f[offs[k]:offs[k] + v.numel()].view(v.shape) for k, v in host.items()} for f in self.d_flat]
        self.copied = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.consumed = [torch.cuda.Event() for _ in range(self.DEPTH)]
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self.h_loss = torch.zeros(1).pin_memory()
        self.h_coords = torch.zeros(dstep.B, w["J"], 3).pin_memory()
        self.h2d = total * 4
        self.d2h = (self.h_loss.numel() + self.h_coords.numel()) * 4
        self.issued = 0   # copies issued so far
        self.done = 0     # steps computed so far
        self.graphs = None

    def _issue_copy(self):
        k = self.issued % self.DEPTH
        with torch.cuda.stream(self.copy_stream):
            if self.issued >= self.DEPTH:
                self.copy_stream.wait_event(self.consumed[k])   # the step that used this set has finished
            self.d_flat[k].copy_(self.h_flat, non_blocking=True)
            self.copied[k].record(self.copy_stream)
        self.issued += 1

    def drain(self):
        """Called between warm-up and the timed region: drop the prefetched set, so that every timed step's copy
        (its own, and the prefetch of its successor) is issued inside the timed region."""
        torch.cuda.synchronize()
        self.done = self.issued

    def _compute(self, k):
        """The step on input set k: the user-facing calls, nothing else."""
        import lt_b200
        w, d = self.w, self.dbuf[k]
        feat = d["feat"].detach().requires_grad_(True)
        # the cuboids travel as 18 floats per item; K1 / K2 evaluate the voxel centres from them in-kernel (lt_b200.CoordSpec,
        # bit-identical to building the reference's coord_volumes tensor with K6 and reading it back four times)
        coord = lt_b200.CoordSpec(d["spec"], self.vol_size)
        vol = lt_b200.unproject_heatmaps(feat, d["proj"], coord, w["agg"], d.get("conf"))
        # The V2V net between the two ops stays in PyTorch and is out of scope: as in the device arm and the CPU arm,
        # its output (x) and the gradient it sends back into the unprojected volume (gvol) are device-resident tensors.
        x = self.x.detach().requires_grad_(True)
        coords, _ = lt_b200.integrate_tensor_3d_with_coordinates(x, coord, softmax=True)
        loss = (coords - d["target"]).abs().mean()
        torch.autograd.backward([loss, vol], [None, self.gvol])
        self.h_loss.copy_(loss.detach().view(1), non_blocking=True)
        self.h_coords.copy_(coords.detach(), non_blocking=True)
        return feat.grad

    def capture(self):
        """Capture the step once per input set into a CUDA graph (the ops enqueue on torch's current stream, allocate
        through torch's caching allocator and never synchronise, so `torch.cuda.graph` can record them, autograd
        included). Falls back to eager launches when capture is refused; `self.graphs` says which one ran."""
        self.graphs = None
        try:
            cur = torch.cuda.current_stream()
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(cur)
            with torch.cuda.stream(side):          # warm-up off the default stream, as torch.cuda.graph requires
                for k in range(self.DEPTH):
                    self.d_flat[k].copy_(self.h_flat, non_blocking=True)
                    self._compute(k)
            cur.wait_stream(side)
            torch.cuda.synchronize()
            graphs = []
            for k in range(self.DEPTH):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    self._compute(k)
                graphs.append(g)
            torch.cuda.synchronize()
            self.graphs = graphs
        except Exception as e:  # noqa: BLE001
            sys.stderr.write("e2e: CUDA-graph capture refused (%s: %s); running the ops eagerly\n" % (type(e).__name__, e))
            torch.cuda.synchronize()
            self.graphs = None
        return self.graphs is not None

    def run(self):
        while self.issued < self.done + self.DEPTH:   # this step's own copy (first step) + the prefetches behind it
            self._issue_copy()
        k = self.done % self.DEPTH
        cur = torch.cuda.current_stream()
        cur.wait_event(self.copied[k])
        if self.graphs is not None:
            self.graphs[k].replay()
        else:
            self._compute(k)
        self.consumed[k].record(cur)
        self.done += 1


# ---------------------------------------------------------------------------------------------------------------------
# CPU reference arm (oracle port of the reference's torch CPU path): one volume (B=1) per step
# ---------------------------------------------------------------------------------------------------------------------
def bench_process_features(step, w, B, peak, iters=20):
    """K7 at this workload's shape: (B*V, 256, H, W) planar features -> the gather layout K1 reads (and back): CUDA events
    around `iters` launches of each entry point on torch's current stream; 256 input channels as in the reference."""
    L = step.L
    BV, C, H, W, Cin = B * w["V"], w["C"], w["H"], w["W"], 256
    dev = step.feat.device
    x = torch.randn(BV, Cin, H, W, device=dev)
    wt = torch.randn(C, Cin, device=dev) / 16.0
    bias = torch.randn(C, device=dev)
    n_hwcp = L.lt_feat_layout_elems(BV, C, H, W)
    out, g = torch.empty(n_hwcp, device=dev), torch.randn(n_hwcp, device=dev)
    gx, gw, gb = torch.empty_like(x), torch.empty_like(wt), torch.empty_like(bias)
    ws = torch.empty(max(L.lt_pwconv_ws_bytes(BV, Cin, C, H, W), 4) // 4, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    calls = {
        "fwd": lambda: L.lt_pwconv_fwd(x.data_ptr(), wt.data_ptr(), bias.data_ptr(), out.data_ptr(), ws.data_ptr(), BV, Cin,
                                       C, H, W, st),
        "bwd_data": lambda: L.lt_pwconv_bwd_data(g.data_ptr(), wt.data_ptr(), gx.data_ptr(), BV, Cin, C, H, W, st),
        "bwd_weight": lambda: L.lt_pwconv_bwd_weight(x.data_ptr(), g.data_ptr(), gw.data_ptr(), gb.data_ptr(), ws.data_ptr(),
                                                     BV, Cin, C, H, W, st),
    }
    nbytes = (x.numel() + n_hwcp) * 4
    res = {"shape": "B*V=%d maps, %d -> %d channels, %dx%d" % (BV, Cin, C, H, W), "algorithmic_bytes": nbytes,
           "flops": 2.0 * BV * H * W * Cin * C, "kernels": {}}
    for name, fn in calls.items():
        for _ in range(3):
            rc = fn()
            if rc != 0:
                raise RuntimeError("%s returned %d" % (name, rc))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(iters):
            fn()
        b.record(); torch.cuda.synchronize()
        us = a.elapsed_time(b) / iters * 1e3
        res["kernels"][name] = {"us": round(us, 1), "gbs": round(nbytes / us * 1e-3, 1),
                                "frac": round(nbytes / us * 1e-3 / peak, 4)}
    return res


def bench_in_kernel_coordinates(step, w, B, peak, iters=30):
    """SURVEY 8(f) row 1 beside the path: the four kernels of the step with the voxel centres evaluated in-kernel from
    the 18-float specs (lt_*_spec entry points) instead of read from the (B,N,N,N,3) tensor. Same launches as the device
    arm otherwise; CUDA events around `iters` launches each. Algorithmic bytes drop by K = nvox*12 per item and kernel."""
    from lt_b200 import volumetric
    L, p = step.L, step.lib.ptr
    h = step.host
    N0, N1, N2 = step.N
    spec = volumetric.coord_volume_specs(h["base_points"], 2500.0, step.N, h["thetas"]).to(step.dev)
    st = torch.cuda.current_stream().cuda_stream
    V, C, H, W, J = w["V"], w["C"], w["H"], w["W"], w["J"]
    calls = {
        "k1_unproject_fwd": lambda: L.lt_unproject_fwd_spec(p(step.feat_hwc), p(step.proj), p(spec), 0, p(step.conf),
                                                            p(step.vol), B, V, C, H, W, N0, N1, N2, step.agg, 0, st),
        "k2_softargmax_fwd": lambda: L.lt_softargmax_fwd_spec(p(step.x), p(spec), 0, p(step.coords), p(step.prob),
                                                              p(step.stats), p(step.ws), B, J, N0, N1, N2, 1, 1.0, st),
        "k2_softargmax_bwd": lambda: L.lt_softargmax_bwd_spec(p(step.prob), p(spec), 0, p(step.coords), p(step.gcoords),
                                                              None, None, p(step.gx), B, J, N0, N1, N2, 1, 1.0, st),
        "k1_unproject_bwd": lambda: L.lt_unproject_bwd_spec(p(step.gvol), p(step.feat_hwc), p(step.proj), p(spec), 0,
                                                            p(step.conf), p(step.gfeat_hwc), p(step.gconf), p(step.ws),
                                                            B, V, C, H, W, N0, N1, N2, step.agg, 0, st),
    }
    ab = algorithmic_bytes(w, B)
    K = B * step.nvox * 12
    res = {"note": "coordinates computed in-kernel (lt_*_spec); algorithmic bytes = the tensor variant's minus K", "kernels": {}}
    for name, fn in calls.items():
        for _ in range(3):
            rc = fn()
            if rc != 0:
                raise RuntimeError("%s returned %d" % (name, rc))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(iters):
            fn()
        b.record(); torch.cuda.synchronize()
        ms = a.elapsed_time(b) / iters
        nbytes = ab[name] - K
        res["kernels"][name] = {"ms": round(ms, 4), "algorithmic_bytes": nbytes, "gbs": round(nbytes / ms * 1e-6, 1),
                                "frac": round(nbytes / ms * 1e-6 / peak, 4)}
    return res


def bench_sharded_softargmax(rank, world, dev, steps=30, N=256, J=17):
    """BASELINE config 5's soft-argmax across the ranks of this launch, the one path of the product with a collective:
    every rank holds a slab of ONE (1, J, N, N, N) volume (split along the first volume axis), computes its (m, S, Sxyz)
    tuples (lt_softargmax_partial), exchanges them with ONE all_gather_into_tensor (NCCL, B*J*5 floats per rank), merges
    in rank order (lt_softargmax_merge) and normalises its slab (lt_softargmax_finalize). Timed with CUDA events on the
    launching stream, max over ranks; the exchange has its own event pair. Parity: the coordinates against the
    unsharded lt_softargmax_fwd of the same volume on this GPU (max |diff| in mm; bar 1e-3), and the slab of
    probabilities against the unsharded ones (max-norm relative; bar 1e-5). Reported beside the path, never in `value`."""
    import torch.distributed as dist
    from lt_b200 import _lib, slab, volumetric
    L, p, chk = _lib.lib(), _lib.ptr, _lib.check
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator(device=dev).manual_seed(1234)              # identical on every rank
    x = torch.randn((1, J, N, N, N), generator=g, device=dev)
    spec = volumetric.coord_volume_specs([[10.0, -20.0, 30.0]], 2500.0, (N, N, N), [0.7])
    coord = volumetric.coord_volumes_from_specs(spec.to(dev), (N, N, N))
    lo, hi = slab.slab_bounds(N, world, rank)
    xs, cs = x[:, :, lo:hi].contiguous(), coord[:, lo:hi].contiguous()
    nvox_s, nvox = (hi - lo) * N * N, N * N * N
    stats, merged = torch.empty((1, J, 5), device=dev), torch.empty((1, J, 5), device=dev)
    parts = torch.empty((world, 1, J, 5), device=dev)
    coords, prob = torch.empty((1, J, 3), device=dev), torch.empty_like(xs)
    ws = torch.empty((max(L.lt_softargmax_ws_bytes(1, J, nvox), 4) // 4,), device=dev)

    def step(ev=None):
        chk(L.lt_softargmax_partial(p(xs), p(cs), p(stats), p(ws), 1, J, nvox_s, 1, 1.0, st))
        if ev is not None:
            ev[0].record()
        dist.all_gather_into_tensor(parts.view(-1), stats.view(-1))
        if ev is not None:
            ev[1].record()
        chk(L.lt_softargmax_merge(p(parts), p(merged), world, J, 1, st))
        chk(L.lt_softargmax_finalize(p(xs), p(merged), p(coords), p(prob), 1, J, nvox_s, 1, 1.0, st))

    for _ in range(3):
        step()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0.record()
    for i in range(steps):
        step(evs[i])
    t1.record()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([t0.elapsed_time(t1), sum(a.elapsed_time(b) for a, b in evs)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # parity against the unsharded op on this GPU
    c_full, p_full, s_full = torch.empty((1, J, 3), device=dev), torch.empty_like(x), torch.empty((1, J, 5), device=dev)
    chk(L.lt_softargmax_fwd(p(x), p(coord), p(c_full), p(p_full), p(s_full), p(ws), 1, J, nvox, 1, 1.0, st))
    err = torch.stack([(coords - c_full).abs().max(),
                       (prob - p_full[:, :, lo:hi]).abs().max() / p_full.abs().max()])
    dist.all_reduce(err, op=dist.ReduceOp.MAX)
    ms = t[0].item() / steps
    return {"workload": "cfg5 K2 forward, ONE (1,%d,%d^3) volume in %d z-slabs" % (J, N, world), "ranks": world,
            "steps": steps, "ms_per_step": round(ms, 4), "volumes_per_s": round(1e3 / ms, 1),
            "exchange": {"op": "all_gather_into_tensor (NCCL)", "bytes_per_rank": J * 5 * 4,
                         "us_per_step": round(t[1].item() / steps * 1e3, 1)},
            "parity": {"max_abs_coord_diff_mm": err[0].item(), "max_rel_prob_diff": err[1].item(),
                       "against": "unsharded lt_softargmax_fwd of the same volume on each rank's GPU"}}


def reference_ops():
    """(module with unproject_heatmaps / integrate_tensor_3d_with_coordinates, kind): the UNMODIFIED reference
    (`mvn.utils.op` from /root/reference or its verbatim copy oracle/_ref, see oracle/fetch_ref.py) when present
    -> "reference"; else the oracle's restatement of the same functions -> "port"."""
    import warnings
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    try:
        from fetch_ref import ref_path
        rp = ref_path()
        if rp is not None:
            sys.path.insert(0, rp)
            try:
                from mvn.utils import op as ref_op
                warnings.filterwarnings("ignore", message=".*grid_sample.*")
                return ref_op, "reference"
            except Exception as e:  # noqa: BLE001
                sys.stderr.write("reference import failed (%s); timing the oracle port instead\n" % e)
    finally:
        sys.path.remove(os.path.join(ROOT, "oracle"))
    from oracle import ref_ops
    return ref_ops, "port"


def cpu_reference_step(w, seed=0, cache={}):
    """fwd+bwd of unproject_heatmaps and integrate_tensor_3d_with_coordinates for ONE batch item on all host cores with
    the reference's own functions (mvn/utils/op.py:84-163, unmodified) or, where its tree is absent, the oracle's
    restatement of them (same ATen calls). Returns seconds."""
    from lt_b200 import synthetic
    ref_ops = reference_ops()[0]
    key = (w["V"], w["C"], w["N"], w["agg"], seed)
    if key not in cache:
        conf = "conf" if w["agg"] == "conf" else None
        n = min(w["N"], 64)  # the 256^3 config needs 8 GiB of per-view temporaries on the CPU: sample it at 64^3
        d = synthetic.make_unproject_inputs(B=1, V=w["V"], C=w["C"], H=w["H"], W=w["W"], N=n, seed=seed, train=True,
                                            conf=conf, dome=w["V"] > 8)
        g = torch.Generator().manual_seed(seed + 5)
        cache[key] = (d, torch.randn(1, w["C"], n, n, n, generator=g), torch.randn(1, w["J"], n, n, n, generator=g),
                      torch.randn(1, w["J"], 3, generator=g))
    d, gvol, x, gc = cache[key]
    t0 = time.perf_counter()
    feat = d["heatmaps"].clone().requires_grad_(True)
    vc = d["vol_confidences"]
    vol = ref_ops.unproject_heatmaps(feat, d["proj_matricies"], d["coord_volumes"], w["agg"], vc)
    vol.backward(gvol)
    xv = x.clone().requires_grad_(True)
    coords, prob = ref_ops.integrate_tensor_3d_with_coordinates(xv, d["coord_volumes"], softmax=True)
    coords.backward(gc)
    return time.perf_counter() - t0


def run_cpu_arm(w, steps, warmup):
    torch.set_num_threads(os.cpu_count() or 1)
    for _ in range(warmup):
        cpu_reference_step(w)
    times = [cpu_reference_step(w) for _ in range(steps)]
    total = sum(times)
    return steps / total, total / steps, torch.get_num_threads(), reference_ops()[1]


# ---------------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--agg", default=None, choices=["sum", "conf", "max", "softmax"],
                    help="A/B runs only: another view aggregation on the workload's shape (named in config.aggregation)")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the bounded CPU sample (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer arm (profiling runs)")
    ap.add_argument("--no-graph", action="store_true", help="e2e arm: launch the python ops eagerly instead of replaying "
                                                            "a CUDA graph captured from them")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.agg is not None and args.agg != w["agg"]:
        w["agg"] = args.agg
        w["desc"] += " [A/B run: aggregation overridden to %s]" % args.agg
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    metric = "volumes/sec fwd+bwd (unproject+aggregate + soft-argmax, %d^3, %d views, %dch)" % (w["N"], w["V"], w["C"])
    config = {"workload": "%s: %s" % (args.workload, w["desc"]), "aggregation": w["agg"], "joints": w["J"],
              "parallelism": ("z-slabs of ONE 256^3 volume x%d ranks, one all_gather_into_tensor of %d bytes per rank and "
                              "step (NCCL) in K2 forward" % (world, 1 * w["J"] * 5 * 4) if args.workload == "cfg5" and world > 1
                              else "batch-sharded x%d, no data-path collective" % world),
              "cache": None}
    _ab = algorithmic_bytes(w, w["B"])
    config["cache"] = ("inputs larger than L2 (per-step working set %.0f MB >> 126 MB L2)" % (sum(_ab.values()) / 2e6)
                       if sum(_ab.values()) / 2 > 3 * 126e6 else
                       "working set fits L2; a 256 MB buffer is rewritten between timed steps (its time is subtracted)")

    if args.impl == "reference":
        # CPU arm: rank 0 only; other ranks exit without work
        if rank != 0:
            return
        steps = max(1, min(args.steps, 8))
        vps, sec, cores, kind = run_cpu_arm(w, steps, max(1, min(args.warmup, 1)))
        what = ("the reference's own mvn/utils/op.py functions, unmodified (oracle/_ref)" if kind == "reference" else
                "oracle port of mvn/utils/op.py (the reference tree is absent)")
        line = {"impl": "reference", "metric": metric, "value": vps, "unit": "volumes/s", "n_gpus": args.gpus,
                "steps": steps, "warmup": 1, "ms_per_step": sec * 1e3, "higher_is_better": True,
                "scaling": w["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": vps, "unit": "volumes/s", "cores": cores, "kind": kind,
                                 "sample": "%d x one batch item (B=1) fwd+bwd on torch CPU, all host threads: %s"
                                           % (steps, what)},
                "e2e": {"value": vps, "unit": "volumes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    assert torch.cuda.is_available(), "bench.py --impl ours needs a CUDA device (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    B = w["B"] if w["scaling"] == "weak" else max(1, w["B"] // world)
    if args.workload == "cfg5":  # z-slabs: this rank's share of the first volume axis
        from lt_b200.slab import slab_bounds
        lo, hi = slab_bounds(w["N"], world, rank)
        w["N0"] = hi - lo
    total_B = w["B"] if args.workload == "cfg5" else B * world   # cfg5: ONE volume shared by all ranks (z-slabs)
    slab_world = world if args.workload == "cfg5" else 1
    step = DeviceStep(w, B, dev, seed=(0 if slab_world > 1 else rank), slab_world=slab_world)
    flush = None if "larger" in config["cache"] else torch.empty(64 * 1024 * 1024, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, per_stage, after_warmup=None):
        for _ in range(warmup):
            fn(None)
        if after_warmup is not None:
            after_warmup()
        evs = [[torch.cuda.Event(enable_timing=True) for _ in range(len(STAGES) + 1)] for _ in range(steps)] \
            if per_stage else None
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sampler = ClockSampler(local_rank)
        sampler.start()
        flush_ms = 0.0
        t0.record()
        for i in range(steps):
            if flush is not None:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); flush.fill_(float(i)); b.record()
                evs and evs[i].append((a, b))
            fn(evs[i] if evs else None)
        t1.record()
        barrier()
        clocks = sampler.stop()
        ms = t0.elapsed_time(t1)
        if flush is not None and evs:
            flush_ms = sum(e[-1][0].elapsed_time(e[-1][1]) for e in evs)
        ms -= flush_ms
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
        return ms, evs, clocks

    ms, evs, clocks = timed(lambda e: step.run(e), args.steps, args.warmup, True)
    value = total_B * args.steps / (ms * 1e-3)

    # per-kernel durations from the events of the timed loop (rank 0's GPU)
    peak, peak_src = measured_peaks()
    abytes = algorithmic_bytes(w, B)
    kernels = {}
    for si, name in enumerate(STAGES):
        dur = sum(e[si].elapsed_time(e[si + 1]) for e in evs) / args.steps
        gbs = abytes[name] / (dur * 1e-3) / 1e9 if dur > 0 else 0.0
        kernels[name] = {"ms": round(dur, 4), "algorithmic_bytes": abytes[name], "gbs": round(gbs, 1),
                         "frac": round(gbs / peak, 4), "share": round(dur / (ms / args.steps), 4)}
    top = max(kernels, key=lambda k: kernels[k]["ms"])
    total_bytes = sum(abytes[k] for k in STAGES if not k.startswith("k0"))  # K0 is our own layout overhead
    path_gbs = total_bytes / (ms / args.steps * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": top, "achieved": kernels[top]["gbs"], "peak": peak, "unit": "GB/s",
                "frac": kernels[top]["frac"], "traffic": None, "peak_source": peak_src,
                "whole_step": {"achieved": round(path_gbs, 1), "frac": round(path_gbs / peak, 4),
                               "note": "algorithmic bytes of K1 fwd/bwd + K2 fwd/bwd over the whole step time"},
                "kernels": kernels}
    # What actually binds K1 on this chip is not HBM: the backward's scatter is limited by the L2
    # reduction units, the forward's gather by the L1 data pipe. Both ceilings were measured with the same address
    # pattern in isolation (tools/microbench/red_paths.cu -> profiles/r1c_microbench_red_paths.txt).
    taps = step.valid_voxel_views()
    if taps is not None:
        tap_bytes = taps * 4 * step.Cp * 4          # 4 bilinear taps x Cp channels x fp32 per valid (voxel, view)
        for name, key, kind, pk in (("k1_unproject_bwd", "scatter", "l2_reduction_units", L2_RED_PEAK_GBS),
                                    ("k1_unproject_fwd", "gather", "l1_gather", L1_GATHER_PEAK_GBS)):
            ach = tap_bytes / (kernels[name]["ms"] * 1e-3) / 1e9
            kernels[name]["binding_limit"] = {"kind": kind, "payload_bytes": tap_bytes, "achieved": round(ach, 1),
                                              "peak": pk, "unit": "GB/s", "frac": round(ach / pk, 4),
                                              "peak_source": "microbenchmark, profiles/r1c_microbench_red_paths.txt"}
        roofline["binding_limit"] = kernels[top].get("binding_limit")
    traffic_file = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_file):
        try:
            roofline["traffic"] = json.load(open(traffic_file)).get(args.workload, {}).get(top)
        except Exception:
            pass

    # SURVEY 8(f) row 4, reported beside the path (never part of `value`): K7 = the process_features 1x1 convolution that
    # feeds K1 (triangulation.py:238-240, 344-346) at this workload's map count, 256 -> C channels, through the C ABI.
    if rank == 0 and world == 1 and not args.no_e2e:
        try:
            roofline["next_rows"] = {"k7_process_features": bench_process_features(step, w, B, peak)}
        except Exception as exc:  # a side measurement must never cost the bench line
            roofline["next_rows"] = {"k7_process_features": {"error": repr(exc)[:200]}}
        try:
            roofline["next_rows"]["in_kernel_coordinates"] = bench_in_kernel_coordinates(step, w, B, peak)
        except Exception as exc:  # noqa: BLE001
            roofline["next_rows"]["in_kernel_coordinates"] = {"error": repr(exc)[:200]}

    # the one collective of the product (cfg5's z-slab soft-argmax), measured on the ranks of THIS launch with a parity
    # number, whatever the workload: the default multi-GPU line (cfg3) has no collective of its own
    if world > 1:
        try:
            rec = bench_sharded_softargmax(rank, world, dev)
        except Exception as exc:  # noqa: BLE001 - a side measurement must never cost the bench line
            rec = {"error": repr(exc)[:200]}
        roofline.setdefault("next_rows", {})["cfg5_sharded_softargmax"] = rec

    e2e = None
    if not args.no_e2e:
        es = E2EStep(step)
        graphed = (not args.no_graph) and es.capture()
        e_steps = max(3, min(args.steps, 50))
        ems, _, _ = timed(lambda e: es.run(), e_steps, 3, False, after_warmup=es.drain)
        e2e = {"value": total_B * e_steps / (ems * 1e-3), "unit": "volumes/s", "h2d_bytes_per_step": es.h2d,
               "d2h_bytes_per_step": es.d2h, "steps": e_steps, "ms_per_step": ems / e_steps, "cuda_graph": graphed,
               "note": "public python ops under autograd (captured once into a CUDA graph per input set when cuda_graph is "
                       "true, replayed every step); pinned-host inputs (feature maps, projection matrices, "
                       "18-float cuboid specs) staged in ONE pinned buffer and copied every step with one cudaMemcpyAsync on a "
                       "second stream (prefetch depth 2), voxel "
                       "coordinates evaluated inside K1/K2 from the specs (the reference builds the coordinate volumes "
                       "on the device too, triangulation.py:280-341), loss + coordinates read back every step"}

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        vps, sec, cores, kind = run_cpu_arm(w, 5, 1)
        cpu_baseline = {"value": vps, "unit": "volumes/s", "cores": cores, "kind": kind,
                        "sample": "5 x one batch item (B=1) fwd+bwd of the same workload on torch CPU, all host threads: "
                                  + ("the reference's own mvn/utils/op.py functions, unmodified (oracle/_ref)"
                                     if kind == "reference" else "oracle port of mvn/utils/op.py")}
    if rank == 0:
        line = {"metric": metric, "value": value, "unit": "volumes/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": w["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config,
                "clocks": clocks, "e2e": e2e, "gpu_launches": sum(launches_per_step(w, B, step.nvox).values()) * args.steps,
                "roofline": roofline, "cpu_baseline": cpu_baseline}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
