"""Diagnostic (GPU box): gradient distances between arms of the unmodified volumetric net: stock fp32 twice (noise
floor), stock fp32 vs stock fp64, patched fp32 vs stock fp64."""
import copy, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lt_b200
import test_gpu_reference_nets as T
sys.path.insert(0, T.REF)
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.deterministic = True
from mvn.models.triangulation import VolumetricTriangulationNet
dev = torch.device("cuda", 0)
torch.manual_seed(0)
net = VolumetricTriangulationNet(T._cfg(32, "softmax"), device=dev).to(dev).eval()
images, batch = T._batch(B=2, V=4, img=128, seed=1)
images = images.to(dev)
target = torch.from_numpy(batch["pred_keypoints_3d"]).float().to(dev)
def dist(ga, gb, top=3):
    scale = max(v.abs().max().item() for v in gb.values())
    rows = []
    for n in gb:
        den = max(gb[n].abs().max().item(), 1e-6 * scale)
        rows.append(((ga[n].double() - gb[n].double()).abs().max().item() / den, n, gb[n].abs().max().item() / scale))
    rows.sort(reverse=True)
    return rows[:top]
kp_a, g_a = T._run(net, images, batch, target)
kp_b, g_b = T._run(net, images, batch, target)
print("stock vs stock (noise floor): kp", (kp_a - kp_b).abs().max().item(), dist(g_a, g_b))
lt_b200.patch_reference()
kp_o, g_o = T._run(net, images, batch, target)
lt_b200.unpatch_reference()
print("ours32 vs stock32: kp", (kp_o - kp_a).abs().max().item(), dist(g_o, g_a))
