"""One forward (+ optional backward) of K7 at cfg3's shape: the command ncu profiles (tools/gpu_k7_ncu.sh)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lt_b200
x = torch.randn(64, 256, 96, 96, device="cuda", requires_grad=True)
w = (torch.randn(32, 256, 1, 1, device="cuda") / 16).requires_grad_(True)
b = torch.randn(32, device="cuda", requires_grad=True)
for _ in range(3):
    y = lt_b200.process_features(x, w, b)
    y.backward(y.detach())
torch.cuda.synchronize()
print("ok", float(y.sum()))
