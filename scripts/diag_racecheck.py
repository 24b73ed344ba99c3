"""Diagnostic: one resident launch, for `compute-sanitizer --tool racecheck`."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch
from hipe_b200 import ops
J, D = 21, 64
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
x = torch.randn(B, J * D, D, D, device="cuda")
gt = torch.rand(B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
out = ops.joint_loss_resident(x, B * J, D, D, D, gt, vis, 1.0 / B)
torch.cuda.synchronize()
print("ok", out[0][0].tolist())
