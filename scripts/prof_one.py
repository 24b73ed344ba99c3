"""Run a handful of launches of one op (for ncu): python scripts/prof_one.py --op fwd --dtype f32"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch
from hipe_b200 import ops

ap = argparse.ArgumentParser()
ap.add_argument("--op", default="fwd")
ap.add_argument("--dtype", default="f32")
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--dim", type=int, default=64)
ap.add_argument("--cluster", type=int, default=0)
ap.add_argument("--variant", type=int, default=0)
ap.add_argument("--iters", type=int, default=4)
a = ap.parse_args()
J, D = 21, a.dim
n = a.B * J
x = (torch.randn(a.B, J * D, D, D, device="cuda") * 2)
x = x.to(torch.float32 if a.dtype == "f32" else torch.bfloat16)
gt = torch.rand(a.B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
g = torch.randn(n, 3, device="cuda")
c, e, s = ops.softargmax_forward(x, n, D, D, D, cluster=a.cluster, variant=a.variant)
for _ in range(a.iters):
    if a.op == "fwd":
        ops.softargmax_forward(x, n, D, D, D, cluster=a.cluster, variant=a.variant)
    elif a.op == "bwd":
        ops.softargmax_backward(x, n, D, D, D, s, e, g, cluster=a.cluster, variant=a.variant)
    elif a.op == "resident":
        ops.joint_loss_resident(x, n, D, D, D, gt, vis, 1.0 / a.B)
    elif a.op == "cluster":
        ops.joint_loss_cluster(x, n, D, D, D, gt, vis, 1.0 / a.B)
    elif a.op == "fused":
        ops.joint_loss_fused(x, n, D, D, D, gt, vis, 1.0 / a.B, cluster=a.cluster, variant=a.variant)
torch.cuda.synchronize()
print("ok")
