"""Diagnostic (GPU): are the launches bit-reproducible run to run on near-one-hot volumes?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch
from hipe_b200 import ops
from tests.test_gpu_fullsize import _logits, J, D

B = 16
x, _ = _logits(B, 100)
g = torch.Generator(device="cuda").manual_seed(101)
gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
n = B * J
def run(kind):
    if kind == "resident":
        c, e, s, t, d = ops.joint_loss_resident(x, n, D, D, D, gt, vis, 1.0 / B)
    elif kind == "stream":
        c, e, s, t, d = ops.joint_loss_fused(x, n, D, D, D, gt, vis, 1.0 / B, mode="stream")
    else:
        c, e, s = ops.softargmax_forward(x, n, D, D, D)
        _, gc = ops.l1_loss(c.view(B, 3 * J), gt, vis, 1.0 / B)
        d = ops.softargmax_backward(x, n, D, D, D, s, e, gc.view(-1, 3))
    return c.clone(), e.clone(), s.clone(), d.clone()
for kind in ("resident", "stream", "two"):
    ref = run(kind)
    for r in range(5):
        out = run(kind)
        diffs = [int((a != b).sum().item()) for a, b in zip(ref, out)]
        if any(diffs):
            k = (ref[1] != out[1]).nonzero()
            print(kind, "repeat", r, "differs: coords/expect/stats/dlogits elements", diffs,
                  "max |dE|", (ref[1] - out[1]).abs().max().item())
            break
    else:
        print(kind, "bit-identical over 5 repeats")
