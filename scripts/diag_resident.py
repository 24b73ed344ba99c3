"""Diagnostic (GPU): resident vs two-launch vs oracle on the cases of tests/test_gpu_resident.py."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from hipe_b200 import ops
from oracle import integral_np as O
from tests import synth

def cuda(a, dt=None):
    t = torch.as_tensor(a).cuda()
    return t if dt is None else t.to(dt)

D = H = W = 64; J = 21
for regime, seed, B in [("sharp", 5, 3), ("peaked", 4, 2)]:
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=seed)
    gt, vis = synth.targets(B, J, seed=seed)
    for dt in (torch.float32, torch.bfloat16):
        x = cuda(hm).to(dt)
        up = x.float().cpu().numpy().astype(np.float64)
        ref_loss, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
        ref_g = O.joint_location_loss_grad(up, gt, vis).reshape(B * J, -1)
        ref_E = (ref_c.reshape(-1, 3) + 0.5) * np.array([W, H, D])
        scale = np.abs(ref_g).max(axis=1, keepdims=True)
        c1, e1, s1, t1, d1 = ops.joint_loss_resident(x, B * J, D, H, W, cuda(gt), cuda(vis), 1.0 / B)
        c0, e0, s0 = ops.softargmax_forward(x, B * J, D, H, W)
        _, g = ops.l1_loss(c0.view(B, 3 * J), cuda(gt), cuda(vis), 1.0 / B)
        d0 = ops.softargmax_backward(x, B * J, D, H, W, s0, e0, g.view(-1, 3))
        for name, e, d in (("resident", e1, d1), ("two-launch", e0, d0)):
            dE = np.abs(e.cpu().numpy() - ref_E).max()
            dd = d.float().cpu().numpy().astype(np.float64).reshape(B * J, -1)
            if dt == torch.bfloat16:
                want = torch.as_tensor(ref_g).to(torch.bfloat16).float().numpy()
                ulp = np.maximum(np.abs(want), 1e-30) * 2.0 ** -7
                bad = (np.abs(dd - want) > ulp + 1e-12)
                print(regime, dt, name, "max|E-E64| %.2e voxel" % dE, "bf16 grads >1ulp: %d of %d, worst %.2f ulp" %
                      (bad.sum(), bad.size, (np.abs(dd - want) / ulp).max()))
            else:
                rel = np.abs(dd - ref_g) / scale
                print(regime, dt, name, "max|E-E64| %.2e voxel" % dE, "grad err/max per volume: worst %.2e" % rel.max())

# ---- where is the two-launch gradient off in the sharp regime? ----
B = 3
hm = synth.heatmaps(B, J, D, H, W, "sharp", seed=5)
gt, vis = synth.targets(B, J, seed=5)
x = cuda(hm)
up = hm.astype(np.float64)
ref_loss, ref_c = O.joint_location_loss(up, gt, vis, return_coords=True)
ref_g = O.joint_location_loss_grad(up, gt, vis).reshape(B * J, -1)
c0, e0, s0 = ops.softargmax_forward(x, B * J, D, H, W)
_, g = ops.l1_loss(c0.view(B, 3 * J), cuda(gt), cuda(vis), 1.0 / B)
for cl in (0, 1, 2, 4, 8):
    for var in (1, 2):
        d0 = ops.softargmax_backward(x, B * J, D, H, W, s0, e0, g.view(-1, 3), cluster=cl, variant=var)
        dd = d0.cpu().numpy().astype(np.float64).reshape(B * J, -1)
        rel = np.abs(dd - ref_g) / np.abs(ref_g).max(axis=1, keepdims=True)
        v = int(rel.max(axis=1).argmax()); i = int(rel[v].argmax())
        z, y, xx = i // (H * W), (i // W) % H, i % W
        print("cluster", cl, "variant", var, "worst %.2e" % rel.max(), "vol", v, "voxel zyx", (z, y, xx), "E", e0[v].cpu().numpy(),
              "ours %.6e ref %.6e max %.3e" % (dd[v, i], ref_g[v, i], np.abs(ref_g[v]).max()),
              "c-gt", (ref_c.reshape(-1, 3)[v] - gt.reshape(-1, 3)[v]))
