"""Diagnostic: where the CUDA path's error against the fp64 oracle comes from (GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch
from hipe_b200 import ops
from oracle import integral_np as O
from tests import synth

J = 21
for geom, regime in [((16, 16, 16), "peaked"), ((64, 64, 64), "peaked"), ((64, 64, 64), "noise"), ((64, 64, 64), "sharp"), ((12, 20, 24), "peaked")]:
    D, H, W = geom
    B = 2
    hm = synth.heatmaps(B, J, D, H, W, regime, seed=5)
    x64 = hm.astype(np.float64)
    n = B * J
    s64 = x64.reshape(n, -1)
    m = s64.max(axis=1); L = np.exp(s64 - m[:, None]).sum(axis=1)
    p = O.volume_softmax(x64, J)
    ex, ey, ez = O.marginal_expectations(p, J, W, H, D)
    E = np.concatenate([ex, ey, ez], axis=2).reshape(n, 3)
    g = np.random.default_rng(3).standard_normal((B, 3 * J))
    ref = O.integral_coords_vjp(x64, J, W, H, D, g).reshape(n, -1)
    scale = np.abs(ref).max(axis=1, keepdims=True)
    x = torch.as_tensor(hm).cuda()
    gg = torch.as_tensor(g.reshape(-1, 3)).float().cuda()
    for cl in (1, 4):
        c, e, s = ops.softargmax_forward(x, n, D, H, W, cluster=cl)
        eE = np.abs(e.cpu().numpy() - E)
        eL = np.abs(s[:, 1].cpu().numpy() - L) / L
        d = ops.softargmax_backward(x, n, D, H, W, s, e, gg, cluster=cl).cpu().numpy().reshape(n, -1)
        err = (np.abs(d - ref) / scale).max()
        # backward with the exact expectations / exact L
        e_exact = torch.as_tensor(E).cuda()
        d2 = ops.softargmax_backward(x, n, D, H, W, s, e_exact, gg, cluster=cl).cpu().numpy().reshape(n, -1)
        s_exact = s.clone(); s_exact[:, 1] = torch.as_tensor(L).float().cuda()
        d3 = ops.softargmax_backward(x, n, D, H, W, s_exact, e_exact, gg, cluster=cl).cpu().numpy().reshape(n, -1)
        print("%s %-6s cl%d  max|E err| %.2e vox (rel %.2e)  max rel L err %.2e | grad err/scale: ours %.2e, exact E %.2e, exact E+L %.2e"
              % (geom, regime, cl, eE.max(), (eE / np.maximum(E, 1)).max(), eL.max(), err, (np.abs(d2 - ref) / scale).max(),
                 (np.abs(d3 - ref) / scale).max()))
