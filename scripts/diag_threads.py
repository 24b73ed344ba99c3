"""Diagnostic (GPU): concurrent launches from several Python threads / streams on ONE device."""
import os, sys, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch
from hipe_b200 import ops
from tests.test_gpu_fullsize import _logits, J, D

B = 16
n = B * J
kind = sys.argv[1] if len(sys.argv) > 1 else "resident"
x, _ = _logits(B, 100)
g = torch.Generator(device="cuda").manual_seed(101)
gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
def run():
    if kind == "resident":
        c, e, s, t, d = ops.joint_loss_resident(x, n, D, D, D, gt, vis, 1.0 / B)
    elif kind == "stream":
        c, e, s, t, d = ops.joint_loss_fused(x, n, D, D, D, gt, vis, 1.0 / B, mode="stream")
    else:
        c, e, s = ops.softargmax_forward(x, n, D, D, D)
        _, gc = ops.l1_loss(c.view(B, 3 * J), gt, vis, 1.0 / B)
        d = ops.softargmax_backward(x, n, D, D, D, s, e, gc.view(-1, 3))
    return e, s, d
ref = [t.clone() for t in run()]
torch.cuda.synchronize()
lock = threading.Lock()
def work(k):
    st = torch.cuda.Stream() if k else torch.cuda.current_stream()
    with torch.cuda.stream(st):
        for r in range(6):
            e, s, d = run()
            st.synchronize()
            diffs = [int((a != b).sum().item()) for a, b in zip(ref, (e, s, d))]
            to = ops.resident_timed_out("cuda:0", n) if kind == "resident" else None
            if any(diffs) or to:
                bad = (ref[2] != d).view(n, -1).any(dim=1).nonzero().flatten().tolist()[:8]
                dd = (ref[2] != d).view(n, -1)
                v0 = bad[0] if bad else 0
                pos = dd[v0].nonzero().flatten()
                rv, nv = ref[2].view(n, -1)[v0][pos[:6]].tolist(), d.view(n, -1)[v0][pos[:6]].tolist()
                # hypothesis: the write team consumed the logits of ANOTHER stage (ring race): then
                # new/ref = exp(x[pos + shift] - x[pos]) for one shift common to a run of positions
                xf = x.view(n, -1)
                rr_, nn_ = ref[2].view(n, -1)[v0].double(), d.view(n, -1)[v0].double()
                p0 = int(pos[0].item()) // 128 * 128
                pp = torch.arange(p0, p0 + 128, device=x.device)
                implied = (xf[v0][pp].double() + torch.log((nn_[pp] / rr_[pp]).abs())).float()   # if ax + rt were right
                blocks = xf.reshape(-1, 128)
                dist = (blocks - implied[None, :]).abs().median(dim=1).values
                bi = int(dist.argmin().item())
                best = (dist[bi].item(), bi // 2048, (bi % 2048) * 128, p0)
                with lock:
                    print("  implied-logit search: median |err| %.3g -> volume %d offset %d (bad block at offset %d of volume %d)" % (best + (v0,)), flush=True)
                    print("  rows bad:", sorted(set((pos // 64).tolist()))[:40], flush=True)
                    print("  vol", v0, "n_bad_in_vol", int(dd[v0].sum()), "first pos", pos[:6].tolist(), "last pos", pos[-3:].tolist(),
                          "ref", rv, "new", nv, "logit there", x.view(n, -1)[v0][pos[:3]].tolist(), flush=True)
                    print("thread", k, "repeat", r, "diff elements (expect, stats, dlogits)", diffs, "timed_out", to,
                          "first bad volumes", bad, "max|dE|", (ref[0] - e).abs().max().item(), flush=True)
ths = [threading.Thread(target=work, args=(k,)) for k in range(3)]
[t.start() for t in ths]; [t.join() for t in ths]
print(kind, "done")
