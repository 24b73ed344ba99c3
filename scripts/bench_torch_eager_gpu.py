"""Same-box comparator: the reference's PyTorch-eager operator sequence for
JointLocationLoss forward + backward (common/nets/loss.py:23-59,74-91; main/train.py:113) run ON THE
B200 with stock ATen kernels, at the bench configuration (64 samples x 21 x 64^3 fp32).  Not part of
the product or of bench.py's contract; the sequence is restated inline (the reference tree does not
travel to the GPU box; on CUDA the reference additionally pays three H2D arange copies per call,
loss.py:35-37, which are NOT reproduced here -- this comparator is therefore slightly flattering).
Prints one JSON line with volumes/s, ms/step and peak memory."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench as B  # noqa: E402  (synthetic logits only)

J = 21


def eager_step(x, gt, vis, D, H, W):
    x = x.detach().requires_grad_(True)
    p = torch.softmax(x.reshape(x.shape[0], J, -1), dim=2)                       # loss.py:48-49
    h = p.reshape(p.shape[0], J, D, H, W)                                        # :25
    mx = h.sum(dim=2).sum(dim=2)                                                 # :27,29
    my = h.sum(dim=2).sum(dim=3)                                                 # :31-32
    mz = h.sum(dim=3).sum(dim=3)                                                 # :33-34
    ex = (mx * torch.arange(W, device=x.device)).sum(dim=2, keepdim=True)        # :35,39
    ey = (my * torch.arange(H, device=x.device)).sum(dim=2, keepdim=True)
    ez = (mz * torch.arange(D, device=x.device)).sum(dim=2, keepdim=True)
    c = torch.cat((ex / float(W) - 0.5, ey / float(H) - 0.5, ez / float(D) - 0.5), dim=2)
    c = c.reshape(c.shape[0], J * 3)                                             # :54-58
    loss = (torch.abs(c - gt) * vis).sum() / len(c)                              # :86-89
    loss.backward()                                                              # train.py:113
    return loss.detach(), x.grad


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    D = a.dim
    xs = [B.synth_logits(torch, a.batch, D, torch.float32, dev, 1234 + i) for i in range(3)]
    g = torch.Generator(device=dev).manual_seed(99)
    gt = torch.rand(a.batch, 3 * J, generator=g, device=dev, dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    for i in range(3):
        eager_step(xs[i % 3], gt, vis, D, D, D)
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(a.steps):
        loss, _ = eager_step(xs[i % 3], gt, vis, D, D, D)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    n_vol = a.batch * J
    print(json.dumps({"impl": "pytorch-eager-on-gpu (comparator)", "volumes_per_s": n_vol / (ms / 1e3), "ms_per_step": ms,
                      "batch": a.batch, "dim": D, "loss": float(loss.item()),
                      "peak_extra_memory_GB": (torch.cuda.max_memory_allocated() - base) / 1e9,
                      "step_GBs_if_3_passes": n_vol * 3 * D ** 3 * 4 / (ms / 1e3) / 1e9}), flush=True)


if __name__ == "__main__":
    main()
