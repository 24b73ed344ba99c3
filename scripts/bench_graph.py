"""The drop-in criterion's fwd+bwd step replayed from CUDA graphs vs issued eagerly (one B200).

Short steps (the reference trains 56^3 heatmaps at 32 samples/GPU) are bound by the host's enqueue
rate when issued eagerly; a caller that captures its training step hides that.  This script checks
that the library's launches are capture-safe (stream-ordered memset + kernels, cooperative launch
included, no host synchronisation) and reports both step times.
    python scripts/bench_graph.py --dim 56 --batch 32"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
from nets import loss as L  # noqa: E402
from tests import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dim", type=int, default=56)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--steps", type=int, default=300)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    J, D, B = 21, a.dim, a.batch
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    nbuf = max(3, int(3 * 200e6 // (B * J * D ** 3 * (4 if a.dtype == "f32" else 2))) + 1)   # > L2 in rotation
    nbuf = min(nbuf, 8)
    xs = [(torch.randn(B, J * D, D, D, device=dev) * 2).to(dt).requires_grad_(True) for _ in range(nbuf)]
    gt, vis = synth.targets(B, J, seed=1)
    gt, vis = torch.as_tensor(gt).to(dev), torch.as_tensor(vis).to(dev)
    crit = L.JointLocationLoss()

    def step(x):
        x.grad = None                     # assign, do not accumulate (what optimizer.zero_grad() does)
        loss = crit(x, gt, vis)
        loss.backward()
        return loss

    def timed(fn, n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(n):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    # eager
    for i in range(10):
        step(xs[i % nbuf])
    eager_loss = step(xs[0]).item()
    eager_grad = xs[0].grad.clone()
    eager_ms = timed(lambda i: step(xs[i % nbuf]), a.steps)

    # graphs: one per input buffer (static addresses), warmed up on a side stream first
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for x in xs:
            x.grad = None
            step(x)
    torch.cuda.current_stream().wait_stream(s)
    graphs, losses = [], []
    for x in xs:
        x.grad = None
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            losses.append(step(x))
        graphs.append(g)
    graphs[0].replay()
    torch.cuda.synchronize()
    # (xs[0].grad is the buffer the capture allocated; replay refills it)
    same_loss = abs(losses[0].item() - eager_loss) <= 1e-12 * abs(eager_loss)
    same_grad = bool(torch.equal(xs[0].grad, eager_grad))
    graph_ms = timed(lambda i: graphs[i % nbuf].replay(), a.steps)
    n_vol = B * J
    print(json.dumps({"workload": "JointLocationLoss fwd+bwd, %s %d^3, batch %d" % (a.dtype, D, B),
                      "eager_ms_per_step": eager_ms, "graph_ms_per_step": graph_ms,
                      "eager_volumes_per_s": n_vol / eager_ms * 1e3, "graph_volumes_per_s": n_vol / graph_ms * 1e3,
                      "graph_matches_eager": {"loss": same_loss, "gradient_bits": same_grad},
                      "rotating_buffers": nbuf}))


if __name__ == "__main__":
    main()
