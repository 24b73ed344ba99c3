"""Eager sharded criterion against hipe_b200.graphed_sharded on the strong-scaling config (bf16 64^3,
global batch 512): ms per step (CUDA events, max over ranks, median of 7 blocks of 20) and host
enqueue per step.   torchrun --nproc-per-node N scripts/bench_graphed_sharded.py [--global-batch 512]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hand-integral-pose-estimation_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--global-batch", type=int, default=512)
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--dim", type=int, default=64)
a = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))))
import hipe_b200  # noqa: E402
from hipe_b200 import parallel  # noqa: E402

J, D = 21, a.dim
B = a.global_batch // world
dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
xs = [(torch.randn(B, J * D, D, D, device="cuda") * 2).to(dt) for _ in range(3)]
gt = torch.rand(B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
crit = parallel.ShardedJointLocationLoss(global_batch=B * world)
xr = [x.clone().requires_grad_(True) for x in xs]


def eager(i):
    x = xr[i % 3]
    x.grad = None
    loss = crit(x, gt, vis)
    loss.backward()
    return loss


step = hipe_b200.graphed_sharded(xs[0], gt, vis, global_batch=B * world)


def graphed(i):
    # (the producer writes into the static input: no copy in the timed region, like the eager arm)
    return step()[0]


def measure(fn, blocks=7, steps=20):
    for i in range(10):
        fn(i)
    torch.cuda.synchronize()
    ts, hs = [], []
    for _ in range(blocks):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        h0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            last = fn(i)
        e1.record()
        hs.append((time.perf_counter() - h0) / steps * 1e3)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(t.item())
    ts.sort(); hs.sort()
    return ts[len(ts) // 2], ts[0], hs[len(hs) // 2], float(last.item())


out = {"world": world, "samples_per_rank": B, "dtype": a.dtype, "dim": D}
for name, fn in (("eager", eager), ("graphed", graphed), ("eager2", eager), ("graphed2", graphed)):
    ms, mn, host, val = measure(fn)
    out[name] = {"ms_per_step": round(ms, 4), "min": round(mn, 4), "host_enqueue_ms": round(host, 4),
                 "volumes_per_s": round(world * B * J / ms * 1e3), "loss": val}
step.check()
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
