"""Where the host time of one sharded training step goes (cProfile of the enqueue path, rank 0):
    torchrun --nproc-per-node 2 scripts/prof_host_sharded.py [--dtype bf16 --B 64]"""
import argparse
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hand-integral-pose-estimation_b200"), ROOT):
    sys.path.insert(0, p)
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--B", type=int, default=64)
ap.add_argument("--steps", type=int, default=300)
a = ap.parse_args()
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
from hipe_b200 import parallel  # noqa: E402
from nets import loss as L  # noqa: E402

J, D = 21, 64
dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
xs = [(torch.randn(a.B, J * D, D, D, device="cuda") * 2).to(dt).requires_grad_(True) for _ in range(3)]
gt = torch.rand(a.B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
vis = torch.ones_like(gt)
crit = parallel.ShardedJointLocationLoss(global_batch=a.B * world) if world > 1 else L.JointLocationLoss()


def step(i):
    x = xs[i % 3]
    x.grad = None
    loss = crit(x, gt, vis)
    loss.backward()
    return loss


for i in range(30):
    step(i)
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(a.steps):
    step(i)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
if rank == 0:
    print("enqueue %.1f us/step, with drain %.1f us/step" % ((t1 - t0) / a.steps * 1e6, (t2 - t0) / a.steps * 1e6))
pr = cProfile.Profile()
pr.enable()
for i in range(a.steps):
    step(i)
pr.disable()
torch.cuda.synchronize()
if rank == 0:
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
