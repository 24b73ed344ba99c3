"""Intrinsic latency of the peer-memory exchange: ranks in lockstep, exchanges back to back.
torchrun --nproc-per-node N scripts/diag_exchange_latency.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from hipe_b200 import parallel as PL

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
b, k = int(os.environ.get("DIAG_B", "64")), 63
px = PL._PeerExchange(world, rank, b, k, dev, None)
loss = torch.tensor([1.0 + rank], dtype=torch.float64, device=dev)
coords = torch.full((b, k), float(rank), dtype=torch.float32, device=dev)
for _ in range(20):
    px.run(loss, 1.0, coords, 1.0)
torch.cuda.synchronize(); dist.barrier()
n = 500
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(n):
    v, _ = px.run(loss, 1.0, coords, 1.0)
e1.record(); torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print("world %d, %d samples/rank: %.2f us per exchange (back to back, max over ranks), value %.3f, err %d"
          % (world, b, 1e3 * t.item(), v.item(), px.err.item()))
dist.barrier(); dist.destroy_process_group()
