"""Diagnostic: run the peer exchange a few steps on every rank and print its raw state.
torchrun --nproc-per-node 2 scripts/diag_exchange.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200")); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from hipe_b200 import parallel as PL

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank); dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
b, k = int(os.environ.get("DIAG_B", "3")), 63
px = PL._PeerExchange(world, rank, b, k, dev, None)
for step in range(4):
    loss = torch.tensor([10.0 * (rank + 1) + step], dtype=torch.float64, device=dev)
    coords = torch.full((b, k), float(rank) + 0.25 * step, dtype=torch.float32, device=dev)
    v, allc = px.run(loss, 1.0, coords, 1.0)
    torch.cuda.synchronize()
    print("rank", rank, "step", step, "value", v.item(), "err", px.err.item(), "scratch", px.scratch.tolist(),
          "coords rows", allc.view(world, -1)[:, 0].tolist(), allc.view(world, -1)[:, -1].tolist(), flush=True)
dist.barrier(); dist.destroy_process_group()
