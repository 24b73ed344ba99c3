"""BASELINE config 5 (eval path): a FreiHAND-sized synthetic evaluation set of 3960 samples through
soft-argmax forward -> patch->camera chain + Procrustes (one fused launch) -> PA-MPJPE / MPJPE ->
EvalUtil measures, everything on the device.  One JSON line; not the driver's bench (bench.py is).

    python scripts/bench_eval.py [--samples 3960] [--batch 32] [--dim 64] [--dtype f32] [--reps 5]
    torchrun --nproc-per-node N scripts/bench_eval.py     # samples sharded across ranks

The heatmaps of a batch are what the network would have just written to HBM: 3 rotating resident
batches stand in for them (working set > L2), the per-batch soft-argmax output is appended to a
(3960, 63) device buffer, and the geometry runs once at the end -- the reference instead gathers the
full heatmaps to GPU 0 (main/test.py:91-92), evaluates the soft-argmax up to 4x per batch and loops
over the samples in numpy (data/FreiHand/FreiHand.py:703-767)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from hipe_b200 import evaluate as E, ops, parallel  # noqa: E402
from nets import loss as L  # noqa: E402
from tests import synth  # noqa: E402
import bench as B  # noqa: E402

J = 21


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=3960)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lo, hi = parallel.shard_bounds(a.samples, world, rank)
    n_local = hi - lo
    D = a.dim
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    xs = [B.synth_logits(torch, a.batch, D, dt, dev, 77 + 3 * rank + i) for i in range(3)]
    c = synth.chain_inputs(a.samples, seed=5)
    cu = {k: torch.as_tensor(np.ascontiguousarray(v[lo:hi])).to(dev) for k, v in c.items()}
    coords_all = torch.empty((max(n_local, 1), 3 * J), dtype=torch.float32, device=dev)
    n_batches = (n_local + a.batch - 1) // a.batch

    def one_pass():
        with torch.no_grad():
            for b in range(n_batches):
                s0 = b * a.batch
                n = min(a.batch, n_local - s0)
                x = xs[b % 3][:n]
                coords_all[s0:s0 + n] = L.softmax_integral_tensor(x, J, D, D, D)
            pa, raw, o = E.evaluate_chain(coords_all[:n_local], cu["gt_cam"], cu["tprime"], cu["K"], bbox=cu["bbox"],
                                          R=cu["R"], bone_len=cu["bone_len"])
            m = E.keypoint_measures(o["pa_dist"], 0.0, 0.05, 100)
            rec = torch.stack([pa * o["pa_dist"].numel(), raw * o["raw_dist"].numel(),
                               torch.tensor(float(o["pa_dist"].numel()), dtype=torch.float64, device=dev)])
            if world > 1:
                pa_g, raw_g = parallel.reduce_error_sums(parallel.gather_records(rec))
            else:
                pa_g, raw_g = pa, raw
        return pa_g, raw_g, m

    for _ in range(2):
        one_pass()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    times = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.KERNEL_EVENTS = {}
        e0.record()
        pa, raw, m = one_pass()
        e1.record()
        torch.cuda.synchronize()
        ev, ops.KERNEL_EVENTS = ops.KERNEL_EVENTS, None
        times.append(e0.elapsed_time(e1))
    kern = {k: sum(s.elapsed_time(e) for s, e in v) for k, v in ev.items()}
    t = torch.tensor([sorted(times)[len(times) // 2]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        esz = 4 if a.dtype == "f32" else 2
        fwd_ms = kern.get("softargmax_fwd", 0.0)
        print(json.dumps({
            "metric": "eval samples/sec (soft-argmax fwd + chain + Procrustes + measures, 21x%d^3)" % D,
            "value": a.samples / (ms / 1e3), "unit": "samples/s", "volumes_per_s": a.samples * J / (ms / 1e3),
            "n_gpus": world, "samples": a.samples, "batch": a.batch, "ms_per_pass": ms, "dtype": a.dtype,
            "kernel_ms_per_pass": kern,
            "softargmax_fwd_GBs": (n_local * J * D ** 3 * esz / (fwd_ms / 1e3) / 1e9) if fwd_ms else None,
            "pa_mpjpe_m": float(pa.item()), "mpjpe_m": float(raw.item()),
            "epe_mean_m": float(m[0].item()), "auc": float(m[2].item())}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
