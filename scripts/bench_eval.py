"""BASELINE config 5 (eval path) on its own: bench.py's `configs.C5` entry as one JSON line.

    python scripts/bench_eval.py [--samples 3960] [--batch 32] [--dim 64] [--dtype f32] [--reps 5]
    torchrun --nproc-per-node N scripts/bench_eval.py     # samples sharded across ranks
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=3960)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    cx = B.Ctx()
    r = B.eval_case(cx, a.samples, a.batch, a.dim, a.dtype, a.reps)
    if cx.rank == 0:
        r["n_gpus"] = cx.world
        print(json.dumps(r), flush=True)
    if cx.world > 1:
        cx.dist.barrier()
        cx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
