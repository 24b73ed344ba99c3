"""Kernel sweep on one B200: time fwd / bwd / fused over variants and cluster sizes with CUDA
events (device buffers, working set >> L2).  Tuning aid; bench.py is the reported number."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from hipe_b200 import ops  # noqa: E402


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in evs)
    return t[len(t) // 2], t[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--clusters", default="0,1,2,4,8")
    ap.add_argument("--variants", default="1,2")
    ap.add_argument("--ops", default="fwd,bwd,fused")
    a = ap.parse_args()
    J, D = 21, a.dim
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    esz = 4 if a.dtype == "f32" else 2
    n_vol = a.B * J
    V = D ** 3
    x = (torch.randn(a.B, J * D, D, D, device="cuda") * 2).to(dt)
    gt = torch.rand(a.B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    g = torch.randn(n_vol, 3, device="cuda")
    peak = 6551.7
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    if "resident" in a.ops:
        med, mn = timeit(lambda: ops.joint_loss_resident(x, n_vol, D, D, D, gt, vis, 1.0 / a.B))
        print(json.dumps({"op": "resident", "B": a.B, "dim": D, "dtype": a.dtype, "ms": round(med, 4), "min_ms": round(mn, 4),
                          "GBs_2V": round(2 * n_vol * V * esz / med / 1e6, 1), "Mvol_s": round(n_vol / med / 1e3, 3),
                          "env": {k: v for k, v in os.environ.items() if k.startswith("HIPE_")}}), flush=True)
        if a.ops == "resident":
            return
    for variant in [int(v) for v in a.variants.split(",")]:
        for cl in [int(c) for c in a.clusters.split(",")]:
            c, e, s = ops.softargmax_forward(x, n_vol, D, D, D, cluster=cl, variant=variant)
            row = {"variant": variant, "cluster": cl, "B": a.B, "dim": D, "dtype": a.dtype}
            if "fwd" in a.ops:
                med, mn = timeit(lambda: ops.softargmax_forward(x, n_vol, D, D, D, cluster=cl, variant=variant))
                row["fwd_ms"] = round(med, 4)
                row["fwd_GBs"] = round(n_vol * V * esz / med / 1e6, 1)
            if "bwd" in a.ops:
                med, mn = timeit(lambda: ops.softargmax_backward(x, n_vol, D, D, D, s, e, g, cluster=cl, variant=variant))
                row["bwd_ms"] = round(med, 4)
                row["bwd_GBs"] = round(2 * n_vol * V * esz / med / 1e6, 1)
            if "fused" in a.ops:
                med, mn = timeit(lambda: ops.joint_loss_fused(x, n_vol, D, D, D, gt, vis, 1.0 / a.B, cluster=cl, variant=variant))
                row["fused_ms"] = round(med, 4)
                row["fused_GBs_2V"] = round(2 * n_vol * V * esz / med / 1e6, 1)
            if "fwd_ms" in row and "bwd_ms" in row:
                t = row["fwd_ms"] + row["bwd_ms"]
                row["fwdbwd_frac_of_peak"] = round(3 * n_vol * V * esz / t / 1e6 / peak, 3)
                row["Mvol_s_unfused"] = round(n_vol / t / 1e3, 3)
            if "fused_ms" in row:
                row["Mvol_s_fused"] = round(n_vol / row["fused_ms"] / 1e3, 3)
            print(json.dumps(row), flush=True)


if __name__ == "__main__":
    main()
