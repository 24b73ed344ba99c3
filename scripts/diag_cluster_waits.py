"""Where the cluster-resident loss+gradient launch waits (diagnostic build, -DHIPE_CLU_PROF).

    HIPE_NVCC_EXTRA=-DHIPE_CLU_PROF python hand-integral-pose-estimation_b200/hipe_b200/_build.py --force
    python scripts/diag_cluster_waits.py --dtype f32 --B 64        (then rebuild without the flag)

Prints, per class of wait, the share of a warp's lifetime spent in it, averaged over the warps of
that role, all CTAs and `--launches` launches."""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
from hipe_b200 import _build, _lib, ops  # noqa: E402

# class -> (what the warp waits for, warps of that role per CTA, class holding the role's lifetime)
CLASSES = {
    0: ("producer: ring slot empty (write team behind)", 1, 8),
    1: ("reduce warps: stage landed (data from HBM)", 8, 9),
    2: ("reduce warps: hand-off buffer to merge warp", 8, 9),
    3: ("merge warp: reduce team's partial", 1, 10),
    4: ("merge warp: peers' partials (DSMEM)", 1, 10),
    5: ("merge warp: record slot free", 1, 10),
    6: ("write warps: scalar record of the volume", 8, 11),
    7: ("write warps: full barrier re-check", 8, 11),
}
TOTALS = {8: ("producer", 1), 9: ("reduce warp", 8), 10: ("merge warp", 1), 11: ("write warp", 8)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--launches", type=int, default=10)
    a = ap.parse_args()
    _lib.load()
    raw = ctypes.CDLL(_build.LIB_PATH)
    if not hasattr(raw, "hipe_cluster_prof"):
        raise SystemExit("library was not built with -DHIPE_CLU_PROF")
    J, D = 21, a.dim
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    xs = [torch.randn(a.B, J * D, D, D, device="cuda").to(dt) for _ in range(3)]
    gt = torch.rand(a.B, 3 * J, device="cuda", dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    n_vol = a.B * J
    plan = ops.cluster_plan(xs[0], n_vol, D, D, D)
    buf = (ctypes.c_ulonglong * 12)()
    for i in range(3):
        ops.joint_loss_cluster(xs[i % 3], n_vol, D, D, D, gt, vis, 1.0 / a.B)
    torch.cuda.synchronize()
    assert raw.hipe_cluster_prof(buf, 1) == 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(a.launches):
        ops.joint_loss_cluster(xs[i % 3], n_vol, D, D, D, gt, vis, 1.0 / a.B)
    ev1.record()
    torch.cuda.synchronize()
    assert raw.hipe_cluster_prof(buf, 1) == 0
    v = list(buf)
    ms = ev0.elapsed_time(ev1) / a.launches
    life_w = v[11] / max(1, 1)  # summed warp-cycles of write warps
    # CTAs per launch from the lifetime of the producers: sum(lifetimes) / (launch cycles) is not known
    # exactly, so normalise every class by its own role's summed lifetime instead
    print("instrumented launch: %.4f ms (%s B=%d, %d^3), plan C=%d slots=%d share=%d env=%s" % (
        ms, a.dtype, a.B, D, plan[0], plan[1], plan[2],
        {k: v_ for k, v_ in os.environ.items() if k.startswith("HIPE_CLUSTER")}))
    for k, (name, per, tot) in CLASSES.items():
        print("  %-50s %5.1f %% of its role's lifetime" % (name, 100.0 * v[k] / max(v[tot], 1)))
    if hasattr(raw, "hipe_cluster_trace"):
        tr = (ctypes.c_longlong * 80)()
        assert raw.hipe_cluster_trace(tr) == 0
        t = [list(tr[i * 10:(i + 1) * 10]) for i in range(8)]
        names = ["issue0", "issueL", "landL", "folded", "m:team", "m:pub", "m:peers", "m:rec", "w:start", "w:end"]
        base = t[0][0]
        print("  CTA 0 event trace, microseconds at 1.965 GHz relative to volume k0's first issue:")
        print("    k   " + " ".join("%8s" % n for n in names))
        for i, row in enumerate(t):
            print("    %2d  " % i + " ".join("%8.2f" % ((v - base) / 1965.0) for v in row))
    print("  mean role lifetime / launch time: " + ", ".join(
        "%s %.2f" % (TOTALS[k][0], (v[k] / TOTALS[k][1]) / max(v[8], 1)) for k in TOTALS))


if __name__ == "__main__":
    main()
