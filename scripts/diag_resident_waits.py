"""Where the persistent loss+gradient launch waits (diagnostic build, -DHIPE_RES_PROF).

Prints, per class of wait, the share of a warp's lifetime spent in it, averaged over the warps of
that role, all CTAs and `--launches` launches.  Run through scripts/diag_resident_waits.sh, which
builds the instrumented library first and restores the normal one afterwards."""
import argparse
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hand-integral-pose-estimation_b200"))
sys.path.insert(0, ROOT)
from hipe_b200 import _build, _lib, ops  # noqa: E402
from tests import synth  # noqa: E402

# class -> (what the warp waits for, warps of that role per CTA, class holding the role's lifetime)
CLASSES = {
    0: ("reduce warps: ring slot full (data from HBM)", 8, 10),
    7: ("reduce warps: hand-off buffer to publish warp", 8, 10),
    1: ("write warps: ring slot full (data from L2)", 8, 11),
    5: ("write warps: scalar record of the item", 8, 11),
    8: ("write warps: team barrier (row terms)", 8, 11),
    2: ("reduce producer: ring slot empty", 1, 12),
    4: ("reduce producer: lag throttle", 1, 12),
    3: ("write producer: ring slot empty", 1, 13),
    6: ("scalar warp: volume counter (other CTAs' partials)", 1, 14),
    9: ("scalar warp: record slot free", 1, 14),
    15: ("publish warp: warp partials ready", 1, 16),
}
TOTALS = {10: ("reduce warp", 8), 11: ("write warp", 8), 12: ("reduce producer", 1), 13: ("write producer", 1),
          14: ("scalar warp", 1), 16: ("publish warp", 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=64)
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--launches", type=int, default=10)
    a = ap.parse_args()
    _lib.load()
    raw = ctypes.CDLL(_build.LIB_PATH)
    if not hasattr(raw, "hipe_resident_prof"):
        raise SystemExit("library was not built with -DHIPE_RES_PROF")
    J, D = 21, a.dim
    xs = [torch.randn(a.B, J * D, D, D, device="cuda") for _ in range(3)]
    gt, vis = synth.targets(a.B, J, seed=1)
    gt, vis = torch.as_tensor(gt).cuda(), torch.as_tensor(vis).cuda()
    buf = (ctypes.c_ulonglong * 20)()
    for i in range(3):
        ops.joint_loss_resident(xs[i % 3], a.B * J, D, D, D, gt, vis, 1.0 / a.B)
    torch.cuda.synchronize()
    assert raw.hipe_resident_prof(buf, 1) == 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(a.launches):
        ops.joint_loss_resident(xs[i % 3], a.B * J, D, D, D, gt, vis, 1.0 / a.B)
    ev1.record()
    torch.cuda.synchronize()
    assert raw.hipe_resident_prof(buf, 1) == 0
    v = list(buf)
    n_cta = torch.cuda.get_device_properties(0).multi_processor_count
    den = n_cta * a.launches
    print("instrumented launch: %.4f ms (B=%d, %d^3)" % (ev0.elapsed_time(ev1) / a.launches, a.B, D))
    for k, (name, per) in TOTALS.items():
        print("  lifetime %-16s %10.0f cycles" % (name, v[k] / (per * den)))
    for k, (name, per, tot) in CLASSES.items():
        cyc = v[k] / (per * den)
        life = v[tot] / (TOTALS[tot][1] * den)
        print("  %-52s %10.0f cycles  %5.1f %% of its lifetime" % (name, cyc, 100.0 * cyc / max(life, 1.0)))


if __name__ == "__main__":
    main()
