"""Same-box A/B of the single-launch loss+gradient kernels (CUDA events, rotating inputs larger
than L2): cluster-resident vs persistent L2-window (resident) vs streaming fused vs the pair.
    python scripts/ab_cluster.py [quick | 56 | str | 96 | lead [lead2] | pf | bfc | peek | sym | wt2]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "hand-integral-pose-estimation_b200"), ROOT):
    sys.path.insert(0, p)
import torch
from hipe_b200 import ops

J = 21
PEAK = 6535.7
try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timeit(fn, bufs, reps=30, warm=5):
    for i in range(warm):
        fn(bufs[i % len(bufs)])
    torch.cuda.synchronize()
    ts = []
    for blk in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            fn(bufs[i % len(bufs)])
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / reps)
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def case(dt, D, B, envs, baselines=True):
    n_vol = B * J
    esz = 4 if dt == torch.float32 else 2
    nbuf = max(3, int(400e6 // (n_vol * D ** 3 * esz)) + 1)
    g = torch.Generator(device="cuda").manual_seed(1)
    bufs = []
    for i in range(nbuf):
        x = torch.randn(B, J * D, D, D, generator=g, device="cuda")
        x.view(n_vol, -1)[:, 12345 % D ** 3] += 12.0
        bufs.append(x.to(dt))
    gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    byt = n_vol * 2 * D ** 3 * esz
    out = {}

    def rec(name, fn):
        try:
            med, mn = timeit(fn, bufs)
            out[name] = med
            print("  %-34s %.4f ms (min %.4f)  %6.0f GB/s  frac %.3f" % (name, med, mn, byt / med / 1e6, byt / med / 1e6 / PEAK), flush=True)
        except Exception as e:
            print("  %-34s FAILED %r" % (name, e), flush=True)

    print("%s %d^3 x %d volumes (B=%d)" % ("fp32" if esz == 4 else "bf16", D, n_vol, B), flush=True)
    loss = torch.empty((), dtype=torch.float64, device="cuda")
    if baselines and ops.resident_supported(bufs[0], n_vol, D, D, D) and n_vol >= 296:
        rec("resident", lambda x: ops.joint_loss_resident(x, n_vol, D, D, D, gt, vis, 1.0 / B, loss_out=loss))
    if baselines:
        rec("stream fused", lambda x: ops.joint_loss_fused(x, n_vol, D, D, D, gt, vis, 1.0 / B, mode="stream"))
    for env in envs:
        os.environ.update(env)
        try:
            plan = ops.cluster_plan(bufs[0], n_vol, D, D, D)
            if plan is None:
                print("  cluster %s: unsupported" % env)
                continue
            rec("cluster %s C=%d share=%d" % (",".join("%s=%s" % (k[13:], v) for k, v in env.items()), plan[0], plan[2]),
                lambda x: ops.joint_loss_cluster(x, n_vol, D, D, D, gt, vis, 1.0 / B, loss_out=loss))
        finally:
            for k in env:
                del os.environ[k]

    def pair(x):
        c, e, s = ops.softargmax_forward(x, n_vol, D, D, D)
        _, gg = ops.l1_loss(c.view(B, 3 * J), gt, vis, 1.0 / B)
        ops.softargmax_backward(x, n_vol, D, D, D, s, e, gg.view(-1, 3))
    if baselines:
        rec("pair fwd+l1+bwd (3V bytes: x1.5)", pair)
    return out


def cross_check(dt, D, B, env=None):
    """cluster launch against the stream launch on the same input (a smoke check before timing;
    the parity tests proper compare with the oracle)"""
    os.environ.update(env or {})
    try:
        n_vol = B * J
        g = torch.Generator(device="cuda").manual_seed(7)
        x = (torch.randn(B, J * D, D, D, generator=g, device="cuda") * 2).to(dt)
        gt = torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) - 0.5
        vis = (torch.rand(B, 3 * J, generator=g, device="cuda", dtype=torch.float64) > 0.2).double()
        la, lb = (torch.zeros((), dtype=torch.float64, device="cuda") for _ in range(2))
        a = ops.joint_loss_cluster(x, n_vol, D, D, D, gt, vis, 1.0 / B, loss_out=la)
        b = ops.joint_loss_fused(x, n_vol, D, D, D, gt, vis, 1.0 / B, mode="stream", loss_out=lb)
        torch.cuda.synchronize()
        names = ("coords", "expect", "stats", "terms", "dlogits")
        d = {n: float((u.double() - v.double()).abs().max()) for n, u, v in zip(names, a, b)}
        d["dlogits_rel_to_max"] = d["dlogits"] / float(b[4].double().abs().max())
        d["loss"] = abs(float(la) - float(lb))
        print("cross-check %s %d^3 B=%d %s plan=%s: %s" % (dt, D, B, env, ops.cluster_plan(x, n_vol, D, D, D), d), flush=True)
    finally:
        for k in (env or {}):
            del os.environ[k]


if __name__ == "__main__":
    quick = "quick" in sys.argv
    f32, bf = torch.float32, torch.bfloat16
    if "lead" in sys.argv:
        # streamed form of the headline shape: cluster size x lead (what the L2 must hold) x ring split
        S, C, L, R = "HIPE_CLUSTER_STREAM", "HIPE_CLUSTER_C", "HIPE_CLUSTER_LEAD", "HIPE_CLUSTER_RSLOTS"
        st = lambda c, l=None, r=None: dict({S: "1", C: str(c)}, **({L: str(l)} if l else {}), **({R: str(r)} if r else {}))
        if "lead2" in sys.argv:
            a6, b6, c6, d6 = st(6, 20, 7), st(6, 22, 7), st(6, 21, 6), st(6, 20, 8)
            case(f32, 64, 64, [{}, a6, {}, b6, {}, c6, {}, d6, {}, a6, {}, b6, {}, c6, {}, d6, {}], False)
            case(f32, 64, 32, [{}, a6, {}, b6, {}, a6, {}], False)
            sys.exit(0)
        case(f32, 64, 64, [{}, st(8, 20), st(8, 22), st(8, 23), {}, st(6, 18), st(6, 19), st(6, 21), st(6, 22), {}, st(6, 20, 7), st(6, 20, 9),
                           st(5, 21), st(5, 23), st(8, 21, 7), {}], False)
        sys.exit(0)
    if "pf" in sys.argv:
        # resident form: L2 bulk prefetch of the stages the producer cannot load yet (HIPE_CLUSTER_PF stages ahead;
        # 0 = off, the library's default is 4 for fp32 56^3 and single-CTA bf16 plans)
        P, S0 = "HIPE_CLUSTER_PF", {"HIPE_CLUSTER_STREAM": "0"}
        case(bf, 64, 128, [{P: "0"}, {P: "2"}, {P: "4"}, {P: "8"}, {P: "0"}], False)
        case(bf, 64, 512, [{P: "0"}, {P: "4"}, {P: "0"}], False)
        case(f32, 56, 32, [{P: "0"}, {P: "2"}, {P: "4"}, {P: "8"}, {P: "0"}], False)
        case(f32, 32, 256, [{P: "0"}, {P: "2"}, {P: "4"}, {P: "0"}], False)
        case(f32, 64, 64, [{}, dict(S0, **{P: "0"}), dict(S0, **{P: "4"}), dict(S0, **{P: "8"}), {}], False)
        case(bf, 32, 256, [{P: "0"}, {P: "2"}, {P: "4"}, {P: "0"}], False)
        sys.exit(0)
    if "bfc" in sys.argv:
        # bf16 64^3: cluster sizes of the resident form
        C = "HIPE_CLUSTER_C"
        case(bf, 64, 128, [{}, {C: "6"}, {C: "8"}, {}, {C: "5"}, {C: "3"}, {}], False)
        case(bf, 64, 512, [{}, {C: "6"}, {C: "8"}, {}], False)
        sys.exit(0)
    if "peek" in sys.argv:
        S8 = {"HIPE_CLUSTER_STREAM": "1", "HIPE_CLUSTER_C": "8"}
        case(f32, 96, 32, [{}, {}], False)
        case(f32, 64, 64, [S8, {}, S8, {}], False)
        sys.exit(0)
    if "str" in sys.argv:
        # streamed shares (two rings, second read through L2) against the shared-memory-resident default
        S, C, L, R = "HIPE_CLUSTER_STREAM", "HIPE_CLUSTER_C", "HIPE_CLUSTER_LEAD", "HIPE_CLUSTER_RSLOTS"
        st = lambda c, l=None, r=None: dict({S: "1", C: str(c)}, **({L: str(l)} if l else {}), **({R: str(r)} if r else {}))
        case(f32, 64, 64, [{}, st(8), st(8, 15), st(8, 18), st(8, 24), st(8, 21, 6), st(8, 21, 9), st(8, 18, 6), st(6, 17), st(6, 20), {}], False)
        case(bf, 64, 128, [{}, st(8), st(8, 12), st(8, 21), st(6), st(6, 14), st(4, 15), st(4, 19), st(4, 19, 6), {}], False)
        case(bf, 64, 512, [{}, st(8), st(4, 17), {}], False)
        case(f32, 56, 32, [{}, st(8), st(8, 14), st(8, 26), st(6, 20), {}], False)
        case(f32, 32, 256, [{}, st(1, 14), st(1, 18), st(1, 12, 7), st(2, 12), {}], False)
        case(f32, 64, 32, [{}, st(8), st(8, 18), {}], False)
        sys.exit(0)
    if "96" in sys.argv:
        cross_check(f32, 96, 4)
        cross_check(f32, 96, 3, {"HIPE_CLUSTER_C": "6"})
        cross_check(f32, 64, 4, {"HIPE_CLUSTER_STREAM": "1"})
        cross_check(bf, 64, 4, {"HIPE_CLUSTER_STREAM": "1"})
        cross_check(f32, 80, 2, {})
        R, L, P, C = "HIPE_CLUSTER_RSLOTS", "HIPE_CLUSTER_LEAD", "HIPE_CLUSTER_POL1", "HIPE_CLUSTER_C"
        case(f32, 96, 32, [{}, {L: "20"}, {L: "22"}, {L: "25"}, {L: "28"}, {L: "36"}, {L: "25", R: "8"}, {L: "25", R: "12"}, {L: "28", P: "0"},
                           {C: "8", L: "40"}, {}])
        case(f32, 96, 64, [{}, {L: "25"}])
        S8 = {"HIPE_CLUSTER_STREAM": "1", C: "8"}
        case(f32, 64, 64, [{}, S8, {}, S8, {}, S8], False)
        sys.exit(0)
    if "56" in sys.argv:
        case(f32, 56, 32, [{}, {"HIPE_CLUSTER_GENERIC": "1"}, {"HIPE_CLUSTER_C": "4"}, {"HIPE_CLUSTER_C": "5"}, {"HIPE_CLUSTER_C": "8"}])
        case(f32, 56, 96, [{}, {"HIPE_CLUSTER_GENERIC": "1"}, {"HIPE_CLUSTER_C": "4"}, {"HIPE_CLUSTER_C": "5"}, {"HIPE_CLUSTER_C": "8"}])
        case(bf, 56, 96, [{}, {"HIPE_CLUSTER_C": "3"}, {"HIPE_CLUSTER_C": "4"}])
        case(f32, 64, 64, [{}])
        sys.exit(0)
    if "wt2" in sys.argv:
        W = {"HIPE_CLUSTER_WT2": "1"}
        case(bf, 64, 128, [{}, W, dict(W, HIPE_CLUSTER_C="3"), dict(W, HIPE_CLUSTER_C="5"), dict(W, HIPE_CLUSTER_C="6")])
        case(bf, 64, 512, [{}, W, dict(W, HIPE_CLUSTER_C="3")])
        case(bf, 64, 64, [{}, W])
        case(f32, 64, 64, [{}, W, dict(W, HIPE_CLUSTER_C="8")])
        case(bf, 56, 96, [{}, W, dict(W, HIPE_CLUSTER_C="3")])
        case(bf, 32, 256, [{}, W])
        sys.exit(0)
    if "sym" in sys.argv:
        S1, S0 = {"HIPE_CLUSTER_SYM": "1"}, {"HIPE_CLUSTER_SYM": "0"}
        case(bf, 64, 128, [S1, S0, dict(S1, HIPE_CLUSTER_C="4"), dict(S0, HIPE_CLUSTER_C="4"), dict(S1, HIPE_CLUSTER_C="5"),
                           dict(S1, HIPE_CLUSTER_C="6"), dict(S1, HIPE_CLUSTER_C="8"), dict(S1, HIPE_CLUSTER_C="4", HIPE_CLUSTER_POLY="1")])
        case(bf, 64, 512, [S1, S0, dict(S1, HIPE_CLUSTER_C="4"), dict(S1, HIPE_CLUSTER_C="6")])
        case(f32, 64, 64, [S0, S1, dict(S1, HIPE_CLUSTER_C="8"), dict(S1, HIPE_CLUSTER_C="7")])
        case(f32, 56, 32, [S0, S1, dict(S1, HIPE_CLUSTER_C="8")])
        case(bf, 56, 96, [S1, S0, dict(S1, HIPE_CLUSTER_C="3"), dict(S1, HIPE_CLUSTER_C="4")])
        sys.exit(0)
    case(f32, 64, 64, [{}, {"HIPE_CLUSTER_STASYNC": "0"}, {"HIPE_CLUSTER_C": "8"}, {"HIPE_CLUSTER_C": "5"}])
    case(bf, 64, 128, [{}, {"HIPE_CLUSTER_POLY": "1"}, {"HIPE_CLUSTER_C": "4"}])
    if not quick:
        case(bf, 64, 512, [{}, {"HIPE_CLUSTER_C": "4"}])
        case(bf, 64, 64, [{}])
        case(f32, 56, 32, [{}, {"HIPE_CLUSTER_C": "4"}, {"HIPE_CLUSTER_C": "8"}])
        case(f32, 56, 96, [{}, {"HIPE_CLUSTER_C": "4"}, {"HIPE_CLUSTER_C": "8"}])
        case(f32, 64, 32, [{}])
        case(f32, 48, 96, [{}])
        case(f32, 32, 256, [{}, {"HIPE_CLUSTER_C": "1"}])
