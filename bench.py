"""Benchmark of the integral-regression hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the path over one batch: `loss = JointLocationLoss()(heatmaps, gt, vis);
loss.backward()` through the drop-in module.  The headline workload is BASELINE.json configs[1]
(fp32, 64 samples x 21 x 64^3 per GPU, weak scaling); the module picks the launch plan itself
(ops.fused_plan).  R back-to-back blocks of exactly K steps are timed (barrier + synchronize on
both sides of each block, CUDA events, max over ranks); the MEDIAN block is the reported one, min
and max stand beside it.  Prints ONE JSON line (rank 0):

  value      volumes/s, whole job, inputs resident in HBM
  roofline   the dominant kernel (the single launch: reads V, writes V per volume), timed live with
             CUDA events on the launching stream inside the timed region
  e2e        the same metric through the C-ABI host entry point hipe_joint_location_loss_host:
             pinned HOST logits/targets in, loss + joints back to the host
  configs    the other BASELINE configs in the same process, same method, each with its dominant
             kernel's fraction of the HBM roofline:
               C3  bf16 logits, global batch 512 STRONG-scaled over the ranks (512/N samples each)
               C4  fp32 21 x 96^3, 32 samples per rank
               C5  eval: 3960 samples sharded over the ranks, soft-argmax forward + patch->camera
                   chain + batched Procrustes + EvalUtil measures, with the reference's per-sample
                   numpy Procrustes loop timed on the host beside it
  cpu_baseline / --impl reference
             the reference's own common/nets/loss.py (unmodified; build() drops it into the
             git-ignored oracle/_ref/ where /root/reference is mounted) on this box's host cores at
             the SAME batch and input regime; the oracle port when that file did not travel.
  sharded_parity_max_err  (N > 1) one step of the sharded criterion against the single-rank
             criterion on the gathered batch, outside the timed region.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "hand-integral-pose-estimation_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

J = 21
METRIC = "heatmap volumes/sec (21x64^3 soft-argmax fwd+bwd)"
UNIT = "volumes/s"
FALLBACK_HBM_GBS = 6650.0          # /opt/skills/guides/B200_PROFILING.md
REGIME = "peaked (A=12, sigma=2) + N(0,1)"
SINGLE = ("joint_loss_cluster", "joint_loss_resident", "joint_loss_fused")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--blocks", type=int, default=9, help="timed blocks of --steps steps; the median is reported")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="samples per GPU (weak scaling)")
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--fused", action="store_true", help="force the single-launch loss+grad path as the primary arm")
    ap.add_argument("--two-launch", action="store_true", help="force forward + recompute backward as the primary arm")
    ap.add_argument("--no-alt", action="store_true", help="skip timing the other launch plan for comparison")
    ap.add_argument("--no-configs", action="store_true", help="skip BASELINE configs C3 / C4 / C5")
    ap.add_argument("--configs", default="C3,C4,C5")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-batch", type=int, default=None, help="CPU leg batch (default: --batch)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    return ap.parse_args()


def workload_name(a):
    return "JointLocationLoss fwd+bwd (soft-argmax + L1 integral loss + heatmap gradient), %s logits, batch %d/GPU, 21x%d^3" % (
        "fp32" if a.dtype == "f32" else "bf16", a.batch, a.dim)


def config_of(a, world):
    """The workload definition; identical for both arms."""
    return {"workload": workload_name(a), "global_batch": a.batch * world, "volumes_per_step": a.batch * J * world,
            "volume": [a.dim, a.dim, a.dim], "regime": REGIME,
            "parallelism": "batch-sharded x%d" % world if world > 1 else "single GPU"}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), \
            "measured (MEASURED_PEAKS.json hbm_gbs, torch copy)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


def median(v):
    s = sorted(v)
    return s[len(s) // 2]


def synth_logits(torch, B, D, dtype, device, seed):
    """Peaked regime of SURVEY.md 8d: N(0,1) noise plus a 12-high, sigma-2 Gaussian bump at a
    uniform random centre per volume.  Works on CUDA and CPU devices."""
    g = torch.Generator(device=device).manual_seed(seed)
    x = torch.randn(B * J, D, D, D, generator=g, device=device)
    c = torch.rand(B * J, 3, generator=g, device=device) * (D - 1)
    ax = torch.arange(D, device=device, dtype=torch.float32)
    for lo in range(0, B * J, 84):           # bounded temporaries
        hi = min(lo + 84, B * J)
        dz = (ax[None, :] - c[lo:hi, 2:3]) ** 2
        dy = (ax[None, :] - c[lo:hi, 1:2]) ** 2
        dx = (ax[None, :] - c[lo:hi, 0:1]) ** 2
        r2 = dz[:, :, None, None] + dy[:, None, :, None] + dx[:, None, None, :]
        x[lo:hi] += 12.0 * torch.exp(-r2 / 8.0)
    return x.reshape(B, J * D, D, D).to(dtype).contiguous()


# ----------------------------------------------------------------------------------------
# CPU leg: the reference's own loss.py on the host cores (oracle/_ref, written by build() where
# /root/reference is mounted), else the oracle's port of its op sequence.  Checker code, timed
# only here.
# ----------------------------------------------------------------------------------------
def load_reference_loss():
    """-> (step_fn(x, gt, vis) -> loss, kind).  kind "reference": the unmodified
    common/nets/loss.py from oracle/_ref/ (shims: stub `config`/`augment`/`config_panet` modules and
    an identity torch.cuda.comm.broadcast for CPU tensors, loss.py:35-37 -- none touches
    arithmetic); kind "port": oracle/integral_torch.py."""
    import torch
    path = os.path.join(ROOT, "oracle", "_ref", "loss.py")
    if os.path.isfile(path):
        try:
            import importlib.util
            import types
            import torch.cuda.comm  # noqa: F401
            cfg_mod = types.ModuleType("config")
            cfg_mod.cfg = types.SimpleNamespace(_lambda=0, patch_width=224, patch_height=224)
            fh = types.ModuleType("FreiHand_config")
            fh.FreiHandConfig = types.SimpleNamespace(num_joints=J, root_idx=9)   # FreiHand_config.py:60-61
            for name, mod in (("config", cfg_mod), ("augment", types.ModuleType("augment")),
                              ("config_panet", types.ModuleType("config_panet")), ("FreiHand_config", fh)):
                sys.modules.setdefault(name, mod)
            orig = torch.cuda.comm.broadcast

            def broadcast(tensor, devices=None, *, out=None):
                return [tensor] if devices == [None] else orig(tensor, devices=devices, out=out)

            torch.cuda.comm.broadcast = broadcast
            spec = importlib.util.spec_from_file_location("_hipe_ref_loss", path)
            mod = importlib.util.module_from_spec(spec)
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                spec.loader.exec_module(mod)
            crit = mod.JointLocationLoss()

            def step(x, gt, vis):
                x = x.detach().requires_grad_(True)
                loss = crit(x, gt, vis)
                loss.backward()
                return loss.detach()

            return step, "reference"
        except Exception as e:                                    # pragma: no cover
            print("bench: oracle/_ref/loss.py could not be loaded (%r); timing the port" % (e,), file=sys.stderr)
    from oracle import integral_torch as OT
    return (lambda x, gt, vis: OT.fwd_bwd_step(x, gt, vis)[0]), "port"


def cpu_reference_rate(a, steps=None, warmup=2, budget_s=None):
    import torch
    cores = host_cores()
    torch.set_num_threads(cores)
    step, kind = load_reference_loss()
    B = a.cpu_batch or a.batch
    D = a.dim
    x = synth_logits(torch, B, D, torch.float32, torch.device("cpu"), 1234)
    g = torch.Generator().manual_seed(99)
    gt = torch.rand(B, 3 * J, generator=g, dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    for _ in range(warmup):
        step(x, gt, vis)
    times = []
    t_begin = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        loss = step(x, gt, vis)
        times.append(time.perf_counter() - t0)
        if steps is not None and len(times) >= steps:
            break
        if steps is None and (time.perf_counter() - t_begin >= budget_s or len(times) >= 200):
            break
    total = sum(times)
    what = ("the unmodified common/nets/loss.py (JointLocationLoss + backward)" if kind == "reference"
            else "the PyTorch-eager op sequence of common/nets/loss.py (oracle port)")
    return {"value": B * J * len(times) / total, "unit": UNIT, "cores": cores, "kind": kind,
            "threads": torch.get_num_threads(),
            "sample": "%d steps of batch %d (%d volumes each) of the same fp32 21x%d^3 fwd+bwd workload, %s regime, "
                      "%s on CPU" % (len(times), B, B * J, D, REGIME, what),
            "ms_per_step": 1e3 * total / len(times), "ms_per_step_median": 1e3 * median(times),
            "loss": float(loss)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    r = cpu_reference_rate(a, steps=a.steps, warmup=a.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_of(a, world),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "loss": r["loss"]}
    print(json.dumps(line), flush=True)


def procrustes_cpu_baseline(X, Y):
    """The reference's per-sample loop (data/FreiHand/FreiHand.py:707-767: one
    compute_similarity_transform call per sample, single-threaded numpy/LAPACK) on this host:
    the function's own source (common/augment.py:18-77, from oracle/_ref/augment.py) when it
    travelled, else the oracle's restatement called one sample at a time."""
    import numpy as np
    fn, kind = None, "port"
    path = os.path.join(ROOT, "oracle", "_ref", "augment.py")
    if os.path.isfile(path):
        try:
            with open(path) as f:
                lines = f.readlines()
            ns = {"np": np}
            exec("".join(lines[17:77]), ns)
            fn, kind = ns["compute_similarity_transform"], "reference"
        except Exception:
            fn = None
    if fn is None:
        from oracle import procrustes_np as P
        fn = lambda x, y, compute_optimal_scale: tuple(v[0] for v in P.similarity_transform(x[None], y[None], compute_optimal_scale))  # noqa: E731
    n = len(X)
    t0 = time.perf_counter()
    err = 0.0
    for i in range(n):
        Z = fn(X[i], Y[i], compute_optimal_scale=True)[1]
        err += float(np.sqrt(((Z - X[i]) ** 2).sum(axis=1)).sum())
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": "samples/s", "cores": 1, "kind": kind, "us_per_sample": 1e6 * dt / n,
            "sample": "%d samples, one compute_similarity_transform call each + PA-MPJPE sum, single-threaded as "
                      "the reference runs it" % n, "pa_mpjpe_m": err / (n * X.shape[1])}


# ----------------------------------------------------------------------------------------
# clocks during the timed region
# ----------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.stop_flag = threading.Event()
        self.sm, self.bits, self.power = [], 0, []
        self.max_mhz = None
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        while not self.stop_flag.is_set():
            try:
                self.sm.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                self.bits |= int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                self.power.append(self.nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0)
            except Exception:
                try:
                    self.bits |= int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                except Exception:
                    pass
            time.sleep(0.004)

    def summary(self):
        if not self.ok or not self.sm:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        s = sorted(self.sm)
        reasons = [n for b, n in self.REASONS.items() if self.bits & b and n != "gpu_idle"]
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(s),
                "power_w_max": max(self.power) if self.power else None}


# ----------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------
class Ctx:
    """Process-wide state of one bench run."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise RuntimeError("bench.py needs a CUDA device: hipe_b200 has no CPU path")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.peak, self.peak_src = peak_gbs()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def time_steps(cx, step, steps, warm, blocks, sampler=None, events_every=4):
    """`warm` untimed steps, then `blocks` blocks of exactly `steps` steps, each bracketed by a
    barrier + synchronize on both sides and CUDA events; block times are the max over ranks.
    Returns (sorted block ms, per-kernel mean ms from the event pairs, launches per step, host
    enqueue ms per step, last loss)."""
    from hipe_b200 import ops
    torch = cx.torch
    for i in range(warm):
        step(i)
    cx.barrier()
    if sampler:
        sampler.start()
    events, times, host = {}, [], []
    launches0 = ops.LAUNCHES
    last = None
    n = 0
    for b in range(blocks):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h0 = time.perf_counter()
        for i in range(steps):
            # per-kernel event pairs on every 4th step only: they cost host time, and at ~0.5 ms of
            # kernels per step the enqueue rate matters
            ops.KERNEL_EVENTS = events if (events_every and n % events_every == 0) else None
            last = step(n)
            n += 1
        ops.KERNEL_EVENTS = None
        host.append(1e3 * (time.perf_counter() - h0) / steps)
        e1.record()
        cx.barrier()
        times.append(cx.max_over_ranks(e0.elapsed_time(e1)))
    if sampler:
        sampler.stop_flag.set()
    launches = (ops.LAUNCHES - launches0) / float(blocks * steps)
    kern = {k: sum(s_.elapsed_time(e_) for s_, e_ in v) / len(v) for k, v in events.items()}
    return sorted(times), kern, launches, median(host), (float(last.item()) if last is not None else None)


def dominant(kern):
    for k in SINGLE:
        if k in kern:
            return k, True
    return ("softargmax_bwd" if "softargmax_bwd" in kern else None), False


def traffic_lookup(kernel, dtype, D, B):
    """Measured DRAM bytes per launch of `kernel` at this shape from profiles/traffic.json (ncu
    --set full captures; the entry names the capture and commit it came from), or None."""
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = prof.get(kernel, {}).get("%s_%d_B%d" % (dtype, D, B))
        if ent is None:
            return None, None
        if isinstance(ent, dict):
            return ent.get("bytes"), {k: ent.get(k) for k in ("capture", "commit")}
        return ent, {"capture": prof.get("_default_capture"), "commit": prof.get("_default_commit")}
    except Exception:
        return None, None


KNAME = {"joint_loss_cluster": "cluster", "joint_loss_resident": "resident", "joint_loss_fused": "fused",
         "softargmax_bwd": "softargmax_bwd", "softargmax_fwd": "softargmax_fwd"}


def train_case(cx, a, B, D, dtype_name, fused_flag, steps, warm, blocks, sample_clocks=False, want_alt=False,
               nbuf=3, seed0=1234):
    """Time `loss = crit(x, gt, vis); loss.backward()` on B samples per rank of 21 x D^3 logits."""
    from hipe_b200 import ops, parallel
    from nets import loss as L
    torch = cx.torch
    dt = torch.float32 if dtype_name == "f32" else torch.bfloat16
    esz = 4 if dtype_name == "f32" else 2
    n_vol, V = B * J, D ** 3
    xs = [synth_logits(torch, B, D, dt, cx.dev, seed0 + cx.rank * 16 + i).requires_grad_(True) for i in range(nbuf)]
    g = torch.Generator(device=cx.dev).manual_seed(99 + cx.rank)
    gt = torch.rand(B, 3 * J, generator=g, device=cx.dev, dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)

    def make_crit(fused):
        if cx.world > 1:
            c = parallel.ShardedJointLocationLoss(global_batch=B * cx.world)
            if fused is False:
                c.local_fn = lambda h, g_, v_: parallel._local_cuda_loss(h, g_, v_)   # never the fused launch
            return c
        return L.JointLocationLoss(fused=fused)

    def stepper(crit):
        def step(i):
            x = xs[i % nbuf]
            x.grad = None
            loss = crit(x, gt, vis)
            loss.backward()
            return loss
        return step

    crit = make_crit(fused_flag)
    sampler = ClockSampler(cx.local) if sample_clocks else None
    times, kern, launches, host_ms, loss_value = time_steps(cx, stepper(crit), steps, warm, blocks, sampler)
    if loss_value != loss_value or abs(loss_value) == float("inf"):
        raise RuntimeError("bench: the timed steps produced a non-finite loss (%r)" % loss_value)
    ms = median(times) / steps
    dom, single = dominant(kern)
    per_vol = 2 * V * esz                        # the dominant launch reads V and writes V per volume
    dom_ms = kern.get(dom) if dom else None
    achieved = n_vol * per_vol / (dom_ms / 1e3) / 1e9 if dom_ms else None
    traffic, tsrc = traffic_lookup(KNAME.get(dom, dom), dtype_name, D, B)
    step_bytes = n_vol * ((2 if single else 3) * V * esz)
    out = {
        "ms_per_step": ms, "ms_per_step_min": times[0] / steps, "ms_per_step_max": times[-1] / steps,
        "blocks": blocks, "steps_per_block": steps,
        "value": cx.world * n_vol / (ms / 1e3), "unit": UNIT, "samples_per_rank": B, "volumes_per_rank": n_vol,
        "dtype": dtype_name, "volume": [D, D, D],
        "launch_plan": ("single launch: loss + gradient, one HBM read + one write per volume (%s)" % dom) if single
        else "two launches: forward (read V) + recompute backward (read V, write V)",
        "algorithmic_bytes_per_volume": (2 if single else 3) * V * esz,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": cx.peak, "unit": "GB/s",
                     "frac": (achieved / cx.peak) if achieved else None, "traffic": traffic, "traffic_source": tsrc,
                     "peak_source": cx.peak_src, "algorithmic_bytes_per_launch": n_vol * per_vol, "launch_ms": dom_ms},
        "step_hbm_GBs": step_bytes / ms / 1e6, "step_frac_of_peak": step_bytes / ms / 1e6 / cx.peak,
        "kernels_ms": kern, "gpu_launches_per_step": launches, "host_enqueue_ms_per_step": host_ms,
        "loss": loss_value, "exchange": getattr(crit, "exchange", None),
        "l2": "inputs larger than L2: %d rotating %.2f GB logit buffers per GPU" % (nbuf, n_vol * V * esz / 1e9),
    }
    if want_alt and (cx.world == 1 or single):
        t2, k2, _, _, loss2 = time_steps(cx, stepper(make_crit(not single)), steps, warm, max(3, blocks // 3))
        ms2 = median(t2) / steps
        out["other_plan"] = {
            "plan": "two launches (forward + recompute backward)" if single else "single launch (loss + gradient)",
            "value": cx.world * n_vol / (ms2 / 1e3), "unit": UNIT, "ms_per_step": ms2, "kernels_ms": k2, "loss": loss2,
            "step_hbm_GBs": n_vol * ((3 if single else 2) * V * esz) / ms2 / 1e6}
    return out, (xs, gt, vis, crit, sampler)


def sharded_parity(cx, xs, gt, vis, crit):
    """One step of the sharded criterion against the single-rank module on the gathered batch
    (rank 0 holds all shards once, outside any timed region): relative loss difference and the
    worst gradient difference of rank 0's shard relative to each volume's largest gradient."""
    from nets import loss as L
    torch, dist = cx.torch, cx.dist
    x = xs[0]
    x.grad = None
    loss = crit(x, gt, vis)
    loss.backward()
    g_shard = x.grad.detach().clone()
    coords_sh = crit.coords.detach().clone() if getattr(crit, "coords", None) is not None else None
    parts = [torch.empty_like(x.detach()) for _ in range(cx.world)] if cx.rank == 0 else None
    dist.gather(x.detach().contiguous(), parts, dst=0)
    gts = [torch.empty_like(gt) for _ in range(cx.world)] if cx.rank == 0 else None
    dist.gather(gt, gts, dst=0)
    res = None
    if cx.rank == 0:
        xa = torch.cat(parts, dim=0).requires_grad_(True)
        del parts
        ga = torch.cat(gts, dim=0)
        crit1 = L.JointLocationLoss()
        l1 = crit1(xa, ga, torch.ones_like(ga))
        l1.backward()
        b = x.shape[0]
        ref = xa.grad[:b].reshape(b * J, -1)
        got = g_shard.reshape(b * J, -1)
        scale = ref.abs().amax(dim=1, keepdim=True)
        gerr = float(((got.float() - ref.float()).abs() / (scale.float() + 1e-30)).max().item())
        lerr = abs(float(loss.item()) - float(l1.item())) / abs(float(l1.item()))
        cerr = None
        if coords_sh is not None:
            with torch.no_grad():
                c1 = L.softmax_integral_tensor(xa.detach(), J, xa.shape[-1], xa.shape[-2], xa.shape[-3] // J)
            cerr = float((coords_sh.to(c1.dtype) - c1).abs().max().item())
        res = {"loss_rel": lerr, "grad_rel_to_volume_max": gerr, "coords_abs": cerr,
               "max": max(lerr, gerr, cerr or 0.0), "global_batch": int(xa.shape[0])}
        del xa, ga
    cx.barrier()
    return res


def eval_case(cx, samples=3960, batch=32, D=64, dtype_name="f32", reps=5, cpu=True):
    """BASELINE configs[4]: the FreiHAND-sized evaluation pass, samples sharded over the ranks:
    soft-argmax forward per batch of 32 (main/config.py:41) -> patch->camera chain + batched
    Procrustes in one launch -> PA-MPJPE / MPJPE -> EvalUtil measures, all on the device."""
    import numpy as np
    from hipe_b200 import evaluate as E, ops, parallel
    from nets import loss as L
    from tests import synth
    torch, dist = cx.torch, cx.dist
    lo, hi = parallel.shard_bounds(samples, cx.world, cx.rank)
    n_local = hi - lo
    dt = torch.float32 if dtype_name == "f32" else torch.bfloat16
    esz = 4 if dtype_name == "f32" else 2
    xs = [synth_logits(torch, batch, D, dt, cx.dev, 77 + 3 * cx.rank + i) for i in range(3)]
    c = synth.chain_inputs(samples, seed=5)
    cu = {k: torch.as_tensor(np.ascontiguousarray(v[lo:hi])).to(cx.dev) for k, v in c.items()}
    coords_all = torch.empty((max(n_local, 1), 3 * J), dtype=torch.float32, device=cx.dev)
    n_batches = (n_local + batch - 1) // batch

    def one_pass():
        with torch.no_grad():
            for b in range(n_batches):
                s0 = b * batch
                n = min(batch, n_local - s0)
                coords_all[s0:s0 + n] = L.softmax_integral_tensor(xs[b % 3][:n], J, D, D, D)
            # the soft-argmax output goes straight into the geometry (no host round trip); the heat maps
            # are synthetic, so the error values only serve as a GPU-vs-CPU agreement check
            pa, raw, o = E.evaluate_chain(coords_all[:n_local], cu["gt_cam"], cu["tprime"], cu["K"], bbox=cu["bbox"],
                                          R=cu["R"], bone_len=cu["bone_len"])
            m = E.keypoint_measures(o["pa_dist"], 0.0, 0.05, 100)
            if cx.world > 1:
                rec = torch.stack([pa * o["pa_dist"].numel(), raw * o["raw_dist"].numel(),
                                   torch.tensor(float(o["pa_dist"].numel()), dtype=torch.float64, device=cx.dev)])
                pa, raw = parallel.reduce_error_sums(parallel.gather_records(rec))
        return pa, raw, m, o

    for _ in range(2):
        one_pass()
    cx.barrier()
    times, ev = [], {}
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.KERNEL_EVENTS = {}
        e0.record()
        pa, raw, m, o = one_pass()
        e1.record()
        cx.barrier()
        ev, ops.KERNEL_EVENTS = ops.KERNEL_EVENTS, None
        times.append(cx.max_over_ranks(e0.elapsed_time(e1)))
    kern = {k: sum(s.elapsed_time(e) for s, e in v) for k, v in ev.items()}
    ms = median(times)
    fwd_ms = kern.get("softargmax_fwd", 0.0)
    fwd_gbs = (n_local * J * D ** 3 * esz / (fwd_ms / 1e3) / 1e9) if fwd_ms else None
    geo = kern.get("eval_chain_procrustes")
    out = {"metric": "eval samples/sec (soft-argmax fwd + chain + Procrustes + measures, 21x%d^3)" % D,
           "value": samples / (ms / 1e3), "unit": "samples/s", "volumes_per_s": samples * J / (ms / 1e3),
           "samples": samples, "samples_per_rank": n_local, "batch": batch, "dtype": dtype_name,
           "ms_per_pass": ms, "ms_per_pass_min": min(times), "ms_per_pass_max": max(times), "passes": reps,
           "kernel_ms_per_pass": {"softargmax_fwd": fwd_ms, "chain+procrustes": geo,
                                  "measures": kern.get("keypoint_measures"), "sum_f64": kern.get("sum_f64")},
           "roofline": {"bound": "hbm", "kernel": "softargmax_fwd", "achieved": fwd_gbs, "peak": cx.peak,
                        "unit": "GB/s", "frac": (fwd_gbs / cx.peak) if fwd_gbs else None,
                        "algorithmic_bytes_per_pass": n_local * J * D ** 3 * esz},
           "procrustes": {"samples_per_s": (n_local / (geo / 1e3)) if geo else None,
                          "note": "latency-bound (%.1f MB of joints): reported as time only" % (n_local * J * 3 * 8 * 3 / 1e6)},
           "pa_mpjpe_m": float(pa.item()), "mpjpe_m": float(raw.item()),
           "epe_mean_m": float(m[0].item()), "auc": float(m[2].item())}
    if cpu and cx.rank == 0:
        # the reference's loop on the same camera-space joints (copied out once, outside the timed region)
        cam_all = o["cam"].cpu().numpy() if cx.world == 1 else None
        if cam_all is not None:
            base = procrustes_cpu_baseline(c["gt_cam"][lo:hi], cam_all)
            base["gpu_pa_mpjpe_m"] = out["pa_mpjpe_m"]
            base["pa_mpjpe_abs_diff_m"] = abs(base["pa_mpjpe_m"] - out["pa_mpjpe_m"])
            out["cpu_baseline"] = base
            if geo:
                out["procrustes"]["speedup_vs_cpu_loop"] = (n_local / (geo / 1e3)) / base["value"]
    del xs
    return out


def run_ours(a):
    cx = Ctx()
    torch, dist = cx.torch, cx.dist
    from hipe_b200 import _lib
    from nets import loss as L
    world, rank = cx.world, cx.rank
    lib = _lib.load()                        # raises if the extension is missing

    B, D = a.batch, a.dim
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    esz = 4 if a.dtype == "f32" else 2
    n_vol = B * J
    fused_flag = True if a.fused else (False if a.two_launch else None)     # None = the library's own choice
    steps, warm = a.steps, max(a.warmup, 3)
    main, (xs, gt, vis, crit, sampler) = train_case(
        cx, a, B, D, a.dtype, fused_flag, steps, warm, a.blocks, sample_clocks=True,
        want_alt=not a.no_alt and not (a.fused or a.two_launch))
    parity = sharded_parity(cx, xs, gt, vis, crit) if world > 1 else None

    # ---- end to end through the C ABI with host buffers ----
    e2e = None
    if not a.no_e2e:
        import ctypes
        k2 = a.e2e_steps if a.e2e_steps is not None else max(3, min(a.steps, 20))
        hx = torch.empty((B, J * D, D, D), dtype=dt, pin_memory=True)
        hx.copy_(xs[0].detach())
        hgt = gt.cpu().pin_memory()
        hvis = vis.cpu().pin_memory()
        hcoords = torch.empty((B, 3 * J), dtype=torch.float32, pin_memory=True)
        dgrad = torch.empty_like(xs[0])
        hloss = ctypes.c_double()
        code = _lib.F32 if a.dtype == "f32" else _lib.BF16

        def host_step():
            rc = lib.hipe_joint_location_loss_host(hx.data_ptr(), code, B, J, D, D, D, hgt.data_ptr(),
                                                   hvis.data_ptr(), _lib.F64, 1, ctypes.byref(hloss),
                                                   hcoords.data_ptr(), None, dgrad.data_ptr(), 0)
            _lib.check(rc)

        for _ in range(3):
            host_step()
        cx.barrier()
        t0 = time.perf_counter()
        for _ in range(k2):
            host_step()
        torch.cuda.synchronize()
        dt_s = cx.max_over_ranks(time.perf_counter() - t0)
        h2d = hx.numel() * esz + 2 * hgt.numel() * 8
        d2h = hcoords.numel() * 4 + 8
        e2e = {"value": world * n_vol * k2 / dt_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": k2, "ms_per_step": 1e3 * dt_s / k2,
               "h2d_GBs": h2d / (dt_s / k2) / 1e9,
               "api": "hipe_joint_location_loss_host (C ABI): pinned host logits -> chunked H2D overlapped with "
                      "the kernels -> loss + joints to host, gradient left on device",
               "loss_check": abs(hloss.value - float(crit_single(L, xs[0], gt, vis))) if world == 1 else None}
        del hx, dgrad
    del xs, gt, vis, crit
    torch.cuda.empty_cache()

    # ---- the other BASELINE configs, same process, same method ----
    configs = {}
    if not a.no_configs:
        want = [c.strip() for c in a.configs.split(",") if c.strip()]
        ksteps, kblocks = max(4, min(a.steps, 10)), max(3, min(a.blocks, 5))
        if "C3" in want:
            b3 = 512 // world
            r, keep = train_case(cx, a, b3, 64, "bf16", None, ksteps, 3, kblocks, nbuf=2, seed0=4321)
            r["workload"] = "BASELINE configs[2]: bf16 logits, fp32 accumulation, GLOBAL batch 512 (%d samples/rank), 21x64^3" % b3
            r["scaling"] = "strong"
            configs["C3"] = r
            del keep
            torch.cuda.empty_cache()
        if "C4" in want:
            r, keep = train_case(cx, a, 32, 96, "f32", None, ksteps, 3, kblocks, nbuf=2, seed0=5321)
            r["workload"] = "BASELINE configs[3]: fp32 21x96^3, 32 samples/rank (batch 256 on 8 GPUs)"
            r["scaling"] = "weak"
            configs["C4"] = r
            del keep
            torch.cuda.empty_cache()
        if "C5" in want:
            r = eval_case(cx, cpu=not a.no_cpu_baseline)
            r["workload"] = "BASELINE configs[4]: 3960-sample eval pass sharded over the ranks"
            r["scaling"] = "strong"
            configs["C5"] = r
            torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_reference_rate(a, budget_s=a.cpu_seconds)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    cfg = config_of(a, world)
    line = {"metric": METRIC, "value": main["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": warm, "ms_per_step": main["ms_per_step"], "ms_per_step_min": main["ms_per_step_min"],
            "ms_per_step_max": main["ms_per_step_max"], "blocks": a.blocks, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": a.dtype, "data": "synthetic", "config": cfg,
            "details": {"l2": main["l2"], "launch_plan": main["launch_plan"],
                        "algorithmic_bytes_per_volume": main["algorithmic_bytes_per_volume"],
                        "exchange": ("hipe_peer_publish + hipe_peer_collect over NVLink peer memory (both deferred on a side stream)" if main["exchange"] == "peer"
                                     else "NCCL all-gather of the packed records") if world > 1 else None,
                        "timing": "median of %d blocks of %d steps, each block bracketed by barrier + synchronize" % (a.blocks, a.steps),
                        "kernels_note": ("peer_publish and peer_collect run on a side stream behind the step's loss kernel, up to two steps "
                                         "deferred: the collect's event time is mostly the wait for the slowest rank and is off the critical path")
                        if world > 1 else None},
            "step_hbm_GBs": main["step_hbm_GBs"], "step_frac_of_peak": main["step_frac_of_peak"],
            "roofline": main["roofline"], "kernels": {k: {"ms": v} for k, v in main["kernels_ms"].items()},
            "other_plan": main.get("other_plan"), "configs": configs, "cpu_baseline": cpu, "e2e": e2e,
            "sharded_parity_max_err": parity["max"] if parity else None, "sharded_parity": parity,
            "gpu_launches": int(round(main["gpu_launches_per_step"] * a.steps)),
            "gpu_launches_per_step": main["gpu_launches_per_step"],
            "host_enqueue_ms_per_step": main["host_enqueue_ms_per_step"], "clocks": sampler.summary(),
            "loss": main["loss"]}
    print(json.dumps(line), flush=True)


def crit_single(L, x, gt, vis):
    import torch
    with torch.no_grad():
        return L.JointLocationLoss()(x.detach(), gt, vis).item()


def main():
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)


if __name__ == "__main__":
    main()
