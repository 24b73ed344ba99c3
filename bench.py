"""Benchmark of the integral-regression hot path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = one pass of the path over one batch: `loss = JointLocationLoss()(heatmaps, gt, vis);
loss.backward()` through the drop-in module on `--batch` samples of 21 x 64^3 fp32 logits
(BASELINE.json configs[1]) per GPU.  The module picks the launch plan itself (ops.fused_plan):
at this geometry ONE persistent launch that reads every volume once and writes its gradient once
(hipe_joint_loss_resident); the forward + recompute-backward pair is timed in the same run and
reported under `other_plan`.  Prints ONE JSON line (rank 0):

  value      volumes/s, whole job, inputs resident in HBM, CUDA events, max over ranks
  e2e        the same metric through the C-ABI host entry point hipe_joint_location_loss_host:
             pinned HOST logits/targets in, loss + joints back to the host, the gradient left in
             a device buffer where the head's backward would consume it
  roofline   the dominant kernel (single launch or backward: reads V, writes V per volume), timed
             live with CUDA events on the launching stream inside the timed region
  cpu_baseline  the reference's PyTorch-eager op sequence (oracle/integral_torch.py, a port:
             /root/reference does not travel) on this box's host cores, bounded sample
  --impl reference  times that CPU port alone on the same workload definition.
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="samples per GPU (weak scaling)")
    ap.add_argument("--dim", type=int, default=64)
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--fused", action="store_true", help="force the single-launch loss+grad path as the primary arm")
    ap.add_argument("--two-launch", action="store_true", help="force forward + recompute backward as the primary arm")
    ap.add_argument("--no-alt", action="store_true", help="skip timing the other launch plan for comparison")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-batch", type=int, default=8)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    return ap.parse_args()


def workload_name(a):
    return "JointLocationLoss fwd+bwd (soft-argmax + L1 integral loss + heatmap gradient), %s logits, batch %d/GPU, 21x%d^3" % (
        "fp32" if a.dtype == "f32" else "bf16", a.batch, a.dim)


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


# ----------------------------------------------------------------------------------------
# CPU leg: the reference's op sequence on the host cores (oracle port; checker code, timed
# only here)
# ----------------------------------------------------------------------------------------
def cpu_reference_rate(a, steps=None, warmup=2, budget_s=None):
    import torch
    from oracle import integral_torch as OT
    cores = host_cores()
    torch.set_num_threads(cores)
    B, D = a.cpu_batch, a.dim
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(B, J * D, D, D, generator=g)
    gt = torch.rand(B, 3 * J, generator=g, dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    for _ in range(warmup):
        OT.fwd_bwd_step(x, gt, vis)
    times = []
    t_begin = time.perf_counter()
    while True:
        t0 = time.perf_counter()
        OT.fwd_bwd_step(x, gt, vis)
        times.append(time.perf_counter() - t0)
        if steps is not None and len(times) >= steps:
            break
        if steps is None and (time.perf_counter() - t_begin >= budget_s or len(times) >= 200):
            break
    total = sum(times)
    return {"value": B * J * len(times) / total, "unit": UNIT, "cores": cores, "kind": "port",
            "threads": torch.get_num_threads(),
            "sample": "%d steps of batch %d (%d volumes each) of the same fp32 21x%d^3 fwd+bwd workload, "
                      "PyTorch-eager op sequence of common/nets/loss.py on CPU" % (len(times), B, B * J, D),
            "ms_per_step": 1e3 * total / len(times)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_rate(a, steps=a.steps, warmup=a.warmup)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(a), "timed_sample": r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


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
            time.sleep(0.002)

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
def synth_logits(torch, B, D, dtype, device, seed):
    """Peaked regime of SURVEY.md 8d generated on the device: N(0,1) noise plus a 12-high, sigma-2
    Gaussian bump at a uniform random centre per volume."""
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


def run_ours(a):
    import torch
    import torch.distributed as dist
    from hipe_b200 import _lib, ops, parallel
    from nets import loss as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: hipe_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()                        # raises if the extension is missing

    B, D = a.batch, a.dim
    dt = torch.float32 if a.dtype == "f32" else torch.bfloat16
    esz = 4 if a.dtype == "f32" else 2
    n_vol = B * J
    V = D ** 3
    NBUF = 3                                 # rotate inputs: no step re-reads what the last one left in L2
    xs = [synth_logits(torch, B, D, dt, dev, 1234 + rank * 16 + i).requires_grad_(True) for i in range(NBUF)]
    g = torch.Generator(device=dev).manual_seed(99 + rank)
    gt = torch.rand(B, 3 * J, generator=g, device=dev, dtype=torch.float64) - 0.5
    vis = torch.ones_like(gt)
    fused_flag = True if a.fused else (False if a.two_launch else None)     # None = the library's own choice

    def make_crit(fused):
        if world > 1:
            c = parallel.ShardedJointLocationLoss(global_batch=B * world)
            if fused is False:
                c.local_fn = lambda h, g_, v_: parallel._local_cuda_loss(h, g_, v_)   # never the fused launch
            return c
        return L.JointLocationLoss(fused=fused)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(crit, steps, warm, sample_clocks):
        """W untimed + K timed steps of crit; CUDA events, barrier on both sides, max over ranks."""
        def step(i):
            x = xs[i % NBUF]
            x.grad = None
            loss = crit(x, gt, vis)
            loss.backward()
            return loss
        for i in range(warm):
            step(i)
        barrier()
        sampler = ClockSampler(local) if sample_clocks else None
        if sampler:
            sampler.start()
        events = {}
        launches0 = ops.LAUNCHES
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        h0 = time.perf_counter()
        for i in range(steps):
            # per-kernel event pairs on every 4th step only: they cost host time, and at ~0.5 ms of
            # kernels per step the enqueue rate matters
            ops.KERNEL_EVENTS = events if i % 4 == 0 else None
            last = step(i)
        ops.KERNEL_EVENTS = None
        host_ms[0] = 1e3 * (time.perf_counter() - h0) / steps      # host time to ENQUEUE one step
        e1.record()
        barrier()
        if sampler:
            sampler.stop_flag.set()
        ms_ = e0.elapsed_time(e1)
        n_launch = ops.LAUNCHES - launches0
        kern_ = {k: sum(s_.elapsed_time(e_) for s_, e_ in v) / len(v) for k, v in events.items()}
        t_ = torch.tensor([ms_], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_, op=dist.ReduceOp.MAX)
        return float(t_.item()), kern_, n_launch, float(last.item()), sampler

    crit = make_crit(fused_flag)
    ms, kern, launches, loss_value, sampler = timed(crit, a.steps, max(a.warmup, 3), True)
    value = world * n_vol * a.steps / (ms / 1e3)
    host_enqueue_ms = host_ms[0]
    if loss_value != loss_value or abs(loss_value) == float("inf"):
        raise RuntimeError("bench: the timed steps produced a non-finite loss (%r)" % loss_value)
    single_launch = ("joint_loss_resident" in kern) or ("joint_loss_fused" in kern)
    # the other launch plan, same buffers, same K: reported beside the primary, never as `value`
    alt = None
    if not a.no_alt and not (a.fused or a.two_launch) and (world == 1 or single_launch):
        ms2, kern2, _, loss2, _ = timed(make_crit(not single_launch), a.steps, max(a.warmup, 3), False)
        alt = {"plan": "two launches (forward + recompute backward)" if single_launch else "single launch (loss + gradient)",
               "value": world * n_vol * a.steps / (ms2 / 1e3), "unit": UNIT, "ms_per_step": ms2 / a.steps,
               "kernels_ms": kern2, "loss": loss2,
               "step_hbm_GBs": n_vol * ((3 if single_launch else 2) * V * esz) / (ms2 / a.steps) / 1e6}

    # ---- roofline of the dominant kernel ----
    peak, peak_src = FALLBACK_HBM_GBS, "fallback"
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs, torch copy)"
    except Exception:
        pass
    traffic = None
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = prof.get(("resident" if "joint_loss_resident" in kern else "fused") if single_launch else "softargmax_bwd", {}).get("%s_%d_B%d" % (a.dtype, D, B))
    except Exception:
        pass
    if single_launch:
        dom = "joint_loss_resident" if "joint_loss_resident" in kern else "joint_loss_fused"
        per_vol = 2 * V * esz                                    # one HBM read + one write per volume
    else:
        dom, per_vol = "softargmax_bwd", 2 * V * esz             # read V, write V
    dom_ms = kern.get(dom)
    achieved = n_vol * per_vol / (dom_ms / 1e3) / 1e9 if dom_ms else None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None, "traffic": traffic,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": n_vol * per_vol,
                "launch_ms": dom_ms}
    step_bytes = n_vol * ((2 if single_launch else 3) * V * esz)
    kernels = {k: {"ms": v} for k, v in kern.items()}
    if "softargmax_fwd" in kern:
        kernels["softargmax_fwd"]["GBs"] = n_vol * V * esz / kern["softargmax_fwd"] / 1e6
    if "softargmax_bwd" in kern:
        kernels["softargmax_bwd"]["GBs"] = n_vol * 2 * V * esz / kern["softargmax_bwd"] / 1e6

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
        barrier()
        t0 = time.perf_counter()
        for _ in range(k2):
            host_step()
        torch.cuda.synchronize()
        dt_s = time.perf_counter() - t0
        t = torch.tensor([dt_s], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt_s = float(t.item())
        h2d = hx.numel() * esz + 2 * hgt.numel() * 8
        d2h = hcoords.numel() * 4 + 8
        e2e = {"value": world * n_vol * k2 / dt_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": k2, "ms_per_step": 1e3 * dt_s / k2,
               "h2d_GBs": h2d / (dt_s / k2) / 1e9,
               "api": "hipe_joint_location_loss_host (C ABI): pinned host logits -> chunked H2D overlapped with "
                      "fwd/L1/bwd kernels -> loss + joints to host, gradient left on device",
               "loss_check": abs(hloss.value - float(crit_single(L, xs[0], gt, vis))) if world == 1 else None}
        # launches inside one host call: 3 per chunk + the final fixed-order loss reduction
    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_reference_rate(a, budget_s=a.cpu_seconds)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": workload_name(a), "global_batch": B * world, "volumes_per_step": n_vol * world,
                       "volume": [D, D, D], "parallelism": ("batch-sharded x%d, one exchange per step (%s)" % (
                           world, "hipe_peer_exchange: one kernel over NVLink peer memory" if getattr(crit, "exchange", None) == "peer"
                           else "NCCL all-gather of the packed records")) if world > 1 else "single GPU",
                       "l2": "inputs larger than L2: %d rotating %.2f GB logit buffers per GPU" % (NBUF, n_vol * V * esz / 1e9),
                       "regime": "peaked (A=12, sigma=2) + N(0,1)", "launch_plan": ("single launch: loss + gradient, one HBM read + one write per volume (%s)" % dom)
                       if single_launch else "two launches: forward (read V) + recompute backward (read V, write V)",
                       "algorithmic_bytes_per_volume": (2 if single_launch else 3) * V * esz},
            "step_hbm_GBs": step_bytes / (ms / a.steps) / 1e6,
            "step_frac_of_peak": step_bytes / (ms / a.steps) / 1e6 / peak,
            "roofline": roofline, "kernels": kernels, "other_plan": alt, "cpu_baseline": cpu, "e2e": e2e,
            "gpu_launches": launches, "host_enqueue_ms_per_step": host_enqueue_ms, "clocks": sampler.summary(), "loss": loss_value}
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
