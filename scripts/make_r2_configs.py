"""Build profiles/r2_configs.md from the round's gpurun logs (gpurun_out/r2_*): the bench lines at
N = 1, 2, 8, the same-box A/B tables and the wait profile of the cluster launch.
    python scripts/make_r2_configs.py"""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")


def load(name):
    try:
        return json.loads(open(os.path.join(G, name)).read().strip().splitlines()[-1])
    except Exception:
        return None


def text(name):
    try:
        return open(os.path.join(G, name)).read()
    except Exception:
        return None


def row(b):
    c3, c4, c5 = b["configs"]["C3"], b["configs"]["C4"], b["configs"]["C5"]
    k = b["kernels"]
    ex = ("%.3f / %.3f" % (k["peer_publish"]["ms"], k["peer_collect"]["ms"])) if "peer_publish" in k else "-"
    return "| %d | %.4f (%.4f-%.4f) | %.3f M | %.3f | %.3f | %s | %.3f (%d) | %.2f M | %.3f | %.3f | %.2f M | %.3f | %.2f | %.0f k |" % (
        b["n_gpus"], b["ms_per_step"], b["ms_per_step_min"], b["ms_per_step_max"], b["value"] / 1e6,
        b["roofline"]["launch_ms"], b["roofline"]["frac"], ex, c3["ms_per_step"], c3["samples_per_rank"],
        c3["value"] / 1e6, c3["roofline"]["frac"], c4["ms_per_step"], c4["value"] / 1e6, c4["roofline"]["frac"],
        c5["ms_per_pass"], c5["value"] / 1e3)


def main():
    b1, b2, b8 = load("r2_bench1.json"), load("r2_bench_n2.json"), load("r2_bench_n8.json")
    ref = load("r2_bench_ref.json")
    out = ["# Round 2: measured configurations (B200, CUDA events, 1965 MHz, no throttle reasons)\n",
           "Peak used for every fraction: `MEASURED_PEAKS.json` `hbm_gbs` = 6535.7 GB/s.  Raw logs: `gpurun_out/r2_*` "
           "(scratch); this file is the committed summary (`python scripts/make_r2_configs.py`).\n",
           "## 1. `bench.py` (the driver's command) at N = 1, 2, 8\n",
           "`python bench.py` / `torchrun --nproc-per-node N bench.py --gpus N`; median of 9 blocks of 20 steps (C3/C4: 5 "
           "blocks of 10).  C2 = fp32 64^3, 64 samples per GPU (weak); C3 = bf16 64^3 global batch 512 (strong: samples "
           "per rank in brackets); C4 = fp32 96^3, 32 samples per rank; C5 = 3960-sample eval pass sharded over the ranks.\n",
           "| N | C2 ms/step (min-max) | C2 volumes/s | dominant launch ms | frac | publish / collect event ms | C3 ms/step "
           "(samples/rank) | C3 volumes/s | C3 frac | C4 ms/step | C4 volumes/s | C4 frac | C5 ms/pass | C5 samples/s |",
           "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    for b in (b1, b2, b8):
        if b:
            out.append(row(b))
    if b1 and b2 and b8:
        out.append("\nWeak-scaling efficiency of C2 (value / (N x value at N=1)): N=2 %.3f, N=8 %.3f (round 1: 0.956 / 0.792).  "
                   "C3 strong scaling: N=2 %.2fx, N=8 %.2fx of the 1-GPU rate; host enqueue per sharded step %.2f ms (N=8), "
                   "against %.2f ms of kernels per step at 64 bf16 samples per rank." % (
                       b2["value"] / (2 * b1["value"]), b8["value"] / (8 * b1["value"]),
                       b2["configs"]["C3"]["value"] / b1["configs"]["C3"]["value"],
                       b8["configs"]["C3"]["value"] / b1["configs"]["C3"]["value"],
                       b8["configs"]["C3"]["host_enqueue_ms_per_step"],
                       b8["configs"]["C3"]["kernels_ms"]["joint_loss_cluster"]))
        out.append("\n`sharded_parity` (one step of the sharded criterion vs the single-rank module on the gathered batch): "
                   "N=2 %s; N=8 %s." % (json.dumps(b2["sharded_parity"]), json.dumps(b8["sharded_parity"])))
        c5 = b1["configs"]["C5"]
        out.append("\nOther keys of the N=1 line: roofline.traffic %.4f GB (%s) against %.4f GB algorithmic; cpu_baseline %s; "
                   "e2e %.1f k volumes/s at %.1f GB/s H2D (N=8: %.1f k at %.1f GB/s per GPU); C5 Procrustes %.3f ms on the GPU "
                   "(%.1f M samples/s) vs CPU baseline %s; host enqueue %.3f ms/step; other plan (forward + recompute backward) "
                   "%.3f ms/step." % (
                       b1["roofline"]["traffic"] / 1e9, json.dumps(b1["roofline"]["traffic_source"]),
                       b1["roofline"]["algorithmic_bytes_per_launch"] / 1e9,
                       json.dumps({k: b1["cpu_baseline"][k] for k in ("value", "cores", "kind")}), b1["e2e"]["value"] / 1e3,
                       b1["e2e"]["h2d_GBs"], b8["e2e"]["value"] / 1e3, b8["e2e"]["h2d_GBs"],
                       c5["kernel_ms_per_pass"]["chain+procrustes"], c5["procrustes"]["samples_per_s"] / 1e6,
                       json.dumps({k: c5["cpu_baseline"][k] for k in ("value", "us_per_sample", "kind", "pa_mpjpe_abs_diff_m")}),
                       b1["host_enqueue_ms_per_step"], b1["other_plan"]["ms_per_step"]))
    if ref and b1:
        out.append("\n`bench.py --impl reference --steps 20 --warmup 5` on the same box: %.1f volumes/s (%.1f ms per step of "
                   "batch 64, kind %s, %d cores); `config` identical to the GPU arm's: %s." % (
                       ref["value"], ref["ms_per_step"], ref["cpu_baseline"]["kind"], ref["cpu_baseline"]["cores"],
                       ref["config"] == b1["config"]))
    b56 = load("r2_bench_56.json")
    if b56:
        out.append("\nThe reference's own training shape through the same script (`bench.py --dim 56 --batch 32`): %.4f ms/step, "
                   "launch %.4f ms (%.3f of peak), host enqueue %.3f ms/step." % (
                       b56["ms_per_step"], b56["roofline"]["launch_ms"], b56["roofline"]["frac"], b56["host_enqueue_ms_per_step"]))
    out.append("\n## 2. Same-box A/B of the single-launch kernels (`scripts/ab_cluster.py`)\n")
    out.append("Median of 5 blocks of 30 launches, rotating inputs larger than L2; GB/s and frac on `V*s + V*g` (the pair moves "
               "1.5x that).  `cluster <env>`: `HIPE_CLUSTER_*` overrides; the first `cluster` line of a block is the library's "
               "own choice.  The build before the streamed form (every `cluster` line is the resident form):\n\n```\n" + (text("r2_ab_cluster_full.log") or "") + "```\n")
    for name, what in (("r2_ab_cluster_str.log", "Streamed shares forced onto every specialised shape (`scripts/ab_cluster.py str`; `STREAM=1,C=..,LEAD=..,RSLOTS=..` "
                        "are the overrides, the plain `cluster` lines the resident form that was the default in that build)"),
                       ("r2_ab_cluster_96c.log", "fp32 96^3: streamed shares on 16-CTA clusters against the stream launch, and the headline shape alternating "
                        "resident / streamed (`scripts/ab_cluster.py 96`; the cross-check lines compare the two launches' outputs)"),
                       ("r2_ab_cluster_peek.log", "Early non-blocking barrier peek off / on, two builds alternating (`scripts/ab_cluster_peek.sh`) -- no effect"),
                       ("r2_ab_cluster_56.log", "56^3 right after its specialisation (3136-byte visits), `scripts/ab_cluster.py 56`"),
                       ("r2_ab_cluster_sym.log", "Symmetric teams (`-DHIPE_CLU_EXPERIMENTS`, `HIPE_CLUSTER_SYM`), `scripts/ab_cluster.py sym` -- negative"),
                       ("r2_ab_cluster_wt2.log", "One reduce team + two write teams (`HIPE_CLUSTER_WT2`), `scripts/ab_cluster.py wt2` -- negative; "
                        "the plain `cluster` lines are the first run after the polynomial / store-hint branches left the loop")):
        t = text(name)
        if t:
            out.append(what + ":\n\n```\n" + t + "```\n")
    w = text("r2_cluster_waits.log")
    if w:
        w = "\n".join(l for l in w.splitlines() if not l.startswith("+"))
        out.append("## 3. Where the cluster launch waits (`scripts/diag_cluster_waits.sh`, `-DHIPE_CLU_PROF` build)\n")
        out.append("Per class of wait: share of the role's lifetime spent in it (all CTAs, 10 launches); then the event trace of "
                   "CTA 0 for 8 consecutive volumes (issue0/issueL: first/last TMA of the volume issued; landL: last stage landed; "
                   "folded: team partial handed over; m:team/m:pub/m:peers/m:rec: merge warp saw the team partial / published / "
                   "saw all peers / record ready; w:start/w:end: gradient write).\n\n```\n" + w + "\n```\n")
    out.append(open(os.path.join(ROOT, "profiles", "r2_configs_notes.md")).read())
    open(os.path.join(ROOT, "profiles", "r2_configs.md"), "w").write("\n".join(out))
    print("\n".join(out)[:2500])


if __name__ == "__main__":
    main()
