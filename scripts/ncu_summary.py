"""Summarise ncu artefacts under gpurun_out/ into profiles/ (run in the build container):
   python scripts/ncu_summary.py r1  -> profiles/r1_launches.md, profiles/r1_kernels.md, profiles/traffic.json"""
import collections, csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
out = os.path.join(ROOT, "profiles")
os.makedirs(out, exist_ok=True)

# ---- launch list (gpu__time_duration.sum per launch; cold-cache, serialised: compare SHARES) ----
lp = os.path.join(ROOT, "gpurun_out", "%s_launches.csv" % tag)
if not os.path.exists(lp):
    lp = os.path.join(ROOT, "gpurun_out", "launches.csv")
if os.path.exists(lp):
    rows = [r for r in csv.reader(open(lp)) if len(r) > 10]
    hdr = rows[0]
    iN, iV, iG, iB = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            agg.setdefault((r[iN], r[iG], r[iB]), []).append(float(r[iV].replace(",", "")))
        except ValueError:
            pass
    tot = sum(sum(v) for v in agg.values())
    with open(os.path.join(out, "%s_launches.md" % tag), "w") as f:
        cmd = ("python bench.py --steps 10 --warmup 3 --blocks 3 --no-e2e --no-cpu-baseline --configs C3,C4" if tag >= "r2"
               else "python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline")
        f.write("# ncu launch list of `%s` (%s)\n\n" % (cmd, tag) +
                "`ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_stream|k_l1|...` on one B200; "
                "per-launch times are cold-cache and serialised, so only the SHARES are comparable with the "
                "CUDA-event numbers in the bench line.\n\n| kernel | grid | block | launches | mean us | share |\n|---|---|---|---|---|---|\n")
        for (k, g, b), v in agg.items():
            f.write("| `%s` | %s | %s | %d | %.1f | %.1f %% |\n" % (k.replace("|", "/"), g, b, len(v), sum(v) / len(v) / 1e3, 100 * sum(v) / tot))
    print(open(os.path.join(out, "%s_launches.md" % tag)).read())

# ---- full captures ----
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "launch__registers_per_thread", "launch__grid_size", "launch__cluster_size",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
traffic_path = os.path.join(out, "traffic.json")
traffic = json.load(open(traffic_path)) if os.path.exists(traffic_path) else {}
md = ["# ncu --set full captures (%s), one launch each, B200, `--clock-control none`\n" % tag]
BF16_B = os.environ.get("HIPE_NCU_BF16_B", "64")     # batch the bf16 forward capture was taken at
try:
    COMMIT = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
except Exception:
    COMMIT = None
CAPTURES = [("%s_resident_f32" % tag, "resident", "f32_64_B64"),
            ("%s_fwd_f32" % tag, "softargmax_fwd", "f32_64_B64"),
            ("%s_fwd_bf16" % tag, "softargmax_fwd", "bf16_64_B%s" % BF16_B),
            ("%s_bwd_f32" % tag, "softargmax_bwd", "f32_64_B64"), ("%s_fused_f32" % tag, "fused", "f32_64_B64"),
            ("%s_fused_bf16" % tag, "fused", "bf16_64_B128"),
            # round 2: <tag>_<kernel>_<dtype>_<dim>_B<batch>
            ("%s_cluster_f32_64_B64" % tag, "cluster", "f32_64_B64"),
            ("%s_cluster_bf16_64_B512" % tag, "cluster", "bf16_64_B512"),
            ("%s_cluster_f32_56_B32" % tag, "cluster", "f32_56_B32"),
            ("%s_cluster_f32_96_B32" % tag, "cluster", "f32_96_B32"),
            ("%s_clusterres_f32_64_B64" % tag, "cluster_resident_form", "f32_64_B64"),
            ("%s_fused_f32_96_B32" % tag, "fused", "f32_96_B32"),
            ("%s_fwd_f32_64_B32" % tag, "softargmax_fwd", "f32_64_B32")]
for name, key, cfgkey in CAPTURES:
    rep = os.path.join(ROOT, "gpurun_out", name + ".ncu-rep")
    if not os.path.exists(rep):
        continue
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    md.append("\n## %s — `%s`\n\n| metric | value |\n|---|---|" % (name, d["Kernel Name"][0]))
    for w in WANT:
        if w in d:
            md.append("| %s | %s %s |" % (w, d[w][0], d[w][1]))
    def to_bytes(x):
        v, u = x
        return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    tr = to_bytes(d["dram__bytes_read.sum"]) + to_bytes(d["dram__bytes_write.sum"])
    md.append("| **dram read + write per launch** | %.4f GB |" % (tr / 1e9))
    traffic.setdefault(key, {})[cfgkey] = {"bytes": tr, "capture": name, "commit": COMMIT}
    # hottest SASS by stall samples
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(src.splitlines()))
    h = srows[1]
    iS, iE, iW = h.index("Source"), h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)")
    data = [r for r in srows[2:] if len(r) > iW]
    totE = sum(int(r[iE]) for r in data)
    totW = sum(int(r[iW]) for r in data) or 1
    ops = collections.Counter()
    for r in data:
        t = r[iS].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        ops[op] += int(r[iE])
    md.append("\nwarp instructions executed: %d; opcode mix: %s" % (
        totE, ", ".join("%s %.1f%%" % (o, 100 * c / totE) for o, c in ops.most_common(12))))
    top = sorted(data, key=lambda r: -int(r[iW]))[:6]
    md.append("\nhottest instructions by stall samples: " + "; ".join(
        "`%s` %.1f%%" % (" ".join(r[iS].split()), 100 * int(r[iW]) / totW) for r in top))
open(os.path.join(out, "%s_kernels.md" % tag), "w").write("\n".join(md) + "\n")
json.dump(traffic, open(traffic_path, "w"), indent=1, sort_keys=True)
print("\n".join(md))
