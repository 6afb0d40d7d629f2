"""Turn the ncu outputs a gpurun call left in gpurun_out/ into the tracked summaries under profiles/.

    python profiles/summarize.py r1      # reads gpurun_out/r1_step_launches.csv, r1_hot_b{1,16}.ncu-rep
"""
import collections
import csv
import os
import re
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO = os.path.join(ROOT, "gpurun_out")
OUT = os.path.join(ROOT, "profiles")
lines = [f"# ncu summary, round {tag}\n"]

# ---- launch list of one full step ----
fn = os.path.join(GO, f"{tag}_step_launches.csv")
if os.path.exists(fn):
    rows = [l for l in open(fn) if not l.startswith("==")]
    recs = [(r["Kernel Name"], float(r["Metric Value"].replace(",", ""))) for r in csv.DictReader(rows)
            if r.get("Metric Name") == "gpu__time_duration.sum"]
    tot = sum(v for _, v in recs)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, v in recs:
        k = re.sub(r"\(.*", "", k)
        k = re.sub(r"<.*", "", k)[:70]
        agg[k][0] += 1
        agg[k][1] += v
    mine = sum(t for k, (n, t) in agg.items() if any(t in k for t in ("raftk", "tc::", "mu::", "lc::")))
    lines += [f"## One raft_large 440x1024 x32-iteration step, eager, every launch (`{os.path.basename(fn)}`)\n",
              f"`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised: compare shares).",
              f"{len(recs)} launches, {tot / 1e6:.2f} ms summed; raftk kernels: {mine / 1e3:.1f} us = {100 * mine / tot:.1f} % of the step.\n",
              "| share | launches | mean us | kernel |", "|---|---|---|---|"]
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:18]:
        lines.append(f"| {100 * t / tot:.1f} % | {n} | {t / n / 1e3:.1f} | `{k}` |")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if (any(t in k for t in ("raftk", "tc::", "mu::", "lc::"))) and all(k not in l for l in lines):
            lines.append(f"| {100 * t / tot:.2f} % | {n} | {t / n / 1e3:.1f} | `{k}` |")
    with open(os.path.join(OUT, f"{tag}_step_launches.csv"), "w") as f:
        f.writelines(rows)
    lines.append("")

# ---- launch list of bench.py itself (kernels inside the CUDA graphs), aggregated per kernel ----
fn = os.path.join(GO, f"{tag}_bench_launches.csv")
if os.path.exists(fn):
    rows = [l for l in open(fn) if not l.startswith("==")]
    recs = [(r["Kernel Name"], float(r["Metric Value"].replace(",", ""))) for r in csv.DictReader(rows)
            if r.get("Metric Name") == "gpu__time_duration.sum"]
    tot = sum(v for _, v in recs)
    agg = collections.defaultdict(lambda: [0, 0.0])
    for k, v in recs:
        k = re.sub(r"\(.*", "", k)[:100]
        agg[k][0] += 1
        agg[k][1] += v
    with open(os.path.join(OUT, f"{tag}_bench_launches_summary.csv"), "w") as f:
        f.write("kernel,launches,total_ns,mean_ns,share\n")
        for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f'"{k}",{n},{t:.0f},{t / n:.0f},{t / tot:.5f}\n')
    mine = sum(t for k, (n, t) in agg.items() if any(t in k for t in ("raftk", "tc::", "mu::", "lc::")))
    lines += [f"## `python bench.py --steps 2 --warmup 3 --skip-cpu-baseline --micro-batch 0` under the same ncu pass "
              f"(`{tag}_bench_launches_summary.csv`)\n",
              f"{len(recs)} launches over the whole command (eager warm-ups, graph capture, replays, event-timed pass); "
              f"raftk kernels = {100 * mine / tot:.1f} % of the summed kernel time "
              f"(of which hot-path rows: {100 * sum(t for k, (n, t) in agg.items() if 'raftk' in k and 'bias_act' not in k and 'instnorm' not in k) / tot:.1f} %).\n",
              "| share | launches | mean us | kernel |", "|---|---|---|---|"]
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        if "raftk" in k:
            lines.append(f"| {100 * t / tot:.2f} % | {n} | {t / n / 1e3:.1f} | `{k[:70]}` |")
    lines.append("")

# ---- full captures ----
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "smsp__inst_executed.sum",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"]
for b in (1, 16):
    rep = os.path.join(GO, f"{tag}_hot_b{b}.ncu-rep")
    if not os.path.exists(rep):
        continue
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    with open(os.path.join(OUT, f"{tag}_hot_b{b}_raw.csv"), "w") as f:
        f.write(raw)
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    lines += [f"## `ncu --set full`, three kernel families alone, batch {b} (55x128, D=256) (`{tag}_hot_b{b}_raw.csv`)\n",
              "| kernel | " + " | ".join(w.split(".")[0].replace("__", ".") for w in WANT) + " |", "|---|" + "---|" * len(WANT)]
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]])[:40]
        vals = [f"{r[idx[w]]} {units[idx[w]]}" if w in idx else "-" for w in WANT]
        lines.append(f"| `{name}` | " + " | ".join(vals) + " |")
    lines.append("")

# ---- f1 (lookup fused with convcorr1) ----
rep = os.path.join(GO, f"{tag}_f1_b1.ncu-rep")
raw_path = os.path.join(OUT, f"{tag}_f1_b1_raw.csv")
if os.path.exists(rep):
    with open(raw_path, "w") as f:
        f.write(subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout)
if os.path.exists(raw_path):
    rows = list(csv.reader(open(raw_path)))
    hdr, units, r = rows[0], rows[1], rows[-1]
    more = ["smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio"]
    lines += [f"## `ncu --set full`, f1 `lookup_conv_kernel<4>` at batch 1 (`{tag}_f1_b1_raw.csv`; `profiles/lookup_conv_time.py 1`)\n",
              "| metric | value |", "|---|---|"]
    for w in WANT + more:
        if w in hdr:
            lines.append(f"| `{w}` | {r[hdr.index(w)]} {units[hdr.index(w)]} |")
    lines.append("")

open(os.path.join(OUT, f"{tag}_summary.md"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
