"""Markdown table of the headline metrics of `ncu --page raw --csv` exports:
    python profiles/summarize_raw.py "title" file.csv ["title" file.csv ...] > profiles/rN_summary.md"""
import csv
import re
import sys

WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM written"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "L1/shared %"), ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("smsp__inst_executed.sum", "warp instructions")]
print("# ncu summary, round 2\n")
print("`ncu --set full --clock-control none --import-source on` under gpurun, one pass per file, after the same command had "
      "exited 0 without ncu; times are cold-cache and serialised (compare shares and counters, not absolutes). "
      "Timed numbers are in DESIGN.md §5 / bench.py.\n")
args = sys.argv[1:]
for title, fn in zip(args[0::2], args[1::2]):
    rows = list(csv.reader(open(fn)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"## {title} (`{fn.split('/')[-1]}`)\n")
    print("| kernel | " + " | ".join(n for _, n in WANT) + " |")
    print("|---|" + "---|" * len(WANT))
    for r in rows[2:]:
        name = re.sub(r"\(.*", "", r[idx["Kernel Name"]]).replace("void ", "")[:60]
        vals = []
        for w, _ in WANT:
            if w not in idx:
                vals.append("-")
                continue
            v, u = r[idx[w]], units[idx[w]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.4g}" if f < 1e6 else f"{f:.4g}"
            except ValueError:
                pass
            vals.append(f"{v} {u}".strip())
        print(f"| `{name}` | " + " | ".join(vals) + " |")
    print()
