"""Readable digest of one bench.py JSON line: python profiles/bench_digest.py gpurun_out/xxx.json"""
import json
import sys

for line in open(sys.argv[1]):
    if line.startswith("{"):
        d = json.loads(line)
print("value %.1f %s  ms/step %.3f  e2e %.1f  launches/step %s  per-rank %s" % (
    d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d.get("gpu_launches_per_step"), d.get("per_rank_ms")))
r = d["roofline"]
print("roofline: %s  %.2f us  frac %.3f (net of events %.3f)  traffic %s" % (
    r["kernel"], r["mean_launch_ms"] * 1e3, r["frac"], r["frac_net_of_event_overhead"], r.get("traffic")))
print("pyramid in step: %.1f us" % (d["kernels_in_step"]["corr_pyramid"]["mean_ms"] * 1e3))
if "fp32_convolutions" in d:
    print("fp32 convolutions:", d["fp32_convolutions"])


def show(e, indent="  "):
    for k, x in e.items():
        if isinstance(x, dict) and "mean_ms" in x:
            extra = ""
            if x.get("frac_of_tensor_peak") is not None:
                extra += "  tensor %.3f (%.0f TF)" % (x["frac_of_tensor_peak"], x.get("tflops_issued_bf16", x.get("tflops_tf32", 0)))
            if "own_frac_of_hbm_peak" in x:
                extra += "  own-bytes %.3f" % x["own_frac_of_hbm_peak"]
            print("%s%-22s %9.2f us  hbm %.3f%s" % (indent, k, x["mean_ms"] * 1e3, x["frac_of_hbm_peak"], extra))
        elif k == "altcorr":
            print("%saltcorr: pool+split %.1f us, lookup %.1f us/iter (%.1f TFLOP/s useful); simt %.1f us/iter" % (
                indent, x["pool_fmap2_ms"] * 1e3, x["lookup_ms_per_iter"] * 1e3, x["tflops"], x.get("simt_lookup_ms_per_iter", 0) * 1e3))


ks = d.get("kernels_saturating", {})
for b in sorted(k for k in ks if k.startswith("batch_")):
    print(b)
    show(ks[b])
for k, c in d.get("configs", {}).items():
    if k == "c5":
        for e in c["sweep"]:
            print("c5 %dx%d batch %d" % (e["h"], e["w"], e["batch"]))
            show(e)
    else:
        print(k, json.dumps({kk: vv for kk, vv in c.items() if kk not in ("workload", "timing", "rest_of_model")}))
if "cpu_baseline" in d:
    print("cpu_baseline %.3f %s on %d cores" % (d["cpu_baseline"]["value"], d["cpu_baseline"]["unit"], d["cpu_baseline"]["cores"]))
