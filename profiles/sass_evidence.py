"""SASS evidence that the tensor-core / TMA kernels are Blackwell-native: per kernel of lib/libraftk.so, the count of
tcgen05 MMA (UTCHMMA / UTC*MMA), TMEM load / store (LDTM / STTM), TMA load/store (UTMALDG / UTMASTG), bulk copy (UBLKCP) and
tcgen05 commit (UTCBAR) instructions.  Runs without a GPU:
    python profiles/sass_evidence.py > profiles/r2_sass_evidence.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "jax_raft_b200", "lib", "libraftk.so")
MNEMONICS = ("UTCHMMA", "UTCQMMA", "UTCIMMA", "UTCOMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "UTCATOMSWS", "HMMA", "FFMA")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()  # noqa: E731
counts, variants = collections.OrderedDict(), collections.Counter()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = demangle(m.group(1))
        cur = re.sub(r"\(.*", "", cur).replace("raftk::", "")
        counts.setdefault(cur, collections.Counter())
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)((?:\.[A-Z0-9_]+)*)", line)
    if m and m.group(1) in MNEMONICS:
        counts[cur][m.group(1)] += 1
        if m.group(1) != "FFMA":
            variants[m.group(1) + m.group(2)] += 1
print("# SASS evidence, round 2 (`cuobjdump -sass jax_raft_b200/lib/libraftk.so`, sm_100a)\n")
print("Kernels that contain tensor-core, TMEM or TMA instructions (other kernels: SIMT only).\n")
cols = [c for c in MNEMONICS if c != "FFMA" and any(v[c] for v in counts.values())]
print("| kernel | " + " | ".join(cols) + " |")
print("|---|" + "---|" * len(cols))
for k, v in counts.items():
    if any(v[c] for c in cols):
        print(f"| `{k[:110]}` | " + " | ".join(str(v[c]) for c in cols) + " |")
print("\nInstruction variants over the whole library:\n")
for k, n in sorted(variants.items()):
    print(f"* `{k}` x {n}")
print(f"\n{len(counts)} kernels in the library; no `HMMA`/`mma.sync` path: " + str(sum(v['HMMA'] for v in counts.values())) + " HMMA instructions.")
