"""Per-kernel SASS opcode histogram of libromtime_b200.so (cuobjdump -sass): the instructions that show which hardware
paths a kernel uses (TMA loads / stores, FP64 tensor MMA, warp reductions, async copies, ...).
    python tools/sass_histogram.py > profiles/<round>_sass_opcodes.md"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "romtime_b200", "libromtime_b200.so")
KEY = ["UTMALDG", "UTMASTG", "DMMA", "DFMA", "REDUX", "CREDUX", "LDGSTS", "SYNCS", "LDS", "STS", "LDG", "STG", "SHFL", "BAR", "MUFU", "ATOM", "RED"]

out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern.replace("(anonymous namespace)::", ""))
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
    if m and kern:
        hist[kern][m.group(1)] += 1
        hist[kern]["_total"] += 1
print("# SASS opcode histogram per kernel (cuobjdump -sass romtime_b200/libromtime_b200.so, sm_100a)\n")
print("Static instruction counts.  `UTMALDG` = TMA tensor load, `DMMA` = FP64 tensor-core MMA (`mma.sync.m8n8k4.f64`; sm_100a has no\n"
      "`tcgen05` FP64 kind), `REDUX` / `CREDUX` = warp reductions (`redux.sync`), `SYNCS` = mbarrier operations, `LDGSTS` = `cp.async`.\n")
print("| kernel | total | " + " | ".join(KEY) + " |")
print("|---|---|" + "---|" * len(KEY))
for k, h in hist.items():
    if h["_total"] < 40:
        continue
    print(f"| `{k[:90]}` | {h['_total']} | " + " | ".join(str(sum(v for o, v in h.items() if o == key or o.startswith(key + '.'))) for key in KEY) + " |")
