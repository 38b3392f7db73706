"""Per-source-line digest of an ncu --set full --import-source on report (needs -lineinfo):
    python tools/ncu_lines.py <report.ncu-rep> <source file> [min non-barrier samples]
Prints, per CUDA source line, the warp-stall samples (barrier waits split off) and the warp instructions executed."""
import csv, io, subprocess, sys

rep, srcfile = sys.argv[1], sys.argv[2]
thr = int(sys.argv[3]) if len(sys.argv) > 3 else 100
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
src = open(srcfile).read().split("\n")
hdr, ix, agg, grid = None, None, {}, 1
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        ix = {}
        for i, h in enumerate(hdr):
            ix.setdefault(h, i)
        continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    ln = int(r[0])
    a = agg.setdefault(ln, [0, 0, 0, {}])
    num = lambda x: int(x) if x.strip().lstrip("-").isdigit() else 0
    a[0] += num(r[ix["# Samples"]]); a[1] += num(r[ix["stall_barrier"]]); a[2] += num(r[ix["Instructions Executed"]])
    for h in hdr:
        if h.startswith("stall_") and "Not" not in h and h != "stall_barrier":
            v = num(r[ix[h]])
            if v:
                a[3][h[6:]] = a[3].get(h[6:], 0) + v
tot = sum(a[0] for a in agg.values()); totb = sum(a[1] for a in agg.values()); tote = sum(a[2] for a in agg.values())
print(f"samples {tot} (barrier {totb}), warp instructions {tote}")
for ln in sorted(agg):
    s, b, ie, st = agg[ln]
    if s - b >= thr:
        top = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{ln:5d} nonbar {s - b:5d} bar {b:5d} winst {ie:10d} | {src[ln - 1].strip()[:90]} | {top}")
