"""Shared-memory wavefronts per source line of an ncu --set full --import-source on report:
    python tools/ncu_smem.py <report.ncu-rep> <source file> <grid size> [min wavefronts per CTA]"""
import csv, io, subprocess, sys
rep, srcfile, grid = sys.argv[1], sys.argv[2], int(sys.argv[3])
thr = float(sys.argv[4]) if len(sys.argv) > 4 else 20
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
src = open(srcfile).read().split("\n")
hdr = None; agg = {}
num = lambda x: int(x) if x.strip().lstrip("-").isdigit() else 0
for r in rows:
    if r and r[0] == "Line No":
        hdr = r; ix = {}
        for i, h in enumerate(hdr): ix.setdefault(h, i)
        continue
    if hdr is None or not r or not r[0].isdigit() or len(r) != len(hdr): continue
    a = agg.setdefault(int(r[0]), [0, 0, 0])
    a[0] += num(r[ix["L1 Wavefronts Shared"]]); a[1] += num(r[ix["L1 Wavefronts Shared Excessive"]]); a[2] += num(r[ix["Instructions Executed"]])
print("wavefronts/CTA", sum(a[0] for a in agg.values()) / grid, "excessive", sum(a[1] for a in agg.values()) / grid, "warp inst/CTA", sum(a[2] for a in agg.values()) / grid)
for ln in sorted(agg):
    w, e, ie = agg[ln]
    if w / grid > thr: print(f"{ln:5d} wf {w/grid:7.1f} excess {e/grid:6.1f} inst {ie/grid:6.1f} | {src[ln-1].strip()[:100]}")
