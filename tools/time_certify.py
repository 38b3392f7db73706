"""Timing of the certification kernels at C5 scale: S-ROM estimator over B x nt (parameter, step) pairs + summary."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from romtime_b200 import kernels

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 500
N = int(sys.argv[3]) if len(sys.argv) > 3 else 64
Nhat = int(sys.argv[4]) if len(sys.argv) > 4 else 64
Nh = 100000
g = torch.Generator(device="cuda").manual_seed(5)
V = torch.randn((Nh, Nhat), dtype=torch.float64, device="cuda", generator=g)
uN = torch.randn((B, nt, N), dtype=torch.float64, device="cuda", generator=g)
uS = torch.randn((B, nt, Nhat), dtype=torch.float64, device="cuda", generator=g)
ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
for _ in range(2):
    err = kernels.rom_difference(uN, uS, V)
    summ = kernels.errors_summary(err)
torch.cuda.synchronize()
e = [ev() for _ in range(3)]
e[0].record(); err = kernels.rom_difference(uN, uS, V); e[1].record(); summ = kernels.errors_summary(err); e[2].record()
torch.cuda.synchronize()
t1, t2 = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])
rows = B * nt
byts = 8.0 * rows * (N + Nhat + 1)
flop = 2.0 * rows * Nhat * Nhat
print(f"rom_difference B={B} nt={nt} N={N} Nhat={Nhat}: {t1:.2f} ms incl. Gram of V -> {byts / t1 * 1e-6:.0f} GB/s, "
      f"{flop / t1 * 1e-9:.2f} TFLOP/s; errors_summary: {t2:.2f} ms ({8.0 * rows / t2 * 1e-6:.0f} GB/s)")
# spot check against the direct formula
b, n = B // 3, nt // 2
d = uS[b, n].clone(); d[:N] -= uN[b, n]
ref = torch.linalg.norm(V @ d) / Nh ** 0.5
print("spot check rel err", float(abs(err[b, n] - ref) / ref), "summary row0", summ[0].tolist())
