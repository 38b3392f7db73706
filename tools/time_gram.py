import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels
Nh = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
for order in ("C", "F"):
    S = torch.randn((Nh, n) if order == "C" else (n, Nh), dtype=torch.float64, device="cuda")
    if order == "F":
        S = S.t()
    for _ in range(2):
        kernels.pod_gram(S, normalize=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(3):
        e0.record(); kernels.pod_gram(S, normalize=True); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = min(ts)
    fl = Nh * n * (n + 1.0)
    print(f"gram {order} {Nh}x{n}: {ms:.2f} ms  {fl/ms*1e-9:.2f} TFLOP/s (sym half)  S {8*Nh*n/ms*1e-6:.0f} GB/s  peak {kernels.fp64_peak_tflops():.2f}")
    del S
