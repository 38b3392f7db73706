import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels
Nh = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
m = int(sys.argv[2]) if len(sys.argv) > 2 else 128
g = torch.Generator(device="cuda").manual_seed(1)
A = torch.randn((m, Nh), dtype=torch.float64, device="cuda", generator=g)
Q = torch.linalg.qr(A.T)[0]            # Nh x m (orthonormal), make column-major
U = Q.T.contiguous().T                 # F-order view
print(U.shape, U.stride())
for _ in range(3):
    kernels.deim_greedy(U)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ts = []
for _ in range(5):
    e0.record(); kernels.deim_greedy(U, sync=False); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = min(ts)
byt = 8 * Nh * m * (m + 1) / 2
print(f"greedy {Nh}x{m}: {ms:.3f} ms  {byt/ms*1e-6:.1f} GB/s algorithmic (all: {ts})")
