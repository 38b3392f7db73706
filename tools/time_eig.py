"""Eigensolve of the C3 Gram matrix alone (n = 4096, rank ~258): CUDA-event time of rt_sym_eig_f64 and of the refinement's
small Jacobi, so the replicated part of the multi-GPU POD can be profiled without the 1M-row Gram.

    python tools/time_eig.py [rows] [n]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from romtime_b200 import kernels
from romtime_b200.distributed import _bench_snapshots

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
dev = torch.device("cuda:0")
S = _bench_snapshots(rows, n, 0, n, dev, noise_seed=None)
G, cn = kernels.pod_gram(S, normalize=True)
reps = 1 if os.environ.get("PROF_NO_WARMUP") == "1" else 4
for it in range(reps):
    Gc = G.clone()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    w, W, sweeps = kernels.sym_eig(Gc)
    e1.record()
    torch.cuda.synchronize()
    print(f"sym_eig n={n}: {e0.elapsed_time(e1):.3f} ms, sweeps {sweeps}, rank {int((w > 0).sum())}")
c = 72
Y = kernels.pod_lift(S, (W[:, :c] / cn[:, None]).contiguous(), order="F")
H = kernels.pod_gram(Y, normalize=False)[0]
for it in range(reps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lam, X, _, defect = kernels.sym_eig_jacobi(H)
    e1.record()
    torch.cuda.synchronize()
    print(f"sym_eig_jacobi c={c}: {e0.elapsed_time(e1):.3f} ms, defect {defect:.2e}")
