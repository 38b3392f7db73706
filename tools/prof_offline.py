"""Offline chain at C4 size for ncu: POD (Gram + eigensolve + lift) of 200,000 x 2048 matrix snapshots, greedy on the
200,000 x 128 basis, DEIM / MDEIM projections.  One pass, no warm-up when PROF_NO_WARMUP=1 (ncu replays every launch).

    python tools/prof_offline.py [rows] [cols]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from romtime_b200 import kernels, synthetic
from romtime_b200.rom.pod import orth_device

nnz = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
m, N = 128, 64
S = synthetic.mdeim_snapshots_device(nnz, n, seed=1004)
passes = 1 if os.environ.get("PROF_NO_WARMUP") == "1" else 2
for _ in range(passes):
    Q, s, energy, VT, _ = orth_device(S, num=m, normalize=False)
    dofs, PT_U = kernels.deim_greedy(Q)
    Nh = (nnz + 2) // 3
    rows = np.repeat(np.arange(Nh), 3)[1:nnz + 1]
    cols = np.clip((np.repeat(np.arange(Nh), 3) + np.tile([-1, 0, 1], Nh))[1:nnz + 1], 0, Nh - 1)
    V = torch.linalg.qr(torch.randn((Nh, N), dtype=torch.float64, device="cuda"))[0]
    basis_rom = kernels.project_mat(Q, rows, cols, V)
    torch.cuda.synchronize()
print("sigma head", s[:3], "sigma_m/sigma_1", s[m - 1] / s[0], "dofs head", dofs[:6].cpu().tolist(),
      "unique", int(torch.unique(dofs).numel()))
