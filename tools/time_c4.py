"""C4 (BASELINE.json configs[3]) on one B200: MDEIM offline chain at the named size --
POD of 200,000 x 2048 matrix snapshots (num = 128, not normalised), greedy point selection on the
200,000 x 128 collateral basis, projection onto a 64-dimensional solution basis.
Prints one JSON line with device times (CUDA events, best of 3) and algorithmic rates."""
import json
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


def timed(fn, reps=3):
    best = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        out = fn()
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1)
        best = t if best is None or t < best else best
    return best, out


S = torch.from_numpy(synthetic.mdeim_snapshots(nnz, n, seed=1004)).cuda()  # F-order, duplicated off-diagonals
res = {"config": f"C4: {nnz} x {n} snapshots, m={m}, N={N}"}
t_pod, (Q, s, energy, VT, _) = timed(lambda: orth_device(S, num=m, normalize=False))
res["pod_ms"] = t_pod
res["pod_GBps_of_S"] = 8.0 * nnz * n / t_pod * 1e-6
U = Q  # (nnz x m) F-order CUDA tensor
kernels.deim_greedy(U)
t_gr, (dofs, PT_U) = timed(lambda: kernels.deim_greedy(U, sync=False))
bytes_alg = 8.0 * nnz * m * (m + 1) / 2
res["greedy_ms"] = t_gr
res["greedy_GBps_algorithmic"] = bytes_alg / t_gr * 1e-6
res["greedy_dofs_head"] = dofs[:6].cpu().tolist()
res["greedy_unique"] = int(torch.unique(dofs).numel())
# tridiagonal topology with nnz entries
Nh = (nnz + 2) // 3
rows = np.repeat(np.arange(Nh), 3)[1:nnz + 1]
cols = (np.repeat(np.arange(Nh), 3) + np.tile([-1, 0, 1], Nh))[1:nnz + 1]
cols = np.clip(cols, 0, Nh - 1)
V = torch.linalg.qr(torch.randn((Nh, N), dtype=torch.float64, device="cuda"))[0]
rd, cd = torch.from_numpy(rows).cuda(), torch.from_numpy(cols).cuda()


def proj():
    return kernels.project_mat(U, rows, cols, V)


t_pr, basis_rom = timed(proj)
res["project_ms_with_host_index_upload"] = t_pr
res["project_flops_T"] = 2.0 * Nh * N * N * m / t_pr * 1e-9
t_pv, _ = timed(lambda: kernels.project_vec(V, torch.randn((Nh, m), dtype=torch.float64, device="cuda")))
res["project_vec_ms"] = t_pv
print(json.dumps(res))
