"""Eigensolver timing on the Gram matrices of the two named POD shapes (rows reduced: the spectrum is what matters).
C3-like: n = 4096, rank 256 graded + 1e-13 noise, normalised;  C4-like: n = 2048 matrix-snapshot family, not normalised."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels, synthetic

dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1003)
k, Nh, n = 256, 60000, 4096
sig = torch.pow(10.0, -torch.arange(k, dtype=torch.float64, device=dev) / 32.0)
S3 = (torch.randn((Nh, k), dtype=torch.float64, device=dev, generator=g) * sig) @ torch.randn((k, n), dtype=torch.float64, device=dev, generator=g)
S3 += 1e-13 * torch.randn((Nh, n), dtype=torch.float64, device=dev, generator=g)
S4 = synthetic.mdeim_snapshots_device(60000, 2048, seed=1004)
for name, S, norm in (("C3-like n=4096", S3, True), ("C4-like n=2048", S4, False)):
    G, cn = kernels.pod_gram(S, normalize=norm)
    G0 = G.clone()
    kernels.sym_eig(G.clone())
    best = 1e9
    for _ in range(3):
        Gc = G0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); w, W, sweeps = kernels.sym_eig(Gc); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    r = int((w > 0).sum())
    res = (G0 @ W[:, :r] - W[:, :r] * w[:r]).abs().max().item() / w[0].item()
    orth = (W[:, :r].t() @ W[:, :r] - torch.eye(r, dtype=torch.float64, device=dev)).abs().max().item()
    wt = torch.linalg.eigvalsh(G0).flip(0)
    rel = ((w[:64] - wt[:64]).abs() / wt[:64]).max().item()
    print(f"{name}: eig {best:.2f} ms, sweeps {sweeps}, rank {r}, residual {res:.2e}, orth {orth:.2e}, lambda[:64] rel vs eigvalsh {rel:.2e}")
