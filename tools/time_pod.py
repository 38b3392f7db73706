import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels
from romtime_b200.rom.pod import orth_device
Nh = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
def ev(): return torch.cuda.Event(enable_timing=True)
g = torch.Generator(device="cuda").manual_seed(1)
# graded spectrum snapshot matrix: low-rank part + noise
k = 256
L = torch.randn((Nh, k), dtype=torch.float64, device="cuda", generator=g)
R = torch.randn((k, n), dtype=torch.float64, device="cuda", generator=g)
sig = torch.pow(10.0, -torch.arange(k, dtype=torch.float64, device="cuda") / 32.0)
S = (L * sig) @ R + 1e-13 * torch.randn((Nh, n), dtype=torch.float64, device="cuda", generator=g)
del L, R
G, cn = kernels.pod_gram(S, normalize=True)
torch.cuda.synchronize()
for rep in range(2):
    G2 = G.clone()
    e0, e1 = ev(), ev()
    e0.record(); w, W, sweeps = kernels.sym_eig(G2); e1.record(); torch.cuda.synchronize()
    print(f"sym_eig n={n}: {e0.elapsed_time(e1):.1f} ms, sweeps={sweeps}")
ref = torch.linalg.eigvalsh(G).flip(0)
print("eig max abs err / w0:", ((w - ref).abs().max() / ref[0]).item(), " orth err:", (W.t() @ W - torch.eye(n, dtype=torch.float64, device='cuda')).abs().max().item())
for rep in range(2):
    e0, e1 = ev(), ev()
    e0.record(); Q, s, energy, VT, _ = orth_device(S, num=64); e1.record(); torch.cuda.synchronize()
    print(f"orth {Nh}x{n} num=64: {e0.elapsed_time(e1):.1f} ms")
print("QtQ err", (Q.t() @ Q - torch.eye(64, dtype=torch.float64, device='cuda')).abs().max().item())
