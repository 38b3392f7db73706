"""Tree-walk POD + greedy for ncu: 64 per-parameter PODs (1000 x 250) through the batched small-POD kernels, then the greedy."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from romtime_b200 import kernels, synthetic
from romtime_b200.rom.pod import orth, orth_batch

mats = [synthetic.treewalk_snapshots(1000, 250, seed=p) for p in range(64)]
per = orth_batch(mats, tol=1.0 - 1e-8)
Q, s, e = orth(np.hstack([q for q, _, _ in per]), tol=1.0 - 1e-8, normalize=False)
dofs, PT_U = kernels.deim_greedy(np.asfortranarray(Q))
torch.cuda.synchronize()
print("widths", [q.shape[1] for q, _, _ in per][:8], "final", Q.shape, "dofs", dofs[:5].cpu().tolist())
