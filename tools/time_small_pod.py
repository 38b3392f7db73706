"""Latency of orth() at the reference's own sizes (C1/C2: Nh = 100..1000, n = 10..250, many calls)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from romtime_b200.rom.pod import orth

rng = np.random.default_rng(0)
for Nh, n in ((100, 50), (1000, 50), (1000, 250), (200, 10), (5000, 100)):
    x = np.linspace(0, 1, Nh)[:, None]
    tt = np.linspace(0, 1, n)[None, :]
    S = sum(10.0 ** (-k / 3) * np.sin((k + 1) * np.pi * x) * np.cos((k + 1) * tt * 3 + k) for k in range(20)) + 1e-12 * rng.standard_normal((Nh, n))
    orth(S, tol=1 - 1e-8)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        Q, s, e = orth(S, tol=1 - 1e-8)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 5
    import scipy.linalg as sl
    t0 = time.perf_counter()
    for _ in range(5):
        sl.svd(S / np.linalg.norm(S, axis=0), full_matrices=False)
    dc = (time.perf_counter() - t0) / 5
    print(f"orth {Nh}x{n}: device path {dt*1e3:.2f} ms per call (r={Q.shape[1]}), scipy gesvd on this host {dc*1e3:.2f} ms")
