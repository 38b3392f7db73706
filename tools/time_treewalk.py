"""Tree-walk POD at C1/C2-like sizes: P per-parameter time-PODs (Nh x nt each, tol-driven ragged widths) + the stacked
mu-POD, through the batched device path, the per-matrix device path and the reference's CPU loop (scipy gesvd per matrix)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import numpy as np
import torch
import romtime_oracle as orc
from romtime_b200.rom.pod import orth, orth_batch
from romtime_b200 import synthetic

def run(P, Nh, nt, tol=1.0 - 1e-8):
    mats = [synthetic.treewalk_snapshots(Nh, nt, seed=p) for p in range(P)]
    def walk(batch):
        if batch:
            per = orth_batch(mats, tol=tol)
        else:
            per = [orth(S, tol=tol) for S in mats]
        stacked = np.hstack([q for q, _, _ in per])
        return orth(stacked, tol=tol, normalize=False), stacked.shape[1]
    out = {}
    for name, batch in (("batched", True), ("per_matrix", False)):
        walk(batch); torch.cuda.synchronize()
        t0 = time.perf_counter(); (Q, s, e), width = walk(batch); torch.cuda.synchronize()
        out[name] = (time.perf_counter() - t0) * 1e3
    t0 = time.perf_counter()
    per = [orc.orth(S, tol=tol) for S in mats]
    Qr, sr, er = orc.orth(np.hstack([q for q, _, _ in per]), tol=tol, normalize=False)
    out["cpu"] = (time.perf_counter() - t0) * 1e3
    r = min(Q.shape[1], Qr.shape[1])
    rel = float(np.max(np.abs(s[:r] - sr[:r]) / sr[:r]))
    print(f"treewalk P={P} {Nh}x{nt}: batched {out['batched']:.1f} ms, per-matrix device {out['per_matrix']:.1f} ms, CPU gesvd loop {out['cpu']:.1f} ms "
          f"(stacked width {width}, final r {Q.shape[1]} vs {Qr.shape[1]}, sigma rel {rel:.1e})")

if __name__ == "__main__":
    for cfg in ((64, 1000, 250), (64, 100, 50), (16, 1000, 50), (128, 500, 100)):
        run(*cfg)
