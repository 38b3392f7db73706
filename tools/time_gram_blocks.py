"""Per-block times of the column-sharded Gram at 8-GPU shapes, on ONE GPU: the diagonal block (1M x 512, symmetric), a full
cross block (512 x 512), the antipodal halves (256 x 512 / 512 x 256), with and without a concurrent 4.1 GB device copy.

    python tools/time_gram_blocks.py [rows] [width]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from romtime_b200 import kernels

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
w = int(sys.argv[2]) if len(sys.argv) > 2 else 512
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
A = torch.randn((rows, w), dtype=torch.float64, device=dev, generator=g)
B = torch.randn((rows, w), dtype=torch.float64, device=dev, generator=g)
C = torch.empty_like(B)
side = torch.cuda.Stream(device=dev)


def timed(name, fn, reps=5, copy=False):
    best = 1e30
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if copy:
            with torch.cuda.stream(side):
                C.copy_(B, non_blocking=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    flops = None
    print(f"{name:44s} {best:8.3f} ms" + ("  (with a concurrent device copy)" if copy else ""))
    return best


full = 2.0 * rows * w * w / 1e9   # GFLOP of a full block
for copy in (False, True):
    t = timed("diagonal block  pod_gram(A)", lambda: kernels.pod_gram(A, normalize=False), copy=copy)
    print(f"    {full * 0.5 / t:8.1f} TFLOP/s on w(w+1)/2 ... counted as half a block")
    t = timed("cross block     gram_cross(A, B)", lambda: kernels.gram_cross(A, B), copy=copy)
    print(f"    {full / t:8.1f} TFLOP/s")
    t = timed("antipodal lower gram_cross(A[:, :w/2], B)", lambda: kernels.gram_cross(A[:, : w // 2], B), copy=copy)
    print(f"    {full * 0.5 / t:8.1f} TFLOP/s")
    t = timed("antipodal upper gram_cross(A, B[:, w/2:])", lambda: kernels.gram_cross(A, B[:, w // 2:]), copy=copy)
    print(f"    {full * 0.5 / t:8.1f} TFLOP/s")
G = torch.zeros((8 * w, 8 * w), dtype=torch.float64, device=dev)
blk = kernels.gram_cross(A, B)


def place():
    G[0:w, w:2 * w] = blk
    G[w:2 * w, 0:w] = blk.t()


timed("placing a block and its transpose in G", place)
timed("torch.zeros G", lambda: torch.zeros((8 * w, 8 * w), dtype=torch.float64, device=dev))

# ---- the lift / refinement phase of the 8-GPU POD on this rank's shapes (everything but the two all-reduces)
c, r = 72, 64
Wl = torch.randn((w, c), dtype=torch.float64, device=dev, generator=g)
Y = kernels.pod_lift(A, Wl, order="C")
X = torch.linalg.qr(torch.randn((c, c), dtype=torch.float64, device=dev, generator=g))[0]
Xr = X[:, :r].contiguous()
H = kernels.pod_gram(Y, normalize=False)[0]
timed(f"lift  S_g ({rows} x {w}) W ({w} x {c}), C-order", lambda: kernels.pod_lift(A, Wl, order="C"))
timed(f"H = Y^T Y ({rows} x {c})", lambda: kernels.pod_gram(Y, normalize=False))
timed(f"Jacobi on H ({c} x {c}) incl. host syncs", lambda: kernels.sym_eig_jacobi(H))
timed(f"Q = Y X_r ({rows} x {c} x {r}), F-order", lambda: kernels.pod_lift(Y, Xr, order="F"))
Wc = torch.randn((8 * w, c), dtype=torch.float64, device=dev, generator=g)
timed(f"VT = (W_c X)^T ({8 * w} x {c} x {r})", lambda: kernels.pod_lift(Wc, Xr, order="C"))
