"""Does the assembly GEMM stay bit-reproducible when other kernels share its SMs?  One sweep alone, then the same sweep while
a side stream runs elementwise torch kernels (128-thread CTAs, ~20 registers, no shared memory: they fit beside a GEMM CTA only
when the GEMM is built with head-room, RT_NVCC_EXTRA="-DRT_GEMM_STAGES=5 -DRT_GEMM_LAUNCH_REGS=144 -DRT_GEMM_CONSUMER_REGS=200").

    python tools/coresident_check.py [B] [nt]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from romtime_b200 import kernels, synthetic

B = int(sys.argv[1]) if len(sys.argv) > 1 else 14208
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 60
model = synthetic.make_piston_model(N=64, m=128, seed=1005)
dm = kernels.DeviceRomModel(model)
mus = torch.from_numpy(synthetic.sample_mus(B, seed=1005)).cuda()
out = torch.empty((B, 64), dtype=torch.float64, device="cuda")
kernels.online_sweep(dm, mus, 3, 1e-3, history=False, out=out)
torch.cuda.synchronize()
kernels.online_sweep(dm, mus, nt, 1e-3, history=False, out=out)
torch.cuda.synchronize()
ref = out.clone()
side = torch.cuda.Stream()
x = torch.randn(1 << 22, device="cuda")
for mode in ("quiet", "elementwise side stream", "quiet"):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if mode != "quiet":
        with torch.cuda.stream(side):
            for _ in range(6000):
                x.mul_(1.0000001)
    e0.record()
    kernels.online_sweep(dm, mus, nt, 1e-3, history=False, out=out)
    e1.record()
    torch.cuda.synchronize()
    d = (out - ref).abs()
    bad = (d.max(dim=1).values > 0).nonzero().flatten()
    print(f"{mode:26s}: {e0.elapsed_time(e1):8.1f} ms, bit-identical {bad.numel() == 0}"
          + (f" ({bad.numel()} parameters differ, first {bad[:8].tolist()}, max abs {d.max().item():.2e})" if bad.numel() else ""))
