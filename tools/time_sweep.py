import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels, synthetic
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 50
m = int(sys.argv[3]) if len(sys.argv) > 3 else 128
N = 64
model = synthetic.make_piston_model(N=N, m=m, seed=1005)
dm = kernels.DeviceRomModel(model)
mus = torch.from_numpy(synthetic.sample_mus(B, seed=1005)).cuda()
kernels.online_sweep(dm, mus, 3, 1e-3, history=False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(2):
    e0.record(); out, info = kernels.online_sweep(dm, mus, nt, 1e-3, history=False); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    fl = 5 * (2.0 * N * N * m + 2.0 * m * m) + 2.0 * m * N + 2.0 / 3 * N**3 + 4.0 * N * N
    print(f"sweep B={B} nt={nt} m={m}: {ms:.1f} ms -> {B*nt/ms*1e3:.3e} param*steps/s, {B*nt*fl/ms*1e-9:.2f} TFLOP/s algorithmic, info any={bool(info.any())}")
kernels.profile_enable(True)
out, info = kernels.online_sweep(dm, mus, nt, 1e-3, history=False)
prof = kernels.profile_read()
tot = sum(v[0] for v in prof.values())
for k, (ms, c) in prof.items():
    print(f"  {k:14s} {ms/c*1e3:9.1f} us/launch  x{c}  {100*ms/tot:5.1f}%")
