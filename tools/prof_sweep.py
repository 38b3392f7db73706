import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from romtime_b200 import kernels, synthetic
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4736
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 4
m = int(sys.argv[3]) if len(sys.argv) > 3 else 128
model = synthetic.make_piston_model(N=64, m=m, seed=1005)
dm = kernels.DeviceRomModel(model)
mus = torch.from_numpy(synthetic.sample_mus(B, seed=1005)).cuda()
out, info = kernels.online_sweep(dm, mus, nt, 1e-3, history=False)
torch.cuda.synchronize()
print("ok", float(out.abs().max()))
