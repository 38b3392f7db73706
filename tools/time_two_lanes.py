"""Feasibility of running two parameter chunks side by side (the Gauss-Jordan solve of one under the assembly GEMM of the
other): the same sweep on ONE stream, then two sweeps (two handles, two streams) queued together.

    python tools/time_two_lanes.py [B per lane] [nt]
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from romtime_b200 import _lib, kernels, synthetic

B = int(sys.argv[1]) if len(sys.argv) > 1 else 14208
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 100
N, m = 64, 128
model = synthetic.make_piston_model(N=N, m=m, seed=1005)
dms = [kernels.DeviceRomModel(model), kernels.DeviceRomModel(model)]
mus = [torch.from_numpy(synthetic.sample_mus(B, seed=1005 + i)).cuda() for i in range(2)]
lib = _lib.load_library()
h1 = _lib.handle(0)
h2 = ctypes.c_void_p()
assert lib.rt_create(0, ctypes.byref(h2)) == 0
handles = [h1, h2]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]
outs = [torch.empty((B, N), dtype=torch.float64, device="cuda") for _ in range(2)]


def lane(i, steps):
    _lib._handles[0] = handles[i]
    with torch.cuda.stream(streams[i]):
        return kernels.online_sweep(dms[i], mus[i], steps, 1e-3, history=False, out=outs[i])


for i in range(2):
    lane(i, 3)
torch.cuda.synchronize()
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(streams[0])
    lane(0, nt)
    e1.record(streams[0])
    torch.cuda.synchronize()
    t1 = e0.elapsed_time(e1)
    ref = outs[0].clone()
    # two lanes: both queued from this thread; the timer spans both streams
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    start.record()
    for s in streams:
        s.wait_event(start)
    lane(0, nt)
    lane(1, nt)
    for s in streams:
        torch.cuda.current_stream().wait_stream(s)
    end.record()
    torch.cuda.synchronize()
    t2 = start.elapsed_time(end)
    same = torch.equal(ref, outs[0])
    if not same:
        d = (ref - outs[0]).abs()
        bad = (d.max(dim=1).values > 0).nonzero().flatten()
        print(f"   lane 0 differs in {bad.numel()} of {B} parameters (first {bad[:8].tolist()}), max abs {d.max().item():.3e}, "
              f"max |ref| {ref.abs().max().item():.3e}")
    print(f"one lane: {t1:.1f} ms = {B * nt / t1 * 1e3:.4e} param*steps/s; two lanes: {t2:.1f} ms = {2 * B * nt / t2 * 1e3:.4e} "
          f"param*steps/s ({2 * t1 / t2:.3f}x); lane 0 bit-identical to the single-lane run: {same}")
_lib._handles[0] = h1
