"""torchrun --nproc-per-node P tools/pod_peer_check.py [Nh] [n]: column-sharded Gram through peer memory against the ring of
send / receive panels (same blocks, bit-identical G expected) and the timing of both."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from romtime_b200 import distributed as rd

Nh = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
world, rank = dist.get_world_size(), dist.get_rank()
dev = torch.device("cuda", local)
lo, hi = rd.shard_range(n, world, rank)
S = rd._bench_snapshots(Nh, n, lo, hi, dev, noise_seed=1)
res = {}
import functools
for name, fn in (("ring", rd.gram_column_sharded_overlapped), ("pull", rd.gram_column_sharded_peer),
                 ("direct", functools.partial(rd.gram_column_sharded_peer, direct=True))):
    G = fn(S, n)
    best = 1e9
    for _ in range(3):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); G = fn(S, n); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, t.item())
    res[name] = (G, best)
# the pull variant again with the mappings opened once (what bench.py times) and this rank's timeline
panels = rd.PeerPanels(S)
best, trace = 1e9, {}
for _ in range(4):
    tr = {}
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); G = rd.gram_column_sharded_peer(S, n, panels=panels, trace=tr); e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if t.item() < best:
        best, trace = t.item(), tr
if rank == 0:
    print(f"pull with cached mappings {best:.2f} ms; rank 0 timeline (ms): {trace}")
same = torch.equal(res["ring"][0], res["pull"][0]) and torch.equal(res["ring"][0], res["direct"][0])
diff = (res["ring"][0] - res["pull"][0]).abs().max().item()
if rank == 0:
    flops = float(Nh) * n * (n + 1)
    print(f"P={world} Nh={Nh} n={n}: ring {res['ring'][1]:.2f} ms, pull {res['pull'][1]:.2f} ms, direct TMA {res['direct'][1]:.2f} ms, ideal share of a single GPU at 35 TF/s {flops / 35e12 / world * 1e3:.2f} ms; G bit-identical {same} (max diff {diff:.2e})")
dist.destroy_process_group()
