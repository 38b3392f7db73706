"""Multi-GPU partitioning of the hot path on one 8 x B200 NVLink/NVSwitch box.

One process per GPU (``torch.distributed``, backend ``nccl`` on the box, ``gloo`` in the CPU tests).

* online sweep -- the parameter batch is split into contiguous shards (``shard_range``); the model is
  replicated; there is NO data-path collective.
* POD -- the snapshot COLUMNS are sharded: rank g owns ``S_g (Nh x n_g)``.  The Gram matrix is built
  block-wise: panels travel round a ring (``batch_isend_irecv`` double-buffered against the block GEMMs,
  NVLink 5 through the switch), rank g computes the blocks ``G[g, (g+h) % P]`` for the hops
  ``h = 0 .. P//2`` only (symmetry: the other half is the transpose, written by the same rank so the
  assembled matrix is exactly symmetric; the antipodal block of an even world is split between the pair), writes them into a zero ``n x n`` buffer and one
  ``all_reduce(SUM)`` of G (134 MB at n = 4096) leaves the full Gram matrix on every rank.  Every entry
  of G is produced by exactly one rank, so the reduction adds exact zeros and the result is
  bit-identical on all ranks; the eigensolve is then replicated.  The lift ``Q = sum_g S_g W_g`` is a
  second ``all_reduce`` (Nh x r).

The block operations are injected (``block_ops``) so that the sharding / ring / symmetry logic runs
unchanged under gloo on CPU tensors in tests/ (where the blocks are computed by the test's own torch
reference); on the GPU the blocks go through the C ABI (rt_pod_gram_f64 / rt_gram_cross_f64 /
rt_pod_lift_f64).
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(total, world, rank):
    """Contiguous balanced shard [lo, hi) of ``total`` units for ``rank`` of ``world``."""
    return rank * total // world, (rank + 1) * total // world


def column_shards(n, world):
    return [shard_range(n, world, g) for g in range(world)]


class DeviceBlockOps:
    """Block operations through libromtime_b200.so (CUDA tensors)."""

    def gram(self, A):  # A^T A, raw (not normalised)
        from . import kernels

        return kernels.pod_gram(A, normalize=False)[0]

    def cross(self, A, B):  # A^T B
        from . import kernels

        return kernels.gram_cross(A, B)

    def lift(self, S, W, order="C"):  # S @ W, C-ordered by default so that it can be all-reduced in place
        from . import kernels

        return kernels.pod_lift(S, W, order=order)

    def jacobi(self, H):  # plain Jacobi on the Gram matrix of nearly orthogonal columns (Hestenes refinement)
        from . import kernels

        lam, X, _, defect = kernels.sym_eig_jacobi(H)
        return lam, X, defect

    def finalize(self, G, normalize):
        from . import kernels

        return kernels.gram_finalize(G, normalize=normalize)

    def eig(self, G):
        from . import kernels

        w, W, _ = kernels.sym_eig(G)
        return w, W


def _ring_exchange(send_buf, recv_buf, group):
    """Post one ring hop: send my current panel to rank-1, receive the next one from rank+1."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    ops = [dist.P2POp(dist.isend, send_buf, (rank - 1) % world, group),
           dist.P2POp(dist.irecv, recv_buf, (rank + 1) % world, group)]
    return dist.batch_isend_irecv(ops)


def _cross_block(ops, G, S_local, panel, rank, src, shards, antipodal):
    """Compute and place the Gram block of (this rank's columns) x (rank ``src``'s columns) and its transpose.
    For an even world the antipodal pair (g, g + P/2) would both compute the same block: the LOWER rank takes the
    first half of ITS columns against all of the partner's, the HIGHER rank all of its columns against the second
    half of the partner's -- the two pieces tile the block exactly and the work stays balanced."""
    lo, hi = shards[rank]
    slo, shi = shards[src]
    if not antipodal:
        blk = ops.cross(S_local, panel)                     # (n_g x n_src) = S_g^T S_src
        G[lo:hi, slo:shi] = blk
        G[slo:shi, lo:hi] = blk.t()
    elif rank < src:
        h1 = (hi - lo) // 2
        if h1 > 0:
            blk = ops.cross(S_local[:, :h1], panel)
            G[lo:lo + h1, slo:shi] = blk
            G[slo:shi, lo:lo + h1] = blk.t()
    else:
        h1s = (shi - slo) // 2
        if shi - slo - h1s > 0:
            blk = ops.cross(S_local, panel[:, h1s:])
            G[lo:hi, slo + h1s:shi] = blk
            G[slo + h1s:shi, lo:hi] = blk.t()


def gram_column_sharded(S_local, n, block_ops=None, group=None):
    """Raw Gram matrix ``G = S^T S`` (n x n) of the column-sharded snapshot matrix, on every rank.

    ``S_local``: this rank's columns ``[lo, hi) = shard_range(n, world, rank)`` as an (Nh x n_g) tensor
    (any strides).  Returns G (n x n, contiguous) after the all-reduce."""
    ops = block_ops or DeviceBlockOps()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    shards = column_shards(n, world)
    lo, hi = shards[rank]
    Nh = S_local.shape[0]
    assert S_local.shape[1] == hi - lo, (S_local.shape, lo, hi)
    G = torch.zeros((n, n), dtype=S_local.dtype, device=S_local.device)
    G[lo:hi, lo:hi] = ops.gram(S_local)
    if world == 1:
        return G
    # panels are shipped column-major (each snapshot contiguous) and padded to the widest shard
    wmax = max(b - a for a, b in shards)
    bufs = [torch.empty((wmax, Nh), dtype=S_local.dtype, device=S_local.device) for _ in range(2)]
    bufs[0][: hi - lo].copy_(S_local.t())
    hops = world // 2
    cur = 0
    for h in range(1, hops + 1):
        reqs = _ring_exchange(bufs[cur], bufs[1 - cur], group)
        for r in reqs:
            r.wait()
        cur = 1 - cur
        src = (rank + h) % world
        slo, shi = shards[src]
        panel = bufs[cur][: shi - slo].t()  # Nh x n_src view, column-major
        _cross_block(ops, G, S_local, panel, rank, src, shards, world % 2 == 0 and h == hops)
    dist.all_reduce(G, op=dist.ReduceOp.SUM, group=group)
    return G


def gram_column_sharded_overlapped(S_local, n, block_ops=None, group=None):
    """Same result as ``gram_column_sharded`` with the ring hop h+1 in flight while block h is computed
    (two panel buffers + the local panel; NCCL runs the transfer on its own stream)."""
    ops = block_ops or DeviceBlockOps()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world <= 2:
        return gram_column_sharded(S_local, n, ops, group)
    rank = dist.get_rank(group)
    shards = column_shards(n, world)
    lo, hi = shards[rank]
    Nh = S_local.shape[0]
    G = torch.zeros((n, n), dtype=S_local.dtype, device=S_local.device)
    wmax = max(b - a for a, b in shards)
    bufs = [torch.empty((wmax, Nh), dtype=S_local.dtype, device=S_local.device) for _ in range(3)]
    bufs[0][: hi - lo].copy_(S_local.t())
    hops = world // 2
    reqs = _ring_exchange(bufs[0], bufs[1], group)       # hop 1 in flight during the diagonal block
    G[lo:hi, lo:hi] = ops.gram(S_local)
    for h in range(1, hops + 1):
        for r in reqs:
            r.wait()
        cur = h % 3
        reqs = _ring_exchange(bufs[cur], bufs[(h + 1) % 3], group) if h < hops else []
        src = (rank + h) % world
        slo, shi = shards[src]
        panel = bufs[cur][: shi - slo].t()
        _cross_block(ops, G, S_local, panel, rank, src, shards, world % 2 == 0 and h == hops)
    dist.all_reduce(G, op=dist.ReduceOp.SUM, group=group)
    return G


def orth_column_sharded(S_local, n, num=None, tol=None, normalize=True, block_ops=None, group=None,
                        overlap=True):
    """Distributed ``orth`` (reference rom/pod.py:7-62) of a column-sharded snapshot matrix.

    Returns ``(Q, s, energy, VT)``: Q (Nh x r) and VT (r x n) as tensors on this rank's device (Q is
    identical on every rank after the lift all-reduce), s / energy as NumPy arrays."""
    from .rom.pod import _energy, _guard_width, _truncation, refine_modes

    ops = block_ops or DeviceBlockOps()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_range(n, world, rank)
    Nh = S_local.shape[0]
    gram = gram_column_sharded_overlapped if overlap else gram_column_sharded
    G = gram(S_local, n, ops, group)
    cn = ops.finalize(G, normalize)
    w, W = ops.eig(G)
    k = min(Nh, n)
    w_host = w[:k].cpu().numpy()
    s = np.sqrt(np.maximum(w_host, 0.0))
    nrank = int(np.count_nonzero(s > 0.0))
    energy = _energy(s)
    r = min(int(np.count_nonzero(_truncation(s, energy, num, tol))), nrank)
    if r == 0:
        return S_local.new_zeros((Nh, 0)), s, energy, W[:, :0].t()
    # lift of the retained + guard modes (this rank's columns' contribution), summed over the ranks; the Hestenes
    # refinement (rom/pod.py) then runs replicated on the Nh x c result -- one level only on the sharded path
    c = _guard_width(s, r, nrank)
    Wc = W[:, :c].contiguous()
    Wl = Wc / cn[:, None] if normalize else Wc
    Y = ops.lift(S_local, Wl[lo:hi].contiguous())
    if world > 1:
        dist.all_reduce(Y, op=dist.ReduceOp.SUM, group=group)
    sc, Q, VT, _, _ = refine_modes(Y, Wc, r, ops=ops)
    s = s.copy()
    s[:c] = sc
    energy = _energy(s)
    r = min(int(np.count_nonzero(_truncation(s, energy, num, tol))), Q.shape[1])
    return Q[:, :r], s, energy, VT[:r]


# ------------------------------------------------------------------------------- bench leg (N > 1)
def bench_pod(a, Nh, n):
    """C3 POD leg of bench.py under torchrun: column-sharded synthetic snapshots, device timing, max over
    ranks.  Every rank generates the same low-rank factors from the same seed and keeps its columns."""
    from . import kernels

    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    lo, hi = shard_range(n, world, rank)
    g = torch.Generator(device=dev).manual_seed(1003)
    k = 256
    R = torch.randn((k, n), dtype=torch.float64, device=dev, generator=g)[:, lo:hi].contiguous()
    sig = torch.pow(10.0, -torch.arange(k, dtype=torch.float64, device=dev) / 32.0)
    S = torch.empty((Nh, hi - lo), dtype=torch.float64, device=dev)
    rows = 65536
    gn = torch.Generator(device=dev).manual_seed(7919 + rank)
    for r0 in range(0, Nh, rows):
        r1 = min(Nh, r0 + rows)
        L = torch.randn((r1 - r0, k), dtype=torch.float64, device=dev, generator=g) * sig
        S[r0:r1] = L @ R
        S[r0:r1] += 1e-13 * torch.randn((r1 - r0, hi - lo), dtype=torch.float64, device=dev, generator=gn)
    del L, R
    best = None
    for it in range(3):
        dist.barrier()
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        G = gram_column_sharded_overlapped(S, n)
        e[1].record()
        del G
        Q, s, energy, VT = orth_column_sharded(S, n, num=64)
        e[2].record()
        torch.cuda.synchronize()
        t = torch.tensor([e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t = t.tolist()
        if best is None or t[1] < best[1]:
            best = t
    peak = kernels.fp64_peak_tflops()
    flops = float(Nh) * n * (n + 1)
    QtQ = (Q.t() @ Q - torch.eye(Q.shape[1], dtype=torch.float64, device=dev)).abs().max().item()
    return {
        "shape": [Nh, n], "num": 64, "sharding": f"snapshot columns over {world} GPUs, ring panels + all_reduce(G) + all_reduce(Q)",
        "svd_time_ms": best[1], "gram_allreduce_ms": best[0],
        "gram_tflops_sym_half_aggregate": flops / best[0] * 1e-9,
        "gram_frac_of_fp64_peak_aggregate": flops / best[0] * 1e-9 / (peak * world),
        "S_GBps_whole_orth": 8.0 * Nh * n / best[1] * 1e-6, "QtQ_minus_I_max": QtQ,
        "energy_last": float(energy[-1]), "singular_values_head": [float(x) for x in s[:4]],
    }
