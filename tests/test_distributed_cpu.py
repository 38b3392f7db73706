"""Host-side multi-GPU logic under gloo on CPU (world_size 2, 3 and 4): parameter sharding, the
column-sharded Gram ring with symmetry exploitation, the Gram/lift all-reduces and the distributed
``orth``.  The per-block arithmetic is supplied by THIS test (plain torch on CPU tensors) -- the
product's block operations are CUDA-only; what is tested here is the partitioning, the ring schedule,
the block placement / transposition and the truncation logic shared with rom/pod.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class TorchCpuBlockOps:
    """Reference block operations for the CPU test (never used by the product)."""

    def gram(self, A):
        return A.t() @ A

    def cross(self, A, B):
        return A.t() @ B

    def lift(self, S, W, order="C"):
        return (S @ W).contiguous()

    def jacobi(self, H):
        import pod_emulation

        lam, X, _, defect = pod_emulation.sym_eig_jacobi(H)
        return lam, X, defect

    def finalize(self, G, normalize):
        cn = torch.sqrt(torch.diagonal(G)).clone()
        if normalize:
            G /= cn[:, None] * cn[None, :]
        return cn

    def eig(self, G):
        w, W = torch.linalg.eigh(G)
        return w.flip(0).contiguous(), W.flip(1).contiguous()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, Nh, n, overlap, out_dir):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from romtime_b200 import distributed as rd

        rng = np.random.default_rng(77)
        S = rng.standard_normal((Nh, 12)) @ np.diag(10.0 ** (-np.arange(12) / 3.0)) @ rng.standard_normal((12, n))
        S += 1e-9 * rng.standard_normal((Nh, n))
        lo, hi = rd.shard_range(n, world, rank)
        S_local = torch.from_numpy(np.asfortranarray(S[:, lo:hi]))
        ops = TorchCpuBlockOps()
        gram = rd.gram_column_sharded_overlapped if overlap else rd.gram_column_sharded
        G = gram(S_local, n, ops)
        Gref = S.T @ S
        assert np.max(np.abs(G.numpy() - Gref)) < 1e-12 * np.max(np.abs(Gref))
        assert torch.equal(G, G.t()), "assembled Gram matrix must be exactly symmetric"
        Q, s, energy, VT = rd.orth_column_sharded(S_local, n, num=6, normalize=True, block_ops=ops, overlap=overlap)
        np.save(os.path.join(out_dir, f"Q{rank}.npy"), Q.numpy())
        if rank == 0:
            np.save(os.path.join(out_dir, "s.npy"), s)
            np.save(os.path.join(out_dir, "S.npy"), S)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,overlap", [(2, False), (2, True), (3, True), (4, True), (4, False)])
def test_column_sharded_pod_gloo(tmp_path, world, overlap):
    import romtime_oracle as orc

    Nh, n = 400, 22  # 22 columns: ragged shards for world = 3 and 4
    mp.spawn(_worker, args=(world, _free_port(), Nh, n, overlap, str(tmp_path)), nprocs=world, join=True)
    S = np.load(tmp_path / "S.npy")
    s = np.load(tmp_path / "s.npy")
    Qr, sr, _ = orc.orth(S, num=6)
    assert np.max(np.abs(s[:6] - sr[:6]) / sr[:6]) < 1e-10
    Qs = [np.load(tmp_path / f"Q{r}.npy") for r in range(world)]
    for Q in Qs[1:]:
        assert np.array_equal(Q, Qs[0]), "every rank must hold the same lifted basis"
    sign = np.sign(np.sum(Qs[0] * Qr, axis=0))
    assert np.max(np.abs(Qs[0] * sign - Qr)) < 1e-10


def test_shard_range_partitions_exactly():
    from romtime_b200.distributed import column_shards, shard_range

    for total in (0, 1, 7, 4096, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            pieces = [shard_range(total, world, r) for r in range(world)]
            assert pieces[0][0] == 0 and pieces[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(pieces, pieces[1:]))
            sizes = [b - a for a, b in pieces]
            assert max(sizes) - min(sizes) <= 1
            assert pieces == column_shards(total, world)
