"""TMA + DMMA GEMM core through its C-ABI users (Gram, cross-Gram, lift): fp64 results against
NumPy (CPU oracle arithmetic) and, at sizes NumPy cannot do in seconds, against a plain torch fp64
matmul on the device.  Tolerance: 1e-13 relative to ||A|| ||B|| (fp64 accumulation, different order)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b, scale):
    return np.max(np.abs(a - b)) / scale


@pytest.mark.parametrize("order", ["C", "F"])
@pytest.mark.parametrize("Nh,n", [(1, 1), (5, 3), (37, 8), (300, 24), (1000, 129), (4096, 256), (10007, 130), (2000, 64)])
@pytest.mark.parametrize("normalize", [False, True])
def test_gram_vs_numpy(Nh, n, order, normalize):
    from romtime_b200 import kernels

    rng = np.random.default_rng(Nh + 7 * n)
    S = rng.standard_normal((Nh, n)) * (1.0 + np.arange(n))
    S = np.asfortranarray(S) if order == "F" else np.ascontiguousarray(S)
    G, cn = kernels.pod_gram(S, normalize=normalize)
    G, cn = G.cpu().numpy(), cn.cpu().numpy()
    ref_cn = np.linalg.norm(S, axis=0)
    Sn = S / ref_cn if normalize else S
    ref = Sn.T @ Sn
    assert np.allclose(cn, ref_cn, rtol=1e-13)
    assert rel(G, ref, np.max(np.abs(ref))) < 1e-13
    assert np.array_equal(G, G.T)  # exactly symmetric


def test_gram_views():
    from romtime_b200 import kernels

    rng = np.random.default_rng(3)
    big = rng.standard_normal((1003, 77))
    for view in (big[1:, 2:41], np.asfortranarray(big)[3:, 5:70], big[::2, ::3]):
        G, cn = kernels.pod_gram(view, normalize=False)
        ref = view.T @ view
        assert rel(G.cpu().numpy(), ref, np.max(np.abs(ref))) < 1e-13


@pytest.mark.parametrize("oa,ob", [("C", "C"), ("C", "F"), ("F", "C"), ("F", "F")])
def test_cross_gram(oa, ob):
    from romtime_b200 import kernels

    rng = np.random.default_rng(9)
    A = rng.standard_normal((5000, 70))
    B = rng.standard_normal((5000, 200))
    A = np.asfortranarray(A) if oa == "F" else A
    B = np.asfortranarray(B) if ob == "F" else B
    C = kernels.gram_cross(A, B).cpu().numpy()
    ref = A.T @ B
    assert rel(C, ref, np.max(np.abs(ref))) < 1e-13


@pytest.mark.parametrize("order", ["C", "F"])
@pytest.mark.parametrize("Nh,n,r", [(300, 24, 6), (5000, 130, 64), (4097, 64, 1), (999, 33, 33), (20000, 256, 200), (40013, 200, 64), (40013, 96, 65)])
def test_lift(Nh, n, r, order):
    from romtime_b200 import kernels

    rng = np.random.default_rng(Nh + n + r)
    S = rng.standard_normal((Nh, n))
    S = np.asfortranarray(S) if order == "F" else S
    W = rng.standard_normal((n, r))
    for qorder in ("F", "C"):
        Q = kernels.pod_lift(S, W, order=qorder).cpu().numpy()
        ref = S @ W
        assert Q.shape == ref.shape
        assert rel(Q, ref, np.max(np.abs(ref))) < 1e-13


def test_gram_large_vs_torch():
    """Split-k + symmetric tile schedule at a size with several waves (device fp64 matmul as check)."""
    import torch

    from romtime_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(7)
    for shape, order in (((60000, 1024), "C"), ((60000, 1024), "F"), ((200000, 512), "C")):
        S = torch.randn(shape, dtype=torch.float64, device="cuda", generator=g)
        if order == "F":
            S = S.t().contiguous().t()
        G, cn = kernels.pod_gram(S, normalize=False)
        ref = S.t() @ S
        err = (G - ref).abs().max().item() / ref.abs().max().item()
        assert err < 1e-13, (shape, order, err)
        assert torch.equal(G, G.t())
