"""POD: eigensolver and ``orth`` against the oracle (NumPy/SciPy gesvd) and the reference-generated
golden vectors.  Tolerances (north_star): singular values 1e-10 relative, bases up to sign 1e-10
-- asserted on EVERY retained mode with s_k/s_0 >= 1e-6 (tests/pod_checks.py; DESIGN.md "POD accuracy")."""
import ast

import numpy as np
import pytest

import romtime_oracle as orc
from pod_checks import check_orth_result
from romtime_b200 import synthetic

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 2, 5, 24, 63, 64, 65, 100, 128, 200, 333])
def test_sym_eig_random_psd(n):
    import torch

    from romtime_b200 import kernels

    rng = np.random.default_rng(n)
    X = rng.standard_normal((n + 5, n)) * (10.0 ** (-np.arange(n) * 3.0 / max(n, 2)))
    G = X.T @ X
    w, W, sweeps = kernels.sym_eig(torch.from_numpy(G.copy()).cuda())
    w, W = w.cpu().numpy(), W.cpu().numpy()
    ref = np.linalg.eigvalsh(G)[::-1]
    assert sweeps > 0
    assert np.all(np.diff(w) <= 0)
    assert np.max(np.abs(w - ref)) <= 1e-13 * ref[0]
    # eigenvectors come back as L X lam^-1/2: orthogonality degrades like eps * sqrt(lam_0/lam_k)
    assert np.max(np.abs(W.T @ W - np.eye(n))) < 5e-12
    assert np.max(np.abs(W @ np.diag(w) @ W.T - G)) < 1e-13 * ref[0]


def test_sym_eig_1024():
    import torch

    from romtime_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(3)
    n = 1024
    X = torch.randn((3000, n), dtype=torch.float64, device="cuda", generator=g)
    X = X * torch.logspace(0, -6, n, dtype=torch.float64, device="cuda")
    G = X.t() @ X
    w, W, sweeps = kernels.sym_eig(G.clone())
    ref = torch.linalg.eigvalsh(G).flip(0)
    assert (w - ref).abs().max().item() <= 1e-13 * ref[0].item()
    assert (W.t() @ W - torch.eye(n, dtype=torch.float64, device="cuda")).abs().max().item() < 1e-10
    assert (W @ torch.diag(w) @ W.t() - G).abs().max().item() < 1e-12 * ref[0].item()


def _check_orth(S, kw, Qr, sr, er, VTr=None, rtol=1e-10):
    from romtime_b200.rom.pod import orth

    res = orth(S, **kw)
    VT = res[3] if len(res) == 4 else None
    check_orth_result(res[0], res[1], res[2], VT, Qr, sr, er, VTr if VT is not None else None, rtol)


def test_golden_orth(golden):
    z = golden("orth")
    for name in z["names"]:
        S = z[f"{name}/S"]
        if str(z[f"{name}/order"]) == "F":
            S = np.asfortranarray(S)
        kw = ast.literal_eval(str(z[f"{name}/kw"]))
        VTr = z[f"{name}/VT"] if f"{name}/VT" in z.files else None
        _check_orth(S, kw, z[f"{name}/Q"], z[f"{name}/s"], z[f"{name}/energy"], VTr)


def test_list_input_raises():
    from romtime_b200.rom.pod import orth

    with pytest.raises(ValueError):
        orth([[1.0, 2.0], [3.0, 4.0]])


@pytest.mark.parametrize("order", ["C", "F"])
def test_orth_c3_like_vs_oracle(order):
    S = synthetic.pod_snapshots(20000, 256, seed=1003, modes=64, decay=16.0, noise=1e-13, order=order)
    Qr, sr, er = orc.orth(S, num=24)
    _check_orth(S, dict(num=24), Qr, sr, er)


def test_orth_mdeim_like_vs_oracle():
    S = synthetic.mdeim_snapshots(30000, 128, seed=1004)
    Qr, sr, er = orc.orth(S, num=16, normalize=False)
    _check_orth(S, dict(num=16, normalize=False), Qr, sr, er)
    Qr, sr, er = orc.orth(S, tol=1 - 1e-8)
    _check_orth(S, dict(tol=1 - 1e-8), Qr, sr, er)


def test_orth_deep_spectrum_every_retained_mode():
    """sigma_r / sigma_1 = 1e-5 with further modes below: guard band + one Hestenes pass (round-1 bar was 1e-3)."""
    rng = np.random.default_rng(5)
    n, Nh = 96, 5000
    U, _ = np.linalg.qr(rng.standard_normal((Nh, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    sig = 10.0 ** (-np.arange(n) / 8.0)
    S = (U * sig) @ V.T
    Qr, sr, er = orc.orth(S, num=41, normalize=False)
    assert sr[40] / sr[0] < 1.1e-5
    _check_orth(S, dict(num=41, normalize=False), Qr, sr, er)


def test_orth_default_rule_below_gram_floor():
    """rom/pod.py:56 keeps s > 1e-7 ABSOLUTE; with sigma_1 = 30 that is 3e-9 relative: the deflated levels deliver them."""
    from romtime_b200.rom.pod import orth_device

    rng = np.random.default_rng(6)
    n, Nh = 24, 3000
    U, _ = np.linalg.qr(rng.standard_normal((Nh, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    sig = np.concatenate([3.0 * 10.0 ** (1.0 - np.arange(12)), np.zeros(12)])
    S = (U * sig) @ V.T
    Qr, sr, er = orc.orth(S, normalize=False)
    assert Qr.shape[1] == 9
    _check_orth(S, dict(normalize=False), Qr, sr, er, rtol=2e-9)
    Q, s, energy, VT, info = orth_device(S, normalize=False)
    assert info["levels"] >= 2
    # against the constructed singular vectors: as close as gesvd itself
    Qh = Q.cpu().numpy()
    for k in range(9):
        ours = min(np.abs(Qh[:, k] - U[:, k]).max(), np.abs(Qh[:, k] + U[:, k]).max())
        theirs = min(np.abs(Qr[:, k] - U[:, k]).max(), np.abs(Qr[:, k] + U[:, k]).max())
        assert ours < 10.0 * theirs + 1e-13, (k, ours, theirs)


def test_num_beyond_numerical_rank_is_orthonormal():
    S = synthetic.mdeim_snapshots(600, 40, seed=1004)
    from romtime_b200.rom.pod import orth

    Q, s, energy = orth(S, num=10, normalize=False)
    assert Q.shape == (S.shape[0], 10)
    assert np.max(np.abs(Q.T @ Q - np.eye(10))) < 1e-8


@pytest.mark.parametrize("n", [1, 3, 40, 64, 65, 130, 200])
def test_sym_eig_jacobi_relative_accuracy(n):
    """Plain Jacobi on the Gram matrix of nearly orthogonal GRADED columns: every eigenvalue to a small relative error."""
    import torch

    from romtime_b200 import kernels

    rng = np.random.default_rng(n)
    Nh = n + 50
    Qo, _ = np.linalg.qr(rng.standard_normal((Nh, n)))
    sig = 10.0 ** (-8.0 * np.arange(n) / max(n - 1, 1))
    mix = np.eye(n) + 1e-4 * rng.standard_normal((n, n))
    Y = (Qo @ mix) * sig                        # nearly orthogonal directions, THEN graded: H = D (I + O(1e-4)) D
    H = Y.T @ Y
    lam, X, sweeps, defect = kernels.sym_eig_jacobi(torch.from_numpy(H).cuda())
    lam, X = lam.cpu().numpy(), X.cpu().numpy()
    import pod_emulation

    ref, _, _ = pod_emulation.jacobi(H)       # scalar cyclic Jacobi: relatively accurate on scaled diagonally dominant H
    assert np.all(np.diff(lam) <= 0) and sweeps > 0
    assert np.max(np.abs(lam - ref) / ref) < 1e-12
    assert np.max(np.abs(X.T @ X - np.eye(n))) < 1e-12
    if n > 1:
        assert 0 < defect < 1e-2


def _smooth_snapshots(Nh, n, seed, modes=18, decay=2.5, noise=1e-13):
    rng = np.random.default_rng(seed)
    x = np.linspace(0.0, 1.0, Nh)[:, None]
    tt = np.linspace(0.0, 1.0, n)[None, :]
    S = sum(10.0 ** (-k / decay) * np.sin((k + 1) * np.pi * x + rng.random()) * np.cos((k + 1) * 3.0 * tt + rng.random() * 6)
            for k in range(modes))
    return S + noise * rng.standard_normal((Nh, n))


@pytest.mark.parametrize("kw", [dict(tol=1.0 - 1e-8), dict(num=6), dict(tol=1.0 - 1e-6, normalize=False), dict(num=9, normalize=False)])
def test_orth_batch_matches_gesvd_per_matrix(kw):
    """The batched tree-walk path (rt_small_gram_eig_f64 + rt_small_refine_f64, two host reads per BATCH): ragged sizes,
    tol-driven ragged widths, and two members that must fall back to the general path (300 snapshots; numerical rank above
    96).  Every member against the reference's per-matrix gesvd at 1e-10."""
    from romtime_b200.rom.pod import orth_batch

    mats = [_smooth_snapshots(Nh, n, seed) for seed, (Nh, n) in enumerate([(100, 50), (1000, 250), (400, 10), (333, 77), (64, 64), (2000, 128)])]
    mats.append(np.asfortranarray(_smooth_snapshots(500, 120, 17)))                       # F-order member (np.array(list).T)
    mats.append(_smooth_snapshots(700, 300, 18))                                          # more than 256 snapshots: general path
    mats.append(np.random.default_rng(19).standard_normal((400, 150)))                    # full numerical rank: general path
    res = orth_batch(mats, **kw)
    assert len(res) == len(mats)
    for S, (Q, s, energy) in zip(mats, res):
        Qr, sr, er = orc.orth(S, **kw)
        check_orth_result(Q, s, energy, None, Qr, sr, er)


def test_small_eig_kernels_against_lapack():
    """rt_small_gram_eig_f64 / rt_small_refine_f64 on their own: eigenvalues to a small relative error, orthonormal vectors."""
    import torch

    from romtime_b200 import kernels

    rng = np.random.default_rng(3)
    sizes = [7, 64, 200, 256]
    nmax = max(sizes)
    G_all = torch.zeros((len(sizes), nmax * nmax), dtype=torch.float64, device="cuda")
    Gs = []
    for j, n in enumerate(sizes):
        X = rng.standard_normal((n + 40, min(n, 30))) * 10.0 ** (-np.arange(min(n, 30)) / 4.0)
        A = X @ rng.standard_normal((min(n, 30), n))
        G = A.T @ A
        Gs.append(G)
        G_all[j, : n * n] = torch.from_numpy(G.reshape(-1)).cuda()
    n_of = torch.tensor(sizes, dtype=torch.int32, device="cuda")
    w_all, W_all, rank, info = kernels.small_gram_eig(G_all, n_of, nmax)
    assert not info.any().item()
    for j, n in enumerate(sizes):
        r = int(rank[j].item())
        w = w_all[j, :n].cpu().numpy()
        W = W_all[j, : n * n].reshape(n, n).cpu().numpy()
        ref = np.linalg.eigvalsh(Gs[j])[::-1]
        assert 0 < r <= min(n, 30) and np.all(w[r:] == 0.0)
        assert np.max(np.abs(w[:r] - ref[:r])) <= 1e-13 * ref[0]
        assert np.max(np.abs(W[:, :r].T @ W[:, :r] - np.eye(r))) < 1e-11
        assert np.max(np.abs((W[:, :r] * w[:r]) @ W[:, :r].T - Gs[j])) < 1e-13 * ref[0]
    # refinement: Gram matrix of nearly orthogonal graded columns -> every eigenvalue to a small RELATIVE error
    cs = [5, 40, 96]
    cmax = max(cs)
    H_all = torch.zeros((len(cs), cmax * cmax), dtype=torch.float64, device="cuda")
    Hs = []
    for j, c in enumerate(cs):
        Qo, _ = np.linalg.qr(rng.standard_normal((300, c)))
        Y = (Qo + 1e-4 * rng.standard_normal((300, c))) * 10.0 ** (-np.arange(c) / 8.0)
        H = Y.T @ Y
        Hs.append((H, Y))
        H_all[j, : c * c] = torch.from_numpy(H.reshape(-1)).cuda()
    lam_all, X_all, defect, info2 = kernels.small_refine(H_all, torch.tensor(cs, dtype=torch.int32, device="cuda"), cmax)
    assert not info2.any().item()
    for j, c in enumerate(cs):
        H, Y = Hs[j]
        lam = lam_all[j, :c].cpu().numpy()
        X = X_all[j, : c * c].reshape(c, c).cpu().numpy()
        sv = np.linalg.svd(Y, compute_uv=False)
        assert np.max(np.abs(np.sqrt(lam) - sv) / sv) < 1e-12          # relative, on a spectrum spanning 1e-12
        assert np.max(np.abs(X.T @ X - np.eye(c))) < 1e-13
        assert 1e-5 < float(defect[j].item()) < 1e-2


@pytest.mark.parametrize("c", [8, 57, 64, 65, 72, 80, 81, 96, 97])
@pytest.mark.parametrize("order", ["C", "F"])
def test_lift_narrow_tiles_against_torch(c, order):
    """The GEMM core's narrow instantiations (one column of tiles, 64 / 80 / 96 columns per warp row) and the full tile next
    to them: Q = S W for ragged row counts against torch's fp64 matmul."""
    import torch

    from romtime_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(c)
    S = torch.randn((20_011, 133), dtype=torch.float64, device="cuda", generator=g)
    W = torch.randn((133, c), dtype=torch.float64, device="cuda", generator=g)
    Q = kernels.pod_lift(S, W, order=order)
    ref = S @ W
    assert Q.shape == ref.shape
    assert (Q - ref).abs().max().item() < 1e-11 * ref.abs().max().item() * 133 ** 0.5


@pytest.mark.parametrize("n", [5, 64, 72, 96, 100])
def test_gram_of_few_columns_deep_split(n):
    """Gram matrix of a tall matrix with ONE output tile: the k-range is split over the whole GPU (deterministic fixed-order
    reduction of the partial tiles), narrow tiles for n <= 96."""
    import torch

    from romtime_b200 import kernels

    g = torch.Generator(device="cuda").manual_seed(n)
    Y = torch.randn((300_007, n), dtype=torch.float64, device="cuda", generator=g) * torch.pow(
        10.0, -torch.arange(n, dtype=torch.float64, device="cuda") / 8.0)
    H = kernels.pod_gram(Y, normalize=False)[0]
    H2 = kernels.pod_gram(Y, normalize=False)[0]
    assert torch.equal(H, H2)                      # bitwise reproducible
    assert torch.equal(H, H.t())                   # exactly symmetric
    ref = Y.t() @ Y
    scale = torch.sqrt(torch.outer(torch.diag(ref), torch.diag(ref)))
    assert ((H - ref).abs() / scale).max().item() < 1e-12
