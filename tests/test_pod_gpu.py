"""POD: eigensolver and ``orth`` against the oracle (NumPy/SciPy gesvd) and the reference-generated
golden vectors.  Tolerances (north_star): singular values 1e-10 relative, bases up to sign 1e-10
-- asserted on the modes the Gram route can resolve (s_k/s_0 >= 1e-3, DESIGN.md "POD accuracy")."""
import ast

import numpy as np
import pytest

import romtime_oracle as orc
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


def _check_orth(S, kw, Qr, sr, er, VTr=None):
    from romtime_b200.rom.pod import orth

    res = orth(S, **kw)
    Q, s, energy = res[0], res[1], res[2]
    assert s.shape == sr.shape and energy.shape == er.shape
    assert Q.shape == Qr.shape, (Q.shape, Qr.shape)
    good = sr >= 1e-3 * sr[0]
    assert np.max(np.abs(s[good] - sr[good]) / sr[good]) < 1e-10
    # unresolved tail: absolute agreement at the Gram noise floor
    assert np.max(np.abs(s - sr)) < 2e-7 * sr[0]
    assert np.max(np.abs(energy - er)) < 1e-12
    r = Q.shape[1]
    gr = good[:r]
    sign = np.sign(np.sum(Q * Qr, axis=0))
    assert np.max(np.abs(Q[:, gr] * sign[gr] - Qr[:, gr])) < 1e-10 * max(1.0, 1.0)  # unit vectors
    assert np.max(np.abs(Q.T @ Q - np.eye(r))[np.ix_(gr, gr)]) < 1e-10
    if VTr is not None:
        VT = res[3]
        assert np.max(np.abs(VT[gr] * sign[gr, None] - VTr[gr])) < 1e-10


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
