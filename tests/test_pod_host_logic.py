"""Host logic of the POD (guard band, Hestenes refinement passes, deflation levels, truncation rules) on CPU, with the
device block operations replaced by NumPy stand-ins of the same numerical contract (tests/pod_emulation.py).  What is
asserted here is what the GPU tests assert through the real kernels: every retained mode of the reference's gesvd
(golden vectors from the unmodified reference, rom/pod.py:7-62) to 1e-10."""
import ast

import numpy as np
import pytest

import pod_emulation
from pod_checks import check_orth_result
import romtime_oracle as orc
from romtime_b200 import synthetic


def check_against(S, kw, Qr, sr, er, VTr=None, rtol=1e-10):
    from romtime_b200.rom.pod import orth_device

    kw = dict(kw)
    kw.pop("return_VT", None)
    Q, s, energy, VT, info = orth_device(S, **kw)
    check_orth_result(Q.numpy(), s, energy, VT.numpy(), Qr, sr, er, VTr, rtol)
    return info


def test_golden_orth_through_host_logic(golden, monkeypatch):
    pod_emulation.install(monkeypatch)
    z = golden("orth")
    for name in z["names"]:
        kw = ast.literal_eval(str(z[f"{name}/kw"]))
        VTr = z[f"{name}/VT"] if f"{name}/VT" in z.files else None
        check_against(z[f"{name}/S"], kw, z[f"{name}/Q"], z[f"{name}/s"], z[f"{name}/energy"], VTr)


def test_guard_band_and_single_pass(monkeypatch):
    """sigma_r / sigma_1 = 3e-4 (the C4 regime): the guard band must widen the rotation, one pass must do."""
    pod_emulation.install(monkeypatch)
    rng = np.random.default_rng(5)
    n, Nh = 60, 400
    U, _ = np.linalg.qr(rng.standard_normal((Nh, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    sig = 10.0 ** (-np.arange(n) / 6.0)
    S = (U * sig) @ V.T
    Qr, sr, er = orc.orth(S, num=22, normalize=False)
    info = check_against(S, dict(num=22, normalize=False), Qr, sr, er)
    assert info["candidates"] > 22 and info["passes"] == 1 and info["levels"] == 1
    assert 1e-12 < info["defect"] < 1e-2


def test_default_rule_reaches_below_the_gram_floor(monkeypatch):
    """rom/pod.py:56 keeps s > 1e-7 ABSOLUTE: with sigma_1 ~ 10 that is 1e-8 relative, below what one Gram matrix
    resolves -- the deflated second level must deliver those modes (and stop at the noise)."""
    pod_emulation.install(monkeypatch)
    rng = np.random.default_rng(6)
    n, Nh = 24, 300
    U, _ = np.linalg.qr(rng.standard_normal((Nh, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    sig = np.concatenate([3.0 * 10.0 ** (1.0 - np.arange(12)), np.zeros(12)])  # 30 ... 3e-10, then exact rank deficiency
    S = (U * sig) @ V.T
    Qr, sr, er = orc.orth(S, normalize=False)
    assert Qr.shape[1] == 9                       # 30, 3, ..., 3e-7 > 1e-7
    info = check_against(S, dict(normalize=False), Qr, sr, er, rtol=2e-9)   # sigma_9/sigma_1 = 1e-8: gesvd's own limit
    assert info["levels"] >= 2


def test_num_beyond_numerical_rank_returns_orthonormal_noise_modes(monkeypatch):
    pod_emulation.install(monkeypatch)
    from romtime_b200.rom.pod import orth_device

    S = synthetic.mdeim_snapshots(600, 40, seed=1004)
    Qr, sr, er = orc.orth(S, num=10, normalize=False)
    Q, s, energy, VT, info = orth_device(S, num=10, normalize=False)
    Q = Q.numpy()
    assert Q.shape == Qr.shape
    assert np.max(np.abs(Q.T @ Q - np.eye(10))) < 1e-8      # the reference's gesvd also returns orthonormal noise vectors
    assert np.max(np.abs(s - sr)) < 1e-10 * sr[0]


@pytest.mark.parametrize("kw", [dict(num=24), dict(tol=1 - 1e-10), dict()])
def test_c3_like(monkeypatch, kw):
    pod_emulation.install(monkeypatch)
    S = synthetic.pod_snapshots(3000, 96, seed=1003, modes=48, decay=8.0, noise=1e-13)
    Qr, sr, er = orc.orth(S, **kw)
    check_against(S, kw, Qr, sr, er, rtol=1e-9 if not kw else 1e-10)
