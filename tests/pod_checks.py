"""Shared assertion helper of the POD tests (CPU host-logic tests with emulated block ops, GPU tests with the real
kernels): the bars of north_star -- singular values and bases (up to sign) within 1e-10 relative of the reference's
gesvd (rom/pod.py:38) on EVERY retained mode with sigma_k >= 1e-6 sigma_1."""
import numpy as np


def check_orth_result(Q, s, energy, VT, Qr, sr, er, VTr=None, rtol=1e-10):
    assert s.shape == sr.shape and energy.shape == er.shape and Q.shape == Qr.shape, (Q.shape, Qr.shape)
    r = Q.shape[1]
    # every RETAINED mode: singular value 1e-10 relative, vector up to sign 1e-10 (unit vectors) wherever gesvd itself
    # defines it that well (sigma_k >= 1e-6 sigma_1; below that two LAPACK builds differ by more than 1e-10)
    tight = sr[:r] >= 1e-6 * sr[0]
    assert np.max(np.abs(s[:r][tight] - sr[:r][tight]) / sr[:r][tight], initial=0.0) < rtol
    # modes that are NOT returned keep the first Gram matrix's resolution: absolute, eps sigma_1^2 / sigma_k, and zero
    # below its numerical rank (sigma_k < 4.2e-8 sigma_1)
    assert np.max(np.abs(s - sr)) < 5e-8 * sr[0]
    assert np.max(np.abs(energy - er)) < 1e-12
    sign = np.sign(np.sum(Q * Qr, axis=0))
    assert np.max(np.abs(Q * sign - Qr)[:, tight], initial=0.0) < rtol
    noise = sr[:r] < 1e-12 * sr[0]                          # gesvd's own vectors are arbitrary there
    ok = ~noise
    assert np.max(np.abs(Q.T @ Q - np.eye(r))[np.ix_(ok, ok)], initial=0.0) < rtol
    loose = ok & ~tight
    if loose.any():
        assert np.max(np.abs(Q * sign - Qr)[:, loose]) < 1e-6
    if VTr is not None:
        assert np.max(np.abs(VT * sign[:, None] - VTr)[tight], initial=0.0) < rtol
