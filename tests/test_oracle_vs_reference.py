"""CPU, build container only: a LIVE re-check of the oracle's pin -- the oracle (NumPy/SciPy restatement) against the
unmodified reference imported through oracle/ref_harness.py on fresh seeded inputs (not the committed fixtures).
Skipped where no reference tree is present (/root/reference or the vendored oracle/_ref)."""
import numpy as np
import pytest

import ref_harness
import romtime_oracle as orc
from romtime_b200 import synthetic

pytestmark = pytest.mark.skipif(not ref_harness.reference_available(), reason="no reference tree (/root/reference or oracle/_ref)")


def test_orth_matches_reference_pod():
    ref_harness.load_reference()
    from romtime.rom.pod import orth as ref_orth  # rom/pod.py:7-62

    rng = np.random.default_rng(5)
    S = rng.standard_normal((300, 12)) @ np.diag(10.0 ** -np.arange(12)) @ rng.standard_normal((12, 40))
    for kw in (dict(num=5), dict(tol=1.0 - 1e-8), dict(num=4, normalize=False, return_VT=True)):
        a, b = orc.orth(S, **kw), ref_orth(S, **kw)
        assert len(a) == len(b) and a[0].shape == b[0].shape
        np.testing.assert_allclose(a[1], b[1], rtol=1e-13, atol=1e-15 * b[1][0])
        np.testing.assert_allclose(a[2], b[2], rtol=1e-13)
        np.testing.assert_allclose(np.abs(np.sum(a[0] * b[0], axis=0)), 1.0, atol=1e-12)


def test_greedy_matches_reference_build_interpolation_mesh():
    ref_harness.load_reference()
    from romtime.deim.deim import DiscreteEmpiricalInterpolation  # deim/deim.py:517-561

    rng = np.random.default_rng(6)
    U, _ = np.linalg.qr(rng.standard_normal((400, 9)))
    U[37] = U[36]  # an exact tie: the first maximum wins
    deim = DiscreteEmpiricalInterpolation(assemble=None, name="x")
    deim.basis_fom = U
    dofs, P = deim.build_interpolation_mesh()
    assert [int(d) for d in dofs] == orc.deim_greedy(U)
    assert np.array_equal(P.T @ U, orc.interpolation_matrix(U, orc.deim_greedy(U)))


@pytest.mark.parametrize("bdf", ["1", "2"])
def test_rom_sweep_matches_reference_solve(bdf):
    """RomConstructorNonlinear.solve (rom/rom.py:430-555), unmodified, with the algebraic-solver hook swapped for a direct
    solve, against the oracle's rom_sweep on the same synthetic hyper-reduced model."""
    model = synthetic.make_piston_model(N=12, m=7, m_rhs=5, seed=77)
    V, _ = np.linalg.qr(np.random.default_rng(7).standard_normal((50, 12)))
    mu_row = synthetic.sample_mus(3, seed=8)[2]
    nt, dt = 40, 1e-3
    rom = ref_harness.build_reference_rom(model, V, bdf, dt, nt, True)
    rom.solve(mu={"a0": mu_row[0], "omega": mu_row[1], "delta": mu_row[2]}, step="online")
    ref = np.array(rom.solutions.rom)
    out = orc.rom_sweep(model, mu_row, nt, dt, bdf2=(bdf == "2"), solver="direct")
    assert np.max(np.abs(out - ref)) <= 1e-12 * np.max(np.abs(ref))
