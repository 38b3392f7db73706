"""DEIM/MDEIM greedy: CUDA kernel (through the C ABI) vs the CPU oracle and the reference-generated
golden vectors.  Acceptance: indices bit-exact, PT_U equal to 0 ulp (it is a gather)."""
import numpy as np
import pytest

import romtime_oracle as orc
from romtime_b200 import synthetic

pytestmark = pytest.mark.gpu


def run(U):
    from romtime_b200 import kernels

    dofs, PT_U = kernels.deim_greedy(U)
    return dofs.cpu().numpy(), PT_U.cpu().numpy()


def test_golden_greedy(golden):
    z = golden("greedy")
    for name in z["names"]:
        basis = z[f"{name}/basis"]
        if name.endswith("_F"):
            basis = np.asfortranarray(basis)
        dofs, PT_U = run(basis)
        assert dofs.tolist() == z[f"{name}/dofs"].tolist(), name
        assert np.array_equal(PT_U, z[f"{name}/PT_U"]), name


@pytest.mark.parametrize("order", ["C", "F"])
@pytest.mark.parametrize("Nh,m", [(1, 1), (7, 3), (1000, 17), (4097, 33), (20000, 64), (50001, 128)])
def test_random_orthonormal_vs_oracle(Nh, m, order):
    rng = np.random.default_rng(Nh * 131 + m)
    m = min(m, Nh)
    Q, _ = np.linalg.qr(rng.standard_normal((Nh, m)))
    Q = np.asfortranarray(Q) if order == "F" else np.ascontiguousarray(Q)
    dofs, PT_U = run(Q)
    ref = orc.deim_greedy(Q)
    assert dofs.tolist() == ref
    assert np.array_equal(PT_U, Q[ref, :])


def test_duplicated_rows_ties():
    """Symmetric FEM matrices give bit-identical rows in the MDEIM basis: the first one wins."""
    S = synthetic.mdeim_snapshots(30000, 96, seed=1004)
    Q, _, _ = orc.orth(S, num=24, normalize=False)
    dofs, PT_U = run(Q)
    ref = orc.deim_greedy(Q)
    assert dofs.tolist() == ref
    # every selected off-diagonal entry must be the FIRST of its duplicated pair
    for d in dofs:
        dup = np.flatnonzero((Q == Q[d]).all(axis=1))
        assert d == dup.min()


def test_unaligned_view_and_strides():
    rng = np.random.default_rng(5)
    big = rng.standard_normal((3001, 40))
    view = big[1:, 3:23]  # C-order view with an offset and a row stride of 40
    dofs, PT_U = run(view)
    assert dofs.tolist() == orc.deim_greedy(view)
    fview = np.asfortranarray(big)[1:, 3:23]  # F-order view, odd leading dimension
    dofs, PT_U = run(fview)
    assert dofs.tolist() == orc.deim_greedy(fview)


def test_singular_raises():
    U = np.zeros((50, 3))
    U[:, 0] = 1.0
    U[:, 1] = 2.0  # second column in the span of the first -> zero residual -> singular at step 2
    U[:, 2] = np.arange(50)
    with pytest.raises(np.linalg.LinAlgError):
        run(U)
    with pytest.raises(np.linalg.LinAlgError):
        orc.deim_greedy(U)


def test_full_size_properties():
    """C4 size (200,000 x 128): properties that do not need the oracle at full size."""
    Nh, m = 200_000, 128
    rng = np.random.default_rng(1004)
    Q, _ = np.linalg.qr(rng.standard_normal((Nh, m)))
    Q = np.asfortranarray(Q)
    dofs, PT_U = run(Q)
    assert len(set(dofs.tolist())) == m
    assert np.array_equal(PT_U, Q[dofs, :])
    # DEIM residual of u_j vanishes at the previously selected points and peaks at p_j
    for j in (1, 17, 64, 127):
        c = np.linalg.solve(Q[dofs[:j], :j], Q[dofs[:j], j])
        res = Q[:, j] - Q[:, :j] @ c
        assert np.max(np.abs(res[dofs[:j]])) < 1e-12
        assert np.abs(res[dofs[j]]) >= np.max(np.abs(res)) * (1 - 1e-12)
