"""CPU-side checks of the reference-facing Python layer: names / signatures mirror the reference's API
(SURVEY.md section 8b), host-only helpers behave like the reference's, and every compute entry point
FAILS LOUDLY without a GPU (there is no CPU fallback)."""
import inspect

import numpy as np
import pytest
import torch

from romtime_b200 import base, conventions, utils
from romtime_b200._lib import RomtimeB200Error
from romtime_b200.deim import (DiscreteEmpiricalInterpolation, MatrixDiscreteEmpiricalInterpolation,
                               MatrixDiscreteEmpiricalInterpolationNonlinear, SelectionMatrix)
from romtime_b200.rom import Reductor, RomConstructor, RomConstructorMoving, RomConstructorNonlinear, orth

NO_GPU = not torch.cuda.is_available()


def params(fn):
    return [p for p in inspect.signature(fn).parameters if p != "self"]


def test_signatures_mirror_the_reference():
    # reference rom/pod.py:7, deim/deim.py:29-35, :133, :175, :517, deim/mdeim.py:28-34, rom/rom.py:38, :430
    assert params(orth) == ["snapshots", "num", "tol", "normalize", "return_VT"]
    assert params(DiscreteEmpiricalInterpolation.__init__) == ["assemble", "grid", "tree_walk_params", "name"]
    assert params(MatrixDiscreteEmpiricalInterpolation.__init__) == ["assemble", "name", "grid", "tree_walk_params"]
    assert params(DiscreteEmpiricalInterpolation.load_fom_basis) == ["keep", "basis"]
    assert params(DiscreteEmpiricalInterpolation.run) == ["normalize", "mu_space"]
    assert params(DiscreteEmpiricalInterpolation.compute_thetas) == ["rhs"]
    assert params(DiscreteEmpiricalInterpolation.project_basis) == ["V"]
    assert params(DiscreteEmpiricalInterpolation.interpolate) == ["mu", "t", "which"]
    assert params(DiscreteEmpiricalInterpolation.evaluate) == ["ts", "num", "mu_space"]
    assert params(DiscreteEmpiricalInterpolation.tree_walk) == ["ts", "normalize", "num_mu", "num_t", "tol_mu", "tol_t",
                                                                "num_snapshots", "mu_space"]
    assert params(MatrixDiscreteEmpiricalInterpolationNonlinear.interpolate) == ["mu", "t", "u_n", "which"]
    assert params(MatrixDiscreteEmpiricalInterpolationNonlinear.setup) == ["rnd", "V"]
    assert params(MatrixDiscreteEmpiricalInterpolationNonlinear.run) == ["u_n", "mu_space"]
    assert params(RomConstructor.__init__) == ["fom", "grid", "name"]
    assert params(RomConstructor.solve) == ["mu", "step"]
    assert params(RomConstructor.add_hyper_reductor) == ["reductor", "which"]
    assert params(RomConstructor.build_reduced_basis) == ["num_snapshots", "mu_space", "num_basis", "tolerances"]
    assert params(RomConstructorNonlinear.assemble_system) == ["mu", "t", "bdf", "uh", "uh_n1"]
    assert params(RomConstructorNonlinear.assemble_system_rhs) == ["mu", "t", "MN_mat", "uN_n", "uN_n1"]
    assert issubclass(RomConstructorNonlinear, RomConstructorMoving) and issubclass(RomConstructorMoving, RomConstructor)
    assert issubclass(RomConstructor, Reductor) and issubclass(DiscreteEmpiricalInterpolation, Reductor)


def test_convention_strings():
    c = conventions
    assert (c.ProblemType.FOM, c.ProblemType.ROM) == ("fom", "rom")
    assert (c.Stage.OFFLINE, c.Stage.ONLINE, c.Stage.VALIDATION) == ("offline", "online", "validation")
    assert (c.BDF.ONE, c.BDF.TWO) == ("1", "2")
    assert c.OperatorType.NONLINEAR_LIFTING == "nonlinear-lifting" and c.OperatorType.ROM == "rom"
    assert c.Treewalk.BASIS_AFTER_WALK == "basis-shape-after-tree-walk"
    assert c.TreewalkNonlinear.SPECTRUM_TIME == "N-spectrum-time"
    assert c.EmpiricalInterpolation.NONLINEAR == "N-MDEIM"
    assert c.RomParameters.NUM_BASIS == "num_phi" and c.RomParameters.TS == "ts"
    assert DiscreteEmpiricalInterpolation(assemble=None, name="Forcing Term").basis_pickle_name == "basis_fom_deim_forcing_term.pkl"
    assert MatrixDiscreteEmpiricalInterpolation(assemble=None, name="mass").basis_pickle_name == "basis_fom_mdeim_mass.pkl"


def test_reductor_bookkeeping():
    r = Reductor(grid=None)
    r.setup(rnd=7)
    assert r.random_state == 7
    assert r.report["offline"]["spectrum-time"] == {} and r.report["offline"]["N-basis-shape-final"] is None
    i0, _ = r.add_mu("offline", {"a": 1.0})
    i1, _ = r.add_mu("offline", {"a": 2.0})
    i2, _ = r.add_mu("offline", {"a": 1.0})  # list.index semantics of the reference: first occurrence
    assert (i0, i1, i2) == (0, 1, 0) and r.mu == {"a": 1.0}
    u, ue = np.arange(4.0), np.zeros(4)
    assert np.isclose(Reductor._compute_error(u, ue), np.linalg.norm(u) / 2.0)
    r.errors_rom[0] = np.array([1.0, 3.0])
    r.create_errors_summary()
    assert r.summary_errors.loc[0, "mean"] == 2.0 and r.summary_errors.loc[0, "max"] == 3.0


def test_selection_matrix_is_a_one_hot_P():
    P = SelectionMatrix(6, [4, 0, 2])
    X = np.arange(12.0).reshape(6, 2)
    assert P.shape == (6, 3) and P.T.shape == (3, 6)
    dense = np.asarray(P)
    assert dense.sum() == 3 and dense[4, 0] == dense[0, 1] == dense[2, 2] == 1.0
    assert np.array_equal(P.T @ X, dense.T @ X)
    assert np.array_equal(np.matmul(P.T, X), X[[4, 0, 2]])


def test_mdeim_store_dofs_and_copy():
    md = MatrixDiscreteEmpiricalInterpolation(assemble=None, name="m")
    md.rows, md.cols = [0, 0, 1, 1, 1, 2], [0, 1, 0, 1, 2, 2]
    md.store_dofs([4, 0])
    assert md.dofs == [(1, 2), (0, 0)] and md.get_entry(3) == (1, 1)
    md.basis_fom = np.ones((6, 2))
    c = md.copy()
    assert c.rows == md.rows and c.rows is not md.rows and c.Nh == 6 and c.N == 2
    d = DiscreteEmpiricalInterpolation(assemble=None, name="v")
    d.store_dofs(np.array([3, 1]))
    assert d.dofs == [(3,), (1,)]


def test_host_helpers_match_reference_formulas():
    import scipy.sparse as sp

    s = np.array([3.0, 2.0, 1.0])
    assert np.allclose(utils.singular_to_energy(s), np.cumsum(s ** 2) / 14.0)
    assert np.allclose(utils.singular_to_pod_error(s), np.sqrt(14.0 - np.cumsum(s ** 2)))
    V = np.eye(4)[:, :3]
    assert np.isclose(utils.compute_rom_difference(np.array([1.0, 2.0]), np.array([1.0, 2.0, 0.5]), V), 0.5 / 2.0)
    A = sp.csr_matrix(np.array([[1.0, 1e-17], [0.0, 2.0]]))
    assert utils.eliminate_zeros(A).nnz == 2
    st = base.RomSolutionsStorage(ts=[0.1, 0.2], mu={"a0": 2.0}, domain=np.array([[1.0, 1.0], [0.0, 0.0]]),
                                  fom=np.array([[3.0, 5.0], [1.0, 1.0]]), rom=np.zeros((1, 2)))
    assert np.allclose(st.compute_at(0.5), [4.0, 6.0])  # midpoint of (1, 3) and (1, 5), scaled by a0 = 2


@pytest.mark.skipif(not NO_GPU, reason="checks the no-GPU failure mode")
def test_compute_paths_fail_loudly_without_gpu():
    d = DiscreteEmpiricalInterpolation(assemble=None, name="v")
    d.basis_fom = np.linalg.qr(np.random.default_rng(0).standard_normal((50, 4)))[0]
    with pytest.raises(RomtimeB200Error):
        d.build_interpolation_mesh()
    with pytest.raises(RomtimeB200Error):
        d.project_basis(np.ones((50, 2)))
    d.PT_U = np.eye(4)
    with pytest.raises(RomtimeB200Error):
        d.compute_thetas(np.ones(4))
    with pytest.raises(RomtimeB200Error):
        orth(np.ones((10, 3)))
    md = MatrixDiscreteEmpiricalInterpolation(assemble=None, name="m")
    md.basis_fom, md.rows, md.cols = np.ones((3, 1)), [0, 1, 2], [0, 1, 2]
    with pytest.raises(RomtimeB200Error):
        md.project_basis(np.ones((3, 2)))
