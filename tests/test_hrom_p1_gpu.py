"""End to end, FEniCS-free, on the two named problems (BASELINE.json configs[0], [1] as math): closed-form P1 full-order
model (oracle/p1_fom.py, pinned by the reference's own operator entries) -> reduced basis (device POD, batched tree walk)
-> MDEIM / DEIM / N-MDEIM offline (device POD + greedy) -> projection -> online through BOTH paths:
  (a) the reference's per-step protocol: host ``assemble(mu, t, entries)`` callbacks + one device step per time step
      (``RomConstructor*.solve``), and
  (b) the device sampled-entry evaluator built from the mesh, the sampled entries and the reduced basis
      (``Mfp1Evaluator.from_mesh`` / ``PistonEvaluator.from_mesh``, romtime_b200/evaluators.py) + one ``solve_batch`` sweep.
(a) and (b) must agree to 1e-10; the ROM must reproduce the FOM (and MFP1's manufactured solution)."""
import numpy as np
import pytest

import p1_fom
import p1_forms

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def _walk(ts, **kw):
    base = {"ts": ts, "num_snapshots": None}
    base.update(kw)
    return base


def test_mfp1_moving_hrom_end_to_end():
    from romtime_b200.deim import DiscreteEmpiricalInterpolation, MatrixDiscreteEmpiricalInterpolation
    from romtime_b200.rom import Mfp1Evaluator, RomConstructorMoving

    nx, L0, tf, nt = 60, 2.0, 1.0, 100
    fom = p1_fom.P1Fom(p1_forms.Mfp1(nx, L0, moving=True), tf=tf, nt=nt, bdf_scheme="2")
    rng = np.random.default_rng(7)
    space = [{"delta": rng.uniform(0.3, 1.2), "beta": rng.uniform(2, 6), "alpha_0": rng.uniform(0.3, 1.5), "omega": rng.uniform(0.3, 1.0)}
             for _ in range(6)]
    from scipy.stats import uniform

    grid = {"delta": uniform(0.3, 0.9), "beta": uniform(2.0, 4.0), "alpha_0": uniform(0.3, 1.2), "omega": uniform(0.3, 0.7)}
    rom = RomConstructorMoving(fom=fom, grid=grid, name="mfp1")
    rom.setup(rnd=0)
    rom.build_reduced_basis(mu_space=[dict(m) for m in space], tolerances={"tol_time": 1.0 - 1e-12, "tol_mu": 1.0 - 1e-12})
    assert 3 <= rom.N <= 30
    ts = [float(t) for t in np.linspace(0.02, tf, 15)]
    reds = {}
    # every operator is affine in a few (mu, t) features (romtime_b200/evaluators.py): mass = [Dirichlet 1's] + L m0, ... -- the
    # collateral bases are asked for exactly that many modes (an energy tolerance never keeps the mode that completes the energy)
    for which, cb, cls, k in (("mass", fom.assemble_mass, MatrixDiscreteEmpiricalInterpolation, 2),
                              ("stiffness", fom.assemble_stiffness, MatrixDiscreteEmpiricalInterpolation, 2),
                              ("convection", fom.assemble_convection, MatrixDiscreteEmpiricalInterpolation, 2),
                              ("rhs", fom.assemble_rhs, DiscreteEmpiricalInterpolation, 3)):
        r = cls(assemble=cb, name=which, grid=grid, tree_walk_params=_walk(ts, num_mu=k, num_time=k))
        r.setup(rnd=0)
        r.run(mu_space=[dict(m) for m in space])
        rom.add_hyper_reductor(reductor=r, which=which)
        reds[which] = r
    assert [reds[w].N for w in ("mass", "stiffness", "convection", "rhs")] == [2, 2, 2, 3]
    rom.project_reductors()
    online = [{"delta": 0.9, "beta": 3.3, "alpha_0": 0.8, "omega": 0.7}, {"delta": 0.5, "beta": 5.1, "alpha_0": 1.2, "omega": 0.4}]
    # (a) reference protocol: host callbacks every step
    traj = []
    for mu in online:
        rom.solve(mu=dict(mu), step="online")
        traj.append((np.array(rom.solutions.rom), np.array(rom.solutions.fom)))
    # (b) device evaluator, one sweep for both parameters
    ev = Mfp1Evaluator.from_mesh(rom, nx=nx, L0=L0)
    batch = rom.solve_batch([dict(m) for m in online], ev, nt=nt, dt=fom.dt, bdf=2)
    uN = batch.rom_device.cpu().numpy()
    assert not batch.info.any().item()
    # lazy FOM reconstruction of the whole batch: ONE lift GEMM, the boundary lifting riding along (rom/rom.py:496-522)
    full = batch.fom_batch(lifting=ev.p1).cpu().numpy()
    for b, mu in enumerate(online):
        assert rel(uN[b].T, traj[b][0]) < 1e-10
        assert rel(full[b].T, traj[b][1]) < 1e-10
        # against the full-order model and the manufactured solution (discretisation error of nx = 60, nt = 100: ~5e-4)
        fom.update_parametrization(mu)
        fom.solve()
        assert rel(traj[b][1], fom.solutions.fom) < 1e-6
        exact = np.stack([fom.exact(mu, (k + 1) * fom.dt) for k in range(nt)], axis=1)
        assert rel(traj[b][1], exact) < 2e-3


@pytest.mark.parametrize("which", ["rest", "sudden"])
def test_piston_hrom_host_callbacks_vs_device_evaluator(which):
    from romtime_b200.deim import (DiscreteEmpiricalInterpolation, MatrixDiscreteEmpiricalInterpolation,
                                   MatrixDiscreteEmpiricalInterpolationNonlinear)
    from romtime_b200.rom import PistonEvaluator, RomConstructorNonlinear

    nx, tf, nt = 80, 0.3, 60
    fom = p1_fom.P1Fom(p1_forms.Piston(nx, 1.0, which=which), tf=tf, nt=nt, bdf_scheme="2")
    rng = np.random.default_rng(11)
    space = [{"a0": rng.uniform(8, 10), "omega": rng.uniform(15, 20), "delta": rng.uniform(0.1, 0.15)} for _ in range(4)]
    from scipy.stats import uniform

    grid = {"a0": uniform(8.0, 2.0), "omega": uniform(15.0, 5.0), "delta": uniform(0.1, 0.05)}
    rom = RomConstructorNonlinear(fom=fom, grid=grid, name="piston")
    rom.setup(rnd=0)
    rom.build_reduced_basis(mu_space=[dict(m) for m in space], num_basis=12, tolerances={"tol_time": 1.0 - 1e-10})
    ts = [float(t) for t in np.linspace(0.01, tf, 10)]
    for name, cb in (("mass", fom.assemble_mass), ("stiffness", fom.assemble_stiffness), ("convection", fom.assemble_convection),
                     ("nonlinear-lifting", fom.assemble_nonlinear_lifting)):
        k = {"mass": 2, "stiffness": 2, "convection": 3, "nonlinear-lifting": 2}[name]   # Dirichlet 1 + the operator's own features
        r = MatrixDiscreteEmpiricalInterpolation(assemble=cb, name=name, grid=grid, tree_walk_params=_walk(ts, num_mu=k, num_time=k))
        r.setup(rnd=0)
        r.run(mu_space=[dict(m) for m in space])
        rom.add_hyper_reductor(reductor=r, which=name)
    d = DiscreteEmpiricalInterpolation(assemble=fom.assemble_rhs, name="lifting", grid=grid, tree_walk_params=_walk(ts, num_mu=5, num_time=5))
    d.setup(rnd=0)
    d.run(mu_space=[dict(m) for m in space])
    rom.add_hyper_reductor(reductor=d, which="lifting")
    nm = MatrixDiscreteEmpiricalInterpolationNonlinear(assemble=fom.assemble_trilinear, name="trilinear", grid=grid,
                                                       tree_walk_params=_walk(ts[:5], num_mu=12, num_time=8))
    nm.setup(rnd=0, V=rom.basis)
    nm.run(u_n=rom.basis[:, :6], mu_space=[dict(m) for m in space[:3]])
    rom.add_hyper_reductor(reductor=nm, which="trilinear")
    rom.project_reductors()
    online = [{"a0": 9.1, "omega": 17.3, "delta": 0.12}, {"a0": 8.4, "omega": 19.0, "delta": 0.14}]
    traj = []
    for mu in online:
        rom.solve(mu=dict(mu), step="online")
        traj.append(np.array(rom.solutions.rom))
    ev = PistonEvaluator.from_mesh(rom, nx=nx, L0=1.0, which=which)
    batch = rom.solve_batch([dict(m) for m in online], ev, nt=nt, dt=fom.dt, bdf=2)
    uN = batch.rom_device.cpu().numpy()
    assert not batch.info.any().item()
    for b in range(len(online)):
        assert rel(uN[b].T, traj[b]) < 1e-10
    # on-disk: the batch as the reference's per-parameter RomSolutionsStorage pickles, FOM fields rebuilt by one GEMM per block
    import pickle
    import tempfile

    from romtime_b200.storage import ResultTree

    with tempfile.TemporaryDirectory() as tmp:
        tree = ResultTree(tmp)
        files = tree.dump_batch(batch, prefix="solutions_online", lifting=ev.p1)
        tree.dump_mu_space(rom)
        tree.dump_reduced_basis(rom)
        tree.dump_collateral_bases(rom)
        st = pickle.load(open(files[1], "rb"))
        assert st.rom.shape == (rom.N, nt) and st.fom.shape == (nx + 1, nt) and st.mu["a0"] == online[1]["a0"]
        rom.solve(mu=dict(online[1]), step="online")
        assert rel(st.fom, np.array(rom.solutions.fom)) < 1e-10
        fresh = rom.mdeim_Mh.copy()
        tree.load_collateral_basis(fresh)
        assert fresh.dofs == rom.mdeim_Mh.dofs and np.array_equal(fresh.PT_U, rom.mdeim_Mh.PT_U)
    # the hyper-reduced model follows the full-order one (N = 12 modes, trilinear term interpolated)
    fom.update_parametrization(online[0])
    fom.solve()
    lifted = rom.basis @ traj[0]
    # The bound is loose on purpose.  The collateral bases of this problem are requested beyond their numerical rank (the mass
    # and stiffness entries of the uniformly scaled mesh span ONE direction besides the Dirichlet rows, the trilinear walk has
    # fewer than 12 independent snapshots directions), so the last modes and the interpolation points they select are decided
    # by rounding: the reference's own classes on this FOM (scipy gesvd, GMRES) give 0.036, this package 0.025 or 0.081
    # depending on the summation order inside the Gram / Jacobi kernels -- with mass / stiffness / trilinear bases that do not
    # even span the same subspace (distance ~ 1), while the reduced basis itself agrees to 5e-7.  The parity checks above
    # (device evaluator against host callbacks at 1e-10) do not depend on that choice.
    print("hrom_vs_fom", which, rel(lifted, fom.solutions.snapshots))
    assert rel(lifted, fom.solutions.snapshots) < 0.2
