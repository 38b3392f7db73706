"""Nonlinear MDEIM (N-MDEIM) on B200 -- drop-in for the reference's deim/nonlinear.py: an MDEIM whose
operator also depends on a state function ``u_n`` (the trilinear Burgers term).  Offline, the snapshots
are taken over (mu, t, reduced-basis function) with a three-level POD; online, ``u_n`` is forwarded to
the ``assemble`` callback when the sampled entries are evaluated."""
from copy import deepcopy

import numpy as np

from ..conventions import EmpiricalInterpolation, RomParameters, Stage, Treewalk
from ..rom.base import Reductor
from ..rom.pod import orth, orth_batch
from ..utils import bilinear_to_csr, eliminate_zeros, vector_to_csr
from .mdeim import MatrixDiscreteEmpiricalInterpolation, _sorted_topology


class MatrixDiscreteEmpiricalInterpolationNonlinear(MatrixDiscreteEmpiricalInterpolation):

    TYPE = EmpiricalInterpolation.NONLINEAR

    def __init__(self, assemble, name=None, grid=None, tree_walk_params=None):
        super().__init__(assemble, name=name, grid=grid, tree_walk_params=tree_walk_params)
        self.u_n = None  # functions the operator is linearised around (columns), set by run()

    def copy(self):
        new = super().copy()
        if self.u_n is not None:
            new.u_n = deepcopy(self.u_n)
        return new

    # ------------------------------------------------------------------ offline
    def truncate(self, n):
        """N-MDEIM with the last ``n`` collateral modes removed and a fresh interpolation mesh
        (reference deim/nonlinear.py:49-103)."""
        N = self.N
        assert n < N, "You want to remove too many modes from S-NonlinearMDEIM to create NonlinearMDEIM."
        truncated = self.__class__(assemble=self.assemble, grid=self.grid, tree_walk_params=self.tree_walk_params,
                                   name="S-" + str(self.name))
        Reductor.setup(self=truncated, rnd=self.random_state)
        truncated.rows, truncated.cols = self.rows, self.cols
        truncated.basis_fom = self.basis_fom[:, : N - n]
        truncated._finalise_basis()
        truncated.mu_space = deepcopy(self.mu_space)
        truncated.report = deepcopy(self.report)
        truncated.report[Stage.OFFLINE][Treewalk.BASIS_FINAL] = truncated.N
        return truncated

    def get_matrix_topology(self, mu, t, u_n):
        return _sorted_topology(eliminate_zeros(self._assemble_matrix(mu=mu, t=t, u_n=u_n)))

    def setup(self, rnd, V):
        """Seed + topology.  The reference linearises around the FE interpolant of ``x`` on the function
        space ``V`` (deim/nonlinear.py:133-157, FEniCS); a plain array is used as ``u_n`` directly."""
        Reductor.setup(self=self, rnd=rnd)
        mu = list(self.build_sampling_space(num=1))[0]
        if isinstance(V, np.ndarray):
            u_n = V
        else:
            import fenics  # only reached with a real FEniCS function space

            u_n = fenics.interpolate(fenics.Expression("x[0]", degree=1), V)
        self.rows, self.cols = self.get_matrix_topology(mu=mu, t=1.0, u_n=u_n)

    def run(self, u_n, mu_space=None):
        """N-MDEIM offline phase (reference deim/nonlinear.py:159-212)."""
        u_n = np.asarray(u_n)
        self.u_n = u_n.reshape(u_n.shape[0], 1) if u_n.ndim == 1 else u_n
        p = self.tree_walk_params
        get = p.get
        Vfh, sigmas = self.tree_walk(
            ts=p[RomParameters.TS], normalize=True, num_mu=get(RomParameters.NUM_MU), num_t=get(RomParameters.NUM_TIME),
            num_basis=get(RomParameters.NUM_BASIS), tol_mu=get(RomParameters.TOL_MU), tol_t=get(RomParameters.TOL_TIME),
            tol_basis=get(RomParameters.TOL_BASIS), num_snapshots=p[RomParameters.NUM_SNAPSHOTS], mu_space=mu_space)
        self.basis_fom = Vfh
        self.sigmas = sigmas
        self._finalise_basis()

    def _assemble_matrix(self, mu, t, u_n):
        return bilinear_to_csr(matrix=self.assemble(mu=mu, t=t, u_n=u_n))

    def assemble_snapshot(self, mu, t, u_n):
        return eliminate_zeros(self._assemble_matrix(mu, t, u_n)).data

    def walk_time(self, mu, ts, normalize=True, num_t=None, tol_t=None, num_basis=None, tol_basis=None):
        """Time branch with an inner POD over the linearisation functions at every instant, then a POD of
        the stacked per-instant bases (reference deim/nonlinear.py:399-466)."""
        return orth(snapshots=self._instant_bases(mu, ts, normalize, num_t, tol_t), num=num_t, tol=tol_t, normalize=normalize)

    def _instant_bases(self, mu, ts, normalize, num_t, tol_t):
        """Inner level of the walk (deim/nonlinear.py:435-458): one POD over the linearisation functions per instant --
        independent problems, one device batch -- stacked side by side."""
        per_instant = []
        for t in ts:
            snaps = np.array([self.assemble_snapshot(mu=mu, t=t, u_n=self.u_n[:, k]) for k in range(self.u_n.shape[1])]).T
            snaps[0, :] = 0.0
            per_instant.append(snaps)
        return np.hstack([b for b, _, _ in orth_batch(per_instant, num=num_t, tol=tol_t, normalize=normalize)])

    def tree_walk(self, ts, normalize=True, num_mu=None, num_t=None, num_basis=None, tol_mu=None, tol_t=None,
                  tol_basis=None, num_snapshots=None, mu_space=None):
        """Parameter branch (reference deim/nonlinear.py:296-397).  Returns (basis, sigmas_mu)."""
        space = mu_space if mu_space else self.build_sampling_space(num=num_snapshots, rnd=self.random_state)
        off = self.report[Stage.OFFLINE]
        for key in (self.SPECTRUM_TIME, self.ENERGY_TIME, self.BASIS_TIME):
            off.setdefault(key, dict())
        # three levels (deim/nonlinear.py:367-397, :435-466): per instant (batched inside _instant_bases), per parameter
        # (one batch over the parameters), then the mu-POD of the stacked bases
        ids, stacks = [], []
        for mu in space:
            mu_idx, mu = self.add_mu(step=Stage.OFFLINE, mu=mu)
            ids.append(mu_idx)
            stacks.append(self._instant_bases(mu, ts, normalize, num_t, tol_t))
        blocks = []
        for mu_idx, (basis_t, sigmas_t, energy_t) in zip(ids, orth_batch(stacks, num=num_t, tol=tol_t, normalize=normalize)):
            off[self.SPECTRUM_TIME][mu_idx] = sigmas_t
            off[self.ENERGY_TIME][mu_idx] = energy_t
            off[self.BASIS_TIME][mu_idx] = basis_t.shape[1]
            blocks.append(basis_t)
        stacked = np.hstack(blocks)
        off[self.BASIS_AFTER_WALK] = stacked.shape[1]
        basis, sigmas_mu, energy_mu = orth(snapshots=stacked, num=num_mu, tol=tol_mu, normalize=normalize)
        off[self.SPECTRUM_MU] = sigmas_mu
        off[self.ENERGY_MU] = energy_mu
        off[self.BASIS_FINAL] = basis.shape[1]
        return basis, sigmas_mu

    # ------------------------------------------------------------------ online
    def _interpolate(self, mu, t, u_n, which=None):
        """Vector-form interpolation with the state forwarded to the callback (reference
        deim/nonlinear.py:247-283); FOM space pins entry 0 to 1 (:280-281)."""
        which = self.FOM if which is None else which
        fh_local = self._local_entries(mu, t, u_n=u_n)
        approximation = self.interpolate_batch(fh_local.reshape(1, -1), which=which)[0].cpu().numpy()
        if which == self.FOM:
            approximation[0] = 1.0
        return approximation

    def interpolate(self, mu, t, u_n, which=None):
        approximation = self._interpolate(mu, t, u_n, which=which)
        if which == self.ROM:
            return approximation.reshape((self.N_V, self.N_V))
        return vector_to_csr(entries=approximation, rows=self.rows, cols=self.cols)

    def evaluate(self, ts, funcs=None, num=None, mu_space=None):
        """Mean RMS interpolation error over the linearisation functions (reference
        deim/nonlinear.py:468-555)."""
        if mu_space:
            space = mu_space
        else:
            assert num, "Provide number of samples to test"
            space = self.build_sampling_space(num=num)
        funcs = self.u_n if funcs is None else funcs
        n_psi = funcs.shape[1]
        for mu in space:
            mu_idx, mu = self.add_mu(step=Stage.ONLINE, mu=mu)
            for t in ts:
                err = 0.0
                for k in range(n_psi):
                    psi = funcs[:, k]
                    exact = self.assemble_snapshot(mu=mu, t=t, u_n=psi)
                    approx = self._interpolate(mu=mu, t=t, u_n=psi, which=self.FOM)
                    err += self._compute_error(u=approx, ue=exact)
                self.errors_rom[mu_idx].append(err / n_psi)
            self.errors_rom[mu_idx] = np.array(self.errors_rom[mu_idx])
