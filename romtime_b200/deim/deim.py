"""Discrete Empirical Interpolation on B200 -- drop-in for the reference's deim/deim.py.

Same class name, constructor, attributes (``basis_fom, basis_rom, PT_U, dofs, sigmas, N_V, N, Nh``) and
methods as the reference; host attributes stay NumPy arrays / Python lists.  The arithmetic goes through
the C ABI (include/romtime_b200.h):

    build_interpolation_mesh  -> rt_deim_greedy_f64      (reference deim/deim.py:517-561)
    compute_thetas            -> rt_invert_f64 once + rt_theta_f64   (:477-493)
    _interpolate              -> rt_theta_f64 + rt_assemble_batch_f64 (:416-452)
    project_basis             -> rt_project_vec_f64      (:495-515)
    tree_walk / walk_time     -> romtime_b200.rom.pod.orth (:255-398)

``interpolate_batch`` is the batched form the online sweep uses (no reference counterpart: the
reference evaluates one (mu, t) at a time).
"""
from copy import deepcopy

import numpy as np

from ..conventions import EmpiricalInterpolation, RomParameters, Stage
from ..rom.base import Reductor
from ..rom.pod import orth, orth_batch
from ..utils import dump_pickle, functional_to_array, read_pickle


class SelectionMatrix:
    """The one-hot matrix ``P`` (Nh x m) that build_interpolation_mesh returns (reference
    deim/deim.py:18-22, :556-557), kept as its index list: ``P.T @ X`` is the row gather ``X[dofs]``;
    ``np.asarray(P)`` materialises the dense matrix for callers that really want it."""

    def __init__(self, Nh, dofs, transposed=False):
        self.Nh = int(Nh)
        self.dofs = np.asarray(dofs, dtype=np.int64)
        self._t = transposed

    @property
    def shape(self):
        return (len(self.dofs), self.Nh) if self._t else (self.Nh, len(self.dofs))

    @property
    def T(self):
        return SelectionMatrix(self.Nh, self.dofs, not self._t)

    def __array__(self, dtype=None, copy=None):
        P = np.zeros((self.Nh, len(self.dofs)), dtype=dtype or float)
        P[self.dofs, np.arange(len(self.dofs))] = 1.0
        return P.T if self._t else P

    def __matmul__(self, other):
        if self._t:
            return np.asarray(other)[self.dofs]
        return np.asarray(self) @ other

    def dot(self, other):
        return self.__matmul__(other)


class DiscreteEmpiricalInterpolation(Reductor):

    TYPE = EmpiricalInterpolation.DEIM

    def __init__(self, assemble, grid=None, tree_walk_params=None, name=None):
        super().__init__(grid=grid)
        self.name = name
        self.assemble = assemble
        self.tree_walk_params = tree_walk_params

        self.N_V = None      # size of the solution reduced basis the collateral basis was projected on
        self.PT_U = None     # interpolation matrix  P^T U  (m x m)
        self.sigmas = None
        self.dofs = None     # list of tuples: (dof,) for vectors, (row, col) for matrices

        self.basis_fom = None
        self.basis_rom = None
        self.snapshots = None

        self.basis_pickle_name = self._define_basis_name()
        self._dev = {}       # device-side mirrors, keyed by attribute name -> (id(host array), tensor)

    # ------------------------------------------------------------------ naming / printing
    def _define_basis_name(self):
        label = "_".join(str(self.name).lower().split())
        return f"basis_fom_{self.TYPE.lower()}_{label}.pkl"

    def __str__(self):
        return f"{self.TYPE} - {self.name}"

    __repr__ = __str__

    @property
    def Nh(self):
        return self.basis_fom.shape[0]

    @property
    def N(self):
        """Collateral basis size."""
        return self.basis_fom.shape[1]

    def __getstate__(self):
        state = dict(self.__dict__)
        state["_dev"] = {}  # CUDA mirrors are rebuilt on demand (deepcopy / pickle safe)
        return state

    def copy(self):
        """New reductor sharing ``assemble`` / ``grid`` with deep copies of the data
        (reference deim/deim.py:110-131)."""
        new = self.__class__(assemble=self.assemble, grid=self.grid, tree_walk_params=self.tree_walk_params,
                             name=self.name)
        for attr in ("basis_fom", "basis_rom", "PT_U", "dofs", "errors_rom"):
            value = getattr(self, attr)
            if value is not None:
                setattr(new, attr, deepcopy(value))
        return new

    # ------------------------------------------------------------------ device mirrors
    def _mirror(self, attr, build):
        host = getattr(self, attr)
        hit = self._dev.get(attr)
        if hit is not None and hit[0] is host:
            return hit[1]
        tensor = build(host)
        self._dev[attr] = (host, tensor)
        return tensor

    def _PT_U_inverse(self):
        from .. import kernels

        if self.PT_U is None:
            raise ValueError(f"{self}: no interpolation matrix yet (run / load_fom_basis first)")
        return self._mirror("PT_U", lambda a: kernels.invert(np.ascontiguousarray(a, dtype=np.float64)))

    def _basis_device(self, which):
        from .. import kernels

        attr = "basis_rom" if which == self.ROM else "basis_fom"
        if getattr(self, attr) is None:
            raise ValueError(f"{self}: {attr} is not available")
        return self._mirror(attr, lambda a: kernels.to_device(np.asarray(a, dtype=np.float64)))

    # ------------------------------------------------------------------ offline: basis + interpolation mesh
    def _finalise_basis(self):
        """Greedy point selection + interpolation matrix for the current ``basis_fom``."""
        from .. import kernels

        dofs_d, PT_U_d = kernels.deim_greedy(self.basis_fom)
        self.store_dofs([int(d) for d in dofs_d.cpu().tolist()])
        self.PT_U = PT_U_d.cpu().numpy()  # == P.T @ basis_fom (reference deim/deim.py:159, :212)

    def load_fom_basis(self, keep=None, basis=None):
        """Load a FOM collateral basis (from ``basis`` or the pickle on disk) and build the interpolation
        mesh for it (reference deim/deim.py:133-164)."""
        if basis is None:
            basis = read_pickle(self.basis_pickle_name)
        if keep:
            basis = basis[:, :keep]
        self.basis_fom = basis
        self._finalise_basis()

    def dump_fom_basis(self, path=None):
        if self.basis_fom is None:
            raise ValueError(f"Trying to dump basis for {self.name} without building it!")
        dump_pickle(self.basis_pickle_name, obj=self.basis_fom)

    def run(self, normalize=True, mu_space=None):
        """DEIM offline phase: tree walk (POD in time per mu, then POD in mu), greedy interpolation mesh
        (reference deim/deim.py:175-215)."""
        p = self.tree_walk_params
        Vfh, sigmas = self.tree_walk(
            ts=p[RomParameters.TS],
            num_snapshots=p[RomParameters.NUM_SNAPSHOTS],
            num_mu=p.get(RomParameters.NUM_MU, None),
            num_t=p.get(RomParameters.NUM_TIME, None),
            tol_mu=p.get(RomParameters.TOL_MU, None),
            tol_t=p.get(RomParameters.TOL_TIME, None),
            normalize=normalize,
            mu_space=mu_space,
        )
        self.basis_fom = Vfh
        self.sigmas = sigmas
        self._finalise_basis()

    def store_dofs(self, dofs):
        """Vector entries as 1-tuples (reference deim/deim.py:217-224)."""
        self.dofs = [(int(dof),) for dof in dofs]

    def build_interpolation_mesh(self):
        """Classic DEIM greedy on ``basis_fom``.  Returns ``(interpolation_dofs, P)`` like the reference
        (deim/deim.py:517-561); ``P`` is a ``SelectionMatrix`` (``P.T @ basis`` works, ``np.asarray(P)``
        gives the dense one-hot matrix).  Indices are bit-exact with the reference, first maximum on ties."""
        from .. import kernels

        dofs_d, _ = kernels.deim_greedy(self.basis_fom)
        dofs = [int(d) for d in dofs_d.cpu().tolist()]
        return dofs, SelectionMatrix(self.basis_fom.shape[0], dofs)

    # ------------------------------------------------------------------ offline: snapshots and tree walk
    def _assemble_functional(self, mu, t):
        return functional_to_array(self.assemble(mu=mu, t=t))

    def assemble_snapshot(self, mu, t):
        return self._assemble_functional(mu, t)

    def _time_snapshots(self, mu, ts):
        """Snapshots over ``ts`` for a frozen ``mu`` (reference deim/deim.py:378-388; MDEIM zeroes the boundary entry 0)."""
        snapshots = np.array([self.assemble_snapshot(mu, t) for t in ts]).T
        if self.TYPE == EmpiricalInterpolation.MDEIM:
            snapshots[0, :] = 0.0
        return snapshots

    def walk_time(self, mu, ts, normalize=True, num=None, tol=None):
        """Time branch: snapshots over ``ts`` for a frozen ``mu`` -> POD (never normalised, reference
        deim/deim.py:355-398)."""
        return orth(snapshots=self._time_snapshots(mu, ts), num=num, tol=tol, normalize=False)

    def tree_walk(self, ts, normalize=True, num_mu=None, num_t=None, tol_mu=None, tol_t=None,
                  num_snapshots=None, mu_space=None):
        """Two-level POD over parameter and time (reference deim/deim.py:255-353).
        Returns (basis, sigmas_mu)."""
        space = mu_space if mu_space else self.build_sampling_space(num=num_snapshots, rnd=self.random_state)
        off = self.report[Stage.OFFLINE]
        for key in (self.SPECTRUM_TIME, self.ENERGY_TIME, self.BASIS_TIME):
            off.setdefault(key, dict())
        # the assemble callbacks run on the host, parameter after parameter (as in the reference); the per-parameter time-PODs
        # (deim/deim.py:390-395) are independent of each other and go to the device as ONE batch (rom/pod.py orth_batch)
        ids, snaps = [], []
        for mu in space:
            mu_idx, mu = self.add_mu(step=Stage.OFFLINE, mu=mu)
            ids.append(mu_idx)
            snaps.append(self._time_snapshots(mu, ts))
        blocks = []
        for mu_idx, (basis_t, sigmas_t, energy_t) in zip(ids, orth_batch(snaps, num=num_t, tol=tol_t, normalize=False)):
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
    def compute_thetas(self, rhs):
        """Interpolation coefficients ``PT_U^-1 rhs`` (reference deim/deim.py:477-493).  The inverse is
        formed once on the device (rt_invert_f64) and reused; ``rhs`` may be (m,) or a batch (B, m)."""
        from .. import kernels

        rhs = np.asarray(rhs, dtype=np.float64)
        single = rhs.ndim == 1
        theta = kernels.theta_batch(self._PT_U_inverse(), rhs.reshape(1, -1) if single else rhs)
        out = theta.cpu().numpy()
        return out[0] if single else out

    def interpolate_batch(self, f_local, which=None):
        """Batched ``_interpolate``: ``f_local`` (B, m) sampled entries (NumPy or CUDA) ->
        CUDA tensor (B, R), R = rows of the chosen basis.  No boundary fix-up, no reshape."""
        from .. import kernels

        which = self.FOM if which is None else which
        theta = kernels.theta_batch(self._PT_U_inverse(), f_local)
        return kernels.assemble_batch(self._basis_device(which), theta)

    def _local_entries(self, mu, t, **extra):
        return np.asarray(self.assemble(mu=mu, t=t, entries=self.dofs, **extra), dtype=np.float64)

    def _interpolate(self, mu, t, which=None, **extra):
        """Interpolated operator in vector form for one (mu, t) (reference deim/deim.py:416-452):
        sampled entries from the ``assemble`` callback, theta solve, ``sum_i theta_i Vf[:, i]``."""
        which = self.FOM if which is None else which
        fh_local = self._local_entries(mu, t, **extra)
        approximation = self.interpolate_batch(fh_local.reshape(1, -1), which=which)[0].cpu().numpy()
        if which == self.FOM and self.TYPE == EmpiricalInterpolation.MDEIM:
            approximation[0] = 1.0  # boundary entry (reference deim/deim.py:449-450)
        return approximation

    def interpolate(self, mu, t, which=None):
        return self._interpolate(mu=mu, t=t, which=which)

    def evaluate(self, ts, num=None, mu_space=None):
        """Online check: RMS error of the FOM-space interpolation against the assembled operator for
        sampled parameters (reference deim/deim.py:226-253)."""
        if mu_space:
            space = mu_space
        else:
            assert num, "Provide number of samples to test"
            space = self.build_sampling_space(num=num)
        for mu in space:
            mu_idx, mu = self.add_mu(step=Stage.ONLINE, mu=mu)
            for t in ts:
                exact = self.assemble_snapshot(mu, t)
                approx = self._interpolate(mu, t, which=self.FOM)
                self.errors_rom[mu_idx].append(self._compute_error(u=approx, ue=exact))
            self.errors_rom[mu_idx] = np.array(self.errors_rom[mu_idx])

    # ------------------------------------------------------------------ projection onto the solution basis
    def project_basis(self, V):
        """``basis_rom = V^T basis_fom``  (N_V x m)  (reference deim/deim.py:495-515)."""
        from .. import kernels

        self.basis_rom = kernels.project_vec(V, self.basis_fom).cpu().numpy()
        self.N_V = V.shape[1]
