"""Reduced-order model constructors on B200 -- drop-in for the reference's rom/rom.py.

``RomConstructor`` / ``RomConstructorMoving`` / ``RomConstructorNonlinear`` keep the reference's names,
attributes (``basis``, ``solutions``, ``mdeim_Mh`` ... ``deim_rhs``) and methods.  What changes is where
the arithmetic runs:

* ``build_reduced_basis``: every ``orth`` is the device POD (rom/pod.py); the FOM solves stay on the host.
* ``solve(mu, step)``: per time step the sampled entries of ALL hyper-reductors are collected from their
  host ``assemble`` callbacks (FEniCS in real runs) into one row, and ONE call to rt_online_step_f64 does
  theta solves, reduced-operator assembly, right-hand side and the dense pivoted-LU solve on the device
  (the reference assembles five N x N operators with Python loops and calls GMRES, rom/rom.py:486-492).
* ``solve_batch(mus, ...)``: B parameters at once -- with a device evaluator (closed-form sampled
  entries, e.g. the piston problem) the whole BDF sweep runs on the device (rt_online_sweep_f64), sharded
  by parameter across GPUs without any collective (romtime_b200/distributed.py).

The linear solve is a direct LU with partial pivoting instead of GMRES(atol = tol = 1e-10); results agree
with the reference's GMRES to its own tolerance and with a direct solve to ~1e-13 (tests/).
"""
from copy import deepcopy
from pprint import pprint

import numpy as np

from ..base import RomBatchSolutions, RomSolutionsStorage
from ..conventions import BDF, OperatorType, PistonParameters, RomParameters, Stage, Treewalk, TreewalkNonlinear
from ..utils import bilinear_to_csr, function_to_array, functional_to_array, project_csr
from .base import Reductor
from .pod import orth, orth_batch

# device operator slots, in the order of rt_rom_model.ops (include/romtime_b200.h)
_SLOTS = ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear", "rhs")


class PistonEvaluator:
    """Closed-form sampled entries for the device sweep: for operator ``op`` the entries at the DEIM
    points are ``C[op] @ phi(mu, t)`` (+ ``b0 * T @ u*`` for the trilinear term), ``phi`` being the 12
    piston features of include/romtime_b200.h (RT_NF).  ``which`` = "rest" | "sudden"
    (reference problems/piston.py:18-60).  ``from_mesh`` derives the tables from the P1 mesh, the sampled entries of the
    attached reductors and the reduced basis (romtime_b200/evaluators.py) -- the device replacement of the per-step
    ``assemble_local`` calls (fom/base.py:523-599)."""

    evaluator_id = None                       # None: 0 / 1 by ``which``
    mu_names = (PistonParameters.A0, PistonParameters.OMEGA, PistonParameters.DELTA)

    def __init__(self, tables, which="rest"):
        self.tables = tables  # {slot name: {"C": (m, 12) array, "T": (m, N) array or None}}
        self.which = which

    @classmethod
    def from_mesh(cls, rom, nx, L0=1.0, which="rest"):
        from ..evaluators import PistonP1

        p1 = PistonP1(nx, L0=L0, which=which)
        ev = cls(_tables_from_p1(rom, p1), which=which)
        ev.p1 = p1      # also knows the boundary lifting at the nodes (RomBatchSolutions.fom_batch(lifting=ev.p1))
        return ev


class Mfp1Evaluator(PistonEvaluator):
    """The same for the MFP1 heat problem on the fixed or moving domain (problems/mfp1.py:27-75; device feature set 2,
    parameters (delta, beta, alpha_0, omega), omega = 0 on the fixed domain)."""

    evaluator_id = 2
    mu_names = ("delta", "beta", "alpha_0", "omega")

    def __init__(self, tables):
        super().__init__(tables, which="rest")

    @classmethod
    def from_mesh(cls, rom, nx, L0):
        from ..evaluators import Mfp1P1

        p1 = Mfp1P1(nx, L0)
        ev = cls(_tables_from_p1(rom, p1))
        ev.p1 = p1
        return ev


def _tables_from_p1(rom, p1):
    """{slot: {"C", "T"}} for every reductor attached to ``rom`` from the closed-form P1 tables of ``p1``; the sampled
    entries are the reductors' own ``dofs`` ((i,) or (row, col) tuples, deim/deim.py:217-224, deim/mdeim.py:102-124)."""
    out = {}
    for name, red in rom._slot_reductors().items():
        reds = [r for r in (red if isinstance(red, list) else [red]) if r]
        if not reds:
            continue
        if name not in p1.OPS:
            raise ValueError(f"the {type(p1).__name__} evaluator has no closed form for the operator slot {name!r}")
        if len(reds) > 1:
            raise ValueError("the closed-form evaluators take ONE right-hand-side reductor (forcing + lifting together)")
        parts = [p1.tables(name, [tuple(d) for d in r.dofs], V=rom.basis) for r in reds]
        tab = {"C": np.vstack([p["C"] for p in parts])}
        if any("T" in p for p in parts):
            tab["T"] = np.vstack([p["T"] for p in parts])
        out[name] = tab
    return out


def _is_sparse_like(op):
    import scipy.sparse as sp

    return sp.issparse(op) or (isinstance(op, np.ndarray) and op.ndim == 2) or type(op).__name__ in ("Matrix", "PETScMatrix")


class RomConstructor(Reductor):

    def __init__(self, fom, grid, name=None):
        super().__init__(grid=grid)
        self.fom = fom
        self.name = name
        self.basis = None
        self.basis_nonlinear = None
        self.solutions = dict()
        self.errors = dict()
        self.exact = dict()
        # vector reductors
        self.deim_fh = None       # forcing
        self.deim_fgh = None      # lifting
        self.deim_rhs = None      # forcing + lifting in one
        # matrix reductors
        self.mdeim_Mh = None      # mass
        self.mdeim_Ah = None      # stiffness
        self.mdeim_Ch = None      # convection
        self.mdeim_Nh = None      # trilinear (N-MDEIM)
        self.mdeim_Nh_hat = None  # nonlinear lifting
        self._dev_model = None
        self._dev_basis = None

    # ------------------------------------------------------------------ shapes
    @property
    def N(self):
        """Number of reduced basis elements."""
        return self.basis.shape[1]

    @property
    def shape(self):
        return self.basis.shape

    @property
    def timesteps(self):
        return self.solutions.ts

    # ------------------------------------------------------------------ FOM <-> ROM transfers (device GEMMs)
    def _basis_device(self):
        from .. import kernels

        if self._dev_basis is None or self._dev_basis[0] is not self.basis:
            self._dev_basis = (self.basis, kernels.to_device(np.asarray(self.basis, dtype=np.float64)))
        return self._dev_basis[1]

    def to_fom_vector(self, uN):
        """``u_h = V u_N``  (reference rom/rom.py:97-113)."""
        from .. import kernels

        col = np.ascontiguousarray(np.asarray(uN, dtype=np.float64).reshape(-1, 1))
        return kernels.pod_lift(self._basis_device(), col, order="C").reshape(-1).cpu().numpy()

    def to_rom_vector(self, uh):
        """``u_N = V^T u_h``  (reference rom/rom.py:115-133).  ``uh``: FE function or plain array."""
        from .. import kernels

        vec = uh if isinstance(uh, np.ndarray) else function_to_array(uh)
        return kernels.project_vec(self._basis_device(), np.asarray(vec, dtype=np.float64).reshape(-1, 1)).reshape(-1).cpu().numpy()

    def to_rom(self, oph):
        """Galerkin projection of a FOM operator: ``V^T A V`` for matrices, ``V^T f`` for vectors
        (reference rom/rom.py:135-158)."""
        if _is_sparse_like(oph):
            return project_csr(bilinear_to_csr(oph), self._basis_device())
        return self.to_rom_vector(functional_to_array(oph))

    # ------------------------------------------------------------------ basis management
    def load_from_basis(self, basis, mu_space):
        self.basis = deepcopy(basis)
        mu_space[Stage.ONLINE] = []
        mu_space[Stage.VALIDATION] = []
        self.mu_space = deepcopy(mu_space)

    def truncate(self, n):
        """ROM with the last ``n`` basis vectors removed (reference rom/rom.py:170-203)."""
        truncated = self.__class__(fom=self.fom, grid=self.grid, name=self.name)
        truncated.setup(rnd=self.random_state)
        N = self.N
        assert n < N, "You want to remove too many modes from S-ROM to create ROM."
        truncated.basis = self.basis[:, : N - n]
        truncated.mu_space = deepcopy(self.mu_space)
        truncated.report = deepcopy(self.report)
        truncated.report[Stage.OFFLINE][Treewalk.BASIS_FINAL] = truncated.N
        return truncated

    def setup(self, rnd):
        super().setup(rnd=rnd)
        self.algebraic_solver = self.create_algebraic_solver()

    def create_algebraic_solver(self):
        """Single-system solver with the reference's call form ``solver(A=K, b=b) -> (x, info)``
        (rom/rom.py:414-425).  Dense LU on the device through rt_online_step_f64 is used inside ``solve``;
        this stand-alone hook solves one system the same way (identity mass trick: K u = b)."""
        def solver(A, b):
            return _dense_solve(np.asarray(A, dtype=np.float64), np.asarray(b, dtype=np.float64)), 0

        return solver

    _ATTR_OF = {
        OperatorType.FORCING: "deim_fh", OperatorType.LIFTING: "deim_fgh", OperatorType.RHS: "deim_rhs",
        OperatorType.MASS: "mdeim_Mh", OperatorType.STIFFNESS: "mdeim_Ah", OperatorType.CONVECTION: "mdeim_Ch",
        OperatorType.TRILINEAR: "mdeim_Nh", OperatorType.NONLINEAR_LIFTING: "mdeim_Nh_hat",
    }

    def add_hyper_reductor(self, reductor, which):
        """Attach a COPY of ``reductor`` for operator ``which`` (reference rom/rom.py:216-256)."""
        attr = self._ATTR_OF.get(which)
        if attr is None:
            raise NotImplementedError(f"Which is this reductor? {which}")
        setattr(self, attr, reductor.copy())
        self._dev_model = None

    def _reductors(self):
        return [getattr(self, a) for a in ("deim_fh", "deim_fgh", "deim_rhs", "mdeim_Mh", "mdeim_Ah", "mdeim_Ch",
                                           "mdeim_Nh", "mdeim_Nh_hat")]

    def project_reductors(self):
        """Project every attached collateral basis onto the solution basis (reference rom/rom.py:258-280)."""
        for red in self._reductors():
            if red:
                red.project_basis(V=self.basis)
        self._dev_model = None

    # ------------------------------------------------------------------ offline: solution reduced basis
    def build_reduced_basis(self, num_snapshots=None, mu_space=None, num_basis=None, tolerances=dict()):
        """Two-level POD of FOM solutions: POD in time per parameter (``tol_time``), stack, POD in mu
        (``num_basis`` / ``tol_mu``, not normalised); the same for the nonlinear-term snapshots
        (reference rom/rom.py:282-412).  Returns the FOM solutions per parameter index."""
        if num_snapshots:
            space = self.build_sampling_space(num=num_snapshots, rnd=self.random_state)
        elif mu_space:
            space = mu_space
        else:
            raise NotImplementedError("You need to provide a number of mu-snapshots or a space.")
        print("Building Reduced Basis ... ")
        pprint(space)
        fom = self.fom
        if fom.is_setup == False:  # noqa: E712
            fom.setup()
        off = self.report[Stage.OFFLINE]
        fom_solutions = dict()
        tol_t = tolerances.get(RomParameters.TOL_TIME, None)
        # the FOM solves stay on the host, parameter after parameter; the per-parameter time-PODs of the solution and of the
        # nonlinear term (rom/rom.py:335, :349-351) are independent and go to the device as two batches after the loop
        ids, snaps, snaps_nl_all = [], [], []
        for mu in space:
            mu_idx, mu = self.add_mu(mu=mu, step=Stage.OFFLINE)
            fom.setup()
            fom.update_parametrization(mu)
            fom.solve()
            fom_solutions[mu_idx] = fom.solutions.fom.copy()
            ids.append(mu_idx)
            snaps.append(np.array(fom.solutions.snapshots, dtype=float, copy=True))
            # nonlinear term: drop the zero initial snapshot, boundary row does not matter
            snaps_nl = np.array(fom.nonlinear_snapshots[1:]).T
            snaps_nl[0, :] = 0.0
            snaps_nl_all.append(snaps_nl)
            if getattr(fom, "RUNTIME_PROCESS", False):
                fom.save_probes(name=f"probes_offline_fom_{mu_idx}.csv")
        blocks, blocks_nl = [], []
        for mu_idx, (basis_t, sig_t, en_t), (basis_nl, sig_nl, en_nl) in zip(ids, orth_batch(snaps, tol=tol_t),
                                                                             orth_batch(snaps_nl_all, tol=tol_t)):
            blocks.append(basis_t)
            off[Treewalk.SPECTRUM_TIME][mu_idx] = sig_t
            off[Treewalk.ENERGY_TIME][mu_idx] = en_t
            off[Treewalk.BASIS_TIME][mu_idx] = basis_t.shape[1]
            blocks_nl.append(basis_nl)
            off[TreewalkNonlinear.SPECTRUM_TIME][mu_idx] = sig_nl
            off[TreewalkNonlinear.ENERGY_TIME][mu_idx] = en_nl
            off[TreewalkNonlinear.BASIS_TIME][mu_idx] = basis_t.shape[1]  # sic (reference rom/rom.py:363-365)
        stacked, stacked_nl = np.hstack(blocks), np.hstack(blocks_nl)
        off[Treewalk.BASIS_AFTER_WALK] = stacked.shape[1]
        off[TreewalkNonlinear.BASIS_AFTER_WALK] = stacked_nl.shape[1]
        tol_mu = tolerances.get(RomParameters.TOL_MU, None)
        basis, sig_mu, en_mu = orth(stacked, num=num_basis, tol=tol_mu, normalize=False)
        off[Treewalk.SPECTRUM_MU] = sig_mu
        off[Treewalk.ENERGY_MU] = en_mu
        off[Treewalk.BASIS_FINAL] = basis.shape[1]
        self.basis = basis
        basis_nl, sig_mu_nl, en_mu_nl = orth(stacked_nl, normalize=False)
        off[TreewalkNonlinear.SPECTRUM_MU] = sig_mu_nl
        off[TreewalkNonlinear.ENERGY_MU] = en_mu_nl
        off[TreewalkNonlinear.BASIS_FINAL] = basis_nl.shape[1]
        self.basis_nonlinear = basis_nl
        assert self.N != 0, "(ROM) There are no basis vectors: check tol_mu against the mu-energy."
        return fom_solutions

    # ------------------------------------------------------------------ single-operator assembly (reference API)
    def _assemble(self, reductor, fom_method, mu, t, **extra):
        if reductor:
            return reductor.interpolate(mu=mu, t=t, which=self.ROM, **extra)
        return self.to_rom(getattr(self.fom, fom_method)(mu, t, **extra))

    def assemble_mass(self, mu, t):
        return self._assemble(self.mdeim_Mh, "assemble_mass", mu, t)

    def assemble_stiffness(self, mu, t):
        return self._assemble(self.mdeim_Ah, "assemble_stiffness", mu, t)

    def assemble_forcing(self, mu, t):
        return self._assemble(self.deim_fh, "assemble_forcing", mu, t)

    def assemble_lifting(self, mu, t):
        return self._assemble(self.deim_fgh, "assemble_lifting", mu, t)

    def assemble_rhs(self, mu, t):
        if self.deim_rhs:
            return self.deim_rhs.interpolate(mu=mu, t=t, which=self.ROM)
        return self.assemble_forcing(mu, t) + self.assemble_lifting(mu, t)

    def assemble_system(self, mu, t, bdf=1.0, uh=None, uh_n1=None):
        """(M_N, K_N) with ``K = bdf M + dt A`` (reference rom/rom.py:565-573; the reference's base
        class drops ``bdf`` and is not callable from its own ``solve`` at HEAD -- see DESIGN.md)."""
        MN = self.assemble_mass(mu=mu, t=t)
        KN = bdf * MN + self.fom.dt * self._stiffness_like(mu, t, uh, uh_n1)
        return MN, KN

    def _stiffness_like(self, mu, t, uh, uh_n1):
        return self.assemble_stiffness(mu=mu, t=t)

    def assemble_system_rhs(self, mu, t, MN_mat, uN_n, uN_n1=None):
        """``b = M w + dt f`` with w = u_n (BDF1) or 2 u_n - u_nm1 / 2 (BDF2)  (reference
        rom/rom.py:911-929, :557-563)."""
        fN = self._rhs_vector(mu, t)
        w = uN_n if uN_n1 is None else 2.0 * uN_n - 0.5 * uN_n1
        return MN_mat.dot(w) + self.fom.dt * fN

    def _rhs_vector(self, mu, t):
        return self.assemble_rhs(mu=mu, t=t)

    # ------------------------------------------------------------------ device model of the attached reductors
    def _slot_reductors(self):
        """slot name -> reductor (or list of vector reductors summed in the rhs) for THIS class."""
        rhs = [self.deim_rhs] if self.deim_rhs else [r for r in (self.deim_fh, self.deim_fgh) if r]
        return {"mass": self.mdeim_Mh, "stiffness": self.mdeim_Ah, "convection": None, "nonlinear_lifting": None,
                "trilinear": None, "rhs": rhs}

    def _fallback_methods(self):
        """FOM assembly methods for slots WITHOUT a reductor: slot -> list of fom method names."""
        rhs = [] if self.deim_rhs else [m for m, r in (("assemble_forcing", self.deim_fh), ("assemble_lifting", self.deim_fgh)) if not r]
        return {"mass": ["assemble_mass"], "stiffness": ["assemble_stiffness"], "rhs": rhs}

    def _device_model(self, evaluator=None):
        from .. import kernels

        slots = self._slot_reductors()
        # the cache entry KEEPS the objects it was built from and compares by identity (ids of freed objects are reused)
        key = [getattr(r, "basis_rom", None) for v in slots.values() for r in (v if isinstance(v, list) else [v])]
        key += [getattr(r, "PT_U", None) for v in slots.values() for r in (v if isinstance(v, list) else [v])]
        key += [evaluator, self.basis]
        hit = self._dev_model
        if hit is not None and len(hit[0]) == len(key) and all(a is b for a, b in zip(hit[0], key)):
            return hit[1]
        model = {"N": self.N, "which": evaluator.which if evaluator else "rest",
                 "evaluator_id": getattr(evaluator, "evaluator_id", None)}
        for name in _SLOTS:
            red = slots[name]
            reds = [r for r in (red if isinstance(red, list) else [red]) if r]
            if not reds:
                model[name] = None
                continue
            for r in reds:
                if r.basis_rom is None:
                    raise ValueError(f"{r}: project_reductors() must run before solving")
            if len(reds) == 1:
                entry = {"basis_rom": reds[0].basis_rom, "PT_U": reds[0].PT_U}
            else:  # several vector reductors summed: side-by-side bases, block-diagonal interpolation matrix
                ms = [r.PT_U.shape[0] for r in reds]
                P = np.zeros((sum(ms), sum(ms)))
                o = 0
                for r, mm in zip(reds, ms):
                    P[o:o + mm, o:o + mm] = r.PT_U
                    o += mm
                entry = {"basis_rom": np.hstack([r.basis_rom for r in reds]), "PT_U": P}
            if evaluator is not None:
                tab = evaluator.tables[name]
                entry["C"], entry["T"] = tab["C"], tab.get("T")
            model[name] = entry
        if all(model[name] is None for name in _SLOTS):
            # pure Galerkin ROM (no hyper-reductor attached, the reference's default RomConstructor path through
            # to_rom(fom.assemble_*), rom/rom.py:575-616): every operator arrives as a dense additive term of
            # rt_online_step_f64; a one-mode zero mass operator keeps the device model non-empty
            N = self.N
            model["mass"] = {"basis_rom": np.zeros((N * N, 1)), "PT_U": np.eye(1)}
        dm = kernels.DeviceRomModel(model)
        self._dev_model = (key, dm)
        return dm

    def _sampled_row(self, mu, t, u_star):
        """Sampled entries of every attached reductor, in slot order (one row of rt_online_step_f64's F)."""
        parts = []
        for name, red in self._slot_reductors().items():
            for r in (red if isinstance(red, list) else [red]):
                if not r:
                    continue
                extra = {"u_n": u_star} if name == "trilinear" else {}
                parts.append(r._local_entries(mu, t, **extra))
        if not parts:
            return np.zeros(1)  # the dummy mass mode of a reductor-free model (_device_model)
        return np.concatenate(parts)

    def _dense_terms(self, mu, t, bdf, u_star):
        """(Kadd, Madd, fadd) for operators without a reductor (FOM assembly + projection), or Nones."""
        dt = self.fom.dt
        N = self.N
        Kadd = Madd = fadd = None
        slots = self._slot_reductors()
        for name, methods in self._fallback_methods().items():
            red = slots[name]
            present = bool(red) if not isinstance(red, list) else False
            if present:
                continue
            for method in methods:
                extra = {"u_n": u_star} if name == "trilinear" else {}
                op = self.to_rom(getattr(self.fom, method)(mu, t, **extra))
                if name == "rhs":
                    fadd = (0.0 if fadd is None else fadd) + dt * op.reshape(1, N)
                elif name == "mass":
                    Madd = (0.0 if Madd is None else Madd) + op.reshape(1, N, N)
                    Kadd = (0.0 if Kadd is None else Kadd) + bdf * op.reshape(1, N, N)
                else:
                    Kadd = (0.0 if Kadd is None else Kadd) + dt * op.reshape(1, N, N)
        return Kadd, Madd, fadd

    # ------------------------------------------------------------------ online: one parameter (reference API)
    def solve(self, mu, step):
        """Solve the ROM for one parameter over ``fom.domain["nt"]`` steps (reference rom/rom.py:430-555).
        Leaves a ``RomSolutionsStorage`` in ``self.solutions`` and returns the index of ``mu``."""
        import torch

        from .. import kernels

        idx_mu, mu = self.add_mu(mu=mu, step=step)
        fom = self.fom
        has_exact = fom.exact_solution is not None
        errors, exact = [], dict()
        dt = fom.dt
        bdf_scheme = 2 if fom.BDF_SCHEME == BDF.TWO else 1
        N = self.N
        dm = self._device_model()
        dev = self._basis_device().device
        Un = torch.zeros((1, N), dtype=torch.float64, device=dev)     # zero initial condition (:451-452)
        Unm1 = torch.zeros((1, N), dtype=torch.float64, device=dev)
        info = torch.zeros(1, dtype=torch.int32, device=dev)
        uh = self.to_fom_vector(np.zeros(N))
        uh_n1 = None
        t = 0.0
        timesteps, fom_coeffs, rom_coeffs, domains = [], [], [], []
        for n in range(fom.domain["nt"]):
            t += dt
            timesteps.append(t)
            bdf = 1.5 if (bdf_scheme == 2 and n > 0) else 1.0
            u_star = uh if uh_n1 is None else 2.0 * uh - uh_n1
            F = self._sampled_row(mu, t, u_star)
            Kadd, Madd, fadd = self._dense_terms(mu, t, bdf, u_star)
            kernels.online_step(dm, F.reshape(1, -1), Un, Unm1, info, dt, bdf_scheme, n, Kadd=Kadd, Madd=Madd, fadd=fadd)
            uN = Un[0].cpu().numpy()
            rom_coeffs.append(uN)
            if bdf_scheme == 2:
                uh_n1 = uh
            uh = self.to_fom_vector(uN)
            # FEM side: moving mesh, lifting, full solution (reference rom/rom.py:505-522)
            fom.move_mesh(mu=mu, t=t)
            domains.append(fom.x.copy())
            g, _, _ = fom.create_lifting_operator(mu=mu, t=t, L=fom.L)
            fom.move_mesh(back=True)
            gh = function_to_array(fom.interpolate_func(g, fom.V, mu, t))
            uc_h = uh + gh
            fom_coeffs.append(uc_h.copy())
            if has_exact:
                import fenics

                ue = fenics.Expression(fom.exact_solution, degree=1, t=t, **mu)
                ue_h = function_to_array(fom.interpolate_func(ue, fom.V, mu, t))
                exact[t] = ue_h.copy()
                errors.append(self._compute_error(u=uc_h, ue=ue_h))
        if int(info.item()) != 0:
            raise np.linalg.LinAlgError(f"Singular matrix: reduced system singular at time step {int(info.item()) - 1}")
        self.solutions = RomSolutionsStorage(ts=timesteps, mu=mu, domain=np.hstack(domains), fom=np.vstack(fom_coeffs).T,
                                             rom=np.vstack(rom_coeffs).T)
        if has_exact:
            self.errors.update({idx_mu: np.array(errors)})
            self.exact.update({idx_mu: exact})
        return idx_mu

    # ------------------------------------------------------------------ online: a batch of parameters
    def solve_batch(self, mus, evaluator, nt=None, dt=None, bdf=None, history=True, step=Stage.ONLINE, to_host=False):
        """Sweep B parameters at once on the device.  ``mus``: (B, n_mu) array in the evaluator's parameter order
        ([a0, omega, delta] for the piston, [delta, beta, alpha_0, omega] for MFP1) or a list of mu dictionaries;
        ``evaluator``: a ``PistonEvaluator`` / ``Mfp1Evaluator`` with the closed-form sampled-entry tables.
        Under ``torch.distributed`` every rank passes ITS shard of ``mus`` (see distributed.shard_range);
        there is no collective.  ``to_host=True`` delivers the history in pinned host memory, the copy of one chunk of
        parameters overlapping the sweep of the next (kernels.online_sweep_to_host).  Returns a ``RomBatchSolutions``."""
        from .. import kernels

        fom = self.fom
        nt = int(nt if nt is not None else fom.domain["nt"])
        dt = float(dt if dt is not None else fom.dt)
        if bdf is None:
            bdf = 2 if fom.BDF_SCHEME == BDF.TWO else 1
        names = evaluator.mu_names
        if isinstance(mus, (list, tuple)) and mus and isinstance(mus[0], dict):
            arr = np.array([[m.get(k, 0.0) for k in names] for m in mus], dtype=np.float64)
            mu_list = list(mus)
        else:
            arr = np.asarray(mus, dtype=np.float64).reshape(-1, len(names))
            mu_list = [dict(zip(names, r)) for r in arr]
        self.mu_space[step].extend(mu_list)  # one operation: add_mu's list.index per parameter is O(B^2)
        if mu_list:
            self.mu = mu_list[-1]
        dm = self._device_model(evaluator)
        if to_host and history:
            out, info = kernels.online_sweep_to_host(dm, arr, nt, dt, bdf=bdf)
        else:
            out, info = kernels.online_sweep(dm, arr, nt, dt, bdf=bdf, history=history)
        ts = dt * np.arange(1, nt + 1)
        if not history:
            out = out.unsqueeze(1)
            ts = ts[-1:]
        return RomBatchSolutions(ts=ts, mus=mu_list, rom_device=out, info=info, basis=self.basis)


class RomConstructorMoving(RomConstructor):
    """ROM on a moving (ALE) domain: adds the mesh-velocity convection operator
    (reference rom/rom.py:688-736)."""

    def assemble_convection(self, mu, t):
        return self._assemble(self.mdeim_Ch, "assemble_convection", mu, t)

    def _stiffness_like(self, mu, t, uh, uh_n1):
        return self.assemble_stiffness(mu=mu, t=t) + self.assemble_convection(mu=mu, t=t)

    def _slot_reductors(self):
        slots = super()._slot_reductors()
        slots["convection"] = self.mdeim_Ch
        return slots

    def _fallback_methods(self):
        methods = super()._fallback_methods()
        methods["convection"] = ["assemble_convection"]
        return methods


class RomConstructorNonlinear(RomConstructorMoving):
    """Burgers / piston ROM: trilinear term linearised around u* (extrapolated state) and the nonlinear
    lifting operator; no forcing, the right-hand side is the lifting only (reference rom/rom.py:739-974)."""

    PISTON_MACH_MIN = 0.15
    PISTON_MACH_MAX = 0.4

    def __init__(self, fom, grid, name=None):
        super().__init__(fom=fom, grid=grid, name=name)
        self.probe_location = fom.probe_location
        self.probes = None

    # ---- ad-hoc sampling of the parameter space by piston Mach number (reference rom/rom.py:752-842)
    @staticmethod
    def compute_piston_mach_number(sample):
        P = PistonParameters
        return sample[P.DELTA] * sample[P.OMEGA] / sample[P.A0]

    @staticmethod
    def compute_piston_mach_number_space(grid, num, mach_min=None, mach_max=None):
        P = PistonParameters
        lim = {}
        for var in (P.A0, P.OMEGA, P.DELTA):
            support = grid[var].support()
            lim[var] = (min(support), max(support))
        if mach_min is None:
            mach_min = lim[P.DELTA][0] * lim[P.OMEGA][0] / lim[P.A0][1]
        if mach_max is None:
            mach_max = lim[P.DELTA][1] * lim[P.OMEGA][1] / lim[P.A0][0]
        return np.linspace(start=mach_min, stop=mach_max, num=num + 1)

    def build_sampling_space(self, num, rnd=None):
        """One parameter per piston-Mach bin, sorted by Mach number (reference rom/rom.py:752-809)."""
        edges = self.compute_piston_mach_number_space(grid=self.grid, num=num, mach_min=self.PISTON_MACH_MIN,
                                                      mach_max=self.PISTON_MACH_MAX)
        bins = list(zip(edges, edges[1:]))
        samples = []
        for sample in super().build_sampling_space(rnd=rnd, num=int(2e4)):
            mach = self.compute_piston_mach_number(sample)
            hit = next((b for b in bins if b[0] <= mach <= b[1]), None)
            if hit is not None:
                sample[PistonParameters.MACH_PISTON] = mach
                samples.append(sample)
                bins.remove(hit)
            if not bins:
                break
        return sorted(samples, key=lambda s: s[PistonParameters.MACH_PISTON])

    def assemble_trilinear(self, mu, t, uh):
        if self.mdeim_Nh:
            return self.mdeim_Nh.interpolate(mu=mu, t=t, u_n=uh, which=self.ROM)
        return self.to_rom(self.fom.assemble_trilinear(mu=mu, t=t, u_n=uh))

    def assemble_nonlinear_lifting(self, mu, t):
        return self._assemble(self.mdeim_Nh_hat, "assemble_nonlinear_lifting", mu, t)

    def _stiffness_like(self, mu, t, uh, uh_n1):
        u_star = uh if uh_n1 is None else 2.0 * uh - uh_n1  # reference rom/rom.py:896-901
        return (self.assemble_stiffness(mu=mu, t=t) + self.assemble_convection(mu=mu, t=t)
                + self.assemble_trilinear(mu=mu, t=t, uh=u_star) + self.assemble_nonlinear_lifting(mu=mu, t=t))

    def _rhs_vector(self, mu, t):
        return self.assemble_lifting(mu=mu, t=t)  # no forcing for Burgers (reference rom/rom.py:913-915)

    def _slot_reductors(self):
        slots = super()._slot_reductors()
        slots["nonlinear_lifting"] = self.mdeim_Nh_hat
        slots["trilinear"] = self.mdeim_Nh
        slots["rhs"] = [self.deim_fgh] if self.deim_fgh else []
        return slots

    def _fallback_methods(self):
        methods = super()._fallback_methods()
        methods["nonlinear_lifting"] = ["assemble_nonlinear_lifting"]
        methods["trilinear"] = ["assemble_trilinear"]
        methods["rhs"] = [] if self.deim_fgh else ["assemble_lifting"]
        return methods


def _dense_solve(K, b):
    """One dense system K u = b on the device: rt_online_step_f64 with no reductors would be empty, so the
    system is passed as the dense additive terms (K via Kadd, b via fadd) of a one-mode dummy model."""
    import torch

    from .. import kernels

    N = K.shape[0]
    model = {"N": N, "mass": {"basis_rom": np.zeros((N * N, 1)), "PT_U": np.eye(1)}}
    dm = kernels.DeviceRomModel(model)
    Un = torch.zeros((1, N), dtype=torch.float64, device="cuda")
    Unm1 = torch.zeros_like(Un)
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    kernels.online_step(dm, np.zeros((1, 1)), Un, Unm1, info, 1.0, 1, 0, Kadd=K.reshape(1, N, N), fadd=b.reshape(1, N))
    if int(info.item()) != 0:
        raise np.linalg.LinAlgError("Singular matrix")
    return Un[0].cpu().numpy()
