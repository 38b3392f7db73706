"""TEST INFRASTRUCTURE -- a FEniCS-free full-order model for the two named problems, built on the closed-form P1 assembly
of oracle/p1_forms.py, with the interface the ROM classes consume (reference fom/base.py: ``setup``,
``update_parametrization``, ``solve``, ``solutions``, ``assemble_*(mu, t, entries=None[, u_n])``, ``move_mesh``,
``create_lifting_operator``, ``interpolate_func``).  It lets the whole pipeline -- FOM snapshots, reduced basis, (M)DEIM
offline, projection, online solve through host callbacks AND through the device evaluator -- run end to end without FEniCS.
Never imported by romtime_b200/."""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

import p1_forms
from synthetic_problems import PetscLikeMatrix, _ArrayFunction, _Solutions


class P1Fom:
    RUNTIME_PROCESS = False

    def __init__(self, problem, tf, nt, bdf_scheme="2"):
        self.problem, self.mesh = problem, problem.mesh
        self.dt = tf / nt
        self.domain = {"nt": nt, "T": tf, "L0": problem.mesh.L0, "nx": problem.mesh.nx}
        self.BDF_SCHEME = bdf_scheme
        self.exact_solution = None
        self.probe_location = [0.0, 0.5]
        self.V = None
        self.mu = None
        self.is_setup = False
        self._scale = 1.0
        self.nonlinear = isinstance(problem, p1_forms.Piston)
        self.matrix_ops = ("mass", "stiffness", "convection") + (("nonlinear_lifting",) if self.nonlinear else ())

    # ---- FEM structures
    @property
    def x(self):
        return (self._scale * self.mesh.x0).reshape(-1, 1)

    @property
    def L(self):
        return float(np.max(self.x))

    def setup(self):
        self.is_setup = True

    def update_parametrization(self, mu):
        self.mu = dict(mu)

    def move_mesh(self, mu=None, t=None, back=False):
        self._scale = 1.0 if back else float(self.problem.scale(mu, t))

    def create_lifting_operator(self, mu, t, L, only_g=False):
        g = ("g", dict(mu), float(t))
        return g if only_g else (g, None, None)

    def lifting_values(self, mu, t):
        """g at the nodes of the moved mesh (fom/base.py:440-465 / fom/nonlinear.py:119-150)."""
        s = self.problem.scale(mu, t)
        x, L = s * self.mesh.x0, s * self.mesh.L0
        if self.nonlinear:
            bL, _ = self.problem._boundary(mu, t)
            return bL * x / L
        b0, bL, _, _, _ = self.problem._boundary(mu, t, L)
        return bL * x / L + b0 * (L - x) / L

    def interpolate_func(self, g, V, mu, t):
        _, mu_, t_ = g
        return _ArrayFunction(self.lifting_values(mu_, t_))

    # ---- assembly callbacks (the contract of fom/base.py:17-43, :651-665)
    def _matrix(self, op, mu, t, entries=None, u_n=None):
        if entries is not None:
            return self.problem.assemble(op, mu, t, [tuple(int(v) for v in e) for e in entries], u_n=u_n)
        return PetscLikeMatrix(sp.csr_matrix(self.problem.full_matrix(op, mu, t, u_n=u_n)))

    def _vector(self, op, mu, t, entries=None):
        if entries is not None:
            return self.problem.assemble(op, mu, t, [tuple(int(v) for v in e) for e in entries])
        return self.problem.full_vector(op, mu, t)

    def assemble_mass(self, mu, t, entries=None):
        return self._matrix("mass", mu, t, entries)

    def assemble_stiffness(self, mu, t, entries=None):
        return self._matrix("stiffness", mu, t, entries)

    def assemble_convection(self, mu, t, entries=None):
        return self._matrix("convection", mu, t, entries)

    def assemble_nonlinear_lifting(self, mu, t, entries=None):
        return self._matrix("nonlinear_lifting", mu, t, entries)

    def assemble_trilinear(self, mu, t, entries=None, u_n=None):
        return self._matrix("trilinear", mu, t, entries, u_n=np.asarray(u_n, dtype=float))

    def assemble_forcing(self, mu, t, entries=None):
        return self._vector("forcing", mu, t, entries)

    def assemble_lifting(self, mu, t, entries=None):
        return self._vector("lifting", mu, t, entries)

    def assemble_rhs(self, mu, t, entries=None):
        return self._vector("rhs", mu, t, entries)

    # ---- time integration of the homogeneous problem (fom/base.py solve loop; fom/heat.py:57-78, fom/nonlinear.py:322-372)
    def solve(self):
        mu, dt, nt, Nh = self.mu, self.dt, self.domain["nt"], self.mesh.Nh
        bdf2 = self.BDF_SCHEME == "2"
        free = np.array([i for i in range(Nh) if i not in self.mesh.dirichlet_dofs])
        u, um1 = np.zeros(Nh), None
        snaps, full, nl = [], [], []   # the first trilinear snapshot (u* = 0) is the zero one the ROM drops (rom/rom.py:346)
        t = 0.0
        for n in range(nt):
            t += dt
            bdf = 1.5 if (bdf2 and n > 0) else 1.0
            M = self.problem.full_matrix("mass", mu, t)
            K = bdf * M
            for op in self.matrix_ops[1:]:
                K = K + dt * self.problem.full_matrix(op, mu, t)
            if self.nonlinear:
                ustar = u if um1 is None else 2.0 * u - um1
                Ct = self.problem.full_matrix("trilinear", mu, t, u_n=ustar)
                K = K + dt * Ct
                nl.append(Ct[self._pattern_mask()].copy())
            w = u if (um1 is None or not bdf2) else 2.0 * u - 0.5 * um1
            b = M @ w + dt * self.problem.full_vector("rhs", mu, t)
            b[list(self.mesh.dirichlet_dofs)] = 0.0
            new = np.zeros(Nh)
            new[free] = spla.spsolve(sp.csc_matrix(K[np.ix_(free, free)]), b[free])
            um1, u = (u, new) if bdf2 else (None, new)
            snaps.append(new.copy())
            full.append(new + self.lifting_values(mu, t))
        self.solutions = _Solutions(np.stack(snaps, axis=1))
        self.solutions.fom = np.stack(full, axis=1)
        # the linear problem has no nonlinear term; build_reduced_basis still runs its (unused) POD on this list (rom/rom.py:349-351)
        self.nonlinear_snapshots = nl if self.nonlinear else [np.zeros(Nh)] + [s_.copy() for s_ in snaps]

    def _pattern_mask(self):
        m = np.zeros((self.mesh.Nh, self.mesh.Nh), dtype=bool)
        for i, j in self.mesh.pattern():
            m[i, j] = True
        return m

    def exact(self, mu, t):
        """MFP1's manufactured solution at the nodes of the moved mesh (problems/mfp1.py:44)."""
        x = self.problem.scale(mu, t) * self.mesh.x0
        return (1.0 - np.exp(-mu["beta"] * t)) * (1.0 + mu["delta"] ** 2 * x * x)
