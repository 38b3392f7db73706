"""TEST INFRASTRUCTURE -- FEniCS-free stand-ins for the ``assemble`` callbacks and the FOM object the
reference's offline stage consumes.  The SAME objects are fed to the unmodified reference
(oracle/gen_golden.py, in the build container) and to romtime_b200 (tests/, on the GPU box), so the
golden files compare like with like.  Nothing under romtime_b200/ imports this module.

Shapes follow the reference's 1-D P1 problems: vectors of Nh dofs, tridiagonal matrices whose non-zeros
are vectorised in CSR order, parameters as dictionaries, non-affine dependence on (mu, t)
(reference problems/mfp1.py:27-75, fom/base.py:17-43 for the callback contract)."""
import numpy as np
import scipy.sparse as sp


class PetscLikeMatrix:
    """scipy CSR matrix behind the two PETSc calls of the reference's utils.bilinear_to_csr
    (utils.py:90-91): ``as_backend_type(m).mat().getValuesCSR()`` and ``.size``."""

    def __init__(self, csr):
        self._csr = csr.tocsr()
        self.size = self._csr.shape

    def mat(self):
        return self

    def getValuesCSR(self):
        return self._csr.indptr, self._csr.indices, self._csr.data


def mu_space(num, seed):
    rng = np.random.default_rng(seed)
    return [{"a": float(0.5 + rng.random()), "w": float(2.0 + 4.0 * rng.random())} for _ in range(num)]


class VectorProblem:
    """Non-affine forcing-like vector: a travelling Gaussian bump plus a polynomial part."""

    def __init__(self, Nh=160):
        self.Nh = Nh
        self.x = np.linspace(0.0, 1.0, Nh)

    def full(self, mu, t):
        x, a, w = self.x, mu["a"], mu["w"]
        centre = 0.5 + 0.3 * np.sin(w * t)
        return np.exp(-((x - centre) ** 2) / (0.02 + 0.05 * a)) * (1.0 + a * x * t) + 0.1 * a * x * (1.0 - x) * np.cos(w * t)

    def assemble(self, mu, t, entries=None):
        f = self.full(mu, t)
        if entries is None:
            return f
        return np.array([f[e[0]] for e in entries])


class MatrixProblem:
    """Tridiagonal stiffness-plus-convection-like operator on a scaled mesh with a non-affine diffusivity;
    all pattern entries stay away from zero so the sparsity pattern is parameter independent."""

    def __init__(self, Nh=90):
        self.Nh = Nh
        self.x = np.linspace(0.0, 1.0, Nh)

    def _bands(self, mu, t, u_n=None):
        x, a, w = self.x, mu["a"], mu["w"]
        L = 1.0 - 0.2 * (1.0 - np.cos(w * t))
        kappa = 1.0 + 0.5 * a * np.exp(-((x - 0.4 - 0.2 * np.sin(w * t)) ** 2) / 0.05)
        diag = 2.0 * kappa / L + 0.3 * L
        lower = -kappa[1:] / L - 0.1 * a * x[1:] * np.sin(w * t) - 0.05
        upper = -kappa[:-1] / L + 0.1 * a * x[:-1] * np.sin(w * t) - 0.05
        if u_n is not None:  # state-dependent (trilinear-like) part, linear in u_n
            diag = diag + 0.5 * u_n
            lower = lower - 0.25 * u_n[1:]
            upper = upper + 0.25 * u_n[:-1] - 0.1 * u_n[1:]
        diag[0] = 1.0  # Dirichlet row like the reference's boundary handling
        upper[0] = 0.0
        return lower, diag, upper

    def matrix(self, mu, t, u_n=None):
        lower, diag, upper = self._bands(mu, t, u_n)
        A = sp.diags([lower, diag, upper], offsets=[-1, 0, 1], format="csr")
        return A

    def assemble(self, mu, t, entries=None, u_n=None):
        A = self.matrix(mu, t, u_n)
        if entries is None:
            return PetscLikeMatrix(A)
        return np.array([A[int(r), int(c)] for (r, c) in entries])


class NonlinearMatrixProblem(MatrixProblem):
    """Same operator with a mandatory state argument (N-MDEIM callback, deim/nonlinear.py:268)."""

    def assemble(self, mu, t, u_n=None, entries=None):
        u = np.asarray(u_n, dtype=float)
        return super().assemble(mu, t, entries=entries, u_n=u)

    def state_basis(self, k=3):
        x = self.x
        return np.stack([0.3 + 0.2 * np.sin((j + 1) * np.pi * x) for j in range(k)], axis=1)


class _Solutions:
    def __init__(self, snapshots):
        self.snapshots = snapshots
        self.fom = snapshots


class _ArrayFunction:
    def __init__(self, arr):
        self._arr = np.asarray(arr, dtype=float)

    def vector(self):
        return self

    def vec(self):
        return self

    @property
    def array(self):
        return self._arr


class SnapshotFom:
    """Duck-typed FOM for RomConstructor.build_reduced_basis (reference rom/rom.py:313-345) and
    RomConstructor*.solve (:441-515): ``solve()`` fills ``solutions.snapshots`` (Nh x nt) and
    ``nonlinear_snapshots`` (nt + 1 vectors, first one zero) from closed forms."""

    RUNTIME_PROCESS = False

    def __init__(self, Nh=200, nt=30, dt=0.02, bdf_scheme="2"):
        self.Nh, self.nt = Nh, nt
        self.x_ref = np.linspace(0.0, 1.0, Nh)
        self.x = self.x_ref.reshape(Nh, 1)
        self.is_setup = False
        self.exact_solution = None
        self.dt = dt
        self.BDF_SCHEME = bdf_scheme
        self.domain = {"nt": nt}
        self.probe_location = [0.0, 0.5]
        self.L = 1.0
        self.V = None
        self.mu = None

    def setup(self):
        self.is_setup = True

    def update_parametrization(self, mu):
        self.mu = mu

    def field(self, mu, t):
        x, a, w = self.x_ref, mu["a"], mu["w"]
        return np.sin(np.pi * x) * np.exp(-a * t) * np.cos(w * t) + 0.3 * a * np.sin(2 * np.pi * x * (1 + 0.2 * t)) * t \
            + 0.05 * np.sin(3 * np.pi * x) * np.sin(w * t) ** 2

    def solve(self):
        ts = self.dt * np.arange(1, self.nt + 1)
        snaps = np.stack([self.field(self.mu, t) for t in ts], axis=1)
        self.solutions = _Solutions(snaps)
        nl = [np.zeros(self.Nh)] + [0.5 * snaps[:, k] ** 2 + 0.1 * np.gradient(snaps[:, k]) for k in range(self.nt)]
        self.nonlinear_snapshots = nl

    # online protocol
    def move_mesh(self, mu=None, t=None, back=False):
        pass

    def create_lifting_operator(self, mu, t, L):
        return (mu, t), None, None

    def interpolate_func(self, g, V, mu, t):
        mu_, t_ = g
        return _ArrayFunction(0.1 * mu_["a"] * self.x_ref * np.sin(mu_["w"] * t_))
