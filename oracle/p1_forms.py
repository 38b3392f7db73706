"""TEST INFRASTRUCTURE -- CPU restatement of the reference's sampled-entry assembly (``assemble_local``,
fom/base.py:523-599) for the two named problems, with the FEniCS element integrals replaced by their closed forms on
the uniformly scaled 1-D P1 mesh.  Never imported by romtime_b200/.

What is restated (every weak form is at most quadratic in x per factor, so FEniCS' degree-2 expression interpolation
and its automatic quadrature are exact and the closed forms below ARE what ``fenics.assemble_local`` returns):

* mesh / dof map: ``IntervalMesh(nx, 0, L0)`` with P1 elements (fom/base.py:269-272); FEniCS numbers the dofs of an
  interval mesh from the right end: dof i sits at vertex nx - i.  ``move_mesh`` scales every coordinate by L(t)
  (fom/base.py:239-255) and ``self.L`` is the current right end (fom/base.py:156-163).
* Dirichlet entries: rows of constrained dofs are the identity (DIRICHLET_ENTRY = 1 on the diagonal, DIRICHLET_VALUE = 0
  in vectors; fom/base.py:50-51, :320-348, :547-552); columns are left alone (``bc.apply`` only touches rows).
* MFP1 / heat (problems/mfp1.py:27-75; fom/heat.py:42-55, :80-169, :242-297; lifting fom/base.py:377-495):
  mass, stiffness alpha_0 grad u . grad v, ALE convection -w u_x v with w = x L'/L, forcing f v, lifting
  -(dg/dt v + alpha_0 g_x v_x), rhs = forcing + lifting.
* piston / Burgers (problems/piston.py:18-60; fom/nonlinear.py:159-166, :203-233, :277-299, :374-581): mass, stiffness
  (ALPHA = 1e-10), convection -(a0 + w) u_x v, nonlinear lifting b0 (g u_x + g_x u) v, trilinear b0 u_n u_x v,
  rhs = lifting = -(dg/dt v + b0 g g_x v + ALPHA g_x v_x) + (a0 + w) g_x v.

Pin: the operator entries hard-coded in the reference's own tests/test_mpf1.py:126-317 (nx = 3; committed as
tests/golden/mfp1_nx3.json) -- see tests/test_p1_forms.py.  The piston forms have no reference-held numbers (every
reference test of them needs FEniCS): parity UNPINNED for those, they are checked through properties only.
"""
import numpy as np

GAMMA = 1.4       # fom/nonlinear.py:31
ALPHA = 1e-10     # fom/nonlinear.py:27


class P1Interval:
    """``IntervalMesh(nx, 0, L0)`` + P1 dof map in FEniCS' numbering (dof i at vertex nx - i)."""

    def __init__(self, nx, L0, dirichlet=("left", "right")):
        self.nx, self.L0, self.Nh = int(nx), float(L0), int(nx) + 1
        self.h0 = self.L0 / self.nx
        self.x0 = self.L0 * (self.nx - np.arange(self.Nh)) / self.nx
        self.dirichlet_dofs = set()
        if "left" in dirichlet:
            self.dirichlet_dofs.add(self.nx)
        if "right" in dirichlet:
            self.dirichlet_dofs.add(0)

    def cells(self, dof):
        """Cells covered by the hat function of ``dof`` (fom/base.py:601-624): cell e spans vertices e, e + 1."""
        v = self.nx - dof
        return [e for e in (v - 1, v) if 0 <= e < self.nx]

    def cell_dofs(self, e):
        """(dof of the left vertex, dof of the right vertex) of cell e."""
        return self.nx - e, self.nx - e - 1

    def pattern(self):
        """Non-zero pattern of an assembled operator with the Dirichlet rows reduced to their diagonal (row-major order)."""
        out = []
        for i in range(self.Nh):
            for j in range(self.Nh):
                if abs(i - j) <= 1 and (i not in self.dirichlet_dofs or i == j):
                    out.append((i, j))
        return out


# ---- element integrals on the cell [xa, xb], h = xb - xa; local index 0 = left vertex, 1 = right vertex -------------------
def _hat(xa, xb):
    return 0.5 * (xb - xa) * np.ones(2)                                    # int phi_i


def _xhat(xa, xb):
    h = xb - xa
    return h * np.array([2.0 * xa + xb, xa + 2.0 * xb]) / 6.0             # int x phi_i


def _x2hat(xa, xb):
    h = xb - xa
    return h * np.array([3 * xa * xa + 2 * xa * xb + xb * xb, xa * xa + 2 * xa * xb + 3 * xb * xb]) / 12.0   # int x^2 phi_i


def _mass(xa, xb):
    return (xb - xa) / 6.0 * np.array([[2.0, 1.0], [1.0, 2.0]])            # int phi_i phi_j


def _dphi(xa, xb):
    return np.array([-1.0, 1.0]) / (xb - xa)                               # phi_j'


class _Problem:
    """Sampled-entry assembly shared by both problems: loop over entries, Dirichlet rule, cells to cover, local index."""

    def __init__(self, mesh):
        self.mesh = mesh

    def scale(self, mu, t):
        return 1.0

    def local_matrix(self, op, mu, t, xa, xb, L, u_loc):
        raise NotImplementedError

    def local_vector(self, op, mu, t, xa, xb, L):
        raise NotImplementedError

    def assemble(self, op, mu, t, entries, u_n=None):
        """fom/base.py:523-599 with closed-form local tensors.  ``entries``: list of (i,) or (i, j) tuples."""
        m, s = self.mesh, self.scale(mu, t)
        L = s * m.L0                       # self.L of the moved mesh
        is_vec = len(entries[0]) == 1
        out = []
        for entry in entries:
            if is_vec and entry[0] in m.dirichlet_dofs:
                out.append(0.0)            # DIRICHLET_VALUE
                continue
            if not is_vec and entry[0] in m.dirichlet_dofs:
                # rows of constrained dofs: 1 on the diagonal; off-diagonal entries are not part of any pattern
                out.append(1.0 if entry[0] == entry[1] else 0.0)
                continue
            cells = set()
            for dof in entry:
                cells.update(m.cells(dof))
            val = 0.0
            for e in sorted(cells):
                da, db = m.cell_dofs(e)
                loc = (da, db)
                if any(d not in loc for d in entry):
                    continue
                xa, xb = s * m.x0[da], s * m.x0[db]
                if is_vec:
                    val += self.local_vector(op, mu, t, xa, xb, L)[loc.index(entry[0])]
                else:
                    u_loc = None if u_n is None else np.array([u_n[da], u_n[db]])
                    val += self.local_matrix(op, mu, t, xa, xb, L, u_loc)[loc.index(entry[0]), loc.index(entry[1])]
            out.append(val)
        return np.array(out)

    def full_matrix(self, op, mu, t, u_n=None):
        m = self.mesh
        pat = m.pattern()
        A = np.zeros((m.Nh, m.Nh))
        for (i, j), v in zip(pat, self.assemble(op, mu, t, pat, u_n=u_n)):
            A[i, j] = v
        return A

    def full_vector(self, op, mu, t):
        return self.assemble(op, mu, t, [(i,) for i in range(self.mesh.Nh)])


class Mfp1(_Problem):
    """MFP1 heat problem, fixed (``moving=False``: HeatEquationSolver) or ALE (HeatEquationMovingSolver, mu has omega)."""

    def __init__(self, nx, L0, moving=False):
        super().__init__(P1Interval(nx, L0, ("left", "right")))
        self.moving = moving

    def scale(self, mu, t):
        return 1.0 - np.sin(mu["omega"] * t) if self.moving else 1.0          # problems/mfp1.py:46-59

    def _dL(self, mu, t):
        return -mu["omega"] * np.cos(mu["omega"] * t) if self.moving else 0.0  # problems/mfp1.py:61-73

    def _boundary(self, mu, t, L):
        """b0, bL, db0/dt, dbL/dt (problems/mfp1.py:27-36); ``dLt_dt`` inside the expressions is dL/dt * L0 (fom/base.py:411-412)."""
        d2, beta = mu["delta"] ** 2, mu["beta"]
        E = np.exp(-beta * t)
        dLt = self._dL(mu, t) * self.mesh.L0
        b0 = 1.0 - E
        bL = (1.0 - E) * (1.0 + d2 * L * L)
        db0 = beta * E
        dbL = beta * E * (1.0 + d2 * L * L) + 2.0 * (1.0 - E) * d2 * L * dLt
        return b0, bL, db0, dbL, dLt

    def local_matrix(self, op, mu, t, xa, xb, L, u_loc):
        if op == "mass":
            return _mass(xa, xb)
        if op == "stiffness":
            dp = _dphi(xa, xb)
            return mu["alpha_0"] * (xb - xa) * np.outer(dp, dp)
        if op == "convection":                                               # fom/heat.py:268-286: -w u_x v, w = x L'/L
            c1 = self._dL(mu, t) / self.scale(mu, t)
            return -np.outer(c1 * _xhat(xa, xb), _dphi(xa, xb))
        raise KeyError(op)

    def local_vector(self, op, mu, t, xa, xb, L):
        d2, beta, a0 = mu["delta"] ** 2, mu["beta"], mu["alpha_0"]
        E = np.exp(-beta * t)
        if op == "forcing":                                                  # problems/mfp1.py:38-39
            return beta * E * (_hat(xa, xb) + d2 * _x2hat(xa, xb)) - 2.0 * d2 * a0 * (1.0 - E) * _hat(xa, xb)
        if op == "lifting":                                                  # fom/heat.py:131-169, fom/base.py:377-495
            b0, bL, db0, dbL, dLt = self._boundary(mu, t, L)
            # dg/dt = dbL x/L + db0 (L - x)/L + (b0 - bL) (x/L) (dLt/L);  g_x = (bL - b0)/L
            c1 = (dbL - db0) / L + (b0 - bL) * dLt / (L * L)
            dgdt = db0 * _hat(xa, xb) + c1 * _xhat(xa, xb)
            gx = (bL - b0) / L
            return -(dgdt + a0 * gx * _dphi(xa, xb) * (xb - xa))
        if op == "rhs":                                                      # fom/heat.py:171-188
            return self.local_vector("forcing", mu, t, xa, xb, L) + self.local_vector("lifting", mu, t, xa, xb, L)
        raise KeyError(op)


class Piston(_Problem):
    """Moving-piston Burgers problem (OneDimensionalBurgers, fom/nonlinear.py); mu = {a0, omega, delta}."""

    def __init__(self, nx, L0=1.0, which="rest"):
        super().__init__(P1Interval(nx, L0, ("right",)))                     # fom/nonlinear.py:236-258: right end only
        self.which = which

    def scale(self, mu, t):
        om, de = mu["omega"], mu["delta"]
        return 1.0 - de * (1.0 - np.cos(om * t)) if self.which == "rest" else 1.0 - de * np.sin(om * t)   # problems/piston.py:43-51

    def _dL(self, mu, t):
        om, de = mu["omega"], mu["delta"]
        return -om * de * np.sin(om * t) if self.which == "rest" else -om * de * np.cos(om * t)           # problems/piston.py:53-60

    def _boundary(self, mu, t):
        a0, om, de = mu["a0"], mu["omega"], mu["delta"]
        if self.which == "rest":                                             # problems/piston.py:22-27
            return -de * (om / a0) * np.sin(om * t), -de * om * (om / a0) * np.cos(om * t)
        return -de * (om / a0) * np.cos(om * t), de * om * (om / a0) * np.sin(om * t)

    def b0(self, mu):
        return (GAMMA + 1.0) / 2.0 * mu["a0"]                                # fom/nonlinear.py:159-166

    def local_matrix(self, op, mu, t, xa, xb, L, u_loc):
        dp = _dphi(xa, xb)
        if op == "mass":
            return _mass(xa, xb)
        if op == "stiffness":
            return ALPHA * (xb - xa) * np.outer(dp, dp)
        if op == "convection":                                               # -(a0 + w) u_x v, w = x L'/L
            c1 = self._dL(mu, t) / self.scale(mu, t)
            return -np.outer(mu["a0"] * _hat(xa, xb) + c1 * _xhat(xa, xb), dp)
        if op == "nonlinear_lifting":                                        # b0 (g u_x + g_x u) v, g = bL x/L
            bL, _ = self._boundary(mu, t)
            return self.b0(mu) * (bL / L) * (np.outer(_xhat(xa, xb), dp) + _mass(xa, xb))
        if op == "trilinear":                                                # b0 u_n u_x v
            return self.b0(mu) * np.outer(_mass(xa, xb) @ u_loc, dp)
        raise KeyError(op)

    def local_vector(self, op, mu, t, xa, xb, L):
        if op in ("lifting", "rhs"):                                         # fom/nonlinear.py:520-581
            bL, dbL = self._boundary(mu, t)
            a0, gx = mu["a0"], bL / L
            c1 = self._dL(mu, t) / self.scale(mu, t)
            time_derivative = (dbL / L) * _xhat(xa, xb)
            nonlinear = self.b0(mu) * (bL / L) * gx * _xhat(xa, xb)
            convection = gx * (a0 * _hat(xa, xb) + c1 * _xhat(xa, xb))
            diffusion = ALPHA * gx * _dphi(xa, xb) * (xb - xa)
            return -(time_derivative + nonlinear + diffusion) + convection
        raise KeyError(op)
