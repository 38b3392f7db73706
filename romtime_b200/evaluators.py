"""Sampled-entry evaluator tables for the two named problems.

Online, the reference evaluates the m sampled entries of every reduced operator with FEniCS local assembly
(``assemble_local``, fom/base.py:523-599) at every (mu, t).  On the uniformly scaled 1-D P1 mesh
(``IntervalMesh(nx, 0, L0)`` moved by ``mesh.scale(L(t))``, fom/base.py:239-272) every weak form of the two problems
factorises into

    entry_k(mu, t[, u*]) = sum_f C[k, f] phi_f(mu, t)  +  b0(mu) (T u*)_k

with a handful of scalar features phi_f (closed forms in L(t), L'(t), the boundary data and the parameters -- evaluated
on the device by k_state, csrc/online.cu) and GEOMETRY-ONLY tables C (m x RT_NF) and T (m x N): element integrals of
hat functions on the reference mesh.  This module derives those tables from (mesh, sampled entries, reduced basis rows)
-- nothing is assembled per step, the sweep (rt_online_sweep_f64) evaluates the entries itself.

* ``PistonP1``  moving-piston Burgers problem (fom/nonlinear.py:374-581, problems/piston.py:18-60); features
  [1, L, 1/L, L', a0, b0 bL, bL' L, b0 bL^2, a0 bL, L' bL, bL/L, b0]; the trilinear term b0 u_n u_x v needs only the rows
  of V its sampled entries touch (the restricted lift):  T = Tfom[:, touched] V[touched, :].
* ``Mfp1P1``    MFP1 heat problem, fixed or ALE (fom/heat.py:80-297, problems/mfp1.py:27-75; omega = 0 is the fixed
  domain); features [1, L, alpha_0 / L, L', beta E L, beta E delta^2 L^3, alpha_0 (1 - E) delta^2 L,
  (1 - E) delta^2 L' L^2], E = exp(-beta t); parameters (delta, beta, alpha_0, omega).

FEniCS numbers the dofs of an interval mesh from the right end (dof i at vertex nx - i); Dirichlet rows are the
identity (fom/base.py:50-51, :547-552).  The CPU restatement that the tables are tested against lives in
oracle/p1_forms.py (pinned by the reference's own tests/test_mpf1.py numbers).
"""
import numpy as np

NF = 12           # RT_NF, include/romtime_b200.h
GAMMA = 1.4       # fom/nonlinear.py:31
ALPHA = 1e-10     # fom/nonlinear.py:27


class _RefMesh:
    """Reference-configuration element integrals of the P1 hat functions, summed over the cells an entry touches."""

    def __init__(self, nx, L0, dirichlet):
        self.nx, self.L0, self.Nh = int(nx), float(L0), int(nx) + 1
        self.h = self.L0 / self.nx
        self.dirichlet = set()
        if "left" in dirichlet:
            self.dirichlet.add(self.nx)
        if "right" in dirichlet:
            self.dirichlet.add(0)

    def _cells(self, i, j=None):
        """Cells shared by the hats of dofs i (and j): (left coordinate, local index of i, local index of j)."""
        vi = self.nx - i
        vj = vi if j is None else self.nx - j
        out = []
        for e in (vi - 1, vi):
            if 0 <= e < self.nx and vj in (e, e + 1):
                out.append((e * self.h, vi - e, vj - e))
        return out

    # per-cell integrals on [xa, xa + h]; local index 0 = left vertex
    def hat(self, i):                      # int phi_i
        return sum(0.5 * self.h for _ in self._cells(i))

    def xhat(self, i):                     # int x phi_i
        return sum(self.h * ((2 * xa + xa + self.h) if a == 0 else (xa + 2 * (xa + self.h))) / 6.0 for xa, a, _ in self._cells(i))

    def x2hat(self, i):                    # int x^2 phi_i
        s = 0.0
        for xa, a, _ in self._cells(i):
            xb = xa + self.h
            s += self.h * ((3 * xa * xa + 2 * xa * xb + xb * xb) if a == 0 else (xa * xa + 2 * xa * xb + 3 * xb * xb)) / 12.0
        return s

    def dsum(self, i):                     # int phi_i'
        return sum(-1.0 if a == 0 else 1.0 for _, a, _ in self._cells(i))

    def mass(self, i, j):                  # int phi_i phi_j
        return sum(self.h / 6.0 * (2.0 if a == b else 1.0) for _, a, b in self._cells(i, j))

    def stiff(self, i, j):                 # int phi_i' phi_j'
        return sum((1.0 if a == b else -1.0) / self.h for _, a, b in self._cells(i, j))

    def hat_d(self, i, j):                 # int phi_i phi_j'
        return sum(0.5 * self.h * (-1.0 if b == 0 else 1.0) / self.h for _, a, b in self._cells(i, j))

    def xhat_d(self, i, j):                # int x phi_i phi_j'
        s = 0.0
        for xa, a, b in self._cells(i, j):
            xb = xa + self.h
            xi = self.h * ((2 * xa + xb) if a == 0 else (xa + 2 * xb)) / 6.0
            s += xi * (-1.0 if b == 0 else 1.0) / self.h
        return s

    def tri_row(self, i, j):
        """Row of the trilinear form over FOM dofs: entry(i, j) = b0 sum_k row[k] u_k = b0 phi_j' int u_n phi_i."""
        row = {}
        for xa, a, b in self._cells(i, j):
            e = int(round(xa / self.h))
            d = (-1.0 if b == 0 else 1.0) / self.h
            for loc, dof in ((0, self.nx - e), (1, self.nx - e - 1)):
                row[dof] = row.get(dof, 0.0) + d * self.h / 6.0 * (2.0 if loc == a else 1.0)
        return row


def _check_entries(entries, vector):
    ent = [tuple(int(x) for x in e) for e in entries]
    if any(len(e) != (1 if vector else 2) for e in ent):
        raise ValueError("entries must be (dof,) tuples for vectors and (row, col) tuples for matrices")
    return ent


class PistonP1:
    """Tables of the moving-piston problem.  ``which``: "rest" / "sudden" (problems/piston.py:18-60) -- selects the device
    feature set (rt_rom_model.evaluator 0 / 1); parameters per sweep row: (a0, omega, delta)."""

    OPS = ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear", "rhs")
    N_MU = 3

    def __init__(self, nx, L0=1.0, which="rest", alpha=ALPHA):
        if which not in ("rest", "sudden"):
            raise NotImplementedError("Which case do you want to solve?")
        self.ref = _RefMesh(nx, L0, ("right",))      # fom/nonlinear.py:236-258: homogeneous Dirichlet at the piston only
        self.which, self.alpha = which, float(alpha)

    @property
    def evaluator_id(self):
        return 0 if self.which == "rest" else 1

    def tables(self, op, entries, V=None):
        """``{"C": (m, NF)[, "T": (m, N)]}`` for the sampled ``entries`` of operator ``op`` (dofs as stored by
        ``store_dofs``: (i,) or (row, col)).  ``V`` (Nh x N): reduced basis, needed by "trilinear"."""
        r, L0 = self.ref, self.ref.L0
        ent = _check_entries(entries, op == "rhs")
        C = np.zeros((len(ent), NF))
        T = None
        if op == "trilinear":
            if V is None:
                raise ValueError("the trilinear tables need the reduced basis V (restricted lift)")
            V = np.asarray(V, dtype=float)
            T = np.zeros((len(ent), V.shape[1]))
        for k, e in enumerate(ent):
            i = e[0]
            if i in r.dirichlet:
                if op != "rhs" and e[1] == i:
                    C[k, 0] = 1.0                     # identity row of a constrained dof
                continue
            if op == "mass":
                C[k, 1] = r.mass(*e)
            elif op == "stiffness":
                C[k, 2] = self.alpha * r.stiff(*e)
            elif op == "convection":                  # -(a0 + x L'/L) u_x v
                C[k, 4] = -r.hat_d(*e)
                C[k, 3] = -r.xhat_d(*e)
            elif op == "nonlinear_lifting":           # b0 (g u_x + g_x u) v, g = bL x / L
                C[k, 5] = (r.xhat_d(*e) + r.mass(*e)) / L0
            elif op == "trilinear":                   # b0 u_n u_x v: only the rows of V under the sampled hats
                for dof, c in r.tri_row(*e).items():
                    T[k] += c * V[dof]
            elif op == "rhs":                         # -(dg/dt v + b0 g g_x v + alpha g_x v_x) + (a0 + w) g_x v
                C[k, 6] = -r.xhat(i) / L0
                C[k, 7] = -r.xhat(i) / (L0 * L0)
                C[k, 10] = -self.alpha * r.dsum(i) / L0
                C[k, 8] = r.hat(i) / L0
                C[k, 9] = r.xhat(i) / L0
            else:
                raise KeyError(op)
        return {"C": C, "T": T} if T is not None else {"C": C}

    def lifting(self, mus, ts):
        """Boundary lifting at the mesh nodes as coef (B, nt, 1) @ shapes (1, Nh): g_h = bL(mu, t) x / L(t), and x / L is the
        reference coordinate x0 / L0 on the scaled mesh (fom/nonlinear.py:119-150, rom/rom.py:505-522)."""
        ts = np.asarray(ts, dtype=float)
        coef = np.empty((len(mus), len(ts), 1))
        for b, mu in enumerate(mus):
            a0, om, de = (mu["a0"], mu["omega"], mu["delta"]) if isinstance(mu, dict) else mu
            coef[b, :, 0] = -de * (om / a0) * (np.sin(om * ts) if self.which == "rest" else np.cos(om * ts))
        x0 = self.ref.L0 * (self.ref.nx - np.arange(self.ref.Nh)) / self.ref.nx
        return coef, (x0 / self.ref.L0)[None, :]

    # host restatement of the device features (k_state), for tests and for callers that want entries on the host
    def features(self, mu, t):
        a0, om, de = (mu["a0"], mu["omega"], mu["delta"]) if isinstance(mu, dict) else mu
        s, c = np.sin(om * t), np.cos(om * t)
        if self.which == "rest":
            L, dL, bL, dbL = 1.0 - de * (1.0 - c), -om * de * s, -de * (om / a0) * s, -de * om * (om / a0) * c
        else:
            L, dL, bL, dbL = 1.0 - de * s, -om * de * c, -de * (om / a0) * c, de * om * (om / a0) * s
        b0 = (GAMMA + 1.0) / 2.0 * a0
        return np.array([1.0, L, 1.0 / L, dL, a0, b0 * bL, dbL * L, b0 * bL * bL, a0 * bL, dL * bL, bL / L, b0])


class Mfp1P1:
    """Tables of the MFP1 heat problem on the fixed (omega = 0) or moving domain L(t) = 1 - sin(omega t); device feature
    set rt_rom_model.evaluator 2; parameters per sweep row: (delta, beta, alpha_0, omega)."""

    OPS = ("mass", "stiffness", "convection", "rhs")
    N_MU = 4
    evaluator_id = 2

    def __init__(self, nx, L0):
        self.ref = _RefMesh(nx, L0, ("left", "right"))

    def tables(self, op, entries, V=None):
        r, L0 = self.ref, self.ref.L0
        ent = _check_entries(entries, op == "rhs")
        C = np.zeros((len(ent), NF))
        for k, e in enumerate(ent):
            i = e[0]
            if i in r.dirichlet:
                if op != "rhs" and e[1] == i:
                    C[k, 0] = 1.0
                continue
            if op == "mass":
                C[k, 1] = r.mass(*e)
            elif op == "stiffness":                   # alpha_0 grad u . grad v
                C[k, 2] = r.stiff(*e)
            elif op == "convection":                  # -x L'/L u_x v
                C[k, 3] = -r.xhat_d(*e)
            elif op == "rhs":                         # forcing + lifting (fom/heat.py:99-188, fom/base.py:377-495)
                C[k, 4] = 0.0                         # beta E L (int phi_i - int phi_i): forcing and dg/dt cancel exactly
                C[k, 5] = r.x2hat(i) - L0 * r.xhat(i)
                C[k, 6] = -2.0 * r.hat(i) - L0 * r.dsum(i)
                C[k, 7] = -L0 * r.xhat(i)
            else:
                raise KeyError(op)
        return {"C": C}

    def lifting(self, mus, ts):
        """g_h = bL x / L + b0 (L - x) / L at the nodes (fom/base.py:440-465) as coef (B, nt, 2) @ shapes (2, Nh):
        shapes = [x0 / L0, 1 - x0 / L0], coef = [bL, b0] with bL = (1 - E)(1 + delta^2 L(t)^2 L0^2), b0 = 1 - E."""
        ts = np.asarray(ts, dtype=float)
        coef = np.empty((len(mus), len(ts), 2))
        for b, mu in enumerate(mus):
            de, be, al, om = (mu["delta"], mu["beta"], mu["alpha_0"], mu.get("omega", 0.0)) if isinstance(mu, dict) else mu
            E = np.exp(-be * ts)
            Lc = (1.0 - np.sin(om * ts)) * self.ref.L0
            coef[b, :, 0] = (1.0 - E) * (1.0 + de * de * Lc * Lc)
            coef[b, :, 1] = 1.0 - E
        x0 = self.ref.L0 * (self.ref.nx - np.arange(self.ref.Nh)) / self.ref.nx
        return coef, np.stack([x0 / self.ref.L0, 1.0 - x0 / self.ref.L0])

    def features(self, mu, t):
        de, be, al, om = (mu["delta"], mu["beta"], mu["alpha_0"], mu.get("omega", 0.0)) if isinstance(mu, dict) else mu
        L, dL, E, d2 = 1.0 - np.sin(om * t), -om * np.cos(om * t), np.exp(-be * t), de * de
        phi = np.zeros(NF)
        phi[:8] = [1.0, L, al / L, dL, be * E * L, be * E * d2 * L ** 3, al * (1.0 - E) * d2 * L, (1.0 - E) * d2 * dL * L * L]
        return phi
