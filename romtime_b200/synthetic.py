"""Seeded synthetic inputs of the shapes named in BASELINE.json.

Host-side NumPy only: these generators produce the *inputs* (snapshot matrices,
collateral bases, hyper-reduced piston models, parameter batches) that tests and
bench.py feed to both the CUDA path and the CPU oracle.  No reference code is involved.

The hyper-reduced "piston" model follows the closed forms of the reference's Burgers /
piston problem on a uniformly scaled P1 mesh (reference problems/piston.py:18-60,
fom/nonlinear.py:374-581): every sampled operator entry is a linear combination of NF
scalar features of (mu, t), the trilinear operator additionally of the extrapolated
state u* (fom/nonlinear.py:400-417).
"""
import numpy as np

NF = 12  # number of (mu, t) features, see csrc/online.cu :: piston_features
OPS_MATRIX = ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear")
OPS = OPS_MATRIX + ("rhs",)

# piston parameter box (reference tests/fom/test_piston.py:89-96): a0, omega, delta
MU_LO = np.array([8.0, 15.0, 0.10])
MU_HI = np.array([10.0, 20.0, 0.15])


def sample_mus(B, seed=1005):
    rng = np.random.default_rng(seed)
    return MU_LO + (MU_HI - MU_LO) * rng.random((B, 3))


def make_piston_model(N=8, m=6, m_rhs=None, seed=1005, which="rest", mass_identity=1.0,
                      coupling=0.05):
    """Random hyper-reduced model with the structure RomConstructorNonlinear consumes
    (reference rom/rom.py:877-929): five matrix reductors (basis_rom N^2 x m, PT_U m x m)
    and one vector reductor (basis_rom N x m_rhs).

    ``C`` (m x NF) maps features to sampled entries, ``T`` (m x N) maps u* to the
    state-dependent part of the trilinear entries.  They are built as PT_U @ D so that the
    exact interpolation coefficients are theta = D phi (+ b0 * Dt u*).
    """
    rng = np.random.default_rng(seed)
    m_rhs = m if m_rhs is None else m_rhs
    model = {"N": N, "m": m, "m_rhs": m_rhs, "which": which, "nf": NF}

    def pt_u(mm):
        return np.eye(mm) + 0.25 * rng.standard_normal((mm, mm)) / np.sqrt(mm)

    eye = np.eye(N).reshape(-1)
    for op in OPS_MATRIX:
        basis = coupling * rng.standard_normal((N * N, m)) / np.sqrt(N)
        D = np.zeros((m, NF))
        T = None
        if op == "mass":
            basis[:, 0] = mass_identity * eye + (1.0 - mass_identity) * basis[:, 0]
            D[0, 1] = 1.0  # theta_0 = L(t)
            D[1:, 0:3] = 0.1 * rng.standard_normal((m - 1, 3))
        elif op == "stiffness":
            basis[:, 0] = eye
            D[0, 2] = 0.5  # ~ alpha / L
            D[1:, 2] = 0.3 * rng.standard_normal(m - 1)
        elif op == "convection":
            basis[:, 0] = eye
            D[0, 4] = 0.05  # ~ a0
            D[1:, 3] = 0.3 * rng.standard_normal(m - 1)
            D[1:, 4] = 0.05 * rng.standard_normal(m - 1)
        elif op == "nonlinear_lifting":
            D[:, 5] = 0.3 * rng.standard_normal(m)
        elif op == "trilinear":
            T_theta = 0.05 * rng.standard_normal((m, N))
        P = pt_u(m)
        entry = {"basis_rom": np.asfortranarray(basis), "PT_U": P, "C": P @ D, "D": D}
        if op == "trilinear":
            entry["T"] = P @ T_theta
        else:
            entry["T"] = None
        model[op] = entry
    basis = rng.standard_normal((N, m_rhs)) / np.sqrt(N)
    D = np.zeros((m_rhs, NF))
    D[:, 6:11] = 0.5 * rng.standard_normal((m_rhs, 5))
    P = pt_u(m_rhs)
    model["rhs"] = {"basis_rom": np.asfortranarray(basis), "PT_U": P, "C": P @ D, "D": D,
                    "T": None}
    return model


def pod_snapshots(Nh, n, seed=1003, modes=256, decay=32.0, noise=1e-13, order="C",
                  row0=0, rows=None):
    """C3 snapshot matrix: S[i,j] = sum_k sigma_k sin((k+1) pi x_i)
    cos((k+1) theta_j + phi_k) + noise * xi_ij, sigma_k = 10^(-k/decay).
    ``row0/rows`` generate a row slab so a 1M-row matrix can be produced in pieces."""
    rows = Nh if rows is None else rows
    rng = np.random.default_rng(seed)
    phase = 2.0 * np.pi * rng.random(modes)
    k = np.arange(1, modes + 1)
    sig = 10.0 ** (-(k - 1) / decay)
    x = (np.arange(row0, row0 + rows) + 0.5) / Nh
    th = 2.0 * np.pi * np.arange(n) / n
    left = np.sin(np.pi * np.outer(x, k)) * sig  # rows x modes
    right = np.cos(np.outer(k, th) + phase[:, None])  # modes x n
    S = left @ right
    if noise:
        S += noise * np.random.default_rng(seed + 7919 * (row0 + 1)).standard_normal(S.shape)
    return np.asfortranarray(S) if order == "F" else np.ascontiguousarray(S)


def treewalk_snapshots(Nh, nt, seed=0, modes=24, decay=2.5, noise=1e-12):
    """Time snapshots of ONE parameter of a tree walk at the reference's own sizes (C1/C2: Nh = 100-1000 dofs, nt = 10-250
    instants): a smooth moving-domain field u(x, t; mu) = sum_k sigma_k sin((k+1) pi x + a_k) cos((k+1) w t + b_k) / L(t),
    L(t) = 1 - delta (1 - cos w t), mu = (a0, w, delta) drawn from the piston box by ``seed``, plus round-off-level noise.
    Numerical rank ~ 20-60, so tol-driven truncation gives ragged widths across parameters."""
    rng = np.random.default_rng(1000 + seed)
    a0, om, de = MU_LO + (MU_HI - MU_LO) * rng.random(3)
    x = np.linspace(0.0, 1.0, Nh)[:, None]
    t = np.linspace(0.0, 1.0, nt)[None, :]
    L = 1.0 - de * (1.0 - np.cos(om * t))
    k = np.arange(modes)
    sig = 10.0 ** (-k / decay)
    pa, pb = rng.random(modes) * np.pi, rng.random(modes) * 2.0 * np.pi
    S = np.zeros((Nh, nt))
    for j in range(modes):
        S += sig[j] * np.sin((j + 1) * np.pi * x + pa[j]) * np.cos((j + 1) * om * t / 5.0 + pb[j])
    return S / L + noise * rng.standard_normal((Nh, nt))


def mdeim_snapshots(nnz, n, seed=1004, order="F"):
    """C4-style MDEIM snapshots: vectorised non-zeros of a tridiagonal-like operator,
    a_k(mu,t) = c0 + c1 L + c2 / L + c3 L'/L + c4 sin(w t + psi_k); every off-diagonal
    pair is duplicated bit-for-bit (symmetric part) so the greedy meets exact ties."""
    rng = np.random.default_rng(seed)
    nrow = (nnz + 2) // 3
    coef = rng.standard_normal((nrow, 2, 5))  # [row][diag/offdiag][term]
    psi = 2.0 * np.pi * rng.random((nrow, 2))
    mu = sample_mus(n, seed + 1)
    t = rng.random(n)
    L = 1.0 - mu[:, 2] * (1.0 - np.cos(mu[:, 1] * t))
    dL = -mu[:, 1] * mu[:, 2] * np.sin(mu[:, 1] * t)
    feats = np.stack([np.ones(n), L, 1.0 / L, dL / L], axis=0)  # 4 x n

    def vals(kind):
        v = coef[:, kind, :4] @ feats
        v += coef[:, kind, 4:5] * np.sin(np.outer(np.ones(nrow), mu[:, 1] * t) + psi[:, kind:kind + 1])
        return v

    diag, off = vals(0), vals(1)
    # CSR order of a symmetric tridiagonal matrix: row i holds (i,i-1),(i,i),(i,i+1);
    # (i,i+1) and (i+1,i) carry the same value.
    out = np.empty((3 * nrow, n))
    out[0::3] = np.vstack([off[:1] * 0.0, off[:-1]])  # (i, i-1) == off[i-1]
    out[1::3] = diag
    out[2::3] = off  # (i, i+1) == off[i]
    out = out[1:nnz + 1]
    return np.asfortranarray(out) if order == "F" else np.ascontiguousarray(out)


def mdeim_snapshots_device(nnz, n, seed=1004, device=None, bump_width=0.01):
    """C4 at full size, evaluated on the device (torch only as the array library): the affine family of
    ``mdeim_snapshots`` plus a non-affine Lorentzian bump 1 / (1 + ((x_k - p_j) / w)^2) whose centre p_j
    moves with the sample, so that the snapshot matrix has more than 128 modes above the Gram route's
    noise floor (sigma_128 / sigma_1 ~ 1e-3 at w = 0.01).  Column-major (nnz x n) CUDA tensor; every
    off-diagonal pair is duplicated bit-for-bit, as in the host version."""
    import torch

    rng = np.random.default_rng(seed)
    nrow = (nnz + 2) // 3
    coef = rng.standard_normal((nrow, 2, 6))
    psi = 2.0 * np.pi * rng.random((nrow, 2))
    mu = sample_mus(n, seed + 1)
    t = rng.random(n)
    p = rng.random(n)
    L = 1.0 - mu[:, 2] * (1.0 - np.cos(mu[:, 1] * t))
    dL = -mu[:, 1] * mu[:, 2] * np.sin(mu[:, 1] * t)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    feats = up(np.stack([np.ones(n), L, 1.0 / L, dL / L], axis=0))      # 4 x n
    wt, pj = up(mu[:, 1] * t), up(p)
    out = torch.empty((n, 3 * nrow), dtype=torch.float64, device=dev)   # transposed storage = column-major snapshots
    x = (torch.arange(nrow, dtype=torch.float64, device=dev) + 0.5) / nrow
    cols = 256
    for c0 in range(0, n, cols):
        c1 = min(n, c0 + cols)
        for kind in (0, 1):
            ck, pk = up(coef[:, kind, :]), up(psi[:, kind])
            xs = x + (0.5 / nrow if kind else 0.0)
            v = feats[:, c0:c1].t() @ ck[:, :4].t()                                          # (cols x nrow)
            v += torch.sin(wt[c0:c1, None] + pk[None, :]) * ck[None, :, 4]
            v += ck[None, :, 5] / (1.0 + torch.square((xs[None, :] - pj[c0:c1, None]) / bump_width))
            if kind == 0:
                out[c0:c1, 1::3] = v
            else:
                out[c0:c1, 2::3] = v                                                         # (i, i+1)
                out[c0:c1, 3::3] = v[:, :-1]                                                 # (i+1, i): same bits
                out[c0:c1, 0] = 0.0
    return out[:, 1:nnz + 1].contiguous().t()   # fresh allocation: 16-byte aligned base for the TMA descriptors
