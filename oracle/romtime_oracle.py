"""CPU ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A NumPy/SciPy restatement of the romtime hot path (reference: KikeM/romtime @ 7bca28a,
paths relative to /root/reference/src/romtime).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module; nothing under ``romtime_b200/`` does.

Pinning: the reference is pure Python, so the restatement is pinned against *outputs of
the reference itself* run in the build container (oracle/gen_golden.py ->
tests/golden/*.npz; checked by tests/test_oracle_golden.py).  The arithmetic of record
lives in un-vendored third parties (LAPACK dgesvd via scipy.linalg.svd, dgesv via
np.linalg.solve, scipy GMRES); the container's NumPy 2.3 / SciPy 1.18 / OpenBLAS 0.3.30
are newer than the reference's pins (environment.yml:77,107,204), so parity is defined
against this container's execution of the reference code.  The reference's own runnable
tests do not touch this path -- "parity unpinned by the
reference's own tests, pinned by reference-generated golden vectors".
"""
import numpy as np
import scipy.linalg
import scipy.sparse
import scipy.sparse.linalg

DROP_TOLERANCE = 1e-7  # rom/pod.py:4


# ------------------------------------------------------------------ POD (rom/pod.py:7-62)
def orth(snapshots, num=None, tol=None, normalize=True, return_VT=False):
    """Column-normalise, thin SVD (gesvd), energy, truncate.  rom/pod.py:7-62.

    Truncation priority: ``tol`` (strict ``energy < tol`` mask, :46-49) beats ``num``
    (:51-53) beats the default ``s > 1e-7`` floor (:55-57).  ``s`` and ``energy`` are
    returned untruncated (:59-62).
    """
    if isinstance(snapshots, list):  # :27-28
        raise ValueError("You should use an array, not a list.")
    A = np.asarray(snapshots, dtype=float)
    if normalize == True:  # noqa: E712  (:31-33)
        A = A / np.linalg.norm(A, axis=0)
    u, s, vt = scipy.linalg.svd(A, full_matrices=False, lapack_driver="gesvd")  # :38
    lam = s ** 2
    energy = np.cumsum(lam) / np.sum(lam)  # :41-43
    if tol:
        keep = energy < tol
        Q, VT = u[:, keep], vt[keep, :]
    elif num:
        Q, VT = u[:, :num], vt[:num, :]
    else:
        keep = s > DROP_TOLERANCE
        Q, VT = u[:, keep], vt[keep, :]
    if return_VT:
        return Q, s, energy, VT
    return Q, s, energy


# ------------------------------------------- DEIM greedy (deim/deim.py:18-22, :517-561)
def deim_greedy(basis):
    """Classic DEIM point selection.  deim/deim.py:517-561.

    p_0 = argmax|u_0| (:530-531).  For j >= 1: c = (P^T U)^-1 P^T u_j with np.linalg.solve
    (:545-547), residual = u_j - U c (:550), p_j = argmax|residual| (:553; np.argmax keeps
    the FIRST maximum).  The reference builds the dense one-hot P (Nh x j) and multiplies
    by it; P^T X is exactly the row gather X[dofs] (adding exact zeros), which is what is
    used here.
    """
    U = np.asarray(basis, dtype=float)
    Nh, m = U.shape
    dofs = [int(np.argmax(np.abs(U[:, 0])))]
    for j in range(1, m):
        uj = U[:, j]
        Uj = U[:, :j]
        coeff = np.linalg.solve(Uj[dofs, :], uj[dofs])
        residual = uj - Uj @ coeff
        dofs.append(int(np.argmax(np.abs(residual))))
    return dofs


def interpolation_matrix(basis, dofs):
    """PT_U = P^T @ basis_fom == basis_fom[dofs, :].  deim/deim.py:159, :212."""
    return np.asarray(basis)[np.asarray(dofs, dtype=np.int64), :].copy()


def one_hot(Nh, dofs):
    """Dense P as returned by build_interpolation_mesh (deim/deim.py:18-22, :556-557)."""
    P = np.zeros((Nh, len(dofs)))
    P[np.asarray(dofs), np.arange(len(dofs))] = 1.0
    return P


# ------------------------------------------------ theta solve (deim/deim.py:477-493)
def compute_thetas(PT_U, rhs):
    """theta = PT_U^-1 rhs via dgesv (LU re-factored each call).  deim/deim.py:491-492."""
    return np.linalg.solve(PT_U, rhs)


# ----------------------------------------------- projections (deim.py:495-515, mdeim.py:153-192)
def deim_project(V, basis_fom):
    """basis_rom = V^T U_f.  deim/deim.py:508-509."""
    return np.matmul(V.T, basis_fom)


def mdeim_project(basis_fom, rows, cols, V):
    """For each collateral mode q: A_q = csr(U[:,q]; rows, cols) (utils.py:136-149),
    A_q^r = V^T (A_q V) (utils.py:96-113), flattened row-major (mdeim.py:183) and stacked
    as columns -> (N_V^2, m)  (mdeim.py:170-186)."""
    m = basis_fom.shape[1]
    out = []
    for q in range(m):
        Aq = scipy.sparse.csr_matrix((basis_fom[:, q], (rows, cols)))
        AN = np.matmul(V.T, Aq.dot(V))
        out.append(AN.flatten())
    return np.array(out).T


# -------------------------------------------------- interpolation (deim/deim.py:416-452)
def interpolate(Vf, PT_U, f_local, mdeim_fom=False):
    """theta solve + sum_i theta_i Vf[:, i]  (deim/deim.py:441-445; N-MDEIM
    deim/nonlinear.py:272-276).  ``mdeim_fom`` forces entry 0 to 1.0 (:449-450)."""
    thetas = compute_thetas(PT_U, f_local)
    m = Vf.shape[1]
    approximation = np.sum([thetas[i] * Vf[:, i] for i in range(m)], axis=0)
    if mdeim_fom:
        approximation[0] = 1.0
    return approximation


def interpolate_rom_matrix(basis_rom, PT_U, f_local, N_V):
    """MDEIM ``interpolate(which="rom")``: reshape to (N_V, N_V).  deim/mdeim.py:247-251."""
    return interpolate(basis_rom, PT_U, f_local).reshape((N_V, N_V))


# ------------------------------------------------ certification helpers (utils.py:173-243)
def compute_rom_difference(uN, uN_srom, V_srom):
    """S-ROM estimator.  utils.py:173-212."""
    pad = np.zeros(len(uN_srom))
    pad[: len(uN)] = uN
    diff = uN_srom - pad
    lincomb = np.sum(diff * V_srom, axis=1)
    return np.linalg.norm(lincomb, ord=2) / np.sqrt(len(lincomb))


def rom_difference_batch(uN, uN_srom, V_srom):
    """compute_rom_difference for every (parameter, step): uN (B, nt, N), uN_srom (B, nt, Nhat) -> (B, nt)."""
    B, nt, _ = uN_srom.shape
    return np.array([[compute_rom_difference(uN[b, n], uN_srom[b, n], V_srom) for n in range(nt)] for b in range(B)])


def errors_summary(err):
    """create_errors_summary per parameter: columns mean, median, max, min.  rom/base.py:154-163."""
    return np.array([[np.mean(e), np.median(e), np.max(e), np.min(e)] for e in err])


def rms_error(u, ue):
    """Reductor._compute_error.  rom/base.py:52-73."""
    return np.linalg.norm(u - ue, ord=2) / np.sqrt(len(u))


def singular_to_energy(sigmas):
    lam = np.power(sigmas, 2)  # utils.py:236-243
    return np.cumsum(lam) / np.sum(lam)


def singular_to_pod_error(sigmas):
    lam = np.power(sigmas, 2)  # utils.py:224-233
    return np.sqrt(np.sum(lam) - np.cumsum(lam))


# --------------------------------------------------------------------------------------
# Synthetic "piston" hyper-reduced model: sampled entries are closed forms in (mu, t).
# Mirrors fom/nonlinear.py weak forms on a uniformly scaled P1 mesh (problems/piston.py
# :18-60 "rest" case): every sampled entry is  sum_f C[k, f] * phi_f(mu, t), plus
# phi_b0 * (T u*)_k for the state-dependent trilinear operator (fom/nonlinear.py:400-417).
# --------------------------------------------------------------------------------------
NF = 12
GAMMA = 1.4  # fom/nonlinear.py:31


def piston_features(a0, omega, delta, t, which="rest"):
    """phi(mu, t), length NF.  problems/piston.py:18-60; fom/nonlinear.py:159-166 (b0),
    :203-233 (lifting g = bL x/L), :277-299 (mesh velocity w = x L'/L)."""
    if which == "rest":
        L = 1.0 - delta * (1.0 - np.cos(omega * t))
        dL = -omega * delta * np.sin(omega * t)
        bL = -delta * (omega / a0) * np.sin(omega * t)
        dbL = -delta * omega * (omega / a0) * np.cos(omega * t)
    else:  # "sudden"
        L = 1.0 - delta * np.sin(omega * t)
        dL = -omega * delta * np.cos(omega * t)
        bL = -delta * (omega / a0) * np.cos(omega * t)
        dbL = delta * omega * (omega / a0) * np.sin(omega * t)
    b0 = (GAMMA + 1.0) / 2.0 * a0
    return np.array(
        [1.0, L, 1.0 / L, dL, a0, b0 * bL, dbL * L, b0 * bL * bL, a0 * bL, dL * bL, bL / L, b0]
    )


OPS = ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear", "rhs")


def sampled_entries(model, op, mu, t, ustar=None):
    """f_local for operator ``op`` (the assemble(mu,t,entries[,u_n]) callback of
    deim/deim.py:437-438 / deim/nonlinear.py:268)."""
    phi = piston_features(mu[0], mu[1], mu[2], t, model.get("which", "rest"))
    f = model[op]["C"] @ phi
    T = model[op].get("T")
    if T is not None:
        f = f + phi[11] * (T @ ustar)
    return f


def online_step_entries(model, F_row, uN_n, uN_n1, dt, bdf_scheme, step):
    """One implicit step of rom/rom.py:486-503 from a row of caller-supplied sampled entries (the f_local of every
    reductor side by side in OPS order, deim/deim.py:437-438): theta solves (:491-492), reduced operators (:445),
    K = bdf M + dt (A + C + N + N^) (rom/rom.py:906-907), b = M w + dt f (:919-925), dense solve.
    ``uN_n1`` is the array of zeros on the first step (:456-460), bdf = 1.0 there (:481-483)."""
    N = model["N"]
    mats, off = {}, 0
    for op in OPS:
        m = model[op]["PT_U"].shape[0]
        f = F_row[off:off + m]
        off += m
        if op == "rhs":
            fg = interpolate(model[op]["basis_rom"], model[op]["PT_U"], f)
        else:
            mats[op] = interpolate_rom_matrix(model[op]["basis_rom"], model[op]["PT_U"], f, N)
    bdf = 1.5 if (bdf_scheme == 2 and step > 0) else 1.0
    K = bdf * mats["mass"] + dt * (mats["stiffness"] + mats["convection"] + mats["trilinear"] + mats["nonlinear_lifting"])
    w = uN_n if bdf_scheme == 1 else 2.0 * uN_n - 0.5 * uN_n1
    return np.linalg.solve(K, mats["mass"].dot(w) + dt * fg)


def rom_sweep(model, mu, nt, dt, bdf2=True, solver="direct"):
    """RomConstructorNonlinear.solve for one parameter.  rom/rom.py:430-555 with
    assemble_system (:877-909) and assemble_system_rhs (:911-929).

    Zero initial condition (:451-452); first step uses bdf=1.0 even under BDF2 (:481-483)
    while uN_n1 is an array of zeros, not None (:456-460), so the BDF2 right-hand side
    formula (:923-925) and the extrapolation u* = 2u^n - u^{n-1} (:899-901, with
    uh_n1=None on the first step -> u* = uh :897-898) are reproduced literally.
    ``solver``: "direct" = np.linalg.solve (the hook create_algebraic_solver :414-425
    swapped for a dense solve), "gmres" = stock GMRES options (:36).
    Returns rom coefficients (N, nt) like RomSolutionsStorage.rom (:538).
    """
    N = model["N"]
    uN_n = np.zeros(N)
    uN_n1 = np.zeros(N) if bdf2 else None
    have_prev = False  # uh_n1 is None until the first history shuffle (:456, :499-502)
    t = 0.0
    out = np.zeros((N, nt))

    def op_matrix(op, t, ustar=None):
        f = sampled_entries(model, op, mu, t, ustar)
        return interpolate_rom_matrix(model[op]["basis_rom"], model[op]["PT_U"], f, N)

    for n in range(nt):
        t += dt
        bdf = 1.5 if (bdf2 and n > 0) else 1.0
        if have_prev:
            ustar = 2.0 * uN_n - uN_n1
        else:
            ustar = uN_n
        M = op_matrix("mass", t)
        A = op_matrix("stiffness", t)
        C = op_matrix("convection", t)
        Nn = op_matrix("trilinear", t, ustar)
        Nh = op_matrix("nonlinear_lifting", t)
        K = bdf * M + dt * (A + C + Nn + Nh)
        f = sampled_entries(model, "rhs", mu, t)
        fg = interpolate(model["rhs"]["basis_rom"], model["rhs"]["PT_U"], f)
        if uN_n1 is None:
            b = M.dot(uN_n) + dt * fg
        else:
            b = M.dot(2.0 * uN_n - 0.5 * uN_n1) + dt * fg
        if solver == "direct":
            uN = np.linalg.solve(K, b)
        else:
            uN, _ = scipy.sparse.linalg.gmres(K, b, rtol=1e-10, atol=1e-10, maxiter=1000000)
        out[:, n] = uN
        if bdf2:
            uN_n1 = uN_n.copy()
            have_prev = True
        uN_n = uN.copy()
    return out
