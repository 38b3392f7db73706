"""TEST INFRASTRUCTURE: NumPy stand-ins for the device block operations of romtime_b200.kernels that the POD host logic
(romtime_b200/rom/pod.py: guard band, Hestenes passes, deflation levels, truncation) is driven with on a box without a
GPU.  Each stand-in mirrors the NUMERICAL CONTRACT of the kernel it replaces, not its code: rt_sym_eig_f64 resolves
eigenvalues to eps * lambda_1 and cuts the rank at 8 eps (here: LAPACK eigh + the same cut); rt_sym_eig_jacobi_f64 is a
plain cyclic Jacobi with the relative rotation criterion (here: the same, scalar)."""
import numpy as np
import torch

EPS = np.finfo(float).eps


def _np(t):
    return t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def to_device(a, dtype=torch.float64):
    return a if isinstance(a, torch.Tensor) else torch.from_numpy(np.array(a, dtype=np.float64))


def pod_gram(S, normalize=True):
    A = _np(S)
    G = A.T @ A
    cn = np.sqrt(np.diag(G)).copy()
    if normalize:
        G = G / np.outer(cn, cn)
    return torch.from_numpy(G), torch.from_numpy(cn)


def gram_cross(A, B):
    return torch.from_numpy(_np(A).T @ _np(B))


def pod_lift(S, Wr, order="F"):
    out = _np(S) @ _np(Wr)
    return torch.from_numpy(np.asfortranarray(out).T).t() if order == "F" else torch.from_numpy(np.ascontiguousarray(out))


def sym_eig(G):
    w, W = np.linalg.eigh(_np(G))
    w, W = w[::-1].copy(), W[:, ::-1].copy()
    dead = ~(w > 8.0 * EPS * w[0])          # pivoted-Cholesky rank cut of rt_sym_eig_f64
    w[dead] = 0.0
    W[:, dead] = 0.0
    return torch.from_numpy(w), torch.from_numpy(W), 1


def jacobi(H, max_sweeps=40):
    A = np.array(H, dtype=float)
    n = A.shape[0]
    V = np.eye(n)
    for sweep in range(max_sweeps):
        rot = 0
        for p in range(n - 1):
            for q in range(p + 1, n):
                apq = A[p, q]
                if abs(apq) > EPS * np.sqrt(abs(A[p, p] * A[q, q])):
                    tau = (A[q, q] - A[p, p]) / (2.0 * apq)
                    t = (1.0 if tau >= 0 else -1.0) / (abs(tau) + np.sqrt(1.0 + tau * tau))
                    c = 1.0 / np.sqrt(1.0 + t * t)
                    s = t * c
                    J = np.array([[c, s], [-s, c]])
                    A[:, [p, q]] = A[:, [p, q]] @ J
                    A[[p, q], :] = J.T @ A[[p, q], :]
                    V[:, [p, q]] = V[:, [p, q]] @ J
                    A[p, q] = A[q, p] = 0.0
                    rot += 1
        if rot == 0:
            break
    w = np.diag(A).copy()
    order = np.argsort(-w, kind="stable")
    return w[order], V[:, order], sweep + 1


def sym_eig_jacobi(H):
    Hn = _np(H)
    d = np.diag(Hn)
    prod = np.outer(d, d)
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio = np.where(prod > 0, np.abs(Hn) / np.sqrt(np.where(prod > 0, prod, 1.0)), 0.0)
    np.fill_diagonal(ratio, 0.0)
    w, V, sweeps = jacobi(Hn)
    return torch.from_numpy(w), torch.from_numpy(V), sweeps, float(ratio.max()) if Hn.shape[0] > 1 else 0.0


def deflate(S, cn, Q):
    A = _np(S)
    if cn is not None:
        A = A / _np(cn)
    Qn = _np(Q)
    return torch.from_numpy(np.ascontiguousarray(A - Qn @ (Qn.T @ A)))


def gram_finalize(G, normalize=True):
    Gn = _np(G)
    cn = np.sqrt(np.diag(Gn)).copy()
    if normalize:
        Gn /= np.outer(cn, cn)
    return torch.from_numpy(cn)


def install(monkeypatch):
    """Route romtime_b200.rom.pod's block operations to the stand-ins above."""
    import sys

    from romtime_b200.rom import pod

    monkeypatch.setattr(pod, "kernels", sys.modules[__name__])
