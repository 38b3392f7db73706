"""POD / ``orth`` on B200 -- drop-in for the reference's rom/pod.py:7-62.

Same signature, same truncation rules, same return types (NumPy arrays on the host); the
arithmetic is the method of snapshots on the device:

    G = S^T S (TMA-fed DMMA tiles, column norms = sqrt(diag G), normalisation D^-1 G D^-1 fused)
    [multi-GPU: snapshot columns sharded, NCCL all-reduce of G -- see romtime_b200/distributed.py]
    G = W diag(w) W^T  (block Jacobi),  s = sqrt(w),  VT = W^T
    Q = S D^-1 W_r diag(1/s_r)          (lift GEMM)

The Gram route squares the condition number: singular values below ~1e-8 * s[0] are noise (the
reference's gesvd resolves them to ~1e-16 * s[0]).  Retained modes (s_k / s_0 >~ 1e-3) agree with
gesvd to 1e-10 relative; see DESIGN.md "POD accuracy".
"""
import numpy as np
import torch

from .. import kernels

DROP_TOLERANCE = 1e-7  # rom/pod.py:4


def _truncation(s, energy, num, tol):
    """Number of retained modes, following rom/pod.py:45-57 literally (``tol`` beats ``num``
    beats the 1e-7 floor; the energy mask is strict)."""
    if tol:
        mask = energy < tol
    elif num:
        mask = np.zeros(len(s), dtype=bool)
        mask[:num] = True
    else:
        mask = s > DROP_TOLERANCE
    return mask


def orth_device(S, num=None, tol=None, normalize=True, allreduce=None):
    """Device-resident POD. ``S`` is a CUDA tensor (Nh x n, any strides) or a NumPy array.
    ``allreduce`` (optional callable) is applied to the raw Gram matrix before finalisation --
    the hook used by the column/row-sharded multi-GPU path.
    Returns (Q cuda Nh x r F-order, s numpy, energy numpy, VT cuda r x n, timings dict)."""
    Sd = kernels.to_device(S)
    Nh, n = Sd.shape
    if allreduce is None:
        G, cn = kernels.pod_gram(Sd, normalize=normalize)
    else:
        G, cn = kernels.pod_gram(Sd, normalize=False)
        allreduce(G)
        cn = kernels.gram_finalize(G, normalize=normalize)
    w, W, _ = kernels.sym_eig(G)
    k = min(Nh, n)
    w_host = w[:k].cpu().numpy()
    s = np.sqrt(np.maximum(w_host, 0.0))
    eigenvalues = np.power(s, 2)
    energy = np.cumsum(eigenvalues) / np.sum(eigenvalues)
    mask = _truncation(s, energy, num, tol)
    keep = np.flatnonzero(mask)
    r = len(keep)
    if r == 0:
        Q = torch.empty((0, Nh), dtype=torch.float64, device=Sd.device).t()
        return Q, s, energy, W[:, :0].t(), {}
    idx = torch.as_tensor(keep, device=Sd.device)
    Wk = W.index_select(1, idx)                      # n x r
    sk = s[keep]
    # modes below the rank-revealing tolerance come back as (0, zero vector): keep them zero
    scale = torch.as_tensor(np.divide(1.0, sk, out=np.zeros_like(sk), where=sk > 0), device=Sd.device)
    Wr = Wk * scale                                   # columns scaled by 1/sigma
    if normalize:
        Wr = Wr / cn[:, None]                         # rows scaled by 1/||S_j||
    Q = kernels.pod_lift(Sd, Wr.contiguous(), order="F")
    return Q, s, energy, Wk.t(), {}


def orth(snapshots, num=None, tol=None, normalize=True, return_VT=False):
    """Compute orthogonalization via SVD (reference rom/pod.py:7-62).

    Parameters
    ----------
    snapshots : np.array
    num : int, optional
        Number of modes to select.

    Returns
    -------
    Q : np.array
        Basis vectors.
    s : np.array
        Singular values (for all basis vectors).
    """
    if isinstance(snapshots, list):
        raise ValueError("You should use an array, not a list.")
    Q, s, energy, VT, _ = orth_device(snapshots, num=num, tol=tol, normalize=(normalize == True))  # noqa: E712
    Qh = Q.cpu().numpy()
    if return_VT:
        return Qh, s, energy, VT.cpu().numpy()
    return Qh, s, energy
