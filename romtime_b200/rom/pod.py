"""POD / ``orth`` on B200 -- drop-in for the reference's rom/pod.py:7-62.

Same signature, same truncation rules, same return types (NumPy arrays on the host).  The arithmetic is the method
of snapshots followed by a one-sided (Hestenes) refinement of the modes that are returned:

    G = S^T S                (TMA-fed DMMA tiles; column norms = sqrt(diag G), normalisation D^-1 G D^-1 fused)
    [multi-GPU: snapshot columns sharded, NCCL all-reduce of G -- romtime_b200/distributed.py]
    G = W diag(w) W^T        (pivoted Cholesky + block Jacobi),  s0 = sqrt(w)
    Y = S D^-1 W_c           (lift of the c = retained + guard modes; columns ~ sigma_k u_k)
    H = Y^T Y = X diag(lam) X^T   (plain Jacobi: H is the Gram matrix of NEARLY ORTHOGONAL graded columns, so its
                                   entries carry errors relative to |y_i||y_j| and Jacobi resolves every lam_k to a
                                   small RELATIVE error)
    s_k = sqrt(lam_k),  Q = Y X_r diag(1/s_r),  VT = (W_c X_r)^T

The first Gram matrix squares the condition number: on its own it resolves sigma_k only to eps sigma_1^2 / sigma_k
and a singular vector to eps (sigma_1 / sigma_k)^2.  The second, small Gram matrix is formed from vectors that are
already orthogonal to ~eps (sigma_1/sigma_k)^2 <= 1e-3, which is all Demmel-Veselic need: after it every retained
mode with sigma_k / sigma_1 >~ 1e-7 agrees with LAPACK gesvd (rom/pod.py:38) to the conditioning of the problem
(values: eps sigma_1/sigma_k; vectors: eps sigma_1/gap).  The guard band keeps the directions just below the
retained ones inside the rotation so that their leakage (eps (sigma_1/sigma_r)^2 sigma_j/sigma_r) stays below 1e-11.
Modes below the Gram route's numerical rank (sigma_k < 4e-8 sigma_1) are only produced by the deflated second
level (``_orth_level``: the same algorithm on S - Q Q^T S), which runs when the caller's truncation rule reaches
that far (``num`` beyond the numerical rank, or the default 1e-7 ABSOLUTE floor).  See DESIGN.md "POD accuracy".
"""
import numpy as np
import torch

from .. import kernels

DROP_TOLERANCE = 1e-7  # rom/pod.py:4
EPS = np.finfo(np.float64).eps
GUARD_TAU = 1e-11      # admissible leakage of an excluded direction into a retained mode
RANK_FLOOR = np.sqrt(8.0 * EPS)   # sigma_k / sigma_1 below which the first Gram matrix is noise (rt_sym_eig_f64's rank cut)
DELIVER_FLOOR = (EPS * RANK_FLOOR / GUARD_TAU) ** (1.0 / 3.0)   # ~1e-4: see _resolve_level


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


def _energy(s):
    lam = np.power(s, 2)
    total = np.sum(lam)
    return np.cumsum(lam) / total if total > 0 else np.ones_like(lam)


def _guard_width(s0, r, rank):
    """Candidate count c >= r: the modes j > r whose leakage into mode r, eps (s_1/s_r)^2 (s_j/s_r), exceeds
    GUARD_TAU are rotated together with the retained ones."""
    if r == 0 or r >= rank:
        return min(r, rank)
    amp = EPS * (s0[0] / s0[r - 1]) ** 2 / s0[r - 1]
    c = r
    while c < rank and amp * s0[c] > GUARD_TAU:
        c += 1
    return c


class _KernelOps:
    """The block operations of the refinement through libromtime_b200.so (the multi-GPU path and the CPU tests inject
    their own: row-sharded Gram + all-reduce, test stand-ins)."""

    @staticmethod
    def gram(Y):
        return kernels.pod_gram(Y, normalize=False)[0]

    @staticmethod
    def jacobi(H):
        lam, X, _, defect = kernels.sym_eig_jacobi(H)
        return lam, X, defect

    @staticmethod
    def lift(S, W, order="F"):
        return kernels.pod_lift(S, W, order=order)


def refine_modes(Y, Wc, n_keep, ops=None, max_passes=4):
    """One-sided Jacobi (Hestenes) refinement of lifted modes.  ``Y`` (Nh x c) holds columns ~ sigma_k u_k that are
    orthogonal up to the Gram route's error, ``Wc`` (n x c) the right vectors they were lifted with (or None).
    Rotates until the columns are orthogonal to working precision; returns (s_c descending numpy, Q (Nh x n_keep,
    orthonormal), VT (n_keep x n) or None, passes, first defect)."""
    ops = ops or _KernelOps
    c = Y.shape[1]
    first_defect = None
    for it in range(max_passes):
        H = ops.gram(Y)
        lam, X, defect = ops.jacobi(H)
        if first_defect is None:
            first_defect = defect
        # kappa of the scaled Gram matrix <= (1 + c mu) / (1 - c mu): one rotation from a 10 %-defect is exact to working precision
        if c * defect <= 0.1 or it == max_passes - 1:
            break
        Y = ops.lift(Y, X, order="F")
        if Wc is not None:
            Wc = ops.lift(Wc, X, order="C")
    sc = np.sqrt(np.maximum(lam.cpu().numpy(), 0.0))
    keep = min(n_keep, c)
    inv = np.divide(1.0, sc[:keep], out=np.zeros(keep), where=sc[:keep] > 0)
    Xr = (X[:, :keep] * torch.as_tensor(inv, device=Y.device)).contiguous()
    Q = ops.lift(Y, Xr, order="F")
    VT = None
    if Wc is not None:
        VT = ops.lift(Wc, X[:, :keep].contiguous(), order="C").t()
    return sc, Q, VT, it + 1, first_defect


def _resolve_level(A, normalize, allreduce, s_prefix, num, tol, k, Q_prev=None):
    """One level of the POD: Gram -> eigensolve -> lift of the (still wanted + guard) modes -> Hestenes refinement.
    ``s_prefix``: singular values already delivered by earlier levels (A is then the deflated remainder).
    Returns (Q_new, VT_new, s_tail (k - len(s_prefix) values, refined where rotated), cn, info)."""
    have = len(s_prefix)
    Nh, n = A.shape
    if allreduce is None:
        G, cn = kernels.pod_gram(A, normalize=normalize)
    else:
        G, cn = kernels.pod_gram(A, normalize=False)
        allreduce(G)
        cn = kernels.gram_finalize(G, normalize=normalize)
    w, W, _ = kernels.sym_eig(G)
    s_tail = np.sqrt(np.maximum(w[: k - have].cpu().numpy(), 0.0))
    rank = int(np.count_nonzero(s_tail > 0.0))
    info = {"rank_gram": rank, "candidates": 0, "passes": 0, "defect": 0.0}
    empty = (torch.empty((0, Nh), dtype=torch.float64, device=A.device).t(), W[:, :0].t())

    def wanted(tail):
        full = np.concatenate([s_prefix, tail])
        return int(np.count_nonzero(_truncation(full, _energy(full), num, tol))) - have

    want = min(max(wanted(s_tail), 0), rank)
    if want > 0 and have + rank < k:
        # directions this Gram matrix cannot see (below RANK_FLOOR * s_tail[0]) cannot be put in the guard band: only modes
        # they cannot contaminate (eps (s_1/s_r)^2 * floor / s_r <= GUARD_TAU, i.e. s_r >= ~1e-4 s_1) are delivered at
        # this level, the rest by the next one (whose Gram matrix starts eight digits lower)
        want = max(1, min(want, int(np.count_nonzero(s_tail >= DELIVER_FLOOR * s_tail[0]))))
    if want == 0:
        return empty[0], empty[1], s_tail, cn, info
    c = _guard_width(s_tail, want, rank)
    Wc = W[:, :c].contiguous()
    Wl = (Wc / cn[:, None]).contiguous() if normalize else Wc   # rows scaled by 1/||S_j||: Y = (S D^-1) W_c
    Y = kernels.pod_lift(A, Wl, order="F")
    if Q_prev is not None and Q_prev.shape[1] > 0:
        # the remainder was formed by a subtraction that leaves eps * sigma_1 of the delivered directions in it; relative
        # to this level's modes that is not small, so the lifted columns are projected once more ("twice is enough")
        Y = kernels.deflate(Y, None, Q_prev)
    sc, Q, VT, passes, defect = refine_modes(Y, Wc, want)
    s_tail = s_tail.copy()
    s_tail[:c] = sc
    info.update(candidates=c, passes=passes, defect=defect)
    keep = min(max(wanted(s_tail), 0), Q.shape[1])   # the rule re-applied to the refined spectrum
    return Q[:, :keep], VT[:keep], s_tail, cn, info


MAX_LEVELS = 4


def orth_device(S, num=None, tol=None, normalize=True, allreduce=None, deflate=True):
    """Device-resident POD. ``S`` is a CUDA tensor (Nh x n, any strides) or a NumPy array.
    ``allreduce`` (optional callable) is applied to the raw Gram matrix before finalisation --
    the hook used by the row-sharded multi-GPU path.
    Returns (Q cuda Nh x r F-order, s numpy, energy numpy, VT cuda r x n, info dict)."""
    Sd = kernels.to_device(S)
    Nh, n = Sd.shape
    k = min(Nh, n)
    default_rule = not tol and not num
    A, norm = Sd, bool(normalize)
    s = np.zeros(0)
    Q = VT = cn1 = None
    info = {"levels": 0}
    for level in range(1, MAX_LEVELS + 1):
        Qn, VTn, s_tail, cn, li = _resolve_level(A, norm, allreduce if level == 1 else None, s, num, tol, k, Q_prev=Q)
        have = len(s)
        s_all = np.concatenate([s, s_tail])
        Q = Qn if Q is None else torch.cat([Q, Qn], dim=1)
        VT = VTn if VT is None else torch.cat([VT, VTn], dim=0)
        info["levels"] = level
        if level == 1:
            info.update(li)
            cn1 = cn if normalize else None
        r_total = int(np.count_nonzero(_truncation(s_all, _energy(s_all), num, tol)))
        got = Q.shape[1]
        # what this level could not see: everything below RANK_FLOOR * (its largest singular value)
        floor_abs = RANK_FLOOR * (s_tail[0] if len(s_tail) else 0.0)
        unseen = have + li["rank_gram"] < k
        more = unseen and ((r_total > got) or (default_rule and floor_abs > DROP_TOLERANCE))
        if not (deflate and allreduce is None and more and level < MAX_LEVELS and got > 0 and (Qn.shape[1] > 0 or level == 1)):
            break
        # deflate: R = S D^-1 - Q (Q^T S D^-1) keeps what the delivered modes leave; its POD resolves the next ~8 digits
        A = kernels.deflate(Sd, cn1, Q)
        norm = False
        s = s_all[:got]
    s = s_all
    energy = _energy(s)
    r_total = int(np.count_nonzero(_truncation(s, energy, num, tol)))
    if r_total > Q.shape[1]:
        pad = r_total - Q.shape[1]            # beyond every level's reach (pure rounding noise): keep the reference's shape
        Q = torch.cat([Q, torch.zeros((Nh, pad), dtype=torch.float64, device=Sd.device)], dim=1)
        VT = torch.cat([VT, torch.zeros((pad, n), dtype=torch.float64, device=Sd.device)], dim=0)
    elif r_total < Q.shape[1]:
        Q, VT = Q[:, :r_total], VT[:r_total]
    return Q, s, energy, VT, info


_STAGING = [None]   # grow-only pinned staging buffer of the packed transfers (cudaHostAlloc costs more than the copy)


def _pinned(count):
    if _STAGING[0] is None or _STAGING[0].numel() < count:
        _STAGING[0] = torch.empty(max(count, 1 << 20), dtype=torch.float64, pin_memory=True)
    return _STAGING[0][:count]


def _to_device_packed(mats):
    """NumPy blocks -> CUDA tensors through ONE pinned staging buffer and ONE host-to-device copy (CUDA tensors pass through).
    Every block keeps C order; block starts are 16-byte aligned so that TMA can address them."""
    idx = [i for i, S in enumerate(mats) if not isinstance(S, torch.Tensor)]
    out = [S if isinstance(S, torch.Tensor) else None for S in mats]
    if not idx:
        return [kernels.to_device(S) for S in out]
    arrs = {i: np.asarray(mats[i], dtype=np.float64) for i in idx}
    offs, total = {}, 0
    for i in idx:
        offs[i] = total
        total += (arrs[i].size + 1) & ~1
    if not torch.cuda.is_available():
        kernels.to_device(arrs[idx[0]])   # raises: there is no CPU fallback
    host = _pinned(max(total, 2))
    hv = host.numpy()
    for i in idx:
        hv[offs[i]: offs[i] + arrs[i].size] = arrs[i].reshape(-1)       # C-order copy whatever the source order
    dev = host.to("cuda", non_blocking=True)
    torch.cuda.current_stream().synchronize()   # the staging buffer is reused by the next batch
    for i in idx:
        out[i] = dev[offs[i]: offs[i] + arrs[i].size].view(arrs[i].shape)
    return [kernels.to_device(S) for S in out]


def orth_device_batch(mats, num=None, tol=None, normalize=True):
    """The same POD for a BATCH of small snapshot matrices (the per-parameter time-PODs of a tree walk), with the host
    out of the inner loop: one packed host-to-device copy, the Gram matrices and lifts of all problems queued by one call
    each (rt_pod_gram_batch_f64 / rt_pod_lift_batch_f64), the eigen-work of the whole batch in two launches
    (rt_small_gram_eig_f64, rt_small_refine_f64: one CTA per problem), and the host reads spectra twice per BATCH instead
    of ~5 times per matrix.  Same algorithm and decisions as ``orth_device`` (Gram -> eigensolve -> lift of the wanted +
    guard modes -> Hestenes refinement); any problem outside the fast path's limits (more than 256 snapshots, numerical
    rank or candidate count above 96, a truncation rule that reaches below the first Gram matrix's floor, a refinement
    that needs a second pass) goes through ``orth_device`` unchanged.
    Returns a list of (Q cuda Nh x r, s numpy, energy numpy)."""
    P = len(mats)
    out = [None] * P
    Sd = _to_device_packed(list(mats))
    default_rule = not tol and not num
    fast = [i for i in range(P) if Sd[i].dim() == 2 and 1 <= Sd[i].shape[1] <= kernels.SMALL_NMAX and Sd[i].shape[0] >= 1]
    if len(fast) >= 2:
        dev = Sd[fast[0]].device
        f64 = torch.float64
        nmax = max(Sd[i].shape[1] for i in fast)
        G_all = torch.zeros((len(fast), nmax * nmax), dtype=f64, device=dev)
        cn_all = torch.empty((len(fast), nmax), dtype=f64, device=dev)
        kernels.pod_gram_batch([Sd[i] for i in fast], G_all, cn_all, nmax * nmax, nmax, normalize)
        n_of = torch.tensor([Sd[i].shape[1] for i in fast], dtype=torch.int32).to(dev)
        w_all, W_all, rank_d, info_d = kernels.small_gram_eig(G_all, n_of, nmax)
        w_h, rank_h, info_h = w_all.cpu().numpy(), rank_d.cpu().numpy(), info_d.cpu().numpy()   # host read 1 of 2
        plan = []  # [slot, problem, s_tail, rank, want, c]
        for j, i in enumerate(fast):
            Nh, n = Sd[i].shape
            k = min(Nh, n)
            if info_h[j] != 0:
                continue
            s_tail = np.sqrt(np.maximum(w_h[j, :k], 0.0))
            rank = int(np.count_nonzero(s_tail > 0.0))
            want = min(max(int(np.count_nonzero(_truncation(s_tail, _energy(s_tail), num, tol))), 0), rank)
            if want > 0 and rank < k:
                want = max(1, min(want, int(np.count_nonzero(s_tail >= DELIVER_FLOOR * s_tail[0]))))
            if want == 0:
                continue
            c = _guard_width(s_tail, want, rank)
            if c > kernels.SMALL_RMAX:
                continue
            plan.append([j, i, s_tail, rank, want, c])
        if plan:
            # lifts Y = S D^-1 W_c of the whole batch: the scaled right vectors packed in one buffer, one call
            cmax = max(pl[5] for pl in plan)
            Wl_all = torch.empty((len(plan), nmax * cmax), dtype=f64, device=dev)
            Ys, items = [], []
            for jj, (j, i, s_tail, rank, want, c) in enumerate(plan):
                Nh, n = Sd[i].shape
                W = W_all[j, : n * n].view(n, n)
                Wl = Wl_all[jj, : n * c].view(n, c)
                if normalize:
                    torch.div(W[:, :c], cn_all[j, :n, None], out=Wl)
                else:
                    Wl.copy_(W[:, :c])
                Y = torch.empty((c, Nh), dtype=f64, device=dev).t()         # F-ordered like pod_lift(order="F")
                Ys.append(Y)
                items.append((Sd[i], Wl, Y))
            kernels.pod_lift_batch(items)
            H_all = torch.zeros((len(plan), cmax * cmax), dtype=f64, device=dev)
            hn_all = torch.empty((len(plan), cmax), dtype=f64, device=dev)
            kernels.pod_gram_batch(Ys, H_all, hn_all, cmax * cmax, cmax, False)
            c_of = torch.tensor([pl[5] for pl in plan], dtype=torch.int32).to(dev)
            lam_all, X_all, defect_d, info2_d = kernels.small_refine(H_all, c_of, cmax)
            lam_h, defect_h, info2_h = lam_all.cpu().numpy(), defect_d.cpu().numpy(), info2_d.cpu().numpy()   # host read 2 of 2
            finals = []
            for jj, (j, i, s_tail, rank, want, c) in enumerate(plan):
                if info2_h[jj] != 0 or c * defect_h[jj] > 0.1:
                    continue   # needs a second rotation pass: general path
                Nh, n = Sd[i].shape
                k = min(Nh, n)
                sc = np.sqrt(np.maximum(lam_h[jj, :c], 0.0))
                s = s_tail.copy()
                s[:c] = sc
                energy = _energy(s)
                r_total = int(np.count_nonzero(_truncation(s, energy, num, tol)))
                floor_abs = RANK_FLOOR * s_tail[0]
                if rank < k and (r_total > want or (default_rule and floor_abs > DROP_TOLERANCE)):
                    continue   # the rule reaches below this Gram matrix's floor: deflated levels of the general path
                keep = min(r_total, want)
                if keep == 0:
                    out[i] = (torch.empty((0, Nh), dtype=f64, device=dev).t(), s, energy)
                    continue
                inv = np.divide(1.0, sc[:keep], out=np.zeros(keep), where=sc[:keep] > 0)
                finals.append((jj, i, keep, inv, s, energy))
            if finals:
                inv_all = torch.zeros((len(finals), cmax), dtype=f64)
                for f, (jj, i, keep, inv, s, energy) in enumerate(finals):
                    inv_all[f, :keep] = torch.from_numpy(inv)
                inv_all = inv_all.to(dev)
                Xr_all = torch.empty((len(finals), cmax * cmax), dtype=f64, device=dev)
                items = []
                for f, (jj, i, keep, inv, s, energy) in enumerate(finals):
                    c = plan[jj][5]
                    Xr = Xr_all[f, : c * keep].view(c, keep)
                    torch.mul(X_all[jj, : c * c].view(c, c)[:, :keep], inv_all[f, None, :keep], out=Xr)
                    Q = torch.empty((keep, Sd[i].shape[0]), dtype=f64, device=dev).t()
                    items.append((Ys[jj], Xr, Q))
                    out[i] = (Q, s, energy)
                kernels.pod_lift_batch(items)
    for i in range(P):
        if out[i] is None:
            Q, s, energy, _, _ = orth_device(Sd[i], num=num, tol=tol, normalize=normalize)
            out[i] = (Q, s, energy)
    return out


def orth_batch(snapshots_list, num=None, tol=None, normalize=True):
    """``[orth(S, num, tol, normalize) for S in snapshots_list]`` (reference rom/pod.py:7-62 per matrix) through the batched
    device path -- what the tree walks call for their per-parameter time-PODs.  Returns a list of (Q, s, energy) NumPy."""
    for S in snapshots_list:
        if isinstance(S, list):
            raise ValueError("You should use an array, not a list.")
    res = orth_device_batch(list(snapshots_list), num=num, tol=tol, normalize=(normalize == True))  # noqa: E712
    if not res:
        return []
    # ONE packed device-to-host copy of all bases (F-ordered blocks, like LAPACK's U)
    sizes = [Q.shape[0] * Q.shape[1] for Q, _, _ in res]
    dev = res[0][0].device
    pack = torch.empty(max(sum(sizes), 1), dtype=torch.float64, device=dev)
    off = 0
    for (Q, _, _), sz in zip(res, sizes):
        if sz:
            pack[off: off + sz].view(Q.shape[1], Q.shape[0]).copy_(Q.t())
        off += sz
    host = pack.cpu().numpy()
    out, off = [], 0
    for (Q, s, energy), sz in zip(res, sizes):
        out.append((host[off: off + sz].reshape(Q.shape[1], Q.shape[0]).T, s, energy))
        off += sz
    return out


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
