"""Thin functional wrappers over the C ABI (include/romtime_b200.h).

Inputs may be NumPy arrays (copied to the current CUDA device through pinned staging by torch)
or CUDA torch tensors (used in place, any 2-D strides).  Outputs are CUDA torch tensors; the
reference-facing classes in romtime_b200/rom and romtime_b200/deim convert them to NumPy.
Nothing here computes on the CPU: without the shared object or a GPU every call raises.
"""
import ctypes

import numpy as np
import torch

from . import _lib
from ._lib import c_vp, check, handle, load_library, stream_ptr

F64 = torch.float64


def to_device(a, dtype=F64):
    """NumPy / torch (any device) -> CUDA tensor, strides preserved (C- and F-order both kept)."""
    if not torch.cuda.is_available():
        raise _lib.RomtimeB200Error("romtime_b200 needs a B200 (sm_100a) GPU: there is no CPU fallback")
    if isinstance(a, torch.Tensor):
        t = a
    else:
        arr = np.asarray(a)
        if arr.dtype != (np.float64 if dtype == F64 else np.int64):
            arr = arr.astype(np.float64 if dtype == F64 else np.int64)
        if not arr.flags.writeable:
            arr = arr.copy(order="K")
        t = torch.from_numpy(arr)
    if t.dtype != dtype:
        t = t.to(dtype)
    if not t.is_cuda:
        t = t.to("cuda", non_blocking=False)
    return t


def _ptr(t):
    return c_vp(t.data_ptr()) if t is not None else c_vp(0)


def _strides2(t):
    if t.dim() == 1:
        return int(t.stride(0)), int(t.shape[0] * t.stride(0))
    return int(t.stride(0)), int(t.stride(1))


def fp64_peak_tflops():
    out = ctypes.c_double(0.0)
    h = handle()
    check(load_library().rt_fp64_peak_tflops(h, ctypes.byref(out)), h, "rt_fp64_peak_tflops")
    return out.value


# ------------------------------------------------------------------ DEIM greedy
def deim_greedy(U, sync=True):
    """rt_deim_greedy_f64: returns (dofs int64[m], PT_U float64[m, m]) as CUDA tensors.

    Raises numpy.linalg.LinAlgError when the interpolation matrix becomes exactly singular
    (the reference raises it from np.linalg.solve, deim/deim.py:547)."""
    Ud = to_device(U)
    if Ud.dim() != 2:
        raise ValueError("basis must be a 2-D array (Nh x m)")
    Nh, m = Ud.shape
    if m > 160:
        raise ValueError(f"collateral basis size m = {m} exceeds the supported maximum 160 (include/romtime_b200.h, Limits)")
    dofs = torch.empty(m, dtype=torch.int64, device=Ud.device)
    PT_U = torch.empty((m, m), dtype=F64, device=Ud.device)
    info = torch.zeros(1, dtype=torch.int32, device=Ud.device)
    h = handle(Ud.device.index)
    with torch.cuda.device(Ud.device):
        rc = load_library().rt_deim_greedy_f64(h, _ptr(Ud), Nh, m, Ud.stride(0), Ud.stride(1),
                                               _ptr(dofs), _ptr(PT_U), _ptr(info), stream_ptr())
    check(rc, h, "rt_deim_greedy_f64")
    if sync:
        step = int(info.item())
        if step != 0:
            raise np.linalg.LinAlgError(f"Singular matrix (interpolation matrix singular after {step} points)")
    return dofs, PT_U


# ------------------------------------------------------------------ POD building blocks
def pod_gram(S, normalize=True):
    """rt_pod_gram_f64: (G n x n, colnorm n) for a snapshot matrix S (Nh x n), any 2-D strides."""
    Sd = to_device(S)
    Nh, n = Sd.shape
    G = torch.empty((n, n), dtype=F64, device=Sd.device)
    cn = torch.empty(n, dtype=F64, device=Sd.device)
    h = handle(Sd.device.index)
    with torch.cuda.device(Sd.device):
        rc = load_library().rt_pod_gram_f64(h, _ptr(Sd), Nh, n, Sd.stride(0), Sd.stride(1), int(bool(normalize)),
                                            _ptr(G), _ptr(cn), stream_ptr())
    check(rc, h, "rt_pod_gram_f64")
    return G, cn


def enable_peer_access(peer, device=None):
    """rt_enable_peer_access: kernels of ``device`` (default: current) may read buffers living on device ``peer``."""
    h = handle(device)
    check(load_library().rt_enable_peer_access(h, int(peer)), h, "rt_enable_peer_access")


def gram_cross(A, B):
    """rt_gram_cross_f64: C = A^T B (na x nb) for A (Nh x na), B (Nh x nb).  The call runs on A's device; B may be peer
    memory (a CUDA tensor of another device after enable_peer_access)."""
    Ad, Bd = to_device(A), to_device(B)
    Nh, na = Ad.shape
    Nh2, nb = Bd.shape
    if Nh != Nh2:
        raise ValueError("row counts differ")
    C = torch.empty((na, nb), dtype=F64, device=Ad.device)
    h = handle(Ad.device.index)
    with torch.cuda.device(Ad.device):
        rc = load_library().rt_gram_cross_f64(h, _ptr(Ad), na, Ad.stride(0), Ad.stride(1), _ptr(Bd), nb,
                                              Bd.stride(0), Bd.stride(1), Nh, _ptr(C), nb, stream_ptr())
    check(rc, h, "rt_gram_cross_f64")
    return C


def gram_finalize(G, normalize=True):
    n = G.shape[0]
    cn = torch.empty(n, dtype=F64, device=G.device)
    h = handle(G.device.index)
    with torch.cuda.device(G.device):
        rc = load_library().rt_gram_finalize_f64(h, _ptr(G), n, int(bool(normalize)), _ptr(cn), stream_ptr())
    check(rc, h, "rt_gram_finalize_f64")
    return cn


def pod_lift(S, Wr, order="F"):
    """rt_pod_lift_f64: Q = S @ Wr (Nh x r); Q is returned F-ordered (LAPACK-like) or C-ordered."""
    Sd, Wd = to_device(S), to_device(Wr)
    if not Wd.is_contiguous():
        Wd = Wd.contiguous()
    Nh, n = Sd.shape
    r = Wd.shape[1]
    if order == "F":
        Q = torch.empty((r, Nh), dtype=F64, device=Sd.device).t()
    else:
        Q = torch.empty((Nh, r), dtype=F64, device=Sd.device)
    h = handle(Sd.device.index)
    with torch.cuda.device(Sd.device):
        rc = load_library().rt_pod_lift_f64(h, _ptr(Sd), Nh, n, Sd.stride(0), Sd.stride(1), _ptr(Wd), Wd.stride(0), r,
                                            _ptr(Q), Q.stride(0), Q.stride(1), stream_ptr())
    check(rc, h, "rt_pod_lift_f64")
    return Q


def sym_eig(G):
    """rt_sym_eig_f64: (w descending, W with eigenvector k in column k, sweeps). G is destroyed."""
    n = G.shape[0]
    if not (G.is_contiguous() and G.dtype == F64):
        G = G.contiguous().to(F64)
    w = torch.empty(n, dtype=F64, device=G.device)
    W = torch.empty((n, n), dtype=F64, device=G.device)
    info = torch.zeros(1, dtype=torch.int32, device=G.device)
    h = handle(G.device.index)
    with torch.cuda.device(G.device):
        rc = load_library().rt_sym_eig_f64(h, _ptr(G), n, _ptr(w), _ptr(W), _ptr(info), stream_ptr())
    check(rc, h, "rt_sym_eig_f64")
    return w, W, int(info.item())


def deflate(S, cn, Q):
    """R = S D^-1 - Q (Q^T S D^-1)  (Nh x n, C-order): the snapshots with the delivered modes Q projected out
    (rt_gram_cross_f64, rt_pod_lift_f64, rt_deflate_f64).  ``cn``: column norms (None: D = I)."""
    Sd = to_device(S)
    Nh, n = Sd.shape
    T = gram_cross(Q, Sd)                       # r x n
    if cn is not None:
        T = T / cn[None, :]
    P = pod_lift(Q, T.contiguous(), order="C")   # Nh x n
    h = handle(Sd.device.index)
    with torch.cuda.device(Sd.device):
        rc = load_library().rt_deflate_f64(h, _ptr(Sd), Nh, n, Sd.stride(0), Sd.stride(1), _ptr(cn), _ptr(P), stream_ptr())
    check(rc, h, "rt_deflate_f64")
    return P


def sym_eig_jacobi(H):
    """rt_sym_eig_jacobi_f64: plain block Jacobi on the Gram matrix of nearly orthogonal columns.
    Returns (w descending, W, sweeps, defect) -- defect = max |h_ij| / sqrt(h_ii h_jj) of the input."""
    n = H.shape[0]
    if not (H.is_contiguous() and H.dtype == F64):
        H = H.contiguous().to(F64)
    w = torch.empty(n, dtype=F64, device=H.device)
    W = torch.empty((n, n), dtype=F64, device=H.device)
    info = torch.zeros(1, dtype=torch.int32, device=H.device)
    defect = torch.zeros(1, dtype=F64, device=H.device)
    h = handle(H.device.index)
    with torch.cuda.device(H.device):
        rc = load_library().rt_sym_eig_jacobi_f64(h, _ptr(H), n, _ptr(w), _ptr(W), _ptr(info), _ptr(defect), stream_ptr())
    check(rc, h, "rt_sym_eig_jacobi_f64")
    return w, W, int(info.item()), float(defect.item())


SMALL_NMAX, SMALL_RMAX = 256, 96  # limits of the batched small-POD kernels (include/romtime_b200.h)


GRAM_DESC = np.dtype([("S", "u8"), ("Nh", "i8"), ("n", "i8"), ("srow", "i8"), ("scol", "i8"), ("G", "u8"), ("cn", "u8")])
LIFT_DESC = np.dtype([("S", "u8"), ("Nh", "i8"), ("n", "i8"), ("srow", "i8"), ("scol", "i8"), ("W", "u8"), ("ldw", "i8"), ("r", "i8"),
                      ("Q", "u8"), ("qrow", "i8"), ("qcol", "i8")])


def pod_gram_batch(mats, G_all, cn_all, stride_g, stride_cn, normalize):
    """rt_pod_gram_batch_f64: Gram matrix of every CUDA matrix in ``mats`` into slot j of G_all (leading dimension n_j) and
    its column norms into slot j of cn_all -- one call, nothing synchronised."""
    d = np.zeros(len(mats), dtype=GRAM_DESC)
    for j, S in enumerate(mats):
        d[j] = (S.data_ptr(), S.shape[0], S.shape[1], S.stride(0), S.stride(1), G_all.data_ptr() + 8 * j * stride_g,
                cn_all.data_ptr() + 8 * j * stride_cn)
    dev = G_all.device
    h = handle(dev.index)
    with torch.cuda.device(dev):
        rc = load_library().rt_pod_gram_batch_f64(h, c_vp(d.ctypes.data), len(mats), int(bool(normalize)), stream_ptr())
    check(rc, h, "rt_pod_gram_batch_f64")


def pod_lift_batch(items):
    """rt_pod_lift_batch_f64: Q_j = S_j @ W_j for every (S_j, W_j contiguous, Q_j) of CUDA tensors -- one call."""
    d = np.zeros(len(items), dtype=LIFT_DESC)
    for j, (S, W, Q) in enumerate(items):
        d[j] = (S.data_ptr(), S.shape[0], S.shape[1], S.stride(0), S.stride(1), W.data_ptr(), W.stride(0), W.shape[1],
                Q.data_ptr(), Q.stride(0), Q.stride(1))
    dev = items[0][0].device
    h = handle(dev.index)
    with torch.cuda.device(dev):
        rc = load_library().rt_pod_lift_batch_f64(h, c_vp(d.ctypes.data), len(items), stream_ptr())
    check(rc, h, "rt_pod_lift_batch_f64")


def small_gram_eig(G_all, n_of, nmax):
    """rt_small_gram_eig_f64 for a batch of Gram matrices packed in G_all (P x nmax*nmax).  Returns device tensors
    (w_all P x nmax, W_all P x nmax*nmax, rank P, info P); nothing is synchronised."""
    P = G_all.shape[0]
    dev = G_all.device
    w_all = torch.empty((P, nmax), dtype=F64, device=dev)
    W_all = torch.empty((P, nmax * nmax), dtype=F64, device=dev)
    rank = torch.empty(P, dtype=torch.int32, device=dev)
    info = torch.empty(P, dtype=torch.int32, device=dev)
    h = handle(dev.index)
    with torch.cuda.device(dev):
        rc = load_library().rt_small_gram_eig_f64(h, _ptr(G_all), nmax * nmax, _ptr(n_of), P, nmax, _ptr(w_all), _ptr(W_all),
                                                  nmax * nmax, _ptr(rank), _ptr(info), stream_ptr())
    check(rc, h, "rt_small_gram_eig_f64")
    return w_all, W_all, rank, info


def small_refine(H_all, c_of, cmax):
    """rt_small_refine_f64 for a batch of refinement Gram matrices packed in H_all (P x cmax*cmax).  Returns device
    tensors (lam_all P x cmax, X_all P x cmax*cmax, defect P, info P); nothing is synchronised."""
    P = H_all.shape[0]
    dev = H_all.device
    lam_all = torch.zeros((P, cmax), dtype=F64, device=dev)
    X_all = torch.zeros((P, cmax * cmax), dtype=F64, device=dev)
    defect = torch.zeros(P, dtype=F64, device=dev)
    info = torch.empty(P, dtype=torch.int32, device=dev)
    h = handle(dev.index)
    with torch.cuda.device(dev):
        rc = load_library().rt_small_refine_f64(h, _ptr(H_all), cmax * cmax, _ptr(c_of), P, cmax, _ptr(lam_all), _ptr(X_all),
                                                cmax * cmax, _ptr(defect), _ptr(info), stream_ptr())
    check(rc, h, "rt_small_refine_f64")
    return lam_all, X_all, defect, info


# ------------------------------------------------------------------ projections onto the reduced basis
def project_vec(V, U):
    """rt_project_vec_f64: DEIM.project_basis, out (N x m) = V^T U."""
    Vd, Ud = to_device(V), to_device(U)
    Nh, N = Vd.shape
    Nh2, m = Ud.shape
    if Nh != Nh2:
        raise ValueError(f"V has {Nh} rows, the collateral basis has {Nh2}")
    out = torch.empty((N, m), dtype=F64, device=Vd.device)
    h = handle(Vd.device.index)
    with torch.cuda.device(Vd.device):
        rc = load_library().rt_project_vec_f64(h, _ptr(Vd), Vd.stride(0), Vd.stride(1), _ptr(Ud), Ud.stride(0),
                                               Ud.stride(1), Nh, N, m, _ptr(out), stream_ptr())
    check(rc, h, "rt_project_vec_f64")
    return out


def project_mat(U, rows, cols, V):
    """rt_project_mat_f64: MDEIM.project_basis, out (N^2 x m, F-order: column q contiguous) with
    out[i*N + j, q] = (V^T csr(U[:, q]; rows, cols) V)[i, j]."""
    Ud, Vd = to_device(U), to_device(V)
    nnz, m = Ud.shape
    Nh, N = Vd.shape
    rows_h = np.asarray(rows, dtype=np.int64).reshape(-1)
    cols_h = np.asarray(cols, dtype=np.int64).reshape(-1)
    if rows_h.shape[0] != nnz or cols_h.shape[0] != nnz:
        raise ValueError("rows/cols must have one entry per row of the collateral basis")
    if nnz and (rows_h.min() < 0 or cols_h.min() < 0 or rows_h.max() >= Nh or cols_h.max() >= Nh):
        raise ValueError("matrix topology does not fit the reduced basis (row/column index out of range)")
    if nnz == 0 or m == 0:  # an operator without non-zeros (e.g. the trilinear term around u* = 0) projects to zero
        return torch.zeros((m, N * N), dtype=F64, device=Vd.device).t()
    rd, cd = to_device(rows_h, torch.int64), to_device(cols_h, torch.int64)
    out = torch.empty((m, N * N), dtype=F64, device=Ud.device).t()
    h = handle(Ud.device.index)
    with torch.cuda.device(Ud.device):
        rc = load_library().rt_project_mat_f64(h, _ptr(Ud), Ud.stride(0), Ud.stride(1), _ptr(rd), _ptr(cd), nnz, _ptr(Vd),
                                               Vd.stride(0), Vd.stride(1), Nh, N, m, _ptr(out), out.stride(0),
                                               out.stride(1), stream_ptr())
    check(rc, h, "rt_project_mat_f64")
    return out


# ------------------------------------------------------------------ theta / assembly / online sweep
def invert(A):
    """rt_invert_f64: inverse of a small dense matrix (m <= 160) by Gauss-Jordan with partial pivoting."""
    Ad = to_device(A).contiguous()
    m = Ad.shape[0]
    if Ad.dim() != 2 or Ad.shape[1] != m or not 1 <= m <= DeviceRomModel.MAX_M:
        raise ValueError(f"interpolation matrix must be square with 1 <= m <= {DeviceRomModel.MAX_M}, got {tuple(Ad.shape)}")
    Ainv = torch.empty_like(Ad)
    info = torch.zeros(1, dtype=torch.int32, device=Ad.device)
    h = handle(Ad.device.index)
    with torch.cuda.device(Ad.device):
        rc = load_library().rt_invert_f64(h, _ptr(Ad), m, _ptr(Ainv), _ptr(info), stream_ptr())
    check(rc, h, "rt_invert_f64")
    if int(info.item()) != 0:
        raise np.linalg.LinAlgError("Singular matrix")
    return Ainv


def theta_batch(PT_Uinv, F):
    """rt_theta_f64: theta[b] = PT_U^-1 F[b] for a batch F (B x m)."""
    Fd = to_device(F).contiguous()
    B, m = Fd.shape
    out = torch.empty_like(Fd)
    h = handle(Fd.device.index)
    with torch.cuda.device(Fd.device):
        rc = load_library().rt_theta_f64(h, _ptr(PT_Uinv), m, _ptr(Fd), B, _ptr(out), stream_ptr())
    check(rc, h, "rt_theta_f64")
    return out


def assemble_batch(basis, theta):
    """rt_assemble_batch_f64: out[b] = sum_q theta[b, q] basis[:, q]  (B x R)."""
    Bd, Td = to_device(basis), to_device(theta).contiguous()
    R, m = Bd.shape
    B = Td.shape[0]
    out = torch.empty((B, R), dtype=F64, device=Bd.device)
    h = handle(Bd.device.index)
    with torch.cuda.device(Bd.device):
        rc = load_library().rt_assemble_batch_f64(h, _ptr(Bd), R, m, Bd.stride(0), Bd.stride(1), _ptr(Td), B, _ptr(out),
                                                  stream_ptr())
    check(rc, h, "rt_assemble_batch_f64")
    return out


OPS = ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear", "rhs")


class DeviceRomModel:
    """Device-resident hyper-reduced model (rt_rom_model).  ``model`` maps every operator name in ``OPS``
    to ``{"basis_rom", "PT_U"[, "C", "T"]}`` or to ``None`` / nothing (operator absent, m = 0):
    basis_rom (column q contiguous), PT_U^-1 (computed once with rt_invert_f64) and, for the device
    evaluator of rt_online_sweep_f64, the tables C (m x 12) and T (m x N)."""

    _generation = 0
    MAX_N = 128   # include/romtime_b200.h "Limits"
    MAX_M = 160

    def __init__(self, model, device=None):
        self.N = int(model["N"])
        if not 1 <= self.N <= self.MAX_N:
            raise ValueError(f"reduced dimension N = {self.N} is outside the supported range 1..{self.MAX_N} "
                             "(the batched dense solve keeps one system per CTA in shared memory)")
        # unique id of this set of device tables: lets the library keep the staged, folded operators between calls
        DeviceRomModel._generation += 1
        # device feature set of the sampled-entry evaluator (include/romtime_b200.h): piston rest / sudden, or MFP1
        self.evaluator = int(model["evaluator_id"]) if model.get("evaluator_id") is not None else (0 if model.get("which", "rest") == "rest" else 1)
        self.n_mu = 4 if self.evaluator == 2 else 3
        self._keep = []
        self.m = []
        self.has_evaluator = True
        self.struct = _lib.RomModel()
        self.struct.N = self.N
        self.struct.n_ops = len(OPS)
        self.struct.evaluator = self.evaluator
        self.struct.n_mu = self.n_mu
        self.struct.generation = DeviceRomModel._generation
        for i, name in enumerate(OPS):
            op = model.get(name)
            so = self.struct.ops[i]
            so.is_vector = 1 if name == "rhs" else 0
            if op is None:
                so.m = 0
                self.m.append(0)
                continue
            basis = np.asfortranarray(np.asarray(op["basis_rom"], dtype=np.float64))  # column q contiguous
            rows_expected = self.N if name == "rhs" else self.N * self.N
            if basis.shape[0] != rows_expected:
                raise ValueError(f"{name}: basis_rom has {basis.shape[0]} rows, expected {rows_expected}")
            if basis.shape[1] > self.MAX_M:
                raise ValueError(f"{name}: collateral basis size m = {basis.shape[1]} exceeds the supported maximum "
                                 f"{self.MAX_M} (interpolation matrix inverse in shared memory)")
            bd = to_device(basis)
            inv = invert(np.ascontiguousarray(op["PT_U"], dtype=np.float64))
            Cd = Td = None
            if op.get("C") is not None:
                Cd = to_device(np.ascontiguousarray(op["C"], dtype=np.float64))
            else:
                self.has_evaluator = False
            if op.get("T") is not None:
                Td = to_device(np.ascontiguousarray(op["T"], dtype=np.float64))
            self._keep += [bd, inv, Cd, Td]
            so.m = basis.shape[1]
            self.m.append(int(basis.shape[1]))
            so.basis_rom = bd.data_ptr()
            so.ld = int(bd.stride(1)) if basis.shape[1] > 1 else int(basis.shape[0])
            so.PT_Uinv = inv.data_ptr()
            so.C = Cd.data_ptr() if Cd is not None else None
            so.T = Td.data_ptr() if Td is not None else None
        self.qsum = sum(self.m)


def online_step(dev_model, F, Un, Unm1, info, dt, bdf, step, Kadd=None, Madd=None, fadd=None):
    """rt_online_step_f64: one implicit step for B parameters from caller-supplied sampled entries
    ``F`` (B x sum(m), CUDA or NumPy).  ``Un``/``Unm1`` (B x N CUDA tensors) are updated in place."""
    Fd = to_device(F).contiguous()
    if Fd.dim() == 1:
        Fd = Fd.reshape(1, -1)
    B = Fd.shape[0]
    if Fd.shape[1] != dev_model.qsum:
        raise ValueError(f"sampled entries have {Fd.shape[1]} columns, the model needs {dev_model.qsum}")
    extras = []
    for x, shape in ((Kadd, (B, dev_model.N, dev_model.N)), (Madd, (B, dev_model.N, dev_model.N)), (fadd, (B, dev_model.N))):
        if x is None:
            extras.append(None)
        else:
            xd = to_device(x).contiguous()
            if tuple(xd.shape) != shape:
                raise ValueError(f"dense additive term has shape {tuple(xd.shape)}, expected {shape}")
            extras.append(xd)
    h = handle(Fd.device.index)
    with torch.cuda.device(Fd.device):
        rc = load_library().rt_online_step_f64(h, ctypes.byref(dev_model.struct), _ptr(Fd), B, float(dt), int(bdf), int(step),
                                               _ptr(extras[0]), _ptr(extras[1]), _ptr(extras[2]), _ptr(Un), _ptr(Unm1),
                                               _ptr(info), stream_ptr())
    check(rc, h, "rt_online_step_f64")


def _mus_on_device(dev_model, mus):
    md = to_device(mus).contiguous()
    if md.dim() != 2 or md.shape[1] != dev_model.n_mu:
        raise ValueError(f"mus must be (B, {dev_model.n_mu}) for this model's evaluator, got {tuple(md.shape)}")
    return md


def online_sweep(dev_model, mus, nt, dt, bdf=2, history=True, out=None):
    """rt_online_sweep_f64 / rt_online_sweep_last_f64.  mus: (B, 3) [a0, omega, delta].  ``out``: optional preallocated
    contiguous CUDA result tensor.  Returns (uN cuda tensor (B, nt, N) or (B, N), info int32 (B,))."""
    if not dev_model.has_evaluator:
        raise _lib.RomtimeB200Error("this model has no device evaluator tables (C): use online_step with host-assembled entries")
    md = _mus_on_device(dev_model, mus)
    B = md.shape[0]
    N = dev_model.N
    info = torch.empty(B, dtype=torch.int32, device=md.device)
    h = handle(md.device.index)
    lib = load_library()
    with torch.cuda.device(md.device):
        shape = (B, nt, N) if history else (B, N)
        if out is not None and not (out.is_cuda and out.is_contiguous() and out.dtype == F64 and tuple(out.shape) == shape):
            raise ValueError(f"out must be a contiguous float64 CUDA tensor of shape {shape}")
        if history:
            out = torch.empty(shape, dtype=F64, device=md.device) if out is None else out
            rc = lib.rt_online_sweep_f64(h, ctypes.byref(dev_model.struct), _ptr(md), B, nt, float(dt), int(bdf), _ptr(out),
                                         _ptr(info), stream_ptr())
        else:
            out = torch.empty(shape, dtype=F64, device=md.device) if out is None else out
            rc = lib.rt_online_sweep_last_f64(h, ctypes.byref(dev_model.struct), _ptr(md), B, nt, float(dt), int(bdf),
                                              _ptr(out), _ptr(info), stream_ptr())
    check(rc, h, "rt_online_sweep_f64")
    return out, info


def online_sweep_segments(dev_model, mus, nt, dt, bdf=2, segment=100, out=None, on_segment=None):
    """rt_online_sweep_segment_f64: the nt-step sweep in segments of ``segment`` steps with the state carried between
    them (bit-identical to the uninterrupted sweep).  ``on_segment(n0, n1)`` is called after segment [n0, n1) has been
    enqueued on the current stream.  Returns (uN cuda tensor (B, nt, N), info int32 (B,))."""
    if not dev_model.has_evaluator:
        raise _lib.RomtimeB200Error("this model has no device evaluator tables (C): use online_step with host-assembled entries")
    md = _mus_on_device(dev_model, mus)
    B, N = md.shape[0], dev_model.N
    if out is None:
        out = torch.empty((B, nt, N), dtype=F64, device=md.device)
    info = torch.empty(B, dtype=torch.int32, device=md.device)
    Un = torch.empty((B, N), dtype=F64, device=md.device)
    Unm1 = torch.empty((B, N), dtype=F64, device=md.device)
    h = handle(md.device.index)
    lib = load_library()
    with torch.cuda.device(md.device):
        for n0 in range(0, nt, segment):
            n1 = min(nt, n0 + segment)
            rc = lib.rt_online_sweep_segment_f64(h, ctypes.byref(dev_model.struct), _ptr(md), B, nt, n0, n1, float(dt), int(bdf),
                                                 _ptr(Un), _ptr(Unm1), _ptr(out), _ptr(info), stream_ptr())
            check(rc, h, "rt_online_sweep_segment_f64")
            if on_segment is not None:
                on_segment(n0, n1)
    return out, info


def online_sweep_to_host(dev_model, mus, nt, dt, bdf=2, out_host=None, device=None, segment=100):
    """The sweep with the history delivered in (pinned) HOST memory.  The sweep runs in segments of ``segment`` time steps
    (rt_online_sweep_segment_f64, state carried on the device); the finished time slab [:, n0:n1, :] of the device
    history goes to the host as one strided copy (rt_memcpy2d_async) on a side stream while the next segment is
    computed, so only the last slab's copy is exposed -- whatever the number of parameters per GPU.
    ``mus``: (B, 3) host array / tensor (pinned for an asynchronous upload).
    Returns (uN host tensor (B, nt, N), info int32 cuda tensor (B,))."""
    mh = mus if isinstance(mus, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(mus, dtype=np.float64))
    mh = mh.reshape(-1, dev_model.n_mu)
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    B, N = mh.shape[0], dev_model.N
    if out_host is None:
        out_host = torch.empty((B, nt, N), dtype=F64, pin_memory=True)
    if not (out_host.is_contiguous() and tuple(out_host.shape) == (B, nt, N)):
        raise ValueError("out_host must be a contiguous (B, nt, N) host tensor")
    lib = load_library()
    h = handle(dev.index)
    with torch.cuda.device(dev):
        main = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        md = mh.to(dev, non_blocking=True)
        info = torch.empty(B, dtype=torch.int32, device=dev)
        # the device history of a parameter block must fit next to the sweep's own scratch: blocks of at most a third of the free memory
        free_bytes, _ = torch.cuda.mem_get_info(dev)
        block = max(1, min(B, int(free_bytes // 3) // (nt * N * 8)))
        pitch = nt * N * 8
        for p0 in range(0, B, block):
            nb = min(block, B - p0)
            out = torch.empty((nb, nt, N), dtype=F64, device=dev)
            host0 = out_host.data_ptr() + p0 * pitch

            def ship(n0, n1, out=out, host0=host0, nb=nb):
                done = torch.cuda.Event()
                done.record(main)
                side.wait_event(done)
                rc = lib.rt_memcpy2d_async(h, c_vp(host0 + n0 * N * 8), pitch, c_vp(out.data_ptr() + n0 * N * 8), pitch,
                                           (n1 - n0) * N * 8, nb, 1, c_vp(side.cuda_stream))
                check(rc, h, "rt_memcpy2d_async")

            _, info_b = online_sweep_segments(dev_model, md[p0:p0 + nb], nt, dt, bdf=bdf, segment=segment, out=out, on_segment=ship)
            info[p0:p0 + nb] = info_b
            out.record_stream(side)
        main.wait_stream(side)
        side.synchronize()  # the history is in HOST memory: it must be complete when this returns
    return out_host, info


def rom_difference(uN, uN_srom, V_srom):
    """rt_rom_difference_f64: the S-ROM estimator of utils.compute_rom_difference for every (parameter, step).
    uN (B, nt, N), uN_srom (B, nt, Nhat) device (or host) histories, V_srom (Nh, Nhat).  Returns err (B, nt) cuda."""
    a, b = to_device(uN).contiguous(), to_device(uN_srom).contiguous()
    if a.dim() != 3 or b.dim() != 3 or a.shape[:2] != b.shape[:2]:
        raise ValueError("uN and uN_srom must be (B, nt, N) and (B, nt, Nhat) histories of the same sweep")
    V = to_device(V_srom)
    Nh, Nhat = V.shape
    if b.shape[2] != Nhat or a.shape[2] > Nhat:
        raise ValueError("V_srom must have one column per S-ROM coefficient and the S-ROM must contain the ROM")
    G, _ = pod_gram(V, normalize=False)
    G = G.contiguous()
    B, nt, N = a.shape
    err = torch.empty((B, nt), dtype=F64, device=a.device)
    h = handle(a.device.index)
    with torch.cuda.device(a.device):
        rc = load_library().rt_rom_difference_f64(h, _ptr(a), B, nt, N, _ptr(b), Nhat, _ptr(G), Nh, _ptr(err), stream_ptr())
    check(rc, h, "rt_rom_difference_f64")
    return err


def errors_summary(err):
    """rt_errors_summary_f64: err (B, nt) -> (B, 4) cuda tensor {mean, median, max, min} per parameter."""
    e = to_device(err).contiguous()
    if e.dim() != 2:
        raise ValueError("err must be (B, nt)")
    B, nt = e.shape
    out = torch.empty((B, 4), dtype=F64, device=e.device)
    h = handle(e.device.index)
    with torch.cuda.device(e.device):
        rc = load_library().rt_errors_summary_f64(h, _ptr(e), B, nt, _ptr(out), stream_ptr())
    check(rc, h, "rt_errors_summary_f64")
    return out


PROFILE_CLASSES = ("state", "gemm_Tu", "entries", "gemm_theta", "gemm_assembly", "gemm_forcing", "lu_solve", "prestep_fused")


def profile_enable(on=True):
    h = handle()
    check(load_library().rt_profile_enable(h, int(bool(on))), h, "rt_profile_enable")


def profile_read():
    """{class: (milliseconds, launches)} accumulated since the last read (first chunk of each sweep)."""
    h = handle()
    ms = (ctypes.c_double * 8)()
    cnt = (ctypes.c_int64 * 8)()
    check(load_library().rt_profile_read(h, ms, cnt), h, "rt_profile_read")
    return {name: (ms[i], int(cnt[i])) for i, name in enumerate(PROFILE_CLASSES) if cnt[i]}
