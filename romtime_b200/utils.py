"""Helpers of the hot path with the reference's names (reference utils.py).  Dense projections run on the
device through the C ABI; FEniCS-typed inputs (dolfin Matrix / Vector) are converted by duck typing so
that this module imports without FEniCS."""
import json
import pickle

import numpy as np


def functional_to_array(operator):
    """Algebraic vector -> dense NumPy array (reference utils.py:57-72)."""
    return np.array(operator)


def function_to_array(coef_func):
    """FE coefficient function -> its dof array (reference utils.py:44-55)."""
    return coef_func.vector().vec().array


def bilinear_to_csr(matrix):
    """Assembled bilinear form -> scipy CSR (reference utils.py:75-93).  Accepts scipy sparse / dense
    input directly; a dolfin matrix goes through its PETSc backend like the reference."""
    import scipy.sparse as sp

    if sp.issparse(matrix):
        return matrix.tocsr()
    if isinstance(matrix, np.ndarray):
        return sp.csr_matrix(matrix)
    backend = matrix
    if not hasattr(matrix, "mat"):
        import fenics  # only reached with real dolfin matrices

        backend = fenics.as_backend_type(matrix)
    mat = backend.mat()  # PETSc Mat protocol: getValuesCSR() -> (indptr, indices, data), .size
    return sp.csr_matrix(mat.getValuesCSR()[::-1], shape=mat.size)


def vector_to_csr(entries, rows, cols):
    """Vectorised non-zeros -> CSR (reference utils.py:136-149)."""
    from scipy.sparse import csr_matrix

    return csr_matrix((entries, (rows, cols)))


def get_nonzero_entries(Ah):
    from scipy.sparse import find

    return find(Ah)


def eliminate_zeros(Ah, tolerance=1e-15):
    """Drop entries that are zero to 1e-15 (reference utils.py:152-170)."""
    mask = np.isclose(Ah.data, 0, rtol=tolerance, atol=tolerance)
    Ah.data[mask] = 0
    Ah.eliminate_zeros()
    return Ah


def project_csr(Ah, V):
    """A_N = V^T A_h V on the device (reference utils.py:96-113): the CSR matrix is handed over as one
    'collateral mode' (its COO triplets) to rt_project_mat_f64."""
    from . import kernels

    coo = Ah.tocoo()
    N = V.shape[1]
    out = kernels.project_mat(coo.data.reshape(-1, 1), coo.row, coo.col, V)
    return out.reshape(N, N).cpu().numpy()


# ------------------------------------------------------------------ certification (host, O(N))
def compute_rom_difference(uN, uN_srom, V_srom):
    """S-ROM error estimator (reference utils.py:173-212): RMS of V_srom (uN_srom - [uN; 0])."""
    pad = np.zeros(len(uN_srom))
    pad[: len(uN)] = uN
    lincomb = np.sum((uN_srom - pad) * V_srom, axis=1)
    return np.linalg.norm(lincomb, ord=2) / np.sqrt(len(lincomb))


def compute_rom_difference_batch(uN, uN_srom, V_srom):
    """compute_rom_difference for every (parameter, step) of two batched sweeps, on the device
    (rt_rom_difference_f64): uN (B, nt, N) and uN_srom (B, nt, Nhat) as ``RomBatchSolutions.rom_device`` holds them,
    V_srom (Nh, Nhat) the S-ROM basis.  Returns a CUDA tensor (B, nt) -- what the certification loop
    (reference rom/hrom.py:504-626) collects per online parameter."""
    from . import kernels

    return kernels.rom_difference(uN, uN_srom, V_srom)


def errors_summary_batch(err, index=None):
    """Reductor.create_errors_summary (reference rom/base.py:154-163) for a (B, nt) error table on the device:
    a pandas DataFrame with one row per parameter and the columns mean / median / max / min."""
    import pandas as pd

    from . import kernels

    summ = kernels.errors_summary(err).cpu().numpy()
    return pd.DataFrame(summ, columns=["mean", "median", "max", "min"], index=index)


def time_average(ts, func):
    integrate = getattr(np, "trapezoid", None) or np.trapz
    return integrate(y=func, x=ts) / np.max(ts)


def singular_to_pod_error(sigmas):
    lam = np.power(sigmas, 2)
    return np.sqrt(np.sum(lam) - np.cumsum(lam))


def singular_to_energy(sigmas):
    lam = np.power(sigmas, 2)
    return np.cumsum(lam) / np.sum(lam)


# ------------------------------------------------------------------ IO (reference utils.py:246-278)
def read_pickle(path):
    with open(path, mode="rb") as fp:
        return pickle.load(fp)


def dump_pickle(path, obj):
    with open(path, mode="wb") as fp:
        pickle.dump(obj, fp)


def dump_json(path, obj):
    with open(path, mode="w") as fp:
        json.dump(obj, fp)


def read_json(path):
    with open(path, mode="r") as fp:
        return json.load(fp)


def dump_csv(path, obj):
    import pandas as pd

    pd.DataFrame(obj).to_csv(path)
