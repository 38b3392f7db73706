"""GPU: rt_sym_eig_jacobi_static_f64 -- the entry point that allocates nothing (work space in a static __device__ array of the
library, n <= RT_STATIC_EIG_MAX_N = 128) -- against rt_sym_eig_jacobi_f64 on the same inputs, through the C ABI, two calls in
a row (the second call re-uses the array with the first call's contents in it).  No counterpart in the reference; collected
last on purpose."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

MAX_N = 128


def _gram_of_nearly_orthogonal(n, seed, rows=4000):
    """H = Y^T Y of graded, nearly orthogonal columns: what the Hestenes refinement hands to the Jacobi solver."""
    rng = np.random.default_rng(seed)
    Q, _ = np.linalg.qr(rng.standard_normal((rows, n)))
    Y = Q * np.power(10.0, -np.arange(n) / 16.0)
    Y = Y + 1e-9 * rng.standard_normal((rows, n)) * np.linalg.norm(Y, axis=0)
    return Y.T @ Y


def _call(name, H):
    """One call of entry point ``name`` through ctypes; every buffer is allocated BEFORE the call."""
    from romtime_b200 import _lib
    from romtime_b200.kernels import stream_ptr

    n = H.shape[0]
    w = torch.empty(n, dtype=torch.float64, device="cuda")
    W = torch.empty((n, n), dtype=torch.float64, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    defect = torch.zeros(1, dtype=torch.float64, device="cuda")
    h = _lib.handle(H.device.index)
    lib = _lib.load_library()
    torch.cuda.synchronize()
    rc = getattr(lib, name)(h, H.data_ptr(), n, w.data_ptr(), W.data_ptr(), info.data_ptr(), defect.data_ptr(), stream_ptr())
    torch.cuda.synchronize()
    return rc, w, W, int(info.item()), float(defect.item())


@pytest.mark.parametrize("n", [72, 128, 17])
def test_static_scratch_jacobi_matches_the_allocating_entry_point(n):
    for call in range(2):   # two calls in a row: the second finds the previous contents in the static array
        H = torch.as_tensor(_gram_of_nearly_orthogonal(n, seed=100 * n + call), device="cuda")
        rc0, w0, W0, sweeps0, defect0 = _call("rt_sym_eig_jacobi_f64", H)
        rc1, w1, W1, sweeps1, defect1 = _call("rt_sym_eig_jacobi_static_f64", H)
        assert rc0 == 0 and rc1 == 0
        # same kernels, same launch order: the results are the same numbers (1e-10 is the tolerance of the POD tests)
        assert torch.max((w1 - w0).abs() / w0.abs()).item() < 1e-10
        assert (W1 - W0).abs().max().item() < 1e-10
        assert sweeps1 == sweeps0 and defect1 == defect0
        # and they are an eigen-decomposition: H W = W diag(w) relative to each eigenvalue's scale, W orthogonal
        Hn, Wn, wn = H.cpu().numpy(), W1.cpu().numpy(), w1.cpu().numpy()
        assert np.max(np.abs(Wn.T @ Wn - np.eye(n))) < 1e-12
        ref = np.linalg.eigvalsh(Hn)[::-1]
        assert np.max(np.abs(wn - ref) / ref) < 1e-10


def test_static_scratch_jacobi_rejects_sizes_beyond_the_array():
    from romtime_b200 import _lib

    n = MAX_N + 1
    H = torch.eye(n, dtype=torch.float64, device="cuda")
    rc, _, _, _, _ = _call("rt_sym_eig_jacobi_static_f64", H)
    assert rc != 0
    msg = _lib.load_library().rt_last_error(_lib.handle(0)).decode()
    assert "128" in msg and "static" in msg
    # the limit is not a property of the method: the allocating entry point takes the same matrix
    rc, w, _, _, _ = _call("rt_sym_eig_jacobi_f64", H)
    assert rc == 0 and torch.allclose(w, torch.ones(n, dtype=torch.float64, device="cuda"))
