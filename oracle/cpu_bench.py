"""TEST / BASELINE INFRASTRUCTURE -- timed CPU runs for bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm.
Never imported by romtime_b200/.

Two implementations of the online loop (rom/rom.py:430-555) can be timed, one worker process per host core with the
BLAS threads pinned to 1 in the workers (the loop is a Python loop of small NumPy calls):

* ``kind = "reference"``: the UNMODIFIED reference -- ``RomConstructorNonlinear.solve`` imported from /root/reference or,
  on the GPU box, from the vendored copy ``oracle/_ref`` (``make -C oracle``), driven by oracle/ref_harness.py (stub
  FEniCS modules, a duck-typed FOM, the reference's own DEIM / MDEIM / N-MDEIM objects).  ``solver = "gmres"`` is the
  stock restarted GMRES (rom/rom.py:36), ``"direct"`` swaps the algebraic-solver hook for ``np.linalg.solve``.
* ``kind = "port"``: the oracle's NumPy restatement ``romtime_oracle.rom_sweep`` (used when no reference tree travels).

The pool is created and warmed ONCE (imports, model build); only ``pool.map`` of the measured work is timed.
"""
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
FOM_ROWS = 1000  # rows of the reduced basis V the reference lifts with every step (rom/rom.py:496): MFP1-like nx


def _paths():
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)


_STATE = {}


def _worker(args):
    """One parameter, nt steps.  Returns (seconds, trajectory (N x nt) or None)."""
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    _paths()
    import numpy as np

    import romtime_oracle as orc
    from romtime_b200 import synthetic

    kind, solver, seed, N, m, nt, dt, idx, keep = args
    key = (seed, N, m)
    if _STATE.get("key") != key:
        _STATE["key"] = key
        _STATE["model"] = synthetic.make_piston_model(N=N, m=m, seed=seed)
        _STATE["V"] = np.linalg.qr(np.random.default_rng(seed).standard_normal((FOM_ROWS, N)))[0]
    model = _STATE["model"]
    mu_row = synthetic.sample_mus(idx + 1, seed=seed)[idx]
    if kind == "reference":
        import contextlib
        import io

        import ref_harness

        rom = ref_harness.build_reference_rom(model, _STATE["V"], "2", dt, nt, solver == "direct")
        mu = {"a0": mu_row[0], "omega": mu_row[1], "delta": mu_row[2]}
        with contextlib.redirect_stderr(io.StringIO()):  # the reference's tqdm bar
            t0 = time.perf_counter()
            rom.solve(mu=mu, step="online")
            wall = time.perf_counter() - t0
        out = np.array(rom.solutions.rom)
    else:
        t0 = time.perf_counter()
        out = orc.rom_sweep(model, mu_row, nt, dt, bdf2=True, solver="direct")
        wall = time.perf_counter() - t0
    return wall, (out if keep else None)


def reference_kind():
    """"reference" when the unmodified reference can be imported here (mounted tree or oracle/_ref), else "port"."""
    _paths()
    import ref_harness

    return "reference" if ref_harness.reference_available() else "port"


class SweepPool:
    """`cores` warmed worker processes; run() times only the measured map."""

    def __init__(self, cores, N, m, seed=1005, kind=None, solver="gmres"):
        import multiprocessing as mp

        self.cores, self.N, self.m, self.seed = cores, N, m, seed
        self.kind = kind or reference_kind()
        self.solver = solver if self.kind == "reference" else "direct"
        self._saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS")}
        os.environ["OMP_NUM_THREADS"] = "1"
        os.environ["OPENBLAS_NUM_THREADS"] = "1"
        self.pool = mp.get_context("spawn").Pool(cores)
        self.pool.map(_worker, [(self.kind, self.solver, seed, N, m, 2, 1e-3, i, False) for i in range(cores)])  # warm-up, untimed

    def run(self, nt, dt=1e-3, keep=False, solver=None):
        """`cores` parameters x nt steps.  Returns (param*steps/s, wall seconds, list of trajectories or None)."""
        solver = solver or self.solver
        t0 = time.perf_counter()
        res = self.pool.map(_worker, [(self.kind, solver, self.seed, self.N, self.m, nt, dt, i, keep) for i in range(self.cores)])
        wall = time.perf_counter() - t0
        return self.cores * nt / wall, wall, ([r[1] for r in res] if keep else None)

    def describe(self, nt, solver=None):
        solver = solver or self.solver
        if self.kind == "reference":
            how = "stock GMRES(20), tol 1e-10 (rom/rom.py:36)" if solver == "gmres" else "algebraic-solver hook = np.linalg.solve"
            return (f"unmodified reference RomConstructorNonlinear.solve (rom/rom.py:430-555) from oracle/_ref, {how}: "
                    f"{self.cores} parameters x {nt} BDF2 steps, N = {self.N}, m = {self.m}, V {FOM_ROWS} x {self.N}, "
                    "one process per core, BLAS threads 1, pool start-up and warm-up not timed")
        return (f"oracle rom_sweep (NumPy port of rom/rom.py:430-555, direct solve): {self.cores} parameters x {nt} BDF2 steps, "
                f"N = {self.N}, m = {self.m}, one process per core, BLAS threads 1, pool start-up and warm-up not timed")

    def close(self):
        self.pool.close()
        self.pool.join()
        for k, v in self._saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def orth_time(Nh, n, seed=1003):
    """Seconds for the oracle's orth (LAPACK gesvd, all BLAS threads) on an Nh x n synthetic snapshot matrix."""
    _paths()
    import romtime_oracle as orc
    from romtime_b200 import synthetic

    S = synthetic.pod_snapshots(Nh, n, seed=seed, modes=min(256, n), decay=32.0)
    t0 = time.perf_counter()
    orc.orth(S, num=min(64, n))
    return time.perf_counter() - t0
