"""TEST / BASELINE INFRASTRUCTURE -- timed CPU runs of the oracle (the NumPy/SciPy restatement of the
reference's hot path) for bench.py's ``cpu_baseline`` leg and ``--impl reference`` arm.
Never imported by romtime_b200/.  One worker process per host core, BLAS threads pinned to 1 in the
workers (the reference's online loop is a Python loop of small NumPy calls; rom/rom.py:430-555)."""
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _worker(args):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import numpy as np

    import romtime_oracle as orc
    from romtime_b200 import synthetic

    seed, N, m, nt, dt, idx = args
    model = synthetic.make_piston_model(N=N, m=m, seed=seed)
    mu = synthetic.sample_mus(idx + 1, seed=seed)[idx]
    t0 = time.perf_counter()
    out = orc.rom_sweep(model, mu, nt, dt, bdf2=True, solver="direct")
    return time.perf_counter() - t0, float(np.abs(out).max())


def online_sweep_rate(cores, N, m, nt, dt=1e-3, seed=1005):
    """Aggregate param*steps/s of the oracle's rom_sweep: `cores` workers, one parameter each, nt steps."""
    import multiprocessing as mp

    saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS")}
    os.environ["OMP_NUM_THREADS"] = "1"
    os.environ["OPENBLAS_NUM_THREADS"] = "1"
    try:
        ctx = mp.get_context("spawn")
        with ctx.Pool(cores) as pool:
            pool.map(_worker, [(seed, N, m, 1, dt, i) for i in range(cores)])  # warm the workers (imports, model build)
            t0 = time.perf_counter()
            res = pool.map(_worker, [(seed, N, m, nt, dt, i) for i in range(cores)])
            wall = time.perf_counter() - t0
    finally:
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return cores * nt / wall, wall, max(r[0] for r in res)


def orth_time(Nh, n, seed=1003):
    """Seconds for the oracle's orth (LAPACK gesvd, all BLAS threads) on an Nh x n synthetic snapshot matrix."""
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    import romtime_oracle as orc
    from romtime_b200 import synthetic

    S = synthetic.pod_snapshots(Nh, n, seed=seed, modes=min(256, n), decay=32.0)
    t0 = time.perf_counter()
    orc.orth(S, num=min(64, n))
    return time.perf_counter() - t0
