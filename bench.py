#!/usr/bin/env python
"""bench.py -- headline benchmark of the romtime hot path on B200 (contract: see the task prompt).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[4] + configs[2]):
  * online: batched ROM sweep, B = 65,536 parameters (sharded over the ranks: strong scaling),
    reduced dimension N = 64, five MDEIM operators + one DEIM vector with m = 128 collateral modes,
    BDF2.  One bench "step" = ONE uninterrupted sweep of the whole batch over --nt-per-step implicit time steps from the
    zero initial condition (default 1000 = the named config: 33.5 GB of history per step, written to HBM).
    metric = ROM param*timesteps/s.  The line also carries a 1000-step parity figure against the CPU reference, an
    m = 16 sweep, the POD in both storage orders and the C4 / certification legs.
  * pod (reported in the same line): orth of a synthetic 1,000,000 x 4096 fp64 snapshot matrix, num=64.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "rom_param_timesteps_per_s"
# dominant kernel's DRAM traffic per launch from the committed ncu capture of this round (K written once, M never stored)
NCU_GEMM_PROFILE = "profiles/r2_sweep_ncu.md"
NCU_GEMM_DRAM_BYTES_PER_14208 = 531.8e6  # 101.95 MB read + 429.86 MB written
UNIT = "param*timesteps/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--params", type=int, default=65536)
    ap.add_argument("--nt-per-step", type=int, default=1000)
    ap.add_argument("--e2e-steps", type=int, default=3, help="timed repetitions of the host-to-host leg (at most --steps)")
    ap.add_argument("--m", type=int, default=128)
    ap.add_argument("--pod-rows", type=int, default=1_000_000)
    ap.add_argument("--pod-cols", type=int, default=4096)
    ap.add_argument("--skip-pod", action="store_true")
    ap.add_argument("--c4-rows", type=int, default=200_000)
    ap.add_argument("--c4-cols", type=int, default=2048)
    ap.add_argument("--skip-cpu", action="store_true")
    return ap.parse_args()


def workload_config(a, world):
    return {
        "workload": "C5 batched online ROM sweep (BASELINE.json configs[4]) + C3 POD (configs[2])",
        "params": a.params, "reduced_dim": 64, "collateral_modes": a.m, "operators": "5 MDEIM + 1 DEIM",
        "bdf": 2, "dt": 1e-3, "timesteps_per_bench_step": a.nt_per_step,
        "sharding": f"parameters over {world} GPU(s), no collective",
        "l2": "working set per time step (K buffer 466 MB per 14,208-parameter chunk, written by the GEMM and read by the solve) exceeds the 126 MB L2; no explicit flush",
        "history": "B x nt x N written to HBM every step (33.5 GB per 1000-step sweep at B = 65,536)",
    }


# ------------------------------------------------------------------------------- reference arm
def reference_arm(a, out_fd):
    """CPU arm: the UNMODIFIED reference's RomConstructorNonlinear.solve (rom/rom.py:430-555, stock GMRES) from the
    vendored copy oracle/_ref, one process per host core (oracle/cpu_bench.py); the oracle port only if no reference
    tree travelled.  Each bench step = `cores` parameters x a bounded number of time steps of the same C5 model; the
    worker pool is created and warmed before the timer starts."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_bench

    cores = os.cpu_count() or 1
    nt = 300  # bounded sample per bench step: a few seconds on every core
    pool = cpu_bench.SweepPool(cores, 64, a.m)
    try:
        for _ in range(min(a.warmup, 1)):
            pool.run(20)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            pool.run(nt)
        total = time.perf_counter() - t0
        direct_rate = pool.run(nt, solver="direct")[0] if pool.kind == "reference" else None
    finally:
        pool.close()
    value = cores * nt * a.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": total / a.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(a, a.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": pool.kind, "sample": pool.describe(nt),
                         "value_with_direct_solve_hook": direct_rate},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line, out_fd)


# ------------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) >= 8:
                for name, v in zip(names, r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        pw = [float(r[3]) for r in self.rows if len(r) >= 8 and r[3].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def hbm_peak_gbs():
    """Measured HBM copy bandwidth of this pool's B200s (driver-written MEASURED_PEAKS.json), else the profiling guide's figure."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6550.0


# ------------------------------------------------------------------------------- POD leg
def pod_leg(a, torch, kernels, world, rank):
    from romtime_b200.rom import pod as podmod

    Nh, n = a.pod_rows, a.pod_cols
    if world > 1:
        from romtime_b200 import distributed as rdist

        return rdist.bench_pod(a, Nh, n)
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev).manual_seed(1003)
    k = 256
    S = torch.empty((Nh, n), dtype=torch.float64, device=dev)
    R = torch.randn((k, n), dtype=torch.float64, device=dev, generator=g)
    sig = torch.pow(10.0, -torch.arange(k, dtype=torch.float64, device=dev) / 32.0)
    rows = 65536
    for r0 in range(0, Nh, rows):  # low-rank graded part + 1e-13 noise, generated in row slabs
        r1 = min(Nh, r0 + rows)
        L = torch.randn((r1 - r0, k), dtype=torch.float64, device=dev, generator=g) * sig
        S[r0:r1] = L @ R
        S[r0:r1] += 1e-13 * torch.randn((r1 - r0, n), dtype=torch.float64, device=dev, generator=g)
    del L, R
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    res = {}
    best = None
    for it in range(3):
        e = [ev() for _ in range(5)]
        e[0].record()
        G, cn = kernels.pod_gram(S, normalize=True)
        e[1].record()
        w, W, sweeps = kernels.sym_eig(G)
        e[2].record()
        Q, s, energy, VT, info = podmod.orth_device(S, num=64)
        e[3].record()
        torch.cuda.synchronize()
        t_gram, t_eig, t_orth = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[2].elapsed_time(e[3])
        if best is None or t_orth < best[2]:
            best = (t_gram, t_eig, t_orth)
    t_gram, t_eig, t_orth = best
    peak = kernels.fp64_peak_tflops()
    flops = float(Nh) * n * (n + 1)
    QtQ = (Q.t() @ Q - torch.eye(Q.shape[1], dtype=torch.float64, device=dev)).abs().max().item()
    rest = max(t_orth - t_gram - t_eig, 0.0)
    res = {
        "shape": [Nh, n], "num": 64, "order": "C (np.hstack, rom/rom.py:368)", "svd_time_ms": t_orth, "gram_ms": t_gram, "eig_ms": t_eig,
        "lift_refinement_and_rest_ms": rest, "lift_refinement_share_of_orth": rest / t_orth,
        "gram_tflops_sym_half": flops / t_gram * 1e-9,
        "gram_frac_of_fp64_peak": flops / t_gram * 1e-9 / peak, "S_GBps_whole_orth": 8.0 * Nh * n / t_orth * 1e-6,
        "eig_sweeps": sweeps, "QtQ_minus_I_max": QtQ, "energy_last": float(energy[-1]),
        "refinement": {k: info[k] for k in ("levels", "rank_gram", "candidates", "passes", "defect") if k in info},
        "singular_values_head": [float(x) for x in s[:4]],
    }
    del G, W, Q
    try:
        # the same matrix stored column-major (np.array(list).T, deim/deim.py:384): the Gram kernel reads it through the MN-major tensor map
        Sf = torch.empty((n, Nh), dtype=torch.float64, device=dev)
        for r0 in range(0, Nh, rows):
            Sf[:, r0:min(Nh, r0 + rows)] = S[r0:min(Nh, r0 + rows)].t()
        del S
        Sf = Sf.t()
        tf = None
        for it in range(2):
            e0, e1 = ev(), ev()
            e0.record()
            Qf, sf, _, _, _ = podmod.orth_device(Sf, num=64)
            e1.record()
            torch.cuda.synchronize()
            tf = e0.elapsed_time(e1) if tf is None else min(tf, e0.elapsed_time(e1))
        res["svd_time_ms_F_order"] = tf
        res["F_order_singular_values_max_rel_vs_C_order"] = float(abs(sf[:64] - s[:64]).max() / s[0])
        del Sf, Qf
    except Exception as exc:  # report, never hide
        res["svd_time_ms_F_order"] = repr(exc)
    torch.cuda.empty_cache()
    if rank == 0 and not a.skip_cpu:
        # CPU baseline (SURVEY.md 8d / BASELINE.md 3): the oracle's orth (rom/pod.py:7-62: column norms + LAPACK gesvd, all BLAS
        # threads) on 1,000,000 x 512 of the same family; the full-size figure is an O(Nh n^2) EXTRAPOLATION (x 64) and labelled so
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import numpy as np
        import romtime_oracle as orc_mod
        from romtime_b200 import synthetic

        nh_s, n_s = min(Nh, 1_000_000), min(n, 512)
        Ss = np.empty((nh_s, n_s))
        for r0 in range(0, nh_s, 100_000):  # generated in row slabs: the closed form needs a rows x modes temporary
            r1 = min(nh_s, r0 + 100_000)
            Ss[r0:r1] = synthetic.pod_snapshots(nh_s, n_s, seed=1003, row0=r0, rows=r1 - r0)
        t0 = time.perf_counter()
        Qc, sc, _ = orc_mod.orth(Ss, num=64)
        wall = time.perf_counter() - t0
        res["cpu_orth_sample_ms"] = wall * 1e3
        res["cpu_orth_sample"] = f"oracle orth (scipy gesvd) on {nh_s} x {n_s}, {os.cpu_count()} host cores, BLAS default threads"
        res["cpu_orth_extrapolated_full_ms"] = wall * 1e3 * (Nh / nh_s) * (n / n_s) ** 2
        res["cpu_orth_note"] = f"full size extrapolated by Nh n^2 (x {(Nh / nh_s) * (n / n_s) ** 2:.0f}) from the sample, not measured"
        # the device path on the same sample, for a like-for-like time and parity figure
        Sd = kernels.to_device(Ss)
        podmod.orth_device(Sd, num=64)
        e0, e1 = ev(), ev()
        e0.record()
        Qd, sd, _, _, _ = podmod.orth_device(Sd, num=64)
        e1.record()
        torch.cuda.synchronize()
        res["gpu_orth_same_sample_ms"] = e0.elapsed_time(e1)
        res["sample_singular_values_max_rel"] = float(np.abs(sd[:64] - sc[:64]).max() / sc[0])
        dots = np.abs(np.sum(Qd.cpu().numpy() * Qc, axis=0))
        res["sample_modes_max_abs_1_minus_dot"] = float(np.abs(dots - 1.0).max())
        del Sd, Qd
    return res


# ------------------------------------------------------------------------------- C4 leg (MDEIM offline chain)
def c4_leg(a, torch, kernels, with_cpu):
    """BASELINE.json configs[3]: POD (num = 128, not normalised) of 200,000 x 2048 matrix snapshots, then the greedy
    point selection on the 200,000 x 128 collateral basis (deim/deim.py:517-561) -- the HBM-streaming kernel."""
    from romtime_b200 import synthetic
    from romtime_b200.rom import pod as podmod

    nnz, n, m = a.c4_rows, a.c4_cols, 128
    S = synthetic.mdeim_snapshots_device(nnz, n, seed=1004)
    ev = lambda: torch.cuda.Event(enable_timing=True)  # noqa: E731
    t_pod = None
    for _ in range(3):
        e0, e1 = ev(), ev()
        e0.record()
        Q, s, energy, VT, _ = podmod.orth_device(S, num=m, normalize=False)
        e1.record()
        torch.cuda.synchronize()
        t_pod = e0.elapsed_time(e1) if t_pod is None else min(t_pod, e0.elapsed_time(e1))
    del S
    U = Q                                       # (nnz x m) column-major CUDA tensor
    kernels.deim_greedy(U)
    t_gr = None
    for _ in range(5):
        e0, e1 = ev(), ev()
        e0.record()
        dofs, PT_U = kernels.deim_greedy(U, sync=False)
        e1.record()
        torch.cuda.synchronize()
        t_gr = e0.elapsed_time(e1) if t_gr is None else min(t_gr, e0.elapsed_time(e1))
    alg = 8.0 * nnz * m * (m + 1) / 2           # unblocked left-looking bytes
    blk = 8.0 * nnz * m * (m / 8 + 1) / 2       # what the 8-column blocked kernel streams from HBM
    hbm = hbm_peak_gbs()
    res = {"shape": [nnz, n], "m": m, "pod_ms": t_pod, "pod_GBps_of_S": 8.0 * nnz * n / t_pod * 1e-6,
           "sigma_m_over_sigma_1": float(s[m - 1] / s[0]), "QtQ_minus_I_max": float((Q.t() @ Q - torch.eye(m, dtype=torch.float64, device=Q.device)).abs().max().item()),
           "greedy_ms": t_gr,
           "greedy_traffic_GBps": blk / t_gr * 1e-6, "greedy_frac_of_hbm_peak": blk / t_gr * 1e-6 / hbm,
           "greedy_traffic_model": "bytes the 8-column blocked kernel streams: 8 Nh m (m / 8 + 1) / 2 = 1.74 GB (ncu dram__bytes_read 1.28 GB, the rest L2 hits)",
           "greedy_us_per_point": 1e3 * t_gr / m, "hbm_peak_GBps": hbm, "unique_dofs": int(torch.unique(dofs).numel()),
           "survey_8d_unblocked_bytes_GBps": alg / t_gr * 1e-6,
           "bound": "latency: m grid-wide arg-max exchanges at ~10 us each; HBM at 0.2 of peak.  SURVEY.md 8d's unblocked byte model "
                    "(8 Nh m (m+1) / 2 = 13.2 GB -> 2.0 ms at the HBM peak) is kept as context only: the blocked kernel streams 8x less"}
    if with_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import romtime_oracle as orc_mod

        Uh = U.cpu().numpy()
        t0 = time.perf_counter()
        ref_dofs = orc_mod.deim_greedy(Uh)
        wall = time.perf_counter() - t0
        res["cpu_greedy_ms"] = wall * 1e3
        res["cpu_sample"] = f"oracle deim_greedy (deim/deim.py:517-561) on the same {nnz} x {m} basis, NumPy/OpenBLAS default threads"
        res["indices_bit_exact"] = bool([int(d) for d in ref_dofs] == dofs.cpu().tolist())
        res["PT_U_max_abs_diff"] = float(abs(PT_U.cpu().numpy() - Uh[ref_dofs, :]).max())  # P^T U is a row gather: exact
    del U, Q
    torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------- tree-walk leg
def treewalk_leg(a, torch, with_cpu):
    """How the reference really consumes orth (rom/rom.py:313-406, deim/deim.py:279-397): P per-parameter time-PODs with
    tol-driven ragged widths, then the mu-POD of the stacked bases -- at C1/C2-like sizes, through the batched device path
    (two host reads per batch), with the reference's CPU loop (one scipy gesvd per matrix) beside it."""
    import numpy as np

    from romtime_b200 import synthetic
    from romtime_b200.rom.pod import orth, orth_batch

    tol = 1.0 - 1e-8
    res = {"tol": "1 - 1e-8 (energy)", "cases": []}
    for P, Nh, nt in ((64, 1000, 250), (64, 100, 50)):
        mats = [synthetic.treewalk_snapshots(Nh, nt, seed=p) for p in range(P)]

        def walk():
            per = orth_batch(mats, tol=tol)
            return per, orth(np.hstack([q for q, _, _ in per]), tol=tol, normalize=False)

        walk()
        torch.cuda.synchronize()
        best = None
        for _ in range(3):
            t0 = time.perf_counter()
            per, (Q, s, e) = walk()
            torch.cuda.synchronize()
            w = (time.perf_counter() - t0) * 1e3
            best = w if best is None else min(best, w)
        case = {"parameters": P, "snapshots_per_parameter": [Nh, nt], "host_to_host_ms": best,
                "widths_min_max": [int(min(q.shape[1] for q, _, _ in per)), int(max(q.shape[1] for q, _, _ in per))],
                "final_modes": int(Q.shape[1])}
        if with_cpu:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import romtime_oracle as orc_mod

            t0 = time.perf_counter()
            per_c = [orc_mod.orth(S, tol=tol) for S in mats]
            Qc, sc, _ = orc_mod.orth(np.hstack([q for q, _, _ in per_c]), tol=tol, normalize=False)
            case["cpu_loop_ms"] = (time.perf_counter() - t0) * 1e3
            case["cpu_sample"] = f"oracle orth (rom/pod.py:7-62, scipy gesvd) per matrix + stacked mu-POD, {os.cpu_count()} host cores, BLAS default threads"
            case["widths_equal_cpu"] = bool([q.shape[1] for q, _, _ in per] == [q.shape[1] for q, _, _ in per_c])
            case["per_parameter_sigma_max_rel"] = float(max(np.max(np.abs(a_[1][: a_[0].shape[1]] - b_[1][: b_[0].shape[1]]) / b_[1][: b_[0].shape[1]])
                                                            for a_, b_ in zip(per, per_c)))
        res["cases"].append(case)
    return res


# ------------------------------------------------------------------------------- main arm
def emit(line, fd):
    os.write(fd, (json.dumps(line) + "\n").encode())


def main():
    a = parse()
    # stdout carries exactly ONE JSON line: everything libraries print while we run (NCCL banner, ...)
    # is diverted to stderr, the line itself goes to the saved descriptor
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)
    if a.impl == "reference":
        return reference_arm(a, out_fd)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    cpu_baseline, cpu_traj, cpu_traj_gmres = None, None, None
    if rank == 0 and world == 1 and not a.skip_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import cpu_bench

        cores = os.cpu_count() or 1
        nt_cpu = 1000  # one full C5 trajectory per core: 10-20 s of CPU work
        pool = cpu_bench.SweepPool(cores, 64, a.m)
        try:
            rate, wall, cpu_traj_gmres = pool.run(nt_cpu, keep=True)
            cpu_baseline = {"value": rate, "unit": UNIT, "cores": cores, "kind": pool.kind,
                            "sample": pool.describe(nt_cpu) + f", {wall:.1f} s"}
            if pool.kind == "reference":
                # the same trajectories with the reference's algebraic-solver hook swapped for a direct solve: what the
                # device path (dense pivoted solve) is compared with at 1e-10; stock GMRES carries its own 1e-10 stopping error
                rate_d, wall_d, cpu_traj = pool.run(nt_cpu, keep=True, solver="direct")
                cpu_baseline["value_with_direct_solve_hook"] = rate_d
            else:
                cpu_traj, cpu_traj_gmres = cpu_traj_gmres, None
        finally:
            pool.close()

    import torch
    import torch.distributed as dist

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from romtime_b200 import kernels, synthetic

    dev = torch.device("cuda", local)
    N, m, nt, dt = 64, a.m, a.nt_per_step, 1e-3
    B = a.params
    lo, hi = rank * B // world, (rank + 1) * B // world
    model = synthetic.make_piston_model(N=N, m=m, seed=1005)
    dm = kernels.DeviceRomModel(model)
    mus_all = synthetic.sample_mus(B, seed=1005)
    mus_host = torch.from_numpy(mus_all[lo:hi].copy()).pin_memory()
    mus = mus_host.to(dev)
    Bl = hi - lo

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- 1000-step parity against the CPU reference on the parameters its workers ran (the first `cores` of the batch)
    parity = None
    if cpu_traj is not None:
        import numpy as np

        nb = len(cpu_traj)
        up, ip = kernels.online_sweep(dm, mus_all[:nb], 1000, dt, bdf=2)
        up = up.cpu().numpy().transpose(0, 2, 1)  # (nb, N, nt) like RomSolutionsStorage.rom
        rel = lambda x, y: float(np.max(np.abs(x - y)) / np.max(np.abs(y)))  # noqa: E731
        parity = {"parameters": nb, "steps": 1000, "parity_1000_steps_max_rel": max(rel(up[i], cpu_traj[i]) for i in range(nb)),
                  "against": "reference RomConstructorNonlinear.solve with the algebraic-solver hook = np.linalg.solve"
                             if cpu_traj_gmres is not None else "oracle rom_sweep (NumPy port, direct solve)",
                  "last_step_max_rel": max(rel(up[i][:, -1], cpu_traj[i][:, -1]) for i in range(nb)),
                  "singular_solves": int(ip.ne(0).sum().item())}
        if cpu_traj_gmres is not None:
            parity["vs_stock_gmres_max_rel"] = max(rel(up[i], cpu_traj_gmres[i]) for i in range(nb))
            parity["reference_direct_vs_reference_gmres_max_rel"] = max(rel(cpu_traj[i], cpu_traj_gmres[i]) for i in range(nb))
        del up

    out = torch.empty((Bl, nt, N), dtype=torch.float64, device=dev)

    def step_resident():
        return kernels.online_sweep(dm, mus, nt, dt, bdf=2, history=True, out=out)

    for _ in range(a.warmup):
        _, info = step_resident()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    kernels.profile_enable(True)
    kernels.profile_read()
    launches0 = kernels._lib.launch_count(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(a.steps):
        _, info = step_resident()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = kernels._lib.launch_count(local) - launches0
    prof = kernels.profile_read()
    kernels.profile_enable(False)
    clocks = sampler.stop() if sampler else None
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms = float(tms.item())
    bad = int(info.ne(0).sum().item())
    value = float(B) * nt * a.steps / (ms * 1e-3)

    # ---- end to end: host mus -> device, sweep, history back to pinned host memory ----
    # public host-to-host call (= RomConstructor*.solve_batch(..., to_host=True)): mus from pinned host memory, the history
    # lands in pinned host memory; the sweep runs in 100-step segments and every finished time slab is copied out while the
    # next segment is computed (rt_online_sweep_segment_f64 + rt_memcpy2d_async)
    e2e_steps = max(1, min(a.steps, a.e2e_steps))
    cert_rows = min(Bl, 8192)
    cert_u = out[:cert_rows].clone()  # kept for the certification leg
    del out
    torch.cuda.empty_cache()
    hist_host = torch.empty((Bl, nt, N), dtype=torch.float64, pin_memory=True)

    def step_e2e():
        kernels.online_sweep_to_host(dm, mus_host, nt, dt, bdf=2, out_host=hist_host)
        torch.cuda.synchronize()

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_e2e()
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = float(B) * nt * e2e_steps / float(te.item())
    e2e_matches = bool(torch.equal(hist_host[:cert_rows], cert_u.cpu()))
    del hist_host

    # ---- the same sweep with m = 16 collateral modes per operator (SURVEY.md 8d names m in {16, 128}): the solve, not the
    # assembly GEMM, dominates there
    m16 = None
    try:
        model16 = synthetic.make_piston_model(N=N, m=16, seed=1005)
        dm16 = kernels.DeviceRomModel(model16)
        nt16 = min(nt, 100)
        out16 = torch.empty((Bl, nt16, N), dtype=torch.float64, device=dev)
        for _ in range(2):
            kernels.online_sweep(dm16, mus, nt16, dt, bdf=2, out=out16)
        kernels.profile_enable(True)
        kernels.profile_read()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(3):
            _, info16 = kernels.online_sweep(dm16, mus, nt16, dt, bdf=2, out=out16)
        f1.record()
        barrier()
        prof16 = kernels.profile_read()
        kernels.profile_enable(False)
        t16 = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t16, op=dist.ReduceOp.MAX)
        tot16 = sum(v[0] for v in prof16.values())
        m16 = {"value": float(B) * nt16 * 3 / (float(t16.item()) * 1e-3), "unit": UNIT, "collateral_modes": 16, "timesteps": nt16,
               "steps": 3, "kernel_shares": {k: v[0] / tot16 for k, v in prof16.items()},
               "kernel_us_per_launch": {k: 1e3 * v[0] / v[1] for k, v in prof16.items()},
               "singular_solves": int(info16.ne(0).sum().item())}
        del out16, dm16
    except Exception as exc:  # report, never hide
        m16 = {"error": repr(exc)}

    # ---- roofline of the dominant kernel (assembly GEMM on the FP64 DMMA pipe) ----
    peak = kernels.fp64_peak_tflops()
    roof = None
    if "gemm_assembly" in prof:
        g_ms, g_n = prof["gemm_assembly"]
        sms = torch.cuda.get_device_properties(dev).multi_processor_count
        chunk = min(Bl, int(kernels.load_library().rt_sweep_chunk(Bl, N, sms)))
        flops = 2.0 * chunk * (N * N) * (5 * ((m + 15) // 16 * 16))
        ach = flops / (g_ms / g_n) * 1e-9
        total_prof = sum(v[0] for v in prof.values())
        roof = {"bound": "tensor", "kernel": "dg::k_dgemm<MN,MN> (K assembly + the M w contraction at the k-boundary, TMA + DMMA fp64)", "achieved": ach, "peak": peak,
                "unit": "TFLOP/s", "frac": ach / peak,
                "peak_source": "FP64 DMMA peak measured in this run (rt_fp64_peak_tflops; MEASURED_PEAKS.json has no fp64 entry)",
                "avg_launch_ms": g_ms / g_n, "launches_timed": g_n, "share_of_step": g_ms / total_prof,
                "traffic": NCU_GEMM_DRAM_BYTES_PER_14208 * chunk / 14208.0,
                "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the ncu --set full capture of the same "
                                  "sweep at one 14,208-parameter chunk (" + NCU_GEMM_PROFILE + "), scaled to this chunk; ncu cannot run inside the timed region",
                "kernel_shares": {k: v[0] / total_prof for k, v in prof.items()}}

    # ---- certification leg: S-ROM estimator over every (parameter, step) of the last sweep
    cert = None
    try:
        g = torch.Generator(device=dev).manual_seed(1006)
        Vs = torch.linalg.qr(torch.randn((20000, N), dtype=torch.float64, device=dev, generator=g))[0]
        out = cert_u  # the first parameters of the last timed sweep (the full 33.5 GB history has been released)
        srom = out + 1e-3 * torch.randn(out.shape, dtype=torch.float64, device=dev, generator=g)
        G, _ = kernels.pod_gram(Vs, normalize=False)
        kernels.rom_difference(out, srom, Vs)
        torch.cuda.synchronize()
        c0, c1, c2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        c0.record()
        err = kernels.rom_difference(out, srom, Vs)
        c1.record()
        summ = kernels.errors_summary(err)
        c2.record()
        torch.cuda.synchronize()
        t_est, t_sum = c0.elapsed_time(c1), c1.elapsed_time(c2)
        rows = float(out.shape[0]) * nt
        cert = {"rows": int(rows), "estimator_ms": t_est, "summary_ms": t_sum,
                "estimator_GBps": 8.0 * rows * (2 * N + 1) / t_est * 1e-6,
                "estimator_frac_of_hbm_peak": 8.0 * rows * (2 * N + 1) / t_est * 1e-6 / hbm_peak_gbs(),
                "bound": "hbm (about Nhat^2 flop on the triangular factor vs 8 (N + Nhat + 1) bytes per row: 4 flop/B, FP64 ridge 5.7); "
                         "time includes the Gram matrix of V_srom and its Cholesky factor",
                "max_err": float(err.max().item())}
        if rank == 0 and world == 1 and not a.skip_cpu:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import romtime_oracle as orc_mod

            nb, ns = 64, min(nt, 100)
            hu, hs, hv = out[:nb, :ns].cpu().numpy(), srom[:nb, :ns].cpu().numpy(), Vs.cpu().numpy()
            t0 = time.perf_counter()
            ref = orc_mod.rom_difference_batch(hu, hs, hv)
            wall = time.perf_counter() - t0
            cert["cpu_rows_per_s"] = nb * ns / wall
            cert["cpu_sample"] = f"oracle rom_difference_batch (utils.py:173-212 per row), {nb} parameters x {ns} steps, 1 core, {wall:.1f} s"
            cert["gpu_rows_per_s"] = rows / (t_est * 1e-3)
            cert["parity_max_rel"] = float(abs(err[:nb, :ns].cpu().numpy() - ref).max() / abs(ref).max())
        del srom, err, summ, Vs, out, cert_u
        torch.cuda.empty_cache()
    except Exception as exc:  # report, never hide
        cert = {"error": repr(exc)}

    pod = None
    if not a.skip_pod:
        try:
            pod = pod_leg(a, torch, kernels, world, rank)
        except Exception as exc:  # report, never hide
            pod = {"error": repr(exc)}

    c4 = None
    if not a.skip_pod and world == 1:
        try:
            c4 = c4_leg(a, torch, kernels, not a.skip_cpu)
        except Exception as exc:  # report, never hide
            c4 = {"error": repr(exc)}

    tw = None
    if not a.skip_pod and world == 1:
        try:
            tw = treewalk_leg(a, torch, not a.skip_cpu)
        except Exception as exc:  # report, never hide
            tw = {"error": repr(exc)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_config(a, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(Bl * 3 * 8) * world,
                    "d2h_bytes_per_step": int(Bl * nt * N * 8) * world, "steps": e2e_steps,
                    "api": "kernels.online_sweep_to_host = RomConstructor*.solve_batch(..., to_host=True): 100-step segments, every "
                           "finished time slab copied to pinned host memory while the next segment runs",
                    "host_history_equals_device_history": e2e_matches},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu_baseline,
            "parity_1000_steps": parity, "sweep_m16": m16,
            "pod": pod, "mdeim_c4": c4, "treewalk": tw, "certification": cert, "singular_solves": bad,
        }
        emit(line, out_fd)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
