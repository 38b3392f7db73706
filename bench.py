#!/usr/bin/env python
"""bench.py -- headline benchmark of the romtime hot path on B200 (contract: see the task prompt).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[4] + configs[2]):
  * online: batched ROM sweep, B = 65,536 parameters (sharded over the ranks: strong scaling),
    reduced dimension N = 64, five MDEIM operators + one DEIM vector with m = 128 collateral modes,
    BDF2.  One bench "step" = the whole batch advanced by --nt-per-step implicit time steps
    (default 100; the 1000-step sweep of the config is ten such steps, cost per time step is constant).
    metric = ROM param*timesteps/s.
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
UNIT = "param*timesteps/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--params", type=int, default=65536)
    ap.add_argument("--nt-per-step", type=int, default=100)
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
    }


# ------------------------------------------------------------------------------- reference arm
def reference_arm(a, out_fd):
    """CPU arm: the oracle port of RomConstructorNonlinear.solve (the reference itself needs FEniCS and
    cannot run here) on all host cores; each step = `cores` parameters x a bounded number of time steps."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_bench

    cores = os.cpu_count() or 1
    nt = 300  # bounded sample per bench step: a few seconds on every core
    for _ in range(max(a.warmup, 0) and 1):
        cpu_bench.online_sweep_rate(cores, 64, a.m, 2)
    t0 = time.perf_counter()
    units = 0
    for _ in range(a.steps):
        rate, wall, _ = cpu_bench.online_sweep_rate(cores, 64, a.m, nt)
        units += cores * nt
    total = time.perf_counter() - t0
    value = units / total
    sample = f"{cores} parameters x {nt} BDF2 steps per bench step, one worker process per core, BLAS threads 1"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": total / a.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": workload_config(a, a.gpus),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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
        Q, s, energy, VT, _ = podmod.orth_device(S, num=64)
        e[3].record()
        torch.cuda.synchronize()
        t_gram, t_eig, t_orth = e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), e[2].elapsed_time(e[3])
        if best is None or t_orth < best[2]:
            best = (t_gram, t_eig, t_orth)
    t_gram, t_eig, t_orth = best
    peak = kernels.fp64_peak_tflops()
    flops = float(Nh) * n * (n + 1)
    QtQ = (Q.t() @ Q - torch.eye(Q.shape[1], dtype=torch.float64, device=dev)).abs().max().item()
    res = {
        "shape": [Nh, n], "num": 64, "svd_time_ms": t_orth, "gram_ms": t_gram, "eig_ms": t_eig,
        "lift_and_rest_ms": max(t_orth - t_gram - t_eig, 0.0), "gram_tflops_sym_half": flops / t_gram * 1e-9,
        "gram_frac_of_fp64_peak": flops / t_gram * 1e-9 / peak, "S_GBps_whole_orth": 8.0 * Nh * n / t_orth * 1e-6,
        "eig_sweeps": sweeps, "QtQ_minus_I_max": QtQ, "energy_last": float(energy[-1]),
        "singular_values_head": [float(x) for x in s[:4]],
    }
    del S, G, W, Q
    torch.cuda.empty_cache()
    if rank == 0 and not a.skip_cpu:
        # CPU baseline: the oracle's orth (rom/pod.py:7-62: column norms + LAPACK gesvd, all BLAS threads) on
        # a bounded sample of the same family; the full-size figure is an O(Nh n^2) EXTRAPOLATION and is labelled as one
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import numpy as np
        import romtime_oracle as orc_mod
        from romtime_b200 import synthetic

        nh_s, n_s = 16000, 1024
        Ss = synthetic.pod_snapshots(nh_s, n_s, seed=1003)
        t0 = time.perf_counter()
        Qc, sc, _ = orc_mod.orth(Ss, num=64)
        wall = time.perf_counter() - t0
        res["cpu_orth_sample_ms"] = wall * 1e3
        res["cpu_orth_sample"] = f"oracle orth (scipy gesvd) on {nh_s} x {n_s}, {os.cpu_count()} host cores, BLAS default threads"
        res["cpu_orth_extrapolated_full_ms"] = wall * 1e3 * (Nh / nh_s) * (n / n_s) ** 2
        res["cpu_orth_note"] = "extrapolated by Nh n^2 from the sample, not measured at full size"
        # the device path on the same sample, for a like-for-like parity figure
        Qd, sd, _, _, _ = podmod.orth_device(Ss, num=64)
        res["sample_singular_values_max_rel"] = float(np.abs(sd[:64] - sc[:64]).max() / sc[0])
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
           "sigma_m_over_sigma_1": float(s[m - 1] / s[0]), "greedy_ms": t_gr,
           "greedy_algorithmic_GBps": alg / t_gr * 1e-6, "greedy_algorithmic_frac_of_hbm_peak": alg / t_gr * 1e-6 / hbm,
           "greedy_blocked_traffic_GBps": blk / t_gr * 1e-6, "greedy_us_per_point": 1e3 * t_gr / m,
           "hbm_peak_GBps": hbm, "unique_dofs": int(torch.unique(dofs).numel()),
           "bound": "hbm by SURVEY.md 8d's algorithmic bytes (8 Nh m (m+1) / 2 = 13.2 GB -> 2.0 ms at peak); the blocked kernel "
                    "streams 8x less and is bound by its m grid-wide arg-max exchanges"}
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

    cpu_baseline = None
    if rank == 0 and world == 1 and not a.skip_cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import cpu_bench

        cores = os.cpu_count() or 1
        nt_cpu = 1000  # one full C5 trajectory per core: 10-20 s of CPU work
        rate, wall, _ = cpu_bench.online_sweep_rate(cores, 64, a.m, nt_cpu)
        cpu_baseline = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"oracle rom_sweep (NumPy port of rom/rom.py:430-555): {cores} parameters x {nt_cpu} "
                                  f"BDF2 steps, one process per core, BLAS threads 1, {wall:.1f} s"}

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
    mus_host = torch.from_numpy(synthetic.sample_mus(B, seed=1005)[lo:hi].copy()).pin_memory()
    mus = mus_host.to(dev)
    Bl = hi - lo

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        return kernels.online_sweep(dm, mus, nt, dt, bdf=2, history=True)

    for _ in range(a.warmup):
        out, info = step_resident()
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
        out, info = step_resident()
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
    hist_host = torch.empty((Bl, nt, N), dtype=torch.float64).pin_memory()

    def step_e2e():
        # public host-to-host call: mus from pinned host memory, the history lands in pinned host memory (the
        # device-to-host copy of a chunk overlaps the next chunk's sweep)
        kernels.online_sweep_to_host(dm, mus_host, nt, dt, bdf=2, out_host=hist_host)
        torch.cuda.synchronize()

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        step_e2e()
    barrier()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = float(B) * nt * a.steps / float(te.item())

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
                "traffic": 532.1e6 * chunk / 14208.0,
                "traffic_source": "dram bytes read+write per launch (532.1 MB at a 14208-parameter chunk, scaled to this chunk: K written once, M never stored), ncu --set full, profiles/r1i_sweep_ncu.md",
                "kernel_shares": {k: v[0] / total_prof for k, v in prof.items()}}

    # ---- certification leg: S-ROM estimator over every (parameter, step) of the last sweep
    cert = None
    try:
        g = torch.Generator(device=dev).manual_seed(1006)
        Vs = torch.linalg.qr(torch.randn((20000, N), dtype=torch.float64, device=dev, generator=g))[0]
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
        rows = float(Bl) * nt
        cert = {"rows": int(rows), "estimator_ms": t_est, "summary_ms": t_sum,
                "estimator_GBps": 8.0 * rows * (2 * N + 1) / t_est * 1e-6,
                "estimator_frac_of_hbm_peak": 8.0 * rows * (2 * N + 1) / t_est * 1e-6 / hbm_peak_gbs(),
                "bound": "hbm (about Nhat^2 flop on the triangular factor vs 8 (N + Nhat + 1) bytes per row: 4 flop/B, FP64 ridge 5.7); "
                         "time includes the Gram matrix of V_srom and its Cholesky factor",
                "max_err": float(err.max().item())}
        if rank == 0 and world == 1 and not a.skip_cpu:
            sys.path.insert(0, os.path.join(ROOT, "oracle"))
            import romtime_oracle as orc_mod

            nb = 64
            hu, hs, hv = out[:nb].cpu().numpy(), srom[:nb].cpu().numpy(), Vs.cpu().numpy()
            t0 = time.perf_counter()
            ref = orc_mod.rom_difference_batch(hu, hs, hv)
            wall = time.perf_counter() - t0
            cert["cpu_rows_per_s"] = nb * nt / wall
            cert["cpu_sample"] = f"oracle rom_difference_batch (utils.py:173-212 per row), {nb} parameters x {nt} steps, 1 core, {wall:.1f} s"
            cert["gpu_rows_per_s"] = rows / (t_est * 1e-3)
            cert["parity_max_rel"] = float(abs(err[:nb].cpu().numpy() - ref).max() / abs(ref).max())
        del srom, err, summ, Vs
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

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_config(a, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(Bl * 3 * 8) * world,
                    "d2h_bytes_per_step": int(Bl * nt * N * 8) * world},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu_baseline,
            "pod": pod, "mdeim_c4": c4, "certification": cert, "singular_solves": bad,
        }
        emit(line, out_fd)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
