"""Online stage: theta solve, reduced-operator assembly and the batched BDF sweep against the
reference-generated golden solutions (RomConstructorNonlinear.solve with the algebraic-solver hook
swapped for a direct solve) and the CPU oracle.  Tolerance (north_star): 1e-10 relative in fp64."""
import ast

import numpy as np
import pytest

import romtime_oracle as orc
from romtime_b200 import synthetic

pytestmark = pytest.mark.gpu


def relerr(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def golden_model(z, name):
    kw = ast.literal_eval(str(z[f"{name}/kw"]))
    model = synthetic.make_piston_model(**kw)
    for op in synthetic.OPS:  # the fixture's own tables are authoritative
        for key in ("basis_rom", "PT_U", "C", "T"):
            k = f"{name}/{op}/{key}"
            if k in z.files:
                assert np.array_equal(model[op][key], z[k]), k
    return model


@pytest.mark.parametrize("m", [1, 5, 16, 33, 128])
def test_invert_and_theta(m):
    from romtime_b200 import kernels

    rng = np.random.default_rng(m)
    A = np.eye(m) + 0.3 * rng.standard_normal((m, m)) / np.sqrt(m)
    A[[0, m - 1]] = A[[m - 1, 0]]  # force pivoting
    Ainv = kernels.invert(A)
    assert relerr(Ainv.cpu().numpy(), np.linalg.inv(A)) < 1e-12
    F = rng.standard_normal((77, m))
    th = kernels.theta_batch(Ainv, F).cpu().numpy()
    ref = np.stack([orc.compute_thetas(A, f) for f in F])
    assert relerr(th, ref) < 1e-12


def test_golden_theta_and_interpolation(golden):
    from romtime_b200 import kernels

    z = golden("project_interp")
    for kind in ("deim", "mdeim"):
        inv = kernels.invert(z[f"{kind}/PT_U"])
        th = kernels.theta_batch(inv, z[f"{kind}/f_local"][None, :])
        if kind == "deim":
            assert relerr(th.cpu().numpy()[0], z["deim/thetas"]) < 1e-12
        out = kernels.assemble_batch(z[f"{kind}/basis_rom"], th).cpu().numpy()[0]
        ref = z[f"{kind}/interp_rom"]
        assert relerr(out.reshape(ref.shape), ref) < 1e-12


def test_singular_inverse_raises():
    from romtime_b200 import kernels

    with pytest.raises(np.linalg.LinAlgError):
        kernels.invert(np.zeros((4, 4)))


@pytest.mark.parametrize("N,m,B", [(8, 6, 5), (64, 16, 300), (20, 7, 130)])
def test_assemble_batch_vs_oracle(N, m, B):
    from romtime_b200 import kernels

    rng = np.random.default_rng(N + m)
    basis = np.asfortranarray(rng.standard_normal((N * N, m)))
    P = np.eye(m) + 0.2 * rng.standard_normal((m, m))
    F = rng.standard_normal((B, m))
    th = kernels.theta_batch(kernels.invert(P), F)
    out = kernels.assemble_batch(basis, th).cpu().numpy()
    for b in (0, B // 2, B - 1):
        ref = orc.interpolate_rom_matrix(basis, P, F[b], N)
        assert relerr(out[b].reshape(N, N), ref) < 1e-12


@pytest.mark.parametrize("bdf", [1, 2])
def test_golden_sweep(golden, bdf):
    from romtime_b200 import kernels

    z = golden("solve")
    for name in z["names"]:
        model = golden_model(z, name)
        mus, dt, nt = z[f"{name}/mus"], float(z[f"{name}/dt"]), int(z[f"{name}/nt"])
        dm = kernels.DeviceRomModel(model)
        u, info = kernels.online_sweep(dm, mus, nt, dt, bdf=bdf)
        u = u.cpu().numpy().transpose(0, 2, 1)  # (B, N, nt) like RomSolutionsStorage.rom
        assert not info.cpu().numpy().any()
        ref = z[f"{name}/bdf{bdf}_direct"]
        assert relerr(u, ref) < 1e-10, (name, relerr(u, ref))
        # stock GMRES(1e-10) reference: reported deviation, looser bound (SURVEY hard part 4)
        assert relerr(u, z[f"{name}/bdf{bdf}_gmres"]) < 1e-6
        last, _ = kernels.online_sweep(dm, mus, nt, dt, bdf=bdf, history=False)
        assert np.array_equal(last.cpu().numpy(), u[:, :, -1])


@pytest.mark.parametrize("N,m,m_rhs,B,nt", [(64, 16, 16, 200, 12), (64, 128, 128, 140, 6), (33, 20, 9, 7, 10), (50, 12, 12, 9, 8),
                                               (63, 10, 10, 5, 6)])
def test_sweep_vs_oracle(N, m, m_rhs, B, nt):
    from romtime_b200 import kernels

    model = synthetic.make_piston_model(N=N, m=m, m_rhs=m_rhs, seed=N + m)
    mus = synthetic.sample_mus(B, seed=7)
    dm = kernels.DeviceRomModel(model)
    u, info = kernels.online_sweep(dm, mus, nt, 1e-3, bdf=2)
    u = u.cpu().numpy()
    assert not info.cpu().numpy().any()
    for b in (0, B // 3, B - 1):
        ref = orc.rom_sweep(model, mus[b], nt, 1e-3, bdf2=True)
        assert relerr(u[b].T, ref) < 1e-10


def test_sweep_multi_chunk_consistency():
    """More parameters than one chunk: every chunk (and both sides of the chunk boundary) must reproduce the
    single-parameter run."""
    import torch

    from romtime_b200 import kernels

    N, B = 64, 18000
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    chunk = int(kernels.load_library().rt_sweep_chunk(B, N, sms))
    assert 0 < chunk < B, chunk
    model = synthetic.make_piston_model(N=N, m=16, seed=3)
    mus = synthetic.sample_mus(B, seed=11)
    dm = kernels.DeviceRomModel(model)
    last, info = kernels.online_sweep(dm, mus, 5, 1e-3, bdf=2, history=False)
    last = last.cpu().numpy()
    assert not info.cpu().numpy().any()
    for b in (0, chunk - 1, chunk, B - 1):
        ref = orc.rom_sweep(model, mus[b], 5, 1e-3, bdf2=True)
        assert relerr(last[b], ref[:, -1]) < 1e-10
    # host-delivered history: chunked sweep with the device-to-host copy of a chunk overlapping the next chunk
    hist, info_h = kernels.online_sweep_to_host(dm, mus, 5, 1e-3, bdf=2)
    assert hist.shape == (B, 5, N) and not hist.is_cuda
    assert not info_h.cpu().numpy().any()
    assert np.array_equal(hist[:, -1, :].numpy(), last)


@pytest.mark.parametrize("N,m,B,nt,seg", [(64, 16, 300, 23, 5), (64, 128, 150, 12, 4), (20, 7, 40, 30, 7), (100, 9, 6, 9, 2)])
def test_segmented_sweep_is_bitwise_the_uninterrupted_one(N, m, B, nt, seg):
    """rt_online_sweep_segment_f64 carries the state (u_n, u_n-1, the accumulated time, the BDF start-up) from one
    segment to the next: any segmentation reproduces the uninterrupted sweep bit for bit -- also what makes ten
    100-step bench segments one 1000-step sweep.  The host-delivered form ships every finished time slab."""
    import torch

    from romtime_b200 import kernels

    model = synthetic.make_piston_model(N=N, m=m, seed=N + 3)
    mus = synthetic.sample_mus(B, seed=13)
    dm = kernels.DeviceRomModel(model)
    whole, info = kernels.online_sweep(dm, mus, nt, 1e-3, bdf=2)
    seen = []
    parts, info_s = kernels.online_sweep_segments(dm, mus, nt, 1e-3, bdf=2, segment=seg, on_segment=lambda a, b: seen.append((a, b)))
    assert seen[0] == (0, seg) and seen[-1][1] == nt and all(x[1] == y[0] for x, y in zip(seen, seen[1:]))
    assert torch.equal(whole, parts) and torch.equal(info, info_s)
    host, _ = kernels.online_sweep_to_host(dm, mus, nt, 1e-3, bdf=2, segment=seg)
    assert not host.is_cuda and torch.equal(host, whole.cpu())
    b1, _ = kernels.online_sweep_segments(dm, mus, nt, 1e-3, bdf=1, segment=seg)
    w1, _ = kernels.online_sweep(dm, mus, nt, 1e-3, bdf=1)
    assert torch.equal(b1, w1)


@pytest.mark.parametrize("N", [5, 32, 33, 40, 63, 64, 100])
def test_dense_solve_needs_pivoting(N):
    """The per-parameter solver (LU for N <= 32 and N > 64, blocked Gauss-Jordan on DMMA tiles in between) must
    pivot: a row-permuted diagonally dominant matrix has zeros (or tiny entries) on its diagonal."""
    from romtime_b200.rom.rom import _dense_solve

    rng = np.random.default_rng(N)
    A = rng.standard_normal((N, N)) + N * np.eye(N)
    perm = rng.permutation(N)
    K = A[perm]  # K[i] = A[perm[i]]: the large entries sit at (i, perm[i])
    b = rng.standard_normal(N)
    x = _dense_solve(K, b)
    assert relerr(x, np.linalg.solve(K, b)) < 1e-10
    # wide dynamic range inside every column: the pivot must be the large entry, not the first one
    K2 = K * np.logspace(-6, 6, N)[rng.permutation(N)][:, None]
    x2 = _dense_solve(K2, b)
    assert relerr(x2, np.linalg.solve(K2, b)) < 1e-10


@pytest.mark.parametrize("N", [20, 48, 64, 90])
def test_dense_solve_singular_raises(N):
    from romtime_b200.rom.rom import _dense_solve

    rng = np.random.default_rng(N + 1)
    K = rng.standard_normal((N, N))
    K[:, N // 2] = 0.0  # exactly singular: a zero pivot column whatever the row order
    with pytest.raises(np.linalg.LinAlgError):
        _dense_solve(K, rng.standard_normal(N))


def test_sweep_is_bitwise_reproducible():
    """Race check without compute-sanitizer (closed on this pool): the solve kernel's look-ahead (panel warp ahead of
    the trailing warps, double-buffered W, one barrier per panel) must give bit-identical results run after run, with
    every SM oversubscribed (more CTAs than resident slots)."""
    import torch

    from romtime_b200 import kernels

    model = synthetic.make_piston_model(N=64, m=16, seed=21)
    mus = synthetic.sample_mus(6000, seed=22)
    dm = kernels.DeviceRomModel(model)
    first, info = kernels.online_sweep(dm, mus, 6, 1e-3, bdf=2)
    assert not info.any().item()
    for _ in range(4):
        again, _ = kernels.online_sweep(dm, mus, 6, 1e-3, bdf=2)
        assert torch.equal(first, again)
    ref = orc.rom_sweep(model, mus[5999], 6, 1e-3, bdf2=True)
    assert relerr(first[5999].cpu().numpy().T, ref) < 1e-10


@pytest.mark.parametrize("sizes", [(70, 96), (100, 128), (65, 96, 70)])
def test_dense_solve_growing_sizes_in_one_process(sizes):
    """The shared-memory LU (N > 64) opts in to more than 48 KB of dynamic shared memory per reduced size: a larger
    N after a smaller one of the same size class must raise the opt-in (round-1 bug: a per-class static flag)."""
    from romtime_b200.rom.rom import _dense_solve

    for N in sizes:
        rng = np.random.default_rng(N)
        K = rng.standard_normal((N, N)) + N * np.eye(N)
        b = rng.standard_normal(N)
        assert relerr(_dense_solve(K, b), np.linalg.solve(K, b)) < 1e-10


def test_step_reuses_staged_model_and_sees_new_tables():
    """rt_online_step_f64 keeps the staged model between calls (rt_rom_model.generation); a NEW DeviceRomModel --
    even one whose tensors land on the same addresses -- must not hit the stale entry."""
    import torch

    from romtime_b200 import kernels

    N, m, B = 16, 6, 5
    outs = []
    for seed in (31, 32):
        model = synthetic.make_piston_model(N=N, m=m, seed=seed)
        dm = kernels.DeviceRomModel(model)
        rng = np.random.default_rng(5)
        F = rng.standard_normal((B, dm.qsum))
        Un = torch.zeros((B, N), dtype=torch.float64, device="cuda")
        Unm1 = torch.zeros_like(Un)
        info = torch.zeros(B, dtype=torch.int32, device="cuda")
        for n in range(3):  # same model three times: steps 1, 2 run from the cached staging
            kernels.online_step(dm, F, Un, Unm1, info, 1e-3, 2, n)
        ref_u, ref_um1 = np.zeros((B, N)), np.zeros((B, N))
        for n in range(3):
            new = np.stack([orc.online_step_entries(model, F[b], ref_u[b], ref_um1[b], 1e-3, 2, n) for b in range(B)])
            ref_um1, ref_u = ref_u, new
        assert relerr(Un.cpu().numpy(), ref_u) < 1e-10
        outs.append(Un.cpu().numpy())
        del dm, Un, Unm1, info
    assert relerr(outs[0], outs[1]) > 1e-3  # different models, different answers


@pytest.mark.parametrize("N,m,m_rhs,B", [(64, 128, 128, 300), (40, 24, 7, 95), (64, 16, 16, 33), (33, 128, 40, 64)])
def test_fused_prestep_matches_the_four_kernel_path(N, m, m_rhs, B, monkeypatch):
    """k_prestep (state + T u* + sampled entries + forcing in one DMMA kernel) against k_state / k_tu / k_entries / k_forcing:
    same sweep, the two pre-step paths; ragged last block of 32 parameters, N below 64, bases below 128 modes."""
    from romtime_b200 import kernels

    model = synthetic.make_piston_model(N=N, m=m, m_rhs=m_rhs, seed=N + m + 7)
    dm = kernels.DeviceRomModel(model)
    mus = synthetic.sample_mus(B, seed=5)
    monkeypatch.setenv("RT_FUSED_PRESTEP", "0")
    ref, info0 = kernels.online_sweep(dm, mus, 12, 2e-3, bdf=2)
    monkeypatch.setenv("RT_FUSED_PRESTEP", "1")
    out, info1 = kernels.online_sweep(dm, mus, 12, 2e-3, bdf=2)
    assert not info0.any().item() and not info1.any().item()
    scale = ref.abs().amax(dim=(1, 2), keepdim=True)
    assert ((out - ref).abs() / scale).max().item() < 1e-11
