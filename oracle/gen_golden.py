"""Generate tests/golden/*.npz by running the UNMODIFIED reference (imported from
/root/reference through oracle/ref_harness.py) on seeded synthetic inputs.

Run in the build container only:   python oracle/gen_golden.py
The fixtures are committed; the GPU box never needs /root/reference.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_harness  # noqa: E402
from romtime_b200 import synthetic  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def tridiag_topology(n):
    rows, cols = [], []
    for i in range(n):
        for j in (i - 1, i, i + 1):
            if 0 <= j < n:
                rows.append(i)
                cols.append(j)
    return np.array(rows), np.array(cols)


def gen_orth():
    from romtime.rom.pod import orth

    rng = np.random.default_rng(11)
    out = {}
    cases = []
    # (name, matrix, kwargs)
    A = rng.standard_normal((300, 24)) @ np.diag(10.0 ** (-np.arange(24) / 4.0)) @ rng.standard_normal((24, 24))
    cases.append(("graded_num", A, dict(num=6)))
    cases.append(("graded_tol", A, dict(tol=1.0 - 1e-6)))
    cases.append(("graded_default", A, dict()))
    cases.append(("graded_nonorm", A, dict(num=5, normalize=False)))
    cases.append(("graded_vt", np.asfortranarray(A), dict(num=4, return_VT=True)))
    B = synthetic.pod_snapshots(2000, 64, seed=1003, modes=24, decay=4.0, noise=0.0)
    cases.append(("c3like_C", B, dict(num=8)))
    cases.append(("c3like_F", np.asfortranarray(B), dict(num=8)))
    Cm = synthetic.mdeim_snapshots(600, 40, seed=1004)
    cases.append(("c4like", Cm, dict(num=10, normalize=False)))
    lowrank = rng.standard_normal((150, 5)) @ rng.standard_normal((5, 30))
    cases.append(("lowrank_default", lowrank, dict()))
    cases.append(("wide_tol", rng.standard_normal((40, 12)), dict(tol=0.9)))
    names = []
    for name, M, kw in cases:
        res = orth(M, **kw)
        names.append(name)
        out[f"{name}/S"] = M
        out[f"{name}/order"] = np.array("F" if M.flags.f_contiguous and not M.flags.c_contiguous else "C")
        out[f"{name}/kw"] = np.array(repr(kw))
        out[f"{name}/Q"] = res[0]
        out[f"{name}/s"] = res[1]
        out[f"{name}/energy"] = res[2]
        if len(res) == 4:
            out[f"{name}/VT"] = res[3]
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(GOLD, "orth.npz"), **out)
    print("orth:", names)


def gen_greedy():
    from romtime.deim.deim import DiscreteEmpiricalInterpolation
    from romtime.rom.pod import orth

    rng = np.random.default_rng(12)
    out, names = {}, []

    def run(name, basis):
        d = DiscreteEmpiricalInterpolation(assemble=None, name=name)
        d.basis_fom = basis
        dofs, P = d.build_interpolation_mesh()
        out[f"{name}/basis"] = basis
        out[f"{name}/dofs"] = np.array(dofs, dtype=np.int64)
        out[f"{name}/PT_U"] = np.matmul(P.T, basis)
        names.append(name)

    Q, _, _ = orth(rng.standard_normal((500, 40)), num=12)
    run("random_orth", np.ascontiguousarray(Q))
    run("random_orth_F", np.asfortranarray(Q))
    # MDEIM-like basis with bit-identical duplicated rows (symmetric off-diagonals)
    S = synthetic.mdeim_snapshots(900, 60, seed=1004)
    Qm, _, _ = orth(S, num=16, normalize=False)
    run("mdeim_dup", Qm)
    # raw (non-orthogonal) basis, single column, and exact ties in column 0
    run("raw", rng.standard_normal((257, 9)))
    run("single", rng.standard_normal((33, 1)))
    tie = rng.standard_normal((64, 5))
    tie[10, :] = tie[3, :]
    tie[3, 0] = tie[10, 0] = 9.0
    run("ties", tie)
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(GOLD, "greedy.npz"), **out)
    print("greedy:", {n: out[f"{n}/dofs"][:6].tolist() for n in names})


def gen_project_interp():
    from romtime.deim.deim import DiscreteEmpiricalInterpolation
    from romtime.deim.mdeim import MatrixDiscreteEmpiricalInterpolation

    rng = np.random.default_rng(13)
    Nh, N, m = 120, 7, 9
    V, _ = np.linalg.qr(rng.standard_normal((Nh, N)))
    out = {"V": V}
    # DEIM (vector)
    Uf, _ = np.linalg.qr(rng.standard_normal((Nh, m)))
    coefs = rng.standard_normal((m, 4))

    def assemble_vec(mu, t, entries=None):
        full = Uf @ (coefs @ np.array([1.0, mu["a"], t, mu["a"] * t]))
        if entries is None:
            return full
        return np.array([full[e[0]] for e in entries])

    d = DiscreteEmpiricalInterpolation(assemble=assemble_vec, name="vec")
    d.basis_fom = Uf
    dofs, P = d.build_interpolation_mesh()
    d.store_dofs(dofs)
    d.PT_U = np.matmul(P.T, Uf)
    d.project_basis(V)
    mu, t = {"a": 0.7}, 0.3
    out["deim/basis_fom"] = Uf
    out["deim/dofs"] = np.array(dofs, dtype=np.int64)
    out["deim/PT_U"] = d.PT_U
    out["deim/basis_rom"] = d.basis_rom
    out["deim/f_local"] = assemble_vec(mu, t, d.dofs)
    out["deim/thetas"] = d.compute_thetas(out["deim/f_local"])
    out["deim/interp_fom"] = d.interpolate(mu, t, which="fom")
    out["deim/interp_rom"] = d.interpolate(mu, t, which="rom")
    # MDEIM (matrix): tridiagonal topology
    rows, cols = tridiag_topology(Nh)
    nnz = len(rows)
    Um, _ = np.linalg.qr(rng.standard_normal((nnz, m)))
    Um[0, :] = 0.0  # reference zeroes snapshot row 0 (deim/deim.py:388-389)

    def assemble_mat(mu, t, entries=None):
        full = Um @ (coefs @ np.array([1.0, mu["a"], t, mu["a"] * t]))
        look = {(int(r), int(c)): k for k, (r, c) in enumerate(zip(rows, cols))}
        return np.array([full[look[(int(e[0]), int(e[1]))]] for e in entries])

    md = MatrixDiscreteEmpiricalInterpolation(assemble=assemble_mat, name="mat")
    md.basis_fom = Um
    md.rows, md.cols = rows, cols
    dofs, P = md.build_interpolation_mesh()
    md.store_dofs(dofs)
    md.PT_U = np.matmul(P.T, Um)
    md.project_basis(V)
    out["mdeim/basis_fom"] = Um
    out["mdeim/rows"] = rows
    out["mdeim/cols"] = cols
    out["mdeim/dofs"] = np.array(dofs, dtype=np.int64)
    out["mdeim/dofs_rc"] = np.array(md.dofs, dtype=np.int64)
    out["mdeim/PT_U"] = md.PT_U
    out["mdeim/basis_rom"] = md.basis_rom
    out["mdeim/f_local"] = assemble_mat(mu, t, md.dofs)
    out["mdeim/interp_rom"] = md.interpolate(mu, t, which="rom")
    np.savez_compressed(os.path.join(GOLD, "project_interp.npz"), **out)
    print("project/interp: deim dofs", dofs[:5], "basis_rom", md.basis_rom.shape)


build_reference_rom = ref_harness.build_reference_rom


def gen_solve():
    rng = np.random.default_rng(14)
    out, names = {}, []
    for name, kw in (
        ("n8m6", dict(N=8, m=6, seed=21)),
        ("n16m10_pivot", dict(N=16, m=10, m_rhs=7, seed=22, mass_identity=0.15, coupling=1.0)),
    ):
        model = synthetic.make_piston_model(**kw)
        N = model["N"]
        V, _ = np.linalg.qr(rng.standard_normal((40, N)))
        mus = synthetic.sample_mus(3, seed=kw["seed"])
        dt, nt = 1e-3, 25
        out[f"{name}/kw"] = np.array(repr(kw))
        out[f"{name}/mus"] = mus
        out[f"{name}/dt"] = np.array(dt)
        out[f"{name}/nt"] = np.array(nt)
        for op in synthetic.OPS:
            for key in ("basis_rom", "PT_U", "C", "T"):
                val = model[op][key]
                if val is not None:
                    out[f"{name}/{op}/{key}"] = val
        for bdf in ("1", "2"):
            for direct in (True, False):
                sols = []
                for mu_row in mus:
                    rom = build_reference_rom(model, V, bdf, dt, nt, direct)
                    mu = {"a0": mu_row[0], "omega": mu_row[1], "delta": mu_row[2]}
                    rom.solve(mu=mu, step="online")
                    sols.append(np.array(rom.solutions.rom))
                tag = f"{name}/bdf{bdf}_{'direct' if direct else 'gmres'}"
                out[tag] = np.stack(sols)  # (B, N, nt)
        names.append(name)
    out["names"] = np.array(names)
    np.savez_compressed(os.path.join(GOLD, "solve.npz"), **out)
    print("solve:", names, {k: v.shape for k, v in out.items() if "bdf" in k})


def _walk_params(ts, **kw):
    base = {"ts": ts, "num_snapshots": None}
    base.update(kw)
    return base


def gen_treewalk():
    """Offline stage through the reference's own run(): two-/three-level POD tree walk, greedy mesh,
    projection and online evaluation, on the FEniCS-free problems of oracle/synthetic_problems.py."""
    from romtime.deim.deim import DiscreteEmpiricalInterpolation
    from romtime.deim.mdeim import MatrixDiscreteEmpiricalInterpolation
    from romtime.deim.nonlinear import MatrixDiscreteEmpiricalInterpolationNonlinear
    from romtime.rom.rom import RomConstructorNonlinear
    import synthetic_problems as sprob

    out = {}
    ts = [float(t) for t in np.linspace(0.05, 1.0, 12)]
    space = sprob.mu_space(5, seed=41)
    online = sprob.mu_space(3, seed=42)
    ts_online = [0.11, 0.52, 0.93]
    rng = np.random.default_rng(15)

    # ---- DEIM: fixed sizes, then tolerances (ragged time-basis widths)
    vp = sprob.VectorProblem()
    V = np.linalg.qr(rng.standard_normal((vp.Nh, 6)))[0]
    out["V_vec"] = V
    for tag, params in (("deim_num", _walk_params(ts, num_mu=7, num_time=5)),
                        ("deim_tol", _walk_params(ts, tol_mu=1.0 - 1e-9, tol_time=1.0 - 1e-7))):
        d = DiscreteEmpiricalInterpolation(assemble=vp.assemble, name=tag, tree_walk_params=params)
        d.setup(rnd=0)
        d.run(mu_space=[dict(m) for m in space])
        d.project_basis(V)
        d.evaluate(ts=ts_online, mu_space=[dict(m) for m in online])
        out[f"{tag}/basis_fom"] = d.basis_fom
        out[f"{tag}/sigmas"] = d.sigmas
        out[f"{tag}/dofs"] = np.array([e[0] for e in d.dofs], dtype=np.int64)
        out[f"{tag}/PT_U"] = d.PT_U
        out[f"{tag}/basis_rom"] = d.basis_rom
        out[f"{tag}/widths_time"] = np.array([d.report["offline"]["basis-shape-time"][i] for i in range(len(space))])
        out[f"{tag}/width_after_walk"] = np.array(d.report["offline"]["basis-shape-after-tree-walk"])
        out[f"{tag}/errors"] = np.array([d.errors_rom[i] for i in range(len(online))])
        out[f"{tag}/interp_fom"] = d.interpolate(online[0], 0.37, which="fom")
        out[f"{tag}/interp_rom"] = d.interpolate(online[0], 0.37, which="rom")

    # ---- MDEIM
    mp = sprob.MatrixProblem()
    Vm = np.linalg.qr(rng.standard_normal((mp.Nh, 5)))[0]
    out["V_mat"] = Vm
    md = MatrixDiscreteEmpiricalInterpolation(assemble=mp.assemble, name="mdeim", tree_walk_params=_walk_params(ts, num_mu=8, num_time=5))
    md.grid = None
    from romtime.rom.base import Reductor
    Reductor.setup(md, rnd=0)
    md.rows, md.cols = md.get_matrix_topology(mu=space[0], t=1.0)
    md.run(mu_space=[dict(m) for m in space])
    md.project_basis(Vm)
    md.evaluate(ts=ts_online, mu_space=[dict(m) for m in online])
    out["mdeim/rows"] = np.array(md.rows, dtype=np.int64)
    out["mdeim/cols"] = np.array(md.cols, dtype=np.int64)
    out["mdeim/basis_fom"] = md.basis_fom
    out["mdeim/sigmas"] = md.sigmas
    out["mdeim/dofs_rc"] = np.array(md.dofs, dtype=np.int64)
    out["mdeim/PT_U"] = md.PT_U
    out["mdeim/basis_rom"] = md.basis_rom
    out["mdeim/errors"] = np.array([md.errors_rom[i] for i in range(len(online))])
    out["mdeim/interp_rom"] = md.interpolate(online[1], 0.41, which="rom")
    out["mdeim/interp_fom_data"] = md.interpolate(online[1], 0.41, which="fom").toarray()

    # ---- N-MDEIM (three-level walk) + truncate
    npb = sprob.NonlinearMatrixProblem()
    psi = npb.state_basis(3)
    nm = MatrixDiscreteEmpiricalInterpolationNonlinear(assemble=npb.assemble, name="nmdeim",
                                                       tree_walk_params=_walk_params(ts[:6], num_mu=9, num_time=6))
    Reductor.setup(nm, rnd=0)
    nm.rows, nm.cols = nm.get_matrix_topology(mu=space[0], t=1.0, u_n=psi[:, 0])
    nm.run(u_n=psi, mu_space=[dict(m) for m in space[:3]])
    nm.project_basis(Vm)
    u_probe = 0.2 + 0.1 * np.cos(2 * np.pi * npb.x)
    out["nmdeim/psi"] = psi
    out["nmdeim/u_probe"] = u_probe
    out["nmdeim/rows"] = np.array(nm.rows, dtype=np.int64)
    out["nmdeim/cols"] = np.array(nm.cols, dtype=np.int64)
    out["nmdeim/basis_fom"] = nm.basis_fom
    out["nmdeim/sigmas"] = nm.sigmas
    out["nmdeim/dofs_rc"] = np.array(nm.dofs, dtype=np.int64)
    out["nmdeim/basis_rom"] = nm.basis_rom
    out["nmdeim/interp_rom"] = nm.interpolate(online[2], 0.3, u_n=u_probe, which="rom")
    tr = nm.truncate(2)
    out["nmdeim/trunc_dofs_rc"] = np.array(tr.dofs, dtype=np.int64)
    out["nmdeim/trunc_PT_U"] = tr.PT_U

    # ---- solution reduced basis (RomConstructor.build_reduced_basis)
    fom = sprob.SnapshotFom()
    rom = RomConstructorNonlinear(fom=fom, grid=None, name="rb")
    rom.setup(rnd=0)
    import io, contextlib
    with contextlib.redirect_stdout(io.StringIO()):
        sols = rom.build_reduced_basis(mu_space=[dict(m) for m in space], num_basis=7,
                                       tolerances={"tol_time": 1.0 - 1e-10})
    out["rb/basis"] = rom.basis
    out["rb/basis_nonlinear"] = rom.basis_nonlinear
    out["rb/sigmas_mu"] = rom.report["offline"]["spectrum-mu"]
    out["rb/energy_mu"] = rom.report["offline"]["energy-mu"]
    out["rb/widths_time"] = np.array([rom.report["offline"]["basis-shape-time"][i] for i in range(len(space))])
    out["rb/width_after_walk"] = np.array(rom.report["offline"]["basis-shape-after-tree-walk"])
    out["rb/fom_solution_0"] = sols[0]
    uh = fom.field(space[1], 0.3)
    out["rb/uh"] = uh
    out["rb/to_rom_vector"] = rom.basis.T.dot(uh)
    out["rb/to_fom_vector"] = rom.to_fom_vector(rom.basis.T.dot(uh))
    A = mp.matrix(space[0], 0.5)
    from romtime.utils import project_csr
    Vrb = np.linalg.qr(rng.standard_normal((mp.Nh, 4)))[0]
    out["rb/V_small"] = Vrb
    out["rb/project_csr"] = project_csr(A, Vrb)
    np.savez_compressed(os.path.join(GOLD, "treewalk.npz"), **out)
    print("treewalk:", {k: v.shape for k, v in out.items() if k.endswith("basis_fom") or k.endswith("basis")})


def gen_certification():
    """S-ROM estimator and error summaries: the reference's own compute_rom_difference
    (utils.py:173-212), Reductor._compute_error (rom/base.py:52-73) and create_errors_summary (:154-163)."""
    ref = ref_harness.load_reference()
    from romtime.rom.base import Reductor
    from romtime.utils import compute_rom_difference

    rng = np.random.default_rng(1006)
    out = {}
    cases = [("ortho", 300, 10, 14, 6, 20), ("general", 257, 33, 40, 5, 17), ("equal", 128, 64, 64, 4, 9),
             ("wide", 500, 20, 96, 3, 12)]
    out["names"] = np.array([c[0] for c in cases])
    for name, Nh, N, Nhat, B, nt in cases:
        V = rng.standard_normal((Nh, Nhat))
        if name == "ortho":
            V, _ = np.linalg.qr(V)
        uN = rng.standard_normal((B, nt, N))
        uS = rng.standard_normal((B, nt, Nhat))
        uS[:, :, :N] += uN  # S-ROM = ROM + a correction, as in the certification loop
        err = np.array([[compute_rom_difference(uN[b, n], uS[b, n], V) for n in range(nt)] for b in range(B)])
        red = Reductor(grid=None)
        for b in range(B):
            red.errors_rom[b] = list(err[b])
        red.create_errors_summary()
        summ = red.summary_errors[["mean", "median", "max", "min"]].to_numpy(dtype=float)
        rms = np.array([Reductor._compute_error(uS[b, n], np.append(uN[b, n], (Nhat - N) * [0.0])) for b in range(B)
                        for n in range(nt)]).reshape(B, nt)
        for k, v in (("V", V), ("uN", uN), ("uS", uS), ("err", err), ("summary", summ), ("rms", rms)):
            out[f"{name}/{k}"] = v
    np.savez_compressed(os.path.join(GOLD, "certification.npz"), **out)
    print("certification:", {k: v.shape for k, v in out.items() if k.endswith("err")})


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    ref_harness.load_reference()
    gen_orth()
    gen_greedy()
    gen_project_interp()
    gen_solve()
    gen_treewalk()
    gen_certification()
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))
