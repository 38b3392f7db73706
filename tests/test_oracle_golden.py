"""CPU: the oracle (NumPy/SciPy restatement) against the golden vectors that the UNMODIFIED reference
produced in the build container (oracle/gen_golden.py).  This is what pins the oracle."""
import ast

import numpy as np

import romtime_oracle as orc
from romtime_b200 import synthetic


def test_orth_golden(golden):
    z = golden("orth")
    for name in z["names"]:
        S = z[f"{name}/S"]
        if str(z[f"{name}/order"]) == "F":
            S = np.asfortranarray(S)
        kw = ast.literal_eval(str(z[f"{name}/kw"]))
        res = orc.orth(S, **kw)
        assert res[0].shape == z[f"{name}/Q"].shape
        np.testing.assert_allclose(res[1], z[f"{name}/s"], rtol=1e-12, atol=1e-14 * z[f"{name}/s"][0])
        np.testing.assert_allclose(res[2], z[f"{name}/energy"], rtol=1e-12)
        sign = np.sign(np.sum(res[0] * z[f"{name}/Q"], axis=0))
        good = z[f"{name}/s"][: res[0].shape[1]] > 1e-6 * z[f"{name}/s"][0]
        np.testing.assert_allclose((res[0] * sign)[:, good], z[f"{name}/Q"][:, good], atol=1e-9)
        if len(res) == 4:
            np.testing.assert_allclose((res[3] * sign[:, None])[good], z[f"{name}/VT"][good], atol=1e-9)


def test_greedy_golden(golden):
    z = golden("greedy")
    for name in z["names"]:
        basis = z[f"{name}/basis"]
        dofs = orc.deim_greedy(basis)
        assert dofs == z[f"{name}/dofs"].tolist(), name
        assert np.array_equal(orc.interpolation_matrix(basis, dofs), z[f"{name}/PT_U"])
        P = orc.one_hot(basis.shape[0], dofs)
        assert np.array_equal(P.T @ basis, z[f"{name}/PT_U"])


def test_project_interp_golden(golden):
    z = golden("project_interp")
    V = z["V"]
    np.testing.assert_allclose(orc.deim_project(V, z["deim/basis_fom"]), z["deim/basis_rom"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(orc.compute_thetas(z["deim/PT_U"], z["deim/f_local"]), z["deim/thetas"], rtol=1e-13)
    np.testing.assert_allclose(orc.interpolate(z["deim/basis_rom"], z["deim/PT_U"], z["deim/f_local"]),
                               z["deim/interp_rom"], rtol=1e-13, atol=1e-15)
    np.testing.assert_allclose(orc.interpolate(z["deim/basis_fom"], z["deim/PT_U"], z["deim/f_local"]),
                               z["deim/interp_fom"], rtol=1e-13, atol=1e-15)
    B = orc.mdeim_project(z["mdeim/basis_fom"], z["mdeim/rows"], z["mdeim/cols"], V)
    np.testing.assert_allclose(B, z["mdeim/basis_rom"], rtol=0, atol=1e-15)
    M = orc.interpolate_rom_matrix(z["mdeim/basis_rom"], z["mdeim/PT_U"], z["mdeim/f_local"], V.shape[1])
    np.testing.assert_allclose(M, z["mdeim/interp_rom"], rtol=1e-13, atol=1e-15)
    assert orc.deim_greedy(z["mdeim/basis_fom"]) == z["mdeim/dofs"].tolist()


def test_solve_golden(golden):
    z = golden("solve")
    for name in z["names"]:
        kw = ast.literal_eval(str(z[f"{name}/kw"]))
        model = synthetic.make_piston_model(**kw)
        for op in synthetic.OPS:
            for key in ("basis_rom", "PT_U", "C", "T"):
                k = f"{name}/{op}/{key}"
                if k in z.files:
                    assert np.array_equal(model[op][key], z[k]), k
        mus, dt, nt = z[f"{name}/mus"], float(z[f"{name}/dt"]), int(z[f"{name}/nt"])
        for bdf in (1, 2):
            for b, mu in enumerate(mus):
                out = orc.rom_sweep(model, mu, nt, dt, bdf2=(bdf == 2), solver="direct")
                ref = z[f"{name}/bdf{bdf}_direct"][b]
                assert np.max(np.abs(out - ref)) <= 1e-12 * np.max(np.abs(ref)), (name, bdf, b)
            out = orc.rom_sweep(model, mus[0], nt, dt, bdf2=(bdf == 2), solver="gmres")
            ref = z[f"{name}/bdf{bdf}_gmres"][0]
            assert np.max(np.abs(out - ref)) <= 1e-9 * np.max(np.abs(ref))


def test_known_answer_properties():
    """Properties the reference's own (FEniCS-bound) tests assert, restated on synthetic inputs:
    DEIM is exact on span(U) (tests/test_deim.py:190,211); Q^T Q = I; energy[-1] = 1."""
    rng = np.random.default_rng(2)
    U, _ = np.linalg.qr(rng.standard_normal((400, 9)))
    dofs = orc.deim_greedy(U)
    PT_U = orc.interpolation_matrix(U, dofs)
    f = U @ rng.standard_normal(9)
    np.testing.assert_allclose(orc.interpolate(U, PT_U, f[dofs]), f, atol=1e-13)
    Q, s, e = orc.orth(rng.standard_normal((300, 20)))
    np.testing.assert_allclose(Q.T @ Q, np.eye(Q.shape[1]), atol=1e-13)
    assert abs(e[-1] - 1.0) < 1e-14
    assert np.isclose(orc.compute_rom_difference(np.ones(2), np.ones(3), np.eye(3)), 1 / np.sqrt(3))


def test_certification_golden(golden):
    """S-ROM estimator, RMS error and the per-parameter error summary against the reference's own functions."""
    z = golden("certification")
    for name in z["names"]:
        V, uN, uS = z[f"{name}/V"], z[f"{name}/uN"], z[f"{name}/uS"]
        np.testing.assert_allclose(orc.rom_difference_batch(uN, uS, V), z[f"{name}/err"], rtol=1e-13)
        np.testing.assert_allclose(orc.errors_summary(z[f"{name}/err"]), z[f"{name}/summary"], rtol=1e-14)
        B, nt, N = uN.shape
        pad = np.zeros_like(uS)
        pad[:, :, :N] = uN
        rms = np.array([[orc.rms_error(uS[b, n], pad[b, n]) for n in range(nt)] for b in range(B)])
        np.testing.assert_allclose(rms, z[f"{name}/rms"], rtol=1e-14)
