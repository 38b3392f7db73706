"""CPU: the closed-form P1 sampled-entry assembly (oracle/p1_forms.py, restating fom/base.py:523-599 with the weak forms of
fom/heat.py and fom/nonlinear.py) against the operator entries the REFERENCE ITSELF holds (tests/test_mpf1.py:126-317,
copied to tests/golden/mfp1_nx3.json), plus the structural properties the piston forms -- which no runnable reference
test pins -- must satisfy."""
import json
import os

import numpy as np

import p1_forms

HERE = os.path.dirname(os.path.abspath(__file__))


def test_mfp1_operators_match_the_reference_held_entries():
    g = json.load(open(os.path.join(HERE, "golden", "mfp1_nx3.json")))
    prob = p1_forms.Mfp1(nx=g["nx"], L0=g["L"], moving=False)
    for k, mu in enumerate(g["mus"]):
        np.testing.assert_allclose(prob.full_matrix("stiffness", mu, g["t"]).flatten(), g["Ah"][k], atol=5e-9)
        np.testing.assert_allclose(prob.full_matrix("mass", mu, g["t"]).flatten(), g["Mh"][k], atol=5e-9)
        np.testing.assert_allclose(prob.full_vector("forcing", mu, g["t"]), g["fh"][k], atol=5e-8)
        np.testing.assert_allclose(prob.full_vector("lifting", mu, g["t"]), g["fgh_time"][k], atol=5e-8)


def test_sampled_entries_are_the_entries_of_the_assembled_operator():
    """assemble(entries=...) must pick the entries of the globally assembled operator (what (M)DEIM relies on), in any
    order, for matrices and vectors, on the moved mesh."""
    rng = np.random.default_rng(0)
    heat = p1_forms.Mfp1(nx=9, L0=2.0, moving=True)
    mu = {"delta": 0.7, "beta": 3.0, "alpha_0": 0.4, "omega": 1.3}
    pist = p1_forms.Piston(nx=11, which="rest")
    mup = {"a0": 9.0, "omega": 17.0, "delta": 0.12}
    u = rng.standard_normal(pist.mesh.Nh)
    for prob, m, ops, t in ((heat, mu, ("mass", "stiffness", "convection"), 0.37),
                            (pist, mup, ("mass", "stiffness", "convection", "nonlinear_lifting", "trilinear"), 0.21)):
        pat = prob.mesh.pattern()
        pick = [pat[i] for i in rng.permutation(len(pat))[:12]]
        for op in ops:
            un = u if op == "trilinear" else None
            A = prob.full_matrix(op, m, t, u_n=un)
            np.testing.assert_allclose(prob.assemble(op, m, t, pick, u_n=un), [A[i, j] for i, j in pick], rtol=0, atol=1e-15)
        f = prob.full_vector("rhs", m, t)
        dofs = [(int(i),) for i in rng.permutation(prob.mesh.Nh)[:5]]
        np.testing.assert_allclose(prob.assemble("rhs", m, t, dofs), [f[i[0]] for i in dofs], rtol=0, atol=1e-15)


def test_p1_identities_on_the_moved_mesh():
    """Partition of unity and exactness on linears: M 1 = int phi_i, stiffness and convection annihilate constants, the
    trilinear form with u_n = 1 is b0 times the plain convection of a unit velocity, mass scales with L(t)."""
    pist = p1_forms.Piston(nx=16, which="sudden")
    mu = {"a0": 8.5, "omega": 19.0, "delta": 0.11}
    t = 0.05
    s = pist.scale(mu, t)
    free = [i for i in range(pist.mesh.Nh) if i not in pist.mesh.dirichlet_dofs]
    one = np.ones(pist.mesh.Nh)
    M = pist.full_matrix("mass", mu, t)
    hat = np.full(pist.mesh.Nh, s * pist.mesh.h0)
    hat[[0, -1]] *= 0.5
    np.testing.assert_allclose((M @ one)[free], hat[free], rtol=1e-14)
    np.testing.assert_allclose(M[np.ix_(free, free)], s * p1_forms.Piston(nx=16).full_matrix("mass", mu, 0.0)[np.ix_(free, free)], rtol=1e-14)
    for op in ("stiffness", "convection"):
        np.testing.assert_allclose((pist.full_matrix(op, mu, t) @ one)[free], 0.0, atol=1e-12)
    T = pist.full_matrix("trilinear", mu, t, u_n=one)
    Cunit = -(pist.full_matrix("convection", {**mu, "omega": 0.0, "a0": 1.0}, 0.0))   # -(1) u_x v with w = 0, negated
    np.testing.assert_allclose(T[np.ix_(free, range(pist.mesh.Nh))], pist.b0(mu) * Cunit[np.ix_(free, range(pist.mesh.Nh))], atol=1e-13)


def test_evaluator_tables_reproduce_the_assembly_for_every_mu_and_t():
    """romtime_b200/evaluators.py (product): geometry-only tables C, T with entry = C phi(mu, t) + b0 T u* -- against the
    direct closed-form assembly (oracle) at random parameters, times, entries and states, for both problems."""
    from romtime_b200 import evaluators

    rng = np.random.default_rng(1)
    for which in ("rest", "sudden"):
        nx = 13
        ev = evaluators.PistonP1(nx, L0=1.0, which=which)
        prob = p1_forms.Piston(nx, 1.0, which=which)
        pat = prob.mesh.pattern()
        pick = [pat[i] for i in rng.permutation(len(pat))[:15]]
        dofs = [(int(i),) for i in rng.permutation(prob.mesh.Nh)[:6]]
        V = rng.standard_normal((prob.mesh.Nh, 5))
        for _ in range(4):
            mu = {"a0": rng.uniform(8, 10), "omega": rng.uniform(15, 20), "delta": rng.uniform(0.1, 0.15)}
            t = rng.uniform(0, 1)
            phi = ev.features(mu, t)
            a = rng.standard_normal(5)
            for op in ("mass", "stiffness", "convection", "nonlinear_lifting"):
                ref = prob.assemble(op, mu, t, pick)
                np.testing.assert_allclose(ev.tables(op, pick)["C"] @ phi, ref, rtol=1e-12, atol=1e-14 * np.abs(ref).max())
            tb = ev.tables("trilinear", pick, V=V)
            ref = prob.assemble("trilinear", mu, t, pick, u_n=V @ a)
            np.testing.assert_allclose(tb["C"] @ phi + phi[11] * (tb["T"] @ a), ref, rtol=1e-12, atol=1e-13 * np.abs(ref).max())
            ref = prob.assemble("rhs", mu, t, dofs)
            np.testing.assert_allclose(ev.tables("rhs", dofs)["C"] @ phi, ref, rtol=1e-11, atol=1e-13 * np.abs(ref).max())
    nx, L0 = 10, 2.0
    ev = evaluators.Mfp1P1(nx, L0)
    for moving in (False, True):
        prob = p1_forms.Mfp1(nx, L0, moving=moving)
        pat = prob.mesh.pattern()
        pick = [pat[i] for i in rng.permutation(len(pat))[:12]]
        dofs = [(int(i),) for i in range(prob.mesh.Nh)]
        for _ in range(4):
            mu = {"delta": rng.uniform(0.01, 2), "beta": rng.uniform(1, 10), "alpha_0": rng.uniform(0.01, 2), "omega": rng.uniform(0.1, 1.0) if moving else 0.0}
            t = rng.uniform(0, 1.2)
            phi = ev.features(mu, t)
            for op in ("mass", "stiffness") + (("convection",) if moving else ()):
                ref = prob.assemble(op, mu, t, pick)
                np.testing.assert_allclose(ev.tables(op, pick)["C"] @ phi, ref, rtol=1e-12, atol=1e-14)
            ref = prob.assemble("rhs", mu, t, dofs)
            np.testing.assert_allclose(ev.tables("rhs", dofs)["C"] @ phi, ref, rtol=1e-10, atol=1e-12 * np.abs(ref).max())
