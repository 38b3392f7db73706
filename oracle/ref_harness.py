"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference (KikeM/romtime) from
/root/reference so that golden vectors can be generated in the build container.

Nothing under romtime_b200/ may import this module.  /root/reference does not exist on
the GPU box; there the vendored copy under oracle/_ref (``make -C oracle``: a plain copy of the pure-Python
package, git-ignored) is imported instead.  Used by oracle/gen_golden.py (writes tests/golden/*.npz), by the
``-m "not gpu"`` cross-check tests/test_oracle_vs_reference.py (skipped when no reference tree is present) and by
oracle/cpu_bench.py (the timed ``--impl reference`` arm of bench.py).

The reference imports FEniCS / matplotlib / ujson at module import time
(romtime/utils.py:4, romtime/rom/__init__.py:3, romtime/rom/rom.py:5-7) and uses three
NumPy/SciPy call forms that newer releases dropped (deim/deim.py:535 ``np.reshape(a=,
newshape=)``; rom/rom.py:36 ``gmres(tol=)``; utils.py:217 ``np.trapz``).  We register
empty stub modules and tiny shims -- the reference source itself is not touched.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# the mounted reference in the build container; on the GPU box the vendored copy that `make -C oracle` put under
# oracle/_ref (git-ignored, travels with the snapshot) -- the package is pure Python, nothing is compiled
_CANDIDATES = ("/root/reference/src", os.path.join(HERE, "_ref"))
REFERENCE_SRC = next((c for c in _CANDIDATES if os.path.isdir(os.path.join(c, "romtime"))), _CANDIDATES[0])


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_SRC, "romtime"))


def _stub(name, **attrs):
    mod = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(mod, k, v)
    sys.modules[name] = mod
    return mod


_LOADED = False


def load_reference():
    """Return the imported reference package (``romtime``) with stubs/shims installed."""
    global _LOADED
    if not reference_available():
        raise RuntimeError("reference tree not present at " + REFERENCE_SRC)
    if _LOADED:
        import romtime

        return romtime

    class _Dummy:  # stands in for dolfin Matrix / Vector in isinstance checks
        pass

    class _Matrix(_Dummy):
        pass

    class _Vector(_Dummy):
        pass

    for name in ("fenics", "dolfin", "dolfin.cpp", "matplotlib", "seaborn", "adjustText"):
        if name not in sys.modules:
            _stub(name)
    _stub("dolfin.cpp.la", Matrix=_Matrix, Vector=_Vector)
    sys.modules["dolfin.cpp"].la = sys.modules["dolfin.cpp.la"]
    sys.modules["dolfin"].cpp = sys.modules["dolfin.cpp"]
    sys.modules["fenics"].DOLFIN_EPS = 3e-16
    # utils.bilinear_to_csr (utils.py:75-93) goes fenics.as_backend_type(matrix).mat().getValuesCSR():
    # synthetic_problems.PetscLikeMatrix answers that protocol for scipy matrices, so MDEIM.run works unmodified.
    sys.modules["fenics"].as_backend_type = lambda matrix: matrix
    _stub("matplotlib.pyplot")
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["seaborn"].set_theme = lambda *a, **k: None
    sys.modules["adjustText"].adjust_text = lambda *a, **k: None
    import json

    _stub("ujson", dump=json.dump, load=json.load, dumps=json.dumps, loads=json.loads)

    # --- NumPy / SciPy API shims (call-form only, semantics unchanged) ---------------
    _np_reshape = np.reshape

    def _reshape(*args, **kwargs):
        if "newshape" in kwargs:
            kwargs["shape"] = kwargs.pop("newshape")
        if "a" in kwargs:
            args = (kwargs.pop("a"),) + args
        return _np_reshape(*args, **kwargs)

    np.reshape = _reshape
    if not hasattr(np, "trapz"):
        np.trapz = np.trapezoid

    import scipy.sparse.linalg as spla

    _gmres = spla.gmres

    def _gmres_legacy(A, b, x0=None, tol=None, atol=None, maxiter=None, **kw):
        # SciPy 1.6 semantics: stop when ||r|| <= max(atol, tol*||b||)  (rom/rom.py:36)
        if maxiter is not None:
            maxiter = int(maxiter)
        return _gmres(A, b, x0=x0, rtol=tol if tol is not None else 1e-5,
                      atol=atol if atol is not None else 0.0, maxiter=maxiter, **kw)

    spla.gmres = _gmres_legacy

    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import romtime  # noqa: F401
    import romtime.rom.pod  # noqa: F401
    import romtime.deim.deim  # noqa: F401
    import romtime.deim.mdeim  # noqa: F401
    import romtime.deim.nonlinear  # noqa: F401
    import romtime.rom.rom  # noqa: F401

    _LOADED = True
    return romtime


# ---------------------------------------------------------------------------------
# Duck-typed FOM for RomConstructorNonlinear.solve (attributes touched at
# rom/rom.py:441-515 and :748).  FEniCS-dependent inputs are replaced by arrays.
# ---------------------------------------------------------------------------------
class _ArrayFunction:
    """Mimics dolfin Function enough for utils.function_to_array (utils.py:44-55)."""

    def __init__(self, arr):
        self._arr = np.asarray(arr, dtype=float)

    def vector(self):
        return self

    def vec(self):
        return self

    @property
    def array(self):
        return self._arr


class DuckFom:
    def __init__(self, Nh, dt, nt, bdf_scheme, lifting_fn=None):
        self.exact_solution = None
        self.dt = dt
        self.BDF_SCHEME = bdf_scheme
        self.domain = {"nt": nt}
        self.probe_location = [0.0, 0.5]
        self.L = 1.0
        self.V = None
        self.x = np.linspace(0.0, 1.0, Nh).reshape(Nh, 1)
        self.is_setup = True
        self._Nh = Nh
        self._lifting_fn = lifting_fn

    def move_mesh(self, mu=None, t=None, back=False):
        pass

    def create_lifting_operator(self, mu, t, L):
        return (mu, t), None, None

    def interpolate_func(self, g, V, mu, t):
        if self._lifting_fn is None:
            return _ArrayFunction(np.zeros(self._Nh))
        return _ArrayFunction(self._lifting_fn(mu, t))


def build_reference_rom(model, V, bdf, dt, nt, direct):
    """The reference's own RomConstructorNonlinear (rom/rom.py:739-974), unmodified, wired to a synthetic
    hyper-reduced piston model (romtime_b200.synthetic.make_piston_model): its six reductors are the reference's own
    DEIM / MDEIM / N-MDEIM classes with ``basis_rom`` / ``PT_U`` set directly and NumPy ``assemble`` callbacks in place
    of FEniCS local assembly.  ``direct``: swap the algebraic-solver hook (rom/rom.py:414-425) for np.linalg.solve;
    otherwise the stock GMRES (rom/rom.py:36) runs."""
    load_reference()
    from romtime.deim.deim import DiscreteEmpiricalInterpolation
    from romtime.deim.mdeim import MatrixDiscreteEmpiricalInterpolation
    from romtime.deim.nonlinear import MatrixDiscreteEmpiricalInterpolationNonlinear
    from romtime.rom.rom import RomConstructorNonlinear
    import romtime_oracle as orc

    N = model["N"]
    fom = DuckFom(Nh=V.shape[0], dt=dt, nt=nt, bdf_scheme=bdf)
    rom = RomConstructorNonlinear(fom=fom, grid=None, name="golden")
    rom.setup(rnd=0)
    rom.basis = V
    if direct:
        rom.algebraic_solver = lambda A, b: (np.linalg.solve(A, b), 0)
    Tfom = model["trilinear"]["T"] @ V.T  # (m, Nh): Tfom @ (V u) == T u

    def feats(mu, t):
        return orc.piston_features(mu["a0"], mu["omega"], mu["delta"], t, model["which"])

    def make(op):
        C = model[op]["C"]
        if op == "trilinear":
            def assemble(mu, t, entries=None, u_n=None):
                phi = feats(mu, t)
                return C @ phi + phi[11] * (Tfom @ u_n)
            r = MatrixDiscreteEmpiricalInterpolationNonlinear(assemble=assemble, name=op)
        elif op == "rhs":
            def assemble(mu, t, entries=None):
                return C @ feats(mu, t)
            r = DiscreteEmpiricalInterpolation(assemble=assemble, name=op)
        else:
            def assemble(mu, t, entries=None):
                return C @ feats(mu, t)
            r = MatrixDiscreteEmpiricalInterpolation(assemble=assemble, name=op)
        mm = model[op]["PT_U"].shape[0]
        r.basis_fom = np.zeros((1, mm))  # only .shape[1] is read online (deim/deim.py:106-108)
        r.basis_rom = model[op]["basis_rom"]
        r.PT_U = model[op]["PT_U"]
        r.dofs = [(k,) for k in range(mm)]
        r.N_V = N
        return r

    rom.mdeim_Mh = make("mass")
    rom.mdeim_Ah = make("stiffness")
    rom.mdeim_Ch = make("convection")
    rom.mdeim_Nh = make("trilinear")
    rom.mdeim_Nh_hat = make("nonlinear_lifting")
    rom.deim_fgh = make("rhs")
    return rom
