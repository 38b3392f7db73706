"""TEST INFRASTRUCTURE ONLY -- imports the *unmodified* reference (KikeM/romtime) from
/root/reference so that golden vectors can be generated in the build container.

Nothing under romtime_b200/ may import this module.  /root/reference does not exist on
the GPU box, so this file is only ever executed by oracle/gen_golden.py (which writes
tests/golden/*.npz) and by the optional ``-m "not gpu"`` cross-check
tests/test_oracle_vs_reference.py (skipped when /root/reference is absent).

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

REFERENCE_SRC = "/root/reference/src"


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
