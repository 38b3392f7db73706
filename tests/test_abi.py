"""CPU: the C-ABI library builds, loads and exports every symbol include/romtime_b200.h declares;
the product path fails loudly without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "romtime_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from romtime_b200 import _lib, build

    build.build_library()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = header_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), name
    assert set(_lib.SIGNATURES) == set(names)
    assert _lib.load_library().rt_version() == 100


def test_structs_match_header_layout():
    from romtime_b200 import _lib

    assert ctypes.sizeof(_lib.RomOperator) == 48
    assert ctypes.sizeof(_lib.RomModel) == 24 + 8 * 48  # N, n_ops, evaluator, generation, n_mu, reserved + RT_MAX_OPS operators


def test_no_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from romtime_b200 import _lib, kernels
    from romtime_b200.rom.pod import orth

    with pytest.raises(_lib.RomtimeB200Error):
        kernels.deim_greedy(np.eye(4))
    with pytest.raises(_lib.RomtimeB200Error):
        orth(np.eye(4))
    out = ctypes.c_void_p()
    assert _lib.load_library().rt_create(0, ctypes.byref(out)) == _lib.RT_ENODEVICE


def test_product_code_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "romtime_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert "romtime_oracle" not in src and "ref_harness" not in src and "import oracle" not in src, f
