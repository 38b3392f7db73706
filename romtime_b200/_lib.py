"""ctypes binding of libromtime_b200.so (include/romtime_b200.h).

There is NO CPU fallback: if the shared object is missing or no sm_100 GPU is present, every
compute entry point raises.  PyTorch is used only to own device memory and streams.
"""
import ctypes
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libromtime_b200.so")

RT_OK, RT_EINVAL, RT_ECUDA, RT_ESINGULAR, RT_ENOMEM, RT_ENOTCONV, RT_ENODEVICE = 0, -1, -2, -3, -4, -5, -6
RT_NF = 12
RT_MAX_OPS = 8

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_vp = ctypes.c_void_p
c_dbl = ctypes.c_double


class RomOperator(ctypes.Structure):
    _fields_ = [
        ("m", ctypes.c_int32),
        ("is_vector", ctypes.c_int32),
        ("basis_rom", c_vp),
        ("ld", c_i64),
        ("PT_Uinv", c_vp),
        ("C", c_vp),
        ("T", c_vp),
    ]


class RomModel(ctypes.Structure):
    _fields_ = [
        ("N", ctypes.c_int32),
        ("n_ops", ctypes.c_int32),
        ("evaluator", ctypes.c_int32),
        ("generation", ctypes.c_int32),
        ("n_mu", ctypes.c_int32),
        ("reserved", ctypes.c_int32),
        ("ops", RomOperator * RT_MAX_OPS),
    ]


# name -> (restype, argtypes); mirrors include/romtime_b200.h one to one
SIGNATURES = {
    "rt_version": (c_int, []),
    "rt_create": (c_int, [c_int, ctypes.POINTER(c_vp)]),
    "rt_destroy": (c_int, [c_vp]),
    "rt_last_error": (ctypes.c_char_p, [c_vp]),
    "rt_enable_peer_access": (c_int, [c_vp, c_int]),
    "rt_fp64_peak_tflops": (c_int, [c_vp, ctypes.POINTER(c_dbl)]),
    "rt_pod_gram_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_int, c_vp, c_vp, c_vp]),
    "rt_gram_cross_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_i64, c_vp]),
    "rt_gram_finalize_f64": (c_int, [c_vp, c_vp, c_i64, c_int, c_vp, c_vp]),
    "rt_sym_eig_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "rt_sym_eig_jacobi_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rt_sym_eig_jacobi_static_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "rt_deflate_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp]),
    "rt_pod_lift_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp]),
    "rt_small_gram_eig_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "rt_pod_gram_batch_f64": (c_int, [c_vp, c_vp, c_i64, c_int, c_vp]),
    "rt_pod_lift_batch_f64": (c_int, [c_vp, c_vp, c_i64, c_vp]),
    "rt_small_refine_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "rt_deim_greedy_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "rt_project_vec_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_i64, c_vp, c_vp]),
    "rt_project_mat_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp]),
    "rt_invert_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp]),
    "rt_theta_f64": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "rt_assemble_batch_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "rt_online_sweep_f64": (c_int, [c_vp, ctypes.POINTER(RomModel), c_vp, c_i64, c_i64, c_dbl, c_int, c_vp, c_vp, c_vp]),
    "rt_online_sweep_last_f64": (c_int, [c_vp, ctypes.POINTER(RomModel), c_vp, c_i64, c_i64, c_dbl, c_int, c_vp, c_vp, c_vp]),
    "rt_online_sweep_segment_f64": (c_int, [c_vp, ctypes.POINTER(RomModel), c_vp, c_i64, c_i64, c_i64, c_i64, c_dbl, c_int, c_vp, c_vp,
                                            c_vp, c_vp, c_vp]),
    "rt_memcpy2d_async": (c_int, [c_vp, c_vp, c_i64, c_vp, c_i64, c_i64, c_i64, c_int, c_vp]),
    "rt_online_step_f64": (c_int, [c_vp, ctypes.POINTER(RomModel), c_vp, c_i64, c_dbl, c_int, c_i64, c_vp, c_vp, c_vp, c_vp,
                                   c_vp, c_vp, c_vp]),
    "rt_rom_difference_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_i64, c_vp, c_i64, c_vp, c_i64, c_vp, c_vp]),
    "rt_errors_summary_f64": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_vp]),
    "rt_sweep_chunk": (c_i64, [c_i64, c_int, c_int]),
    "rt_profile_enable": (c_int, [c_vp, c_int]),
    "rt_profile_read": (c_int, [c_vp, ctypes.POINTER(c_dbl), ctypes.POINTER(c_i64)]),
    "rt_launch_count": (c_i64, [c_vp]),
}


class RomtimeB200Error(RuntimeError):
    pass


_lib = None
_lock = threading.Lock()
_handles = {}


def load_library():
    """dlopen the in-tree shared object and declare every prototype. Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RomtimeB200Error(
                    f"{LIB_PATH} is missing: run `python -m romtime_b200.build` (there is no CPU fallback)")
            lib = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)  # AttributeError if the header and the library disagree
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def handle(device=None):
    """One rt_handle per CUDA device, created on first use."""
    import torch

    if not torch.cuda.is_available():
        raise RomtimeB200Error("romtime_b200 needs a B200 (sm_100a) GPU: there is no CPU fallback")
    if device is None:
        device = torch.cuda.current_device()
    device = int(device)
    h = _handles.get(device)
    if h is None:
        lib = load_library()
        out = c_vp()
        rc = lib.rt_create(device, ctypes.byref(out))
        if rc != RT_OK:
            raise RomtimeB200Error(f"rt_create(device={device}) failed with status {rc}")
        h = _handles[device] = out
    return h


def check(rc, h=None, what=""):
    if rc == RT_OK:
        return
    msg = ""
    if h is not None:
        msg = load_library().rt_last_error(h).decode(errors="replace")
    if rc == RT_ESINGULAR:
        raise np.linalg.LinAlgError("Singular matrix" + (f" ({msg})" if msg else ""))
    raise RomtimeB200Error(f"{what} failed with status {rc}: {msg}")


def stream_ptr():
    import torch

    return c_vp(torch.cuda.current_stream().cuda_stream)


def launch_count(device=None):
    return int(load_library().rt_launch_count(handle(device)))
