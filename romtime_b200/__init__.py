"""romtime_b200 -- B200-native (sm_100a) offline-compression + online-solve hot path of romtime.

Keeps the reference's Python ROM API (``orth``, DEIM/MDEIM builders, reduced-operator assembly,
``RomConstructor*`` time integrator); the arithmetic runs in hand-written CUDA kernels behind
the C ABI declared in include/romtime_b200.h.
"""
__version__ = "0.1.0"
