"""Build libromtime_b200.so in-tree with nvcc for sm_100a (called by __graft_entry__.build()).

    python -m romtime_b200.build [--force]

The library is plain CUDA C++ behind a C ABI (include/romtime_b200.h); it is linked against the
static CUDA runtime so that it loads next to torch's own runtime without symbol clashes.
"""
import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libromtime_b200.so")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3",
    "--expt-relaxed-constexpr",
    "-cudart", "static",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libromtime_b200.so cannot be built")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + \
        glob.glob(os.path.join(HERE, "..", "include", "*.h")) + [os.path.abspath(__file__)]
    return any(os.path.getmtime(p) > t for p in deps)


# The shipped assembly GEMM (k_dgemm<MN, MN, full tile>) carries ONE 4-byte spill slot and is bitwise reproducible in every
# parity test.  Builds of the same kernel with a few hundred bytes of spills have produced corrupted 8-row blocks twice (the
# deferred-store experiment of round 1; the 5-stage / 200-register variant of round 2 when it shared an SM with the solve
# kernel), so such a build is not linked.
MAX_GEMM_SPILL_BYTES = 16


def _spilling_gemm_kernels(ptxas_log):
    """dg::k_dgemm instantiations for which ptxas -v reports more than MAX_GEMM_SPILL_BYTES of spill stores."""
    bad, current = [], None
    for line in ptxas_log.splitlines():
        if "Compiling entry function" in line:
            current = line.split("'")[1] if "'" in line else line
        elif "bytes spill stores" in line and current and "k_dgemm" in current:
            nums = [int(tok) for tok in line.replace(",", " ").split() if tok.isdigit()]
            if len(nums) >= 3 and nums[1] > MAX_GEMM_SPILL_BYTES:
                bad.append(f"{current}: {nums[1]} bytes of spill stores")
    return bad


def build_library(force=False, verbose=False):
    """Compile every csrc/*.cu into one shared object. Returns the library path."""
    if not force and not _stale():
        return LIB
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get("RT_NVCC_EXTRA", "").split() + ["-Xptxas", "-v", "-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"nvcc failed on {src}:\n{out}\n")
        else:
            if verbose:
                sys.stderr.write(out)
            else:
                sys.stderr.write("".join(l + "\n" for l in out.splitlines() if "warning" in l and "ptxas info" not in l))
            spilled = _spilling_gemm_kernels(out)
            if spilled and os.environ.get("RT_ALLOW_GEMM_SPILLS") != "1":
                failed = True
                sys.stderr.write(f"{src}: register spills in the GEMM core ({', '.join(spilled)}): heavily spilling builds of "
                                 "dg::k_dgemm have produced corrupted 8-row blocks (DESIGN.md 4.1); refusing to link it "
                                 "(RT_ALLOW_GEMM_SPILLS=1 overrides, for experiments)\n")
    if failed:
        raise RuntimeError("nvcc compilation failed")
    link = [nvcc, "-shared", "-cudart", "static", "-o", LIB] + objs + ["-ldl", "-lpthread", "-lrt"]
    subprocess.run(link, check=True)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
