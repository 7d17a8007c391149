"""Builds libptb.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")
OUT = os.path.join(_HERE, "libptb.so")

SOURCES = ["ptb_api.cu", "ptb_bvh_build.cpp", "ptb_grid_build.cpp"]
HEADERS = sorted(f for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))

# -fmad=false / -prec-div / -prec-sqrt / -ftz=false: the arithmetic contract of ptb_device.cuh
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-pthread",
    "-shared",
]


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.isfile(c):
            return c
    raise RuntimeError("nvcc not found")


def _stale():
    if not os.path.isfile(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(INCLUDE, "ptb.h"), os.path.join(INCLUDE, "ptbs_format.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_native(force=False, verbose=False, defines=(), out=None):
    """Compile csrc/ into libptb.so if sources are newer than the library.  `defines`/`out` build a
    tuning variant beside it (loaded with PTB_LIB=<path>; scripts/ only)."""
    out = out or OUT
    if not force and not defines and not _stale():
        return OUT
    cmd = [_nvcc()] + NVCC_FLAGS + [f"-D{d}" for d in defines] + (["-Xptxas", "-v"] if verbose else []) + \
        ["-I", INCLUDE, "-o", out] + [os.path.join(CSRC, s) for s in SOURCES]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout[-6000:])
    if verbose:
        print(r.stdout)
    return out
