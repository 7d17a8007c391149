"""Builds libptb.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo).
One object per translation unit under csrc/build/ (git-ignored), recompiled only when the unit or a
header it includes changed, then one link."""
import os
import re
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")
OUT = os.path.join(_HERE, "libptb.so")
OBJDIR = os.path.join(CSRC, "build")

SOURCES = ["ptb_api.cu", "ptb_lbvh.cu", "ptb_grid_gpu.cu", "ptb_beam_gpu.cu", "ptb_group.cu", "ptb_bvh_build.cpp", "ptb_grid_build.cpp", "ptb_scene_loader.cpp", "ptb_png.cpp"]

# -fmad=false / -prec-div / -prec-sqrt / -ftz=false: the arithmetic contract of ptb_device.cuh
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-O2,-ffp-contract=off,-pthread",
]


def _nvcc():
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.isfile(c):
            return c
    raise RuntimeError("nvcc not found")


_INC = re.compile(r'^\s*#\s*include\s+"([^"]+)"', re.M)


def _deps(path, seen=None):
    """the file plus every quoted include reachable from it (csrc/ and include/)"""
    seen = set() if seen is None else seen
    if path in seen or not os.path.isfile(path):
        return seen
    seen.add(path)
    with open(path) as f:
        for name in _INC.findall(f.read()):
            for d in (CSRC, INCLUDE):
                _deps(os.path.join(d, name), seen)
    return seen


def _newer(deps, target):
    if not os.path.isfile(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout[-6000:])
    return r.stdout


def _have_zlib():
    """zlib header + link library present?  (PNG textures need them; without, ptb_png.cpp is built with PTB_NO_ZLIB.)"""
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "z.cpp")
        with open(src, "w") as f:
            f.write("#include <zlib.h>\nint main() { return zlibVersion() == nullptr; }\n")
        r = subprocess.run(["g++", src, "-o", os.path.join(d, "z"), "-lz"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        return r.returncode == 0


def build_native(force=False, verbose=False, defines=(), out=None):
    """Compile csrc/ into libptb.so if sources are newer than the library.  `defines`/`out` build a
    tuning variant beside it (loaded with PTB_LIB=<path>; scripts/ only)."""
    variant = bool(defines) or bool(out)
    out = out or OUT
    objdir = OBJDIR + ("_variant" if variant else "")
    os.makedirs(objdir, exist_ok=True)
    zlib = _have_zlib()
    flags = NVCC_FLAGS + [f"-D{d}" for d in defines] + ([] if zlib else ["-DPTB_NO_ZLIB"]) + (["-Xptxas", "-v"] if verbose else [])
    jobs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(objdir, os.path.splitext(s)[0] + ".o")
        if force or variant or _newer(_deps(src) | {os.path.abspath(__file__)}, obj):
            jobs.append([_nvcc()] + flags + ["-I", INCLUDE, "-c", "-o", obj, src])
    objs = [os.path.join(objdir, os.path.splitext(s)[0] + ".o") for s in SOURCES]
    if not jobs and not _newer(objs, out):
        return out
    with ThreadPoolExecutor(max_workers=4) as ex:
        logs = list(ex.map(_run, jobs))
    if verbose:
        print("\n".join(logs))
    _run([_nvcc(), "-shared", "-Xcompiler", "-pthread", "-o", out] + objs + (["-lz"] if zlib else []))  # zlib: PNG textures (ptb_png.cpp)
    return out
