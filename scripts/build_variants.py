#!/usr/bin/env python
"""Tuning variants of libptb.so for A/B runs on the GPU box: only csrc/ptb_api.cu (the one unit that includes the device
headers) is recompiled with the given -D defines; the other objects are the default build's.
usage: scripts/build_variants.py name=DEF1,DEF2=3 [name2=...]   ->  bachelor-thesis-pathtracer_b200/variants/libptb_<name>.so
       (load with PTB_LIB=<path>)"""
import os
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "bachelor-thesis-pathtracer_b200"))
import build as B  # noqa: E402

WATCH = ("ILi1ELb0ELb0ELb1ELb0", "ILi2ELb0ELb0ELb0ELb0", "ILi1ELb0ELb0ELb0ELb0", "ILi3ELb0ELb0ELb0ELb0")


def one(spec):
    name, _, defs = spec.partition("=")
    defs = [d for d in defs.split(",") if d]
    vdir = os.path.join(B._HERE, "variants")
    odir = os.path.join(B.CSRC, "build_variant", name)
    os.makedirs(vdir, exist_ok=True)
    os.makedirs(odir, exist_ok=True)
    obj = os.path.join(odir, "ptb_api.o")
    zlib = B._have_zlib()
    log = B._run([B._nvcc()] + B.NVCC_FLAGS + [f"-D{d}" for d in defs] + ["-Xptxas", "-v", "-I", B.INCLUDE, "-c", "-o", obj,
                                                                          os.path.join(B.CSRC, "ptb_api.cu")])
    objs = [obj if s == "ptb_api.cu" else os.path.join(B.OBJDIR, os.path.splitext(s)[0] + ".o") for s in B.SOURCES]
    out = os.path.join(vdir, f"libptb_{name}.so")
    B._run([B._nvcc(), "-shared", "-Xcompiler", "-pthread", "-o", out] + objs + (["-lz"] if zlib else []))
    lines = log.splitlines()
    info = []
    for i, l in enumerate(lines):
        if "Compiling entry function" in l and "k_mega" in l and any(w in l for w in WATCH):
            info.append(l.split("'")[1][9:32] + " | " + lines[i + 2].strip() + " | " + lines[i + 3].strip().replace("ptxas info    : ", ""))
    return name, out, info


if __name__ == "__main__":
    B.build_native()  # default objects up to date
    with ThreadPoolExecutor(max_workers=int(os.environ.get("JOBS", "3"))) as ex:
        for name, out, info in ex.map(one, sys.argv[1:]):
            print(name, out)
            for l in info:
                print("   ", l)
