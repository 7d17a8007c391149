"""ctypes binding of the CPU oracle (oracle/libptoracle.so) and a runner for the compiled
reference (oracle/_ref/ptref).  TEST INFRASTRUCTURE: imported by tests/, smoke() and
bench.py's cpu_baseline / --impl reference legs only."""
import ctypes
import json
import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "libptoracle.so")
PTREF = os.path.join(ORACLE_DIR, "_ref", "ptref")

MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS = 0, 1, 2, 3


def build_port():
    """Compile the restatement (plain g++, seconds)."""
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "port"])
    return PORT_SO


def have_ref():
    return os.path.isfile(PTREF) and os.access(PTREF, os.X_OK)


_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")


class Oracle:
    """One scene loaded into the restatement."""

    def __init__(self, ptbs_path):
        if not os.path.isfile(PORT_SO):
            build_port()
        L = ctypes.CDLL(PORT_SO)
        L.pto_create.restype = ctypes.c_void_p
        L.pto_destroy.argtypes = [ctypes.c_void_p]
        L.pto_last_error.restype = ctypes.c_char_p
        L.pto_last_error.argtypes = [ctypes.c_void_p]
        L.pto_load_ptbs.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.pto_build.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 8
        L.pto_configure.argtypes = [ctypes.c_void_p] + [ctypes.c_int] * 6
        L.pto_primary_hits.argtypes = [ctypes.c_void_p, _f32p, _i32p, _i32p, ctypes.c_void_p]
        L.pto_trace_rays.argtypes = [ctypes.c_void_p, _f32p, ctypes.c_uint64, _f32p, _i32p, ctypes.c_int]
        L.pto_render.argtypes = [ctypes.c_void_p, ctypes.c_int, _f32p, _i32p, _u32p, _u64p, ctypes.c_int, ctypes.c_int]
        for f in ("pto_bvh_nodes", "pto_num_grids", "pto_num_beams", "pto_grid_entries"):
            getattr(L, f).restype = ctypes.c_uint64
            getattr(L, f).argtypes = [ctypes.c_void_p]
        self.L = L
        self.h = L.pto_create()
        if L.pto_load_ptbs(self.h, ptbs_path.encode()) != 0:
            raise RuntimeError(L.pto_last_error(self.h).decode())
        self.W = self.H = 0

    def close(self):
        if self.h:
            self.L.pto_destroy(self.h)
            self.h = None

    def build(self, grid=False, beams=False, gs=10, os_=5, m=40, bgs=10, bm=40, max_beam_depth=0):
        self.L.pto_build(self.h, int(grid), int(beams), gs, os_, m, bgs, bm, max_beam_depth)

    def configure(self, W, H, mode, nee=0, max_depth=0, primary_only=0):
        self.W, self.H = W, H
        self.L.pto_configure(self.h, W, H, mode, nee, max_depth, primary_only)

    def primary_hits(self, want_rays=False):
        n = self.W * self.H
        t = np.empty(n, np.float32)
        tri = np.empty(n, np.int32)
        mat = np.empty(n, np.int32)
        rays = np.empty((n, 6), np.float32) if want_rays else None
        self.L.pto_primary_hits(self.h, t, tri, mat, rays.ctypes.data if want_rays else None)
        return (t, tri, mat, rays) if want_rays else (t, tri, mat)

    def primary_counts(self):
        """(AS tests, triangle tests) of every pixel's first-sample primary ray — the debug view's input."""
        n = self.W * self.H
        a = np.empty(n, np.uint32)
        t = np.empty(n, np.uint32)
        self.L.pto_primary_counts.argtypes = [ctypes.c_void_p, _u32p, _u32p]
        self.L.pto_primary_counts(self.h, a, t)
        return a, t

    def export_bvh(self):
        """The oracle's (= the reference's) two-level BVH as ptb_upload_ref_bvh takes it:
        (nodes[48-byte records], indices, object_roots, scene_root)."""
        self.L.pto_export_bvh.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * 5
        sizes = np.zeros(3, np.uint64)
        root = ctypes.c_int32(0)
        self.L.pto_export_bvh(self.h, None, None, None, sizes.ctypes.data, ctypes.byref(root))
        nodes = np.zeros(int(sizes[0]) * 48, np.uint8)
        idx = np.zeros(int(sizes[1]), np.uint32)
        roots = np.zeros(int(sizes[2]), np.int32)
        self.L.pto_export_bvh(self.h, nodes.ctypes.data, idx.ctypes.data, roots.ctypes.data, sizes.ctypes.data, ctypes.byref(root))
        return nodes, idx, roots, int(root.value)

    def export_grids(self):
        """The oracle's (= the reference's) Subgrid hierarchy as ptb_upload_subgrid takes it:
        (grids[80-byte records], lut_start, lut_end, indices)."""
        self.L.pto_export_grids.argtypes = [ctypes.c_void_p] * 6
        sizes = np.zeros(3, np.uint64)
        self.L.pto_export_grids(self.h, None, None, None, None, sizes.ctypes.data)
        grids = np.zeros(int(sizes[0]) * 80, np.uint8)
        ls = np.zeros(int(sizes[1]), np.uint32)
        le = np.zeros(int(sizes[1]), np.uint32)
        idx = np.zeros(int(sizes[2]), np.uint32)
        self.L.pto_export_grids(self.h, grids.ctypes.data, ls.ctypes.data, le.ctypes.data, idx.ctypes.data, sizes.ctypes.data)
        return grids, ls, le, idx

    def trace_rays(self, od, threads=0):
        od = np.ascontiguousarray(od, np.float32).reshape(-1, 6)
        n = od.shape[0]
        t = np.empty(n, np.float32)
        tri = np.empty(n, np.int32)
        self.L.pto_trace_rays(self.h, od.reshape(-1), n, t, tri, threads)
        return t, tri

    def render(self, spp, state=None, threads=0):
        """Returns (rgba[H,W,4], samples, rng, stats dict); pass `state` to continue."""
        n = self.W * self.H
        if state is None:
            rgba = np.zeros(n * 4, np.float32)
            samples = np.zeros(n, np.int32)
            rng = np.zeros(n, np.uint32)
            resume = 0
        else:
            rgba, samples, rng = state
            rgba = rgba.reshape(-1)
            resume = 1
        stats = np.zeros(6, np.uint64)
        self.L.pto_render(self.h, spp, rgba, samples, rng, stats, resume, threads)
        keys = ("tri_tests", "as_tests", "find_calls", "rays_traversed", "rays_unique", "paths")
        return rgba.reshape(self.H, self.W, 4), samples, rng, dict(zip(keys, (int(v) for v in stats)))


def run_ptref(args, cwd=None, timeout=3600):
    """Run the compiled reference; returns the list of PTREF_JSON records it printed."""
    out = subprocess.run([PTREF] + [str(a) for a in args], cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                         text=True, timeout=timeout)
    if out.returncode != 0:
        raise RuntimeError("ptref failed: " + out.stderr[-2000:])
    recs = [json.loads(l[len("PTREF_JSON "):]) for l in out.stdout.splitlines() if l.startswith("PTREF_JSON ")]
    return recs


def ptref_primary_hits(ptbs, W, H, mode, extra=()):
    with tempfile.TemporaryDirectory() as d:
        o = os.path.join(d, "hits.bin")
        recs = run_ptref(["--ptbs", ptbs, "--mode", mode, "--width", W, "--height", H, *extra, "primary-hits", o])
        n = W * H
        raw = np.fromfile(o, np.uint8)
        t = raw[: 4 * n].view(np.float32).copy()
        tri = raw[4 * n: 8 * n].view(np.int32).copy()
        mat = raw[8 * n: 12 * n].view(np.int32).copy()
        rays = raw[12 * n:].view(np.float32).reshape(n, 6).copy()
    return t, tri, mat, rays, recs


def ptref_render(ptbs, W, H, mode, spp, max_depth=0, nee=0, extra=(), want_image=True):
    with tempfile.TemporaryDirectory() as d:
        o = os.path.join(d, "img.bin") if want_image else "-"
        recs = run_ptref(["--ptbs", ptbs, "--mode", mode, "--width", W, "--height", H, "--spp", spp, "--max-depth", max_depth,
                          "--nee", nee, *extra, "render", o])
        img = np.fromfile(o, np.float32).reshape(H, W, 4) if want_image else None
    return img, [r for r in recs if r["cmd"] == "render"][0], recs


def ptref_trace_rays(ptbs, od, mode, extra=()):
    with tempfile.TemporaryDirectory() as d:
        i = os.path.join(d, "rays.bin")
        o = os.path.join(d, "out.bin")
        od = np.ascontiguousarray(od, np.float32).reshape(-1, 6)
        od.tofile(i)
        recs = run_ptref(["--ptbs", ptbs, "--mode", mode, *extra, "trace-rays", i, o])
        n = od.shape[0]
        raw = np.fromfile(o, np.uint8)
        t = raw[: 4 * n].view(np.float32).copy()
        tri = raw[4 * n:].view(np.int32).copy()
    return t, tri, recs


PTFACADE = os.path.join(ORACLE_DIR, "_ref", "ptfacade")


def have_facade():
    return os.path.isfile(PTFACADE) and os.access(PTFACADE, os.X_OK)


PTFACADE_GPU = os.path.join(ORACLE_DIR, "_ref", "ptfacade_gpu")


def have_facade_gpu():
    return os.path.isfile(PTFACADE_GPU) and os.access(PTFACADE_GPU, os.X_OK)


def ptfacade_render(ptbs, W, H, mode, spp, max_depth=0, nee=0, extra=(), structures=False, want_build=False):
    """The reference's own objects with src/tracer.cpp replaced by host/tracer_facade.cpp: trace() on the GPU.
    structures=True: the flavour that also swaps bvh.cpp / lightFieldGrid.cpp / lightFieldGridSubBeams.cpp for
    host/structures_facade.cpp (oracle/_ref/ptfacade_gpu): initScene's builds run on the GPU too."""
    with tempfile.TemporaryDirectory() as d:
        o = os.path.join(d, "img.bin")
        out = subprocess.run([PTFACADE_GPU if structures else PTFACADE] + [str(a) for a in ["--ptbs", ptbs, "--mode", mode, "--width", W, "--height", H, "--spp", spp,
                                                           "--max-depth", max_depth, "--nee", nee, *extra, "render", o]],
                             stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=1800)
        if out.returncode != 0:
            raise RuntimeError("ptfacade failed: " + out.stderr[-2000:])
        recs = [json.loads(l[len("PTREF_JSON "):]) for l in out.stdout.splitlines() if l.startswith("PTREF_JSON ")]
        img = np.fromfile(o, np.float32).reshape(H, W, 4)
    if want_build:
        return img, [r for r in recs if r["cmd"] == "render"][0], [r for r in recs if r["cmd"] == "build"][0]
    return img, [r for r in recs if r["cmd"] == "render"][0]
