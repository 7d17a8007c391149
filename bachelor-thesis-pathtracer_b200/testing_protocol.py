"""The reference's `--testing` protocol, headless: for mode BVH (unless --2plane_only) and Subgrids,
three passes each — the full render at -S samples with tone mapping, then the two heat maps (1 sample, debug view:
structure tests, then triangle tests) — NEE off, 1280x720 (main.cpp:92,99-141).  After the first pass the five lines
the reference prints (main.cpp:132-136: Rendering Time / Triangle Intersections / AS Intersections / Total Intersections /
Memory Consumption) go to stdout with the same keys, so showcase/test.py's regexes read them; after every pass the frame is
written as render/render_[heatmap_as_|heatmap_tris_]{BVH,Subgrids}.png — the names window.cpp:540-546 uses — through the
display transform of window.cpp:507-531 (ptb_read_rgba8).  The window, SDL and saveImage themselves are graphics-stack
code (unavailable on this box); the PNG here is written by `write_png` below.

"Rendering Time" brackets what main.cpp:116-125 brackets: initTracer (reset of image / samples / RNG) and the trace;
structure construction happens before, in initScene (main.cpp:97).  With `ref_tree` (the reference's own BVH as exported
by its accessors, ptb_upload_ref_bvh) the BVH counters are the CPU tracer's numbers exactly; otherwise they
count visits of this library's 64-byte nodes.  Likewise with `ref_grids` (the reference's own std::vector<Grid>,
ptb_upload_subgrid) the Subgrids block carries the CPU tracer's numbers exactly, and its "Memory Consumption" is
getMemory2Plane of those grids (lightFieldGrid.cpp:60-68).  Otherwise "Memory Consumption" is the device footprint of the
structure (ptb_structure_bytes), the counterpart of getMemoryBVH / getMemory2Plane (bvh.cpp:30-32).
"""
import os
import struct
import time
import zlib

import numpy as np

from .api import MODE_BVH, MODE_TWO_PLANE

ITEMS = ("ALL.png", "BVH.png", "Subgrids.png", "Subbeams.png")  # window.cpp:541


def write_png(path, rgba8):
    """8-bit RGBA, rows top to bottom; no filter, zlib level 6"""
    a = np.ascontiguousarray(rgba8, np.uint8)
    if a.ndim != 3 or a.shape[2] != 4:
        raise ValueError("write_png wants an (H, W, 4) uint8 array")
    h, w, _ = a.shape
    raw = np.empty((h, 1 + 4 * w), np.uint8)
    raw[:, 0] = 0
    raw[:, 1:] = a.reshape(h, 4 * w)

    def chunk(tag, data):
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 6, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)) + chunk(b"IEND", b""))


def read_png(path):
    """the inverse of write_png (filter 0 rows only) — for the tests"""
    with open(path, "rb") as f:
        b = f.read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    off, w, h, data = 8, 0, 0, b""
    while off < len(b):
        n, tag = struct.unpack(">I", b[off:off + 4])[0], b[off + 4:off + 8]
        body = b[off + 8:off + 8 + n]
        assert struct.unpack(">I", b[off + 8 + n:off + 12 + n])[0] == zlib.crc32(tag + body) & 0xFFFFFFFF
        if tag == b"IHDR":
            w, h, depth, colour = struct.unpack(">IIBB", body[:10])
            assert (depth, colour) == (8, 6)
        elif tag == b"IDAT":
            data += body
        off += 12 + n
    raw = np.frombuffer(zlib.decompress(data), np.uint8).reshape(h, 1 + 4 * w)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4).copy()


def file_name(mode, debug_view, show_tris):  # window.cpp:542-543
    return "render_" + (("heatmap_tris_" if show_tris else "heatmap_as_") if debug_view else "") + ITEMS[mode]


def run(pt, samples=100, gs=10, os_=5, m=40, two_plane_only=False, width=1280, height=720, debug_scale=10, max_depth=0,
        out_dir="./render", ref_tree=None, ref_grids=None, gpu_builders=True, emit=print):
    """`pt`: a PathTracer with the scene uploaded.  Returns the per-mode records it printed."""
    os.makedirs(out_dir, exist_ok=True)
    # initScene (main.cpp:97): both structures are built before the clock starts
    pt.build_bvh()
    if ref_tree is not None:
        pt.upload_ref_bvh(*ref_tree)
    if ref_grids is not None:
        pt.upload_subgrid(*ref_grids)
    else:
        pt.build_subgrid(gs, os_, m, gpu=gpu_builders)
    emit("starting test")
    records = []
    for mode in ((MODE_TWO_PLANE,) if two_plane_only else (MODE_BVH, MODE_TWO_PLANE)):
        pt.use_ref_bvh(ref_tree is not None and mode == MODE_BVH)
        for j in range(3):
            spp = samples if j == 0 else 1
            pt.set_params(width=width, height=height, mode=mode, nee=0, max_depth=max_depth, primary_only=0, count_tests=1,
                          debug_view=0 if j == 0 else j, debug_scale=debug_scale)
            t0 = time.perf_counter()
            pt.reset()  # initTracer
            pt.render(spp)  # returns when the frame is complete, like trace(true)
            ms = (time.perf_counter() - t0) * 1e3
            if j == 0:
                c = pt.counters()
                rec = {"mode": mode, "rendering_time_ms": ms, "device_ms": pt.last_ms(), "tri_tests": c["tri_tests"],
                       "as_tests": c["as_tests"], "memory": pt.structure_bytes(mode), "files": []}
                if mode == MODE_TWO_PLANE and ref_grids is not None:
                    rec["memory"] = int(pt.L.ptb_ref_subgrid_bytes(pt.h))
                emit("Rendering Time: %f" % ms)
                emit("Triangle Intersections: %d " % c["tri_tests"])
                emit("AS Intersections: %d " % c["as_tests"])
                emit("Total Intersections: %d " % (c["as_tests"] + c["tri_tests"]))
                emit("Memory Consumption: %d " % rec["memory"])
                records.append(rec)
            path = os.path.join(out_dir, file_name(mode, j > 0, j == 2))
            write_png(path, pt.read_rgba8(tone_mapping=(j == 0)))  # getToneMapping(): on for the render, off for the heat maps
            records[-1]["files"].append(path)
    pt.use_ref_bvh(False)
    return records
