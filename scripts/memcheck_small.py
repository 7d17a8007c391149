#!/usr/bin/env python
"""Small end-to-end exercise of the device code for `compute-sanitizer --tool memcheck` (one tool, one call; B200_PROFILING.md):
every mode of the path kernel with and without NEE on Cornell at 48x32, the reference's grids uploaded, the wide tree
(PTB_WIDE=1 in the environment), primary hits, trace_rays + the static-scratch entry point, the display transform, state
dump / restore (non-uniform).  Prints OK at the end; any sanitizer report shows up in its own output."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import ptb_b200  # noqa: E402
from oracle_lib import Oracle  # noqa: E402

scene = os.path.join(ROOT, "tests", "golden", "cornel.ptbs")
W, H = 48, 32
pt = ptb_b200.PathTracer(0)
pt.load_ptbs(scene)
pt.build_bvh()
pt.build_subgrid(10, 5, 40, gpu=True)
pt.build_subbeams(6, 40, gpu=True)
for mode in (0, 1, 2, 3):
    for nee in (0, 1):
        pt.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=4, count_tests=nee)
        pt.reset()
        pt.render(2)
        pt.render(1)
        pt.primary_hits()
o = Oracle(scene)
o.build(grid=True, gs=10, os_=5, m=40)
g, ls, le, idx = o.export_grids()
o.close()
pt.upload_subgrid(g.view(ptb_b200.REF_GRID), ls, le, idx)
pt.set_params(width=W, height=H, mode=2, nee=1, max_depth=0, count_tests=1, debug_view=1, debug_scale=10)
pt.reset()
pt.render(1)
pt.set_params(width=W, height=H, mode=1, nee=1, max_depth=0, count_tests=0, debug_view=0)
pt.reset()
pt.render(3)
img = pt.read_image().copy()
s, r = pt.read_state()
s = s.copy()
s[::2] -= 1  # non-uniform restore
pt.write_state(img, s, r)
pt.render(2)
pt.read_rgba8(True)
pt.read_rgba8(False)
rng = np.random.default_rng(1)
od = np.concatenate([rng.normal(size=(5000, 3)) * 400, rng.normal(size=(5000, 3))], axis=1).astype(np.float32)
for mode in (0, 1, 2, 3):
    pt.trace_rays(od[:500] if mode == 0 else od, mode)
pt.trace_rays_static(od)
pt.close()
print("OK")
