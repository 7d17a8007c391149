#!/usr/bin/env python
"""ptb_group_* scaling inside ONE process: the Lucy stand-in at 1920x1080, 8 bounces, tile-sharded over the first
n GPUs of the box (n = 1, 2, 4, 8 as available), gathered on member 0; every n-GPU frame must equal the 1-GPU frame
bit for bit; then the same with sample-range sharding (every member all pixels, spp/n samples from its own seeds:
statistical parity, RMSE against the tile frame reported).  usage: group_scaling.py [spp=64].  Writes
gpurun_out/group_scaling.json."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
sc = ptb_b200.scenes
scene = sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests", "golden", "cornel.ptbs")))
W, H = 1920, 1080
avail = torch.cuda.device_count()
res, ref = [], None
for samples_mode, n in [(False, k) for k in (1, 2, 4, 8)] + [(True, k) for k in (2, 4, 8)]:
    if n > avail:
        continue
    g = ptb_b200.PathTracerGroup(list(range(n)))
    g.set_sharding(samples_mode)

    def prep(m):
        sc.upload_scene(m, scene)
        m.build_bvh()
    g.each(prep)
    g.set_params(width=W, height=H, mode=1, nee=0, max_depth=9)
    g.reset()
    g.render(2)  # warm-up
    ms, wall, gather = [], [], []
    for _ in range(3):
        g.reset()
        t0 = time.perf_counter()
        ms.append(g.render(spp))
        t1 = time.perf_counter()
        img = g.read_image()
        t2 = time.perf_counter()
        wall.append((t1 - t0) * 1e3)
        gather.append((t2 - t1) * 1e3)
    c = g.counters()
    if ref is None:
        ref = img
    rmse = float(np.sqrt(np.mean((np.minimum(img[..., :3], 4.0) - np.minimum(ref[..., :3], 4.0)) ** 2)))
    rec = {"gpus": n, "sharding": "samples" if samples_mode else "tiles", "rmse_vs_1gpu_frame_clamped4": rmse, "spp": spp, "device_ms_max": min(ms), "wall_ms": min(wall), "gather_and_d2h_ms": min(gather), "rays": c["rays"],
           "mrays_s_device": c["rays"] / min(ms) / 1e3, "mrays_s_wall": c["rays"] / min(wall) / 1e3, "peer_direct": g.peer_direct(),
           "identical_to_1gpu": bool((img.view(np.uint32) == ref.view(np.uint32)).all())}
    res.append(rec)
    print(rec, flush=True)
    g.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "group_scaling.json"), "w"), indent=1)
