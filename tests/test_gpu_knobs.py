"""Every tuning knob changes scheduling or the shape of the structure, never the result: closest hits do not depend on the tree or
on the order of visits (ties go to the larger reference id), and a pixel's samples are serial whatever lane runs them.  Each
setting runs in its own process (the knobs are read from the environment at ptb_create / build time) and must reproduce the default
build's frame, RNG states and ray counts bit for bit — including the opt-in 4-wide BVH (PTB_WIDE=1)."""
import hashlib
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import hashlib, json, os, sys
sys.path.insert(0, %r)
import numpy as np, ptb_b200
sc = ptb_b200.scenes
cornel = os.path.join(%r, "tests/golden/cornel.ptbs")
out = {}
for name in ("cornell", "lucy5"):
    pt = ptb_b200.PathTracer(0)
    if name == "cornell":
        pt.load_ptbs(cornel)
    else:
        sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(cornel), level=5))
    pt.build_bvh()
    pt.build_subgrid(6, 4, 40, gpu=True)
    for mode, nee in ((1, 1), (2, 0)):
        pt.set_params(width=200, height=150, mode=mode, nee=nee, max_depth=6)
        pt.reset(); pt.render(3); pt.render(2)
        s, r = pt.read_state()
        c = pt.counters()
        h = hashlib.sha256(pt.read_image().tobytes() + s.tobytes() + r.tobytes()).hexdigest()
        out["%%s_m%%d" %% (name, mode)] = [h, c["rays"], c["ref_rays"], c["paths"]]
    t, tri, _ = pt.primary_hits()
    out[name + "_hits"] = hashlib.sha256(t.tobytes() + tri.tobytes()).hexdigest()
    pt.close()
print("REC " + json.dumps(out))
''' % (ROOT, ROOT)

KNOBS = [
    {"PTB_WIDE": "1"},
    {"PTB_WIDE": "1", "PTB_MEGA_LEAF": "0"},
    {"PTB_BVH_BINS": "8"},
    {"PTB_BVH_BINS": "64", "PTB_BVH_RELAYOUT": "0"},
    {"PTB_LEAF_MAX": "5"},
    {"PTB_LOGIC_KEEP": "12", "PTB_MEGA_LEAF": "24"},
    {"PTB_MEGA_CTAS": "5", "PTB_MEGA_CHUNK": "3"},
    {"PTB_MEGA": "0"},
]


def _run(env_extra):
    env = {k: v for k, v in os.environ.items() if not k.startswith("PTB_")}
    env.update(env_extra)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    rec = [l for l in r.stdout.splitlines() if l.startswith("REC ")]
    assert r.returncode == 0 and rec, r.stderr[-2000:]
    return json.loads(rec[0][4:])


@pytest.fixture(scope="module")
def default_frames():
    return _run({})


@pytest.mark.parametrize("knobs", KNOBS, ids=[" ".join(f"{k}={v}" for k, v in kn.items()) for kn in KNOBS])
def test_knob_settings_reproduce_the_default_frames(default_frames, knobs):
    got = _run(knobs)
    assert got == default_frames
