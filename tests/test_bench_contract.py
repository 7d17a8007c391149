"""bench.py's contract without a GPU: the default workload is the scene north_star names, both arms print the same `config`
(the driver compares them), the auto sharding rule, and the reference arm's JSON line (the unmodified reference on the host cores
through oracle/_ref/ptref, or the oracle port where it is absent)."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _args(**kw):
    d = dict(workload="lucy256", structure="bvh", builder="host", shard="auto", gpus=1)
    d.update(kw)
    return argparse.Namespace(**d)


def test_default_workload_is_the_lucy_frame_of_config_3():
    wl = bench.WORKLOADS["lucy256"]
    assert (wl["W"], wl["H"], wl["spp"], wl["max_depth"], wl["mode"]) == (1920, 1080, 256, 9, 1)
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert 'ap.add_argument("--workload", default="lucy256"' in src


def test_auto_sharding_rule():
    lucy, cornell = bench.WORKLOADS["lucy256"], bench.WORKLOADS["cornell"]
    assert [bench.choose_shard(_args(), lucy, n) for n in (1, 2, 4, 8)] == ["tiles", "tiles", "samples", "samples"]
    assert [bench.choose_shard(_args(), cornell, n) for n in (1, 2, 8)] == ["tiles", "samples", "samples"]
    assert bench.choose_shard(_args(shard="tiles"), lucy, 8) == "tiles"
    assert bench.choose_shard(_args(shard="samples"), lucy, 1) == "tiles"  # one GPU: nothing to shard


def test_both_arms_print_the_same_config():
    a = _args()
    for world in (1, 2, 8):
        wl = bench.WORKLOADS[a.workload]
        ours = bench.workload_config(a, wl, world, bench.choose_shard(a, wl, world), 524310)
        ref = bench.workload_config(a, wl, world, bench.choose_shard(a, wl, world), 524310)
        assert ours == ref and {"workload", "width", "height", "spp", "max_depth", "structure", "triangles", "parallelism"} <= set(ours)


def test_reference_arm_line_on_cornell():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "cornell", "--steps", "1",
                          "--warmup", "0", "--ref-spp", "1"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    a = _args(workload="cornell")
    wl = bench.WORKLOADS["cornell"]
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["config"] == bench.workload_config(a, wl, 1, "tiles", 202)
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["reference_counted_rays_per_step"] >= d["rays_per_step"] > 0
