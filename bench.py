#!/usr/bin/env python
"""bench.py — the hot path of the thesis path tracer on N B200s, one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cornell|lucy|lucy256]
                    [--structure bvh|subgrid|subbeams] [--builder host|gpu] [--shard auto|tiles|samples]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Default workload = the scene BASELINE.json's north_star states its target on: the Lucy stand-in (524,310 triangles: the
Cornell room with a displaced 8x-subdivided octahedron sphere; the real lucy.obj is not shipped),
BVH, 1920x1080, 256 spp, primary + 8 bounces (max_depth 9) — the frame of BASELINE configs[3] — NEE off like the
reference's own --testing protocol (src/main.cpp:101).  --workload cornell = configs[1] (Cornell box, BVH, 1024x1024,
64 spp, 4 bounces: the round-1 default); --workload lucy = the Lucy frame at 16 spp (profiling runs).
A "step" is one full frame: initTracer (ptb_reset) + trace (ptb_render).  With N > 1 the same frame is sharded over
the ranks (strong scaling: total work fixed; tiles while every GPU keeps >= 600k pixels, else sample ranges) and the
radiance is sum-reduced to rank 0 over NCCL inside the timed region.  Outside the timed region the line also gets
`parity`: primary-ray hit ids / hit distances of this very process against the golden written by the unmodified reference
(tests/golden/lucy_standin_ref_1080p.npz), and for a tile-sharded run the bitwise comparison of the reduced frame with a
1-GPU frame rendered by rank 0.

  value      Mrays/s, rays = traversals executed (primary + bounce; the reference's re-traced
             misses are NOT counted: SURVEY.md §8d "unique rays"), state resident in HBM, device time
             (CUDA events on the context's stream, max over ranks)
  e2e        same metric through the C ABI with host buffers: every step uploads the material table
             and camera (what the reference's UI does before reset()), renders, and reads the fp32
             frame back to host memory; wall clock around synchronous calls, max over ranks
  roofline   dominant kernel k_mega<BVH> (one fused launch per render: traversal + shade + accumulate;
             k_extend_bvh with PTB_MEGA=0): algorithmic bytes (48 B ray+hit, 32 B per box
             test, 36 B per triangle test, from the instrumented counters) / its CUDA-event time inside
             the timed region, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference
             the UNMODIFIED reference compiled into oracle/_ref/ptref (trace(true), its 16
             std::threads) on the host cores, bounded sample = same frame at fewer spp
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

CORNEL = os.path.join(ROOT, "tests", "golden", "cornel.ptbs")
METRIC = "Mrays/s (primary+bounce), full path tracing"

WORKLOADS = {
    # name: (W, H, spp, max_depth, mode, description)
    "cornell": dict(W=1024, H=1024, spp=64, max_depth=5, mode=1,
                    name="Cornell box, 1024x1024, 64 spp, 4 bounces (BASELINE configs[1])"),
    "lucy": dict(W=1920, H=1080, spp=16, max_depth=9, mode=1,
                 name="Lucy stand-in (524,310 triangles), 1920x1080, 16 spp, 8 bounces (configs[3] at 1/16 of the samples)"),
    "lucy256": dict(W=1920, H=1080, spp=256, max_depth=9, mode=1,
                    name="Lucy stand-in (524,310 triangles), 1920x1080, 256 spp, 8 bounces (BASELINE configs[3] frame; north_star target scene)"),
}


def workload_config(args, wl, world, shard, triangles):
    """`config` of the JSON line: the same keys and values in both arms (ours / reference)."""
    structure = {"bvh": "BVH" + (" (ours: BVH2 built on the GPU, Morton radix tree)" if args.builder == "gpu" else ""),
                 "subgrid": "Subgrid / 2Plane hierarchy gs 8 os 8 m 50", "subbeams": "Subbeams gs 6 m 50 depth guard 2"}[args.structure]
    if world == 1:
        par = "1 GPU"
    elif shard == "tiles":
        par = f"tiles32x8 round-robin over {world} GPUs, NCCL sum-reduce to rank 0 (bit-identical to 1 GPU)"
    else:
        par = f"sample ranges: {world} GPUs x all pixels x spp/{world} from per-rank seeds, NCCL sum-reduce (statistical parity)"
    return {"workload": wl["name"], "width": wl["W"], "height": wl["H"], "spp": wl["spp"], "max_depth": wl["max_depth"], "nee": 0,
            "structure": structure, "triangles": int(triangles), "parallelism": par,
            "l2": "flushed between timed steps (512 MiB fill)"}


def choose_shard(args, wl, world):
    shard = args.shard
    if shard == "auto":
        shard = "tiles" if world == 1 or (wl["W"] * wl["H"]) // world >= 600_000 else "samples"
    return shard if world > 1 else "tiles"


def ncu_counters(args):
    """Measured bound of the path kernel from the committed ncu capture of the same workload / structure
    (profiles/r2_ncu_counters.json, written by scripts/ncu_counters.py from the .ncu-rep of `bench.py --spp ...`)."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "r2_ncu_counters.json")))
        key = ("cornell" if args.workload == "cornell" else "lucy") + "_" + args.structure
        return tab.get(key)
    except Exception:
        return None


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.idx}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); pw.append(float(r[3]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the unmodified reference on the host cores
# --------------------------------------------------------------------------------------------------
def scene_file(workload, tmpdir):
    """Path of the workload's .ptbs (Cornell ships as a fixture; the Lucy stand-in is generated)."""
    if workload == "cornell":
        return CORNEL
    workload = "lucy"
    import ptb_b200
    path = os.path.join(tmpdir, "lucy_standin.ptbs")
    if not os.path.isfile(path):
        ptb_b200.scenes.write_ptbs(path, ptb_b200.scenes.lucy_standin(ptb_b200.scenes.read_ptbs(CORNEL)))
    return path


def cpu_reference_passes(ptbs, wl, spp, timed_passes):
    """`timed_passes` uninstrumented renders + one counted render of the same deterministic frame in ONE ptref process
    (oracle/ref/ref_driver.cpp --repeat): the scene load and the reference's constructBVH (26 s on the stand-in) are paid once.
    Returns ([trace seconds per timed pass], counted record)."""
    import oracle_lib
    args = ["--ptbs", ptbs, "--mode", wl["mode"], "--width", wl["W"], "--height", wl["H"], "--spp", spp, "--max-depth", wl["max_depth"],
            "--nee", 0, "--repeat", timed_passes + 1, "render", "-"]
    recs = [r for r in oracle_lib.run_ptref(args, timeout=6 * 3600) if r["cmd"] == "render"]
    return [r["trace_ms"] / 1e3 for r in recs[:-1]], recs[-1]


def cpu_reference_run(ptbs, wl, spp):
    """One bounded sample on the CPU.  Returns (unique rays, seconds, kind, threads).
    kind "reference": oracle/_ref/ptref = the reference's own sources; timing from an uninstrumented
    run (--no-count), ray count from a counted run of the same deterministic sample.
    kind "port": the C++ restatement (oracle/libptoracle.so) when the compiled reference is absent."""
    import oracle_lib
    ncpu = os.cpu_count() or 1
    if oracle_lib.have_ref():
        secs, counted = cpu_reference_passes(ptbs, wl, spp, 1)
        return counted["rays_unique"], secs[0], "reference", min(16, ncpu), counted["rays_traversed"]
    o = oracle_lib.Oracle(ptbs)
    o.build()
    o.configure(wl["W"], wl["H"], wl["mode"], nee=0, max_depth=wl["max_depth"])
    t0 = time.perf_counter()
    _, _, _, st = o.render(spp, threads=ncpu)
    dt = time.perf_counter() - t0
    o.close()
    return st["rays_unique"], dt, "port", ncpu, st["rays_traversed"]


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import oracle_lib
    wl = WORKLOADS[args.workload]
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    import tempfile
    ptbs = scene_file(args.workload, tempfile.gettempdir())
    spp = args.ref_spp
    ncpu = os.cpu_count() or 1
    if oracle_lib.have_ref():
        secs, counted = cpu_reference_passes(ptbs, wl, spp, args.warmup + args.steps)
        secs = secs[args.warmup:]
        rays, counted_rays, kind, cores = counted["rays_unique"], counted["rays_traversed"], "reference", min(16, ncpu)
    else:
        secs = []
        for k in range(args.warmup + args.steps):
            rays, dt, kind, cores, counted_rays = cpu_reference_run(ptbs, wl, spp if k >= args.warmup else 1)
            if k >= args.warmup:
                secs.append(dt)
    total = sum(secs)
    v = rays * len(secs) / total / 1e6
    sample = (f"{wl['W']}x{wl['H']} frame at {spp} spp per step (of {wl['spp']}), max_depth {wl['max_depth']}, NEE off; "
              f"{'reference trace(true), 16 std::threads' if kind == 'reference' else 'oracle port'}; structure build outside the timed region "
              "like src/main.cpp:97 vs :116")
    triangles = 202 if args.workload == "cornell" else 524310
    emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total / len(secs) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, wl, world, choose_shard(args, wl, world), triangles),
        "spp_per_s": spp * len(secs) / total,
        "rays_per_step": rays, "reference_counted_rays_per_step": counted_rays,
        "value_reference_counted": counted_rays * len(secs) / total / 1e6,
        "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample, "host_cpus": ncpu,
                         "threads": 16 if kind == "reference" else cores},
        "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))
    return 0


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import ptb_b200
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libptb has no CPU fallback; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = WORKLOADS[args.workload]
    W, H, spp, mode = wl["W"], wl["H"], wl["spp"], wl["mode"]
    sc = ptb_b200.scenes
    scene = sc.read_ptbs(CORNEL)
    if args.workload.startswith("lucy"):
        scene = sc.lucy_standin(scene)
    mats_bytes, n_mats = sc.pack_materials(scene["materials"]), len(scene["materials"])
    cam = scene["camera"]

    pt = ptb_b200.PathTracer(local)
    sc.upload_scene(pt, scene)
    t0 = time.perf_counter()
    pt.build_bvh(gpu=args.builder == "gpu")
    if args.structure == "subgrid":
        pt.build_subgrid(8, 8, 50)
        mode = 2
    elif args.structure == "subbeams":
        pt.build_subbeams(6, 50, max_depth=1)
        mode = 3
    build_s = time.perf_counter() - t0
    # Sharding: tiles (bit-identical to the 1-GPU frame) while every GPU keeps >= 600k pixels in flight; below
    # that the serial samples of a pixel starve the lanes and the frame is sharded by sample ranges instead
    # (every rank renders all pixels, spp/N samples from its own seeds; statistical parity)
    shard = choose_shard(args, wl, world)
    if shard == "samples" and world > 1:
        base = dict(width=W, height=H, mode=mode, nee=0, max_depth=wl["max_depth"], rank=0, world=1, tile_w=32, tile_h=8, seed_offset=rank)
        my_spp = ptb_b200.multigpu.sample_share(spp, rank, world)
    else:
        shard = "tiles"
        base = dict(width=W, height=H, mode=mode, nee=0, max_depth=wl["max_depth"], rank=rank, world=world, tile_w=32, tile_h=8)
        my_spp = spp
    stream = torch.cuda.ExternalStream(pt.stream_handle(), device=local)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=f"cuda:{local}")  # > 126 MB L2

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_frame():
        if dist is None:
            return
        img = ptb_b200.multigpu.device_image_tensor(pt)
        with torch.cuda.stream(stream):
            if shard == "samples":
                ptb_b200.multigpu.combine_sample_shards(img, my_spp, spp, dst=0)
            else:
                ptb_b200.multigpu.reduce_image(img, dst=0)

    # instrumented pass (outside the timed region): the per-ray visit counts behind the algorithmic bytes
    pt.set_params(count_tests=1, **base)
    pt.reset()
    pt.render(2)
    c = pt.counters()
    boxes_per_ray = c["as_tests"] / max(1, c["rays"])
    tris_per_ray = c["tri_tests"] / max(1, c["rays"])
    bytes_per_ray = 48.0 + 32.0 * boxes_per_ray + 36.0 * tris_per_ray
    if args.structure != "bvh":  # per structure test ~ half a 64-B header + half an 8-B LUT cell; 4-B index + 36-B triangle
        bytes_per_ray = 48.0 + 36.0 * boxes_per_ray + 40.0 * tris_per_ray

    pt.set_params(count_tests=0, **base)
    pt.set_profiling(True)

    def step():
        pt.reset()
        pt.render(my_spp)
        reduce_frame()

    # clocks + throttle reasons are sampled from the warm-up to the end of the e2e leg: the timed region can be
    # shorter than nvidia-smi's sampling period (8 GPUs: 6 ms per step), the load before and after it is the same
    clocks = ClockSampler(local)
    clocks.start()
    for _ in range(args.warmup):
        step()
    # ---- value: device time, state resident in HBM --------------------------------------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    rays = paths = kernels = ref_rays = 0
    ext_ms = logic_ms = 0.0
    ext_launches = 0
    barrier()
    for k in range(args.steps):
        flush.fill_(k & 0xFF)  # L2 flush between timed iterations (torch stream; not inside the event pair)
        torch.cuda.synchronize()
        ev[k][0].record(stream)
        step()
        ev[k][1].record(stream)
        cc = pt.counters()
        st = pt.stage_ms()
        rays += cc["rays"]; paths += cc["paths"]; kernels += cc["kernels"]; ref_rays += cc["ref_rays"]
        ext_ms += st["extend_ms"]; logic_ms += st["logic_ms"]; ext_launches += st["extend_launches"]
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)

    # ---- e2e: through the C ABI with host buffers, copies inside the timed region ---------------
    host_img = torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True).numpy()
    e2e_rays = 0
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        pt.upload_materials(mats_bytes, n_mats)                    # H2D: material table (UI edit + reset flow)
        pt.set_camera(cam["pos"], cam["forward"], cam["focus"], cam["dof"])
        pt.set_params(count_tests=0, **base)
        pt.reset()
        pt.render(my_spp)
        reduce_frame()
        if rank == 0:
            pt.L.ptb_read_image(pt.h, host_img.ctypes.data)        # D2H: the fp32 frame
        e2e_rays += pt.counters()["rays"]
    barrier()
    e2e_s = time.perf_counter() - t0
    clk = clocks.stop()
    h2d = len(mats_bytes) + 36 + 56
    d2h = W * H * 16

    # ---- aggregate over ranks -------------------------------------------------------------------
    # ---- parity of THIS process against the reference's golden, outside the timed region -------------------------------
    parity = None
    tiles_equal = None
    if world > 1 and shard == "tiles":
        # the reduced N-GPU frame (rank 0 holds it after the last e2e step) against a 1-GPU frame of rank 0, bit for bit
        reduced = host_img.copy() if rank == 0 else None
        barrier()
        if rank == 0:
            pt.set_params(count_tests=0, **dict(base, rank=0, world=1))
            pt.reset()
            pt.render(spp)
            single = pt.read_image()
            tiles_equal = bool((single.view(np.uint32) == reduced.view(np.uint32)).all())
        barrier()
    if rank == 0 and args.workload.startswith("lucy"):
        import hashlib
        gpath = os.path.join(ROOT, "tests", "golden", "lucy_standin_ref_1080p.npz")
        if os.path.isfile(gpath) and W == 1920 and H == 1080:
            g = np.load(gpath)
            pt.set_params(count_tests=0, **dict(base, rank=0, world=1, seed_offset=0))
            t_hit, tri_hit, _ = pt.primary_hits()
            agree = float((tri_hit == g["tri"]).mean())
            parity = {"against": "tests/golden/lucy_standin_ref_1080p.npz (unmodified reference, BVH mode, primary rays of every pixel)",
                      "hit_id_agreement": agree, "hit_ids_differing": int((tri_hit != g["tri"]).sum()),
                      "t_digest_equal": hashlib.sha256(t_hit.tobytes()).hexdigest() == str(g["t_sha256"]),
                      "structure_checked": args.structure}
    elif rank == 0:
        import hashlib
        g = np.load(os.path.join(ROOT, "tests", "golden", "configs_ref.npz"))
        pt.set_params(count_tests=0, **dict(base, rank=0, world=1, seed_offset=0, width=512, height=512))
        t_hit, tri_hit, _ = pt.primary_hits()
        parity = {"against": "tests/golden/configs_ref.npz c0 (unmodified reference, Cornell 512x512 primary rays)",
                  "hit_id_agreement": float((tri_hit == g["c0_tri"]).mean()), "hit_ids_differing": int((tri_hit != g["c0_tri"]).sum()),
                  "t_digest_equal": hashlib.sha256(t_hit.tobytes()).hexdigest() == str(g["c0_t_sha256"]), "structure_checked": args.structure}
    if parity is not None and tiles_equal is not None:
        parity["n_gpu_tile_frame_equals_1_gpu_frame"] = tiles_equal

    vals = torch.tensor([float(rays), float(paths), float(kernels), float(e2e_rays), float(ref_rays)], dtype=torch.float64, device=f"cuda:{local}")
    tmax = torch.tensor([dev_ms, e2e_s, ext_ms, logic_ms], dtype=torch.float64, device=f"cuda:{local}")
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.SUM)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    rays_all, paths_all, kernels_all, e2e_rays_all, ref_rays_all = (float(x) for x in vals.tolist())
    dev_ms_max, e2e_s_max, ext_ms_max, logic_ms_max = (float(x) for x in tmax.tolist())

    result = None
    if rank == 0:
        peak, peak_kind = peaks()
        traffic = traffic_src = None
        try:  # dram__bytes_read.sum + dram__bytes_write.sum of the same kernel from the committed ncu capture, scaled to this launch
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")))["lucy" if args.workload.startswith("lucy") else args.workload]
            traffic = tr["dram_bytes_per_ray"] * rays / max(1, ext_launches)
            traffic_src = tr["source"]
        except Exception:
            pass
        value = rays_all / (dev_ms_max / 1e3) / 1e6
        # roofline of the dominant kernel on rank 0's own launches.  The waves of a render overlap on the device,
        # so the per-launch CUDA-event durations of concurrent kernels overlap too: the step's device time is
        # attributed to extend / logic in proportion to their summed event durations (with one wave this is
        # the plain sum of the extend launches).
        ext_share = ext_ms / (ext_ms + logic_ms) if ext_ms + logic_ms > 0 else 0.0
        ext_time_ms = dev_ms * ext_share
        ach = bytes_per_ray * rays / (ext_time_ms / 1e3) / 1e9 if ext_time_ms > 0 else None
        kname = {"bvh": "k_mega<BVH>", "subgrid": "k_mega<TWO_PLANE>", "subbeams": "k_mega<SUB_BEAMS>"}[args.structure]
        roofline = {"bound": "hbm", "kernel": (kname + " (one fused launch per render: raygen + traversal + shade + accumulate)") if ext_launches == args.steps else "k_extend_bvh",
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None,
                    "frac_is": "ALGORITHMIC bytes (SURVEY.md 8d) / kernel time / measured HBM copy peak: nominal, the structure is L1/L2 resident; "
                               "the bound that binds is `lane_issue_frac` below",
                    "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_kind,
                    "bytes_per_ray": bytes_per_ray, "boxes_per_ray": boxes_per_ray, "tris_per_ray": tris_per_ray,
                    "avg_launch_ms": ext_time_ms / max(1, ext_launches), "launches": ext_launches,
                    "avg_launch_event_ms": ext_ms / max(1, ext_launches),
                    "share_of_step": ext_share, "logic_share_of_step": 1.0 - ext_share,
                    "overlap": (ext_ms + logic_ms) / dev_ms if dev_ms > 0 else None}
        nc = ncu_counters(args)
        if nc:  # measured by ncu on the same workload (committed capture; refreshed whenever the kernel changes)
            roofline.update({"binding_bound": "issue slots at the lane occupancy divergence leaves (not HBM)",
                             "lanes_per_inst": nc["lanes_per_inst"], "issue_active_pct": nc["issue_active_pct"],
                             "lane_issue_frac": nc["issue_active_pct"] / 100.0 * nc["lanes_per_inst"] / 32.0,
                             "dram_gbs": nc["dram_gbs"], "dram_frac_of_peak": nc["dram_gbs"] / peak, "l2_gbs": nc["l2_gbs"],
                             "l1_hit_pct": nc.get("l1_hit_pct"), "l2_hit_pct": nc.get("l2_hit_pct"), "ncu_source": nc["source"]})
            traffic = nc["dram_bytes_per_ray"] * rays / max(1, ext_launches)
            roofline["traffic"], roofline["traffic_source"] = traffic, nc["source"]
        result = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, wl, world, shard, scene["pos"].shape[0]),
            "structure_build_s": round(build_s, 3),
            "spp_per_s": spp * args.steps / (dev_ms_max / 1e3),
            "mpaths_per_s": paths_all / (dev_ms_max / 1e3) / 1e6,
            "rays_per_step": rays_all / args.steps,
            "reference_counted_rays_per_step": ref_rays_all / args.steps,  # incl. the reference's re-traced misses
            "value_reference_counted": ref_rays_all / (dev_ms_max / 1e3) / 1e6,
            "parity": parity,
            "gpu_launches": int(kernels_all),
            "clocks": clk,
            "e2e": {"value": e2e_rays_all / e2e_s_max / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_s_max / args.steps * 1e3},
            "roofline": roofline,
        }
    pt.close()
    if rank == 0 and world == 1 and not args.no_cpu:
        import tempfile
        ptbs = scene_file(args.workload, tempfile.gettempdir())
        r, s, kind, cores, r_counted = cpu_reference_run(ptbs, wl, args.cpu_spp)
        result["cpu_baseline"] = {
            "value": r / s / 1e6, "value_reference_counted": r_counted / s / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind, "host_cpus": os.cpu_count(),
            "threads": 16 if kind == "reference" else cores, "seconds": s,
            "sample": f"same {W}x{H} frame at {args.cpu_spp} spp (of {spp}), max_depth {wl['max_depth']}, NEE off, "
                      f"{'reference trace(true), 16 std::threads' if kind == 'reference' else 'oracle port'}"}
    if rank == 0:
        emit(json.dumps(result))
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


_RESULT_FD = None


def _claim_stdout():
    """stdout carries ONE JSON line.  Libraries write there too (NCCL prints its version banner to stdout when the box
    exports NCCL_DEBUG=VERSION), so file descriptor 1 is pointed at stderr for the run and the result line goes to the
    original stdout through emit()."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    if _RESULT_FD is None:
        print(line, flush=True)
    else:
        os.write(_RESULT_FD, (line + "\n").encode())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="lucy256", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-spp", type=int, default=0, help="spp of the bounded cpu_baseline sample (0: 8 for Cornell, 4 for the Lucy frame)")
    ap.add_argument("--ref-spp", type=int, default=0, help="spp per step of the --impl reference arm (0: 4 for Cornell, 2 for the Lucy frame = 3 s per step)")
    ap.add_argument("--structure", default="bvh", choices=["bvh", "subgrid", "subbeams"],
                    help="acceleration structure (subgrid: gs 8 / os 8 / m 50 like the published runs; subbeams: gs 6 / m 50 / depth guard 2)")
    ap.add_argument("--builder", default="host", choices=["host", "gpu"],
                    help="BVH construction: host binned SAH (default, the better tree) or ptb_build_bvh_gpu (Morton radix tree built in HBM)")
    ap.add_argument("--shard", default="auto", choices=["auto", "tiles", "samples"], help="multi-GPU partition of the frame")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--res", default="", help="override the frame size WxH (experiments only)")
    ap.add_argument("--spp", type=int, default=0, help="override the workload's samples per pixel (profiling runs only)")
    args = ap.parse_args()
    if args.spp > 0:
        WORKLOADS[args.workload] = dict(WORKLOADS[args.workload], spp=args.spp,
                                        name=WORKLOADS[args.workload]["name"] + f" [spp overridden to {args.spp}]")
    if args.res:
        rw, rh = (int(v) for v in args.res.lower().split("x"))
        WORKLOADS[args.workload] = dict(WORKLOADS[args.workload], W=rw, H=rh,
                                        name=WORKLOADS[args.workload]["name"] + f" [frame overridden to {rw}x{rh}]")
    if args.ref_spp <= 0:
        args.ref_spp = 4 if args.workload == "cornell" else 2
    if args.cpu_spp <= 0:
        args.cpu_spp = 8 if args.workload == "cornell" else 4
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not (args.impl == "ours" and args.gpus > 1 and world == 1):  # (the re-launching parent passes the children's stdout through)
        _claim_stdout()
    if args.impl == "reference":
        return run_reference(args)
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun like the driver does
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29517", os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
