#!/bin/bash
# BASELINE.json configs[3] as written: Lucy stand-in, Hybrid (Subgrid gs 8 / os 8 / m 50) and Subbeam structures (+ the BVH for
# comparison), 1920x1080, 256 spp, 8 bounces, TILE-sharded at 1/2/4/8 GPUs of one box, with the reduced N-GPU frame compared
# bit for bit with a 1-GPU frame inside the run (bench.py `parity.n_gpu_tile_frame_equals_1_gpu_frame`).
# Subbeams at this scene size trace at single-digit Mrays/s (depth guard 1), so they run at SUBBEAM_SPP (default 2).
# usage (under gpurun --gpus 8):  scripts/config3_run.sh [out.jsonl]   ; GPUS="1 2 4 8" STRUCTS="subgrid bvh subbeams"
OUT=${1:-gpurun_out/r2_config3_lines.jsonl}
: > "$OUT"
for st in ${STRUCTS:-subgrid bvh subbeams}; do
  spp_arg=""
  steps=${STEPS:-2}
  [ "$st" = subbeams ] && spp_arg="--spp ${SUBBEAM_SPP:-2}" && steps=1
  for n in ${GPUS:-1 2 4 8}; do
    echo "== $st x $n GPUs" >&2
    if [ "$n" = 1 ]; then
      timeout 1500 python bench.py --workload lucy256 --structure $st --shard tiles --steps $steps --warmup 3 --no-cpu $spp_arg >> "$OUT" 2>> gpurun_out/config3.err
    else
      timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
        bench.py --gpus $n --workload lucy256 --structure $st --shard tiles --steps $steps --warmup 3 --no-cpu $spp_arg >> "$OUT" 2>> gpurun_out/config3.err
    fi
  done
done
python - "$OUT" <<'P'
import json, sys
base = {}
for l in open(sys.argv[1]):
    try: d = json.loads(l)
    except Exception: continue
    st = d["config"]["structure"].split()[0]
    if d["n_gpus"] == 1: base[st] = d["value"]
    p = d.get("parity") or {}
    print(f"{st:9s} N={d['n_gpus']}  {d['value']:9.1f} Mrays/s  x{d['value'] / base.get(st, d['value']):.2f}  {d['spp_per_s']:8.2f} spp/s  ms/step {d['ms_per_step']:9.2f}  "
          f"tiles==1gpu {p.get('n_gpu_tile_frame_equals_1_gpu_frame')}  hit-id agreement {p.get('hit_id_agreement')}  t digest {p.get('t_digest_equal')}")
P
