#!/bin/bash
# usage: scripts/sweep_env.sh VAR "v1 v2 ..." [workloads]   — one bench line per value of an env tuning knob
VAR=$1; VALS=$2; WLS=${3:-"cornell lucy"}
mkdir -p gpurun_out
for wl in $WLS; do for v in $VALS; do
  env $VAR=$v python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl $VAR=$v', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'], 'extend_share=%.3f'%r['roofline']['share_of_step'], 'overlap=%.2f'%r['roofline']['overlap'], 'frac=%.3f'%r['roofline']['frac'])"
done; done | tee -a gpurun_out/sweep_$VAR.txt
