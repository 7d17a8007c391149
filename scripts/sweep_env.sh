#!/bin/bash
# generic one-knob sweep: scripts/sweep_env.sh VAR "v1 v2 ..." [workloads] [extra bench args]
VAR=$1; VALS=$2; WLS=${3:-cornell lucy}; if [ $# -ge 3 ]; then shift 3; else shift $#; fi
mkdir -p gpurun_out
for wl in $WLS; do
for v in $VALS; do
  env $VAR=$v python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu "$@" 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl $VAR=$v', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done | tee -a gpurun_out/sweep_env_$VAR.txt
