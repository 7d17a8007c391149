#!/bin/bash
# k_duo (PTB_MEGA=3): scheduler thresholds (PTB_DUO_LOGIC/_SAMPLE/_WALK/_REFILL/_LEAF) and resident CTAs
mkdir -p gpurun_out
run() { # workload, env...
  wl=$1; shift
  env "$@" python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl $*', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
}
for wl in cornell lucy; do
  run $wl PTB_MEGA=1
  for lf in ${LFS:-1 2 4 8 12 16 24 33}; do run $wl PTB_MEGA=3 PTB_DUO_SAMPLE=24 PTB_DUO_LEAF=$lf; done
  for lk in ${LKS:-}; do for wm in ${WMS:-12 20 26}; do
    run $wl PTB_MEGA=3 PTB_DUO_LOGIC=$lk PTB_DUO_WALK=$wm
  done; done
  for sk in ${SKS:-}; do run $wl PTB_MEGA=3 PTB_DUO_SAMPLE=$sk; done
  for rf in ${RFS:-}; do run $wl PTB_MEGA=3 PTB_DUO_REFILL=$rf; done
  for ct in ${CTS:-}; do run $wl PTB_MEGA=3 PTB_DUO_CTAS=$ct; done
done | tee gpurun_out/sweep_duo_${TAG:-x}.txt
