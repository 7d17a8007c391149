#!/bin/bash
mkdir -p gpurun_out
P=bachelor-thesis-pathtracer_b200
for wl in cornell lucy; do
for v in "8 64" "6 64" "10 48" "8 48"; do set -- $v; mb=$1; np=$2
for keep in ${KEEPS:-24 32}; do
  lib=libptb_pool${mb}_${np}.so; [ "$v" = "8 64" ] && lib=libptb.so
  PTB_LIB=$PWD/$P/$lib PTB_POOL_CTAS=$mb PTB_LOGIC_KEEP=$keep python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl minblocks=$mb NP=$np keep=$keep', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done; done | tee gpurun_out/sweep_pool.txt
