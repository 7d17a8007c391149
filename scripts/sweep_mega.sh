#!/bin/bash
# k_mega: register cap variants (libptb_mb{4,6,8}.so = -DPTB_MEGA_MINBLOCKS) x logic_keep
mkdir -p gpurun_out
P=bachelor-thesis-pathtracer_b200
for wl in cornell lucy; do
for mb in ${MBS:-8 10 12}; do
for keep in ${KEEPS:-24 28 32}; do
  lib=libptb_mb$mb.so; [ $mb = 10 ] && lib=libptb.so
  PTB_LIB=$PWD/$P/$lib PTB_MEGA_CTAS=$mb PTB_LOGIC_KEEP=$keep python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl minblocks=$mb keep=$keep', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done; done | tee gpurun_out/sweep_mega.txt
