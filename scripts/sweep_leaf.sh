#!/bin/bash
# k_mega: straggler threshold of the node loop (PTB_MEGA_LEAF, in 1/32 of the walking lanes; 0 = off)
mkdir -p gpurun_out
for wl in cornell lucy; do
for lf in ${LFS:-0 4 8 12 16 20 24}; do
  PTB_MEGA_LEAF=$lf python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl mega leaf=$lf', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done | tee gpurun_out/sweep_mega_leaf.txt
