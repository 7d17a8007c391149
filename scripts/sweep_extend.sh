#!/bin/bash
# tuning sweep of the persistent BVH extend (PTB_EXTEND_CTAS x PTB_FETCH_KEEP); prints one line per point
mkdir -p gpurun_out
for wl in cornell lucy; do
for ctas in ${CTAS:-8 12}; do
for keep in ${KEEPS:-1 12 20 26 32}; do
  PTB_EXTEND_CTAS=$ctas PTB_FETCH_KEEP=$keep python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl ctas=$ctas keep=$keep', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'], 'extend_share=%.3f'%r['roofline']['share_of_step'], 'ext_launch_ms=%.4f'%r['roofline']['avg_launch_ms'], 'frac=%.3f'%r['roofline']['frac'])"
done; done; done | tee gpurun_out/sweep_extend.txt
