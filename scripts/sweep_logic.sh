#!/bin/bash
# k_logic register cap variants (libptb_lb{6,8}.so built with -DPTB_LOGIC_MINBLOCKS)
mkdir -p gpurun_out
P=bachelor-thesis-pathtracer_b200
for wl in cornell lucy; do
for lib in libptb.so libptb_lb6.so libptb_lb8.so; do
  PTB_LIB=$PWD/$P/$lib python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl $lib', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'], 'extend_share=%.3f'%r['roofline']['share_of_step'], 'logic_share=%.3f'%r['roofline']['logic_share_of_step'], 'ext_launch_ms=%.4f'%r['roofline']['avg_launch_ms'])"
done; done | tee gpurun_out/sweep_logic.txt
