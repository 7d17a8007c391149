#!/bin/bash
# k_mega work-item size vs pixels per GPU (frame sizes emulate 1/N of the frame on one GPU)
mkdir -p gpurun_out
for res in 1024x1024 512x512 512x256; do for ch in 1 2 4 8 16; do
  PTB_MEGA_CHUNK=$ch python bench.py --workload cornell --res $res --steps 5 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('cornell $res chunk=$ch', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done | tee gpurun_out/sweep_chunk.txt
for res in 1920x1080 960x540; do for ch in 1 4 8; do
  PTB_MEGA_CHUNK=$ch python bench.py --workload lucy --res $res --steps 3 --warmup 3 --no-cpu 2>/dev/null | \
    python -c "import sys,json; r=json.loads(sys.stdin.read()); print('lucy $res chunk=$ch', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'])"
done; done | tee -a gpurun_out/sweep_chunk.txt
