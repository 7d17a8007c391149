#!/bin/bash
# strong scaling on one box: prints one line per (workload, N)
mkdir -p gpurun_out
run() { wl=$1; n=$2; steps=$3; shift 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps $steps --warmup 3 --workload $wl --no-cpu "$@" 2>/dev/null | grep "^{" | tee -a gpurun_out/scale8_lines.jsonl | \
  python -c "import sys,json; r=json.loads(sys.stdin.read()); print('$wl N=$n $*', 'Mrays/s=%.0f'%r['value'], 'ms/step=%.2f'%r['ms_per_step'], 'spp/s=%.1f'%r['spp_per_s'], 'e2e=%.0f'%r['e2e']['value'], r['config']['parallelism'][:28], r['clocks']['sm_mhz'], r['clocks']['reasons'])"; }
: > gpurun_out/scale8_lines.jsonl
for n in ${NS:-1 4 8}; do run cornell $n 10; done
for n in ${NL:-1 8}; do run lucy256 $n 3; done
run cornell 8 10 --shard tiles
run lucy256 8 3 --shard tiles
