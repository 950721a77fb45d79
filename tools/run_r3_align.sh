#!/bin/bash
# c5 (27-point variable-coefficient 512^3) with shard cuts aligned to the aggregates (--align-planes K) against the floor rule (K = 0)
# usage: run_r3_align.sh N K...
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-8}; shift
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29560"
for k in "$@"; do
  timeout 300 $TR bench.py --gpus $N --steps 3 --warmup 3 --config c5 --no-cpu --no-e2e --no-parity --align-planes $k > gpurun_out/r03_bench_c5_n${N}_align$k.json 2> gpurun_out/r03_bench_c5_n${N}_align$k.err; echo "rc=$?"
  tail -1 gpurun_out/r03_bench_c5_n${N}_align$k.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('N $N align $k', 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), d['config']['planes_per_rank'], d['final_relative_residual'])"
done
