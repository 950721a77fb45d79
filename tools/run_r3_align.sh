#!/bin/bash
# 8-GPU box: c5 (27-point variable-coefficient 512^3) with shard cuts aligned to the aggregates (--align-planes 6 / 12) against the floor rule
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29560"
for k in 6 12; do
  timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --config c5 --no-cpu --no-e2e --no-parity --align-planes $k > gpurun_out/r03_bench_c5_n8_align$k.json 2> gpurun_out/r03_bench_c5_n8_align$k.err; echo "rc=$?"
  tail -1 gpurun_out/r03_bench_c5_n8_align$k.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('align $k', 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), d['config']['planes_per_rank'], d['final_relative_residual'])"
done
