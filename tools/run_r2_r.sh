#!/bin/bash
# eight GPUs: how the 256^3 loop is launched (WHILE graph / one graph per iteration / eager) and level 2 sharded instead of replicated
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { tag=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e --no-parity "$@" > gpurun_out/r_bench_$tag.json 2> gpurun_out/r_bench_$tag.err; echo "$tag rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r_bench_$tag.json').read().strip().splitlines()[-1]); print('   ms', round(d['value'],2), 'it', d['iterations'], 'graph', d.get('graph'), 'levels', d.get('levels'), 'dist levels', d.get('distributed_levels'), 'launches', d.get('gpu_launches'))
except Exception as e: print('   parse failed', e)
PY
}
run g0 --graph 0
run g1 --graph 1
run rb65k --replicate-below 65536
run rb65k_g0 --replicate-below 65536 --graph 0
run rb8k_g0 --replicate-below 8192 --graph 0
