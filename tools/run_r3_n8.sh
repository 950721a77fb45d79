#!/bin/bash
# 8-GPU box, round 3: world-4 / world-8 dist cases with the oracle leg, c4 at 8 ranks (full line with dist_vs_twin), 512^3 7-point and
# c5 at 8 ranks, the sharded set-up phase trace, and the coarse-tail options of VERDICT r02 item 3 measured at 8 ranks.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29560"
show() { tail -1 "$1" | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read())
    p=d.get('parity') or {}
    print('$1', 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'e2e', (d.get('e2e') or {}).get('value'), 'dist_x', p.get('dist_x_rel_err'), 'twin', {k:(v.get('it_dist'),v.get('it_twin'),v.get('x_rel_err')) for k,v in (p.get('dist_vs_twin') or {}).items() if isinstance(v,dict)})
except Exception as e: print('$1', 'unreadable', e)
"; }
timeout 500 python -m pytest tests/test_gpu_dist.py -m gpu -q -k "w4 or w8" > gpurun_out/r03_pytest_dist_n8box.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r03_pytest_dist_n8box.log
timeout 300 $TR bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r03_bench_c4_n8.json 2> gpurun_out/r03_bench_c4_n8.err; echo "rc=$?"; show gpurun_out/r03_bench_c4_n8.json
for v in "rb65536 --replicate-below 65536" "rb8192 --replicate-below 8192" "sl1 --opt mg_smoothed_levels=1" "cgcg --opt cg_variant=1"; do
  set -- $v; tag=$1; shift
  timeout 200 $TR bench.py --gpus 8 --steps 5 --warmup 3 --no-parity --no-e2e --no-cpu "$@" > gpurun_out/r03_bench_c4_n8_$tag.json 2> gpurun_out/r03_bench_c4_n8_$tag.err; echo "rc=$?"; show gpurun_out/r03_bench_c4_n8_$tag.json
done
MGCG_TRACE=1 timeout 200 $TR bench.py --gpus 8 --steps 1 --warmup 1 --no-parity --no-e2e --no-cpu > gpurun_out/r03_trace_n8.json 2> gpurun_out/r03_setup_trace_256_n8_raw.txt; echo "trace rc=$?"
timeout 300 $TR bench.py --gpus 8 --steps 3 --warmup 3 --size 512 --no-cpu > gpurun_out/r03_bench_512_n8.json 2> gpurun_out/r03_bench_512_n8.err; echo "rc=$?"; show gpurun_out/r03_bench_512_n8.json
timeout 400 $TR bench.py --gpus 8 --steps 3 --warmup 3 --config c5 --no-cpu > gpurun_out/r03_bench_c5_n8.json 2> gpurun_out/r03_bench_c5_n8.err; echo "rc=$?"; show gpurun_out/r03_bench_c5_n8.json
