#!/bin/bash
# one GPU: shape of the coarse tail (where the dense solve starts, aggregate size below level 0) against iterations / solve / setup
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { tag=$1; shift
  timeout 300 python bench.py --no-cpu --no-e2e --no-parity "$@" > gpurun_out/s_bench_$tag.json 2> gpurun_out/s_bench_$tag.err; echo "$tag rc=$? [$*]"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/s_bench_$tag.json').read().strip().splitlines()[-1]); print('   ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'levels', d.get('levels'), 'opc', round(d.get('operator_complexity',0),3), 'launches', d.get('gpu_launches'))
except Exception as e: print('   parse failed', e)
PY
}
run base
run cs1400 --opt mg_coarse_size=1400
run cs11000 --opt mg_coarse_size=3000
run ac3 --opt mg_aggregate_size_coarse=3
run ac3_cs1400 --opt mg_aggregate_size_coarse=3 --opt mg_coarse_size=1400
run ac4 --opt mg_aggregate_size_coarse=4
run om09 --opt mg_omega=0.9
