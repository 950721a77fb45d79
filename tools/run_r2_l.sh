#!/bin/bash
# one GPU: the opt-in modes (Chebyshev smoother, single-reduce CG) — their tests, then c4 / c2 bench lines in each mode; e2e after the
# vector-upload overlap
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mg.py tests/test_gpu_cg.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/l_pytest.log
for mode in "" "--opt mg_smoother=2" "--opt mg_smoother=2 --opt mg_cheby_degree=3" "--opt cg_variant=1" "--opt mg_smoother=2 --opt cg_variant=1" "--opt mg_smoother=1"; do
  tag=$(echo "$mode" | tr -d ' -' | tr '=' '_'); [ -z "$tag" ] && tag=default
  timeout 300 python bench.py --no-cpu --no-parity $mode > gpurun_out/l_bench_c4_$tag.json 2> gpurun_out/l_bench_c4_$tag.err; echo "c4 [$mode] rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/l_bench_c4_$tag.json').read().strip().splitlines()[-1])
    print('   ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'e2e', round(d['e2e']['value'],1) if d.get('e2e') else None, 'res', d['true_relative_residual'])
except Exception as e: print('   parse failed', e)
PY
done
