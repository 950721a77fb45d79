#!/bin/bash
# one GPU: blocked Gauss-Jordan + dense solve from 1331 rows; the whole GPU suite; c4 / c2 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider -rf -x > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 8 gpurun_out/t_pytest.log | cut -c 1-300
MGCG_TRACE=1 timeout 300 python bench.py --no-cpu --no-parity > gpurun_out/t_bench_c4.json 2> gpurun_out/t_bench_c4.err; echo "c4 rc=$?"
grep -E "dense-inverse|mgcg host" gpurun_out/t_bench_c4.err | tail -6
python - <<PY
import json
d=json.loads(open('gpurun_out/t_bench_c4.json').read().strip().splitlines()[-1])
print('c4 ms', round(d['value'],2), 'it', d['iterations'], 'levels', d['levels'], 'setup', round(d['setup_ms'],1), 'e2e', round(d['e2e']['value'],1), 'cold', round(d['e2e']['e2e_cold_ms'],1))
PY
timeout 300 python bench.py --no-cpu --no-parity --config c2 > gpurun_out/t_bench_c2.json 2>/dev/null; python - <<PY
import json
d=json.loads(open('gpurun_out/t_bench_c2.json').read().strip().splitlines()[-1])
print('c2 ms', round(d['value'],2), 'it', d['iterations'], 'levels', d['levels'], 'setup', round(d['setup_ms'],1), 'e2e', round(d['e2e']['value'],1))
PY
timeout 300 python bench.py --no-cpu --no-parity --no-e2e --config c3 > gpurun_out/t_bench_c3.json 2>/dev/null; python - <<PY
import json
d=json.loads(open('gpurun_out/t_bench_c3.json').read().strip().splitlines()[-1])
print('c3 ms', round(d['value'],2), 'it', d['iterations'], 'levels', d['levels'], 'setup', round(d['setup_ms'],1))
PY
