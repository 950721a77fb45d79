#!/bin/bash
# one GPU: the WIDE TMA stream kernel on the 27-point operator, kernel / multigrid tests, c4 line, host-path phase trace
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/spmv_27pt.py 256 > gpurun_out/n_spmv27.txt 2>&1; echo "spmv27 rc=$?"; cat gpurun_out/n_spmv27.txt
MGCG_SPMV_NO_WIDE=1 timeout 300 python tools/spmv_27pt.py 256 128 > gpurun_out/n_spmv27_nowide.txt 2>&1; grep -E "auto" gpurun_out/n_spmv27_nowide.txt
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_mg.py tests/test_gpu_cg.py tests/test_gpu_parity_c2.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/n_pytest.log
MGCG_TRACE=1 timeout 300 python bench.py --no-cpu --no-parity > gpurun_out/n_bench_c4.json 2> gpurun_out/n_bench_c4.err; echo "c4 rc=$?"
grep "mgcg host" gpurun_out/n_bench_c4.err | tail -8
python - <<PY
import json
d=json.loads(open('gpurun_out/n_bench_c4.json').read().strip().splitlines()[-1])
print('c4 ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'e2e', round(d['e2e']['value'],1), 'cold', round(d['e2e']['e2e_cold_ms'],1))
PY
MGCG_SPMV_NO_WIDE=1 timeout 300 python bench.py --no-cpu --no-parity --no-e2e > gpurun_out/n_bench_c4_nowide.json 2> /dev/null; python - <<PY
import json
d=json.loads(open('gpurun_out/n_bench_c4_nowide.json').read().strip().splitlines()[-1])
print('c4 no-wide ms', round(d['value'],2), 'it', d['iterations'])
PY
