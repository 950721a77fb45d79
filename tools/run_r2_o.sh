#!/bin/bash
# one GPU: WIDE stream kernel vs ROWS on the 27-point operator (two sizes), allocation cost model, owner-lane SpGEMM (bit-exact tests,
# setup trace), c4 line
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/spmv_27pt.py 256 > gpurun_out/o_spmv27_256.txt 2>&1; echo "spmv27 rc=$?"; cat gpurun_out/o_spmv27_256.txt
timeout 300 python tools/spmv_27pt.py 86 > gpurun_out/o_spmv27_86.txt 2>&1; cat gpurun_out/o_spmv27_86.txt
timeout 120 python tools/malloc_cost.py 0 > gpurun_out/o_malloc_0.txt 2>&1; cat gpurun_out/o_malloc_0.txt
timeout 120 python tools/malloc_cost.py 40 > gpurun_out/o_malloc_40.txt 2>&1; cat gpurun_out/o_malloc_40.txt
timeout 900 python -m pytest tests/test_gpu_mg.py tests/test_gpu_parity_c2.py tests/test_gpu_kernels.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/o_pytest.log
MGCG_TRACE=1 timeout 300 python bench.py --no-cpu --no-parity > gpurun_out/o_bench_c4.json 2> gpurun_out/o_bench_c4.err; echo "c4 rc=$?"
grep -E "mgcg host|level 0|owner|level 1 R" gpurun_out/o_bench_c4.err | tail -14
python - <<PY
import json
d=json.loads(open('gpurun_out/o_bench_c4.json').read().strip().splitlines()[-1])
print('c4 ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'e2e', round(d['e2e']['value'],1), 'cold', round(d['e2e']['e2e_cold_ms'],1))
PY
