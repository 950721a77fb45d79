#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SWEEP_QUICK=1 timeout 1200 python tools/spmv_sweep.py 160 > gpurun_out/g_spmv_sweep.txt 2> gpurun_out/g_spmv_sweep.err; echo "sweep rc=$?"
grep -E "==|best|vector|stream|library|carve|rows L" gpurun_out/g_spmv_sweep.txt
tail -n 5 gpurun_out/g_spmv_sweep.err
timeout 300 python bench.py --no-cpu --no-e2e > gpurun_out/g_bench_c4.json 2> gpurun_out/g_bench_c4.err; echo "c4 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/g_bench_c4.json').read().strip().splitlines()[-1]); print('c4 ms', round(d['value'],2), 'it', d['iterations'])"
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_mg.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/g_pytest.log
