#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python tools/spmv_sweep.py 160 > gpurun_out/e_spmv_sweep.txt 2> gpurun_out/e_spmv_sweep.err; echo "sweep rc=$?"
grep -E "==|best|vector|stream|library|carve" gpurun_out/e_spmv_sweep.txt
tail -n 5 gpurun_out/e_spmv_sweep.err
timeout 300 python bench.py --no-cpu --no-e2e > gpurun_out/e_bench_c4.json 2> gpurun_out/e_bench_c4.err; echo "c4 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/e_bench_c4.json').read().strip().splitlines()[-1]); print('c4 ms', round(d['value'],2), 'it', d['iterations'])"
python tools/profile_spmv.py 2.1 128 3 > gpurun_out/e_prof_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_spmv_rows -s 2 -c 2 -o gpurun_out/e_ncu_rows_r21 python tools/profile_spmv.py 2.1 128 3 > gpurun_out/e_ncu_r21.log 2>&1; echo "ncu r21 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_spmv_rows -s 2 -c 2 -o gpurun_out/e_ncu_rows_r31 python tools/profile_spmv.py 3.1 112 3 > gpurun_out/e_ncu_r31.log 2>&1; echo "ncu r31 rc=$?"
tail -n 3 gpurun_out/e_ncu_r21.log gpurun_out/e_ncu_r31.log
ls -la gpurun_out/*.ncu-rep
