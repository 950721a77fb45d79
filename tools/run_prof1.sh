#!/bin/bash
# 1 GPU: tests, bench, ncu launch list, one ncu --set full capture of the fine-level SpMV kernels
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu1.log
tail -3 gpurun_out/pytest_gpu1.log
python bench.py > gpurun_out/bench1.log 2>&1; tail -1 gpurun_out/bench1.log | cut -c1-150
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --graph 0"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2200 --csv --log-file gpurun_out/launches_r01d.csv $CMD > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_spmv_tma -s 40 -c 4 -f -o gpurun_out/prof_spmv_tma $CMD > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep
