#!/bin/bash
# 1 GPU: tests, setup trace, bench 256 (with e2e, cpu), bench 512, other configs, ncu launch list
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu1.log
tail -3 gpurun_out/pytest_gpu1.log
MGCG_TRACE=1 python tools/setup_only.py 256 2>&1 | tail -48 > gpurun_out/setup_trace_256d.log; grep -v spgemm gpurun_out/setup_trace_256d.log | tail -31
python bench.py > gpurun_out/bench1.log 2>&1; tail -1 gpurun_out/bench1.log | cut -c1-150
python bench.py --size 512 --steps 3 > gpurun_out/bench1_512.log 2>&1; tail -1 gpurun_out/bench1_512.log | cut -c1-150
timeout 600 python tools/run_configs.py c2 c5 > gpurun_out/configs.log 2>&1; tail -3 gpurun_out/configs.log
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --graph 0"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/launches_r01e.csv $CMD > gpurun_out/ncu_list.log 2>&1
