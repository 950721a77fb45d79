#!/bin/bash
# 1 GPU: what the driver runs at round end (gpu tests, smoke, bench) + the setup phase trace
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu1.log
tail -3 gpurun_out/pytest_gpu1.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
MGCG_TRACE=1 timeout 300 python tools/setup_only.py 256 2>&1 | tail -48 > gpurun_out/setup_trace_256e.log; grep -v spgemm gpurun_out/setup_trace_256e.log | tail -30
timeout 600 python bench.py > gpurun_out/bench1.log 2>&1; tail -1 gpurun_out/bench1.log | cut -c1-150; grep -o '"e2e": {[^}]*}' gpurun_out/bench1.log | cut -c1-260
