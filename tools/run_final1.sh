#!/bin/bash
# 1 GPU: what the driver runs at round end (gpu tests, smoke, bench)
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu1.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu1.log
tail -25 gpurun_out/pytest_gpu1.log | cut -c1-220
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE_OK')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench1.log 2>&1; tail -1 gpurun_out/bench1.log | cut -c1-150; grep -o '"e2e": {[^}]*}' gpurun_out/bench1.log | cut -c1-260
