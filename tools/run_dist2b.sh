#!/bin/bash
# 2-GPU box: sharded tests, then the N=2 bench with the setup phase trace of the e2e leg
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/pytest_dist2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dist2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
timeout 300 $TR bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench2.log 2>&1
MGCG_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 2 --warmup 1 > gpurun_out/bench2_trace.log 2>&1
tail -4 gpurun_out/pytest_dist2.log; tail -1 gpurun_out/bench2.log | cut -c1-100; grep -o '"e2e": {[^}]*}' gpurun_out/bench2.log | cut -c1-300
