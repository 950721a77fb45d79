#!/bin/bash
# 2-GPU check of the sharded path: parity tests, phase traces, then the bench
set -x
timeout 600 python -m pytest tests/test_gpu_dist.py -x -q > gpurun_out/pytest_dist2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_dist2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
MGCG_DIST_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 1 --warmup 1 --no-e2e --transport 2 > gpurun_out/trace2_peer.log 2>&1
MGCG_DIST_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 1 --warmup 1 --no-e2e --transport 2 --replicate-below 200000 > gpurun_out/trace2_peer_rb200k.log 2>&1
MGCG_DIST_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 1 --warmup 1 --no-e2e --transport 1 > gpurun_out/trace2_nccl.log 2>&1
timeout 300 $TR bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e --transport 2 > gpurun_out/bench2_peer.log 2>&1
tail -3 gpurun_out/pytest_dist2.log; tail -1 gpurun_out/bench2_peer.log | cut -c1-200
