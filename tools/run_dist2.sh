#!/bin/bash
# 2-GPU box: all gpu tests, bench at N=1 and N=2
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu2.log
timeout 600 python bench.py > gpurun_out/bench1.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
timeout 300 $TR bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench2.log 2>&1
timeout 300 $TR bench.py --gpus 2 --steps 5 --warmup 3 --device-gen > gpurun_out/bench2_devgen.log 2>&1
tail -3 gpurun_out/pytest_gpu2.log; for f in bench1 bench2 bench2_devgen; do tail -1 gpurun_out/$f.log | cut -c1-120; done
