#!/bin/bash
# two-GPU validation of the sharded path: parity tests, then the bench with the parity block and cold timings
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout 1500 python -m pytest tests/test_gpu_dist.py -m gpu -q --timeout 900 -p no:cacheprovider -rf -x > gpurun_out/i_pytest_dist.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/i_pytest_dist.log
MGCG_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/i_bench_n2.json 2> gpurun_out/i_bench_n2.err; echo "bench2 rc=$?"
tail -n 1 gpurun_out/i_bench_n2.json | cut -c 1-1800
grep -E "setup\]|d:" gpurun_out/i_bench_n2.err | head -50
tail -n 5 gpurun_out/i_bench_n2.err
