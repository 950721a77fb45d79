#!/bin/bash
# two GPUs: the sharded suite (with the coded-column cases), c4 and c5 lines at N=2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 1500 python -m pytest tests/test_gpu_dist.py tests/test_gpu_coded.py -m gpu -q --timeout 900 -p no:cacheprovider -rf > gpurun_out/ab_pytest_dist.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/ab_pytest_dist.log | cut -c 1-600
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577 bench.py --gpus 2 "$@"; }
run --no-cpu > gpurun_out/ab_bench_c4_n2.json 2> gpurun_out/ab_bench_c4_n2.err; echo "c4 n2 rc=$?"; tail -n 1 gpurun_out/ab_bench_c4_n2.json | cut -c 1-1800
run --no-cpu --no-parity --no-e2e --config c5 > gpurun_out/ab_bench_c5_n2.json 2> gpurun_out/ab_bench_c5_n2.err; echo "c5 n2 rc=$?"; tail -n 1 gpurun_out/ab_bench_c5_n2.json | cut -c 1-1500
