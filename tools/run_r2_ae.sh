#!/bin/bash
# two GPUs: the C5 operator (27-point, contrast 1e6) at 192^3 sharded over 2 ranks on coded columns / STENCIL kernel, incl. bench.py's isolated-kernel leg
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 75 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 bench.py --gpus 2 --config c5 --size 192 --steps 2 --warmup 1 --no-cpu --no-parity --no-e2e > gpurun_out/ae_bench_c5_192_n2.json 2> gpurun_out/ae_bench_c5_192_n2.err; echo "rc=$?"
tail -n 1 gpurun_out/ae_bench_c5_192_n2.json | cut -c 1-1600; grep -i "error\|illegal" gpurun_out/ae_bench_c5_192_n2.err | head -3
