#!/bin/bash
# scaling runs on an N-GPU box: bench at 256^3 (host generator, with e2e) and 512^3 (device generator), C5, phase trace
N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555"
timeout 400 $TR bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/scale_256_n$N.log 2>&1
timeout 600 $TR bench.py --gpus $N --steps 3 --warmup 3 --size 512 > gpurun_out/scale_512_n$N.log 2>&1
timeout 600 $TR tools/dist_c5.py 512 > gpurun_out/c5_512_n$N.log 2>&1
MGCG_DIST_TRACE=1 timeout 400 $TR bench.py --gpus $N --steps 1 --warmup 1 --no-e2e > gpurun_out/scale_trace_256_n$N.log 2>&1
for f in scale_256_n$N scale_512_n$N scale_trace_256_n$N; do grep '^{"metric' gpurun_out/$f.log | tail -1 | cut -c1-160; done; grep '^{"config' gpurun_out/c5_512_n$N.log; tail -3 gpurun_out/c5_512_n$N.log | cut -c1-300
