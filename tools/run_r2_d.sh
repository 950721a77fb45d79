#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python tools/spmv_sweep.py 160 > gpurun_out/d_spmv_sweep.txt 2> gpurun_out/d_spmv_sweep.err; echo "sweep rc=$?"
grep -E "==|best|vector|stream" gpurun_out/d_spmv_sweep.txt
tail -n 5 gpurun_out/d_spmv_sweep.err
