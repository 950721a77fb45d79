#!/bin/bash
# round 2, GPU call A: full GPU test suite, the new bench lines (c4 default, c2, c3 at both radii), hierarchy traces
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
tail -5 gpurun_out/a_pytest.log
timeout 600 python bench.py > gpurun_out/a_bench_c4.json 2> gpurun_out/a_bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --config c2 > gpurun_out/a_bench_c2.json 2> gpurun_out/a_bench_c2.err; echo "c2 rc=$?"
MGCG_TRACE=1 timeout 600 python bench.py --config c3 --re 2.1 --steps 3 > gpurun_out/a_bench_c3_r21.json 2> gpurun_out/a_bench_c3_r21.err; echo "c3 2.1 rc=$?"
MGCG_TRACE=1 timeout 900 python bench.py --config c3 --re 3.1 --steps 3 --no-cpu > gpurun_out/a_bench_c3_r31.json 2> gpurun_out/a_bench_c3_r31.err; echo "c3 3.1 rc=$?"
timeout 600 python tools/parity_c4.py 256 > gpurun_out/a_parity_c4.json 2> gpurun_out/a_parity_c4.err; echo "parity rc=$?"
cut -c1-600 gpurun_out/a_bench_c4.json gpurun_out/a_bench_c2.json gpurun_out/a_bench_c3_r21.json gpurun_out/a_bench_c3_r31.json
tail -3 gpurun_out/*.err
