#!/bin/bash
# round 2, GPU call B: full GPU test suite (failure summary), bench lines c4 / c2 / c3 (rows kernel), 256^3 twin parity
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -n 15 gpurun_out/b_pytest.log
timeout 600 python bench.py > gpurun_out/b_bench_c4.json 2> gpurun_out/b_bench_c4.err; echo "c4 rc=$?"
timeout 300 python bench.py --config c2 > gpurun_out/b_bench_c2.json 2> gpurun_out/b_bench_c2.err; echo "c2 rc=$?"
timeout 600 python bench.py --config c3 --re 2.1 > gpurun_out/b_bench_c3_r21.json 2> gpurun_out/b_bench_c3_r21.err; echo "c3 2.1 rc=$?"
MGCG_TRACE=1 timeout 900 python bench.py --config c3 --re 3.1 --no-cpu > gpurun_out/b_bench_c3_r31.json 2> gpurun_out/b_bench_c3_r31.err; echo "c3 3.1 rc=$?"
MGCG_SPMV_NO_ROWS=1 timeout 600 python bench.py --no-cpu --no-e2e > gpurun_out/b_bench_c4_norows.json 2> gpurun_out/b_bench_c4_norows.err; echo "c4 norows rc=$?"
timeout 600 python tools/parity_c4.py 256 > gpurun_out/b_parity_c4.json 2> gpurun_out/b_parity_c4.err; echo "parity rc=$?"
for f in gpurun_out/b_bench_c4.json gpurun_out/b_bench_c2.json gpurun_out/b_bench_c3_r21.json gpurun_out/b_bench_c3_r31.json gpurun_out/b_bench_c4_norows.json gpurun_out/b_parity_c4.json; do cut -c1-300 $f; done
for f in gpurun_out/b_*.err; do echo "== $f"; tail -n 3 $f; done
