#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python tools/spmv_sweep.py 160 > gpurun_out/d_spmv_sweep.txt 2> gpurun_out/d_spmv_sweep.err; echo "sweep rc=$?"
grep -E "==|best|vector|stream" gpurun_out/d_spmv_sweep.txt
tail -n 5 gpurun_out/d_spmv_sweep.err
timeout 900 python -m pytest tests/test_gpu_mg.py tests/test_gpu_kernels.py -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
tail -n 8 gpurun_out/d_pytest.log
for sm in 0 1; do timeout 300 python bench.py --no-cpu --no-e2e --smoother $sm > gpurun_out/d_bench_c4_sm$sm.json 2> gpurun_out/d_bench_c4_sm$sm.err; echo "c4 smoother $sm rc=$?"; done
python - <<'PY'
import json
for f in ("d_bench_c4_sm0","d_bench_c4_sm1"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "ms", round(d["value"],2), "it", d["iterations"], "setup", round(d["setup_ms"],1), "launches/solve", d["gpu_launches"]/d["steps"])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -n 3 gpurun_out/d_bench_c4_sm1.err
