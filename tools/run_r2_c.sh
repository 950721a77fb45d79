#!/bin/bash
# round 2, GPU call C: rows kernel v2 (runtime stage ring), 8-list warp-merge SpGEMM, C3 parity at 64^3 particles
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_mg.py tests/test_gpu_mps.py tests/test_gpu_formats.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -s > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
grep -E "C3 64|passed|failed|rc=" gpurun_out/c_pytest.log | tail -n 12
timeout 600 python bench.py --config c3 --re 2.1 --no-cpu > gpurun_out/c_bench_c3_r21.json 2> gpurun_out/c_bench_c3_r21.err; echo "c3 2.1 rc=$?"
MGCG_TRACE=1 timeout 900 python bench.py --config c3 --re 3.1 --no-cpu > gpurun_out/c_bench_c3_r31.json 2> gpurun_out/c_bench_c3_r31.err; echo "c3 3.1 rc=$?"
timeout 600 python bench.py --no-cpu --no-e2e > gpurun_out/c_bench_c4.json 2> gpurun_out/c_bench_c4.err; echo "c4 rc=$?"
python - <<'PY'
import json
for f in ("c_bench_c3_r21","c_bench_c3_r31","c_bench_c4"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        r=d["roofline"]
        print(f, "ms", round(d["value"],2), "it", d["iterations"], "setup", round(d["setup_ms"],1), "spmv frac", round(r["frac"],3), "iso", round(r["isolated_gbs"]), r["kernel"][:24])
    except Exception as e:
        print(f, "ERR", e)
PY
for f in gpurun_out/c_*.err; do echo "== $f"; grep -v "mgcg" $f | tail -n 3; done
