#!/bin/bash
# one GPU: coded columns — tests, c4 with the layout on / off, ncu --set full of the fine-level kernels on coded and on int32 columns
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_coded.py -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/x_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/x_pytest.log | cut -c 1-300
show() { python - "$1" <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r=d['roofline']
print(sys.argv[1], 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'spmv ms', round(r['ms_per_launch'],4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3),
      'csr-eq', round(r.get('csr_equivalent_gbs',0)), 'iso', round(r['isolated_ms_per_launch'],4))
PY
}
timeout 300 python bench.py --no-cpu --no-parity --no-e2e --opt spmv_coding=1 > gpurun_out/x_c4_coded.json 2> gpurun_out/x_c4_coded.err; echo "rc=$?"; show gpurun_out/x_c4_coded.json
timeout 300 python bench.py --no-cpu --no-parity --no-e2e > gpurun_out/x_c4_plain.json 2> gpurun_out/x_c4_plain.err; echo "rc=$?"; show gpurun_out/x_c4_plain.json
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --graph 0"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_spmv_tma --launch-skip 40 --launch-count 8 -o gpurun_out/x_ncu_coded $CMD --opt spmv_coding=1 > gpurun_out/x_ncu_coded.log 2>&1; echo "ncu coded rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_spmv_tma --launch-skip 40 --launch-count 8 -o gpurun_out/x_ncu_plain $CMD > gpurun_out/x_ncu_plain.log 2>&1; echo "ncu plain rc=$?"
ls -la gpurun_out/*.ncu-rep
