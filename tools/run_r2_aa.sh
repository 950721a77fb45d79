#!/bin/bash
# one GPU: generalized STENCIL kernel (runs of distances, 256- and 64-row tiles) — tests, 27-point SpMV, c4 and c5@256^3 lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_coded.py tests/test_gpu_kernels.py tests/test_gpu_mg.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/aa_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/aa_pytest.log | cut -c 1-400
show() { python - "$1" <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r=d['roofline']
print(sys.argv[1], 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'spmv ms', round(r['ms_per_launch'],4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3),
      'csr-eq', round(r.get('csr_equivalent_gbs',0)), 'iso', round(r['isolated_ms_per_launch'],4), r['kernel'][:24])
PY
}
timeout 300 python bench.py --no-cpu --no-parity --no-e2e > gpurun_out/aa_c4.json 2> gpurun_out/aa_c4.err; echo "rc=$?"; show gpurun_out/aa_c4.json
timeout 300 python tools/spmv_27pt.py 256 > gpurun_out/aa_spmv27.txt 2>&1; echo "rc=$?"; grep -v "^$" gpurun_out/aa_spmv27.txt | tail -12
for s in 2 3; do
MGCG_STENCIL_STAGES=$s timeout 400 python bench.py --config c5 --size 256 --no-cpu --no-parity --no-e2e > gpurun_out/aa_c5_256_st$s.json 2> gpurun_out/aa_c5_256_st$s.err; echo "rc=$?"; show gpurun_out/aa_c5_256_st$s.json
done
timeout 400 python bench.py --config c5 --size 256 --no-cpu --no-parity --no-e2e --opt spmv_coding=1 > gpurun_out/aa_c5_256_plain.json 2> gpurun_out/aa_c5_256_plain.err; echo "rc=$?"; show gpurun_out/aa_c5_256_plain.json
