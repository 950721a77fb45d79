#!/bin/bash
# one GPU: STENCIL kernel (coded columns, x runs staged by TMA) — tests, c4 stage sweep against the coded gather kernel and int32 columns
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_coded.py -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/y_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/y_pytest.log | cut -c 1-400
show() { python - "$1" <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
r=d['roofline']
print(sys.argv[1], 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'spmv ms', round(r['ms_per_launch'],4), 'GB/s', round(r['achieved']), 'frac', round(r['frac'],3),
      'csr-eq', round(r.get('csr_equivalent_gbs',0)), 'iso', round(r['isolated_ms_per_launch'],4))
PY
}
for s in 3 2 4; do
  MGCG_STENCIL_STAGES=$s timeout 300 python bench.py --no-cpu --no-parity --no-e2e --opt spmv_coding=1 > gpurun_out/y_c4_st$s.json 2> gpurun_out/y_c4_st$s.err; echo "rc=$?"; show gpurun_out/y_c4_st$s.json
done
MGCG_STENCIL_STAGES=2 MGCG_STENCIL_CTAS=2 timeout 300 python bench.py --no-cpu --no-parity --no-e2e --opt spmv_coding=1 > gpurun_out/y_c4_st2n2.json 2> gpurun_out/y_c4_st2n2.err; echo "rc=$?"; show gpurun_out/y_c4_st2n2.json
MGCG_NO_STENCIL=1 timeout 300 python bench.py --no-cpu --no-parity --no-e2e --opt spmv_coding=1 > gpurun_out/y_c4_gather.json 2> gpurun_out/y_c4_gather.err; echo "rc=$?"; show gpurun_out/y_c4_gather.json
timeout 300 python bench.py --no-cpu --no-parity --no-e2e > gpurun_out/y_c4_plain.json 2> gpurun_out/y_c4_plain.err; echo "rc=$?"; show gpurun_out/y_c4_plain.json
timeout 300 python bench.py --no-cpu --no-parity --no-e2e --config c2 --opt spmv_coding=1 > gpurun_out/y_c2_st.json 2> gpurun_out/y_c2_st.err; echo "rc=$?"; show gpurun_out/y_c2_st.json
timeout 300 python bench.py --no-cpu --no-parity --no-e2e --config c2 > gpurun_out/y_c2_plain.json 2> gpurun_out/y_c2_plain.err; echo "rc=$?"; show gpurun_out/y_c2_plain.json
