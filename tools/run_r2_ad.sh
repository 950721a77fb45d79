#!/bin/bash
# one GPU: STENCIL kernel with 4 resident blocks per SM (what ncu reports as the shared-memory limit) against the default 3
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
MGCG_STENCIL_CTAS=4 timeout 120 python bench.py --no-cpu --no-parity --no-e2e > gpurun_out/ad_c4_ctas4.json 2> gpurun_out/ad_c4_ctas4.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/ad_c4_ctas4.json').read().strip().splitlines()[-1]); r=d['roofline']
print('ctas4 ms', round(d['value'],2), 'it', d['iterations'], 'spmv', round(r['ms_per_launch'],4), 'frac', round(r['frac'],3), 'iso', round(r['isolated_ms_per_launch'],4))
PY
