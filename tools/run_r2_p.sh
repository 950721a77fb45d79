#!/bin/bash
# one GPU: programmatic dependent launch on/off, slab allocator (cold start of a fresh process), staged owner-lane SpGEMM; tests
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mg.py tests/test_gpu_cg.py tests/test_gpu_kernels.py tests/test_gpu_parity_c2.py tests/test_gpu_mps.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -x > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 5 gpurun_out/p_pytest.log
timeout 200 python tools/cold_start.py 256 > gpurun_out/p_cold.txt 2>&1; cat gpurun_out/p_cold.txt | tail -2
MGCG_POOL_NO_SLAB=1 timeout 200 python tools/cold_start.py 256 > gpurun_out/p_cold_noslab.txt 2>&1; echo "no slab:"; tail -1 gpurun_out/p_cold_noslab.txt
MGCG_TRACE=1 timeout 200 python tools/cold_start.py 256 2>&1 | grep -E "level 0|mgcg host|owner" | head -24
for pdl in 1 0; do
  MGCG_PDL=$pdl timeout 300 python bench.py --no-cpu --no-parity > gpurun_out/p_bench_c4_pdl$pdl.json 2> gpurun_out/p_bench_c4_pdl$pdl.err; echo "c4 pdl=$pdl rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/p_bench_c4_pdl$pdl.json').read().strip().splitlines()[-1])
print('   ms', round(d['value'],3), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'cold', round(d['setup_cold_ms'],1), 'e2e', round(d['e2e']['value'],1), 'e2e cold', round(d['e2e']['e2e_cold_ms'],1), 'spmv', round(d['roofline']['achieved']), 'res', d['true_relative_residual'])
PY
done
MGCG_PDL=1 timeout 300 python bench.py --no-cpu --no-parity --no-e2e --config c2 > gpurun_out/p_bench_c2_pdl1.json 2>/dev/null; MGCG_PDL=0 timeout 300 python bench.py --no-cpu --no-parity --no-e2e --config c2 > gpurun_out/p_bench_c2_pdl0.json 2>/dev/null
python - <<PY
import json
for p in (1,0):
    d=json.loads(open(f'gpurun_out/p_bench_c2_pdl{p}.json').read().strip().splitlines()[-1]); print('c2 pdl',p,'ms', round(d['value'],3), 'it', d['iterations'])
PY
