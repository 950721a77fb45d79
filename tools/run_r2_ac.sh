#!/bin/bash
# one GPU, final build: the whole GPU suite, c4 / c2 / c5@256^3 lines, launch list, ncu capture of the fine-level kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/ac_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 6 gpurun_out/ac_pytest.log | cut -c 1-400
timeout 300 python bench.py > gpurun_out/ac_bench_c4.json 2> gpurun_out/ac_bench_c4.err; echo "c4 rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/ac_bench_c4.json').read().strip().splitlines()[-1]); r=d['roofline']; e=d['e2e']
print('c4 ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'spmv', round(r['ms_per_launch'],4), 'frac', round(r['frac'],3), 'traffic', r['traffic'], 'e2e', round(e['value'],1), 'cold', round(e['e2e_cold_ms'],1), 'cpu', d['cpu_baseline']['value'])
PY
timeout 200 python bench.py --config c2 --no-cpu > gpurun_out/ac_bench_c2.json 2> gpurun_out/ac_bench_c2.err; echo "c2 rc=$?"
timeout 200 python bench.py --config c5 --size 256 --no-cpu --no-e2e > gpurun_out/ac_bench_c5_256.json 2> gpurun_out/ac_bench_c5_256.err; echo "c5@256 rc=$?"
python - <<PY
import json
for f in ('ac_bench_c2','ac_bench_c5_256'):
    d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1]); r=d['roofline']
    print(f, 'ms', round(d['value'],2), 'it', d['iterations'], 'setup', round(d['setup_ms'],1), 'spmv', round(r['ms_per_launch'],4), 'frac', round(r['frac'],3), (d.get('e2e') or {}).get('value'))
PY
CMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity --graph 0"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/ac_launches_c4.csv $CMD > gpurun_out/ac_ncu_list.log 2>&1; echo "ncu list rc=$?"
python tools/launch_summary.py gpurun_out/ac_launches_c4.csv 40 > gpurun_out/ac_launches_c4.txt; head -n 14 gpurun_out/ac_launches_c4.txt
ncu --set full --clock-control none --import-source on -k regex:k_spmv_stencil --launch-skip 12 --launch-count 6 -o gpurun_out/ac_ncu_full_c4 $CMD > gpurun_out/ac_ncu_full.log 2>&1; echo "ncu full rc=$?"
