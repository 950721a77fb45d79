#!/bin/bash
# two GPUs: sharded parity tests (fused halo exchange on by default), then the c4 bench fused / unfused with the phase trace,
# the opt-in modes, and c5 (27-point 512^3) on two GPUs
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | head -4
timeout 1200 python -m pytest tests/test_gpu_dist.py -m gpu -q --timeout 600 -p no:cacheprovider -rf > gpurun_out/m_pytest_dist.log 2>&1; echo "pytest rc=$?"
tail -n 25 gpurun_out/m_pytest_dist.log | cut -c 1-600
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
run() { # tag, env, extra args
  tag=$1; shift; envs=$1; shift
  env $envs timeout 400 $TR bench.py --gpus 2 --steps 5 --warmup 3 "$@" > gpurun_out/m_bench_$tag.json 2> gpurun_out/m_bench_$tag.err; echo "bench $tag rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/m_bench_$tag.json').read().strip().splitlines()[-1])
    print('   ms', round(d['value'],2), 'it', d['iterations'], 'setup', d.get('setup_ms'), 'cold', d.get('setup_cold_ms'), 'e2e', d.get('e2e',{}) and round(d['e2e']['value'],1), 'parity', d.get('parity',{}) and d['parity'].get('dist_x_rel_err'), 'xchg', d.get('exchange_us'))
except Exception as e: print('   parse failed', e)
PY
}
run c4_fused "A=1"
run c4_unfused "MGCG_DIST_FUSED=0" --no-e2e --no-parity
run c4_fused_rb65k "A=1" --no-e2e --no-parity --replicate-below 65536
run c4_fused_cgcg "A=1" --no-e2e --opt cg_variant=1
run c4_fused_cheby "A=1" --no-e2e --opt mg_smoother=2
MGCG_DIST_TRACE=1 timeout 300 $TR bench.py --gpus 2 --steps 2 --warmup 1 --no-e2e --no-parity > gpurun_out/m_trace_fused.json 2> gpurun_out/m_trace_fused.err; echo "trace rc=$?"
grep -A 40 "dist trace" gpurun_out/m_trace_fused.err | head -45
run c5_n2 "A=1" --config c5 --steps 2 --warmup 1 --no-e2e
tail -n 3 gpurun_out/m_bench_c5_n2.err
