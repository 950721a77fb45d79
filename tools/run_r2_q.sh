#!/bin/bash
# eight GPUs: world-4 parity tests of the sharded path, strong scaling of c4 at 1/2/4/8 (fused halo exchange; unfused at 8 beside it),
# phase trace at 8, c5 (27-point 512^3) at 8
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | wc -l; nproc
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q --timeout 600 -p no:cacheprovider -rf -k "w4 or 27pt_chebyshev" > gpurun_out/q_pytest_dist.log 2>&1; echo "pytest rc=$?"
tail -n 12 gpurun_out/q_pytest_dist.log | cut -c 1-400
summ() { python - "$1" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    p=d.get('parity') or {}
    e=d.get('e2e') or {}
    print('   ms', round(d['value'],2), 'it', d['iterations'], 'setup', d.get('setup_ms') and round(d['setup_ms'],1), 'setup cold', d.get('setup_cold_ms') and round(d['setup_cold_ms'],1), 'comm init', d.get('comm_init_ms') and round(d['comm_init_ms'],1), 'e2e', e.get('value') and round(e['value'],1), 'e2e cold', e.get('e2e_cold_ms') and round(e['e2e_cold_ms'],1), 'dist_x_rel_err', p.get('dist_x_rel_err'), 'at1e-8', p.get('dist_x_rel_err_at_1e-8'), 'spmv frac', d['roofline'].get('frac') and round(d['roofline']['frac'],3), 'xchg', d.get('exchange_us'))
except Exception as ex: print('   parse failed', ex)
PY
}
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e --no-parity > gpurun_out/q_bench_c4_n1.json 2> gpurun_out/q_bench_c4_n1.err; echo "c4 n1 rc=$?"; summ gpurun_out/q_bench_c4_n1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29554 bench.py --gpus 4 --steps 5 --warmup 3 --no-e2e > gpurun_out/q_bench_c4_n4.json 2> gpurun_out/q_bench_c4_n4.err; echo "c4 n4 rc=$?"; summ gpurun_out/q_bench_c4_n4.json
for n in 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/q_bench_c4_n$n.json 2> gpurun_out/q_bench_c4_n$n.err; echo "c4 n$n rc=$?"; summ gpurun_out/q_bench_c4_n$n.json
done
MGCG_DIST_FUSED=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e --no-parity > gpurun_out/q_bench_c4_n8_unfused.json 2> gpurun_out/q_bench_c4_n8_unfused.err; echo "c4 n8 unfused rc=$?"; summ gpurun_out/q_bench_c4_n8_unfused.json
MGCG_DIST_TRACE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29564 bench.py --gpus 8 --steps 2 --warmup 1 --no-e2e --no-parity > gpurun_out/q_trace_n8.json 2> gpurun_out/q_trace_n8.err; echo "trace rc=$?"
grep -A 34 "dist trace" gpurun_out/q_trace_n8.err | tail -35
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29565 bench.py --gpus 8 --config c5 --steps 3 --warmup 2 --no-e2e > gpurun_out/q_bench_c5_n8.json 2> gpurun_out/q_bench_c5_n8.err; echo "c5 n8 rc=$?"; summ gpurun_out/q_bench_c5_n8.json
tail -n 2 gpurun_out/q_bench_c5_n8.err | cut -c 1-300
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29566 bench.py --gpus 8 --size 512 --steps 2 --warmup 1 --no-e2e --no-parity > gpurun_out/q_bench_512_n8.json 2> gpurun_out/q_bench_512_n8.err; echo "512 n8 rc=$?"; summ gpurun_out/q_bench_512_n8.json
timeout 300 python bench.py --size 512 --steps 2 --warmup 1 --no-cpu --no-e2e --no-parity > gpurun_out/q_bench_512_n1.json 2> gpurun_out/q_bench_512_n1.err; echo "512 n1 rc=$?"; summ gpurun_out/q_bench_512_n1.json
